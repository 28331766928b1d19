"""ctypes binding of build/libbert.so -- the host-side mirror of the reference's Python client
(/root/reference/examples/sample_dylib.py:17-57, benchmarks/run_mteb.py:39-89): same library symbols,
same call sequence (bert_load_from_file -> bert_n_embd -> bert_encode_batch / bert_eval_batch -> bert_free).

The library is the product; this file only declares argtypes and moves NumPy buffers.  It raises when
the shared library is missing or when no CUDA device is usable -- there is no CPU path behind it.
"""
from __future__ import annotations

import ctypes
import os

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("BERT_B200_LIB") or os.path.join(ROOT, "build", "libbert.so")   # BERT_B200_LIB: another build of the library (A/B of compile-time knobs)

_i32p = ctypes.POINTER(ctypes.c_int32)
_f32p = ctypes.POINTER(ctypes.c_float)
_u16p = ctypes.POINTER(ctypes.c_uint16)

BK_F32, BK_F16, BK_Q4_0, BK_Q4_1 = 0, 1, 2, 3
EPI_BIAS, EPI_BIAS_RESID, EPI_BIAS_GELU, EPI_BIAS_GELU_Q8_0, EPI_BIAS_GELU_Q8_1 = range(5)
KERNEL_NAMES = ["pack", "embed_ln", "gemm_qkv", "attention", "gemm_o", "add_ln", "gemm_ff1", "gemm_ff2", "pool_norm"]

BERT_SYMBOLS = ["bert_load_from_file", "bert_free", "bert_encode", "bert_encode_batch", "bert_tokenize", "bert_eval",
                "bert_eval_batch", "bert_n_embd", "bert_n_max_tokens", "bert_vocab_id_to_token", "bert_model_quantize",
                "bert_params_parse"]
BK_SYMBOLS = ["bk_device_count", "bk_last_error", "bk_engine_create", "bk_engine_destroy", "bk_engine_set_tensor",
              "bk_engine_finalize", "bk_engine_forward", "bk_engine_submit", "bk_engine_collect", "bk_engine_submit_ragged", "bk_host_alloc", "bk_host_free",
              "bk_engine_graph_stats", "bk_bert_graph_stats", "bk_engine_stage", "bk_engine_run_staged", "bk_engine_read_output",
              "bk_engine_kernel_ms", "bk_engine_read_hidden", "bk_engine_set_keep_hidden", "bk_engine_launch_count", "bk_engine_check_guards", "bk_profiler_range",
              "bk_test_gemm", "bk_test_gemm_ln", "bk_test_gelu", "bk_test_add_ln", "bk_test_attention", "bk_test_attention_packed", "bk_test_pool", "bk_test_batcher", "bk_test_embed_ln", "bk_batcher_static_scratch",
              "bk_tokenizer_create", "bk_tokenizer_destroy", "bk_tokenizer_encode"]


class bk_hparams(ctypes.Structure):
    _fields_ = [("n_vocab", ctypes.c_int32), ("n_max_tokens", ctypes.c_int32), ("n_embd", ctypes.c_int32),
                ("n_intermediate", ctypes.c_int32), ("n_head", ctypes.c_int32), ("n_layer", ctypes.c_int32),
                ("eps", ctypes.c_float), ("wtype", ctypes.c_int32)]


_lib = None


def load_library(path: str | None = None):
    """dlopen build/libbert.so and declare every exported signature."""
    global _lib
    if _lib is not None and path is None:
        return _lib
    p = path or LIB_PATH
    if not os.path.exists(p):
        raise FileNotFoundError(f"{p} not built: run `make` (or __graft_entry__.build()) first")
    L = ctypes.CDLL(p)
    vp, i32, i64 = ctypes.c_void_p, ctypes.c_int32, ctypes.c_int64
    # bert.h
    L.bert_load_from_file.restype = vp
    L.bert_load_from_file.argtypes = [ctypes.c_char_p]
    L.bert_free.restype = None
    L.bert_free.argtypes = [vp]
    L.bert_n_embd.restype = i32
    L.bert_n_embd.argtypes = [vp]
    L.bert_n_max_tokens.restype = i32
    L.bert_n_max_tokens.argtypes = [vp]
    L.bert_vocab_id_to_token.restype = ctypes.c_char_p
    L.bert_vocab_id_to_token.argtypes = [vp, i32]
    L.bert_eval.restype = None
    L.bert_eval.argtypes = [vp, i32, _i32p, i32, _f32p]
    L.bert_eval_batch.restype = None
    L.bert_eval_batch.argtypes = [vp, i32, i32, ctypes.POINTER(_i32p), _i32p, ctypes.POINTER(_f32p)]
    L.bert_tokenize.restype = None
    L.bert_tokenize.argtypes = [vp, ctypes.c_char_p, _i32p, _i32p, i32]
    L.bert_encode.restype = None
    L.bert_encode.argtypes = [vp, i32, ctypes.c_char_p, _f32p]
    L.bert_encode_batch.restype = None
    L.bert_encode_batch.argtypes = [vp, i32, i32, i32, ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(_f32p)]
    L.bert_model_quantize.restype = ctypes.c_bool
    L.bert_model_quantize.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]
    # bk_api.h
    L.bk_device_count.restype = ctypes.c_int
    L.bk_last_error.restype = ctypes.c_char_p
    L.bk_engine_create.restype = vp
    L.bk_engine_create.argtypes = [ctypes.c_int, ctypes.POINTER(bk_hparams), i32, i32]
    L.bk_engine_destroy.restype = None
    L.bk_engine_destroy.argtypes = [vp]
    L.bk_engine_set_tensor.argtypes = [vp, i32, i32, vp, i64, i64]
    L.bk_engine_finalize.argtypes = [vp]
    L.bk_engine_forward.argtypes = [vp, _i32p, _i32p, i32, _f32p]
    L.bk_engine_submit.argtypes = [vp, i32, _i32p, _i32p, i32]
    L.bk_engine_collect.argtypes = [vp, i32, _f32p]
    L.bk_engine_submit_ragged.argtypes = [vp, i32, ctypes.POINTER(_i32p), _i32p, i32, _f32p]
    L.bk_host_alloc.restype = vp
    L.bk_host_alloc.argtypes = [ctypes.c_size_t]
    L.bk_host_free.restype = None
    L.bk_host_free.argtypes = [vp]
    _i64p = ctypes.POINTER(ctypes.c_int64)
    L.bk_engine_graph_stats.argtypes = [vp, _i64p, _i64p, _i64p]
    L.bk_bert_graph_stats.argtypes = [vp, _i64p, _i64p, _i64p]
    L.bk_engine_stage.argtypes = [vp, _i32p, _i32p, i32]
    L.bk_engine_run_staged.argtypes = [vp, i32, i32, _f32p]
    L.bk_engine_read_output.argtypes = [vp, _f32p, i32]
    L.bk_engine_kernel_ms.argtypes = [vp, _f32p, _i32p]
    L.bk_engine_read_hidden.argtypes = [vp, i32, _f32p]
    L.bk_engine_set_keep_hidden.argtypes = [vp, i32]
    L.bk_engine_launch_count.restype = i64
    L.bk_engine_launch_count.argtypes = [vp]
    L.bk_engine_check_guards.restype = i64
    L.bk_engine_check_guards.argtypes = [vp]
    L.bk_profiler_range.argtypes = [i32]
    L.bk_test_gemm.argtypes = [ctypes.c_int, i32, i32, i32, _u16p, vp, i32, _f32p, _u16p, i32, _u16p, _f32p]
    L.bk_test_gemm_ln.argtypes = [ctypes.c_int, i32, i32, i32, _u16p, _u16p, _f32p, _u16p, _f32p, _f32p, ctypes.c_float, i32, _u16p, _u16p, _f32p]
    L.bk_test_gelu.argtypes = [ctypes.c_int, i32, _f32p, _f32p]
    L.bk_test_add_ln.argtypes = [ctypes.c_int, i32, i32, ctypes.c_float, _u16p, _f32p, _f32p, i32, _u16p, _u16p]
    L.bk_test_attention.argtypes = [ctypes.c_int, i32, _i32p, i32, i32, _u16p, i32, _u16p]
    L.bk_test_attention_packed.argtypes = [ctypes.c_int, i32, i32, i32, _i32p, i32, i32, _u16p, i32, _u16p]
    L.bk_test_pool.argtypes = [ctypes.c_int, i32, _i32p, i32, _u16p, _f32p]
    L.bk_test_batcher.argtypes = [ctypes.c_int, i32, _i32p, _i32p, _i32p, _i32p, _i32p, _i32p]
    L.bk_test_embed_ln.argtypes = [ctypes.c_int, i32, i32, i32, i32, i32, vp, vp, vp, _f32p, _f32p, ctypes.c_float, _i32p, _i32p, i32, _u16p, _u16p]
    L.bk_batcher_static_scratch.argtypes = [ctypes.c_int, i32, i32, vp, vp, vp, vp, vp, vp, vp]  # device pointers + stream
    L.bk_tokenizer_create.restype = vp
    L.bk_tokenizer_create.argtypes = [ctypes.c_char_p, i64]
    L.bk_tokenizer_destroy.restype = None
    L.bk_tokenizer_destroy.argtypes = [vp]
    L.bk_tokenizer_encode.restype = i32
    L.bk_tokenizer_encode.argtypes = [vp, ctypes.c_char_p, _i32p, i32]
    if path is None:
        _lib = L
    return L


class Tokenizer:
    """The C++ WordPiece tokenizer alone (host only)."""

    def __init__(self, tokenizer_json: str, lib=None):
        self.lib = lib or load_library()
        b = tokenizer_json.encode("utf-8")
        self.t = self.lib.bk_tokenizer_create(b, len(b))
        if not self.t:
            raise RuntimeError("bk_tokenizer_create failed")

    def encode(self, text: str):
        buf = np.zeros(max(16, 4 * len(text) + 16), np.int32)
        n = self.lib.bk_tokenizer_encode(self.t, text.encode("utf-8"), _ptr(buf, _i32p), len(buf))
        assert 0 <= n <= len(buf)
        return buf[:n].tolist()

    def close(self):
        if self.t:
            self.lib.bk_tokenizer_destroy(self.t)
            self.t = None


def _ptr(a: np.ndarray, t):
    return a.ctypes.data_as(t)


class BertB200:
    """bert.h through ctypes, NumPy in/out (what sample_dylib.py's BertModel does for the reference)."""

    def __init__(self, fname: str, lib=None):
        self.lib = lib or load_library()
        self.ctx = self.lib.bert_load_from_file(fname.encode())
        if not self.ctx:
            raise RuntimeError(f"bert_load_from_file({fname}) failed (see stderr); no CPU fallback exists")
        self.n_embd = self.lib.bert_n_embd(self.ctx)
        self.n_max_tokens = self.lib.bert_n_max_tokens(self.ctx)

    def close(self):
        if self.ctx:
            self.lib.bert_free(self.ctx)
            self.ctx = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def eval(self, tokens) -> np.ndarray:
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.zeros(self.n_embd, np.float32)
        self.lib.bert_eval(self.ctx, 1, _ptr(tok, _i32p), len(tok), _ptr(out, _f32p))
        return out

    def make_batch_args(self, batch):
        """ragged list of id arrays -> the pointer arrays bert_eval_batch takes (kept alive by the caller)."""
        toks = [np.ascontiguousarray(t, dtype=np.int32) for t in batch]
        n = len(toks)
        out = np.zeros((n, self.n_embd), np.float32)
        tp = (_i32p * n)(*[_ptr(t, _i32p) for t in toks])
        lens = np.array([len(t) for t in toks], np.int32)
        op = (_f32p * n)(*[ctypes.cast(out.ctypes.data + i * out.strides[0], _f32p) for i in range(n)])
        return toks, tp, lens, op, out

    def eval_batch(self, batch) -> np.ndarray:
        toks, tp, lens, op, out = self.make_batch_args(batch)
        self.lib.bert_eval_batch(self.ctx, 1, len(toks), tp, _ptr(lens, _i32p), op)
        return out

    def graph_stats(self) -> dict:
        """latency path bookkeeping summed over the context's engines: launches replayed from a captured CUDA graph, shapes
        captured, launches done directly (first sighting of a shape)"""
        a, b, c = ctypes.c_int64(0), ctypes.c_int64(0), ctypes.c_int64(0)
        self.lib.bk_bert_graph_stats(self.ctx, ctypes.byref(a), ctypes.byref(b), ctypes.byref(c))
        return {"replayed": a.value, "captured": b.value, "direct": c.value}

    def tokenize(self, text: str) -> np.ndarray:
        buf = np.zeros(self.n_max_tokens, np.int32)
        n = ctypes.c_int32(0)
        self.lib.bert_tokenize(self.ctx, text.encode("utf-8"), _ptr(buf, _i32p), ctypes.byref(n), self.n_max_tokens)
        return buf[: n.value].copy()

    def encode_batch(self, texts) -> np.ndarray:
        n = len(texts)
        out = np.zeros((n, self.n_embd), np.float32)
        tp = (ctypes.c_char_p * n)(*[t.encode("utf-8") for t in texts])
        op = (_f32p * n)(*[ctypes.cast(out.ctypes.data + i * out.strides[0], _f32p) for i in range(n)])
        self.lib.bert_encode_batch(self.ctx, 1, 16, n, tp, op)
        return out


def _sibling(name: str):
    """import a sibling file of this directory (the directory name has a dot, so it is not a package)."""
    import importlib.util
    import sys
    mod_name = "embedding_cpp_b200_" + name
    if mod_name in sys.modules:
        return sys.modules[mod_name]
    spec = importlib.util.spec_from_file_location(mod_name, os.path.join(os.path.dirname(os.path.abspath(__file__)), name + ".py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[mod_name] = mod
    spec.loader.exec_module(mod)
    return mod


class Engine:
    """One-GPU engine through bk_api.h (bench.py's device-resident leg, hidden-state tests).  The GGUF
    file is parsed on the Python side (gguf_io.read_gguf) and handed over tensor by tensor."""

    def __init__(self, gguf_path: str, device: int = 0, cap_tokens: int = 65536, cap_seqs: int = 8192, lib=None):
        self.lib = lib or load_library()
        g = _sibling("gguf_io")
        f = g.read_gguf(gguf_path)
        hp = f.hparams()
        names = g.tensor_names(hp)
        types = [f.tensors[n][0] for n, s in names if len(s) == 2]
        wtype = max(set(types), key=types.count)
        self.hp = hp
        self.chp = bk_hparams(hp.n_vocab, hp.n_max_tokens, hp.n_embd, hp.n_intermediate, hp.n_head, hp.n_layer, hp.eps, wtype)
        self.e = self.lib.bk_engine_create(device, ctypes.byref(self.chp), cap_tokens, cap_seqs)
        if not self.e:
            raise RuntimeError("bk_engine_create: " + self.lib.bk_last_error().decode())
        for i, (name, shape) in enumerate(names):
            t, fshape, raw = f.tensors[name]
            raw = np.ascontiguousarray(raw)
            ne0, ne1 = shape[-1], (shape[0] if len(shape) == 2 else 1)
            self._ck(self.lib.bk_engine_set_tensor(self.e, i, t, raw.ctypes.data, ne0, ne1))
        self._ck(self.lib.bk_engine_finalize(self.e))

    def _ck(self, rc):
        if rc != 0:
            raise RuntimeError(self.lib.bk_last_error().decode())

    def close(self):
        if self.e:
            self.lib.bk_engine_destroy(self.e)
            self.e = None

    @staticmethod
    def _pack(batch):
        lens = np.array([len(t) for t in batch], np.int32)
        toks = np.ascontiguousarray(np.concatenate([np.asarray(t, np.int32) for t in batch]), dtype=np.int32)
        return toks, lens

    def forward(self, batch) -> np.ndarray:
        toks, lens = self._pack(batch)
        out = np.zeros((len(lens), self.hp.n_embd), np.float32)
        self._ck(self.lib.bk_engine_forward(self.e, _ptr(toks, _i32p), _ptr(lens, _i32p), len(lens), _ptr(out, _f32p)))
        return out

    def stage(self, batch):
        toks, lens = self._pack(batch)
        self._n_staged = len(lens)
        self._ck(self.lib.bk_engine_stage(self.e, _ptr(toks, _i32p), _ptr(lens, _i32p), len(lens)))

    def run_staged(self, iters: int = 1, time_kernels: bool = False) -> float:
        ms = ctypes.c_float(0)
        self._ck(self.lib.bk_engine_run_staged(self.e, iters, 1 if time_kernels else 0, ctypes.byref(ms)))
        return ms.value

    def read_output(self) -> np.ndarray:
        out = np.zeros((self._n_staged, self.hp.n_embd), np.float32)
        self._ck(self.lib.bk_engine_read_output(self.e, _ptr(out, _f32p), self._n_staged))
        return out

    def kernel_ms(self):
        ms = np.zeros(len(KERNEL_NAMES), np.float32)
        n = np.zeros(len(KERNEL_NAMES), np.int32)
        self._ck(self.lib.bk_engine_kernel_ms(self.e, _ptr(ms, _f32p), _ptr(n, _i32p)))
        return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(KERNEL_NAMES)}

    def keep_hidden(self, on: bool = True):
        self._ck(self.lib.bk_engine_set_keep_hidden(self.e, 1 if on else 0))

    def hidden(self, which: int, n_tokens: int) -> np.ndarray:
        out = np.zeros((n_tokens, self.hp.n_embd), np.float32)
        self._ck(self.lib.bk_engine_read_hidden(self.e, which, _ptr(out, _f32p)))
        return out

    def launch_count(self) -> int:
        return int(self.lib.bk_engine_launch_count(self.e))

    def check_guards(self) -> int:
        n = int(self.lib.bk_engine_check_guards(self.e))
        if n < 0:
            raise RuntimeError(self.lib.bk_last_error().decode())
        return n
