"""ctypes drivers for the two CPU checkers (TEST INFRASTRUCTURE -- see oracle/bert_oracle.c header).

* `Oracle`     : oracle/liboracle.so, the C restatement of the reference algorithm ("port").
* `Reference`  : oracle/_ref/libbert_ref.so, the reference's own bert.cpp + ggml compiled by
                 oracle/build_ref.sh ("reference"); present wherever it was prebuilt.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
from __future__ import annotations

import ctypes
import importlib.util
import os
import subprocess
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
LIB_ORACLE = os.path.join(HERE, "liboracle.so")
LIB_REF = os.path.join(HERE, "_ref", "libbert_ref.so")

_i32p = ctypes.POINTER(ctypes.c_int32)
_f32p = ctypes.POINTER(ctypes.c_float)


def load_gguf_io():
    name = "embedding_cpp_b200_gguf_io"
    if name in sys.modules:
        return sys.modules[name]
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, "embedding.cpp_b200", "gguf_io.py"))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def build(ref: bool = True) -> None:
    subprocess.check_call(["make", "-s", "-C", HERE, "liboracle.so"])
    if ref and os.path.exists("/root/reference/bert.cpp") and not os.path.exists(LIB_REF):
        subprocess.check_call(["make", "-s", "-C", HERE, "ref"])


class _Tensor(ctypes.Structure):
    _fields_ = [("data", ctypes.c_void_p), ("type", ctypes.c_int32), ("_pad", ctypes.c_int32),
                ("ne0", ctypes.c_int64), ("ne1", ctypes.c_int64)]


class _Model(ctypes.Structure):
    _fields_ = [("n_vocab", ctypes.c_int32), ("n_max_tokens", ctypes.c_int32), ("n_embd", ctypes.c_int32),
                ("n_intermediate", ctypes.c_int32), ("n_head", ctypes.c_int32), ("n_layer", ctypes.c_int32),
                ("eps", ctypes.c_float), ("_pad", ctypes.c_int32), ("t", ctypes.POINTER(_Tensor))]


def _fp(a: np.ndarray):
    return a.ctypes.data_as(_f32p)


class Oracle:
    """C restatement, one sentence at a time (like the reference)."""

    def __init__(self, gguf_path: str, threads: int | None = None):
        if not os.path.exists(LIB_ORACLE):
            build(ref=False)
        if threads is not None:
            os.environ["OMP_NUM_THREADS"] = str(threads)
        self.lib = ctypes.CDLL(LIB_ORACLE)
        g = load_gguf_io()
        self.file = g.read_gguf(gguf_path)
        self.hp = self.file.hparams()
        names = g.tensor_names(self.hp)
        self._keep = []
        arr = (_Tensor * len(names))()
        for i, (name, shape) in enumerate(names):
            t, fshape, raw = self.file.tensors[name]
            assert tuple(fshape) == tuple(shape), (name, fshape, shape)
            raw = np.ascontiguousarray(raw)
            self._keep.append(raw)
            arr[i] = _Tensor(raw.ctypes.data, t, 0, shape[-1], shape[0] if len(shape) == 2 else 1)
        self._arr = arr
        hp = self.hp
        self.model = _Model(hp.n_vocab, hp.n_max_tokens, hp.n_embd, hp.n_intermediate, hp.n_head, hp.n_layer,
                            hp.eps, 0, ctypes.cast(arr, ctypes.POINTER(_Tensor)))
        L = self.lib
        L.oracle_forward.restype = ctypes.c_int
        L.oracle_forward.argtypes = [ctypes.POINTER(_Model), _i32p, ctypes.c_int, _f32p, _f32p]
        L.oracle_linear.argtypes = [ctypes.POINTER(_Tensor), ctypes.POINTER(_Tensor), ctypes.c_int, _f32p, _f32p]
        L.oracle_embed.argtypes = [ctypes.POINTER(_Model), _i32p, ctypes.c_int, _f32p]
        L.oracle_layernorm.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_float, _f32p, _f32p, _f32p, _f32p]
        L.oracle_attention.argtypes = [ctypes.c_int, ctypes.c_int, ctypes.c_int, _f32p, _f32p, _f32p, _f32p]
        L.oracle_gelu.argtypes = [ctypes.c_int, _f32p, _f32p]
        L.oracle_pool.argtypes = [ctypes.c_int, ctypes.c_int, _f32p, _f32p]
        L.oracle_f2h.restype = ctypes.c_uint16
        L.oracle_f2h.argtypes = [ctypes.c_float]
        L.oracle_h2f.restype = ctypes.c_float
        L.oracle_h2f.argtypes = [ctypes.c_uint16]
        L.oracle_quant_q8_0_roundtrip.argtypes = [_f32p, _f32p, ctypes.c_int]
        L.oracle_quant_q8_1_roundtrip.argtypes = [_f32p, _f32p, ctypes.c_int]
        L.oracle_init()

    # whole path ----------------------------------------------------------------
    def eval(self, tokens, layers: bool = False):
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.zeros(self.hp.n_embd, np.float32)
        lo = np.zeros((self.hp.n_layer + 1, len(tok), self.hp.n_embd), np.float32) if layers else None
        rc = self.lib.oracle_forward(ctypes.byref(self.model), tok.ctypes.data_as(_i32p), len(tok), _fp(out),
                                     _fp(lo) if layers else None)
        if rc != 0:
            raise ValueError(f"oracle_forward rc={rc}")
        return (out, lo) if layers else out

    def eval_batch(self, batch):
        return np.stack([self.eval(t) for t in batch])

    # single ops (kernel unit tests) -----------------------------------------------
    def tensor(self, idx: int):
        return ctypes.byref(self._arr[idx])

    def linear(self, w_idx: int, b_idx: int | None, x: np.ndarray) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float32)
        n = x.shape[0]
        y = np.zeros((n, self._arr[w_idx].ne1), np.float32)
        self.lib.oracle_linear(self.tensor(w_idx), self.tensor(b_idx) if b_idx is not None else None, n, _fp(x), _fp(y))
        return y

    def embed(self, tokens) -> np.ndarray:
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        x = np.zeros((len(tok), self.hp.n_embd), np.float32)
        self.lib.oracle_embed(ctypes.byref(self.model), tok.ctypes.data_as(_i32p), len(tok), _fp(x))
        return x

    def layernorm(self, x, w, b) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float32)
        w = np.ascontiguousarray(w, np.float32)
        b = np.ascontiguousarray(b, np.float32)
        y = np.zeros_like(x)
        self.lib.oracle_layernorm(x.shape[0], x.shape[1], self.hp.eps, _fp(x), _fp(w), _fp(b), _fp(y))
        return y

    def attention(self, q, k, v, n_head: int) -> np.ndarray:
        q, k, v = (np.ascontiguousarray(a, np.float32) for a in (q, k, v))
        o = np.zeros_like(q)
        self.lib.oracle_attention(q.shape[0], q.shape[1], n_head, _fp(q), _fp(k), _fp(v), _fp(o))
        return o

    def gelu(self, x) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float32)
        y = np.zeros_like(x)
        self.lib.oracle_gelu(x.size, _fp(x), _fp(y))
        return y

    def pool(self, x) -> np.ndarray:
        x = np.ascontiguousarray(x, np.float32)
        out = np.zeros(x.shape[1], np.float32)
        self.lib.oracle_pool(x.shape[0], x.shape[1], _fp(x), _fp(out))
        return out


def bind_bert_api(lib) -> None:
    """argtypes for the bert.h symbols (same on the reference build and on build/libbert.so)."""
    lib.bert_load_from_file.restype = ctypes.c_void_p
    lib.bert_load_from_file.argtypes = [ctypes.c_char_p]
    lib.bert_free.argtypes = [ctypes.c_void_p]
    lib.bert_n_embd.restype = ctypes.c_int32
    lib.bert_n_embd.argtypes = [ctypes.c_void_p]
    lib.bert_n_max_tokens.restype = ctypes.c_int32
    lib.bert_n_max_tokens.argtypes = [ctypes.c_void_p]
    lib.bert_eval.argtypes = [ctypes.c_void_p, ctypes.c_int32, _i32p, ctypes.c_int32, _f32p]
    lib.bert_eval_batch.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.POINTER(_i32p),
                                    _i32p, ctypes.POINTER(_f32p)]
    lib.bert_tokenize.argtypes = [ctypes.c_void_p, ctypes.c_char_p, _i32p, _i32p, ctypes.c_int32]
    lib.bert_encode_batch.argtypes = [ctypes.c_void_p, ctypes.c_int32, ctypes.c_int32, ctypes.c_int32,
                                      ctypes.POINTER(ctypes.c_char_p), ctypes.POINTER(_f32p)]
    lib.bert_model_quantize.restype = ctypes.c_bool
    lib.bert_model_quantize.argtypes = [ctypes.c_char_p, ctypes.c_char_p, ctypes.c_int]


class Reference:
    """The reference's own CPU path (oracle/_ref/libbert_ref.so), one sentence per bert_eval call
    (the reference has no batching, bert.cpp:1064-1065)."""

    def __init__(self, gguf_path: str, n_threads: int = 1):
        if not os.path.exists(LIB_REF):
            raise FileNotFoundError(LIB_REF)
        self.lib = ctypes.CDLL(LIB_REF)
        bind_bert_api(self.lib)
        self.n_threads = n_threads
        # the reference printf()s progress to stdout; silence fd 1 while loading
        sys.stdout.flush()
        saved = os.dup(1)
        devnull = os.open(os.devnull, os.O_WRONLY)
        os.dup2(devnull, 1)
        try:
            self.ctx = self.lib.bert_load_from_file(gguf_path.encode())
        finally:
            os.dup2(saved, 1)
            os.close(saved)
            os.close(devnull)
        self.n_embd = self.lib.bert_n_embd(self.ctx)

    def eval(self, tokens) -> np.ndarray:
        tok = np.ascontiguousarray(tokens, dtype=np.int32)
        out = np.zeros(self.n_embd, np.float32)
        self.lib.bert_eval(self.ctx, self.n_threads, tok.ctypes.data_as(_i32p), len(tok), _fp(out))
        return out

    def eval_batch(self, batch):
        return np.stack([self.eval(t) for t in batch])

    def close(self):
        if self.ctx:
            self.lib.bert_free(self.ctx)
            self.ctx = None


def available_reference() -> bool:
    return os.path.exists(LIB_REF)
