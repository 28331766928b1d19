"""GGUF producer/consumer for BERT encoder files (host-side tooling, NumPy only).

Mirrors the on-disk contract of the reference so that the same file loads in
the reference (`bert_load_from_file`, /root/reference/bert.cpp:783) and in this
engine's C++ loader (csrc/gguf.cpp):

* container layout ........ /root/reference/ggml/src/ggml.c:18017-18306 (reader),
                            ggml/docs/gguf.md (spec)
* key names ............... /root/reference/gguf.h:256-302, arch string "bert"
* tensor names / shapes ... /root/reference/bert.cpp:623-652
* which tensors are F16 ... /root/reference/models/convert-to-gguf.py:314-321
  (2-D tensors whose name ends in ".weight"; everything else stays F32)
* which tensors are Q4 .... /root/reference/bert.cpp:1431-1436 (same rule)
* Q4_0 / Q4_1 block codec . /root/reference/ggml/src/ggml-quants.c:428-504,
                            block structs ggml-quants.h:10-30

Nothing here runs on the timed path; it exists so tests, smoke() and bench.py
can make random-init model files offline (no network for real checkpoints).
"""
from __future__ import annotations

import json
import struct
from dataclasses import dataclass, field

import numpy as np

GGUF_MAGIC = b"GGUF"
GGUF_VERSION = 3
ALIGNMENT = 32  # GGUF_DEFAULT_ALIGNMENT, ggml.h:238

# ggml_type ids, ggml.h:310-313
GGML_F32, GGML_F16, GGML_Q4_0, GGML_Q4_1 = 0, 1, 2, 3
TYPE_NAME = {GGML_F32: "f32", GGML_F16: "f16", GGML_Q4_0: "q4_0", GGML_Q4_1: "q4_1"}
FTYPE_BY_NAME = {v: k for k, v in TYPE_NAME.items()}
BLOCK_BYTES = {GGML_Q4_0: 18, GGML_Q4_1: 20}
QK = 32

# gguf_type ids, ggml.c:17878-17909
T_U8, T_I8, T_U16, T_I16, T_U32, T_I32, T_F32, T_BOOL, T_STR, T_ARR, T_U64, T_I64, T_F64 = range(13)
_SCALAR_FMT = {T_U8: "<B", T_I8: "<b", T_U16: "<H", T_I16: "<h", T_U32: "<I", T_I32: "<i",
               T_F32: "<f", T_BOOL: "<?", T_U64: "<Q", T_I64: "<q", T_F64: "<d"}


# --------------------------------------------------------------------------- codecs
def quantize_q4_0(x: np.ndarray) -> np.ndarray:
    """f32 [..., k] -> uint8 [..., k/32*18].  Restates quantize_row_q4_0_reference
    (ggml-quants.c:428-463): d = max/-8 where max is the signed value of the first
    element with the largest magnitude; q = min(15, (int8)(x/d + 8.5)); element j in
    the low nibble of byte j, element j+16 in the high nibble."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    assert x.shape[-1] % QK == 0
    b = x.reshape(-1, QK)
    idx = np.argmax(np.abs(b), axis=1)  # first occurrence == strict '<' update rule
    mx = b[np.arange(b.shape[0]), idx]
    d = (mx / np.float32(-8.0)).astype(np.float32)
    with np.errstate(divide="ignore"):
        inv = np.where(d != 0, np.float32(1.0) / d, np.float32(0.0)).astype(np.float32)
    q = (b * inv[:, None]).astype(np.float32) + np.float32(8.5)
    q = np.minimum(15, q.astype(np.int8)).astype(np.uint8)  # (int8_t) truncates toward zero
    out = np.empty((b.shape[0], 18), dtype=np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:] = q[:, :16] | (q[:, 16:] << 4)
    return out.reshape(*x.shape[:-1], x.shape[-1] // QK * 18)


def quantize_q4_1(x: np.ndarray) -> np.ndarray:
    """f32 [..., k] -> uint8 [..., k/32*20].  Restates quantize_row_q4_1_reference
    (ggml-quants.c:469-504): d = (max-min)/15, m = min, q = min(15, (int8)((x-min)/d + 0.5))."""
    x = np.ascontiguousarray(x, dtype=np.float32)
    assert x.shape[-1] % QK == 0
    b = x.reshape(-1, QK)
    mn = b.min(axis=1)
    mxv = b.max(axis=1)
    d = ((mxv - mn) / np.float32(15.0)).astype(np.float32)
    with np.errstate(divide="ignore"):
        inv = np.where(d != 0, np.float32(1.0) / d, np.float32(0.0)).astype(np.float32)
    q = ((b - mn[:, None]).astype(np.float32) * inv[:, None]).astype(np.float32) + np.float32(0.5)
    q = np.minimum(15, q.astype(np.int8)).astype(np.uint8)
    out = np.empty((b.shape[0], 20), dtype=np.uint8)
    out[:, 0:2] = d.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 2:4] = mn.astype(np.float16).view(np.uint8).reshape(-1, 2)
    out[:, 4:] = q[:, :16] | (q[:, 16:] << 4)
    return out.reshape(*x.shape[:-1], x.shape[-1] // QK * 20)


def dequantize(raw: np.ndarray, ggml_type: int, shape: tuple) -> np.ndarray:
    """Decode a tensor's raw bytes to f32 (dequantize_row_q4_0/q4_1, ggml-quants.c:1074-1113)."""
    if ggml_type == GGML_F32:
        return raw.view(np.float32).reshape(shape).copy()
    if ggml_type == GGML_F16:
        return raw.view(np.float16).reshape(shape).astype(np.float32)
    bb = BLOCK_BYTES[ggml_type]
    blk = raw.reshape(-1, bb)
    d = blk[:, 0:2].copy().view(np.float16).astype(np.float32)  # [nb,1]
    if ggml_type == GGML_Q4_0:
        qs = blk[:, 2:]
        lo = (qs & 0x0F).astype(np.int32) - 8
        hi = (qs >> 4).astype(np.int32) - 8
        vals = np.concatenate([lo, hi], axis=1).astype(np.float32) * d
    else:
        m = blk[:, 2:4].copy().view(np.float16).astype(np.float32)
        qs = blk[:, 4:]
        lo = (qs & 0x0F).astype(np.float32)
        hi = (qs >> 4).astype(np.float32)
        vals = np.concatenate([lo, hi], axis=1) * d + m
    return vals.astype(np.float32).reshape(shape)


# --------------------------------------------------------------------------- writer
def _s(b: bytes | str) -> bytes:
    if isinstance(b, str):
        b = b.encode("utf-8")
    return struct.pack("<Q", len(b)) + b


def _kv(key: str, vtype: int, value) -> bytes:
    out = _s(key) + struct.pack("<I", vtype)
    if vtype == T_STR:
        return out + _s(value)
    if vtype == T_ARR:
        etype, items = value
        out += struct.pack("<IQ", etype, len(items))
        if etype == T_STR:
            return out + b"".join(_s(i) for i in items)
        return out + np.asarray(items, dtype=np.dtype(_SCALAR_FMT[etype][1]).newbyteorder("<")).tobytes()
    return out + struct.pack(_SCALAR_FMT[vtype], value)


@dataclass
class BertHParams:
    n_vocab: int = 30522
    n_max_tokens: int = 512
    n_embd: int = 384
    n_intermediate: int = 1536
    n_head: int = 12
    n_layer: int = 6
    eps: float = 1e-12
    cls_id: int = 101
    sep_id: int = 102
    pad_id: int = 0
    unk_id: int = 100

    @property
    def d_head(self) -> int:
        return self.n_embd // self.n_head


SHAPES = {  # name -> BertHParams overrides (BASELINE.json configs)
    "minilm-l6": dict(n_embd=384, n_intermediate=1536, n_head=12, n_layer=6),
    "bert-base": dict(n_embd=768, n_intermediate=3072, n_head=12, n_layer=12),
    "bert-large": dict(n_embd=1024, n_intermediate=4096, n_head=16, n_layer=24),
    # small shapes for CPU-side tests (same structure, seconds on the oracle)
    "tiny": dict(n_embd=64, n_intermediate=128, n_head=2, n_layer=2, n_max_tokens=64),
    "mini": dict(n_embd=128, n_intermediate=512, n_head=4, n_layer=2, n_max_tokens=512),
    "mini64": dict(n_embd=256, n_intermediate=1024, n_head=4, n_layer=2, n_max_tokens=512),  # head size 64
    "mini384": dict(n_embd=384, n_intermediate=768, n_head=6, n_layer=2, n_max_tokens=512),  # n_embd the fused GEMM+LN kernel takes
}


def tensor_names(hp: BertHParams) -> list[tuple[str, tuple]]:
    """(name, numpy shape) in the order of /root/reference/bert.cpp:623-652.  GGUF `ne` is the
    reverse of the NumPy shape (ne[0] = contiguous dim)."""
    d, f = hp.n_embd, hp.n_intermediate
    out = [
        ("embeddings.word_embeddings.weight", (hp.n_vocab, d)),
        ("embeddings.token_type_embeddings.weight", (2, d)),
        ("embeddings.position_embeddings.weight", (hp.n_max_tokens, d)),
        ("embeddings.LayerNorm.weight", (d,)),
        ("embeddings.LayerNorm.bias", (d,)),
    ]
    for i in range(hp.n_layer):
        p = f"encoder.layer.{i}."
        out += [
            (p + "attention.self.query.weight", (d, d)), (p + "attention.self.query.bias", (d,)),
            (p + "attention.self.key.weight", (d, d)), (p + "attention.self.key.bias", (d,)),
            (p + "attention.self.value.weight", (d, d)), (p + "attention.self.value.bias", (d,)),
            (p + "attention.output.dense.weight", (d, d)), (p + "attention.output.dense.bias", (d,)),
            (p + "attention.output.LayerNorm.weight", (d,)), (p + "attention.output.LayerNorm.bias", (d,)),
            (p + "intermediate.dense.weight", (f, d)), (p + "intermediate.dense.bias", (f,)),
            (p + "output.dense.weight", (d, f)), (p + "output.dense.bias", (d,)),
            (p + "output.LayerNorm.weight", (d,)), (p + "output.LayerNorm.bias", (d,)),
        ]
    return out


def random_weights(hp: BertHParams, seed: int, dist: str = "bert-init") -> dict[str, np.ndarray]:
    """Synthetic weights.  "bert-init": matrices/embedding tables N(0,0.02^2), biases and
    LN beta N(0,0.1^2), LN gamma 1+N(0,0.1^2).  "stress": sigma 0.1 matrices, sigma 1 embeddings."""
    sw, se = (0.02, 0.02) if dist == "bert-init" else (0.1, 1.0)
    out = {}
    for i, (name, shape) in enumerate(tensor_names(hp)):
        rng = np.random.default_rng(1000 * seed + i)
        if len(shape) == 2:
            sigma = se if name.startswith("embeddings.") else sw
            w = rng.standard_normal(shape, dtype=np.float32) * np.float32(sigma)
        elif name.endswith("LayerNorm.weight"):
            w = np.float32(1.0) + rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        else:
            w = rng.standard_normal(shape, dtype=np.float32) * np.float32(0.1)
        out[name] = w.astype(np.float32)
    return out


def stub_tokenizer_json(vocab: list[str]) -> str:
    """A minimal HF tokenizer.json (BertNormalizer + BertPreTokenizer + WordPiece) over `vocab`."""
    return json.dumps({
        "version": "1.0", "truncation": None, "padding": None,
        "added_tokens": [
            {"id": i, "content": t, "single_word": False, "lstrip": False, "rstrip": False,
             "normalized": False, "special": True}
            for i, t in enumerate(vocab) if t in ("[PAD]", "[UNK]", "[CLS]", "[SEP]", "[MASK]")],
        "normalizer": {"type": "BertNormalizer", "clean_text": True, "handle_chinese_chars": True,
                       "strip_accents": None, "lowercase": True},
        "pre_tokenizer": {"type": "BertPreTokenizer"},
        "post_processor": None,
        "decoder": {"type": "WordPiece", "prefix": "##", "cleanup": True},
        "model": {"type": "WordPiece", "unk_token": "[UNK]", "continuing_subword_prefix": "##",
                  "max_input_chars_per_word": 100, "vocab": {t: i for i, t in enumerate(vocab)}},
    }, ensure_ascii=False)


def synthetic_vocab(n: int) -> list[str]:
    """n distinct tokens with BERT's special ids in the usual places when n > 102."""
    vocab = [f"[unused{i}]" for i in range(n)]
    specials = {0: "[PAD]", 100: "[UNK]", 101: "[CLS]", 102: "[SEP]", 103: "[MASK]"}
    base = list("abcdefghijklmnopqrstuvwxyz0123456789.,!?'-")
    words = ["the", "of", "and", "to", "in", "is", "that", "for", "it", "as", "was", "with", "be", "by",
             "on", "not", "he", "this", "are", "or", "his", "from", "at", "which", "but", "have", "an",
             "had", "they", "you", "were", "their", "one", "all", "we", "can", "her", "has", "there",
             "been", "if", "more", "when", "will", "would", "who", "so", "no", "embed", "##ding",
             "##s", "##ed", "##ing", "##ly", "##er", "##est", "quebec", "paris", "model", "token"]
    words += ["##" + c for c in "abcdefghijklmnopqrstuvwxyz0123456789"]
    for i, t in specials.items():
        if i < n:
            vocab[i] = t
    slot = 104 if n > 104 else 0
    for t in base + words:
        while slot < n and slot in specials:
            slot += 1
        if slot >= n:
            break
        vocab[slot] = t
        slot += 1
    return vocab


def write_bert_gguf(path: str, hp: BertHParams, weights: dict[str, np.ndarray], ftype: str = "f32",
                    vocab: list[str] | None = None, tokenizer_json: str | None = None) -> None:
    """Write a BERT GGUF.  `weights` holds f32 arrays by reference tensor name; `ftype` in
    f32|f16|q4_0|q4_1.  Q4 files quantise the f16-rounded weights, i.e. what the reference's
    quantize tool sees when it is fed the F16 file (models/run_conversions.sh:7-8)."""
    wt = FTYPE_BY_NAME[ftype]
    vocab = vocab if vocab is not None else synthetic_vocab(hp.n_vocab)
    assert len(vocab) == hp.n_vocab
    tokenizer_json = tokenizer_json if tokenizer_json is not None else stub_tokenizer_json(vocab)

    kvs = [
        _kv("general.architecture", T_STR, "bert"),
        _kv("general.name", T_STR, "embedding.cpp_b200 synthetic"),
        _kv("bert.context_length", T_U32, hp.n_max_tokens),
        _kv("bert.embedding_length", T_U32, hp.n_embd),
        _kv("bert.feed_forward_length", T_U32, hp.n_intermediate),
        _kv("bert.attention.head_count", T_U32, hp.n_head),
        _kv("bert.block_count", T_U32, hp.n_layer),
        _kv("bert.attention.layer_norm_epsilon", T_F32, hp.eps),
        _kv("tokenizer.ggml.model", T_STR, "bert"),
        _kv("tokenizer.ggml.tokens", T_ARR, (T_STR, vocab)),
        _kv("tokenizer.ggml.scores", T_ARR, (T_F32, [0.0] * hp.n_vocab)),
        _kv("tokenizer.ggml.token_type", T_ARR, (T_I32, [1] * hp.n_vocab)),
        _kv("tokenizer.ggml.unknown_token_id", T_U32, hp.unk_id),
        _kv("tokenizer.ggml.seperator_token_id", T_U32, hp.sep_id),  # (sic) gguf.h:297
        _kv("tokenizer.ggml.padding_token_id", T_U32, hp.pad_id),
        _kv("tokenizer.ggml.cls_token_id", T_U32, hp.cls_id),
        _kv("blob.tokenizer.json", T_STR, tokenizer_json),
    ]
    if wt in (GGML_Q4_0, GGML_Q4_1):
        kvs += [_kv("general.quantization_version", T_U32, 2), _kv("general.file_type", T_U32, wt)]

    tensors = []
    for name, shape in tensor_names(hp):
        w = np.ascontiguousarray(weights[name], dtype=np.float32)
        assert w.shape == shape, (name, w.shape, shape)
        t = wt if (w.ndim == 2 and name.endswith("weight")) else GGML_F32
        tensors.append((name, shape, t, encode_tensor(w, t)))
    write_gguf(path, kvs, tensors)


def encode_tensor(w: np.ndarray, t: int) -> bytes:
    """f32 array -> raw bytes of ggml type `t`.  Q4 blocks are taken from the f16-rounded values, i.e. what the
    reference's quantize tool sees when it is fed the F16 file (models/run_conversions.sh:7-8)."""
    w = np.ascontiguousarray(w, dtype=np.float32)
    if t == GGML_F32:
        return w.tobytes()
    if t == GGML_F16:
        return w.astype(np.float16).tobytes()
    w16 = w.astype(np.float16).astype(np.float32)
    return (quantize_q4_0 if t == GGML_Q4_0 else quantize_q4_1)(w16).tobytes()


def write_gguf(path: str, kvs: list[bytes], tensors: list[tuple]) -> None:
    """Container writer (ggml.c:18017-18306 reads it back): header, KV section (`kvs` = records made by _kv), tensor
    infos, padding to ALIGNMENT, tensor data each padded to ALIGNMENT.  `tensors` = (name, numpy shape, ggml type, raw bytes)."""
    infos, blobs, off = [], [], 0
    for name, shape, t, raw in tensors:
        ne = tuple(reversed(shape))
        infos.append(_s(name) + struct.pack("<I", len(ne)) + struct.pack(f"<{len(ne)}Q", *ne)
                     + struct.pack("<IQ", t, off))
        pad = (-len(raw)) % ALIGNMENT
        blobs.append(raw + b"\0" * pad)
        off += len(raw) + pad

    head = GGUF_MAGIC + struct.pack("<IQQ", GGUF_VERSION, len(infos), len(kvs)) + b"".join(kvs) + b"".join(infos)
    head += b"\0" * ((-len(head)) % ALIGNMENT)
    with open(path, "wb") as f:
        f.write(head)
        for b in blobs:
            f.write(b)


def make_model(path: str, shape: str = "minilm-l6", ftype: str = "f32", seed: int = 1, n_vocab: int = 1000,
               dist: str = "bert-init", **overrides) -> BertHParams:
    """Convenience: random-init model of a named shape written to `path`."""
    kw = dict(SHAPES[shape])
    kw.update(overrides)
    hp = BertHParams(n_vocab=n_vocab, **kw)
    write_bert_gguf(path, hp, random_weights(hp, seed, dist), ftype)
    return hp


# --------------------------------------------------------------------------- reader
@dataclass
class GGUFFile:
    kv: dict = field(default_factory=dict)
    tensors: dict = field(default_factory=dict)  # name -> (ggml_type, numpy shape, raw uint8 array)

    def hparams(self) -> BertHParams:
        k = self.kv
        return BertHParams(
            n_vocab=len(k["tokenizer.ggml.tokens"]), n_max_tokens=k["bert.context_length"],
            n_embd=k["bert.embedding_length"], n_intermediate=k["bert.feed_forward_length"],
            n_head=k["bert.attention.head_count"], n_layer=k["bert.block_count"],
            eps=k["bert.attention.layer_norm_epsilon"],
            cls_id=k.get("tokenizer.ggml.cls_token_id", -1), sep_id=k.get("tokenizer.ggml.seperator_token_id", -1),
            pad_id=k.get("tokenizer.ggml.padding_token_id", -1), unk_id=k.get("tokenizer.ggml.unknown_token_id", 0))

    def f32(self, name: str) -> np.ndarray:
        t, shape, raw = self.tensors[name]
        return dequantize(raw, t, shape)


def read_gguf(path: str) -> GGUFFile:
    """Independent GGUF v2/v3 reader (used by tests and by the oracle driver)."""
    data = np.fromfile(path, dtype=np.uint8)
    buf = data.tobytes()
    pos = 0

    def rd(fmt):
        nonlocal pos
        v = struct.unpack_from(fmt, buf, pos)
        pos += struct.calcsize(fmt)
        return v[0] if len(v) == 1 else v

    def rstr():
        nonlocal pos
        n = rd("<Q")
        s = buf[pos:pos + n]
        pos += n
        return s.decode("utf-8", errors="replace")

    def rval(t):
        nonlocal pos
        if t == T_STR:
            return rstr()
        if t == T_ARR:
            et, n = rd("<I"), rd("<Q")
            if et == T_STR:
                return [rstr() for _ in range(n)]
            dt = np.dtype(_SCALAR_FMT[et][1]).newbyteorder("<")
            arr = np.frombuffer(buf, dtype=dt, count=n, offset=pos)
            pos += n * dt.itemsize
            return arr
        return rd(_SCALAR_FMT[t])

    assert buf[:4] == GGUF_MAGIC, "not a GGUF file"
    pos = 4
    version, n_tensors, n_kv = rd("<I"), rd("<Q"), rd("<Q")
    assert version in (2, 3), f"unsupported GGUF version {version}"
    out = GGUFFile()
    for _ in range(n_kv):
        key = rstr()
        out.kv[key] = rval(rd("<I"))
    infos = []
    for _ in range(n_tensors):
        name = rstr()
        nd = rd("<I")
        ne = [rd("<Q") for _ in range(nd)]
        t, off = rd("<I"), rd("<Q")
        infos.append((name, tuple(reversed(ne)), t, off))
    align = out.kv.get("general.alignment", ALIGNMENT)
    base = (pos + align - 1) // align * align
    for name, shape, t, off in infos:
        n = int(np.prod(shape))
        nbytes = {GGML_F32: 4 * n, GGML_F16: 2 * n}.get(t, n // QK * BLOCK_BYTES.get(t, 0))
        out.tensors[name] = (t, shape, data[base + off: base + off + nbytes])
    return out
