#!/usr/bin/env python
"""HF BertModel directory -> GGUF (F32 or F16), the producer side of the path.

Counterpart of the reference's models/convert-to-gguf.py:145-346, written against the on-disk contract rather
than against gguf-py / transformers (NumPy + this package's own GGUF writer; no torch needed for .safetensors):

    python embedding.cpp_b200/convert_to_gguf.py MODEL_DIR [0|1] [--outfile out.gguf]       (0 = f32, 1 = f16, default 1)

What the reference does and this mirrors (file:line under /root/reference/models/convert-to-gguf.py):
  * :145-151   only `architectures[0] == "BertModel"` is accepted
  * :160-198   KVs: general.name = directory name, general.source_hf_repo = config `_name_or_path` (if any),
               bert.context_length (max_sequence_length | max_position_embeddings | model_max_length),
               embedding_length, block_count, feed_forward_length, rope.dimension_count (= head size, unread),
               attention.head_count, head_count_kv, layer_norm_epsilon
  * :218-220   blob.tokenizer.json = the tokenizer.json bytes, verbatim (what bert_load_from_file builds its tokenizer from)
  * :226-250   tokenizer.ggml.model "bert"; tokens = vocabulary by id for ids 0..vocab_size-1; dummy scores 0.0 and token types 1
  * :253-257   special token ids (gguf.SpecialVocab: tokenizer_config.json names looked up in tokenizer.json's added_tokens /
               vocabulary; bos <- cls and eos <- sep when undefined, as gguf-py 0.19 does) + tokenizer.ggml.cls_token_id
  * :294-300   embeddings.position_ids and the pooler are dropped
  * :314-321   ftype 1: float32 2-D tensors whose name ends in ".weight" become float16; everything else stays float32
  * tensor names are the AutoModel state_dict names (a leading "bert." of a BertFor* checkpoint is stripped), in file order

tests/test_convert.py runs the reference's script itself (where /root/reference exists) on a random-init BertModel and
compares KVs and tensor bytes with this converter's output."""
from __future__ import annotations

import argparse
import json
import os
import struct
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))
import gguf_io as g  # noqa: E402

_ST_DTYPES = {"F32": np.float32, "F16": np.float16, "F64": np.float64, "I64": np.int64, "I32": np.int32, "BF16": None}


def read_safetensors(path: str) -> dict[str, np.ndarray]:
    """safetensors container: u64 header length, JSON header {name: {dtype, shape, data_offsets}}, raw little-endian data."""
    with open(path, "rb") as f:
        (n,) = struct.unpack("<Q", f.read(8))
        header = json.loads(f.read(n))
        data = f.read()
    out = {}
    for name, info in header.items():
        if name == "__metadata__":
            continue
        a, b = info["data_offsets"]
        if info["dtype"] == "BF16":   # widen bf16 -> f32 (exact)
            u = np.frombuffer(data[a:b], dtype=np.uint16).astype(np.uint32) << 16
            arr = u.view(np.float32)
        else:
            arr = np.frombuffer(data[a:b], dtype=_ST_DTYPES[info["dtype"]])
        out[name] = arr.reshape(info["shape"])
    return out


def read_state_dict(dir_model: str) -> dict[str, np.ndarray]:
    st = os.path.join(dir_model, "model.safetensors")
    if os.path.exists(st):
        return read_safetensors(st)
    pt = os.path.join(dir_model, "pytorch_model.bin")
    if os.path.exists(pt):
        import torch  # only for legacy pickled checkpoints
        return {k: v.float().numpy() if v.is_floating_point() else v.numpy() for k, v in torch.load(pt, map_location="cpu", weights_only=True).items()}
    raise FileNotFoundError(f"{dir_model}: neither model.safetensors nor pytorch_model.bin")


def special_token_ids(dir_model: str, tok: dict, vocab: dict[str, int]) -> dict[str, int]:
    """{bos,eos,unk,sep,pad,cls,mask} -> id, following gguf.SpecialVocab (tokenizer_config.json `<typ>_token` entries resolved
    through tokenizer.json's added_tokens; then config.json `<typ>_token_id`)."""
    cfg_path = os.path.join(dir_model, "tokenizer_config.json")
    tcfg = json.load(open(cfg_path, encoding="utf-8")) if os.path.exists(cfg_path) else {}
    if tok is not None and tcfg:
        if not tcfg.get("bos_token") and tcfg.get("cls_token"):
            tcfg["bos_token"] = tcfg["cls_token"]
        if not tcfg.get("eos_token") and tcfg.get("sep_token"):
            tcfg["eos_token"] = tcfg["sep_token"]
    added = {t["content"]: t["id"] for t in (tok or {}).get("added_tokens", [])}
    out = {}
    for typ in ("bos", "eos", "unk", "sep", "pad", "cls", "mask"):
        entry = tcfg.get(f"{typ}_token")
        content = entry.get("content") if isinstance(entry, dict) else entry
        if isinstance(content, str) and content in added:
            out[typ] = added[content]
    cfg = json.load(open(os.path.join(dir_model, "config.json"), encoding="utf-8"))
    for typ in ("bos", "eos", "unk", "sep", "pad", "cls", "mask"):
        tid = cfg.get(f"{typ}_token_id")
        if typ not in out and isinstance(tid, int) and tid >= 0:
            out[typ] = tid
    return out


def module_order(state: dict, n_layer: int) -> list[str]:
    """tensor names in BertModel's module order -- the order of the state_dict the reference iterates (:286-293); a
    safetensors header is sorted by name instead.  Names outside the encoder follow in file order."""
    names = ["embeddings.word_embeddings.weight", "embeddings.position_embeddings.weight", "embeddings.token_type_embeddings.weight",
             "embeddings.LayerNorm.weight", "embeddings.LayerNorm.bias"]
    per = ["attention.self.query", "attention.self.key", "attention.self.value", "attention.output.dense", "attention.output.LayerNorm",
           "intermediate.dense", "output.dense", "output.LayerNorm"]
    for i in range(n_layer):
        names += [f"encoder.layer.{i}.{m}.{wb}" for m in per for wb in ("weight", "bias")]
    known = [n for n in names if n in state]
    return known + [n for n in state if n not in set(known)]


_SPECIAL_KEY = {"bos": "bos_token_id", "eos": "eos_token_id", "unk": "unknown_token_id", "sep": "seperator_token_id",  # (sic) gguf.h:297
                "pad": "padding_token_id", "cls": "cls_token_id", "mask": "mask_token_id"}


def convert(dir_model: str, ftype: int, outfile: str | None = None) -> str:
    dir_model = os.path.normpath(dir_model)
    if not os.path.isdir(dir_model):
        raise SystemExit(f"Error: {dir_model} is not a directory")
    hp = json.load(open(os.path.join(dir_model, "config.json"), encoding="utf-8"))
    if (hp.get("architectures") or [""])[0] != "BertModel":
        raise SystemExit("Model architecture not supported: " + str((hp.get("architectures") or [""])[0]))
    outfile = outfile or os.path.join(dir_model, f"ggml-model-{['f32', 'f16'][ftype]}.gguf")
    for key in ("max_sequence_length", "max_position_embeddings", "model_max_length"):
        if key in hp:
            ctx_length = hp[key]
            break
    else:
        raise SystemExit("gguf: can not find ctx length parameter.")
    heads = hp["num_attention_heads"]
    kvs = [g._kv("general.architecture", g.T_STR, "bert"), g._kv("general.name", g.T_STR, os.path.basename(dir_model))]
    if hp.get("_name_or_path"):
        kvs.append(g._kv("general.source_hf_repo", g.T_STR, hp["_name_or_path"]))
    kvs += [g._kv("bert.context_length", g.T_U32, ctx_length),
            g._kv("bert.embedding_length", g.T_U32, hp["hidden_size"]),
            g._kv("bert.block_count", g.T_U32, hp["num_hidden_layers"]),
            g._kv("bert.feed_forward_length", g.T_U32, hp["intermediate_size"]),
            g._kv("bert.rope.dimension_count", g.T_U32, hp["hidden_size"] // heads),
            g._kv("bert.attention.head_count", g.T_U32, heads),
            g._kv("bert.attention.head_count_kv", g.T_U32, hp.get("num_key_value_heads", heads)),
            g._kv("bert.attention.layer_norm_epsilon", g.T_F32, hp["layer_norm_eps"])]

    tok_path = os.path.join(dir_model, "tokenizer.json")
    if not os.path.isfile(tok_path):
        raise SystemExit(f"Error: Missing {tok_path}")
    blob = open(tok_path, "rb").read()
    tok = json.loads(blob.decode("utf-8"))
    vocab = dict(tok["model"]["vocab"])
    for t in tok.get("added_tokens", []):
        vocab.setdefault(t["content"], t["id"])
    by_id = {i: t for t, i in vocab.items()}
    n_vocab = hp["vocab_size"]
    missing = [i for i in range(n_vocab) if i not in by_id]
    if missing:
        raise SystemExit(f"tokenizer.json has no token for ids {missing[:8]} (vocab_size {n_vocab})")
    kvs += [g._kv("blob.tokenizer.json", g.T_STR, blob),
            g._kv("tokenizer.ggml.model", g.T_STR, "bert"),
            g._kv("tokenizer.ggml.tokens", g.T_ARR, (g.T_STR, [by_id[i] for i in range(n_vocab)])),
            g._kv("tokenizer.ggml.scores", g.T_ARR, (g.T_F32, [0.0] * n_vocab)),
            g._kv("tokenizer.ggml.token_type", g.T_ARR, (g.T_I32, [1] * n_vocab))]
    special = special_token_ids(dir_model, tok, vocab)
    for typ, tid in special.items():
        if typ != "cls":
            kvs.append(g._kv("tokenizer.ggml." + _SPECIAL_KEY[typ], g.T_U32, tid))
    if "cls" not in special:
        raise SystemExit("tokenizer has no cls token (convert-to-gguf.py:257 requires tokenizer.cls_token_id)")
    kvs.append(g._kv("tokenizer.ggml.cls_token_id", g.T_U32, special["cls"]))

    tensors = []
    state = {(k[5:] if k.startswith("bert.") else k): v for k, v in read_state_dict(dir_model).items()}
    for name in module_order(state, hp["num_hidden_layers"]):
        data = state[name]
        if name in ("embeddings.position_ids", "pooler.dense.weight", "pooler.dense.bias") or name.startswith("cls."):
            continue
        data = np.squeeze(np.asarray(data))
        if data.dtype not in (np.float32, np.float16):
            data = data.astype(np.float32)
        if ftype == 1 and data.dtype == np.float32 and name.endswith(".weight") and data.ndim == 2:
            data = data.astype(np.float16)
        t = g.GGML_F16 if data.dtype == np.float16 else g.GGML_F32
        tensors.append((name, data.shape, t, np.ascontiguousarray(data).tobytes()))
    g.write_gguf(outfile, kvs, tensors)
    return outfile


def main():
    ap = argparse.ArgumentParser(description="Convert a HuggingFace BertModel directory to a GGUF file for libbert.so")
    ap.add_argument("--outfile", default=None, help="path to write to; default: MODEL_DIR/ggml-model-{f32,f16}.gguf")
    ap.add_argument("model", help="directory containing config.json, tokenizer.json and model.safetensors / pytorch_model.bin")
    ap.add_argument("ftype", type=int, choices=[0, 1], default=1, nargs="?", help="0 = float32, 1 = float16")
    a = ap.parse_args()
    print("gguf: model successfully exported to '%s'" % convert(a.model, a.ftype, a.outfile))


if __name__ == "__main__":
    main()
