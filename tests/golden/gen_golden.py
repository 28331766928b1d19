"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/libbert_ref.so = /root/reference's
bert.cpp + ggml compiled by oracle/build_ref.sh), in the authoring container.  The GPU box has no
/root/reference; it regenerates the same model files from the same seeds (gguf_io.make_model is
deterministic) and compares against the committed outputs.

    python tests/golden/gen_golden.py

Each case: model (shape, ftype, seed, n_vocab, dist), token lists, reference embeddings (f32).
The reference ships no golden embeddings of its own (examples/test_embedding.cpp only prints), so
these are the pin for both oracle/bert_oracle.c (bit-exact) and the CUDA path (tolerance)."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle as O  # noqa: E402

LENS = [2, 3, 16, 31, 32, 33, 128, 384, 512]
CASES = [  # name, shape, ftype, seed, n_vocab, dist, lens
    ("mini_f32", "mini", "f32", 11, 500, "bert-init", LENS),
    ("mini_f16", "mini", "f16", 11, 500, "bert-init", LENS),
    ("mini_q4_0", "mini", "q4_0", 11, 500, "bert-init", LENS),
    ("mini_q4_1", "mini", "q4_1", 11, 500, "bert-init", LENS),
    ("mini64_f16", "mini64", "f16", 12, 500, "bert-init", LENS),
    ("mini64_q4_0", "mini64", "q4_0", 12, 500, "bert-init", LENS),
    ("mini64_q4_1", "mini64", "q4_1", 12, 500, "bert-init", LENS),
    ("tiny_f32", "tiny", "f32", 13, 300, "bert-init", [1, 2, 5, 17, 64]),
    ("minilm_f16", "minilm-l6", "f16", 14, 1000, "bert-init", [7, 128, 200]),
    ("minilm_q4_0", "minilm-l6", "q4_0", 14, 1000, "bert-init", [7, 128, 200]),
    ("mini_q4_0_stress", "mini", "q4_0", 15, 500, "stress", [16, 128]),
    ("mini_f16_stress", "mini", "f16", 15, 500, "stress", [16, 128]),
]


def tokens_for(name, n_vocab, lens):
    rng = np.random.default_rng(sum(map(ord, name)))
    out = []
    for L in lens:
        t = rng.integers(0, n_vocab, L).astype(np.int32)
        if n_vocab > 102:
            t[0] = 101
            if L >= 2:
                t[-1] = 102
        out.append(t)
    return out


def main():
    O.build(ref=True)
    g = O.load_gguf_io()
    with tempfile.TemporaryDirectory() as td:
        for name, shape, ftype, seed, n_vocab, dist, lens in CASES:
            path = os.path.join(td, name + ".gguf")
            g.make_model(path, shape, ftype, seed=seed, n_vocab=n_vocab, dist=dist)
            ref = O.Reference(path, n_threads=4)
            toks = tokens_for(name, n_vocab, lens)
            emb = np.stack([ref.eval(t) for t in toks])
            ref.close()
            np.savez_compressed(os.path.join(HERE, name + ".npz"), shape=shape, ftype=ftype, seed=seed, n_vocab=n_vocab,
                                dist=dist, lens=np.array(lens, np.int32), tokens=np.concatenate(toks), emb=emb)
            print(name, emb.shape, float(np.linalg.norm(emb, axis=1).min()))


if __name__ == "__main__":
    main()
