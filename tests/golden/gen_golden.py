"""Generates tests/golden/*.npz from the REFERENCE ITSELF (oracle/_ref/libbert_ref.so = /root/reference's
bert.cpp + ggml compiled by oracle/build_ref.sh), in the authoring container.  The GPU box has no
/root/reference; it regenerates the same model files from the same seeds (gguf_io.make_model is
deterministic) and compares against the committed outputs.

    python tests/golden/gen_golden.py            # everything (minutes: BERT-large L=512 is ~4 s per sentence on 8 threads)
    python tests/golden/gen_golden.py base_q4_0  # only the named cases

Each case: model (shape, ftype, seed, n_vocab, dist), token lists, reference embeddings (f32).
The reference ships no golden embeddings of its own (examples/test_embedding.cpp only prints), so
these are the pin for both oracle/bert_oracle.c (bit-exact) and the CUDA path (tolerance).

Q4_0 / Q4_1 model files are made the way SURVEY.md section 8d specifies: the REFERENCE'S OWN bert_model_quantize
(bert.cpp:1313-1599) run on the F16 file.  The generator then checks that gguf_io.make_model(ftype=q4_*) -- what the GPU
box uses to regenerate the file from the seed -- holds byte-identical tensors, so the committed outputs are the
reference's outputs on the reference's own quantised weights."""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, os.path.join(ROOT, "oracle"))
import oracle as O  # noqa: E402

LENS = [2, 3, 16, 31, 32, 33, 128, 384, 512]
# BASELINE.json configs[2..4] shapes: every length class of the attention kernels (<= 48 persistent, <= 128 per item,
# streaming above), block boundaries of 32 (Q8 blocks), 128 (query blocks) and the context limit
LENS_BASE = [2, 3, 16, 31, 32, 33, 48, 49, 64, 100, 128, 128, 128, 129, 200, 255, 256, 257, 300, 384, 384, 448, 511, 512, 512, 512]
LENS_LARGE = [2, 16, 33, 128, 128, 128, 129, 256, 384, 512, 512, 512, 512]
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
    # round 2: the shapes the benchmark is quoted on (BASELINE.json configs[0..4])
    ("minilm_f32", "minilm-l6", "f32", 16, 1000, "bert-init", [2, 16, 33, 128, 128, 300, 512, 512]),                  # configs[0]
    ("minilm_q4_0_c1", "minilm-l6", "q4_0", 16, 1000, "bert-init", [128] * 12 + [2, 16, 33, 384, 512, 512]),          # configs[1]
    ("minilm_q4_1", "minilm-l6", "q4_1", 16, 1000, "bert-init", [2, 16, 33, 128, 128, 300, 512, 512]),
    ("base_f32", "bert-base", "f32", 21, 1000, "bert-init", [2, 16, 33, 128, 128, 384, 512, 512]),
    ("base_f16", "bert-base", "f16", 21, 1000, "bert-init", LENS_BASE),                                               # configs[2]
    ("base_q4_0", "bert-base", "q4_0", 21, 1000, "bert-init", LENS_BASE),                                             # the Target dtype
    ("base_q4_1", "bert-base", "q4_1", 21, 1000, "bert-init", LENS_BASE),                                             # configs[3]
    ("large_f16", "bert-large", "f16", 22, 1000, "bert-init", [16, 128, 128, 384, 512, 512]),
    ("large_q4_0", "bert-large", "q4_0", 22, 1000, "bert-init", LENS_LARGE),                                          # configs[4]
    ("large_q4_1", "bert-large", "q4_1", 22, 1000, "bert-init", [16, 128, 384, 512, 512]),
    ("base_q4_0_stress", "bert-base", "q4_0", 23, 1000, "stress", [16, 128, 512]),                                    # reported, not gating
    ("base_f16_stress", "bert-base", "f16", 23, 1000, "stress", [16, 128, 512]),
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


def reference_quantized(g, td, name, shape, ftype, seed, n_vocab, dist):
    """Q4 file written by the reference's own bert_model_quantize from the F16 file; asserts that gguf_io's writer
    (the one the GPU box regenerates the model with) produces the same tensor bytes."""
    import ctypes
    lib = ctypes.CDLL(O.LIB_REF)
    O.bind_bert_api(lib)
    f16 = os.path.join(td, name + ".f16.gguf")
    out = os.path.join(td, name + ".gguf")
    g.make_model(f16, shape, "f16", seed=seed, n_vocab=n_vocab, dist=dist)
    sys.stdout.flush()
    saved, devnull = os.dup(1), os.open(os.devnull, os.O_WRONLY)
    os.dup2(devnull, 1)
    try:
        ok = lib.bert_model_quantize(f16.encode(), out.encode(), {"q4_0": 2, "q4_1": 3}[ftype])
    finally:
        os.dup2(saved, 1)
        os.close(saved)
        os.close(devnull)
    assert ok, name
    os.remove(f16)
    mine = os.path.join(td, name + ".numpy.gguf")
    g.make_model(mine, shape, ftype, seed=seed, n_vocab=n_vocab, dist=dist)
    a, b = g.read_gguf(out), g.read_gguf(mine)
    assert set(a.tensors) == set(b.tensors)
    for k in a.tensors:
        assert a.tensors[k][0] == b.tensors[k][0] and np.array_equal(a.tensors[k][2], b.tensors[k][2]), (name, k)
    os.remove(mine)
    return out


def main():
    only = set(sys.argv[1:])
    O.build(ref=True)
    if not os.path.exists(os.path.join(os.path.dirname(O.LIB_REF), "libbert_ref_portable.so")):
        import subprocess
        subprocess.check_call([os.path.join(ROOT, "oracle", "build_ref.sh")], env=dict(os.environ, VARIANT="portable"))
    g = O.load_gguf_io()
    with tempfile.TemporaryDirectory() as td:
        for name, shape, ftype, seed, n_vocab, dist, lens in CASES:
            if only and name not in only:
                continue
            if ftype.startswith("q4"):
                path = reference_quantized(g, td, name, shape, ftype, seed, n_vocab, dist)
            else:
                path = os.path.join(td, name + ".gguf")
                g.make_model(path, shape, ftype, seed=seed, n_vocab=n_vocab, dist=dist)
            ref = O.Reference(path, n_threads=8)
            toks = tokens_for(name, n_vocab, lens)
            emb = np.stack([ref.eval(t) for t in toks])
            ref.close()
            # the same file and ids through the reference's PORTABLE build (oracle/build_ref.sh VARIANT=portable: its generic-C
            # code paths instead of the AVX2 ones).  Not a second gate: the distance between the two builds of the reference
            # itself is the reproducibility floor of the model (deep Q4 models are chaotic under the Q8 activation rounding,
            # SURVEY.md finding 4), and the parity tests report the CUDA path's distance next to it.
            old, O.LIB_REF = O.LIB_REF, os.path.join(os.path.dirname(O.LIB_REF), "libbert_ref_portable.so")
            try:
                ref = O.Reference(path, n_threads=8)
                emb_portable = np.stack([ref.eval(t) for t in toks])
                ref.close()
            finally:
                O.LIB_REF = old
            os.remove(path)
            np.savez_compressed(os.path.join(HERE, name + ".npz"), shape=shape, ftype=ftype, seed=seed, n_vocab=n_vocab,
                                dist=dist, lens=np.array(lens, np.int32), tokens=np.concatenate(toks), emb=emb, emb_portable=emb_portable)
            e64, p64 = emb.astype(np.float64), emb_portable.astype(np.float64)
            spread = 1.0 - (e64 * p64).sum(1) / (np.linalg.norm(e64, axis=1) * np.linalg.norm(p64, axis=1))
            print(name, emb.shape, float(np.linalg.norm(emb, axis=1).min()), "reference avx2-vs-portable max 1-cos %.2e" % spread.max(), flush=True)


if __name__ == "__main__":
    main()
