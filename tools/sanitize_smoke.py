"""Small pass over EVERY kernel of the path for compute-sanitizer (one tool per gpurun call, B200_PROFILING.md):

    compute-sanitizer --tool memcheck|racecheck|synccheck python tools/sanitize_smoke.py        (`make sanitize TOOL=...`)

Engine forwards of mini384 models (F16, Q4_0, Q4_1) over a ragged batch that reaches the three attention kernels (persistent
<= 48 tokens, per item <= 128, streaming above), both LayerNorm paths (GEMM + add_ln, and the fused cluster kernel), the packed
Q4 GEMM (decode in shared memory), the on-device batcher, the CUDA-graph replay of the latency path; then the single-kernel
test entries.  Sizes are small: the tools slow a launch down 10-100x.  Exit code 0 = the program itself saw no error; the
sanitizer's own report is its summary line."""
import ctypes
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402  (for _load)

gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
lib = binding.load_library()
rng = np.random.default_rng(0)
os.environ.setdefault("BERT_B200_DEVICES", "0")
os.environ["BERT_B200_ENGINES_PER_GPU"] = "1"


def toks(L, n_vocab=500):
    return rng.integers(0, n_vocab, L).astype(np.int32)


with tempfile.TemporaryDirectory() as td:
    for ftype in ("f16", "q4_0", "q4_1"):
        path = os.path.join(td, f"m_{ftype}.gguf")
        gio.make_model(path, "mini384", ftype, seed=3, n_vocab=500)
        for fused in ("-1", "0"):                                  # GEMM + add_ln | fused projection + LayerNorm clusters
            os.environ["BERT_B200_FUSED_LN_TOKENS"] = fused
            m = binding.BertB200(path)
            short = [toks(L) for L in (1, 5, 17, 48)]
            mixed = [toks(L) for L in (1, 33, 100, 128, 129, 300, 512, 64)]
            for batch in (short, mixed, short):                    # the repeat replays the captured graph
                out = m.eval_batch(batch)
                assert np.allclose(np.linalg.norm(out, axis=1), 1.0, atol=1e-3), (ftype, fused)
            one = m.eval(toks(128))
            one = m.eval(toks(128))
            assert abs(float(np.linalg.norm(one)) - 1.0) < 1e-3
            m.close()
        print("engine ok:", ftype, flush=True)

u16 = lambda a: a.view(np.uint16).ctypes.data_as(ctypes.POINTER(ctypes.c_uint16))
f32 = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
i32 = lambda a: a.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
for M, N, K, epi in ((300, 384, 384, 0), (129, 256, 128, 1), (200, 768, 256, 2), (256, 192, 64, 3)):
    A = (rng.standard_normal((M, K)) * 0.5).astype(np.float16)
    W = (rng.standard_normal((N, K)) * 0.05).astype(np.float16)
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    resid = rng.standard_normal((M, N)).astype(np.float16)
    out = np.zeros((M, N), np.float16)
    assert lib.bk_test_gemm(0, M, N, K, u16(A), W.ctypes.data, binding.BK_F16, f32(bias), u16(resid) if epi == 1 else None, epi, u16(out), None) == 0
for qmode in (0, 1):
    M, N, K = 300, 768, 256
    A = (rng.standard_normal((M, K)) * 0.5).astype(np.float16)
    W = (rng.standard_normal((N, K)) * 0.05).astype(np.float16)
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    resid = rng.standard_normal((M, N)).astype(np.float16)
    g, b = np.ones(N, np.float32), np.zeros(N, np.float32)
    y, yq = np.zeros((M, N), np.float16), np.zeros((M, N), np.float16)
    assert lib.bk_test_gemm_ln(0, M, N, K, u16(A), u16(W), f32(bias), u16(resid), f32(g), f32(b), 1e-12, qmode, u16(y), u16(yq), None) == 0
lens = np.array([5, 128, 1, 33, 200, 512], np.int32)
d, H = 256, 4
qkv = (rng.standard_normal((int(lens.sum()), 3 * d))).astype(np.float16)
out = np.zeros((int(lens.sum()), d), np.float16)
assert lib.bk_test_attention(0, len(lens), i32(lens), d, H, u16(qkv), 0, u16(out)) == 0
n = 3000
lens = rng.integers(1, 513, n).astype(np.int32)
tk = rng.integers(0, 30000, int(lens.sum())).astype(np.int32)
order, cu = np.zeros(n, np.int32), np.zeros(n + 1, np.int32)
tok, pos = np.zeros(len(tk), np.int32), np.zeros(len(tk), np.int32)
assert lib.bk_test_batcher(0, n, i32(lens), i32(tk), i32(order), i32(cu), i32(tok), i32(pos)) == 0
assert np.array_equal(order, np.argsort(-lens.astype(np.int64), kind="stable"))
print("sanitize_smoke: all launches returned without error", flush=True)
