"""Stand-alone timing of the weight-stationary Q4 GEMM (gemm_ws.cu) against the same GEMM from f16 operand copies and from the
per-stage decode kernel, at the shapes of the two projections it serves:  python tools/gemm_ws_probe.py [M]"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
lib = binding.load_library()
_u16p, _f32p = ctypes.POINTER(ctypes.c_uint16), ctypes.POINTER(ctypes.c_float)
M = int(sys.argv[1]) if len(sys.argv) > 1 else 65536
rng = np.random.default_rng(0)
for name, N, K, epi in (("qkv base", 2304, 768, 0), ("ff1 base", 3072, 768, 3), ("qkv large", 3072, 1024, 0), ("ff1 large", 4096, 1024, 3), ("qkv minilm", 1152, 384, 0)):
    A = (rng.standard_normal((M, K)) * 0.5).astype(np.float16)
    raw = np.ascontiguousarray(gio.quantize_q4_0((rng.standard_normal((N, K)) * 0.05).astype(np.float32)))
    bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
    out = np.zeros((M, N), np.float16)
    row = []
    cands = [("f16 copy", 0), ("per-stage decode", 1 << 28)]
    for bn in (128, 192, 256):
        if N % bn == 0 and (K // 64) * (bn // 2) * 128 <= 227 * 1024 - 34304 - 3 * 16384 and 74 // (N // bn) >= 1:
            cands.append((f"ws{bn}", (1 << 29) | (bn << 8)))
    for label, flag in cands:
        ms = ctypes.c_float(0)
        rc = lib.bk_test_gemm(0, M, N, K, A.view(np.uint16).ctypes.data_as(_u16p), raw.ctypes.data, 2, bias.ctypes.data_as(_f32p), None,
                              epi | flag, out.view(np.uint16).ctypes.data_as(_u16p), ctypes.byref(ms))
        assert rc == 0
        row.append(f"{label} {ms.value * 1e3:6.1f} us {2.0 * M * N * K / (ms.value * 1e-3) / 1e12:6.0f} TF")
    print(f"{name:12s} M={M} N={N} K={K} epi={epi}: " + " | ".join(row), flush=True)
