#!/usr/bin/env python
"""Run the fused GEMM + LayerNorm kernel alone (bk_test_gemm_ln) on one shape: correctness vs NumPy and time.
usage: ln_probe.py M N K [qmode]"""
import ctypes, importlib.util, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("binding", os.path.join(ROOT, "embedding.cpp_b200", "binding.py"))
binding = importlib.util.module_from_spec(spec); sys.modules["binding"] = binding; spec.loader.exec_module(binding)
lib = binding.load_library()
M, N, K = (int(x) for x in sys.argv[1:4])
qmode = int(sys.argv[4]) if len(sys.argv) > 4 else 0
u16p, f32p = ctypes.POINTER(ctypes.c_uint16), ctypes.POINTER(ctypes.c_float)
u16 = lambda a: a.view(np.uint16).ctypes.data_as(u16p)
rng = np.random.default_rng(1)
A = (rng.standard_normal((M, K)) * 0.5).astype(np.float16)
W = (rng.standard_normal((N, K)) * 0.05).astype(np.float16)
bias = (rng.standard_normal(N) * 0.1).astype(np.float32)
resid = (rng.standard_normal((M, N)) + 0.3).astype(np.float16)
gamma = (1 + 0.1 * rng.standard_normal(N)).astype(np.float32)
beta = (0.1 * rng.standard_normal(N)).astype(np.float32)
y = np.zeros((M, N), np.float16); yq = np.zeros((M, N), np.float16)
ms = ctypes.c_float(0)
rc = lib.bk_test_gemm_ln(0, M, N, K, u16(A), u16(W), bias.ctypes.data_as(f32p), u16(resid), gamma.ctypes.data_as(f32p),
                         beta.ctypes.data_as(f32p), 1e-12, qmode, u16(y), u16(yq), ctypes.byref(ms))
print("rc", rc, "ms", ms.value, "TFLOP/s", 2.0 * M * N * K / (ms.value * 1e-3) / 1e12 if ms.value else None)
if rc == 0 and M <= 40000:
    pre = A.astype(np.float32) @ W.astype(np.float32).T + bias + resid.astype(np.float32)
    mu = pre.mean(-1, keepdims=True); var = ((pre - mu) ** 2).mean(-1, keepdims=True)
    ref = (pre - mu) / np.sqrt(var + 1e-12) * gamma + beta
    err = np.abs(y.astype(np.float32) - ref)
    print("max err", float(err.max()), "bad rows", np.unique(np.argwhere(err > 2e-3 * np.abs(ref) + 2e-3)[:, 0])[:20])
