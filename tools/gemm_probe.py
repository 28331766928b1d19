"""Time one projection GEMM alone through the C-ABI test entry (bk_test_gemm):
    python tools/gemm_probe.py M N K [epi] [bn] [cg] [wtype] [in_smem]
epi: 0 bias, 1 bias+resid, 2 bias+GELU; wtype 1 f16 (default), 2 Q4_0, 3 Q4_1; in_smem 1: decode Q4 in shared memory.
Prints ms and TFLOP/s (5 back-to-back launches, CUDA events)."""
import ctypes
import importlib.util
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("binding", os.path.join(ROOT, "embedding.cpp_b200", "binding.py"))
binding = importlib.util.module_from_spec(spec)
sys.modules["binding"] = binding
spec.loader.exec_module(binding)


def main():
    a = [int(x) for x in sys.argv[1:]]
    M, N, K = a[0], a[1], a[2]
    epi = a[3] if len(a) > 3 else 0
    bn = a[4] if len(a) > 4 else 0
    cg = a[5] if len(a) > 5 else 0
    wtype = a[6] if len(a) > 6 else 1
    in_smem = a[7] if len(a) > 7 else 0
    spec2 = importlib.util.spec_from_file_location("gio", os.path.join(ROOT, "embedding.cpp_b200", "gguf_io.py"))
    gio = importlib.util.module_from_spec(spec2)
    sys.modules["gio"] = gio
    spec2.loader.exec_module(gio)
    lib = binding.load_library()
    rng = np.random.default_rng(0)
    A = (rng.standard_normal((M, K)) * 0.5).astype(np.float16)
    W = (rng.standard_normal((N, K)) * 0.05).astype(np.float16)
    if wtype in (2, 3):
        W = np.ascontiguousarray((gio.quantize_q4_0 if wtype == 2 else gio.quantize_q4_1)(W.astype(np.float32)))
    bias = np.zeros(N, np.float32)
    resid = (rng.standard_normal((M, N))).astype(np.float16) if epi == 1 else None
    out = np.zeros((M, N), np.float16)
    ms = ctypes.c_float(0)
    u16 = lambda x: x.view(np.uint16).ctypes.data_as(ctypes.POINTER(ctypes.c_uint16))
    rc = lib.bk_test_gemm(0, M, N, K, u16(A), W.ctypes.data, wtype, bias.ctypes.data_as(ctypes.POINTER(ctypes.c_float)),
                          u16(resid) if resid is not None else None, epi | (bn << 8) | (cg << 20) | (in_smem << 28), u16(out), ctypes.byref(ms))
    assert rc == 0
    print(f"M={M} N={N} K={K} epi={epi} bn={bn} cg={cg} wtype={wtype} in_smem={in_smem}: {ms.value * 1e3:.1f} us  {2.0 * M * N * K / ms.value / 1e9:.0f} TFLOP/s")


if __name__ == "__main__":
    main()
