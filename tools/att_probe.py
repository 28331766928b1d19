#!/usr/bin/env python
"""Run the attention kernel alone (bk_test_attention): usage att_probe.py n_seq L d n_head [qmode]"""
import ctypes, importlib.util, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
spec = importlib.util.spec_from_file_location("binding", os.path.join(ROOT, "embedding.cpp_b200", "binding.py"))
binding = importlib.util.module_from_spec(spec); sys.modules["binding"] = binding; spec.loader.exec_module(binding)
lib = binding.load_library()
n_seq, L, d, H = (int(x) for x in sys.argv[1:5])
qmode = int(sys.argv[5]) if len(sys.argv) > 5 else 0
u16p, i32p = ctypes.POINTER(ctypes.c_uint16), ctypes.POINTER(ctypes.c_int32)
rng = np.random.default_rng(1)
T = n_seq * L
qkv = (rng.standard_normal((T, 3 * d)) * 1.2).astype(np.float16)
out = np.zeros((T, d), np.float16)
lens = np.full(n_seq, L, np.int32)
t0 = time.perf_counter()
rc = lib.bk_test_attention(0, n_seq, lens.ctypes.data_as(i32p), d, H, qkv.view(np.uint16).ctypes.data_as(u16p), qmode, out.view(np.uint16).ctypes.data_as(u16p))
print("rc", rc, "wall ms", (time.perf_counter() - t0) * 1e3, "finite", bool(np.isfinite(out.astype(np.float32)).all()))
