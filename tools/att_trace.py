"""Per-warp timeline of one CTA of the attention kernel for sentences > 128 tokens (attention_pt.cuh built with -DBK_EXP_TRACE:
`make BUILD=build_hintTRACE DEFS=-DBK_EXP_TRACE`), clock64 stamps at every barrier wait / hand-over of (sentence 64, head 0):
    BERT_B200_LIB=$PWD/build_hintTRACE/libbert.so python tools/att_trace.py [L] > gpurun_out/att_trace.txt"""
import ctypes
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
lib = binding.load_library()
L = int(sys.argv[1]) if len(sys.argv) > 1 else 512
n_seq, d, H = 160, 768, 12
rng = np.random.default_rng(1)
lens = np.full(n_seq, L, np.int32)
qkv = (rng.standard_normal((n_seq * L, 3 * d)) * 1.2).astype(np.float16)
out = np.zeros((n_seq * L, d), np.float16)
_i32p, _u16p = ctypes.POINTER(ctypes.c_int32), ctypes.POINTER(ctypes.c_uint16)
for _ in range(2):
    assert lib.bk_test_attention(0, n_seq, lens.ctypes.data_as(_i32p), d, H, qkv.view(np.uint16).ctypes.data_as(_u16p), 0,
                                 out.view(np.uint16).ctypes.data_as(_u16p)) == 0
tr = np.zeros((8, 1024), np.uint64)
lib.bk_debug_att_trace.argtypes = [ctypes.c_void_p]
assert lib.bk_debug_att_trace(tr.ctypes.data) == 0
NAMES = {0: "entry", 13: "after-setup-sync", 14: "done", 15: "after-exit-sync", 1: "B:wait-S", 2: "B:S-ready", 3: "B:loaded", 4: "B:computed", 5: "B:P-stored",
         6: "B:handed", 7: "A:wait-S", 8: "A:S-ready", 9: "A:done", 10: "max:publish", 11: "max:synced", 12: "epilogue-done", 19: "ITEM",
         16: "epi:O-complete", 17: "epi:O-loaded", 18: "epi:scaled", 20: "S:job", 21: "S:waits-done", 22: "S:committed", 30: "PV:wait0", 31: "PV:wait1", 32: "PV:P0-ready", 33: "PV:P1-ready", 34: "PV:committed"}
t0 = min(int(tr[w][0] & 0xFFFFFFFFFFFF) for w in range(7) if tr[w][0])
for w, name in enumerate(os.environ.get("TRACE_NAMES", "softmax warp 0 (g0),softmax warp 5 (g1),softmax warp 10 (g2),softmax warp 15 (g3),score warp,P.V warp,output warp 20").split(",")):
    print(f"== {name}")
    prev = None
    for v in tr[w]:
        v = int(v)
        if v == 0 and prev is not None:
            break
        tag, t = v >> 48, (v & 0xFFFFFFFFFFFF) - t0
        print(f"  {t:8d} {'' if prev is None else '+%d' % (t - prev):>8s}  {NAMES.get(tag, tag)}")
        prev = t
