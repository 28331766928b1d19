"""batch-1 latency of bert_eval on one sentence: python tools/latency_probe.py bert-base f16 128 [reps]
   (run under `ncu --metrics gpu__time_duration.sum` with BERT_B200_GRAPH_TOKENS=0 to list the kernels of one call)"""
import ctypes, os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

shape, ftype, L = sys.argv[1], sys.argv[2], int(sys.argv[3])
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 300
gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
class _One:
    rank = 0; world = 1; local_rank = 0
    def barrier(self): pass
path = bench.model_file(gio, shape, ftype, _One(), n_vocab=30522)
m = binding.BertB200(path)
rng = np.random.default_rng(1)
toks = rng.integers(1000, 30000, L).astype(np.int32)
for _ in range(5):
    m.eval(toks)
ts = []
for _ in range(reps):
    t0 = time.perf_counter(); m.eval(toks); ts.append(time.perf_counter() - t0)
ts = np.sort(np.array(ts)) * 1e3
print(f"{shape} {ftype} L={L}: p50 {ts[len(ts)//2]:.3f} ms  p90 {ts[int(len(ts)*0.9)]:.3f} ms  min {ts[0]:.3f} ms")
m.close()
