"""run bert_eval_batch repeatedly under several micro-batch sizes / engine counts and compare the outputs bit for bit
   python tools/determinism_probe.py bert-base f16 1024 128"""
import ctypes, os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

shape, ftype, n_sent, seq_len = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
reps = int(sys.argv[5]) if len(sys.argv) > 5 else 4
gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
lib = binding.load_library()
class _One:
    rank = 0; world = 1; local_rank = 0
    def barrier(self): pass
path = bench.model_file(gio, shape, ftype, _One(), n_vocab=30522)
hp = gio.read_gguf(path).hparams()
batch, lens = bench.make_batch(hp, n_sent, seq_len, seed=1234)
ref = None
for devs, mb in (("0", "65536"), ("0,0", "65536"), ("0", "32768"), ("0", "16384"), ("0", "8192")):
    os.environ["BERT_B200_DEVICES"] = devs
    os.environ["BERT_B200_MB_TOKENS"] = mb
    model = binding.BertB200(path)
    toks, tp, lens_c, op, out = model.make_batch_args(batch)
    lp = lens_c.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
    for r in range(reps):
        out[:] = 0
        lib.bert_eval_batch(model.ctx, 1, n_sent, tp, lp, op)
        o = out.copy()
        if ref is None:
            ref = o
        diff = np.flatnonzero((o != ref).any(axis=1))
        msg = "identical" if diff.size == 0 else f"{diff.size} sentences differ, first {diff[:8].tolist()} max-abs {np.abs(o - ref).max():.3e}"
        print(f"devs={devs} mb={mb} rep {r}: {msg}", flush=True)
    model.close()
