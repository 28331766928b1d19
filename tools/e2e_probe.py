"""time bert_eval_batch (host buffers, the e2e leg of bench.py) under a given environment, e.g.
   BERT_B200_DEVICES=0,0 BERT_B200_MB_TOKENS=32768 python tools/e2e_probe.py bert-base f16 1024 128"""
import ctypes, importlib.util, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

shape, ftype, n_sent, seq_len = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
steps = int(sys.argv[5]) if len(sys.argv) > 5 else 10
gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
lib = binding.load_library()
class _One:
    rank = 0; world = 1; local_rank = 0
    def barrier(self): pass
path = bench.model_file(gio, shape, ftype, _One(), n_vocab=30522)
hp = gio.read_gguf(path).hparams()
batch, lens = bench.make_batch(hp, n_sent, seq_len, seed=1234)
model = binding.BertB200(path)
toks, tp, lens_c, op, out = model.make_batch_args(batch)
lp = lens_c.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
for _ in range(3):
    lib.bert_eval_batch(model.ctx, 1, n_sent, tp, lp, op)
t0 = time.perf_counter()
for _ in range(steps):
    lib.bert_eval_batch(model.ctx, 1, n_sent, tp, lp, op)
dt = time.perf_counter() - t0
print(f"DEVICES={os.environ.get('BERT_B200_DEVICES')} MB={os.environ.get('BERT_B200_MB_TOKENS')} {shape} {ftype} {n_sent}x{seq_len}: "
      f"{n_sent * steps / dt:.0f} emb/s  {dt / steps * 1e3:.2f} ms/step  checksum {float(out.sum()):.6f}")
model.close()
