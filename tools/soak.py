"""randomised soak of bert_eval_batch: random batch compositions (1..96 sentences of 1..512 tokens) must reproduce, sentence by
sentence, the embedding of that sentence evaluated alone -- bit for bit where both evaluations take the same kernels, to
cos >= 0.99999 otherwise (fused / unfused LayerNorm input rounding) -- and stay finite and unit-norm.
   python tools/soak.py mini384 f16 60      (shape, weights, seconds)"""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

shape, ftype, secs = sys.argv[1], sys.argv[2], float(sys.argv[3])
gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
class _One:
    rank = 0; world = 1; local_rank = 0
    def barrier(self): pass
path = bench.model_file(gio, shape, ftype, _One(), n_vocab=30522)
hp = gio.read_gguf(path).hparams()
m = binding.BertB200(path)
rng = np.random.default_rng(int(time.time()) & 0xffff)
pool_lens = [1, 2, 15, 16, 17, 31, 33, 47, 48, 49, 63, 64, 65, 100, 127, 128, 129, 200, 255, 256, 257, 300, 384, 385, 511, 512]
pool_lens += [int(v) for v in rng.integers(1, min(513, hp.n_max_tokens + 1), 40)]
pool_lens = [L for L in pool_lens if L <= hp.n_max_tokens]
pool = [rng.integers(1000, 30000, L).astype(np.int32) for L in pool_lens]
alone = [m.eval(t).copy() for t in pool]
t_end = time.time() + secs
n_batches = n_sent = n_exact = 0
worst = 1.0
while time.time() < t_end:
    k = int(rng.integers(1, 97))
    idx = rng.integers(0, len(pool), k)
    out = m.eval_batch([pool[i] for i in idx])
    assert np.isfinite(out).all(), "non-finite output"
    assert np.allclose(np.linalg.norm(out, axis=1), 1.0, atol=1e-3), "norm"
    for row, i in zip(out, idx):
        if np.array_equal(row, alone[i]):
            n_exact += 1
        else:
            c = float(np.dot(row, alone[i]))
            worst = min(worst, c)
            assert c >= 0.99999 or ftype.startswith("q4") and c >= 0.9995, (len(pool[i]), c)
    n_batches += 1; n_sent += k
print(f"{shape} {ftype}: {n_batches} batches, {n_sent} sentences, {n_exact} bit-identical to the stand-alone evaluation, worst cosine of the rest {worst:.7f}")
m.close()
