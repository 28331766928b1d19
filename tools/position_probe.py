"""position invariance: the first n sentences evaluated alone and as the head of a 2n batch must give the same hidden states
   python tools/position_probe.py bert-base f16 64 128"""
import os, sys
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench

shape, ftype, n_sent, seq_len = sys.argv[1], sys.argv[2], int(sys.argv[3]), int(sys.argv[4])
gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
class _One:
    rank = 0; world = 1; local_rank = 0
    def barrier(self): pass
path = bench.model_file(gio, shape, ftype, _One(), n_vocab=30522)
hp = gio.read_gguf(path).hparams()
batch, lens = bench.make_batch(hp, 2 * n_sent, seq_len, seed=1234)
T = n_sent * seq_len
for fused in ("default", "-1"):
    if fused != "default":
        os.environ["BERT_B200_FUSED_LN_TOKENS"] = fused
    e = binding.Engine(path, cap_tokens=2 * T, cap_seqs=8192)
    e.keep_hidden(True)
    a = e.forward(batch[:n_sent]).copy()
    ha = [e.hidden(l, T).copy() for l in range(hp.n_layer)]
    b = e.forward(batch).copy()
    hb = [e.hidden(l, 2 * T)[:T].copy() for l in range(hp.n_layer)]
    e.close()
    print(f"fused_ln={fused}: output rows differing {np.flatnonzero((a != b[:n_sent]).any(axis=1)).tolist()[:12]}")
    for l in range(min(hp.n_layer, 3)):
        rows = np.flatnonzero((ha[l] != hb[l]).any(axis=1))
        if rows.size:
            cols = np.flatnonzero((ha[l] != hb[l]).any(axis=0))
            print(f"  layer {l}: {rows.size} rows differ [{rows.min()}..{rows.max()}], {cols.size} cols [{cols.min()}..{cols.max()}], max-abs {np.abs(ha[l] - hb[l]).max():.3e}")
        else:
            print(f"  layer {l}: identical")
