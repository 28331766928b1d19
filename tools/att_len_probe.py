"""Attention kernel class time per 65 536-token micro-batch at fixed sentence lengths (engine staged run, CUDA events per class):
    [BERT_B200_ATT_LOOP=..] [BERT_B200_ATT_PERSIST=..] python tools/att_len_probe.py [ftype]"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
binding = bench._load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
ftype = sys.argv[1] if len(sys.argv) > 1 else "f16"
path = os.path.join(tempfile.gettempdir(), f"attprobe_{ftype}.gguf")
if not os.path.exists(path):
    gio.make_model(path, "bert-base", ftype, seed=3, n_vocab=1000)
hp = gio.read_gguf(path).hparams()
eng = binding.Engine(path, cap_tokens=65536, cap_seqs=8192)
row = []
LENS = [int(x) for x in sys.argv[2].split(",")] if len(sys.argv) > 2 else [16, 32, 48, 49, 64, 96, 128]
for L in LENS:
    batch, _ = bench.make_batch(hp, 65536 // L if 65536 // L <= 8192 else 8192, L, seed=L)
    eng.stage(batch)
    eng.run_staged(2)
    eng.run_staged(4, time_kernels=True)
    ms, n = eng.kernel_ms()["attention"]
    row.append(f"L={L}: {1e3 * ms / n:.1f} us")
print(f"{ftype} LOOP={os.environ.get('BERT_B200_ATT_LOOP', 'default')} PERSIST={os.environ.get('BERT_B200_ATT_PERSIST', 'default')} STREAM={os.environ.get('BERT_B200_ATT_STREAM', 'default')} PACK={os.environ.get('BERT_B200_ATT_PACK', 'default')} DUO={os.environ.get('BERT_B200_ATT_DUO', 'default')} PT={os.environ.get('BERT_B200_ATT_PT', 'default')} XS={os.environ.get('BERT_B200_ATT_XS', 'default')}: " + "  ".join(row), flush=True)
eng.close()
