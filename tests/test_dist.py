"""N>1 host logic on CPU: two `gloo` ranks run bench.py's sharding / reduction plumbing (no GPU, no CUDA calls).

The data path has no collective (sentences are independent, bert.cpp:1065): ranks take disjoint shards and
torch.distributed is used only for the barrier and the max / sum over ranks of the device time."""
import importlib.util
import json
import os
import socket
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))

WORKER = r'''
import importlib.util, json, os, sys
import numpy as np
ROOT = sys.argv[1]
spec = importlib.util.spec_from_file_location("bench", os.path.join(ROOT, "bench.py"))
bench = importlib.util.module_from_spec(spec); sys.modules["bench"] = bench; spec.loader.exec_module(bench)
gio = bench._load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
dist = bench.Dist(want_gpu=False)                      # gloo on CPU
hp = gio.BertHParams(n_vocab=500)
batch, lens = bench.make_batch(hp, 16, 0, seed=1234 + dist.rank)      # this rank's shard, as bench.py main() draws it
dist.barrier()
out = {"rank": dist.rank, "world": dist.world, "max": dist.max(10.0 + dist.rank), "sum": dist.sum(float(lens.sum())),
       "tokens": int(lens.sum()), "first": [int(x) for x in batch[0][:8]], "sorted": bool((np.diff(lens) <= 0).all())}
print("RESULT " + json.dumps(out), flush=True)
dist.close()
'''


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def test_two_rank_gloo_sharding_and_reductions(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(WORKER)
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), CUDA_VISIBLE_DEVICES="")
        procs.append(subprocess.Popen([sys.executable, str(script), ROOT], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    res = []
    for p in procs:
        out, err = p.communicate(timeout=240)
        assert p.returncode == 0, err[-2000:]
        res.append(json.loads([l for l in out.splitlines() if l.startswith("RESULT ")][0][7:]))
    res.sort(key=lambda r: r["rank"])
    assert [r["world"] for r in res] == [2, 2]
    assert res[0]["max"] == res[1]["max"] == 11.0                        # time = max over ranks
    assert res[0]["sum"] == res[1]["sum"] == res[0]["tokens"] + res[1]["tokens"]   # units = sum over ranks
    assert res[0]["first"] != res[1]["first"]                            # disjoint synthetic shards (seed + rank)
    assert all(r["sorted"] for r in res)                                 # varlen shards are length-sorted per rank


def test_reference_arm_runs_on_rank0_only(tmp_path):
    """--impl reference under torchrun: rank 0 prints the line, the other rank exits 0 silently."""
    port = _free_port()
    procs = []
    for rank in range(2):
        env = dict(os.environ, RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE="2", MASTER_ADDR="127.0.0.1",
                   MASTER_PORT=str(port), CUDA_VISIBLE_DEVICES="")
        procs.append(subprocess.Popen([sys.executable, os.path.join(ROOT, "bench.py"), "--impl", "reference", "--gpus", "2", "--steps", "1",
                                       "--warmup", "1", "--workload", "minilm-l6-q4_0-b256xs128", "--vocab", "500", "--ref-sentences", "1"],
                                      env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = []
    for p in procs:
        out, err = p.communicate(timeout=600)
        assert p.returncode == 0, err[-2000:]
        outs.append(out)
    lines = [l for l in outs[0].splitlines() if l.startswith("{")]
    assert len(lines) == 1 and not [l for l in outs[1].splitlines() if l.startswith("{")]
    j = json.loads(lines[0])
    assert j["impl"] == "reference" and j["n_gpus"] == 2 and j["value"] > 0
    assert j["e2e"]["h2d_bytes_per_step"] == 0 and j["cpu_baseline"]["kind"] in ("reference", "port")
