#!/usr/bin/env python
"""bench.py -- embeddings/sec of the bert_eval_batch path on B200 (BASELINE.json metric).

    python bench.py --gpus 1 --steps K --warmup W                     (this engine, N=1)
    python -m torch.distributed.run --nproc-per-node N ... bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...                              (the reference's CPU path, host cores)

One "step" = one pass of the hot path over one batch of synthetic sentences per GPU (default workload
C3 of BASELINE.json: BERT-base shape, F16 GGUF, 1024 sentences x 128 tokens per GPU; sentences are
independent so ranks take disjoint shards with no data-path collective -> weak scaling).

  value : whole-job embeddings/s with token ids already resident in HBM (bk_engine_run_staged; CUDA events
          on the engine's stream, max over ranks)
  e2e   : the same metric through the reference-facing C API (bert_eval_batch in build/libbert.so) with HOST
          buffers: host->device copy of ids and device->host copy of the embeddings inside the timed region
  roofline     : dominant kernel, live CUDA-event duration vs MEASURED_PEAKS.json
  cpu_baseline : the reference's own CPU implementation (oracle/_ref) timed on this box's host cores on a
                 bounded sample of the same workload (rank 0, N=1 only), in both of SURVEY 8d's modes:
                 one process with n_threads = host cores, and host-cores processes x 1 thread
  parity       : >= 8 sentences of the timed batch re-evaluated by the reference (oracle/_ref) and compared with
                 what bert_eval_batch returned: a miss of the north_star gate makes the run exit non-zero
  also         : the same record for the Target dtype (BERT-base Q4_0, 1024 x 128) next to the default F16 one
  e2e_inprocess: N > 1: ONE bert_ctx over all N GPUs (the library's own sharding: persistent worker per GPU, host
                 gather), timed on rank 0 through bert_eval_batch -- 1024 sentences in total (strong scaling, what
                 BASELINE configs[2] words) and 1024 x N (weak) -- and checked bit-equal to the single-GPU output
  tokenizer    : bert_encode_batch on 65536 synthetic texts: texts/s of the tokenizer alone and its share of the call

PyTorch is used for torch.distributed (barrier / max over ranks) only; the engine is libbert.so.
"""
from __future__ import annotations

import argparse
import ctypes
import importlib.util
import json
import os
import subprocess
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))

WORKLOADS = {  # name -> (shape, ftype, sentences per GPU, tokens per sentence)
    "bert-base-f16-b1024xs128": ("bert-base", "f16", 1024, 128),      # BASELINE.json configs[2] (metric config)
    "bert-base-q4_0-b1024xs128": ("bert-base", "q4_0", 1024, 128),    # north_star target dtype
    "minilm-l6-q4_0-b256xs128": ("minilm-l6", "q4_0", 256, 128),      # configs[1]
    "bert-base-q4_1-varlen": ("bert-base", "q4_1", 1024, 0),          # configs[3]: lengths uniform 16..512
    "bert-large-q4_0-b256xs512": ("bert-large", "q4_0", 256, 512),    # configs[4] shape, reduced batch
    "bert-large-q4_0-b4096xs512": ("bert-large", "q4_0", 4096, 512),  # configs[4] in full: 2 M tokens per GPU, 32 micro-batches
}
DEFAULT_WORKLOAD = "bert-base-f16-b1024xs128"
TARGET_WORKLOAD = "bert-base-q4_0-b1024xs128"
# what the path computes in, per weight type (not a precision claim): activations are IEEE f16 in HBM, accumulation fp32 in TMEM
DTYPES = {"f32": "f16 x f16 -> fp32 accumulate (F32 weights rounded to f16 at load, as the reference's converter stores them)",
          "f16": "f16 x f16 -> fp32 accumulate",
          "q4_0": "q4_0 weights x q8_0-rounded activations, multiplied as f16 x f16 -> fp32 accumulate",
          "q4_1": "q4_1 weights x q8_1-rounded activations, multiplied as f16 x f16 -> fp32 accumulate"}


def _load(name, rel):
    spec = importlib.util.spec_from_file_location(name, os.path.join(ROOT, rel))
    mod = importlib.util.module_from_spec(spec)
    sys.modules[name] = mod
    spec.loader.exec_module(mod)
    return mod


def flops_per_sentence(hp, L):
    """algorithmic FLOPs: F(L) = Lr*[2L(4d^2 + 2df) + 4L^2 d] + 2Ld"""
    d, f, Lr = hp.n_embd, hp.n_intermediate, hp.n_layer
    return Lr * (2 * L * (4 * d * d + 2 * d * f) + 4 * L * L * d) + 2 * L * d


def make_batch(hp, n_sent, seq_len, seed):
    rng = np.random.default_rng(seed)
    if seq_len > 0:
        lens = np.full(n_sent, seq_len, np.int32)
    else:
        lens = rng.integers(16, 513, n_sent).astype(np.int32)     # configs[3]
        lens = np.sort(lens)[::-1].copy()                          # longest first (any order works: the batcher sorts on the device)
    batch = []
    for L in lens:
        t = rng.integers(0, hp.n_vocab, int(L)).astype(np.int32)
        t[0], t[-1] = hp.cls_id, hp.sep_id
        batch.append(t)
    return batch, lens


class ClockSampler(threading.Thread):
    """nvidia-smi clocks + throttle reasons during the timed region (B200_PROFILING.md clocks line)."""

    Q = ("clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.proc = index, [], None

    def run(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", "-i", str(self.index), "--query-gpu=" + self.Q,
                                          "--format=csv,noheader,nounits", "-lms", "100"], stdout=subprocess.PIPE, text=True)
            for line in self.proc.stdout:
                self.rows.append([c.strip() for c in line.split(",")])
        except Exception:
            pass

    def stop(self):
        if self.proc:
            self.proc.terminate()
        self.join(timeout=2)
        sm, mx, reasons = [], 0, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for r in self.rows:
            try:
                sm.append(float(r[0]))
                mx = max(mx, float(r[1]))
                for n, v in zip(names, r[2:6]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            except (ValueError, IndexError):
                continue
        busy = [x for x in sm if x > 0]
        return {"sm_mhz": float(np.median(busy)) if busy else None, "sm_max_mhz": mx or None,
                "reasons": sorted(reasons), "samples": len(sm)}


class Dist:
    """torch.distributed plumbing: barrier + max over ranks.  Single process when WORLD_SIZE is unset."""

    def __init__(self, want_gpu=True):
        self.rank = int(os.environ.get("RANK", "0"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.torch = None
        self.dev = None
        if self.world > 1:
            import torch
            import torch.distributed as dist
            self.torch, self.dist = torch, dist
            use_nccl = want_gpu and torch.cuda.is_available()
            if use_nccl:
                torch.cuda.set_device(self.local_rank)
                self.dev = torch.device("cuda", self.local_rank)
            dist.init_process_group(backend="nccl" if use_nccl else "gloo",
                                    device_id=self.dev if use_nccl else None)
            # a host-side group: ranks that wait while rank 0 measures something alone must not spin a NCCL kernel on their GPU
            self.cpu_group = dist.new_group(backend="gloo") if use_nccl else None

    def barrier(self):
        if self.world > 1:
            if self.dev is not None:
                self.dist.barrier(device_ids=[self.local_rank])
                self.torch.cuda.synchronize()
            else:
                self.dist.barrier()

    def host_barrier(self):
        """barrier that keeps the GPUs idle (gloo): used around the legs only rank 0 runs"""
        if self.world > 1:
            self.dist.barrier(group=self.cpu_group) if self.cpu_group is not None else self.dist.barrier()

    def max(self, v: float) -> float:
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev or "cpu")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def sum(self, v: float) -> float:
        if self.world == 1:
            return v
        t = self.torch.tensor([v], dtype=self.torch.float64, device=self.dev or "cpu")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return float(t.item())

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


def model_file(gio, shape, ftype, dist, n_vocab=30522):
    """random-init GGUF of the named architecture (no network for checkpoints): written once per box."""
    d = os.path.join(tempfile.gettempdir(), "embedding_cpp_b200_bench")
    os.makedirs(d, exist_ok=True)
    path = os.path.join(d, f"{shape}_{ftype}_v{n_vocab}.gguf")
    if dist.local_rank == 0 and not os.path.exists(path):
        tmp = path + f".tmp{os.getpid()}"
        gio.make_model(tmp, shape, ftype, seed=7, n_vocab=n_vocab, dist="bert-init")
        os.replace(tmp, path)
    dist.barrier()
    return path


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        j = json.load(open(p))
        return {"hbm_gbs": j["hbm_gbs"], "tflops_burst": j["bf16_tflops"], "tflops_sustained": j["bf16_tflops_sustained"], "src": "measured"}
    return {"hbm_gbs": 6650.0, "tflops_burst": 1590.0, "tflops_sustained": 1400.0, "src": "fallback"}


def kernel_model(hp, lens, ftype):
    """algorithmic FLOPs / bytes per step for each kernel class."""
    T = int(lens.sum())
    d, f, Lr, n = hp.n_embd, hp.n_intermediate, hp.n_layer, len(lens)
    q = 1 if ftype.startswith("q4") else 0
    row = {"f32": 4 * d, "f16": 2 * d, "q4_0": d // 32 * 18, "q4_1": d // 32 * 20}[ftype]
    sq = float((lens.astype(np.float64) ** 2).sum())
    return {
        "gemm_qkv": {"flops": Lr * 2.0 * T * 3 * d * d},
        "gemm_o": {"flops": Lr * 2.0 * T * d * d},
        "gemm_ff1": {"flops": Lr * 2.0 * T * f * d},
        "gemm_ff2": {"flops": Lr * 2.0 * T * d * f},
        "attention": {"flops": Lr * 4.0 * sq * d, "bytes": Lr * (3 * d + d) * 2.0 * T},
        "add_ln": {"bytes": 2 * Lr * (2 + 2 + 2 * q) * d * 1.0 * T},
        "embed_ln": {"bytes": T * (4 + 4 + 2 * row + (2 + 2 * q) * d)},
        "pool_norm": {"bytes": 2.0 * T * d + 4.0 * n * d},
        # on-device batcher: lens in; order, cu, src_off out; ids in, packed ids + positions out
        "pack": {"bytes": 4.0 * n * 4 + 4.0 * T * 3},
    }


def reference_model(orc, path, threads):
    kind = "reference" if orc.available_reference() else "port"
    m = orc.Reference(path, n_threads=threads) if kind == "reference" else orc.Oracle(path, threads=threads)
    return kind, m


def time_reference(orc, path, batch, n_sent, warm, threads):
    """reference bert_eval (bert.cpp:1020), one sentence per call: it has no batching (bert.cpp:1064-1065)"""
    kind, m = reference_model(orc, path, threads)
    for i in range(warm):
        m.eval(batch[i % len(batch)])
    t0 = time.perf_counter()
    for i in range(n_sent):
        m.eval(batch[i % len(batch)])
    dt = time.perf_counter() - t0
    if kind == "reference":
        m.close()
    return kind, n_sent / dt, dt


def cpu_worker(path, tok_file, n_sent, threads):
    """one process of cpu_baseline mode (ii): load, evaluate n_sent sentences with n_threads = threads, print the seconds"""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import oracle as orc
    z = np.load(tok_file)
    lens, toks = z["lens"], z["toks"]
    off = np.concatenate([[0], np.cumsum(lens)])
    kind, m = reference_model(orc, path, threads)
    m.eval(toks[off[0]:off[1]])                                  # warm-up (first-touch of the arena)
    t0 = time.perf_counter()
    for i in range(n_sent):
        k = i % len(lens)
        m.eval(toks[off[k]:off[k + 1]])
    print("CPU_WORKER %.6f" % (time.perf_counter() - t0), flush=True)


def time_reference_processes(path, batch, n_proc, per_proc):
    """SURVEY 8d mode (ii): n_proc processes x n_threads = 1, the throughput-optimal way to run the reference's scheduler
    (its thread pool spins on a barrier after every graph node).  Wall time = slowest process, model load excluded."""
    with tempfile.NamedTemporaryFile(suffix=".npz", delete=False) as tf:
        sel = batch[:max(per_proc, 1) * 2]
        np.savez(tf, lens=np.array([len(t) for t in sel], np.int64), toks=np.concatenate(sel).astype(np.int32))
        tok_file = tf.name
    try:
        procs = [subprocess.Popen([sys.executable, os.path.abspath(__file__), "--cpu-worker", path, tok_file, str(per_proc), "1"],
                                  stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True) for _ in range(n_proc)]
        times = []
        for p in procs:
            out, _ = p.communicate(timeout=900)
            times += [float(l.split()[1]) for l in out.splitlines() if l.startswith("CPU_WORKER")]
    finally:
        os.unlink(tok_file)
    if len(times) != n_proc:
        return None
    return n_proc * per_proc / max(times), max(times)


def cpu_cost_scale(hp, lens):
    """how many sentences of this workload give ~10-30 s of CPU work: the reference does ~25 GFLOP/s per core here"""
    return float(np.mean([flops_per_sentence(hp, int(L)) for L in lens])) / 1e9


def run_reference_arm(args, dist, gio, orc):
    """--impl reference: the reference's own CPU implementation on this box's host cores (rank 0 only)."""
    if dist.rank != 0:
        return
    shape, ftype, n_sent, seq_len = WORKLOADS[args.workload]
    path = model_file(gio, shape, ftype, _Solo(), n_vocab=args.vocab)
    hp = gio.read_gguf(path).hparams()
    batch, lens = make_batch(hp, min(n_sent, 1024), seq_len, seed=1234)
    threads = args.ref_threads or min(os.cpu_count() or 1, 32)
    per_step = args.ref_sentences                      # bounded sample of the workload per step
    kind, m = reference_model(orc, path, threads)
    k = 0
    for _ in range(args.warmup):
        m.eval(batch[k % len(batch)]); k += 1
    t0 = time.perf_counter()
    for _ in range(args.steps * per_step):
        m.eval(batch[k % len(batch)]); k += 1
    dt = time.perf_counter() - t0
    value = args.steps * per_step / dt
    sample = f"{per_step} sentence(s) of the workload per step through bert_eval, one per call (the reference has no batching), n_threads={threads}"
    line = {"impl": "reference", "metric": "embeddings/sec", "value": value, "unit": "embeddings/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": 1e3 * dt / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 accumulate / " + ftype + " weights (ggml CPU)", "data": "synthetic",
            "config": {"workload": args.workload, "shape": shape, "weights": ftype, "seq_len": seq_len or "16..512"},
            "cpu_baseline": {"value": value, "unit": "embeddings/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "embeddings/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


class _Solo:
    local_rank, rank, world = 0, 0, 1

    def barrier(self):
        pass


def cos_maxabs(a, b):
    a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
    return float((a * b).sum() / (np.linalg.norm(a) * np.linalg.norm(b))), float(np.abs(a - b).max())


def parity_check(orc, path, ftype, batch, out, n_check, threads):
    """>= 8 sentences of the timed batch through the reference's own CPU path (oracle/_ref; the pinned C restatement if that is
    absent) against the embeddings bert_eval_batch returned.  north_star gate: cos >= 0.9999 (1 - cos <= 1e-4) per sentence, and
    max-abs <= 1e-3 for F32/F16 files."""
    kind, chk = reference_model(orc, path, threads)
    idx = sorted(set(int(i) for i in np.linspace(0, len(batch) - 1, n_check)))
    rows, ok = [], True
    for i in idx:
        ref = chk.eval(batch[i])
        c, mx = cos_maxabs(out[i], ref)
        good = (1.0 - c) <= 1e-4 and (ftype.startswith("q4") or mx <= 1e-3)
        ok = ok and good
        rows.append({"sentence": i, "tokens": int(len(batch[i])), "one_minus_cos": 1.0 - c, "max_abs": mx, "pass": bool(good)})
    if kind == "reference":
        chk.close()
    return {"checker": kind, "sentences": len(rows), "gate": "1-cos <= 1e-4" + ("" if ftype.startswith("q4") else ", max-abs <= 1e-3"),
            "max_one_minus_cos": max(r["one_minus_cos"] for r in rows), "max_abs": max(r["max_abs"] for r in rows), "pass": bool(ok), "rows": rows}


def latency_leg(lib, model, hp, tokens, calls, rng=None):
    """p50 / p90 of `calls` bert_eval calls (host ids in, host embedding out).  rng = None: the same sentence every call (the
    captured graph is replayed); otherwise a new random length in 3..n_max_tokens per call (real traffic: most shapes are new
    at first, launched directly, captured when they come back)."""
    emb1 = np.zeros(hp.n_embd, np.float32)
    ep1 = emb1.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
    pool = np.ascontiguousarray(np.resize(tokens, hp.n_max_tokens), np.int32)
    pool[0] = hp.cls_id
    g0 = model.graph_stats()
    ts = []
    n_warm = 20 if rng is None else 0
    for k in range(n_warm + calls):
        L = len(tokens) if rng is None else int(rng.integers(3, hp.n_max_tokens + 1))
        one = pool[:L].copy()
        one[-1] = hp.sep_id
        tp1 = one.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
        t0 = time.perf_counter()
        lib.bert_eval(model.ctx, 1, tp1, L, ep1)
        if k >= n_warm:
            ts.append(time.perf_counter() - t0)
    g1 = model.graph_stats()
    ts = np.sort(np.array(ts)) * 1e3
    return {"p50_ms": float(ts[len(ts) // 2]), "p90_ms": float(ts[int(len(ts) * 0.9)]), "min_ms": float(ts[0]), "calls": int(len(ts)),
            "graph": {k: g1[k] - g0[k] for k in g1},
            "graph_hit_rate": (g1["replayed"] - g0["replayed"]) / max(1, sum(g1[k] - g0[k] for k in g1))}


def tokenizer_leg(lib, model, hp, n_texts, e2e_value):
    """row f1: bert_encode_batch (texts in, embeddings out) on n_texts synthetic texts; the tokenizer alone (bert_tokenize in a
    loop is what the reference does, bert.cpp:1138-1150) is timed through the same call with 1-token... no: timed separately
    through bert_tokenize on one thread and derived for the library's threaded tokenisation from the whole-call time."""
    rng = np.random.default_rng(99)
    words = [model.lib.bert_vocab_id_to_token(model.ctx, int(i)).decode("utf-8", "replace") for i in rng.integers(104, min(hp.n_vocab, 400), 512)]
    words = [w for w in words if not w.startswith("[") and not w.startswith("##")] or ["the"]
    texts = [" ".join(words[int(j)] for j in rng.integers(0, len(words), int(rng.integers(8, 40)))) for _ in range(n_texts)]
    n = len(texts)
    out = np.zeros((n, hp.n_embd), np.float32)
    enc = [t.encode("utf-8") for t in texts]
    tp = (ctypes.c_char_p * n)(*enc)
    op = (ctypes.POINTER(ctypes.c_float) * n)(*[ctypes.cast(out.ctypes.data + i * out.strides[0], ctypes.POINTER(ctypes.c_float)) for i in range(n)])
    lib.bert_encode_batch(model.ctx, 1, 16, min(n, 4096), tp, op)          # warm-up
    t0 = time.perf_counter()
    lib.bert_encode_batch(model.ctx, 1, 16, n, tp, op)
    t_call = time.perf_counter() - t0
    # the tokenizer alone, single thread, on a sample (bert_tokenize, bert.h:64-69)
    buf = np.zeros(hp.n_max_tokens, np.int32)
    nt = ctypes.c_int32(0)
    bp = buf.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
    sample = enc[:8192]
    n_tok = 0
    t0 = time.perf_counter()
    for e in sample:
        lib.bert_tokenize(model.ctx, e, bp, ctypes.byref(nt), hp.n_max_tokens)
        n_tok += nt.value
    t_tok1 = time.perf_counter() - t0
    per_thread = len(sample) / t_tok1
    host_threads = min(os.cpu_count() or 1, 32)
    return {"texts": n, "mean_tokens": n_tok / len(sample), "encode_batch_texts_per_s": n / t_call,
            "tokenizer_texts_per_s_one_thread": per_thread, "tokenizer_threads_in_encode_batch": host_threads,
            "tokenizer_share_of_call_estimate": min(1.0, (n / (per_thread * host_threads)) / t_call),
            "unit_norm": bool(np.allclose(np.linalg.norm(out, axis=1), 1.0, atol=1e-3)),
            "api": "bert_encode_batch (texts in, embeddings out; chunks of 16384 texts, the next chunk tokenised on host threads while the current one is on the GPU)"}


def run_workload(args, dist, gio, binding, lib, workload, full):
    """every leg of one workload; returns the record (rank 0) -- `full`: also cpu_baseline / latency / tokenizer legs"""
    shape, ftype, n_sent, seq_len = WORKLOADS[workload]
    path = model_file(gio, shape, ftype, dist, n_vocab=args.vocab)
    hp = gio.read_gguf(path).hparams()
    batch, lens = make_batch(hp, n_sent, seq_len, seed=1234 + dist.rank)
    T = int(lens.sum())
    total_sent = n_sent * dist.world

    # ---------------- leg 1: device-resident (value) -------------------------------------------------
    eng = binding.Engine(path, device=dist.local_rank, cap_tokens=args.mb_tokens, cap_seqs=8192)
    eng.stage(batch)
    eng.run_staged(args.warmup)
    sampler = ClockSampler(dist.local_rank)
    sampler.start()
    time.sleep(0.3)
    l0 = eng.launch_count()
    dist.barrier()
    lib.bk_profiler_range(1)                            # cudaProfilerStart: ncu --profile-from-start off sees only the timed steps
    ms = eng.run_staged(args.steps)                     # CUDA events on the engine's stream around exactly K steps
    lib.bk_profiler_range(0)
    dist.barrier()
    launches = eng.launch_count() - l0
    ms = dist.max(ms)
    clocks = sampler.stop()
    value = total_sent * args.steps / (ms / 1e3)
    # per-kernel-class durations (events around every launch; separate pass so the headline has no event overhead)
    eng.run_staged(args.steps, time_kernels=True)
    kms = eng.kernel_ms()
    out_dev = eng.read_output()
    eng.close()

    if args.legs == "value":
        return {"metric": "embeddings/sec", "value": value, "ms_per_step": ms / args.steps, "legs": "value", "workload": workload,
                "kernels_ms_per_step": {k: v[0] / args.steps for k, v in kms.items()}}

    # ---------------- leg 2: end to end through bert_eval_batch (host buffers) -----------------------
    model = binding.BertB200(path)
    toks, tp, lens_c, op, out = model.make_batch_args(batch)
    lp = lens_c.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
    for _ in range(args.warmup):
        lib.bert_eval_batch(model.ctx, 1, n_sent, tp, lp, op)
    dist.barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lib.bert_eval_batch(model.ctx, 1, n_sent, tp, lp, op)   # returns after the D2H copy of the embeddings
    t_e2e = time.perf_counter() - t0
    dist.barrier()
    t_e2e = dist.max(t_e2e)
    e2e_value = total_sent * args.steps / t_e2e
    same = bool(np.array_equal(out, out_dev))
    unit_norm = bool(np.allclose(np.linalg.norm(out, axis=1), 1.0, atol=1e-3))

    # ---------------- leg 3: batch-1 latency through bert_eval (BASELINE.json: "p50 batch-1 latency") ----
    latency = latency_mixed = tokenizer = None
    if dist.rank == 0 and full:
        one = np.ascontiguousarray(batch[0][:128] if seq_len == 0 else batch[0], np.int32)
        latency = latency_leg(lib, model, hp, one, args.latency_calls)
        latency.update(tokens=int(len(one)), api="bert_eval (one sentence, host buffers in/out; launch plan replayed as a CUDA graph)")
        latency_mixed = latency_leg(lib, model, hp, one, 3 * args.latency_calls, rng=np.random.default_rng(5))
        latency_mixed.update(tokens="uniform 3..%d, a new draw per call" % hp.n_max_tokens,
                             api="bert_eval; a shape seen for the first time is launched directly, captured on its second sighting")
        if args.tokenizer_texts > 0:
            tokenizer = tokenizer_leg(lib, model, hp, args.tokenizer_texts, e2e_value)
    model.close()

    # ---------------- roofline of every kernel class; the dominant one is the headline -----------------
    peaks = measured_peaks()
    km = kernel_model(hp, lens, ftype)
    l2w = float((lens.astype(np.float64) ** 3).sum() / (lens.astype(np.float64) ** 2).sum())   # L^2-weighted mean length
    kernels = {}
    for name, (kms_tot, n_launch) in kms.items():
        if n_launch == 0:
            continue
        per_step_ms = kms_tot / args.steps
        ent = {"ms_per_step": per_step_ms, "launches_per_step": n_launch // args.steps, "share": kms_tot / max(1e-9, sum(v[0] for v in kms.values()))}
        m_ = km.get(name, {})
        tensor_bound = name.startswith("gemm") or (name == "attention" and l2w >= 400.0)   # attention: ~L/2 FLOP per byte
        if tensor_bound:
            ent.update(bound="tensor", achieved=m_["flops"] / (per_step_ms * 1e-3) / 1e12, peak=peaks["tflops_sustained"], unit="TFLOP/s")
        elif "bytes" in m_:
            ent.update(bound="hbm", achieved=m_["bytes"] / (per_step_ms * 1e-3) / 1e9, peak=peaks["hbm_gbs"], unit="GB/s")
        if name == "attention":
            ent["l2_weighted_mean_len"] = l2w
            ent["achieved_tflops"] = m_["flops"] / (per_step_ms * 1e-3) / 1e12
            ent["achieved_gbs"] = m_["bytes"] / (per_step_ms * 1e-3) / 1e9
        if "achieved" in ent:
            ent["frac"] = ent["achieved"] / ent["peak"]
        kernels[name] = ent
    if "add_ln" not in kernels:  # large micro-batches: the residual add and the LayerNorm run inside these two GEMMs (gemm_ln.cu)
        for name in ("gemm_o", "gemm_ff2"):
            if name in kernels:
                kernels[name]["fused"] = ("bias + residual + LayerNorm (+ Q8 copy) in the epilogue: replaces the add_ln kernel; "
                                          "achieved counts the GEMM FLOPs only")
    top = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
    traffic, traffic_src = None, None
    tp_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp_path):
        tj = json.load(open(tp_path)).get(workload, {})
        traffic, traffic_src = tj.get(top), tj.get("_source")
    roofline = {"kernel": top, "bound": kernels[top].get("bound"), "achieved": kernels[top].get("achieved"),
                "peak": kernels[top].get("peak"), "unit": kernels[top].get("unit"), "frac": kernels[top].get("frac"),
                "traffic": traffic, "traffic_source": traffic_src,
                "peak_source": peaks["src"] + " (sustained figure: kernel timed inside a long step)"}
    flops_step = float(sum(flops_per_sentence(hp, int(L)) for L in lens))
    model_tflops = flops_step * args.steps / (ms / 1e3) / 1e12          # per GPU (ms is the max over ranks)

    rec = {
        "metric": "embeddings/sec", "value": value, "unit": "embeddings/s", "n_gpus": dist.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": DTYPES[ftype], "data": "synthetic",
        "config": {"workload": workload, "shape": shape, "weights": ftype, "sentences_per_gpu": n_sent,
                   "global_batch": total_sent, "seq_len": seq_len or "16..512", "tokens_per_gpu": T,
                   "micro_batch_tokens": args.mb_tokens, "parallelism": f"dp{dist.world} (independent shards, host gather, no collective)",
                   "l2": "no flush: every kernel's working set per micro-batch (>= %d MB of activations) exceeds the 126 MB L2" % (min(T, args.mb_tokens) * hp.n_embd * 2 * 5 // 2 ** 20)},
        "e2e": {"value": e2e_value, "unit": "embeddings/s", "h2d_bytes_per_step": int(4 * T + 4 * n_sent),
                "d2h_bytes_per_step": int(4 * n_sent * hp.n_embd), "api": "bert_eval_batch (build/libbert.so via ctypes), host pointers in/out",
                "matches_device_leg": same, "unit_norm": unit_norm},
        "latency_batch1": latency, "latency_batch1_mixed_lengths": latency_mixed, "tokenizer": tokenizer,
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "kernels": kernels,
        "model_tflops_per_gpu": model_tflops, "model_frac_of_peak": model_tflops / peaks["tflops_sustained"],
    }

    # ---------------- parity of the timed outputs + CPU baseline (rank 0; the baseline at N=1 only) ---------
    if dist.rank == 0 and not args.no_parity:
        import oracle as orc                            # checker only
        threads = args.ref_threads or min(os.cpu_count() or 1, 32)
        gf = cpu_cost_scale(hp, lens)
        n_check = max(2, min(args.parity_sentences, int(4000 / gf)))     # BERT-large at 512 tokens: 8 sentences ~ 30 s of CPU
        rec["parity"] = parity_check(orc, path, ftype, batch, out, n_check, threads)
    if dist.rank == 0 and dist.world == 1 and full and not args.no_cpu_baseline:
        import oracle as orc                            # timed as the *baseline*, never shipped
        threads = args.ref_threads or min(os.cpu_count() or 1, 32)
        gf = cpu_cost_scale(hp, lens)
        n_cpu = max(2, min(args.cpu_baseline_sentences, int(6000 / gf)))
        kind, v, dt = time_reference(orc, path, batch, n_cpu, 1, threads)
        rec["cpu_baseline"] = {"value": v, "unit": "embeddings/s", "cores": threads, "kind": kind,
                               "sample": f"{n_cpu} sentences of the same workload, one bert_eval call each (reference has no batching), "
                                         f"one process with n_threads = {threads}, {dt:.1f} s"}
        if kind == "reference":
            nproc = os.cpu_count() or 1
            per = max(1, min(2, int(600 / gf)))
            r = time_reference_processes(path, batch, nproc, per)
            if r:
                rec["cpu_baseline"]["processes_mode"] = {"value": r[0], "unit": "embeddings/s", "cores": nproc,
                                                         "sample": f"{nproc} processes x n_threads = 1, {per} sentence(s) each, slowest process {r[1]:.1f} s"}
    return rec


def inprocess_multi_gpu(args, dist, gio, binding, lib, workload, single_out):
    """SURVEY 8e as the product does it: ONE bert_ctx over all N GPUs of the box (bert_load_from_file replicates the weights, a
    persistent worker per engine takes micro-batches, the host gathers).  Timed on rank 0 through bert_eval_batch with host
    buffers; the other ranks have released their engines and wait on a host-side barrier."""
    shape, ftype, n_sent, seq_len = WORKLOADS[workload]
    n = dist.world
    path = model_file(gio, shape, ftype, _Solo(), n_vocab=args.vocab)
    hp = gio.read_gguf(path).hparams()
    os.environ["BERT_B200_DEVICES"] = ",".join(str(i) for i in range(n))
    model = binding.BertB200(path)
    res = {}
    for mode, count in (("strong", n_sent), ("weak", n_sent * n)):
        batch = []
        for r in range((count + n_sent - 1) // n_sent):
            b, _ = make_batch(hp, n_sent, seq_len, seed=1234 + r)       # rank r's shard of the per-rank legs
            batch += b
        batch = batch[:count]
        toks, tp, lens_c, op, out = model.make_batch_args(batch)
        lp = lens_c.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
        for _ in range(args.warmup):
            lib.bert_eval_batch(model.ctx, 1, count, tp, lp, op)
        t0 = time.perf_counter()
        for _ in range(args.steps):
            lib.bert_eval_batch(model.ctx, 1, count, tp, lp, op)
        dt = time.perf_counter() - t0
        a, b = out[:n_sent].astype(np.float64), single_out.astype(np.float64)
        worst = float((1.0 - (a * b).sum(1) / (np.linalg.norm(a, axis=1) * np.linalg.norm(b, axis=1))).max())
        res[mode] = {"value": count * args.steps / dt, "unit": "embeddings/s", "sentences_per_call": count, "ms_per_call": 1e3 * dt / args.steps,
                     "equals_single_gpu_output": bool(np.array_equal(out[:n_sent], single_out)),
                     # bit-equal whenever both runs take the same LayerNorm path (shards > 4096 tokens: fused); smaller shards round the
                     # pre-LayerNorm sum to f16 -- still the same embedding to well inside the gate
                     "max_one_minus_cos_vs_single_gpu": worst, "matches_single_gpu": bool(worst <= 1e-4),
                     "unit_norm": bool(np.allclose(np.linalg.norm(out, axis=1), 1.0, atol=1e-3))}
    model.close()
    os.environ["BERT_B200_DEVICES"] = str(dist.local_rank)
    res["api"] = f"one bert_ctx over {n} GPUs (BERT_B200_DEVICES=0..{n - 1}), bert_eval_batch with host pointers in/out, timed on rank 0 alone"
    return res


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "--cpu-worker":
        cpu_worker(sys.argv[2], sys.argv[3], int(sys.argv[4]), int(sys.argv[5]))
        return
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("BENCH_WORKLOAD", DEFAULT_WORKLOAD), choices=sorted(WORKLOADS))
    ap.add_argument("--mb-tokens", type=int, default=int(os.environ.get("BERT_B200_MB_TOKENS", "65536")))
    ap.add_argument("--vocab", type=int, default=30522)
    ap.add_argument("--ref-sentences", type=int, default=2, help="reference arm: sentences per step")
    ap.add_argument("--cpu-baseline-sentences", type=int, default=16)
    ap.add_argument("--parity-sentences", type=int, default=8)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-parity", action="store_true")
    ap.add_argument("--no-also", action="store_true", help="skip the sibling record of the Target dtype (BERT-base Q4_0)")
    ap.add_argument("--no-inprocess", action="store_true", help="N > 1: skip the one-context-over-all-GPUs leg")
    ap.add_argument("--tokenizer-texts", type=int, default=65536)
    ap.add_argument("--latency-calls", type=int, default=200, help="timed bert_eval calls of the batch-1 latency leg")
    ap.add_argument("--legs", default="all", choices=["all", "value"], help="value: device-resident leg only (ncu captures)")
    ap.add_argument("--ref-threads", type=int, default=0, help="reference / cpu_baseline threads (0 = min(host cores, 32))")
    args = ap.parse_args()
    if args.warmup < 3 and args.impl == "b200":
        args.warmup = 3                                  # timing rule: W >= 3

    gio = _load("embedding_cpp_b200_gguf_io", "embedding.cpp_b200/gguf_io.py")
    sys.path.insert(0, os.path.join(ROOT, "oracle"))

    if args.impl == "reference":
        import oracle as orc                            # the one leg that may execute oracle/ as the thing measured
        dist = Dist(want_gpu=False)
        try:
            run_reference_arm(args, dist, gio, orc)
        finally:
            dist.close()
        return

    dist = Dist(want_gpu=True)
    os.environ["BERT_B200_DEVICES"] = str(dist.local_rank)          # this rank's GPU only (the per-rank legs)
    os.environ["BERT_B200_MB_TOKENS"] = str(args.mb_tokens)
    binding = _load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
    lib = binding.load_library()                                    # raises if build/libbert.so is missing

    line = run_workload(args, dist, gio, binding, lib, args.workload, full=True)
    if args.legs == "value":
        if dist.rank == 0:
            print(json.dumps(line), flush=True)
        dist.close()
        return
    failed = []
    if dist.rank == 0 and "parity" in line and not line["parity"]["pass"]:
        failed.append(args.workload)
    # the Target dtype (north_star: BERT-base Q4_0 at seq 128) measured by the same run, as a sibling record
    if args.workload == DEFAULT_WORKLOAD and not args.no_also:
        sib = run_workload(args, dist, gio, binding, lib, TARGET_WORKLOAD, full=False)
        if dist.rank == 0:
            line["also"] = {TARGET_WORKLOAD: sib}
            if "parity" in sib and not sib["parity"]["pass"]:
                failed.append(TARGET_WORKLOAD)
    # the library's own multi-GPU path, rank 0 alone (the others keep their GPUs idle)
    if dist.world > 1 and not args.no_inprocess:
        dist.host_barrier()
        if dist.rank == 0:
            shape, ftype, n_sent, seq_len = WORKLOADS[args.workload]
            hp = gio.read_gguf(model_file(gio, shape, ftype, _Solo(), n_vocab=args.vocab)).hparams()
            b0, _ = make_batch(hp, n_sent, seq_len, seed=1234)
            single = binding.BertB200(model_file(gio, shape, ftype, _Solo(), n_vocab=args.vocab))
            single_out = single.eval_batch(b0)
            single.close()
            line["e2e_inprocess"] = inprocess_multi_gpu(args, dist, gio, binding, lib, args.workload, single_out)
            for mode in ("strong", "weak"):
                if not line["e2e_inprocess"][mode]["matches_single_gpu"]:
                    failed.append("e2e_inprocess." + mode)
        dist.host_barrier()
    if dist.rank == 0:
        if failed:
            line["failed"] = failed
        print(json.dumps(line), flush=True)
    dist.close()
    if failed:
        print("bench.py: parity / equality check FAILED for " + ", ".join(failed), file=sys.stderr)
        sys.exit(3)


if __name__ == "__main__":
    main()
