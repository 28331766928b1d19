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
                 bounded sample of the same workload (rank 0, N=1 only)

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
}
DEFAULT_WORKLOAD = "bert-base-f16-b1024xs128"


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
        lens = np.sort(lens)[::-1].copy()                          # length-sorted, as bert_eval_batch orders them
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

    def barrier(self):
        if self.world > 1:
            if self.dev is not None:
                self.dist.barrier(device_ids=[self.local_rank])
                self.torch.cuda.synchronize()
            else:
                self.dist.barrier()

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


def model_file(gio, shape, ftype, dist: Dist, n_vocab=30522):
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


def kernel_model(hp, lens, mb_tokens):
    """algorithmic FLOPs / bytes per step for each kernel class."""
    T = int(lens.sum())
    d, f, Lr, n = hp.n_embd, hp.n_intermediate, hp.n_layer, len(lens)
    q = 1 if hp_is_q4(hp) else 0
    sq = float((lens.astype(np.float64) ** 2).sum())
    return {
        "gemm_qkv": {"flops": Lr * 2.0 * T * 3 * d * d},
        "gemm_o": {"flops": Lr * 2.0 * T * d * d},
        "gemm_ff1": {"flops": Lr * 2.0 * T * f * d},
        "gemm_ff2": {"flops": Lr * 2.0 * T * d * f},
        "attention": {"flops": Lr * 4.0 * sq * d, "bytes": Lr * (3 * d + d) * 2.0 * T},
        "add_ln": {"bytes": 2 * Lr * (2 + 2 + 2 * q) * d * 1.0 * T},
        "embed_ln": {"bytes": T * (4 + 4 + 2 * hp_row_bytes(hp) + (2 + 2 * q) * d)},
        "pool_norm": {"bytes": 2.0 * T * d + 4.0 * n * d},
        "pack": {"bytes": 4.0 * n + 4.0 * (n + 1) + 4.0 * T},
    }


def hp_is_q4(hp):
    return getattr(hp, "_ftype", "f16").startswith("q4")


def hp_row_bytes(hp):
    ft = getattr(hp, "_ftype", "f16")
    d = hp.n_embd
    return {"f32": 4 * d, "f16": 2 * d, "q4_0": d // 32 * 18, "q4_1": d // 32 * 20}[ft]


def time_reference(orc, path, batch, n_sent, warm, threads):
    """reference bert_eval (bert.cpp:1020), one sentence per call: it has no batching (bert.cpp:1064-1065)"""
    kind = "reference" if orc.available_reference() else "port"
    m = orc.Reference(path, n_threads=threads) if kind == "reference" else orc.Oracle(path, threads=threads)
    for i in range(warm):
        m.eval(batch[i % len(batch)])
    t0 = time.perf_counter()
    for i in range(n_sent):
        m.eval(batch[i % len(batch)])
    dt = time.perf_counter() - t0
    if kind == "reference":
        m.close()
    return kind, n_sent / dt, dt


def run_reference_arm(args, dist, gio, orc):
    """--impl reference: the reference's own CPU implementation on this box's host cores (rank 0 only)."""
    if dist.rank != 0:
        return
    shape, ftype, n_sent, seq_len = WORKLOADS[args.workload]
    path = model_file(gio, shape, ftype, _Solo(), n_vocab=args.vocab)
    hp = gio.read_gguf(path).hparams()
    batch, lens = make_batch(hp, n_sent, seq_len, seed=1234)
    threads = args.ref_threads or min(os.cpu_count() or 1, 32)
    per_step = args.ref_sentences                      # bounded sample of the workload per step
    kind = "reference" if orc.available_reference() else "port"
    m = orc.Reference(path, n_threads=threads) if kind == "reference" else orc.Oracle(path, threads=threads)
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
            "scaling": "weak", "vs_baseline": None, "dtype": "f32 accum / " + ftype + " weights (ggml CPU)", "data": "synthetic",
            "config": {"workload": args.workload, "shape": shape, "weights": ftype, "seq_len": seq_len or "16..512"},
            "cpu_baseline": {"value": value, "unit": "embeddings/s", "cores": threads, "kind": kind, "sample": sample},
            "e2e": {"value": value, "unit": "embeddings/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    print(json.dumps(line), flush=True)


class _Solo:
    local_rank, rank, world = 0, 0, 1

    def barrier(self):
        pass


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=os.environ.get("BENCH_WORKLOAD", DEFAULT_WORKLOAD), choices=sorted(WORKLOADS))
    ap.add_argument("--mb-tokens", type=int, default=int(os.environ.get("BERT_B200_MB_TOKENS", "65536")))
    ap.add_argument("--vocab", type=int, default=30522)
    ap.add_argument("--ref-sentences", type=int, default=2, help="reference arm: sentences per step")
    ap.add_argument("--cpu-baseline-sentences", type=int, default=12)
    ap.add_argument("--no-cpu-baseline", action="store_true")
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
    os.environ["BERT_B200_DEVICES"] = str(dist.local_rank)          # this rank's GPU only
    os.environ["BERT_B200_MB_TOKENS"] = str(args.mb_tokens)
    binding = _load("embedding_cpp_b200_binding", "embedding.cpp_b200/binding.py")
    lib = binding.load_library()                                    # raises if build/libbert.so is missing
    shape, ftype, n_sent, seq_len = WORKLOADS[args.workload]
    path = model_file(gio, shape, ftype, dist, n_vocab=args.vocab)
    hp = gio.read_gguf(path).hparams()
    hp._ftype = ftype
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
        if dist.rank == 0:
            print(json.dumps({"metric": "embeddings/sec", "value": value, "ms_per_step": ms / args.steps, "legs": "value",
                              "kernels_ms_per_step": {k: v[0] / args.steps for k, v in kms.items()}}), flush=True)
        dist.close()
        return

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
    latency = None
    if dist.rank == 0:
        one = np.ascontiguousarray(batch[0][:128] if seq_len == 0 else batch[0], np.int32)
        emb1 = np.zeros(hp.n_embd, np.float32)
        tp1 = one.ctypes.data_as(ctypes.POINTER(ctypes.c_int32))
        ep1 = emb1.ctypes.data_as(ctypes.POINTER(ctypes.c_float))
        for _ in range(20):
            lib.bert_eval(model.ctx, 1, tp1, len(one), ep1)
        ts = []
        for _ in range(args.latency_calls):
            t0 = time.perf_counter()
            lib.bert_eval(model.ctx, 1, tp1, len(one), ep1)      # host ids in, host embedding out, one sentence
            ts.append(time.perf_counter() - t0)
        ts = np.sort(np.array(ts)) * 1e3
        latency = {"p50_ms": float(ts[len(ts) // 2]), "p90_ms": float(ts[int(len(ts) * 0.9)]), "min_ms": float(ts[0]),
                   "calls": int(len(ts)), "tokens": int(len(one)),
                   "api": "bert_eval (one sentence, host buffers in/out; launch plan replayed as a CUDA graph)"}
    model.close()

    # ---------------- roofline of the dominant kernel -------------------------------------------------
    peaks = measured_peaks()
    km = kernel_model(hp, lens, args.mb_tokens)
    kernels = {}
    for name, (kms_tot, n_launch) in kms.items():
        if n_launch == 0:
            continue
        per_step_ms = kms_tot / args.steps
        ent = {"ms_per_step": per_step_ms, "launches_per_step": n_launch // args.steps, "share": kms_tot / max(1e-9, sum(v[0] for v in kms.values()))}
        m_ = km.get(name, {})
        if name.startswith("gemm"):
            ent.update(bound="tensor", achieved=m_["flops"] / (per_step_ms * 1e-3) / 1e12, peak=peaks["tflops_sustained"], unit="TFLOP/s")
        elif "bytes" in m_:
            ent.update(bound="hbm", achieved=m_["bytes"] / (per_step_ms * 1e-3) / 1e9, peak=peaks["hbm_gbs"], unit="GB/s")
        if "achieved" in ent:
            ent["frac"] = ent["achieved"] / ent["peak"]
        kernels[name] = ent
    if "add_ln" not in kernels:  # large micro-batches: the residual add and the LayerNorm run inside these two GEMMs (gemm_ln.cu)
        for name in ("gemm_o", "gemm_ff2"):
            if name in kernels:
                kernels[name]["fused"] = ("bias + residual + LayerNorm (+ Q8 copy) in the epilogue: replaces the add_ln kernel; "
                                          "achieved counts the GEMM FLOPs only")
    top = max(kernels, key=lambda k: kernels[k]["ms_per_step"])
    traffic = None
    tp_path = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tp_path):
        traffic = json.load(open(tp_path)).get(args.workload, {}).get(top)
    roofline = {"kernel": top, "bound": kernels[top].get("bound"), "achieved": kernels[top].get("achieved"),
                "peak": kernels[top].get("peak"), "unit": kernels[top].get("unit"), "frac": kernels[top].get("frac"),
                "traffic": traffic, "peak_source": peaks["src"] + " (sustained figure: kernel timed inside a long step)"}
    flops_step = float(sum(flops_per_sentence(hp, int(L)) for L in lens))
    model_tflops = flops_step * args.steps / (ms / 1e3) / 1e12          # per GPU (ms is the max over ranks)

    line = {
        "metric": "embeddings/sec", "value": value, "unit": "embeddings/s", "n_gpus": dist.world, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms / args.steps, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f16 (fp32 accumulate)", "data": "synthetic",
        "config": {"workload": args.workload, "shape": shape, "weights": ftype, "sentences_per_gpu": n_sent,
                   "global_batch": total_sent, "seq_len": seq_len or "16..512", "tokens_per_gpu": T,
                   "micro_batch_tokens": args.mb_tokens, "parallelism": f"dp{dist.world} (independent shards, host gather, no collective)",
                   "l2": "no flush: every kernel's working set per micro-batch (>= %d MB of activations) exceeds the 126 MB L2" % (args.mb_tokens * hp.n_embd * 2 * 5 // 2 ** 20)},
        "e2e": {"value": e2e_value, "unit": "embeddings/s", "h2d_bytes_per_step": int(4 * T + 4 * n_sent),
                "d2h_bytes_per_step": int(4 * n_sent * hp.n_embd), "api": "bert_eval_batch (build/libbert.so via ctypes), host pointers in/out",
                "matches_device_leg": same, "unit_norm": unit_norm},
        "latency_batch1": latency,
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roofline, "kernels": kernels,
        "model_tflops_per_gpu": model_tflops, "model_frac_of_peak": model_tflops / peaks["tflops_sustained"],
    }

    # ---------------- CPU baseline: the reference itself on this box's host cores (rank 0, N=1) --------
    if dist.rank == 0 and dist.world == 1 and not args.no_cpu_baseline:
        import oracle as orc                            # checker only: timed as the *baseline*, never shipped
        threads = args.ref_threads or min(os.cpu_count() or 1, 32)
        n_cpu = args.cpu_baseline_sentences
        kind, v, dt = time_reference(orc, path, batch, n_cpu, 1, threads)
        line["cpu_baseline"] = {"value": v, "unit": "embeddings/s", "cores": threads, "kind": kind,
                                "sample": f"{n_cpu} sentences of the same workload, one bert_eval call each (reference has no batching), {dt:.1f} s"}
        # parity spot-check of the bench outputs themselves against the CPU reference
        chk = orc.Reference(path, n_threads=threads) if kind == "reference" else orc.Oracle(path, threads=threads)
        ref0 = chk.eval(batch[0])
        line["parity_spot_check"] = {"cos": float(ref0 @ out[0] / (np.linalg.norm(ref0) * np.linalg.norm(out[0]))),
                                     "max_abs": float(np.abs(ref0 - out[0]).max())}
    if dist.rank == 0:
        print(json.dumps(line), flush=True)
    dist.close()


if __name__ == "__main__":
    main()
