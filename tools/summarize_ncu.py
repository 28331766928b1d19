"""Summaries of one tools/profile_r02.sh capture set (run where ncu is installed, no GPU needed):
    python tools/summarize_ncu.py <tag> <commit>
  gpurun_out/<tag>_full_*.ncu-rep      -> profiles/<tag>_ncu_full_summary.csv   (one row per captured launch, the judged metrics)
  gpurun_out/<tag>_launches_*.csv      -> profiles/<tag>_launches_*.csv + profiles/<tag>_launches_summary.txt
  and profiles/traffic.json (DRAM bytes per launch per kernel class, what bench.py reports as roofline.traffic)."""
import csv
import io
import json
import os
import shutil
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
WANT = ['Kernel Name', 'Grid Size', 'Block Size', 'gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum',
        'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed',
        'lts__throughput.avg.pct_of_peak_sustained_elapsed', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread', 'sm__inst_executed.avg.per_cycle_active',
        'sm__cycles_elapsed.avg.per_second', 'launch__shared_mem_per_block_dynamic', 'lts__t_sector_hit_rate.pct']


def raw(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    return rows[0], rows[1], rows[2:]


def main(tag, commit):
    go, pr = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
    out, units = [], None
    for cap in ("full_f16", "full_f16_gemm", "full_q4_0", "full_varlen"):
        rep = os.path.join(go, f"{tag}_{cap}.ncu-rep")
        if not os.path.exists(rep):
            continue
        hdr, u, rows = raw(rep)
        idx = [hdr.index(w) for w in WANT]
        units = units or [f"{w} [{u[i]}]" if u[i] else w for w, i in zip(WANT, idx)]
        out += [[f"{tag}_{cap}"] + [r[i] for i in idx] for r in rows]
    with open(os.path.join(pr, f"{tag}_ncu_full_summary.csv"), "w", newline="") as f:
        w = csv.writer(f)
        w.writerow(["capture"] + units)
        w.writerows(out)
    mb = lambda r: (float(r[5]) + float(r[6])) * 1e6
    pick = lambda cap, pat: [r for r in out if r[0].endswith(cap) and pat in r[1]]
    src = (f"profiles/{tag}_ncu_full_summary.csv (ncu --set full --clock-control none of `python bench.py --legs value --steps 2 --warmup 3 "
           f"--workload %s`, tools/profile_r02.sh; dram__bytes_read.sum + dram__bytes_write.sum per launch, 65536-token micro-batch, first layer; "
           f"kernels of commit {commit})")
    t = {}
    for wl, cap in (("bert-base-f16-b1024xs128", "full_f16"), ("bert-base-q4_0-b1024xs128", "full_q4_0")):
        d = {"_source": src % wl}
        ln = pick(cap, "gemm_ln")
        gm = pick("full_f16_gemm" if cap == "full_f16" else cap, "gemm_tcgen05")
        if len(ln) >= 2:
            d["gemm_o"], d["gemm_ff2"] = mb(ln[0]), mb(ln[1])
        if len(gm) >= 2:
            d["gemm_qkv"], d["gemm_ff1"] = mb(gm[0]), mb(gm[1])
        for k, pat in (("embed_ln", "embed_ln"), ("attention", "attention"), ("pool_norm", "pool_norm")):
            rows = pick(cap, pat) or pick("full_f16", pat)
            if rows:
                d[k] = mb(rows[0])
        t[wl] = d
    json.dump(t, open(os.path.join(pr, "traffic.json"), "w"), indent=1)
    with open(os.path.join(pr, f"{tag}_launches_summary.txt"), "w") as f:
        for wl in ("bert-base-f16", "bert-base-q4_0"):
            p = os.path.join(go, f"{tag}_launches_{wl}.csv")
            if os.path.exists(p):
                shutil.copy(p, pr)
                f.write(subprocess.run([sys.executable, os.path.join(ROOT, "tools", "summarize_launches.py"), p], capture_output=True, text=True).stdout + "\n")
    print(open(os.path.join(pr, f"{tag}_launches_summary.txt")).read())


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
