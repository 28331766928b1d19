#!/usr/bin/env bash
# A/B of tuning knobs on the device-resident leg (per-kernel-class ms per step): tools/ab_probe.sh <tag> "<ENV=..>" ...
TAG=$1; shift
for cfg in "$@"; do
  for wl in bert-base-f16-b1024xs128 bert-base-q4_0-b1024xs128; do
    echo "== $cfg $wl"
    env $cfg python bench.py --legs value --steps 5 --warmup 3 --workload $wl 2>&1 | tail -1
  done
done > gpurun_out/${TAG}_ab.log 2>&1
cat gpurun_out/${TAG}_ab.log
