#!/usr/bin/env bash
# A/B of compile-time variants of the library (make BUILD=<dir> DEFS=...): tools/ab_lib.sh <rounds> <steps> <dir> [<dir> ...]
R=$1; K=$2; shift 2
for r in $(seq 1 $R); do
  for lib in "$@"; do
    BERT_B200_LIB=$PWD/$lib/libbert.so python bench.py --legs value --steps $K --warmup 3 2>/dev/null | tail -1 | python -c "
import sys, json
j = json.loads(sys.stdin.read()); k = j['kernels_ms_per_step']
print('$lib', round(j['value']), 'ms %.2f' % j['ms_per_step'], {a: round(b, 2) for a, b in k.items() if b > 1})"
  done
done
