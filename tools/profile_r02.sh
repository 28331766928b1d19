#!/usr/bin/env bash
# ncu evidence of one build (run under gpurun, one GPU):  tools/profile_r02.sh <tag>
#   1. plain run of the profiled command (must exit 0)            -> gpurun_out/<tag>_plain_*.json
#   2. launch list (gpu__time_duration.sum per launch, F16 + Q4_0) -> gpurun_out/<tag>_launches_*.csv
#   3. --set full captures of the kernel classes of the step       -> gpurun_out/<tag>_full_*.ncu-rep
set -uo pipefail
TAG=${1:-r02}
OUT=gpurun_out
CMD_F16="python bench.py --legs value --steps 2 --warmup 3 --workload bert-base-f16-b1024xs128"
CMD_Q4="python bench.py --legs value --steps 2 --warmup 3 --workload bert-base-q4_0-b1024xs128"
CMD_VAR="python bench.py --legs value --steps 1 --warmup 3 --workload bert-base-q4_1-varlen"
$CMD_F16 > $OUT/${TAG}_plain_f16.json 2> $OUT/${TAG}_plain_f16.err || { echo "plain f16 run failed"; exit 1; }
$CMD_Q4 > $OUT/${TAG}_plain_q4_0.json 2> $OUT/${TAG}_plain_q4_0.err || { echo "plain q4 run failed"; exit 1; }
$CMD_VAR > $OUT/${TAG}_plain_varlen.json 2> $OUT/${TAG}_plain_varlen.err || { echo "plain varlen run failed"; exit 1; }
# launch lists: the timed region only (bench.py brackets it with cudaProfilerStart/Stop)
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $OUT/${TAG}_launches_bert-base-f16.csv $CMD_F16 > $OUT/${TAG}_ncu_f16.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off -c 400 --csv --log-file $OUT/${TAG}_launches_bert-base-q4_0.csv $CMD_Q4 > $OUT/${TAG}_ncu_q4.log 2>&1
# full captures: first layer of the first micro-batch of the timed region (one launch per kernel class; gemm_ln: K = 768 and K = 3072)
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'attention|gemm_ln|embed_ln|pool_norm|batch_' -c 8 -o $OUT/${TAG}_full_f16 -f $CMD_F16 > $OUT/${TAG}_ncu_full_f16.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'gemm_tcgen05' -c 2 -o $OUT/${TAG}_full_f16_gemm -f $CMD_F16 > $OUT/${TAG}_ncu_full_f16_gemm.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'gemm_ln|embed_ln|gemm_tcgen05' -c 5 -o $OUT/${TAG}_full_q4_0 -f $CMD_Q4 > $OUT/${TAG}_ncu_full_q4.log 2>&1
ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:'attention_pt|attention_stream' -c 1 -o $OUT/${TAG}_full_varlen -f $CMD_VAR > $OUT/${TAG}_ncu_full_varlen.log 2>&1
ls -la $OUT/${TAG}_*
