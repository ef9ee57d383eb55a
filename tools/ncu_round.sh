#!/bin/bash
# ncu evidence for one round, ONE ncu invocation per call (each after the same command has run plain):
#   tools/ncu_round.sh <tag> list   launch list of the bench's timed steps (gpu__time_duration per launch)
#   tools/ncu_round.sh <tag> full   --set full capture of one launch each of the sample / gather / update kernels
#   tools/ncu_round.sh <tag> aql    --set full capture of the AQL candidate-Q kernel
# usage (under gpurun): tools/ncu_round.sh r01d list
R=${1:-r01}; STAGE=${2:-list}
mkdir -p gpurun_out
ARGS="--steps 48 --warmup 5 --no-cpu-baseline --no-extras --eager"
KERNELS='k_sample|k_gather_bal|k_gather_ldg|k_gather_tma|k_update_cluster'
case $STAGE in
list)
  python bench.py $ARGS > gpurun_out/plain_$R.log 2>&1 &&
  ncu --metrics gpu__time_duration.sum --clock-control none -k regex:"$KERNELS" -s 30 -c 120 \
      --csv --log-file gpurun_out/launches_$R.csv python bench.py $ARGS > gpurun_out/ncu_list_$R.log 2>&1
  tail -1 gpurun_out/plain_$R.log | cut -c1-200; tail -2 gpurun_out/ncu_list_$R.log | cut -c1-200 ;;
full)
  python bench.py $ARGS > gpurun_out/plain2_$R.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:"$KERNELS" -s 60 -c 3 \
      -o gpurun_out/prof_$R python bench.py $ARGS > gpurun_out/ncu_full_$R.log 2>&1
  tail -2 gpurun_out/ncu_full_$R.log | cut -c1-200 ;;
aql)
  python tools/aql_bench.py > gpurun_out/aql_plain_$R.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_aql_q -s 3 -c 1 -o gpurun_out/prof_aql_$R \
      python tools/aql_bench.py > gpurun_out/ncu_aql_$R.log 2>&1
  tail -3 gpurun_out/aql_plain_$R.log; tail -2 gpurun_out/ncu_aql_$R.log | cut -c1-200 ;;
esac
