#!/bin/bash
# ncu evidence for one round: launch list of the bench's timed steps + one full capture of each hot kernel.
# usage (under gpurun): tools/ncu_round.sh r01
R=${1:-r01}
mkdir -p gpurun_out
ARGS="--steps 48 --warmup 5 --no-cpu-baseline --eager"
python bench.py $ARGS > gpurun_out/plain_$R.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -k regex:'k_sample|k_gather_tma|k_update_cluster' -s 30 -c 120 \
    --csv --log-file gpurun_out/launches_$R.csv python bench.py $ARGS > gpurun_out/ncu_list_$R.log 2>&1
python bench.py $ARGS > gpurun_out/plain2_$R.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:'k_sample|k_gather_tma|k_update_cluster' -s 60 -c 3 \
    -o gpurun_out/prof_$R python bench.py $ARGS > gpurun_out/ncu_full_$R.log 2>&1
python tools/aql_bench.py > gpurun_out/aql_plain_$R.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_aql_q -s 3 -c 1 -o gpurun_out/prof_aql_$R \
    python tools/aql_bench.py > gpurun_out/ncu_aql_$R.log 2>&1
tail -1 gpurun_out/plain_$R.log | cut -c1-200; tail -2 gpurun_out/ncu_list_$R.log | cut -c1-200; tail -2 gpurun_out/ncu_full_$R.log | cut -c1-200; tail -3 gpurun_out/aql_plain_$R.log
