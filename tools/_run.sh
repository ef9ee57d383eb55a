#!/bin/bash
python tools/tree_sweep.py --caps 20 --batches 65536 2>&1 | tail -1
ncu --set full --clock-control none --import-source on -k regex:k_upd_dense_sorted -s 6 -c 2 -o gpurun_out/dense_upd_after -f python tools/tree_sweep.py --caps 20 --batches 65536 > gpurun_out/ncu_dense_after.log 2>&1
tail -2 gpurun_out/ncu_dense_after.log
