#!/bin/bash
timeout 1500 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python tools/tree_sweep.py --caps 16 18 20 22 24 26 --cpu > gpurun_out/tree_sweep.txt 2>&1; tail -37 gpurun_out/tree_sweep.txt | cut -c1-76
python tools/update_phases.py --batch 512 1024 4096 > gpurun_out/update_phases.txt 2>&1
