#!/bin/bash
timeout 900 python -m pytest tests -x -q -m gpu 2>&1 | tail -3
python tools/tree_sweep.py --caps 16 20 21 22 --batches 512 4096 16384 65536 2>&1 | tail -17
python tools/update_phases.py --batch 2048 4096 2>&1 | tail -2
python bench.py --steps 500 --warmup 5 --timeline --no-cpu-baseline --no-extras 2>&1 >/dev/null | grep -E "rank 0 timeline" | cut -c1-800
