#!/bin/bash
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_sorted_update.py tests/test_gpu_sharded.py tests/test_gpu_dropin.py -x -q -m gpu 2>&1 | tail -2
python tools/update_phases.py --batch 512 1024 2>&1 | tail -2 | cut -c1-200
