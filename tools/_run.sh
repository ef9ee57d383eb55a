#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_sorted_update.py -x -q -m gpu 2>&1 | tail -8
