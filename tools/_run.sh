#!/bin/bash
for t in 32 64 128 256; do echo "threads $t"; PER_WIDE_THREADS=$t python tools/tree_sweep.py --caps 20 --batches 4096 16384 65536 2>&1 | tail -3 | cut -c1-36; done
