#!/bin/bash
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus 8 --steps 20 --warmup 5 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; grep -o '"value": [0-9.]*' gpurun_out/bench_n8.json | head -2; grep -o '"period": [0-9.]*' gpurun_out/bench_n8.json; grep -c "parity ok" gpurun_out/bench_n8.err
