timeout 600 python -m pytest tests/test_gpu_sharded.py tests/test_gpu_multi.py -x -q 2>&1 | tail -2
PER_GATHER_PROBE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 300 --warmup 5 --no-extras --no-parity-check --timeline 2>&1 >gpurun_out/probe.json | grep -E "gather probe|rank 0 timeline|Error|error" | cut -c1-420
grep -o '"value": [0-9.]*' gpurun_out/probe.json | head -1
