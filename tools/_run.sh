set -x
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_t1.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/r2_t1.log
timeout 600 python tools/gather_sweep.py > gpurun_out/r2_sweep1.log 2>&1; echo "sweep rc=$?"; tail -3 gpurun_out/r2_sweep1.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r2_bench1.json 2> gpurun_out/r2_bench1.err; echo "bench rc=$?"; cut -c1-1500 gpurun_out/r2_bench1.json; tail -5 gpurun_out/r2_bench1.err
