set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_t2.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/r2_t2.log
timeout 600 python bench.py --steps 20 --warmup 5 --no-cpu-baseline --timeline > gpurun_out/r2_bench2.json 2> gpurun_out/r2_bench2.err; echo "bench rc=$?"; cut -c1-300 gpurun_out/r2_bench2.json; tail -3 gpurun_out/r2_bench2.err | cut -c1-900
