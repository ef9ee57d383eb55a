timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r2_t8.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/r2_t8.log
python tools/update_phases.py --batch 512 2>&1 | tail -1
for V in 8 16 8 16; do
PER_UPDATE_CTAS=$V timeout 600 python bench.py --steps 500 --warmup 5 --timeline --no-cpu-baseline --no-extras > gpurun_out/ab_c$V.json 2> gpurun_out/ab_c$V.err; echo "N=1 ctas $V rc=$? $(cut -c50-110 gpurun_out/ab_c$V.json)"; grep -E "rank 0 timeline" gpurun_out/ab_c$V.err | cut -c1-330
done
for V in 8 16 8 16; do
PER_UPDATE_CTAS=$V timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 500 --warmup 5 --timeline --no-parity-check --no-extras > gpurun_out/ab2_c$V.json 2> gpurun_out/ab2_c$V.err; echo "N=2 ctas $V rc=$? $(grep -o '"value": [0-9.]*' gpurun_out/ab2_c$V.json | head -1)"; grep -E "rank 0 timeline" gpurun_out/ab2_c$V.err | cut -c1-330
done
