for K in 20 200; do
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 8 --steps $K --warmup 5 --timeline > gpurun_out/r2_bench_n8_k$K.json 2> gpurun_out/r2_bench_n8_k$K.err; echo "bench rc=$?"; cut -c1-220 gpurun_out/r2_bench_n8_k$K.json; grep -E "parity ok on rank 0|rank 0 timeline|rank 7 timeline" gpurun_out/r2_bench_n8_k$K.err | cut -c1-800
done
