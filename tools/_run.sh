for V in 3 7 3 7; do
PER_PDL=$V timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 500 --warmup 5 --timeline --no-parity-check > gpurun_out/ab_pdl$V.json 2> gpurun_out/ab_pdl$V.err; echo "pdl $V rc=$? $(cut -c1-120 gpurun_out/ab_pdl$V.json)"; grep -E "rank 0 timeline" gpurun_out/ab_pdl$V.err | cut -c1-520
done
for V in 3 7; do
PER_PDL=$V timeout 600 python bench.py --steps 500 --warmup 5 --timeline --no-cpu-baseline --no-extras > gpurun_out/ab1_pdl$V.json 2> gpurun_out/ab1_pdl$V.err; echo "N=1 pdl $V rc=$? $(cut -c1-120 gpurun_out/ab1_pdl$V.json)"; grep -E "rank 0 timeline" gpurun_out/ab1_pdl$V.err | cut -c1-520
done
