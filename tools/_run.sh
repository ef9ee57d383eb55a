timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 20 --warmup 5 > gpurun_out/r2_bench_n2_p2p.json 2> gpurun_out/r2_bench_n2_p2p.err; echo "rc=$?"; tail -3 gpurun_out/r2_bench_n2_p2p.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_n2_p2p.json') if l.startswith('{')][0])
print(d['value'], d['ms_per_step'], d['roofline']['frac'], d['roofline']['avg_launch_us'], d['e2e']['value'])
print(d['p2p'])
PY
