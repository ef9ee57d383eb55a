#!/bin/bash
# usage under gpurun --gpus N: tools/run_scale.sh N tag [extra bench args]
N=$1; TAG=$2; shift 2
mkdir -p gpurun_out
(timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 2000 --warmup 20 "$@" 2>&1 | tail -8) > gpurun_out/bench_$TAG.log
python - <<PY
import json
for l in open("gpurun_out/bench_$TAG.log"):
    if l.startswith("{"):
        d = json.loads(l)
        print("N=%d value=%.3g eager=%s us/step=%.1f gather_frac=%.3f e2e=%.3g launch=%s" % (d["n_gpus"], d["value"], d["value_eager_launches"], 1e3 * d["ms_per_step"], d["roofline"]["frac"], d["e2e"]["value"], d["launch"][:30]), "exchange_us", d.get("record_exchange_us"))
    elif l.strip() and not l.startswith("*") and "OMP_NUM" not in l:
        print(l[:300].rstrip())
PY
