#!/bin/bash
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 20 --warmup 5 --no-eager-ref > gpurun_out/c6_bench2.json 2> gpurun_out/c6_bench2.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/c6_bench2.json').read().strip().splitlines()[-1])
print(json.dumps(d['parity'],indent=0))
PY
