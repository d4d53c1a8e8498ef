#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29721 bench.py --gpus 2 --steps 200 --warmup 10 > gpurun_out/r02_bench_n2_final.json 2> gpurun_out/r02_bench_n2_final.err; echo "rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n2_final.json').read().strip().splitlines()[-1])
print("N=2", round(d['value']), 'img/s', round(d['ms_per_step']*1e3,1), 'us/step e2e', round(d['e2e']['value']), 'parity', d['parity']['ddp_vs_oracle'], d['parity']['ranks_identical'], 'eager', d['gpu_eager_reference'].get('ms_per_step'))
PY
