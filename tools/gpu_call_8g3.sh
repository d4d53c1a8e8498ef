#!/bin/bash
mkdir -p gpurun_out
for B in 2048 64; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 2971$((B % 10)) bench.py --gpus 8 --steps 100 --warmup 10 --batch $B --no-eager-ref --no-cpu-baseline > gpurun_out/r02_bench_n8_b$B.json 2> gpurun_out/r02_bench_n8_b$B.err; echo "bench N=8 B=$B rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n8_b$B.json').read().strip().splitlines()[-1])
print("N=8 B=$B", round(d['value']), 'img/s', round(d['ms_per_step']*1e3,1), 'us/step e2e', round(d['e2e']['value']), 'parity', (d.get('parity') or {}).get('ddp_vs_oracle'), (d.get('parity') or {}).get('ranks_identical'), 'step_roofline', (d.get('step_roofline') or {}).get('frac'))
PY
done
