#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_ddp.py -q -s -v > gpurun_out/r02_ddp_8gpu_pytest.log 2>&1; echo "pytest rc=$?"
grep -a "PASSED\|FAILED\|SKIPPED\|passed\|failed\|^E  " gpurun_out/r02_ddp_8gpu_pytest.log | cut -c1-300 | head -30
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29655 tools/stamps_ddp.py 2>&1 | grep -v "^\*\*\*\|OMP" | tail -10 | tee gpurun_out/r02_ddp_tail_n8.txt
for N in 8 4; do
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2961$N bench.py --gpus $N --steps 100 --warmup 10 > gpurun_out/r02_bench_n$N.json 2> gpurun_out/r02_bench_n$N.err; echo "bench N=$N rc=$?"
python - <<PY
import json
d=json.loads(open('gpurun_out/r02_bench_n$N.json').read().strip().splitlines()[-1])
print("N=$N", round(d['value']), 'img/s', round(d['ms_per_step']*1e3,1), 'us/step e2e', round(d['e2e']['value']), d['e2e']['warmup_steps'], 'parity', d['parity']['ddp_vs_oracle'], d['parity']['ranks_identical'], d['parity']['bf16']['per_step'][0], 'eager', d['gpu_eager_reference'].get('ms_per_step'))
PY
done
