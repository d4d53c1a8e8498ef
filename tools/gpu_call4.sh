#!/bin/bash
# 2 GPUs: DDP parity tests (W=2) and the bench parity block
mkdir -p gpurun_out
python -m pytest tests/test_gpu_ddp.py -q -s > gpurun_out/c4_ddp.log 2>&1; echo "pytest rc=$?"
grep -a "passed\|failed\|skipped\|^E  \|Error" gpurun_out/c4_ddp.log | cut -c1-300 | head -30
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611 bench.py --gpus 2 --steps 50 --warmup 5 > gpurun_out/c4_bench2.json 2> gpurun_out/c4_bench2.err; echo "bench rc=$?"
tail -c 3000 gpurun_out/c4_bench2.json; tail -5 gpurun_out/c4_bench2.err
