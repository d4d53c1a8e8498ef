#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_ops.py tests/test_gpu_step.py -q -x 2>&1 | tail -2
python tools/stamps.py 256 2>&1 | grep -A2 "stem bwd"
python bench.py --steps 100 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py base
for B in 64 128 512 1024 2048; do
python bench.py --steps 50 --warmup 5 --batch $B --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py "C32 B=$B"
done
python bench.py --steps 30 --warmup 5 --chans 64 --blocks 20 --batch 512 --no-cpu-baseline --no-eager-ref 2>gpurun_out/c14_wide.err | python tools/brief.py "wide C64 NB20 B512"; tail -3 gpurun_out/c14_wide.err
