#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_tc.py tests/test_gpu_parity_bench.py -q -k "streaming or 500" 2>&1 | grep -v "^$" | tail -8
for B in 512 1024 2048 4096; do
  timeout 300 python bench.py --steps 100 --warmup 10 --batch $B --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py B$B
done
