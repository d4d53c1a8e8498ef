#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_tc.py -q -k "fused_bn_backward" 2>&1 | grep -v "^$" | tail -5
timeout 300 python bench.py --steps 50 --warmup 5 --batch 512 --chans 64 --blocks 20 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py C64
