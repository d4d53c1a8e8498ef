#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_parity_bench.py tests/test_gpu_tc.py -q -s -k "benchmarked or fused_stem_backward" 2>&1 | grep -E "passed|failed|FAILED|h1_relu_flips" | sed -e 's/.*\(.h1_relu_flips.: [0-9]*\).*/\1/' | tail -12
