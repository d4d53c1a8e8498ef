#!/bin/bash
mkdir -p gpurun_out
NRD_RESERVE_SMS=8 python -m pytest tests/test_gpu_parity_bench.py -q -s -k "256" 2>&1 | grep -a "parity B=\|passed\|failed" | cut -c1-1500
