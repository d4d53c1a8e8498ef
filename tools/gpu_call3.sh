#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_gpu_boundary.py tests/test_gpu_parity_bench.py -q -s > gpurun_out/c3_pytest.log 2>&1; echo "pytest rc=$?"
grep -a "parity B=\|^FAILED\|passed\|failed" gpurun_out/c3_pytest.log
grep -a "^E  " gpurun_out/c3_pytest.log | cut -c1-400 | head -40
