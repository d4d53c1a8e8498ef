#!/bin/bash
# round 2, call 2: the full GPU suite with the new parity / boundary tests (1 GPU)
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -s --deselect tests/test_gpu_ddp.py > gpurun_out/c2_pytest.log 2>&1; echo "pytest rc=$?"
tail -30 gpurun_out/c2_pytest.log
grep -a "parity B=" gpurun_out/c2_pytest.log
