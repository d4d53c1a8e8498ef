#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_ddp.py > gpurun_out/c12_pytest.log 2>&1; echo "pytest rc=$?"
tail -3 gpurun_out/c12_pytest.log; grep -a "^E  " gpurun_out/c12_pytest.log | cut -c1-300 | head -20
python tools/stamps.py 256 2>&1 | tail -16
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py new
