#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --deselect tests/test_gpu_ddp.py > gpurun_out/ev_pytest.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/ev_pytest.log
python bench.py --steps 200 --warmup 10 > gpurun_out/ev_bench.json 2> gpurun_out/ev_bench.err; echo "bench rc=$?"
python tools/brief.py final < gpurun_out/ev_bench.json
