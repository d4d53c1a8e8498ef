#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q -x --deselect tests/test_gpu_ddp.py > gpurun_out/c7_pytest.log 2>&1; echo "pytest rc=$?"
tail -5 gpurun_out/c7_pytest.log; grep -a "^E  " gpurun_out/c7_pytest.log | cut -c1-300 | head -20
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref > gpurun_out/c7_bench.json 2> gpurun_out/c7_bench.err; echo "bench rc=$?"; tail -3 gpurun_out/c7_bench.err
python tools/brief.py new < gpurun_out/c7_bench.json
NRD_FUSED_UPDATE=0 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py unfused-update
NRD_STEM_CTAS_PER_SM=1 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py stem-1-per-sm
