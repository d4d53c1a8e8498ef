#!/bin/bash
python -m pytest tests/test_gpu_step.py tests/test_gpu_boundary.py -q -x 2>&1 | tail -3
python tools/e2e_probe.py 2>&1 | tail -12
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py mirror
