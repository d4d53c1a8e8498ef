#!/bin/bash
python -m pytest tests/test_gpu_tc.py -q -x -k "fused_stem_backward" 2>&1 | tail -3
python -m pytest tests/test_gpu_tc.py -q -x -k "fused_stem_backward" 2>&1 | grep -a "^E  " | cut -c1-200 | head
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py fused-stem-bwd
NRD_FUSE_STEM_BWD=0 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py unfused
