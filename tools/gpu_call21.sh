#!/bin/bash
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py fused-stem-bwd
NRD_FUSE_STEM_BWD=0 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py unfused
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py fused-stem-bwd
