#!/bin/bash
python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py new
python tools/stamps.py 256 2>&1 | grep -A2 "stem fwd"
NRD_FUSE_STEM=1 python tools/trace_persist.py 256 2>&1 | grep -a "fused stem\|kernel entry" | head -5
NRD_FUSE_STEM=1 python bench.py --steps 200 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py fuse-stem
