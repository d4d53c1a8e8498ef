#!/bin/bash
python -m pytest tests/test_gpu_tc.py -q -x -k "fused_stem_backward" 2>&1 | tail -2
python tools/trace_persist.py 256 2>&1 | grep -a "BWD CTA\|it=9\|last phase C" | head -8
NRD_FUSE_STEM_BWD=0 python tools/trace_persist.py 256 2>&1 | grep -a "last phase C" | tail -2
