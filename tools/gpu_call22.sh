#!/bin/bash
# streaming kernels: parity first (own short timeout: a hang must not take the box), then the timelines
timeout 300 python -m pytest tests/test_gpu_tc.py -x -q -s -k "streaming" 2>&1 | grep -v "^$" | tail -25
for B in 512 1024 2048; do
  NRD_TIMELINE=1 timeout 120 python tools/timeline.py $B 2>&1 | grep -E "step|fwd_blocks|bwd_blocks|sum of"
done
NRD_STREAM=1 NRD_TIMELINE=1 timeout 120 python tools/timeline.py 256 2>&1 | grep -E "step|fwd_blocks|bwd_blocks|sum of"
NRD_TIMELINE=1 timeout 120 python tools/timeline.py 256 2>&1 | grep -E "step|fwd_blocks|bwd_blocks|sum of"
