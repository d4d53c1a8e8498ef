#!/bin/bash
for B in 512 2048; do NRD_TIMELINE=1 timeout 120 python tools/timeline.py $B 2>&1 | grep -E "step|fwd_blocks|bwd_blocks"; done
timeout 600 python -m pytest tests/test_gpu_tc.py -q -k "streaming" 2>&1 | grep -v "^$" | tail -3
