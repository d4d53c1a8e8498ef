#!/bin/bash
mkdir -p gpurun_out
python tools/timeline.py 256 > gpurun_out/c8_timeline.txt 2>&1; cat gpurun_out/c8_timeline.txt
NO_FLUSH=1 python tools/timeline.py 256 2>&1 | tail -12
