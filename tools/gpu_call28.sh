#!/bin/bash
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --batch 2048 --no-graph --no-cpu-baseline --no-eager-ref"
$CMD > gpurun_out/s_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_tc_bwd_stream|k_tc_fwd_stream' -s 4 -c 2 -o gpurun_out/stream_full $CMD > gpurun_out/s_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out | grep stream
