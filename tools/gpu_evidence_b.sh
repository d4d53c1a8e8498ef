#!/bin/bash
# second half of the final evidence (one profiler invocation per GPU call): ncu --set full of every kernel class of the step
mkdir -p gpurun_out
CMD="python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --no-eager-ref"
$CMD > gpurun_out/ev_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_tc_|k_stem_tc|k_head_fused|k_body_reduce|k_head_wgrad|k_fc1_wgrad|k_sgd_pack' -s 36 -c 12 -o gpurun_out/ev_full $CMD > gpurun_out/ev_ncu2.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | grep ev_full
