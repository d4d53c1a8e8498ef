#!/bin/bash
# round 2, call 1: baseline of the round-1 build + ncu --set full of every small kernel around the persistent ones
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/c1_smi.txt
python bench.py --steps 200 --warmup 10 > gpurun_out/c1_bench.json 2> gpurun_out/c1_bench.err; echo "bench rc=$?"
tail -c 1500 gpurun_out/c1_bench.json
TRACE_WAITS=1 python tools/trace_persist.py 256 > gpurun_out/c1_trace.txt 2>&1; echo "trace rc=$?"
CMD="python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline"
$CMD > gpurun_out/c1_plain.log 2>&1 && ncu --set full --clock-control none --import-source on \
  -k regex:'k_stem_tc|k_pool_fc1|k_head_tail|k_body_reduce|k_sgd_pack|k_head_wgrad|k_fc1_wgrad' -s 27 -c 9 \
  -o gpurun_out/c1_small $CMD > gpurun_out/c1_ncu.log 2>&1; echo "ncu rc=$?"
ls -la gpurun_out
