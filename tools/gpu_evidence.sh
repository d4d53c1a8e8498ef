#!/bin/bash
# final evidence of a build: full GPU suite, bench line, ncu launch list + ncu --set full of every kernel of the step
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --deselect tests/test_gpu_ddp.py > gpurun_out/ev_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/ev_pytest.log
python bench.py --steps 200 --warmup 10 > gpurun_out/ev_bench.json 2> gpurun_out/ev_bench.err; echo "bench rc=$?"
python tools/brief.py final < gpurun_out/ev_bench.json
python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/ev_bench_ref.json 2>/dev/null; tail -c 600 gpurun_out/ev_bench_ref.json
CMD="python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --no-eager-ref"
$CMD > gpurun_out/ev_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 24 --csv --log-file gpurun_out/ev_launches.csv $CMD > gpurun_out/ev_ncu1.log 2>&1; echo "ncu launches rc=$?"
$CMD > gpurun_out/ev_plain2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:'k_tc_|k_stem_tc|k_head_fused|k_body_reduce|k_head_wgrad|k_fc1_wgrad|k_sgd_pack' -s 36 -c 12 -o gpurun_out/ev_full $CMD > gpurun_out/ev_ncu2.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | grep ev_
