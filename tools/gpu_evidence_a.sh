#!/bin/bash
# first half of the final evidence: full GPU suite, bench line (ours + reference arm), ncu launch list of the step
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --deselect tests/test_gpu_ddp.py > gpurun_out/ev_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/ev_pytest.log
python bench.py --steps 200 --warmup 10 > gpurun_out/ev_bench.json 2> gpurun_out/ev_bench.err; echo "bench rc=$?"
python tools/brief.py final < gpurun_out/ev_bench.json
python bench.py --impl reference --steps 10 --warmup 2 > gpurun_out/ev_bench_ref.json 2>/dev/null; tail -c 600 gpurun_out/ev_bench_ref.json
CMD="python bench.py --steps 3 --warmup 3 --no-graph --no-cpu-baseline --no-eager-ref"
$CMD > gpurun_out/ev_plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 60 -c 24 --csv --log-file gpurun_out/ev_launches.csv $CMD > gpurun_out/ev_ncu1.log 2>&1; echo "ncu launches rc=$?"
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -2
