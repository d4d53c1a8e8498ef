#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --deselect tests/test_gpu_ddp.py > gpurun_out/ev_pytest.log 2>&1; echo "pytest rc=$?"; tail -2 gpurun_out/ev_pytest.log
python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" 2>&1 | tail -1
