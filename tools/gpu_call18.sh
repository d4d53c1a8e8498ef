#!/bin/bash
python -m pytest tests/test_gpu_data.py tests/test_gpu_step.py -q 2>&1 | tail -4
python -m pytest tests/test_gpu_data.py tests/test_gpu_step.py -q 2>&1 | grep -a "^E  " | cut -c1-300 | head -20
