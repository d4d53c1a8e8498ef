#!/bin/bash
python -m pytest tests/test_gpu_step.py -q -x -k dropin 2>&1 | grep -a "^E  \|Error\|error" | head -20
