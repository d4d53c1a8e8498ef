#!/bin/bash
timeout 600 python -m pytest tests/test_gpu_tc.py -q -k "streaming" 2>&1 | grep -v "^$" | tail -25
