#!/bin/bash
timeout 600 python tools/parity_reserve.py 2>&1 | grep -v Warn | tail -12
