#!/bin/bash
python tools/stamps.py 256 2>&1 | head -8
