#!/bin/bash
python tools/stamps.py 256 2>&1 | tail -14
