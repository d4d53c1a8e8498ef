#!/bin/bash
NRD_TIMELINE=1 timeout 120 python tools/timeline.py 256 2>&1 | tail -10
NRD_RESERVE_SMS=0 NRD_TIMELINE=1 timeout 120 python tools/timeline.py 256 2>&1 | tail -10
