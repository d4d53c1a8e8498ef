#!/bin/bash
for R in 1 2 4; do
NRD_RESERVE_SMS=$R python bench.py --steps 300 --warmup 10 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py B256-reserve$R | head -1 | cut -c1-80
done
for R in 0 2 4 8; do
NRD_RESERVE_SMS=$R python bench.py --steps 100 --warmup 10 --batch 2048 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py B2048-reserve$R | head -1 | cut -c1-80
done
for R in 0 4 8; do
NRD_RESERVE_SMS=$R python bench.py --steps 300 --warmup 10 --batch 64 --no-cpu-baseline --no-eager-ref 2>/dev/null | python tools/brief.py B64-reserve$R | head -1 | cut -c1-80
done
