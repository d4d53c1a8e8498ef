#!/bin/bash
timeout 900 python -m pytest tests/test_gpu_ddp.py -x -q 2>&1 | grep -v "^$" | tail -8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29711 bench.py --gpus 2 --steps 100 --warmup 10 --batch 1024 --no-eager-ref 2>/dev/null | python tools/brief.py N2-B1024
