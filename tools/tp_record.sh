#!/bin/bash
# The 70B tensor-parallel leg alone at N ranks (developer shortcut of bench.py): one JSON line
N=${1:-8}
B200_BENCH_ONLY_TP70=1 timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29515 bench.py --gpus $N --steps 32 --warmup 3 2>/dev/null | grep tp_70b
