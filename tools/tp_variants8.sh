#!/bin/bash
N=8
run() { echo "== $*"; env "$@" B200_BENCH_ONLY_TP70=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29514 bench.py --gpus $N --steps 20 --warmup 3 2>/dev/null | grep tp_70b | python -c "import sys,json; d=json.loads(sys.stdin.read())['tp_70b_only']; print(d['value'], d['ms_per_step'], d['e2e']['value'], d['pp512_tokens_per_s'])"; }
run B200_AR_SMEM=150000
run B200_AR_SMEM=150000 B200_MMVQ_ONEROW_BELOW=16
