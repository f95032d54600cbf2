#!/bin/bash
mkdir -p gpurun_out
{
for pf in 0 1 2 3 4; do echo "== PREFILL=$pf (row segments per warp before griddepcontrol.wait)"; B200_MMVQ_PREFILL=$pf python tools/gemv_seq.py 30; done
for pf in 0 1 2; do echo "== PREFILL=$pf"; B200_MMVQ_PREFILL=$pf python tools/gemv_chain.py; done
} > gpurun_out/gemv_iter2.log 2>&1
cat gpurun_out/gemv_iter2.log
