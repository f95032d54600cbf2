#!/bin/bash
# Developer diagnosis of the decode GEMV fixed cost: hot/cold chains under the debug knobs, then per-CTA timelines.
mkdir -p gpurun_out
for dbg in 0 1 2 3; do
  echo "=== B200_MMVQ_DBG=$dbg (1: no activation phase, 2: no dot)"; B200_MMVQ_DBG=$dbg python tools/gemv_chain.py
done > gpurun_out/diag_chain.log 2>&1
for s in "12 4096 4096" "12 4096 22016" "12 11008 4096" "14 11008 4096"; do
  python tools/trace_mmvq.py $s; python tools/trace_mmvq.py $s hot
done > gpurun_out/diag_trace.log 2>&1
tail -40 gpurun_out/diag_chain.log
