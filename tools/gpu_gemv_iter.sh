#!/bin/bash
# Developer loop for the decode GEMV: parity tests of the kernels that call it, then the chain / sequence timings.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_fused.py tests/test_gpu_llama.py -m gpu -x -q > gpurun_out/gemv_tests.log 2>&1; echo "tests rc=$?"; tail -4 gpurun_out/gemv_tests.log
{
for pf in 1; do echo "== PREFILL=$pf"; B200_MMVQ_PREFILL=$pf python tools/gemv_seq.py 30; done
echo "== one-row kernel"; B200_MMVQ_NO_MMVQ3=1 python tools/gemv_seq.py 30
python tools/gemv_chain.py
} > gpurun_out/gemv_iter.log 2>&1
cat gpurun_out/gemv_iter.log
