#!/bin/bash
# Developer loop for the attention kernels: parity tests, then the decode shapes timed alone.
mkdir -p gpurun_out
python -m pytest tests/test_gpu_kernels.py tests/test_gpu_llama.py tests/test_gpu_dropin.py -m gpu -x -q -k "flash or fattn or attn or llama or dropin" > gpurun_out/fattn_tests.log 2>&1; echo "tests rc=$?"; tail -6 gpurun_out/fattn_tests.log
{
for cfg in "1 32 32 128 4096" "1 64 8 128 4096" "1 32 8 128 32768" "8 32 8 128 32768" "1 8 1 128 4096" "8 32 32 128 4096"; do
  python tools/bench_kernels.py fattn $cfg 50
  B200_FA_NO_MMA=1 python tools/bench_kernels.py fattn $cfg 50
done
} > gpurun_out/fattn_iter.log 2>&1
cat gpurun_out/fattn_iter.log
