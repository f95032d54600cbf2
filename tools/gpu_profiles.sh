#!/bin/bash
# ncu evidence for the round: full captures of the decode GEMV and the attention kernel, the pp512 launch list, the other configurations
mkdir -p gpurun_out
timeout 120 python tools/prof_gemv_one.py 12 4096 22016 > gpurun_out/p1.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:mmvq3 -s 20 -c 2 -f -o gpurun_out/r02_prof_mmvq3_q4k python tools/prof_gemv_one.py 12 4096 22016 > gpurun_out/p1n.log 2>&1
timeout 120 python tools/bench_kernels.py fattn 1 32 32 128 4096 10 > gpurun_out/p2.log 2>&1 && \
timeout 300 ncu --set full --clock-control none --import-source on -k regex:fattn_mma -s 6 -c 2 -f -o gpurun_out/r02_prof_fattn_mma_f16 python tools/bench_kernels.py fattn 1 32 32 128 4096 10 > gpurun_out/p2n.log 2>&1
timeout 120 python tools/prof_pp512.py llama2-7b-q4_k_m 4 > gpurun_out/p3.log 2>&1 && \
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 3000 --csv --log-file gpurun_out/r02_launches_pp512_7b_4layers.csv python tools/prof_pp512.py llama2-7b-q4_k_m 4 > gpurun_out/p3n.log 2>&1
cat gpurun_out/p1.log gpurun_out/p2.log gpurun_out/p3.log
timeout 600 python tools/bench_configs.py c1 c3 c4 > gpurun_out/r02_configs_c1_c3_c4.jsonl 2> gpurun_out/configs.err; tail -c 2500 gpurun_out/r02_configs_c1_c3_c4.jsonl
