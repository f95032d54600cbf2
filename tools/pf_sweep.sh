#!/bin/bash
# developer tool: 7B decode step time as a function of the cross-kernel L2 prefetch budget (B200_PF_MB, MB per issuing kernel)
mkdir -p gpurun_out
for mb in ${PF_LIST:-0 2 4 8 16 24 40}; do
  B200_PF_MB=$mb python bench.py --steps ${STEPS:-40} --warmup 3 --no-cpu --no-tp70 2>/dev/null | python -c "
import sys, json
j = json.loads(sys.stdin.read().strip().splitlines()[-1])
print('PF_MB=$mb', 'tok/s %.1f' % j['value'], 'ms %.3f' % j['ms_per_step'], 'e2e %.1f' % j['e2e']['value'], 'gemv GB/s %.0f' % j['roofline']['achieved'], 'pp512 %.0f' % j['prefill']['value'])
"
done
