#!/bin/bash
# Developer A/B of the 7B decode step under kernel switches (1 GPU)
run() { echo "== $*"; env "$@" python bench.py --steps 30 --warmup 5 --no-cpu --no-llamacpp --no-tp70 2>/dev/null | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['achieved'], d['prefill']['value'])"; }
run X=1
run B200_MMVQ_DBG=4
run B200_MMVQ_PREFILL=0
run B200_MMVQ_PREFILL=2
run B200_MMVQ_NO_MMVQ3=1
run B200_FA_NO_CLUSTER=1
run B200_MMQ_NO_SPLITK=1
