#!/bin/bash
# k_stack_tma tunables on the default workload (4K 8-layer composite)
for cfg in "2 128" "1 128" "4 128" "2 64" "2 96" "1 64" "1 96"; do
  set -- $cfg
  CV_TMA_PPL=$1 CV_TMA_THREADS=$2 python bench.py --steps 20 --no-cpu-baseline 2>&1 | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('ppl $1 threads $2', round(d['value'],1), 'fps frac', round(d['roofline']['frac'],4))
"
done
