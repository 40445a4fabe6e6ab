#!/bin/bash
# k_stack_tma run-time tunables (pairs per lane, stages, threads) on the default workload, 100-step runs on one box
for cfg in "2 2 128" "2 3 128" "2 4 128" "1 2 128" "1 4 128" "2 2 64" "2 3 64" "4 2 128" "2 2 128"; do
  set -- $cfg
  CV_TMA_PPL=$1 CV_TMA_STAGES=$2 CV_TMA_THREADS=$3 python bench.py --steps 100 --no-cpu-baseline 2>/dev/null | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('ppl $1 stages $2 threads $3', round(d['value'],1), 'fps frac', round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])
"
done
