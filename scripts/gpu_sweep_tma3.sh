#!/bin/bash
# k_stack_tma CTA size (warps are autonomous; shared-memory packing decides how many stay resident), default workload:
# the shipped library against variants/libcanvas_b200.tmaauto.so (CV_TMA_THREADS: 0 = pick the best packing)
cp canvas_b200/libcanvas_b200.so /tmp/shipped.so
run() {
  CV_TMA_THREADS=$2 python bench.py --steps 100 --no-cpu-baseline 2>/dev/null | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$1 threads $2', round(d['value'],1), 'fps frac', round(d['roofline']['frac'],4), d['clocks']['sm_mhz'], d['clocks']['reasons'])
"
}
run shipped 128
cp variants/libcanvas_b200.tmaauto.so canvas_b200/libcanvas_b200.so
python -m pytest tests -m gpu -x -q 2>&1 | tail -1
for t in 0 128 224 96 32 0; do run tmaauto $t; done
cp /tmp/shipped.so canvas_b200/libcanvas_b200.so
run shipped 128
