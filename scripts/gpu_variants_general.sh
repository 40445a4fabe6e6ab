#!/bin/bash
# window-classifying (GENERAL) kernels: the 1080p windows workload and the partial-window timeline of the scheduler bench
cp canvas_b200/libcanvas_b200.so /tmp/shipped.so
for v in shipped "$@"; do
  if [ $v = shipped ]; then cp /tmp/shipped.so canvas_b200/libcanvas_b200.so; else cp variants/libcanvas_b200.$v.so canvas_b200/libcanvas_b200.so; fi
  python bench.py --workload 1080p_4layer_windows --steps 10 --no-cpu-baseline 2>&1 | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$v', d['config']['workload'], round(d['value'],1), 'fps frac', round(d['roofline']['frac'],3))
"
  python scripts/bench_scheduler.py --gpus 1 2>&1 | tail -1 | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$v scheduler 4k timeline (windows)', round(d['value'],1), 'fps', round(d['algorithmic_gb_per_s']), 'GB/s')
"
done
cp /tmp/shipped.so canvas_b200/libcanvas_b200.so
