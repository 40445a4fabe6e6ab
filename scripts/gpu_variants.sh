#!/bin/bash
# runs a few workloads with each tuning variant of the library (scripts/build_variant.sh)
cp canvas_b200/libcanvas_b200.so /tmp/shipped.so
for v in shipped "$@"; do
  if [ $v = shipped ]; then cp /tmp/shipped.so canvas_b200/libcanvas_b200.so; else cp variants/libcanvas_b200.$v.so canvas_b200/libcanvas_b200.so; fi
  for w in ${WORKLOADS:-4k_timeline 4k_composite_convert 1080p_cross_u8 1080p_4layer_windows}; do
    python bench.py --workload $w --steps 10 --no-cpu-baseline 2>&1 | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$v', d['config']['workload'], round(d['value'],1), 'fps frac', round(d['roofline']['frac'],3))
"
  done
done
cp /tmp/shipped.so canvas_b200/libcanvas_b200.so
