#!/bin/bash
# scripts/gpu_ab_multi.sh "VARIANT ..." WORKLOAD...: the shipped library, each variants/libcanvas_b200.VARIANT.so, the shipped one again
variants=$1; shift
cp canvas_b200/libcanvas_b200.so /tmp/shipped.so
for v in shipped $variants shipped; do
  if [ $v = shipped ]; then cp /tmp/shipped.so canvas_b200/libcanvas_b200.so; else cp variants/libcanvas_b200.$v.so canvas_b200/libcanvas_b200.so; fi
  for w in "$@"; do
    python bench.py --workload $w --steps 8 --warmup 3 --no-cpu-baseline --no-others --no-plugin --no-scheduler 2>/dev/null | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('%-8s %-22s %9.0f fps  %.3f  ulp %s' % ('$v', d['config']['workload'], d['value'], d['roofline']['frac'], d.get('max_ulp')))
"
  done
done
cp /tmp/shipped.so canvas_b200/libcanvas_b200.so
