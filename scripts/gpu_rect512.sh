#!/bin/bash
# k_rect at 512 threads (16 warps, 128 registers, small spills) against the shipped 384-thread build, per PPL
cd "$(dirname "$0")/.."
cp canvas_b200/libcanvas_b200.so /tmp/shipped.so
run() { env $3 python bench.py --workload $2 --steps 8 --warmup 3 --no-cpu-baseline --no-others --no-plugin --no-scheduler 2>/dev/null | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('%-8s %-14s %-22s %9.0f fps  %.3f  ulp %s' % ('$1', '$3', d['config']['workload'], d['value'], d['roofline']['frac'], d.get('max_ulp')))
"; }
for v in shipped r512 shipped; do
  if [ $v = shipped ]; then cp /tmp/shipped.so canvas_b200/libcanvas_b200.so; else cp variants/libcanvas_b200.$v.so canvas_b200/libcanvas_b200.so; fi
  for w in 1080p_4layer_windows 4k_timeline_windows 4k_4layer_filters; do
    run $v $w CV_X=0
    run $v $w CV_RECT_PPL=2
    [ $w != 4k_4layer_filters ] && run $v $w CV_RECT_PPL=3
  done
done
cp /tmp/shipped.so canvas_b200/libcanvas_b200.so
