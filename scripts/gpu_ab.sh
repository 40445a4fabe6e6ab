#!/bin/bash
# A/B of the shipped library against variants/libcanvas_b200.NAME.so on ONE box: scripts/gpu_ab.sh NAME WORKLOAD...
# (two passes, alternating, so that clock drift shows up as disagreement between the passes)
name=$1; shift
cp canvas_b200/libcanvas_b200.so /tmp/shipped.so
for pass in 1 2; do
  for v in shipped $name; do
    if [ $v = shipped ]; then cp /tmp/shipped.so canvas_b200/libcanvas_b200.so; else cp variants/libcanvas_b200.$v.so canvas_b200/libcanvas_b200.so; fi
    for w in "$@"; do
      python bench.py --workload $w --steps 8 --warmup 3 --no-cpu-baseline --no-others --no-plugin --no-scheduler --no-check 2>/dev/null | grep "^{" | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('pass $pass %-8s %-22s %9.0f fps  %.3f' % ('$v', d['config']['workload'], d['value'], d['roofline']['frac']))
"
    done
  done
done
cp /tmp/shipped.so canvas_b200/libcanvas_b200.so
