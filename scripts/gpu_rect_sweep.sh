#!/bin/bash
# k_rect tuning on one box: scripts/gpu_rect_sweep.sh "VARIANT..." "PPL..." "ROWS..." WORKLOAD...   (uniform jobs go through k_rect too)
variants=$1; ppls=$2; rowss=$3; shift 3
run() { python bench.py --workload $1 --steps 8 --no-others --no-cpu-baseline 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('%-26s %-22s %9.0f fps  %.3f  ulp %s  %s MHz' % ('$2', d['config']['workload'], d['value'], d['roofline']['frac'], d.get('max_ulp'), d['clocks']['sm_mhz']))
    elif 'rror' in l: print(l.rstrip()[-300:])
"; }
cp canvas_b200/libcanvas_b200.so /tmp/shipped.so
for v in $variants; do
  if [ $v = shipped ]; then cp /tmp/shipped.so canvas_b200/libcanvas_b200.so; else cp variants/libcanvas_b200.$v.so canvas_b200/libcanvas_b200.so; fi
  for ppl in $ppls; do for rows in $rowss; do for w in "$@"; do
      CV_RECT_UNIFORM=1 CV_RECT_PPL=$ppl CV_RECT_ROWS=$rows run $w "$v ppl=$ppl rows=$rows"
  done; done; done
done
cp /tmp/shipped.so canvas_b200/libcanvas_b200.so
