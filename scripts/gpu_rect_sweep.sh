#!/bin/bash
# k_rect tuning on one box: thread cap (shipped 512 vs variants), pairs per lane, rows per unit; prints fps and roofline fraction
run() { python bench.py --workload $1 --steps 8 --no-others --no-cpu-baseline $2 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('$3', d['config']['workload'], round(d['value']), round(d['roofline']['frac'],3), 'ulp', d.get('max_ulp'), d['clocks']['sm_mhz'])
    elif 'Error' in l or 'error' in l: print(l.rstrip()[-300:])
"; }
cp canvas_b200/libcanvas_b200.so /tmp/shipped.so
for v in shipped "$@"; do
  if [ $v = shipped ]; then cp /tmp/shipped.so canvas_b200/libcanvas_b200.so; else cp variants/libcanvas_b200.$v.so canvas_b200/libcanvas_b200.so; fi
  for ppl in 2 3; do for rows in 4 8 16; do
    for w in 1080p_4layer_windows 4k_timeline_windows; do
      CV_RECT_PPL=$ppl CV_RECT_ROWS=$rows run $w "" "$v ppl=$ppl rows=$rows"
    done
  done; done
done
cp /tmp/shipped.so canvas_b200/libcanvas_b200.so
