#!/bin/bash
# other_kernels of bench.py with and without the staged-table variants (CV_NO_STAGED_TABLES=1)
cd "$(dirname "$0")/.."
for v in 0 1; do
  CV_NO_STAGED_TABLES=$v python bench.py --steps 3 --warmup 3 --no-plugin --no-scheduler --no-cpu-baseline --no-others --only-kernels > gpurun_out/small_$v.json 2> gpurun_out/small_$v.err || tail -3 gpurun_out/small_$v.err
  python - <<P
import json
d=json.loads(open('gpurun_out/small_$v.json').read().strip().splitlines()[-1])
print("CV_NO_STAGED_TABLES=$v", [(k['op'], round(k['value']), round(k['roofline_frac'],3)) for k in d.get('other_kernels',[])])
P
done
