# usage: bash scripts/gpu_sweep2.sh "<workloads>" "<ENV=val ...>" ...   one bench line per (workload, env set)
WL="$1"; shift
for envs in "$@"; do
  for w in $WL; do
    env $envs python bench.py --workload $w --steps 5 --no-cpu-baseline 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print('$envs', d['config']['workload'], round(d['value'],1), 'fps frac', round(d['roofline']['frac'],3))
"
  done
done
