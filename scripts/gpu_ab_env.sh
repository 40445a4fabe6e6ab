#!/bin/bash
# scripts/gpu_ab_env.sh "ENV=.. ENV=.." WORKLOAD...   -- same library, with and without the environment settings, two passes
envs=$1; shift
run() { env $2 python bench.py --workload $1 --steps 8 --no-others --no-cpu-baseline --no-plugin --no-scheduler 2>&1 | python -c "
import sys,json
for l in sys.stdin:
    if l.startswith('{'):
        d=json.loads(l); print('%-28s %-22s %9.0f fps  %.3f  ulp %s  %s MHz %s' % ('$3', d['config']['workload'], d['value'], d['roofline']['frac'], d.get('max_ulp'), d['clocks']['sm_mhz'], ','.join(d['clocks']['reasons'])))
    elif 'rror' in l: print(l.rstrip()[-300:])
"; }
for pass in 1 2; do for w in "$@"; do run $w "CV_NOTHING=1" "base"; run $w "$envs" "$envs"; done; done
