#!/bin/bash
# full GPU check: the gpu test-suite, then the bench on every workload (no CPU baseline except for the default one)
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/gpu_tests.log 2>&1; echo "gpu tests exit $?" >> gpurun_out/gpu_tests.log
tail -3 gpurun_out/gpu_tests.log
: > gpurun_out/bench_all.jsonl
for w in 1080p_convert 1080p_convert_planar 1080p_cross_u8 1080p_4layer_windows 4k_convert 4k_composite_convert 4k_timeline; do
  python bench.py --workload $w --steps 10 --warmup 3 --no-cpu-baseline >> gpurun_out/bench_all.jsonl 2>> gpurun_out/bench_err.log
done
python bench.py >> gpurun_out/bench_all.jsonl 2>> gpurun_out/bench_err.log
python - <<'PY'
import json
for l in open("gpurun_out/bench_all.jsonl"):
    try: d = json.loads(l)
    except Exception: continue
    print("%-24s %10.0f fps  %5.1f%%  e2e %8.1f  launches %d  cpu %s" % (d["config"]["workload"], d["value"], 100 * d["roofline"]["frac"], d["e2e"]["value"], d["gpu_launches"], d.get("cpu_baseline", {}).get("value")))
PY
