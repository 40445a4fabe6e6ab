rm -f gpurun_out/s2_tests.log gpurun_out/s2_bench.log
python -m pytest tests -m gpu -x -q 2>&1 | tail -15 >> gpurun_out/s2_tests.log
CV_SEG_NO_UNIFORM=1 python -m pytest tests -m gpu -x -q 2>&1 | tail -3 >> gpurun_out/s2_tests.log
for w in 4k_8layer_composite 4k_composite_convert 4k_convert 1080p_convert 1080p_convert_planar 1080p_cross_u8 1080p_4layer_windows 4k_timeline; do python bench.py --workload $w --steps 5 --no-cpu-baseline 2>&1 | grep '^{' | python -c "
import sys,json
for l in sys.stdin:
    d=json.loads(l); print(d['config']['workload'], round(d['value'],1), 'fps frac', round(d['roofline']['frac'],3), 'e2e', round(d['e2e']['value'],1))
" >> gpurun_out/s2_bench.log 2>&1; done
cat gpurun_out/s2_tests.log gpurun_out/s2_bench.log
