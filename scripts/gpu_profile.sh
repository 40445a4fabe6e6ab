# Round-1 ncu evidence: launch list of the default bench command + one full capture per hot kernel.
set -x
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/p_plain_default.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/p_launches_default.csv $B > gpurun_out/p_ncu1.log 2>&1
$B > gpurun_out/p_plain_default2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_stack_tma -s 30 -c 1 -o gpurun_out/p_k_stack_tma $B > gpurun_out/p_ncu2.log 2>&1
for w in 4k_timeline 4k_convert 1080p_cross_u8 4k_composite_convert 1080p_4layer_windows; do
  $B --workload $w > gpurun_out/p_plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_seg -s 4 -c 1 -o gpurun_out/p_k_seg_$w $B --workload $w > gpurun_out/p_ncu_$w.log 2>&1
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/p_launches_$w.csv $B --workload $w > /dev/null 2>&1
done
ls -la gpurun_out/
