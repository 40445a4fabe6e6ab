# ncu evidence for the row-segment kernels as shipped at the end of round 1 (compact conversion table, windowed
# batches, lazy planar groups, 4-pair conversion segments).  Plain run first, ncu only if it exits 0.
# usage: gpu_profile2.sh WORKLOAD...   (<= 3 full captures per gpurun call: gpurun_out/ is capped at 64 MiB)
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
for w in "$@"; do
  K=k_seg; S=4
  if [ $w = 4k_8layer_composite ]; then K=k_stack_tma; S=30; fi
  $B --workload $w > gpurun_out/r_plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$K -s $S -c 1 -o gpurun_out/r_${K}_$w $B --workload $w > gpurun_out/r_ncu_$w.log 2>&1
  ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/r_launches_$w.csv $B --workload $w > /dev/null 2>&1
done
ls -la gpurun_out | grep " r_" | wc -l
