# Round-1 (second half) ncu evidence for the kernels that changed after profiles/r1_kernels.md was first written:
# compact conversion table, windowed batches, lazy planar groups.  Plain run first, ncu only if it exits 0.
# usage: gpu_profile2.sh WORKLOAD...   (<= 3 full captures per gpurun call: gpurun_out/ is capped at 64 MiB)
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline"
for w in "$@"; do
  $B --workload $w > gpurun_out/r_plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:k_seg -s 4 -c 1 -o gpurun_out/r_k_seg_$w $B --workload $w > gpurun_out/r_ncu_$w.log 2>&1
  ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/r_launches_$w.csv $B --workload $w > /dev/null 2>&1
  python bench.py --workload $w --content video --steps 10 --no-cpu-baseline 2>&1 | grep "^{" > gpurun_out/r_video_$w.json
done
ls -la gpurun_out | grep " r_" | wc -l
