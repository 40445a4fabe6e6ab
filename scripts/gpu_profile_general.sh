# ncu evidence for the window-classifying (GENERAL) kernels: 1080p 4-layer windows (bench.py) and the 4K partial-window
# timeline of the scheduler bench.  Plain run first, ncu only if it exits 0.
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --workload 1080p_4layer_windows"
$B > gpurun_out/g_plain_windows.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_seg -s 4 -c 1 -o gpurun_out/g_k_seg_windows $B > gpurun_out/g_ncu_windows.log 2>&1
ncu --metrics gpu__time_duration.sum --clock-control none -c 120 --csv --log-file gpurun_out/g_launches_windows.csv $B > /dev/null 2>&1
S="python scripts/bench_scheduler.py --gpus 1"
$S > gpurun_out/g_plain_sched.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_seg -s 4 -c 1 -o gpurun_out/g_k_seg_sched $S > gpurun_out/g_ncu_sched.log 2>&1
ls -la gpurun_out | grep " g_"
