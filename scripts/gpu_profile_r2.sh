#!/bin/bash
# Round-2 ncu evidence, one box: the launch list of the default bench command, then one `--set full` capture of the
# dominant kernel of every bench workload (each only after its plain run exited 0).  The raw metric page of every
# capture is exported on the box (gpurun_out/p2_<workload>_raw.csv); only the three reports named in KEEP travel back
# (gpurun_out/ is limited to 64 MiB).
set -x
KEEP="4k_8layer_composite 4k_timeline 4k_timeline_windows"
B="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-others --no-plugin --no-scheduler --no-check"
$B > gpurun_out/p2_plain_default.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/p2_launches_4k_8layer_composite.csv $B > gpurun_out/p2_ncu_list.log 2>&1
for spec in 4k_8layer_composite:k_stack_tma:40 4k_composite_convert:k_seg:2 4k_convert:k_seg:4 1080p_convert:k_seg:6 1080p_convert_planar:k_seg:6 \
            1080p_cross_u8:k_seg:6 4k_timeline:k_seg:2 4k_timeline_windows:k_rect:2 1080p_4layer_windows:k_rect:4 4k_4layer_filters:k_rect:2; do
  w=${spec%%:*}; rest=${spec#*:}; k=${rest%%:*}; skip=${rest#*:}
  $B --workload $w > gpurun_out/p2_plain_$w.log 2>&1 &&
  ncu --set full --clock-control none --import-source on -k regex:$k -s $skip -c 1 -o gpurun_out/p2_$w $B --workload $w > gpurun_out/p2_ncu_$w.log 2>&1
  ncu -i gpurun_out/p2_$w.ncu-rep --page raw --csv > gpurun_out/p2_${w}_raw.csv 2>/dev/null
  python scripts/ncu_digest.py gpurun_out/p2_$w.ncu-rep 12 > gpurun_out/p2_${w}_digest.txt 2>&1
  case " $KEEP " in *" $w "*) ;; *) rm -f gpurun_out/p2_$w.ncu-rep ;; esac
done
rm -f gpurun_out/p2_plain_*.log gpurun_out/p2_ncu_*.log
du -sh gpurun_out; ls gpurun_out | head -50
