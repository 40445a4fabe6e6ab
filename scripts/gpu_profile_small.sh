#!/bin/bash
# ncu --set full of the small single-pass kernels (one launch each, after a few warm launches); digests into gpurun_out/
cd "$(dirname "$0")/.."
for k in k_to_argb32 k_color k_encode_u8 k_subsample k_subsample_mpeg2; do
  ncu --set full --clock-control none --import-source on -k regex:"^$k(<.*>)?\$" --launch-skip 5 -c 1 -o gpurun_out/r2_small_$k -f \
      python bench.py --steps 3 --warmup 3 --no-others --only-kernels --no-plugin --no-scheduler --no-cpu-baseline --no-check > gpurun_out/ncu_small_$k.log 2>&1
  python scripts/ncu_digest.py gpurun_out/r2_small_$k.ncu-rep 14 > gpurun_out/r2_small_${k}_digest.txt 2>&1
  rm -f gpurun_out/r2_small_$k.ncu-rep
done
