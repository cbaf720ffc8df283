#!/bin/bash
# T = 2^22: level-group size of the encode passes as a function of the assumed L2 size (groups fit 55 % of it)
for mb in 48 64 126 192 256 512; do
  echo "LONER_B200_L2_BYTES=${mb}M"
  LONER_B200_L2_BYTES=$((mb*1024*1024)) python profiles/bench_kernels.py 22 2>&1 | grep -E "encode_fwd|encode_bwd table only  |sigma_fwd \(rays\) \+|sigma_bwd params  "
done
