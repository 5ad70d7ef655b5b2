#!/bin/bash
mkdir -p gpurun_out
for lib in "" variants/librt_asm4.so; do
  for L in 148 200; do
    RT_B200_LIB=${lib:-radiative_transfer_b200/librt_b200.so} timeout 200 python scripts/time_solve.py h2o150 $L 3 2>&1 | tail -1 | cut -c1-260
  done
done
