#!/bin/bash
for pf in 0 148 296 592 1184; do
  echo "== PF $pf"; RT_B200_GEMM_PF=$pf python scripts/time_step.py ch3oh 200 5 2>&1 | tail -2
done
for pf in 0 592; do
  echo "== PF $pf L=60"; RT_B200_GEMM_PF=$pf python scripts/time_step.py ch3oh 60 5 2>&1 | tail -2
  echo "== PF $pf solve"; RT_B200_GEMM_PF=$pf python scripts/time_solve.py ch3oh 200 3 2>&1 | tail -1
done
