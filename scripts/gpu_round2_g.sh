#!/bin/bash
for L in 8 25 60 200; do
  for pdl in 0 100000; do
    echo "== L=$L PDL=$pdl"; RT_B200_PDL_MAX_LAYERS=$pdl python scripts/time_solve.py ch3oh $L 3 2>&1 | tail -1
  done
done
for L in 8 25; do
  echo "== L=$L lookahead 0 PDL on"; RT_B200_LU_LOOKAHEAD=0 RT_B200_PDL_MAX_LAYERS=100000 python scripts/time_solve.py ch3oh $L 3 2>&1 | tail -1
  echo "== L=$L lookahead 0 PDL off"; RT_B200_LU_LOOKAHEAD=0 python scripts/time_solve.py ch3oh $L 3 2>&1 | tail -1
done
echo "== h2o PDL 0/on"; python scripts/time_solve.py h2o 200 3 2>&1 | tail -1; RT_B200_PDL_MAX_LAYERS=100000 python scripts/time_solve.py h2o 200 3 2>&1 | tail -1
