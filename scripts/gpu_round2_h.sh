#!/bin/bash
for L in 200 120 60; do
  echo "== L=$L prio on";  python scripts/time_solve.py ch3oh $L 3 2>&1 | tail -1
  echo "== L=$L prio off"; RT_B200_NO_STREAM_PRIO=1 python scripts/time_solve.py ch3oh $L 3 2>&1 | tail -1
done
echo "== h2o prio on"; python scripts/time_solve.py h2o 200 3 2>&1 | tail -1
echo "== h2o prio off"; RT_B200_NO_STREAM_PRIO=1 python scripts/time_solve.py h2o 200 3 2>&1 | tail -1
for G in 3 4; do echo "== L=200 groups $G prio on"; RT_B200_LU_GROUPS=$G python scripts/time_solve.py ch3oh 200 3 2>&1 | tail -1; done
