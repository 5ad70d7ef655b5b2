#!/bin/bash
for pdl in 64 96 128 160 200; do echo "== PDL_MAX $pdl"; RT_B200_PDL_MAX_LAYERS=$pdl python scripts/time_solve.py ch3oh 200 5 2>&1 | tail -1 | cut -c1-140; done
for gm in 24 32 48 64; do echo "== GROUP_MIN $gm"; RT_B200_LU_GROUP_MIN=$gm python scripts/time_solve.py ch3oh 200 5 2>&1 | tail -1 | cut -c1-140; done
for pdl in 96 200 400; do echo "== h2o PDL_MAX $pdl"; RT_B200_PDL_MAX_LAYERS=$pdl python scripts/time_solve.py h2o 200 5 2>&1 | tail -1 | cut -c1-140; done
