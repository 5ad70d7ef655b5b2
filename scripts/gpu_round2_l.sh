#!/bin/bash
mkdir -p gpurun_out
run() { tag=$1; shift; env "$@" python bench.py --steps 1 --warmup 1 --no-cpu-baseline --sweep-workers 3 > gpurun_out/r02_det_$tag.json 2> /dev/null; python -c "
import json; d=json.load(open('gpurun_out/r02_det_$tag.json')); s=d['sweep']; print('$tag', s['answer_digest'], s['layers_converged'], round(s['value'],3))"; }
run la0_g2 RT_B200_LU_LOOKAHEAD=0 RT_B200_LU_GROUPS=2
run la1_g1 RT_B200_LU_LOOKAHEAD=1 RT_B200_LU_GROUPS=1
run la1_g1_noprio RT_B200_LU_LOOKAHEAD=1 RT_B200_LU_GROUPS=1 RT_B200_NO_STREAM_PRIO=1
run la1_g2_noprio_nopdl RT_B200_NO_STREAM_PRIO=1 RT_B200_PDL_MAX_LAYERS=0
