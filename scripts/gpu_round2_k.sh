#!/bin/bash
mkdir -p gpurun_out
for run in a b; do for W in 1 3; do
  python bench.py --steps 1 --warmup 3 --no-cpu-baseline --sweep-workers $W > gpurun_out/r02_det_w${W}${run}.json 2> gpurun_out/r02_det_w${W}${run}.err
  python -c "
import json; d=json.load(open('gpurun_out/r02_det_w${W}${run}.json')); s=d['sweep']; print('W=$W run $run', s['answer_digest'], s['layers_converged'], s['value'])"
done; done
RT_B200_PDL_MAX_LAYERS=0 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --sweep-workers 3 > gpurun_out/r02_det_nopdl.json 2> /dev/null; python -c "
import json; d=json.load(open('gpurun_out/r02_det_nopdl.json')); s=d['sweep']; print('W=3 no PDL', s['answer_digest'], s['layers_converged'], s['value'])"
RT_B200_LU_LOOKAHEAD=0 RT_B200_LU_GROUPS=1 python bench.py --steps 1 --warmup 3 --no-cpu-baseline --sweep-workers 3 > gpurun_out/r02_det_single.json 2> /dev/null; python -c "
import json; d=json.load(open('gpurun_out/r02_det_single.json')); s=d['sweep']; print('W=3 single stream', s['answer_digest'], s['layers_converged'], s['value'])"
