#!/bin/bash
# After ordering the H2D uploads on the model's stream: the sweep's answer digest with 3 models in flight per GPU, default
# schedule (groups 2, look-ahead, stream priorities, PDL), five repetitions + one single-model run
mkdir -p gpurun_out
for rep in 1 2 3 4 5; do python bench.py --steps 1 --warmup 1 --no-cpu-baseline --sweep-workers 3 > gpurun_out/r02_det_x.json 2> gpurun_out/r02_det_x.err; python -c "
import json; d=json.load(open('gpurun_out/r02_det_x.json')); s=d['sweep']; print('default W=3 rep $rep', s['answer_digest'], s['layers_converged'], round(s['value'],3))"; done
python bench.py --steps 1 --warmup 1 --no-cpu-baseline --sweep-workers 1 > gpurun_out/r02_det_x.json 2> /dev/null; python -c "
import json; d=json.load(open('gpurun_out/r02_det_x.json')); s=d['sweep']; print('default W=1', s['answer_digest'], s['layers_converged'], round(s['value'],3))"
python -m pytest tests -m gpu -x -q -p no:cacheprovider 2>&1 | tail -3
