#!/bin/bash
mkdir -p gpurun_out
for W in 1 2 3; do
  python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sweep-workers $W > gpurun_out/r02_sweepw$W.json 2> gpurun_out/r02_sweepw$W.err; echo "W=$W rc=$?"
  python - <<PY
import json
d=json.load(open('gpurun_out/r02_sweepw$W.json')); s=d['sweep']
print('W=$W', 'value', d['value'], 'sweep', s['value'], s['wall_s'], s['balance'], s['layers_converged'], s['layers'])
PY
done
python bench.py --steps 2 --warmup 3 --no-cpu-baseline --sweep-workers 2 --sweep-batch 2 > gpurun_out/r02_sweepw2b2.json 2> gpurun_out/r02_sweepw2b2.err; python -c "
import json; d=json.load(open('gpurun_out/r02_sweepw2b2.json')); s=d['sweep']; print('W=2 M=2 sweep', s['value'], s['wall_s'])"
tail -3 gpurun_out/r02_sweepw2.err
