#!/bin/bash
# the driver's own multi-GPU command line with the final defaults (125 distinct grid points, 3 models in flight per GPU)
N=$1
mkdir -p gpurun_out
python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_ch3oh_n${N}_sweep.json 2> gpurun_out/r02_bench_ch3oh_n${N}_sweep.err; echo "sweep N=$N rc=$?"
python - <<PY
import json
d=json.load(open('gpurun_out/r02_bench_ch3oh_n${N}_sweep.json')); s=d['sweep']
print('N=$N value', d['value'], 'e2e', d['e2e']['value'], 'sweep', s['value'], s['wall_s'], 'balance', s['balance'], 'tail', s['tail_idle_s'], s['points'], s['models_in_flight_per_gpu'], s['layers_converged'], s['layers'])
PY
tail -n 2 gpurun_out/r02_bench_ch3oh_n${N}_sweep.err | cut -c1-200
