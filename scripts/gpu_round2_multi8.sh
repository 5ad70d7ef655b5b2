#!/bin/bash
# 8-GPU re-check of the driver's command line with the final tree
N=8
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_ch3oh_n${N}_sweep.json 2> gpurun_out/r02_bench_ch3oh_n${N}_sweep.err; echo "sweep N=$N rc=$?"
python - <<'PY'
import json
d=json.load(open("gpurun_out/r02_bench_ch3oh_n8_sweep.json"))
print({k:d.get(k) for k in ("value","ms_per_step","n_gpus","scaling")}, "e2e", d["e2e"]["value"], "sweep", {k:d["sweep"][k] for k in ("value","answer_digest")} if d.get("sweep") else None)
PY
