#!/bin/bash
# N = 2 re-check after the fused-solver changes: the driver's command line and the sharded OH-overlap slab
N=2
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_ch3oh_n${N}_sweep.json 2> gpurun_out/r02_bench_ch3oh_n${N}_sweep.err; echo "sweep N=$N rc=$?"
timeout 300 $TR --master-port 29513 bench.py --gpus $N --mode shard --config oh_overlap --steps 5 --warmup 3 > gpurun_out/r02_bench_oh_overlap_n${N}_shard.json 2> gpurun_out/r02_bench_oh_overlap_n${N}_shard.err; echo "shard oh_overlap N=$N rc=$?"
for f in gpurun_out/r02_bench_ch3oh_n2_sweep.json gpurun_out/r02_bench_oh_overlap_n2_shard.json; do python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    print({k:d.get(k) for k in ("value","ms_per_step","n_gpus","scaling")}, "e2e", d["e2e"]["value"], "all_gather", d.get("all_gather"), "sweep", {k:d["sweep"][k] for k in ("value","answer_digest")} if d.get("sweep") else None)
except Exception as e:
    print("unreadable", e)
PY
done
