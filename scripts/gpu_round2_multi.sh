#!/bin/bash
# Multi-GPU evidence (run under gpurun --gpus N): bash scripts/gpu_round2_multi.sh N
N=$1
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
# (1) the driver's own command line: fixed-work line + the sweep of 64 DISTINCT grid points through SweepQueue
$TR --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r02_bench_ch3oh_n${N}_sweep.json 2> gpurun_out/r02_bench_ch3oh_n${N}_sweep.err; echo "sweep N=$N rc=$?"
# (2) the same sweep with four grid points stacked per solve
$TR --master-port 29512 bench.py --gpus $N --steps 2 --warmup 3 --sweep-batch 4 > gpurun_out/r02_bench_ch3oh_n${N}_sweep_batch4.json 2> gpurun_out/r02_bench_ch3oh_n${N}_sweep_batch4.err; echo "sweep batch4 N=$N rc=$?"
# (3) ONE slab sharded by layers (config 3: OH overlap 56 x 400; and the headline CH3OH slab), NCCL all_gather timed
$TR --master-port 29513 bench.py --gpus $N --mode shard --config oh_overlap --steps 5 --warmup 3 > gpurun_out/r02_bench_oh_overlap_n${N}_shard.json 2> gpurun_out/r02_bench_oh_overlap_n${N}_shard.err; echo "shard oh_overlap N=$N rc=$?"
$TR --master-port 29514 bench.py --gpus $N --mode shard --config ch3oh --steps 5 --warmup 3 > gpurun_out/r02_bench_ch3oh_n${N}_shard.json 2> gpurun_out/r02_bench_ch3oh_n${N}_shard.err; echo "shard ch3oh N=$N rc=$?"
for f in gpurun_out/r02_bench_*_n${N}_*.json; do echo $f; python - "$f" <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1]))
    print({k:d.get(k) for k in ("value","ms_per_step","n_gpus","scaling")}, "e2e", d["e2e"]["value"], "all_gather", d.get("all_gather"), "sweep", {k:d["sweep"][k] for k in ("value","wall_s","balance","tail_idle_s")} if d.get("sweep") else None)
except Exception as e:
    print("unreadable", e)
PY
done
for f in gpurun_out/r02_bench_*_n${N}_*.err; do tail -n 2 $f | cut -c1-200; done
