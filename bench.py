#!/usr/bin/env python
"""bench.py -- converged slab models / s (and ms per ALI iteration) of the level-population hot path on B200.

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path (one rank per GPU under torchrun)
    python bench.py --impl reference --gpus N --steps K ...  # the reference's own CPU solver on the host cores

Workload (BASELINE.json configs[1]): CH3OH-A class I/II maser slab, N = 768 levels, L = 200 layers, 3 collisional
tables mixed by the reference's vt / J<=9 rule, dust continuum, Ng acceleration, 150-iteration cap, convergence
1e-5, every layer started from boundary_layer_populations.  Synthetic molecule / cloud (the reference ships no data).

One STEP = one slab model taken from "cloud resident in HBM" to "all layers converged or capped":
    value leg : rt_cloud_rebuild (collisional matrices of all layers) + rt_solve_device, populations stay in HBM
    e2e leg   : rt_cloud_upload (pinned host -> device) + rt_calc_molecular_populations (populations + per-layer
                status back on the host) -- the call a user of the C ABI makes
N GPUs: model-per-GPU (BASELINE.json configs[4], no collective): every rank solves its own slab of the same shape
(same molecule and cloud on every rank: fixed per-GPU work), `value` = N models / max-over-ranks time.  scaling = weak.
Beside it, in the same run, the `sweep` object: a parameter-grid sweep of DISTINCT models (density x temperature x
specific column density, --sweep-grid G -> G^3 points of synthetic.sweep_layers; the SAME G^3 points at every N) handed
to the ranks dynamically by sharding.SweepQueue -- the analogue of the reference's `omp for schedule(dynamic,1)` over
shock models (radiative_transfer.cpp:152-159): one rt_model per GPU (tables uploaded once), rt_cloud_upload +
rt_calc_molecular_populations per draw with host buffers, `sweep.value` = points / max-over-ranks wall time.
`--mode shard` instead partitions ONE slab's layers across ranks (strong scaling; cyclic layer assignment, one NCCL
all_gather of the populations per solve, timed separately).
`--start chained` measures the reference's own start mode (warm-start chain, radiative_transfer.cpp:247-253) with
--batch M stacked models advancing chain step l together (rt_layer_soa::model_first).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

# stdout carries exactly one JSON line.  Libraries print there too (NCCL's "NCCL version ..." banner ignores
# NCCL_DEBUG_FILE), so file descriptor 1 is pointed at stderr for the whole run and the JSON line goes to the saved one.
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
_REAL_STDOUT = os.dup(1)
os.dup2(2, 1)


def emit(line: str) -> None:
    os.write(_REAL_STDOUT, (line + "\n").encode())

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "converged_slab_models_per_s"
UNIT = "models/s"
# CPU arm: layers per host thread in one step's sample.  One layer per thread would time the SLOWEST sampled layer (iteration
# counts along the CH3OH slab: 16 .. 67) and bias the extrapolated models/s low by ~1.35x; three per thread under
# schedule(dynamic,1) average that out while 25 driver steps still end within a few minutes.
CPU_LAYERS_PER_CORE = 3


def log(*a):
    print(*a, file=sys.stderr, flush=True)


# ----------------------------------------------------------------------------------------------------------------
# workload
# ----------------------------------------------------------------------------------------------------------------
def make_problem(args, rank: int):
    from radiative_transfer_b200 import synthetic
    if args.config == "ch3oh":
        p = synthetic.config_ch3oh(L=args.layers or 200)
        name = f"CH3OH-A slab N={p.N} L={p.L} (BASELINE configs[1])"
    elif args.config == "h2o":
        p = synthetic.config_h2o(L=args.layers or 200)
        name = f"o-H2O slab N={p.N} L={p.L} (BASELINE configs[3])"
    elif args.config == "h2o150":
        p = synthetic.config_h2o(L=args.layers or 200, N=150)
        name = f"o-H2O slab N={p.N} L={p.L} (the level count the reference driver uses, radiative_transfer.cpp:901,1035)"
    elif args.config == "oh_overlap":
        p = synthetic.config_oh(L=args.layers or 400, overlap=True)
        name = f"OH hyperfine-overlap slab N={p.N} L={p.L} (BASELINE configs[2])"
    elif args.config == "oh":
        p = synthetic.config_oh(L=args.layers or 100, overlap=False)
        name = f"OH slab N={p.N} L={p.L} (BASELINE configs[0])"
    else:
        raise SystemExit(f"unknown --config {args.config}")
    # model-per-GPU (mode sweep): every rank solves a slab of the SAME shape and content, so per-GPU work is fixed
    # as N grows (weak scaling); the models are independent -- no collective anywhere on the path
    return p, name


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled every 200 ms during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device: int):
        self.device = device
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.device)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except subprocess.TimeoutExpired:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for nm, v in zip(names, f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(nm)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


def measured_peaks():
    f = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(f):
        d = json.load(open(f))
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md)"


def pinned_layers(p):
    """Layer arrays in pinned host memory (torch owns the allocation; numpy views go through the C ABI)."""
    import torch
    lay, keep = {}, []
    for k, v in p.layers.items():
        if k == "height":
            lay[k] = v
            continue
        t = torch.empty(v.shape, dtype=torch.float64).pin_memory()
        t.numpy()[...] = v
        keep.append(t)
        lay[k] = t.numpy()
    return lay, keep


# ----------------------------------------------------------------------------------------------------------------
# CPU baseline / reference arm: the reference's own solver (oracle/_ref) on the host cores
# ----------------------------------------------------------------------------------------------------------------
def cpu_solve_sample(p, cores: int, sample_layers: int, which: int = 0, retry_cap: int = 15000):
    """Solves `sample_layers` evenly spaced layers of the slab on `cores` threads with the compiled reference
    (oracle/_ref, else the C port), under the same policy as this repo's e2e leg: accelerated solve, then the
    reference's un-accelerated retry of the layers that failed, capped at `retry_cap` iterations
    (radiative_transfer.cpp:259-275; a no-op for CH3OH and for un-accelerated molecules).
    Returns (models/s, seconds, kind, checker description)."""
    from oracle.loader import Checker, ref_available
    kind = "reference" if ref_available() else "port"
    chk = Checker("ref" if kind == "reference" else "port")
    L = p.L
    n = min(sample_layers, L)
    # evenly spaced over the slab (iteration counts vary smoothly along it), shifted from step to step so that successive
    # steps cover different layers; several layers per thread under schedule(dynamic,1), so a step's wall time is the
    # load-balanced time of the sample, not its slowest layer
    idx = np.unique((np.linspace(0, L - 1, n, endpoint=True) + which * max(1, L // (4 * n))).astype(int) % L)
    n = len(idx)
    lay = {k: (v if k == "height" else np.ascontiguousarray(v[idx])) for k, v in p.layers.items()}
    import copy
    q = copy.copy(p)
    q.layers = lay
    t0 = time.perf_counter()
    pop, st, sec = chk.solve(q, nthreads=cores, retry=True, max_iter_ext=retry_cap)
    wall = time.perf_counter() - t0
    models_per_s = (n / sec) / L
    return models_per_s, sec, kind, dict(layers=n, iters_mean=float(st["iters"].mean()), converged=int(st["converged"].sum()), wall_s=wall)


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    p, name = make_problem(args, 0)
    cores = os.cpu_count() or 1
    vals, secs = [], []
    info = None
    for s in range(args.warmup + args.steps):
        v, sec, kind, info = cpu_solve_sample(p, cores, CPU_LAYERS_PER_CORE * cores, which=s, retry_cap=args.retry_cap)
        log(f"[reference] step {s}: {info['layers']} layers on {cores} threads in {sec:.2f}s -> {v:.5f} models/s")
        if s >= args.warmup:
            vals.append(v); secs.append(sec)
    value = float(len(vals) / sum(1.0 / v for v in vals))  # total models / total time
    ms_per_step = float(np.mean(secs)) * 1e3
    out = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f64", "data": "synthetic", "config": {"workload": name, "start": "boundary", "accel": bool(p.accel), "retry_cap": args.retry_cap},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": cores, "kind": kind,
                         "sample": f"each step: {info['layers']} evenly spaced layers of the {p.L}-layer slab ({CPU_LAYERS_PER_CORE} per thread, schedule(dynamic,1)) solved under the reference's policy "
                                   f"(retry of failed layers included) on {cores} threads (layers are independent); models/s = layers/s / {p.L}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(json.dumps(out))
    return 0


# ----------------------------------------------------------------------------------------------------------------
# this repo's arm
# ----------------------------------------------------------------------------------------------------------------
def run_sweep_leg(args, m, base, dist, world, rank, abi, device):
    """Config 5 for real: G^3 DISTINCT grid points drawn from a shared queue, the per-GPU rt_model reused (tables stay
    resident), every draw = rt_cloud_upload + rt_calc_molecular_populations through host buffers.  The same G^3 points
    run at every N (strong scaling of one sweep), so sweep.value(N) / (N sweep.value(1)) is the sweep's efficiency.
    --sweep-workers W: W models in flight per GPU (W host threads, each with its own rt_model handle and streams, drawing
    from the same queue; the C-ABI calls release the GIL): while one model is in the latency-bound tail of its solve (a
    handful of active layers) the other's full-batch iterations fill the SMs -- SURVEY.md section 8e: "batch several models
    per GPU concurrently (streams) to fill the machine"."""
    import copy
    import ctypes as C
    import hashlib
    import threading
    import torch
    from radiative_transfer_b200 import sharding, solver, synthetic
    G = args.sweep_grid
    npts = G ** 3
    M = max(1, args.sweep_batch)
    W = max(1, args.sweep_workers)
    ncomp = len(base.dust)
    models = [m] + [solver.Model(base, device=device, upload_cloud=False) for _ in range(W - 1)]

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def run_once(tag):
        q = sharding.SweepQueue(npts, name="rt_bench_sweep_" + tag)
        lock = threading.Lock()
        res = dict(points=[], busy_s=0.0, iters_max_per_point=[], layers_converged=0, layers=0, digests={})
        errs = []

        def draw():
            pts = []
            with lock:  # one draw at a time per rank: the store round trip is atomic, the local fallback counter is not
                for _ in range(M):
                    try:
                        pts.append(next(q))
                    except StopIteration:
                        break
            return pts

        def worker(mod):
            try:
                torch.cuda.set_device(device)
                while True:
                    pts = draw()
                    if not pts:
                        return
                    t1 = time.perf_counter()
                    probs = []
                    for pt in pts:
                        qq = copy.copy(base)
                        qq.layers = synthetic.sweep_layers(pt, G, L=base.L, ncomp=ncomp)
                        probs.append(qq)
                    stk = probs[0] if len(probs) == 1 else sharding.stack_clouds(probs)[0]
                    mod.cloud_upload(stk)
                    pop = np.empty((stk.L, stk.N))
                    stat = np.zeros(stk.L, dtype=abi.STATUS_DTYPE)
                    mod._ck(mod.L_.rt_calc_molecular_populations(mod.h, abi.dptr(pop), abi.RT_START_BOUNDARY, int(base.accel), 150,
                                                                 int(args.retry_cap), C.c_double(1e-5),
                                                                 stat.ctypes.data_as(C.POINTER(abi.rt_layer_status))),
                            "rt_calc_molecular_populations")
                    dt = time.perf_counter() - t1
                    with lock:
                        res["busy_s"] += dt
                        res["points"] += pts
                        for k in range(len(pts)):
                            res["iters_max_per_point"].append(int(stat["iters"][k * base.L:(k + 1) * base.L].max()))
                            lo_, hi_ = k * base.L, (k + 1) * base.L
                            res["digests"][int(pts[k])] = hashlib.sha256(pop[lo_:hi_].tobytes() + stat["iters"][lo_:hi_].tobytes()).hexdigest()
                        res["layers_converged"] += int(stat["converged"].sum())
                        res["layers"] += stk.L
            except Exception as e:  # noqa: BLE001 -- reported by the main thread
                errs.append(e)

        barrier()
        t0 = time.perf_counter()
        ths = [threading.Thread(target=worker, args=(mod,)) for mod in models]
        for t in ths:
            t.start()
        for t in ths:
            t.join()
        res["end_s"] = time.perf_counter() - t0
        if errs:
            raise errs[0]
        barrier()
        return res

    run_once("warm") if args.sweep_warm else None
    r = run_once("timed")
    for mod in models[1:]:
        mod.close()
    if dist is not None:
        allr = [None] * world
        dist.all_gather_object(allr, r)
    else:
        allr = [r]
    wall = max(x["end_s"] for x in allr)
    done = sorted(sum((x["points"] for x in allr), []))
    assert done == list(range(npts)), "the queue must hand every grid point out exactly once"
    its = sum((x["iters_max_per_point"] for x in allr), [])
    busy = [x["busy_s"] / W for x in allr]
    dig = {}
    for x in allr:
        dig.update(x["digests"])
    # one digest over the populations + iteration counts of every grid point in point order: the answer of the sweep does not
    # depend on how many GPUs, workers or draws produced it (layers are independent, kernels deterministic)
    digest = hashlib.sha256("".join(dig[k] for k in sorted(dig)).encode()).hexdigest()[:16]
    return {
        "value": npts / wall, "unit": "distinct models/s", "points": npts, "grid": f"{G}x{G}x{G} (n_H2 x T x n_mol/|dv/dz|)",
        "models_per_solve": M, "models_in_flight_per_gpu": W, "wall_s": wall, "answer_digest": digest, "scaling": "strong (the same %d points at every N)" % npts,
        "per_rank": [{"points": len(x["points"]), "busy_s": round(x["busy_s"] / W, 4), "end_s": round(x["end_s"], 4)} for x in allr],
        "balance": (sum(busy) / len(busy)) / wall,
        "tail_idle_s": wall - min(x["end_s"] for x in allr),
        "iters_max_per_point": {"min": int(min(its)), "mean": float(np.mean(its)), "max": int(max(its))},
        "layers_converged": int(sum(x["layers_converged"] for x in allr)), "layers": int(sum(x["layers"] for x in allr)),
        "schedule": "sharding.SweepQueue: atomic counter in the process group's store, %d point(s) per draw, %d worker thread(s) "
                    "(rt_model handles) per GPU, no collective" % (M, W),
        "timing": "host wall clock from a barrier to each rank's last result on the host; max over ranks",
    }


def run_gpu(args):
    import ctypes as C
    import torch
    from radiative_transfer_b200 import abi, sharding, solver

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device -- this framework has no CPU path (use --impl reference for the CPU baseline)")
    torch.cuda.set_device(local)
    dist = None
    if world > 1:
        import torch.distributed as dist
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if args.gpus != world and rank == 0 and world > 1:
        log(f"note: --gpus {args.gpus} but WORLD_SIZE {world}; using WORLD_SIZE")
    n_gpus = world
    chained = args.start == "chained"
    shard = args.mode == "shard" and world > 1

    p_full, name = make_problem(args, rank)
    p = p_full
    if shard:
        # independent start: layers dealt out round-robin so that every rank gets the same mix of fast and slow layers
        # (iteration counts grow along the slab); the chained start keeps contiguous blocks (the chain stays inside a rank)
        mine = sharding.layer_indices(p_full.L, world, rank, cyclic=not chained)
        p = p_full.take(mine)
        name += f" sharded by layers over {world} GPUs ({'contiguous blocks' if chained else 'cyclic'})"
    if args.batch > 1:
        # sweep throughput mode: --batch M models per solve as ONE cloud of M*L independent layers (sharding.stack_clouds);
        # the headline `value` is quoted at the default M = 1, the slab of BASELINE configs[1]
        if chained and args.config == "ch3oh":
            import copy
            from radiative_transfer_b200 import synthetic
            probs = []
            for k in range(args.batch):  # distinct grid points: the chains of a sweep differ in length of convergence
                q = copy.copy(p_full)
                q.layers = synthetic.sweep_layers((k * 7) % 64, 4, L=p_full.L, ncomp=len(p_full.dust))
                probs.append(q)
            p, _ = sharding.stack_clouds(probs)
            name += f" x {args.batch} distinct grid points per solve"
        else:
            p, _ = sharding.stack_clouds([p_full] * args.batch)
            name += f" x {args.batch} models per solve"
    N, L = p.N, p.L
    start_mode = abi.RT_START_CHAINED if chained else p.start_mode
    m = solver.Model(p, device=local, upload_cloud=False)
    lay_pinned, keep = pinned_layers(p)
    p.layers = lay_pinned
    m.cloud_upload(p)
    pop_host = torch.empty((L, N), dtype=torch.float64).pin_memory()
    pop_np = pop_host.numpy()
    pop_dev = torch.empty((L, N), dtype=torch.float64, device="cuda") if shard else None
    gathered = {}

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def calc_host():
        stat = np.zeros(L, dtype=abi.STATUS_DTYPE)
        m._ck(m.L_.rt_calc_molecular_populations(m.h, abi.dptr(pop_np), int(start_mode), int(p.accel), 150, int(args.retry_cap),
                                                 C.c_double(1e-5), stat.ctypes.data_as(C.POINTER(abi.rt_layer_status))),
              "rt_calc_molecular_populations")
        return stat

    def gather():
        # one all_gather of the populations per sharded solve, device to device over NCCL
        m.copy_populations_device(pop_dev.data_ptr())
        gathered["pop"] = sharding.all_gather_layers(pop_dev, p_full.L, world, rank, cyclic=not chained)

    def step_device():
        m.cloud_rebuild()
        if chained:
            calc_host()  # the chain is driven by rt_calc_molecular_populations (one host read of the flags per chain step)
        else:
            m.solve_device(start_mode=start_mode, max_iter=p.max_iter, accel=p.accel)
        if shard:
            gather()

    # ---- value leg: device-resident ------------------------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = solver.launch_count()
    m.profile_reset(timing=False)
    t_wall0 = time.perf_counter()
    m.timer_begin()
    for _ in range(args.steps):
        step_device()
    ms_total = m.timer_end()
    torch.cuda.synchronize()
    ms_wall = (time.perf_counter() - t_wall0) * 1e3
    if shard:
        ms_total = ms_wall  # the NCCL all_gather runs on torch's stream: the step is timed on the host around device syncs
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = solver.launch_count() - launches0
    iters_total = m.iterations_enqueued()
    st = m.get_status()
    nccl_ms = None
    if shard:
        ev = [torch.cuda.Event(enable_timing=True) for _ in range(2)]
        for _ in range(3):
            gather()
        torch.cuda.synchronize()
        t_g = time.perf_counter()
        for _ in range(10):
            gather()
        torch.cuda.synchronize()
        nccl_ms = (time.perf_counter() - t_g) * 1e3 / 10
    # the same K steps once more with a CUDA-event pair around every kernel class (the event pairs cost ~10 % of a step
    # made of ~5000 short launches, so they stay out of the `value` timing); roofline numbers come from this pass
    m.profile_reset(timing=True)
    m.timer_begin()
    for _ in range(args.steps):
        step_device()
    ms_profiled = m.timer_end()
    prof = m.profile_get()
    m.profile_reset(timing=False)
    t = torch.tensor([ms_total], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_total_max = float(t.item())
    ms_per_step = ms_total_max / args.steps
    models_per_step = (n_gpus if args.mode == "sweep" else 1) * args.batch
    value = models_per_step / (ms_per_step * 1e-3)

    # ---- e2e leg: host buffers through the C ABI ---------------------------------------------------------------------
    def step_e2e():
        m.cloud_upload(p)  # pinned host -> device + collisional matrices
        stat = calc_host()
        if shard:
            gather()
            torch.cuda.synchronize()
        return stat

    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        stat = step_e2e()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = models_per_step / (float(t.item()) / args.steps)
    h2d = sum(v.nbytes for k, v in p.layers.items() if k != "height")
    d2h = pop_np.nbytes + stat.nbytes
    conv_all, lay_all = int(st["converged"].sum()), L
    if dist is not None:
        t = torch.tensor([conv_all, lay_all], dtype=torch.float64, device="cuda")
        dist.all_reduce(t)
        conv_all, lay_all = int(t[0].item()), int(t[1].item())

    # ---- config 5: the sweep of distinct models -------------------------------------------------------------------------
    sweep = None
    if args.mode == "sweep" and args.config == "ch3oh" and args.sweep_grid > 0 and args.batch == 1 and not chained:
        sweep = run_sweep_leg(args, m, p_full, dist, world, rank, abi, local)

    if rank != 0:
        m.close()
        if dist is not None:
            dist.barrier()
            dist.destroy_process_group()
        return 0

    # ---- rooflines ---------------------------------------------------------------------------------------------------
    hbm_peak, hbm_src = measured_peaks()
    dfma, dmma = solver.device_fp64_peak(local)
    fp64_peak = max(dfma, dmma)
    kern = {}
    for k, v in prof.items():
        if v["launches"] == 0 or v["ms"] <= 0:
            continue
        sec = v["ms"] * 1e-3
        kern[k] = {"ms_per_step": v["ms"] / args.steps, "launches_per_step": v["launches"] / args.steps,
                   "gbs": v["bytes"] / sec * 1e-9, "tflops": v["flops"] / sec * 1e-12,
                   "share": v["ms"] / ms_profiled}
    traffic = {}
    tf = os.path.join(ROOT, "profiles", "traffic.json")
    if os.path.exists(tf) and args.config == "ch3oh" and args.layers in (None, 0, 200) and args.batch == 1 and not shard and not chained:
        # DRAM bytes of one launch of each class from a committed `ncu --set full` capture of THIS workload.  The capture is
        # static; it is only meaningful while the classes it names are the classes that ran.
        traffic = json.load(open(tf))
        stale = [k for k in traffic if not k.startswith("_") and k not in prof]
        if stale:
            raise SystemExit(f"bench.py: profiles/traffic.json names kernel classes that did not run in this step: {stale} -- "
                             "re-capture it (ncu --set full) for the current kernels")
    top = max(kern, key=lambda k: kern[k]["ms_per_step"])

    def tbytes(k):
        t = traffic.get(k)
        return t.get("bytes") if isinstance(t, dict) else t

    def roof(k):
        v, pv = kern[k], prof[k]
        ai = pv["flops"] / max(pv["bytes"], 1.0)
        src = {"traffic_capture": traffic.get("_source"), "traffic_capture_commit": traffic.get("_commit")} if traffic else {}
        if ai > fp64_peak * 1e3 / hbm_peak:  # flop/byte above the machine balance -> FP64 pipe bound
            return {"kernel": k, "bound": "tensor", "achieved": v["tflops"], "peak": fp64_peak, "unit": "TFLOP/s",
                    "frac": v["tflops"] / fp64_peak, "traffic": tbytes(k), "traffic_detail": traffic.get(k),
                    "peak_source": f"FP64 measured in this run by rt_device_fp64_peak: DFMA {dfma:.1f}, DMMA {dmma:.1f} TFLOP/s "
                                   "(MEASURED_PEAKS.json has no FP64 figure; the LU runs on the FP64 pipe, not bf16 tcgen05)",
                    "share_of_step": v["share"], **src}
        return {"kernel": k, "bound": "hbm", "achieved": v["gbs"], "peak": hbm_peak, "unit": "GB/s", "frac": v["gbs"] / hbm_peak,
                "traffic": tbytes(k), "traffic_detail": traffic.get(k), "peak_source": hbm_src, "share_of_step": v["share"], **src}

    lu_classes = [k for k in ("lu_panel", "lu_trsm", "lu_gemm", "lu_backsolve", "lu_small") if k in kern]
    lu_all = None
    if lu_classes:
        fl = sum(prof[k]["flops"] for k in lu_classes)
        msl = sum(prof[k]["ms"] for k in lu_classes)
        lu_all = {"tflops": fl / (msl * 1e-3) * 1e-12, "frac_of_fp64_peak": fl / (msl * 1e-3) * 1e-12 / fp64_peak,
                  "ms_per_step": msl / args.steps, "classes": lu_classes,
                  "note": "algorithmic flops of the whole LU (panel + TRSM + trailing update + back substitution) over the sum of their event times"}

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak" if args.mode == "sweep" else "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": name, "start": args.start, "accel": bool(p.accel), "max_iter": p.max_iter, "min_err": 1e-5,
                   "retry_cap": args.retry_cap,
                   "mode": args.mode, "models_per_solve": args.batch, "l2": "per-step working set (C and M matrices, 2 x %.0f MB) exceeds the 126 MB L2"
                   % (L * ((N + 31) // 32 * 32) ** 2 * 8 / 1e6),
                   "timing": ("host wall clock around K steps (device work + NCCL all_gather on torch's stream, synchronised); max over ranks"
                              if shard else "CUDA events on the library's stream around K steps; max over ranks") +
                             "; per-kernel times from a second pass of the same K steps with an event pair around every kernel class"},
        "ms_per_step_profiled": ms_profiled / args.steps,
        "ms_per_iteration": ms_total / max(1, iters_total), "iterations_per_step": iters_total / args.steps,
        "layers_converged": conv_all, "layers": lay_all, "iters_mean": float(st["iters"].mean()),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "timing": "host wall clock around rt_cloud_upload + rt_calc_molecular_populations (synchronous calls)" +
                          (" + the all_gather of the populations" if shard else "")},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof(top),
        "roofline_coupling": roof("assemble") if "assemble" in kern else None,
        "lu_whole": lu_all,
        "kernels": kern,
        "fp64_peak_tflops": {"dfma": dfma, "dmma": dmma},
    }
    if shard:
        out["all_gather"] = {"ms": nccl_ms, "bytes_per_rank": int(pop_dev.numel() * 8), "backend": "nccl",
                             "what": "rt_copy_populations_device + one all_gather_into_tensor of the padded [L/P, N] blocks + scatter into slab order"}
    if sweep is not None:
        out["sweep"] = sweep
    if n_gpus == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        nl = p_full.L if args.cpu_full_slab else min(p_full.L, max(4 * cores, 16))
        v, sec, kind, info = cpu_solve_sample(p_full, cores, nl, retry_cap=args.retry_cap)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                               "sample": f"{info['layers']} evenly spaced layers of the {p_full.L}-layer slab, schedule(dynamic,1) over {cores} threads, solved under the reference's "
                                         f"policy (retry of failed layers included) in {sec:.1f} s (mean {info['iters_mean']:.0f} iterations); models/s = layers/s / {p_full.L}"}
    emit(json.dumps(out))
    m.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--batch", type=int, default=1, help="models per solve (sweep mode): M slabs stacked into one cloud")
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="ch3oh", choices=["ch3oh", "h2o", "h2o150", "oh", "oh_overlap"])
    ap.add_argument("--retry-cap", type=int, default=15000,
                    help="iteration cap of the un-accelerated retry of failed layers (the reference's 15000), used by the "
                         "e2e leg and by the CPU arm alike; at N=411 one failing layer costs the CPU minutes at 15000")
    ap.add_argument("--layers", type=int, default=0)
    ap.add_argument("--mode", default="sweep", choices=["sweep", "shard"])
    ap.add_argument("--start", default="boundary", choices=["boundary", "chained"],
                    help="boundary: every layer from its own start vector, all layers at once (the measured configuration); chained: "
                         "the reference's warm-start chain, radiative_transfer.cpp:247-253 (use with --batch M)")
    ap.add_argument("--sweep-grid", type=int, default=5, help="G: the sweep leg solves G^3 distinct grid points (0: skip the sweep leg)")
    ap.add_argument("--sweep-batch", type=int, default=1, help="grid points stacked into one solve per queue draw")
    ap.add_argument("--sweep-workers", type=int, default=3, help="models in flight per GPU in the sweep leg (host threads, one rt_model each)")
    ap.add_argument("--sweep-warm", type=int, default=0, help="1: run the whole sweep once untimed first")
    ap.add_argument("--cpu-full-slab", action="store_true", help="cpu_baseline solves the WHOLE slab once instead of a sample")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
