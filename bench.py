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
`--mode shard` instead partitions ONE slab's layers across ranks (strong scaling; one all_gather of the
populations at the end).
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
    idx = (np.linspace(0, L - 1, n, endpoint=True) + which * max(1, L // (4 * n))).astype(int) % L
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
        v, sec, kind, info = cpu_solve_sample(p, cores, cores, which=s, retry_cap=args.retry_cap)
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
                         "sample": f"each step: {info['layers']} evenly spaced layers of the {p.L}-layer slab solved under the reference's policy (retry of failed layers included) on "
                                   f"{cores} threads (layers are independent); models/s = layers/s / {p.L}"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(json.dumps(out))
    return 0


# ----------------------------------------------------------------------------------------------------------------
# this repo's arm
# ----------------------------------------------------------------------------------------------------------------
def run_gpu(args):
    import torch
    from radiative_transfer_b200 import abi, solver

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

    p_full, name = make_problem(args, rank)
    p = p_full
    if args.mode == "shard" and world > 1:
        from radiative_transfer_b200 import sharding
        lo, hi = sharding.layer_range(p_full.L, world, rank)
        p = p_full.sliced(lo, hi)
    if args.batch > 1:
        # sweep throughput mode: --batch M models per solve as ONE cloud of M*L independent layers (sharding.stack_clouds);
        # the headline `value` is quoted at the default M = 1, the slab of BASELINE configs[1]
        from radiative_transfer_b200 import sharding
        p, _ = sharding.stack_clouds([p_full] * args.batch)
        name += f" x {args.batch} models per solve"
    N, L = p.N, p.L
    m = solver.Model(p, device=local, upload_cloud=False)
    lay_pinned, keep = pinned_layers(p)
    p.layers = lay_pinned
    m.cloud_upload(p)
    pop_host = torch.empty((L, N), dtype=torch.float64).pin_memory()
    pop_np = pop_host.numpy()

    def barrier():
        if dist is not None:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        m.cloud_rebuild()
        m.solve_device(start_mode=abi.RT_START_BOUNDARY, max_iter=p.max_iter, accel=p.accel)

    # ---- value leg: device-resident ------------------------------------------------------------------------------
    for _ in range(args.warmup):
        step_device()
    barrier()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    launches0 = solver.launch_count()
    m.profile_reset(timing=False)
    m.timer_begin()
    for _ in range(args.steps):
        step_device()
    ms_total = m.timer_end()
    barrier()
    clocks = sampler.stop() if rank == 0 else None
    launches = solver.launch_count() - launches0
    iters_total = m.iterations_enqueued()
    st = m.get_status()
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
    models_per_step = n_gpus * args.batch if args.mode == "sweep" else 1
    value = models_per_step / (ms_per_step * 1e-3)

    # ---- e2e leg: host buffers through the C ABI ---------------------------------------------------------------------
    def step_e2e():
        m.cloud_upload(p)  # pinned host -> device + collisional matrices
        stat = np.zeros(L, dtype=abi.STATUS_DTYPE)
        import ctypes as C
        m._ck(m.L_.rt_calc_molecular_populations(m.h, abi.dptr(pop_np), abi.RT_START_BOUNDARY, int(p.accel), 150, int(args.retry_cap),
                                                 C.c_double(1e-5), stat.ctypes.data_as(C.POINTER(abi.rt_layer_status))),
              "rt_calc_molecular_populations")
        return stat

    step_e2e()
    barrier()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        stat = step_e2e()
    if args.mode == "shard" and dist is not None:
        # one all_gather of the populations per solve
        mine = pop_host.cuda(non_blocking=True)
        from radiative_transfer_b200 import sharding
        full = sharding.all_gather_layers(mine, p_full.L, world, rank)
        torch.cuda.synchronize()
    e2e_s = time.perf_counter() - t0
    t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
    if dist is not None:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_value = models_per_step / (float(t.item()) / args.steps)
    h2d = sum(v.nbytes for k, v in p.layers.items() if k != "height")
    d2h = pop_np.nbytes + stat.nbytes

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
    if os.path.exists(tf) and args.config == "ch3oh" and args.layers in (None, 0, 200) and args.batch == 1:  # captured on this workload only
        traffic = json.load(open(tf))
    top = max(kern, key=lambda k: kern[k]["ms_per_step"])

    def tbytes(k):
        t = traffic.get(k)
        return t.get("bytes") if isinstance(t, dict) else t

    def roof(k):
        v, pv = kern[k], prof[k]
        ai = pv["flops"] / max(pv["bytes"], 1.0)
        if ai > fp64_peak * 1e3 / hbm_peak:  # flop/byte above the machine balance -> FP64 pipe bound
            return {"kernel": k, "bound": "tensor", "achieved": v["tflops"], "peak": fp64_peak, "unit": "TFLOP/s",
                    "frac": v["tflops"] / fp64_peak, "traffic": tbytes(k), "traffic_detail": traffic.get(k),
                    "peak_source": f"FP64 measured in this run by rt_device_fp64_peak: DFMA {dfma:.1f}, DMMA {dmma:.1f} TFLOP/s "
                                   "(MEASURED_PEAKS.json has no FP64 figure; the LU runs on the FP64 pipe, not bf16 tcgen05)",
                    "share_of_step": v["share"]}
        return {"kernel": k, "bound": "hbm", "achieved": v["gbs"], "peak": hbm_peak, "unit": "GB/s", "frac": v["gbs"] / hbm_peak,
                "traffic": tbytes(k), "traffic_detail": traffic.get(k), "peak_source": hbm_src, "share_of_step": v["share"]}

    out = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": n_gpus, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": ms_per_step, "higher_is_better": True, "scaling": "weak" if args.mode == "sweep" else "strong",
        "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": name, "start": "boundary", "accel": bool(p.accel), "max_iter": p.max_iter, "min_err": 1e-5,
                   "retry_cap": args.retry_cap,
                   "mode": args.mode, "models_per_solve": args.batch, "l2": "per-step working set (C and M matrices, 2 x %.0f MB) exceeds the 126 MB L2"
                   % (L * ((N + 31) // 32 * 32) ** 2 * 8 / 1e6),
                   "timing": "CUDA events on the library's stream around K steps; max over ranks; per-kernel times from a second pass "
                             "of the same K steps with an event pair around every kernel class"},
        "ms_per_step_profiled": ms_profiled / args.steps,
        "ms_per_iteration": ms_total / max(1, iters_total), "iterations_per_step": iters_total / args.steps,
        "layers_converged": int(st["converged"].sum()), "layers": L, "iters_mean": float(st["iters"].mean()),
        "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": int(h2d), "d2h_bytes_per_step": int(d2h),
                "timing": "host wall clock around rt_cloud_upload + rt_calc_molecular_populations (synchronous calls)"},
        "gpu_launches": int(launches),
        "clocks": clocks,
        "roofline": roof(top),
        "roofline_coupling": roof("assemble") if "assemble" in kern else None,
        "kernels": kern,
        "fp64_peak_tflops": {"dfma": dfma, "dmma": dmma},
    }
    if n_gpus == 1 and not args.no_cpu_baseline:
        cores = os.cpu_count() or 1
        v, sec, kind, info = cpu_solve_sample(p_full, cores, cores, retry_cap=args.retry_cap)
        out["cpu_baseline"] = {"value": v, "unit": UNIT, "cores": cores, "kind": kind,
                               "sample": f"{info['layers']} evenly spaced layers of the slab solved under the reference's policy (retry of failed layers included) on {cores} threads in "
                                         f"{sec:.1f} s (mean {info['iters_mean']:.0f} iterations); models/s = layers/s / {p_full.L}"}
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
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    if args.impl == "reference":
        return run_reference(args)
    return run_gpu(args)


if __name__ == "__main__":
    sys.exit(main())
