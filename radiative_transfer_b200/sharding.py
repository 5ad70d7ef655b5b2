"""Multi-GPU host logic: one process per GPU (torch.distributed; NCCL on the GPUs, gloo in the CPU tests).

The reference solves every layer independently: a layer's mean intensities depend only on
its own populations, so a slab shards by LAYERS with no per-iteration exchange, and a parameter-grid sweep shards
model-per-rank with no collective at all.  The only communication is one all_gather of the
converged populations (+ per-layer status) at the end of a sharded solve.

    layer_range        contiguous block partition of L layers over `world` ranks
    all_gather_layers  ragged all_gather of [l_r, N] blocks into the full [L, N] array (every rank gets it)
    solve_sharded      the N>1 analogue of calc_molecular_populations for ONE slab
    sweep_assign       static round-robin assignment of grid points to ranks (model-per-GPU)
"""
from __future__ import annotations

from typing import Callable, List, Tuple

import numpy as np


def layer_range(L: int, world: int, rank: int) -> Tuple[int, int]:
    """Contiguous blocks, sizes differing by at most one, larger blocks first (so that chained-start mode would
    only need a one-vector hand-off at block boundaries)."""
    base, rem = divmod(L, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def layer_indices(L: int, world: int, rank: int, cyclic: bool = False) -> np.ndarray:
    """Layers owned by `rank`.  Contiguous blocks (layer_range) keep the warm-start chain of the reference inside a rank
    (radiative_transfer.cpp:247-249).  CYCLIC (layer l -> rank l mod world) is for the independent-start mode: iteration
    counts vary smoothly along a slab (CH3OH: 16 -> 66), so a contiguous split leaves the rank that owns the slow end
    working while the others idle; dealing the layers out round-robin gives every rank the same mix."""
    if cyclic:
        return np.arange(rank, L, world)
    lo, hi = layer_range(L, world, rank)
    return np.arange(lo, hi)


def all_gather_layers(local, L: int, world: int, rank: int, cyclic: bool = False):
    """local: torch tensor [l_rank, ...] on the backend's device (rows in the order of layer_indices).  Returns the
    [L, ...] tensor in slab order on every rank: ONE all_gather of equal-size (padded) blocks."""
    import torch
    import torch.distributed as dist
    idx = [layer_indices(L, world, r, cyclic) for r in range(world)]
    sizes = [len(i) for i in idx]
    assert local.shape[0] == sizes[rank]
    if world == 1:
        return local
    mx = max(sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: sizes[rank]] = local
    out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, pad)
    out = out.view((world, mx) + tuple(local.shape[1:]))
    full = torch.empty((L,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    for r in range(world):
        if sizes[r]:
            full[torch.as_tensor(idx[r], device=local.device)] = out[r, : sizes[r]]
    return full


def solve_sharded(problem, solve_fn: Callable, world: int, rank: int, device=None, cyclic: bool = False):
    """Solves ONE slab with its layers partitioned over `world` ranks.

    solve_fn(sub_problem) -> (pop [l, N] float64 ndarray, status structured ndarray [l]) is the rank-local solver
    (radiative_transfer_b200.solver.Model(...).calc_molecular_populations on a GPU rank).  Returns the full
    (pop [L, N], status [L]) on every rank.  cyclic: see layer_indices (independent-start modes only)."""
    import torch
    from . import abi
    mine = layer_indices(problem.L, world, rank, cyclic)
    lo, hi = 0, len(mine)
    if hi > lo:
        pop, st = solve_fn(problem.take(mine))
    else:
        pop, st = np.zeros((0, problem.N)), np.zeros(0, dtype=abi.STATUS_DTYPE)
    if world == 1:
        return pop, st
    dev = device if device is not None else "cpu"
    pop_t = torch.from_numpy(np.ascontiguousarray(pop)).to(dev)
    # status as a float64 [l, 5] block so that one dtype travels
    st_t = torch.from_numpy(np.stack([st["iters"].astype(np.float64), st["converged"].astype(np.float64), st["eq_error"],
                                      st["pop_error"], st["rel_error"]], axis=1) if hi > lo else np.zeros((0, 5))).to(dev)
    pop_all = all_gather_layers(pop_t, problem.L, world, rank, cyclic).cpu().numpy()
    st_all = all_gather_layers(st_t, problem.L, world, rank, cyclic).cpu().numpy()
    out = np.zeros(problem.L, dtype=abi.STATUS_DTYPE)
    out["iters"] = st_all[:, 0].astype(np.int32); out["converged"] = st_all[:, 1].astype(np.int32)
    out["eq_error"], out["pop_error"], out["rel_error"] = st_all[:, 2], st_all[:, 3], st_all[:, 4]
    return pop_all, out


def sweep_assign(n_points: int, world: int, rank: int) -> List[int]:
    """Grid points of a parameter sweep (density x temperature x specific column density) owned by `rank`:
    round-robin, the static analogue of the reference's `omp for schedule(dynamic,1)` over shock models
    (radiative_transfer.cpp:152-159).  No collective: every model is independent."""
    return list(range(rank, n_points, world))


class SweepQueue:
    """Dynamic hand-out of grid points to ranks: the direct analogue of the reference's
    `#pragma omp parallel for schedule(dynamic,1)` over shock models (radiative_transfer.cpp:152-159).  Iteration counts
    differ from model to model, so a static split leaves ranks idle at the end of a sweep; here every rank draws the next
    unclaimed index from a shared counter.  The counter lives in the process group's key-value store (the rendezvous
    TCPStore of torch.distributed: `add` is atomic), so no collective and no GPU traffic is involved.

        for i in SweepQueue(n_points):        # on every rank
            solve(grid_point(i))
    """

    _seq = {}  # sweeps constructed so far in this process, per name

    def __init__(self, n_points: int, name: str = "rt_sweep", store=None):
        import torch.distributed as dist
        self.n = int(n_points)
        # One counter per SWEEP: the k-th queue of a given name gets its own store key.  Every rank runs the same program
        # (SPMD) and so constructs its queues in the same order; the k-th queue of every rank therefore shares one key,
        # and a second sweep in the same process group starts from zero instead of from the first sweep's final count.
        k = SweepQueue._seq.get(name, 0)
        SweepQueue._seq[name] = k + 1
        self.key = f"{name}/{k}/next"
        self._local = 0
        self.store = store
        if store is None and dist.is_available() and dist.is_initialized():
            self.store = dist.distributed_c10d._get_default_store()

    def __iter__(self):
        return self

    def __next__(self) -> int:
        if self.store is None:  # single process
            i, self._local = self._local, self._local + 1
        else:
            i = int(self.store.add(self.key, 1)) - 1
        if i >= self.n:
            raise StopIteration
        return i


def stack_clouds(problems):
    """Several models of the SAME molecule as one cloud: layers are independent, so the layers of
    M slabs concatenated are one batch of M*L layers for the solver -- the way to fill a GPU whose 148 SMs a single
    200-layer slab leaves partly idle.  Returns (stacked problem, [(lo, hi) layer range of each model])."""
    import copy
    base = copy.copy(problems[0])
    keys = list(base.layers.keys())
    lay, ranges, lo = {}, [], 0
    for k in keys:
        v0 = base.layers[k]
        if np.ndim(v0) == 0:
            lay[k] = v0
        else:
            lay[k] = np.concatenate([np.asarray(q.layers[k]) for q in problems], axis=0)
    for q in problems:
        ranges.append((lo, lo + q.L))
        lo += q.L
    base.layers = lay
    base.model_first = np.array([r[0] for r in ranges] + [lo], dtype=np.int32)  # rt_layer_soa::model_first
    return base, ranges
