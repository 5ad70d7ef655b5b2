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


def all_gather_layers(local, L: int, world: int, rank: int):
    """local: torch tensor [l_rank, ...] on the backend's device.  Returns the [L, ...] tensor on every rank."""
    import torch
    import torch.distributed as dist
    sizes = [layer_range(L, world, r)[1] - layer_range(L, world, r)[0] for r in range(world)]
    assert local.shape[0] == sizes[rank]
    if world == 1:
        return local
    mx = max(sizes)
    pad = torch.zeros((mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: sizes[rank]] = local
    out = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(out, pad)
    return torch.cat([o[:s] for o, s in zip(out, sizes)], dim=0)


def solve_sharded(problem, solve_fn: Callable, world: int, rank: int, device=None):
    """Solves ONE slab with its layers partitioned over `world` ranks.

    solve_fn(sub_problem) -> (pop [l, N] float64 ndarray, status structured ndarray [l]) is the rank-local solver
    (radiative_transfer_b200.solver.Model(...).calc_molecular_populations on a GPU rank).  Returns the full
    (pop [L, N], status [L]) on every rank."""
    import torch
    from . import abi
    lo, hi = layer_range(problem.L, world, rank)
    if hi > lo:
        pop, st = solve_fn(problem.sliced(lo, hi))
    else:
        pop, st = np.zeros((0, problem.N)), np.zeros(0, dtype=abi.STATUS_DTYPE)
    if world == 1:
        return pop, st
    dev = device if device is not None else "cpu"
    pop_t = torch.from_numpy(np.ascontiguousarray(pop)).to(dev)
    # status as a float64 [l, 5] block so that one dtype travels
    st_t = torch.from_numpy(np.stack([st["iters"].astype(np.float64), st["converged"].astype(np.float64), st["eq_error"],
                                      st["pop_error"], st["rel_error"]], axis=1) if hi > lo else np.zeros((0, 5))).to(dev)
    pop_all = all_gather_layers(pop_t, problem.L, world, rank).cpu().numpy()
    st_all = all_gather_layers(st_t, problem.L, world, rank).cpu().numpy()
    out = np.zeros(problem.L, dtype=abi.STATUS_DTYPE)
    out["iters"] = st_all[:, 0].astype(np.int32); out["converged"] = st_all[:, 1].astype(np.int32)
    out["eq_error"], out["pop_error"], out["rel_error"] = st_all[:, 2], st_all[:, 3], st_all[:, 4]
    return pop_all, out


def sweep_assign(n_points: int, world: int, rank: int) -> List[int]:
    """Grid points of a parameter sweep (density x temperature x specific column density) owned by `rank`:
    round-robin, the static analogue of the reference's `omp for schedule(dynamic,1)` over shock models
    (radiative_transfer.cpp:152-159).  No collective: every model is independent."""
    return list(range(rank, n_points, world))
