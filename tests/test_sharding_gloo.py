"""CPU, world_size 2, gloo: the N>1 host path (layer partition + ragged all_gather of populations / status).
The rank-local solver is the oracle here (no GPU in this container); on the GPU box the same code runs over NCCL with
radiative_transfer_b200.solver.Model as the rank-local solver (bench.py --mode shard, tests/test_gpu_parity.py)."""
import os
import socket

import numpy as np
import pytest


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, L, q, cyclic=False):
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from radiative_transfer_b200 import sharding, synthetic
    from oracle.loader import Checker
    chk = Checker("port")
    p = synthetic.small("ch3oh_small", L=L)

    def solve_fn(sub):
        pop, st, _ = chk.solve(sub)
        return pop, st
    pop, st = sharding.solve_sharded(p, solve_fn, world, rank, cyclic=cyclic)
    if rank == 0:
        q.put((pop, st["iters"].copy(), st["converged"].copy()))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("L,cyclic", [(5, False), (2, False), (5, True), (1, True)])
def test_sharded_equals_unsharded(L, cyclic, port):
    import torch.multiprocessing as mp
    from radiative_transfer_b200 import synthetic
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world, tcp = 2, _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, tcp, L, q, cyclic)) for r in range(world)]
    for pr in procs:
        pr.start()
    pop, iters, conv = q.get(timeout=180)
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    p = synthetic.small("ch3oh_small", L=L)
    pop1, st1, _ = port.solve(p)
    assert np.array_equal(pop, pop1)  # layers are independent: bitwise
    assert np.array_equal(iters, st1["iters"]) and np.array_equal(conv, st1["converged"])


def test_layer_range_and_sweep_assign():
    from radiative_transfer_b200 import sharding
    for L in (1, 7, 200, 400):
        for w in (1, 2, 3, 8):
            r = [sharding.layer_range(L, w, k) for k in range(w)]
            assert r[0][0] == 0 and r[-1][1] == L
            assert all(r[k][1] == r[k + 1][0] for k in range(w - 1))
            sizes = [b - a for a, b in r]
            assert max(sizes) - min(sizes) <= 1
    pts = sorted(sum((sharding.sweep_assign(512, 8, k) for k in range(8)), []))
    assert pts == list(range(512))
    for L in (1, 7, 400):
        for w in (1, 3, 8):
            for cyc in (False, True):
                got = np.sort(np.concatenate([sharding.layer_indices(L, w, k, cyc) for k in range(w)]))
                assert np.array_equal(got, np.arange(L))
    assert sharding.layer_indices(10, 4, 1, True).tolist() == [1, 5, 9]


def _queue_worker(rank, world, port, n, q):
    import sys, time
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    sys.path.insert(0, root)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    import torch.distributed as dist
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from radiative_transfer_b200 import sharding
    mine = []
    for i in sharding.SweepQueue(n):
        mine.append(i)
        time.sleep(0.002 * (1 + 3 * rank))  # rank 1 is the slow one: it must end up with fewer points
    # a SECOND sweep in the same process group starts from zero again (own counter per sweep)
    second = [i for i in sharding.SweepQueue(n // 2)]
    q.put((rank, mine, second))
    dist.barrier()
    dist.destroy_process_group()


def test_sweep_queue_hands_every_point_out_once():
    """Dynamic model-per-rank scheduling (the analogue of `omp for schedule(dynamic,1)`, radiative_transfer.cpp:152-159)."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    world, tcp, n = 2, _free_port(), 40
    procs = [ctx.Process(target=_queue_worker, args=(r, world, tcp, n, q)) for r in range(world)]
    for pr in procs:
        pr.start()
    res = [q.get(timeout=120) for _ in range(world)]
    got = {r: a for r, a, b in res}
    second = {r: b for r, a, b in res}
    for pr in procs:
        pr.join(timeout=60)
        assert pr.exitcode == 0
    assert sorted(got[0] + got[1]) == list(range(n))
    assert len(got[0]) > len(got[1]) > 0
    assert sorted(second[0] + second[1]) == list(range(n // 2))


def test_sweep_queue_single_process_and_stack_clouds():
    from radiative_transfer_b200 import sharding, synthetic
    assert list(sharding.SweepQueue(5)) == [0, 1, 2, 3, 4]
    assert list(sharding.SweepQueue(3)) == [0, 1, 2]
    a = synthetic.small("ch3oh_small", L=3)
    b = synthetic.config_ch3oh(L=2, n_vt=2, n_k=6, n_j=8, cloud_kw=dict(t_hi=200.0))
    s, ranges = sharding.stack_clouds([a, b])
    assert s.L == 5 and ranges == [(0, 3), (3, 5)]
    assert np.array_equal(s.layers["temp_n"][:3], a.layers["temp_n"]) and np.array_equal(s.layers["temp_n"][3:], b.layers["temp_n"])
    assert s.layers["dust_conc"].shape == (5, a.layers["dust_conc"].shape[1])
