"""The C++ drop-in adapter include/rt_b200_adapter.hpp, compiled INSIDE the reference's tree (oracle/Makefile ->
oracle/_ref/libref_adapter.so = unmodified reference objects + harness + adapter, linked against librt_b200.so).
CPU: the adapter reads the reference's objects back into exactly the flat problem they were built from.
GPU: reference objects -> adapter -> C ABI -> B200, against the reference's own calc_molecular_populations."""
import ctypes as C
import os

import numpy as np
import pytest

from conftest import rel_err

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "oracle", "_ref", "libref_adapter.so")


@pytest.fixture(scope="module")
def adapter():
    if not os.path.exists(SO):
        pytest.skip("oracle/_ref/libref_adapter.so not built (needs /root/reference at build time)")
    return C.CDLL(SO)


@pytest.mark.parametrize("name", ["generic", "oh", "oh_overlap", "ch3oh_small", "h2o_small"])
def test_flatten_round_trip(adapter, problems, name):
    p = problems(name, 3)
    d, h1 = p.desc()
    s, h2 = p.layer_soa()
    assert adapter.ref_adapter_roundtrip(C.byref(d), C.byref(s)) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("name,chained", [("ch3oh_small", True), ("oh_overlap", False), ("generic", True), ("h2o_small", False)])
def test_reference_objects_through_adapter_on_gpu(adapter, ref, problems, name, chained):
    from radiative_transfer_b200 import abi
    p = problems(name, 3)
    d, h1 = p.desc()
    s, h2 = p.layer_soa()
    pop = np.zeros((p.L, p.N))
    st = np.zeros(p.L, dtype=abi.STATUS_DTYPE)
    mode = abi.RT_START_CHAINED if chained else abi.RT_START_BOUNDARY
    rc = adapter.ref_adapter_calc_molecular_populations(C.byref(d), C.byref(s), abi.dptr(pop), int(p.accel), mode, 0,
                                                        st.ctypes.data_as(C.POINTER(abi.rt_layer_status)))
    assert rc == 0
    if chained:
        pop_o, st_o, _ = ref.calc_molecular_populations(p)   # the reference function itself
    else:
        pop_o, st_o, _ = ref.solve(p, retry=True, nthreads=4)  # independent boundary start + retry policy
    solid = (st_o["converged"] == 1) & (st_o["iters"] < p.max_iter - 20)
    assert solid.any()
    assert np.all(st["converged"][solid] == 1)
    assert np.all(np.abs(st["iters"] - st_o["iters"])[solid] <= 1)
    assert rel_err(pop[solid], pop_o[solid], 1e-30) < 1e-6
