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


@pytest.mark.parametrize("name", ["generic", "generic_spline", "oh", "oh_overlap", "ch3oh_small", "h2o_small", "nh3", "h2co"])
def test_flatten_round_trip(adapter, problems, name):
    p = problems(name, 3)
    d, h1 = p.desc()
    s, h2 = p.layer_soa()
    assert adapter.ref_adapter_roundtrip(C.byref(d), C.byref(s)) == 0


@pytest.mark.gpu
@pytest.mark.parametrize("name,chained", [("ch3oh_small", True), ("oh_overlap", False), ("generic", True), ("generic_spline", False), ("h2o_small", False)])
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


@pytest.mark.gpu
def test_adapter_default_call_at_full_level_count(adapter):
    """The drop-in as a maintainer would call it -- calc_molecular_populations_gpu with the reference's OWN objects
    (energy_diagram, einstein_coeff, collisional_transitions, cloud_data, iteration_scheme_lvg built by the harness from the
    flat problem) and its DEFAULT start mode, the reference's warm-start chain -- on the headline molecule (N = 768, 40
    layers), against the committed answer of the reference's calc_molecular_populations (fixture ch3oh_chained): every layer,
    with the same rules as tests/test_gpu_full_slab.py."""
    import full_slab_cases as F
    import golden_cases as G
    from radiative_transfer_b200 import abi
    p = F.problem("ch3oh_chained")
    gold = F.load("ch3oh_chained")
    assert G.problem_digest(p) == bytes(gold["digest"]).decode()
    d, h1 = p.desc()
    s, h2 = p.layer_soa()
    pop = np.zeros((p.L, p.N))
    st = np.zeros(p.L, dtype=abi.STATUS_DTYPE)
    rc = adapter.ref_adapter_calc_molecular_populations(C.byref(d), C.byref(s), abi.dptr(pop), int(p.accel), abi.RT_START_CHAINED, 0,
                                                        st.ctypes.data_as(C.POINTER(abi.rt_layer_status)))
    assert rc == 0
    rep, fails = F.compare(p, pop, st, gold, F.ALLOW.get("ch3oh_chained", {}))
    assert not fails, fails[:10]
    assert rep["flags_equal"] == p.L


# ---- rt_b200::scheme_gpu: the per-call drop-in (a scheme object whose step runs on the GPU) ----------------------------
@pytest.mark.gpu
@pytest.mark.parametrize("name", ["ch3oh_small", "oh_overlap", "generic"])
def test_reference_controller_drives_gpu_scheme(adapter, ref, problems, name):
    """iteration_control<T> of the reference, instantiated on the GPU scheme type: every calc_new_pop is one rt_step."""
    from radiative_transfer_b200 import abi
    p = problems(name, 3)
    d, h1 = p.desc()
    s, h2 = p.layer_soa()
    pop = np.zeros((p.L, p.N))
    st = np.zeros(p.L, dtype=abi.STATUS_DTYPE)
    max_iter = min(p.max_iter, 400)
    adapter.ref_adapter_scheme_solve.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_double, C.c_int,
                                                 C.c_void_p, C.c_int, C.c_int, C.c_int]
    rc = adapter.ref_adapter_scheme_solve(C.addressof(d), C.addressof(s), pop.ctypes.data, abi.RT_START_BOUNDARY, max_iter, 1e-5,
                                          int(p.accel), st.ctypes.data, 0, -1, 0)
    assert rc == 0
    pop_o, st_o, _ = ref.solve(p, start_mode=abi.RT_START_BOUNDARY, max_iter=max_iter, retry=False)
    solid = (st_o["converged"] == 1) & (st_o["iters"] < max_iter - 20)
    assert solid.any()
    assert np.all(st["converged"][solid] == 1)
    assert np.all(np.abs(st["iters"] - st_o["iters"])[solid] <= 1)
    assert rel_err(pop[solid], pop_o[solid], 1e-30) < 1e-6


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["ch3oh_small", "oh_overlap", "h2o_small"])
def test_gpu_scheme_virtuals_through_base_pointer(adapter, ref, problems, name):
    """operator(), calc_new_pop (base LU over the GPU-assembled matrix) and intensity_calc via iteration_scheme_lvg*."""
    p = problems(name, 3)
    d, h1 = p.desc()
    s, h2 = p.layer_soa()
    lay = 1
    pop = ref.boundary(p)
    N = p.N
    iu, il = np.nonzero(np.tril(p.einst, -1) > 0)
    sel = np.linspace(0, iu.shape[0] - 1, min(40, iu.shape[0])).astype(int)
    up = np.ascontiguousarray(iu[sel], dtype=np.int32); low = np.ascontiguousarray(il[sel], dtype=np.int32)
    M = np.zeros((N, N)); df = np.zeros(N); new = np.zeros(N); eq = C.c_double(); J = np.zeros(len(up))
    adapter.ref_adapter_scheme_probe.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p,
                                                 C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    rc = adapter.ref_adapter_scheme_probe(C.addressof(d), C.addressof(s), lay, pop[lay].ctypes.data, M.ctypes.data, df.ctypes.data,
                                          new.ctypes.data, C.addressof(eq), len(up), up.ctypes.data, low.ctypes.data, J.ctypes.data, 0)
    assert rc == 0
    M_o, df_o = ref.assemble(p, pop)
    new_o, eq_o = ref.step(p, pop)
    J_o = ref.intensity(p, lay, up, low, pop[lay])
    scale = np.abs(M_o[lay]).max()
    assert np.abs(M - M_o[lay]).max() <= 1e-12 * scale
    assert np.abs(df - df_o[lay]).max() <= 1e-12 * max(np.abs(df_o[lay]).max(), 1e-300) + 1e-13 * scale
    assert rel_err(new, new_o[lay], 1e-30) < 1e-9
    assert abs(eq.value - eq_o[lay]) <= 1e-9 * abs(eq_o[lay]) + 1e-13 * scale
    assert np.allclose(J, J_o, rtol=1e-10, atol=1e-300)


@pytest.mark.gpu
@pytest.mark.parametrize("case", ["ch3oh_small", "oh_overlap", "h2o_small"])
def test_lim_luminosity_through_adapter(adapter, case):
    """lim_luminosity_lvg_gpu fills the reference's transition_data_container; its own calc_saturation_depth follows."""
    import golden_cases as G
    p = G.problem(case)
    gold = G.load(case)
    d, h1 = p.desc()
    s, h2 = p.layer_soa()
    b = np.ascontiguousarray(gold["boundary"])
    up = np.ascontiguousarray(gold["post_up"], dtype=np.int32); low = np.ascontiguousarray(gold["post_low"], dtype=np.int32)
    T = len(up)
    lum_arr = np.zeros((T, p.L)); loss = np.zeros((T, p.L)); pump = np.zeros((T, p.L)); lum = np.zeros(T); tau = np.zeros(T)
    adapter.ref_adapter_lim_luminosity.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p, C.c_void_p, C.c_double,
                                                   C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int]
    rc = adapter.ref_adapter_lim_luminosity(C.addressof(d), C.addressof(s), b.ctypes.data, T, up.ctypes.data, low.ctypes.data, 1.0,
                                            lum_arr.ctypes.data, loss.ctypes.data, pump.ctypes.data, lum.ctypes.data, tau.ctypes.data, 0)
    assert rc == 0
    for got, key in ((lum_arr, "lim_lum_arr"), (loss, "lim_loss_rate_arr"), (pump, "lim_pump_rate_arr"), (lum, "lim_lum"),
                     (tau, "lim_tau_sat")):
        assert np.allclose(got, gold[key], rtol=1e-9, atol=0.0, equal_nan=True), key
