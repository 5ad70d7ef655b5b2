"""CPU: analytic known answers derivable from the reference code, checked on the oracle."""
import copy

import numpy as np

from conftest import rel_err


def test_lu_solve_random_rate_like_matrices(port, ref):
    rng = np.random.default_rng(0)
    for n in (4, 17, 56, 150):
        A = rng.lognormal(0, 3, (n, n)) * np.where(rng.random((n, n)) < 0.4, 1.0, 1e-6)
        A -= np.diag(A.sum(axis=0))
        A[0, :] = 1.0
        b = np.zeros(n); b[0] = 1.0
        x_np = np.linalg.solve(A, b)
        for chk in (port, ref):
            x = chk.lu_solve(A, b)
            assert np.max(np.abs(A @ x - b)) < 1e-12
            assert rel_err(x, x_np, 1e-12) < 1e-8
        assert rel_err(port.lu_solve(A, b), ref.lu_solve(A, b), 0.0) == 0.0  # same algorithm, same bits


def test_no_radiation_gives_boltzmann(port, problems):
    """All A = 0, one partner table covering all levels with detailed balance -> LTE at T_n
    (coll_rates.cpp:193-195 vs spectroscopy.cpp:1245-1259)."""
    p = copy.copy(problems("generic", 3))
    p.einst = np.zeros_like(p.einst)
    pop, st, _ = port.solve(p, max_iter=50, accel=False)
    assert np.all(st["converged"] == 1)
    assert rel_err(pop, port.boltzmann(p), 1e-30) < 1e-9


def test_normalisation_and_column_sums(port, problems):
    for name in ("generic", "oh_overlap", "ch3oh_small"):
        p = problems(name, 3)
        b = port.boundary(p)
        assert np.allclose(b.sum(axis=1), 1.0, atol=1e-12)
        M, df = port.assemble(p, b)
        assert np.all(M[:, 0, :] == 1.0)  # iteration_lvg.cpp:149-151
        # every rate leaves its source column through the diagonal (iteration_lvg.cpp:127-144): below row 0 the
        # off-diagonal entries are gains (>= 0) and the diagonal is minus the column's total loss rate
        off = M[:, 1:, :].copy()
        for l in range(p.L):
            np.fill_diagonal(off[l, :, 1:], 0.0)
        assert np.all(off >= 0.0)
        diag = np.einsum("lii->li", M)[:, 1:]
        assert np.all(diag < 0.0)
        assert np.all(-diag >= off[:, :, 1:].sum(axis=1) * (1 - 1e-12))
        assert np.allclose(df[:, 0], 1.0 - b.sum(axis=1), atol=1e-14)
        new, eq = port.step(p, b)
        assert np.allclose(new.sum(axis=1), 1.0, atol=1e-10)


def test_esc_func_reproduces_nodes_and_clamps(port, problems):
    p = problems("generic", 2)
    gi = np.array([0, 5, len(p.lvg_gamma) - 2]); di = np.array([0, 7, len(p.lvg_delta) - 2])
    g, d = p.lvg_gamma[gi], p.lvg_delta[di]
    got = port.esc_func(p, g, d)
    assert np.allclose(got, p.lvg_p[di, gi], rtol=1e-14, atol=0)
    lo = port.esc_func(p, np.array([p.lvg_gamma[0] / 10]), np.array([p.lvg_delta[0] / 10]))
    hi = port.esc_func(p, np.array([p.lvg_gamma[-1] * 10]), np.array([p.lvg_delta[-1] * 10]))
    assert lo[0] == p.lvg_p[0, 0] and hi[0] == p.lvg_p[-1, -1]


def test_h2o_22ghz_gain_width(port, ref, problems):
    """calc_gain's oH2O 6_16 -> 5_23 special line width (transition_data.cpp:249-260): the compiled reference takes that
    branch when the two levels carry the quantum numbers it tests for (the harness sets them on the named line); the
    restatement names the line by level indices.  Only the gain side changes (gain, tau_eff), not the line profile."""
    import golden_cases as G
    p = problems("h2o_small", 4)
    assert p.mol_name == "oH2O"
    b = port.boundary(p)
    up, low = G.maser_lines(p, b)
    line = (int(up[0]), int(low[0]))
    plain = ref.maser_postpass(p, b, up, low)
    special = ref.maser_postpass(p, b, up, low, gain_width_line=line)
    mine = port.maser_postpass(p, b, up, low, gain_width_line=line)
    assert not np.allclose(special["gain"][0], plain["gain"][0], rtol=1e-3, atol=0)   # the branch was really taken
    assert np.array_equal(special["gain"][1:], plain["gain"][1:])                      # other lines untouched
    assert np.array_equal(special["tau_vs_freq"], plain["tau_vs_freq"])                # calc_line_profile unaffected
    for k in ("inv", "gain", "scalars", "tau_vs_aspect", "tau_vs_freq"):
        scale = np.max(np.abs(special[k])) + 1e-300
        assert np.max(np.abs(mine[k] - special[k])) / scale < 1e-12, k
    # and the setting does not leak into the next call
    again = port.maser_postpass(p, b, up, low)
    assert np.max(np.abs(again["gain"] - plain["gain"])) / (np.max(np.abs(plain["gain"])) + 1e-300) < 1e-12
