"""ctypes loaders for the two CHECKERS.  Test infrastructure: import only from tests/, bench.py's cpu_baseline /
--impl reference legs and __graft_entry__.smoke() -- never from the product package.

  ref()  -> oracle/_ref/libref_oracle.so   the unmodified reference sources + oracle/ref_harness.cpp
  port() -> oracle/_build/librt_oracle.so  the plain-C restatement oracle/rt_oracle.c
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF_SO = os.path.join(HERE, "_ref", "libref_oracle.so")
PORT_SO = os.path.join(HERE, "_build", "librt_oracle.so")
_cache = {}


def build(which: str = "all") -> None:
    subprocess.run(["make", "-C", HERE, "-j8", which], check=True, stdout=subprocess.DEVNULL)


def ref_available() -> bool:
    return os.path.exists(REF_SO)


def ref():
    if "ref" not in _cache:
        if not os.path.exists(REF_SO):
            raise FileNotFoundError(f"{REF_SO}: build it with `make -C oracle ref` where /root/reference exists")
        _cache["ref"] = C.CDLL(REF_SO)
    return _cache["ref"]


def port():
    if "port" not in _cache:
        if not os.path.exists(PORT_SO):
            build("port")
        _cache["port"] = C.CDLL(PORT_SO)
    return _cache["port"]


class Checker:
    """Uniform Python face of either checker library (entry points ref_* / ora_* have identical signatures)."""

    def __init__(self, kind: str):
        from radiative_transfer_b200 import abi
        self.abi = abi
        self.kind = kind
        self.lib = ref() if kind == "ref" else port()
        self.pfx = "ref_" if kind == "ref" else "ora_"

    def _f(self, name):
        return getattr(self.lib, self.pfx + name)

    def _views(self, p):
        d, h1 = p.desc()
        s, h2 = p.layer_soa()
        return d, s, (h1, h2)

    def set_lu_variant(self, v: int):
        """Rounding-sensitivity probe of the restated lu_matrix_solve (oracle/shim/linear_algebra.h; ref only); 0 = off."""
        assert self.kind == "ref"
        self.lib.ref_set_lu_variant(int(v))

    def lu_solve(self, a, b):
        a = np.array(a, dtype=np.float64, order="C"); b = np.array(b, dtype=np.float64)
        self._f("lu_solve")(self.abi.dptr(a), self.abi.dptr(b), C.c_int(b.shape[0]))
        return b

    def esc_func(self, p, gamma, delta):
        d, s, hold = self._views(p)
        gamma = np.ascontiguousarray(gamma, dtype=np.float64); delta = np.ascontiguousarray(delta, dtype=np.float64)
        out = np.empty_like(gamma)
        rc = self._f("esc_func")(C.byref(d), gamma.shape[0], self.abi.dptr(gamma), self.abi.dptr(delta), self.abi.dptr(out))
        assert rc == 0
        return out

    def overlap_esc_func(self, p, which, gamma, delta, gratio, dx):
        d, s, hold = self._views(p)
        arrs = [np.ascontiguousarray(x, dtype=np.float64) for x in (gamma, delta, gratio, dx)]
        out = np.empty_like(arrs[0])
        rc = self._f("overlap_esc_func")(C.byref(d), which, arrs[0].shape[0], *[self.abi.dptr(x) for x in arrs], self.abi.dptr(out))
        assert rc == 0
        return out

    def dust_absorption(self, p, lay, energy):
        d, s, hold = self._views(p)
        energy = np.ascontiguousarray(energy, dtype=np.float64); out = np.empty_like(energy)
        rc = self._f("dust_absorption")(C.byref(d), C.byref(s), lay, energy.shape[0], self.abi.dptr(energy), self.abi.dptr(out))
        assert rc == 0, "dust wavelength round trip not exact" if rc == 2 else rc
        return out

    def line_groups(self, p):
        d, s, hold = self._views(p)
        g = np.zeros((4 * p.N * p.N // 4 + 8, 5), dtype=np.int32); n = C.c_int()
        rc = self._f("line_groups")(C.byref(d), g.shape[0], self.abi.iptr(g), C.byref(n))
        assert rc == 0
        return g[: n.value].copy()

    def collision_rates(self, p):
        d, s, hold = self._views(p)
        out = np.zeros((p.L, p.N, p.N))
        assert self._f("collision_rates")(C.byref(d), C.byref(s), self.abi.dptr(out)) == 0
        return out

    def assemble(self, p, pop):
        d, s, hold = self._views(p)
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        M = np.zeros((p.L, p.N, p.N)); df = np.zeros((p.L, p.N))
        assert self._f("assemble")(C.byref(d), C.byref(s), self.abi.dptr(pop), self.abi.dptr(M), self.abi.dptr(df)) == 0
        return M, df

    def step(self, p, pop):
        d, s, hold = self._views(p)
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        new = np.zeros((p.L, p.N)); eq = np.zeros(p.L)
        assert self._f("step")(C.byref(d), C.byref(s), self.abi.dptr(pop), self.abi.dptr(new), self.abi.dptr(eq)) == 0
        return new, eq

    def intensity(self, p, lay, up, low, pop):
        d, s, hold = self._views(p)
        up = np.ascontiguousarray(up, dtype=np.int32); low = np.ascontiguousarray(low, dtype=np.int32)
        pop = np.ascontiguousarray(pop, dtype=np.float64); out = np.zeros(up.shape[0])
        assert self._f("intensity")(C.byref(d), C.byref(s), lay, up.shape[0], self.abi.iptr(up), self.abi.iptr(low),
                                    self.abi.dptr(pop), self.abi.dptr(out)) == 0
        return out

    def boundary(self, p):
        d, s, hold = self._views(p)
        pop = np.zeros((p.L, p.N))
        assert self._f("boundary")(C.byref(d), C.byref(s), self.abi.dptr(pop)) == 0
        return pop

    def boltzmann(self, p):
        d, s, hold = self._views(p)
        pop = np.zeros((p.L, p.N))
        assert self._f("boltzmann")(C.byref(d), C.byref(s), self.abi.dptr(pop)) == 0
        return pop

    def solve(self, p, start_mode=None, max_iter=None, min_err=1e-5, accel=None, retry=False, max_iter_ext=15000,
              pop0=None, lay=(0, -1), nthreads=1):
        """Returns (pop [L,N], status recarray [L], seconds)."""
        d, s, hold = self._views(p)
        start_mode = p.start_mode if start_mode is None else start_mode
        max_iter = p.max_iter if max_iter is None else max_iter
        accel = p.accel if accel is None else accel
        pop = np.zeros((p.L, p.N)) if pop0 is None else np.array(pop0, dtype=np.float64, order="C")
        st = np.zeros(p.L, dtype=self.abi.STATUS_DTYPE); sec = C.c_double()
        rc = self._f("solve")(C.byref(d), C.byref(s), self.abi.dptr(pop), start_mode, max_iter, C.c_double(min_err),
                              int(accel), int(retry), max_iter_ext, st.ctypes.data_as(C.POINTER(self.abi.rt_layer_status)),
                              lay[0], lay[1], nthreads, C.byref(sec))
        assert rc == 0, rc
        return pop, st, sec.value

    def maser_postpass(self, p, pop, up, low, nb_freq=300, nb_aspect=37, delta_aspect=0.25, velocity_shift=5.e5,
                       gain_width_line=None):
        """transition_data_container::calc_inv / calc_gain / calc_line_profile.  The reference fixes
        nb_freq=300, nb_aspect=37, delta_aspect=0.25, velocity_shift=5e5 (transition_data.cpp:16-18,191).
        gain_width_line = (up, low): that line gets the oH2O 22 GHz special width of calc_gain (:249-260; the molecule
        must be named "oH2O" for the compiled reference to take that branch)."""
        gw = (-1, -1) if gain_width_line is None else (int(gain_width_line[0]), int(gain_width_line[1]))
        if self.kind == "ref":
            self.lib.ref_set_gain_width_line(gw[0], gw[1])
        else:
            self.lib.ora_set_gain_width_line(gw[0], gw[1], C.c_double(5.e4))
        try:
            return self._maser_postpass(p, pop, up, low, nb_freq, nb_aspect, delta_aspect, velocity_shift)
        finally:
            if self.kind == "ref":
                self.lib.ref_set_gain_width_line(-1, -1)
            else:
                self.lib.ora_set_gain_width_line(-1, -1, C.c_double(0.))

    def _maser_postpass(self, p, pop, up, low, nb_freq, nb_aspect, delta_aspect, velocity_shift):
        d, s, hold = self._views(p)
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        up = np.ascontiguousarray(up, dtype=np.int32); low = np.ascontiguousarray(low, dtype=np.int32)
        T = up.shape[0]
        out = dict(inv=np.zeros((T, p.L)), gain=np.zeros((T, p.L)), scalars=np.zeros((T, 5)),
                   tau_vs_aspect=np.zeros((T, nb_aspect)), tau_vs_freq=np.zeros((T, nb_freq)))
        tail = [self.abi.dptr(out[k]) for k in ("inv", "gain", "scalars", "tau_vs_aspect", "tau_vs_freq")]
        if self.kind == "ref":
            assert (nb_freq, nb_aspect, delta_aspect, velocity_shift) == (300, 37, 0.25, 5.e5)
            rc = self.lib.ref_maser_postpass(C.byref(d), C.byref(s), self.abi.dptr(pop), T, self.abi.iptr(up),
                                             self.abi.iptr(low), *tail)
        else:
            rc = self.lib.ora_maser_postpass(C.byref(d), C.byref(s), self.abi.dptr(pop), T, self.abi.iptr(up),
                                             self.abi.iptr(low), nb_freq, nb_aspect, C.c_double(delta_aspect),
                                             C.c_double(velocity_shift), *tail)
        assert rc == 0
        return out

    LIM_KEYS = ("lum_arr", "pump_eff_arr", "loss_rate_arr", "pump_rate_arr", "emiss_coeff_arr", "exc_temp_arr")

    def lim_luminosity(self, p, pop, up, low, beaming_factor=1.0, lay_nb_hg=None, layer0_intensity=True):
        """lim_luminosity_lvg (maser_luminosity.cpp:7-106) + calc_exc_temp + calc_saturation_depth for listed transitions.
        ref: the reference functions themselves (lay_nb_hg comes out of its own calc_gain; layer0_intensity is what the
        reference does).  port: lay_nb_hg must be given (maser_postpass scalars[:, 4]) for the saturation part."""
        d, s, hold = self._views(p)
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        up = np.ascontiguousarray(up, dtype=np.int32); low = np.ascontiguousarray(low, dtype=np.int32)
        T = up.shape[0]
        out = {k: np.zeros((T, p.L)) for k in self.LIM_KEYS}
        out.update(lum=np.zeros(T), tau_sat=np.zeros(T), temp_sat=np.zeros(T), lay_nb_hg=np.zeros(T, dtype=np.int32))
        arrs = [self.abi.dptr(out[k]) for k in self.LIM_KEYS]
        if self.kind == "ref":
            assert layer0_intensity
            rc = self.lib.ref_lim_luminosity(C.byref(d), C.byref(s), self.abi.dptr(pop), T, self.abi.iptr(up), self.abi.iptr(low),
                                             C.c_double(beaming_factor), *arrs, self.abi.dptr(out["lum"]),
                                             self.abi.iptr(out["lay_nb_hg"]), self.abi.dptr(out["tau_sat"]),
                                             self.abi.dptr(out["temp_sat"]))
        else:
            if lay_nb_hg is None:
                lay_nb_hg = self.maser_postpass(p, pop, up, low)["scalars"][:, 4]
            out["lay_nb_hg"][:] = np.asarray(lay_nb_hg, dtype=np.int32)
            rc = self.lib.ora_lim_luminosity(C.byref(d), C.byref(s), self.abi.dptr(pop), T, self.abi.iptr(up), self.abi.iptr(low),
                                             int(layer0_intensity), *arrs, self.abi.dptr(out["lum"]),
                                             self.abi.iptr(out["lay_nb_hg"]), C.c_double(beaming_factor),
                                             self.abi.dptr(out["tau_sat"]), self.abi.dptr(out["temp_sat"]))
        assert rc == 0
        return out

    def calc_molecular_populations(self, p, acceleration=None):
        """The warm-start chain + retry policy of calc_molecular_populations (radiative_transfer.cpp:219-289).
        ref: the reference function itself; port: ora_solve with RT_START_CHAINED."""
        accel = p.accel if acceleration is None else acceleration
        if self.kind != "ref":
            return self.solve(p, start_mode=self.abi.RT_START_CHAINED, accel=accel, retry=True,
                              max_iter=150 if accel else 15000)
        d, s, hold = self._views(p)
        pop = np.zeros((p.L, p.N))
        st = np.zeros(p.L, dtype=self.abi.STATUS_DTYPE); sec = C.c_double()
        rc = self.lib.ref_calc_molecular_populations(C.byref(d), C.byref(s), self.abi.dptr(pop), int(accel),
                                                     st.ctypes.data_as(C.POINTER(self.abi.rt_layer_status)), C.byref(sec))
        assert rc == 0
        return pop, st, sec.value
