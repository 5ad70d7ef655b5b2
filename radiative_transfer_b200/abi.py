"""ctypes mirror of include/rt_b200.h (the C ABI of librt_b200.so) and the flat problem container.

The structs are laid out field for field like the header; `Problem.desc()` / `Problem.layers()` build
`rt_model_desc` / `rt_layer_soa` views over numpy arrays (kept alive by the returned holder object).
The same structs are what the reference harness (oracle/ref_harness.cpp) and the C restatement
(oracle/rt_oracle.c) take, so one seeded problem feeds all three implementations byte for byte.
"""
from __future__ import annotations

import ctypes as C
from dataclasses import dataclass, field
from typing import List, Optional

import numpy as np

RT_OK = 0
RT_RULE_GENERIC, RT_RULE_CH3OH, RT_RULE_OH_HF, RT_RULE_H2O, RT_RULE_NH3, RT_RULE_H2CO = 0, 1, 2, 3, 4, 5
RT_SCHEME_LVG, RT_SCHEME_LINE_OVERLAP = 0, 1
RT_START_GIVEN, RT_START_BOUNDARY, RT_START_LTE, RT_START_CHAINED = 0, 1, 2, 3
RT_INTERP_LINEAR, RT_INTERP_CUBIC_SPLINE = 0, 1

c_double_p = C.POINTER(C.c_double)
c_int_p = C.POINTER(C.c_int)


class rt_params(C.Structure):
    _fields_ = [
        ("inv_trans_factor", C.c_double), ("min_line_opacity", C.c_double), ("min_collision_rate", C.c_double),
        ("max_dx", C.c_double),
        ("accel_start", C.c_int), ("accel_period", C.c_int), ("nb_prev_steps", C.c_int), ("reserved0", C.c_int),
        ("boltzmann_constant", C.c_double), ("cm_inverse_to_kelvins", C.c_double), ("speed_of_light", C.c_double),
        ("eight_pi", C.c_double), ("onedivby_sqrt_pi", C.c_double), ("planck_constant", C.c_double),
    ]


class rt_coll_table(C.Structure):
    _fields_ = [("nb_lev", C.c_int), ("jmax", C.c_int), ("tgrid", c_double_p), ("coeff", c_double_p),
                ("interp", C.c_int), ("reserved0", C.c_int)]


class rt_dust_comp(C.Structure):
    _fields_ = [("nb_ph_en", C.c_int), ("reserved0", C.c_int), ("ph_en", c_double_p), ("abs_coeff", c_double_p),
                ("wvl_exp", C.c_double)]


class rt_lvg_table(C.Structure):
    _fields_ = [("nb_d", C.c_int), ("nb_g", C.c_int), ("delta", c_double_p), ("gamma", c_double_p), ("p", c_double_p)]


class rt_overlap_table(C.Structure):
    _fields_ = [("nb_d", C.c_int), ("nb_dx", C.c_int), ("nb_gr", C.c_int), ("nb_g", C.c_int),
                ("delta", c_double_p), ("dx", c_double_p), ("gratio", c_double_p), ("gamma", c_double_p),
                ("p", c_double_p), ("delta_is_log10", C.c_int), ("reserved0", C.c_int)]


class rt_model_desc(C.Structure):
    _fields_ = [
        ("nb_lev", C.c_int), ("species_rule", C.c_int),
        ("energy", c_double_p), ("g", c_int_p), ("v", c_int_p), ("j", c_double_p),
        ("mol_mass", C.c_double), ("mol_name", C.c_char_p), ("einst", c_double_p),
        ("nb_neutral_tables", C.c_int), ("nb_electron_tables", C.c_int), ("tables", C.POINTER(rt_coll_table)),
        ("scheme", C.c_int), ("nb_line_groups", C.c_int), ("line_groups", c_int_p),
        ("lvg", rt_lvg_table), ("overlap1", C.POINTER(rt_overlap_table)), ("overlap2", C.POINTER(rt_overlap_table)),
        ("dust_ncomp", C.c_int), ("reserved0", C.c_int), ("dust", C.POINTER(rt_dust_comp)),
        ("params", rt_params),
    ]


class rt_layer_soa(C.Structure):
    _fields_ = [("nb_lay", C.c_int), ("dust_ncomp", C.c_int)] + [
        (n, c_double_p) for n in ("temp_n", "temp_el", "el_conc", "h_conc", "ph2_conc", "oh2_conc", "he_conc",
                                  "mol_conc", "vel_turb", "velg_n", "dz", "vel_n", "dust_conc")
    ] + [("height", C.c_double), ("vel_width", c_double_p), ("partner_conc", c_double_p),
         ("nb_models", C.c_int), ("reserved0", C.c_int), ("model_first", c_int_p)]


class rt_layer_status(C.Structure):
    _fields_ = [("iters", C.c_int), ("converged", C.c_int), ("eq_error", C.c_double), ("pop_error", C.c_double),
                ("rel_error", C.c_double)]


STATUS_DTYPE = np.dtype([("iters", "<i4"), ("converged", "<i4"), ("eq_error", "<f8"), ("pop_error", "<f8"),
                         ("rel_error", "<f8")])
assert STATUS_DTYPE.itemsize == C.sizeof(rt_layer_status)


def default_params() -> rt_params:
    """rt_params_default() of the library, restated (values: SURVEY.md section 5, oracle/shim/constants.h)."""
    p = rt_params()
    p.inv_trans_factor = -0.1
    p.min_line_opacity = 1.e-99
    p.min_collision_rate = 1.e-99
    p.max_dx = 4.0
    p.accel_start, p.accel_period, p.nb_prev_steps = 40, 5, 5
    p.boltzmann_constant = 1.38064852e-16
    p.cm_inverse_to_kelvins = 1.438777364
    p.speed_of_light = 2.99792458e+10
    p.eight_pi = 25.132741228718345
    p.onedivby_sqrt_pi = 0.56418958354775629
    p.planck_constant = 6.62607004e-27
    return p


def dptr(a: np.ndarray):
    assert a.dtype == np.float64 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_double_p)


def iptr(a: np.ndarray):
    assert a.dtype == np.int32 and a.flags["C_CONTIGUOUS"]
    return a.ctypes.data_as(c_int_p)


@dataclass
class CollTable:
    nb_lev: int
    tgrid: np.ndarray  # [jmax]
    coeff: np.ndarray  # [imax, jmax]
    interp: int = 0    # RT_INTERP_LINEAR / RT_INTERP_CUBIC_SPLINE (collision_data_cub_spline, coll_rates.cpp:89-132)


@dataclass
class DustComp:
    ph_en: np.ndarray
    abs_coeff: np.ndarray
    wvl_exp: float


@dataclass
class OverlapTable:
    delta: np.ndarray
    dx: np.ndarray
    gratio: np.ndarray
    gamma: np.ndarray
    p: np.ndarray  # [nb_d*nb_dx, nb_gr*nb_g]


@dataclass
class Problem:
    """One molecule + physics tables + cloud, in the flat layout of include/rt_b200.h."""
    name: str
    mol_name: str
    mol_mass: float
    species_rule: int
    scheme: int
    energy: np.ndarray
    g: np.ndarray
    v: np.ndarray
    j: np.ndarray
    einst: np.ndarray  # [N, N]
    tables: List[CollTable]
    nb_neutral_tables: int
    lvg_delta: np.ndarray
    lvg_gamma: np.ndarray
    lvg_p: np.ndarray  # [nb_d, nb_g]
    dust: List[DustComp]
    layers: dict  # name -> [L] arrays (+ dust_conc [L, ncomp], height float)
    overlap1: Optional[OverlapTable] = None
    overlap2: Optional[OverlapTable] = None
    line_groups: Optional[np.ndarray] = None  # [ngroups, 5] int32
    accel: bool = True
    max_iter: int = 150
    start_mode: int = RT_START_BOUNDARY
    params: rt_params = field(default_factory=default_params)

    @property
    def N(self) -> int:
        return int(self.energy.shape[0])

    @property
    def L(self) -> int:
        return int(self.layers["temp_n"].shape[0])

    @property
    def nb_rad(self) -> int:
        return int(np.count_nonzero(np.tril(self.einst, -1) > 0))

    def sliced(self, lo: int, hi: int) -> "Problem":
        """Same model, layers [lo, hi) only (layer sharding / bounded CPU samples)."""
        import copy
        q = copy.copy(self)
        lay = {}
        for k, a in self.layers.items():
            lay[k] = a if k == "height" else np.ascontiguousarray(a[lo:hi])
        q.layers = lay
        q.model_first = None  # a subset of a stack is a plain list of layers
        return q

    def take(self, idx) -> "Problem":
        """Same model, the listed layers only, in the listed order (cyclic layer sharding)."""
        import copy
        idx = np.asarray(idx, dtype=np.int64)
        q = copy.copy(self)
        q.layers = {k: (a if k == "height" else np.ascontiguousarray(a[idx])) for k, a in self.layers.items()}
        q.model_first = None
        return q

    # -- ctypes views ------------------------------------------------------------------------------------
    def desc(self):
        """Returns (rt_model_desc, holder).  Keep `holder` alive while the desc is in use."""
        hold = []
        d = rt_model_desc()
        d.nb_lev = self.N
        d.species_rule = self.species_rule
        for nm in ("energy", "j"):
            a = np.ascontiguousarray(getattr(self, nm), dtype=np.float64); hold.append(a); setattr(d, nm, dptr(a))
        for nm in ("g", "v"):
            a = np.ascontiguousarray(getattr(self, nm), dtype=np.int32); hold.append(a); setattr(d, nm, iptr(a))
        d.mol_mass = self.mol_mass
        nm_b = self.mol_name.encode(); hold.append(nm_b); d.mol_name = nm_b
        e = np.ascontiguousarray(self.einst, dtype=np.float64); hold.append(e); d.einst = dptr(e)
        d.nb_neutral_tables = self.nb_neutral_tables
        d.nb_electron_tables = len(self.tables) - self.nb_neutral_tables
        tarr = (rt_coll_table * max(1, len(self.tables)))()
        for k, t in enumerate(self.tables):
            tg = np.ascontiguousarray(t.tgrid, dtype=np.float64); co = np.ascontiguousarray(t.coeff, dtype=np.float64)
            assert co.shape == (t.nb_lev * (t.nb_lev - 1) // 2, tg.shape[0])
            hold += [tg, co]
            tarr[k].nb_lev, tarr[k].jmax, tarr[k].tgrid, tarr[k].coeff = t.nb_lev, tg.shape[0], dptr(tg), dptr(co)
            tarr[k].interp = int(getattr(t, "interp", 0))
        hold.append(tarr); d.tables = tarr
        d.scheme = self.scheme
        if self.line_groups is not None:
            lg = np.ascontiguousarray(self.line_groups, dtype=np.int32); hold.append(lg)
            d.nb_line_groups, d.line_groups = lg.shape[0], iptr(lg)
        else:
            d.nb_line_groups = 0
        ld, lg_, lp = (np.ascontiguousarray(x, dtype=np.float64) for x in (self.lvg_delta, self.lvg_gamma, self.lvg_p))
        hold += [ld, lg_, lp]
        d.lvg.nb_d, d.lvg.nb_g, d.lvg.delta, d.lvg.gamma, d.lvg.p = ld.shape[0], lg_.shape[0], dptr(ld), dptr(lg_), dptr(lp)
        for nm in ("overlap1", "overlap2"):
            o = getattr(self, nm)
            if o is not None:
                s = rt_overlap_table()
                arrs = [np.ascontiguousarray(x, dtype=np.float64) for x in (o.delta, o.dx, o.gratio, o.gamma, o.p)]
                hold += arrs
                s.nb_d, s.nb_dx, s.nb_gr, s.nb_g = (arrs[0].shape[0], arrs[1].shape[0], arrs[2].shape[0], arrs[3].shape[0])
                s.delta, s.dx, s.gratio, s.gamma, s.p = (dptr(x) for x in arrs)
                hold.append(s); setattr(d, nm, C.pointer(s))
        d.dust_ncomp = len(self.dust)
        darr = (rt_dust_comp * max(1, len(self.dust)))()
        for k, c in enumerate(self.dust):
            pe = np.ascontiguousarray(c.ph_en, dtype=np.float64); ab = np.ascontiguousarray(c.abs_coeff, dtype=np.float64)
            hold += [pe, ab]
            darr[k].nb_ph_en, darr[k].ph_en, darr[k].abs_coeff, darr[k].wvl_exp = pe.shape[0], dptr(pe), dptr(ab), c.wvl_exp
        hold.append(darr); d.dust = darr
        d.params = self.params
        return d, hold

    def layer_soa(self):
        """Returns (rt_layer_soa, holder)."""
        hold = []
        s = rt_layer_soa()
        s.nb_lay = self.L
        s.dust_ncomp = len(self.dust)
        for nm in ("temp_n", "temp_el", "el_conc", "h_conc", "ph2_conc", "oh2_conc", "he_conc", "mol_conc", "vel_turb",
                   "velg_n", "dz", "vel_n", "dust_conc"):
            a = np.ascontiguousarray(self.layers[nm], dtype=np.float64); hold.append(a); setattr(s, nm, dptr(a))
        s.height = float(self.layers["height"])
        mf = getattr(self, "model_first", None)
        if mf is not None:  # a stack of independent slabs (sharding.stack_clouds): one warm-start chain per model
            a = np.ascontiguousarray(mf, dtype=np.int32); hold.append(a)
            s.nb_models = a.shape[0] - 1
            s.model_first = iptr(a)
        return s, hold
