"""Host-side mirror of the reference's solver entry points over the C ABI (include/rt_b200.h).

`Model` plays the role of the objects the reference driver builds around one molecule
(`iteration_scheme_lvg` + `iteration_control`, radiative_transfer.cpp:664-670): create it from a `Problem`,
upload a cloud, call `calc_molecular_populations` / `calculate_populations` / `calc_new_pop` / `intensity_calc`.
Everything is computed by librt_b200.so on the GPU; there is no CPU path and a missing library is an error.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

from . import abi

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("RT_B200_LIB", os.path.join(HERE, "librt_b200.so"))  # override: kernel-variant experiments only
_lib = None
ABI_VERSION = 5

ENTRY_POINTS = [
    "rt_params_default", "rt_status_string", "rt_last_error", "rt_abi_version", "rt_model_create", "rt_model_destroy",
    "rt_cloud_upload", "rt_boundary_populations", "rt_step", "rt_assemble", "rt_intensity", "rt_solve",
    "rt_calc_molecular_populations", "rt_solve_device", "rt_get_status", "rt_get_populations", "rt_set_populations",
    "rt_profile_reset", "rt_profile_get", "rt_iterations_enqueued", "rt_maser_postpass", "rt_cloud_rebuild",
    "rt_profile_get_work", "rt_launch_count", "rt_timer_begin", "rt_timer_end", "rt_device_fp64_peak",
    "rt_lim_luminosity", "rt_set_gain_width_line", "rt_copy_populations_device",
    "rt_intensity_noalloc",
]


class RtError(RuntimeError):
    pass


def build_library() -> None:
    """nvcc -gencode arch=compute_100a,code=sm_100a ... (radiative_transfer_b200/csrc/Makefile); works without a GPU."""
    subprocess.run(["make", "-C", os.path.join(HERE, "csrc"), "-j8"], check=True, stdout=subprocess.DEVNULL)


def lib():
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RtError(f"{LIB_PATH} is missing: build it with __graft_entry__.build() (there is no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        L.rt_status_string.restype = C.c_char_p
        L.rt_last_error.restype = C.c_char_p
        L.rt_last_error.argtypes = [C.c_void_p]
        L.rt_iterations_enqueued.restype = C.c_longlong
        L.rt_iterations_enqueued.argtypes = [C.c_void_p]
        L.rt_launch_count.restype = C.c_longlong
        _lib = L
    return _lib


class Model:
    def __init__(self, problem: abi.Problem, device: int = 0, upload_cloud: bool = True):
        self.p = problem
        self.L_ = lib()
        self.h = C.c_void_p()
        d, hold = problem.desc()
        rc = self.L_.rt_model_create(C.byref(d), int(device), C.byref(self.h))
        if rc != abi.RT_OK:
            raise RtError(f"rt_model_create: {self.L_.rt_status_string(rc).decode()}: {self.L_.rt_last_error(None).decode()}")
        self.N = problem.N
        self.L = 0
        if upload_cloud:
            self.cloud_upload(problem)

    # -- helpers -------------------------------------------------------------------------------------------
    def _ck(self, rc, what):
        if rc != abi.RT_OK:
            raise RtError(f"{what}: {self.L_.rt_status_string(rc).decode()}: {self.L_.rt_last_error(self.h).decode()}")

    def close(self):
        if self.h:
            self.L_.rt_model_destroy(self.h)
            self.h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _st(self):
        return np.zeros(self.L, dtype=abi.STATUS_DTYPE)

    @staticmethod
    def _stp(st):
        return st.ctypes.data_as(C.POINTER(abi.rt_layer_status))

    # -- the reference's per-layer setters, for all layers at once (radiative_transfer.cpp:238-242) ----------
    def cloud_upload(self, problem: abi.Problem):
        s, hold = problem.layer_soa()
        self._ck(self.L_.rt_cloud_upload(self.h, C.byref(s)), "rt_cloud_upload")
        self.L = problem.L

    # -- boundary_layer_populations (iteration_control.cpp:52-91) ---------------------------------------------
    def boundary_layer_populations(self):
        pop = np.zeros((self.L, self.N))
        self._ck(self.L_.rt_boundary_populations(self.h, abi.dptr(pop)), "rt_boundary_populations")
        return pop

    # -- iteration_scheme_lvg::operator() (iteration_lvg.cpp:112-161) -----------------------------------------
    def assemble(self, pop):
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        M = np.zeros((self.L, self.N, self.N)); df = np.zeros((self.L, self.N))
        self._ck(self.L_.rt_assemble(self.h, abi.dptr(pop), abi.dptr(M), abi.dptr(df)), "rt_assemble")
        return M, df

    # -- iteration_scheme_lvg::calc_new_pop (iteration_lvg.cpp:87-110) ----------------------------------------
    def calc_new_pop(self, old_pop):
        old_pop = np.ascontiguousarray(old_pop, dtype=np.float64)
        new = np.zeros((self.L, self.N)); eq = np.zeros(self.L)
        self._ck(self.L_.rt_step(self.h, abi.dptr(old_pop), abi.dptr(new), abi.dptr(eq)), "rt_step")
        return new, eq

    # -- intensity_calc (iteration_lvg.cpp:163-185 / :503-526) ------------------------------------------------
    def intensity_calc(self, lay, up, low, pop):
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        out = C.c_double()
        self._ck(self.L_.rt_intensity(self.h, int(lay), int(up), int(low), abi.dptr(pop), C.byref(out)), "rt_intensity")
        return out.value

    def intensity_calc_noalloc(self, lay, up, low, pop):
        """rt_intensity_noalloc: intensity_calc through the fixed device scratch array (no allocation; N <= 1024)."""
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        out = C.c_double()
        self._ck(self.L_.rt_intensity_noalloc(self.h, int(lay), int(up), int(low), abi.dptr(pop), C.byref(out)), "rt_intensity_noalloc")
        return out.value

    # -- iteration_control::calculate_populations (iteration_control.h:196-242), every layer ------------------
    def calculate_populations(self, pop0=None, start_mode=None, max_iter=None, min_err=1e-5, accel=None):
        p = self.p
        start_mode = p.start_mode if start_mode is None else start_mode
        if pop0 is not None:
            start_mode = abi.RT_START_GIVEN
        pop = np.zeros((self.L, self.N)) if pop0 is None else np.array(pop0, dtype=np.float64, order="C")
        st = self._st()
        self._ck(self.L_.rt_solve(self.h, abi.dptr(pop), int(start_mode), int(p.max_iter if max_iter is None else max_iter),
                                  C.c_double(min_err), int(p.accel if accel is None else accel), self._stp(st)), "rt_solve")
        return pop, st

    # -- calc_molecular_populations (radiative_transfer.cpp:219-289) incl. retry policy -----------------------
    def calc_molecular_populations(self, acceleration=None, start_mode=None, pop0=None, max_iter_acc=150,
                                   max_iter_ext=15000, min_err=1e-5):
        p = self.p
        start_mode = p.start_mode if start_mode is None else start_mode
        if pop0 is not None:
            start_mode = abi.RT_START_GIVEN
        pop = np.zeros((self.L, self.N)) if pop0 is None else np.array(pop0, dtype=np.float64, order="C")
        st = self._st()
        self._ck(self.L_.rt_calc_molecular_populations(self.h, abi.dptr(pop), int(start_mode),
                                                       int(p.accel if acceleration is None else acceleration),
                                                       int(max_iter_acc), int(max_iter_ext), C.c_double(min_err),
                                                       self._stp(st)), "rt_calc_molecular_populations")
        return pop, st

    # -- device-resident variant (bench `value`) ---------------------------------------------------------------
    def solve_device(self, start_mode=None, max_iter=None, min_err=1e-5, accel=None):
        p = self.p
        self._ck(self.L_.rt_solve_device(self.h, int(p.start_mode if start_mode is None else start_mode),
                                         int(p.max_iter if max_iter is None else max_iter), C.c_double(min_err),
                                         int(p.accel if accel is None else accel)), "rt_solve_device")

    def get_status(self):
        st = self._st()
        self._ck(self.L_.rt_get_status(self.h, self._stp(st)), "rt_get_status")
        return st

    def get_populations(self):
        pop = np.zeros((self.L, self.N))
        self._ck(self.L_.rt_get_populations(self.h, abi.dptr(pop)), "rt_get_populations")
        return pop

    def copy_populations_device(self, dst_ptr: int):
        """Populations [L, N] device-to-device into a caller-owned CUDA buffer (e.g. a torch tensor's data_ptr())."""
        self._ck(self.L_.rt_copy_populations_device(self.h, C.c_void_p(int(dst_ptr))), "rt_copy_populations_device")

    def set_populations(self, pop):
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        self._ck(self.L_.rt_set_populations(self.h, abi.dptr(pop)), "rt_set_populations")

    # -- transition_data_container::calc_inv / calc_gain / calc_line_profile ------------------------------------
    def set_gain_width_line(self, up: int, low: int, thermal_offset: float = 5.e4):
        """The oH2O 6_16 -> 5_23 special line width of calc_gain (transition_data.cpp:249-260); up < 0 clears it."""
        self._ck(lib().rt_set_gain_width_line(self.h, int(up), int(low), C.c_double(thermal_offset)), "rt_set_gain_width_line")

    def maser_postpass(self, pop, up, low, nb_freq=300, nb_aspect=37, delta_aspect=0.25, velocity_shift=5.e5):
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        up = np.ascontiguousarray(up, dtype=np.int32); low = np.ascontiguousarray(low, dtype=np.int32)
        T = up.shape[0]
        out = dict(inv=np.zeros((T, self.L)), gain=np.zeros((T, self.L)), scalars=np.zeros((T, 5)),
                   tau_vs_aspect=np.zeros((T, nb_aspect)), tau_vs_freq=np.zeros((T, nb_freq)))
        self._ck(self.L_.rt_maser_postpass(self.h, abi.dptr(pop), T, abi.iptr(up), abi.iptr(low), nb_freq, nb_aspect,
                                           C.c_double(delta_aspect), C.c_double(velocity_shift), abi.dptr(out["inv"]),
                                           abi.dptr(out["gain"]), abi.dptr(out["scalars"]), abi.dptr(out["tau_vs_aspect"]),
                                           abi.dptr(out["tau_vs_freq"])), "rt_maser_postpass")
        return out

    # -- lim_luminosity_lvg + calc_exc_temp + calc_saturation_depth (maser_luminosity.cpp:7-106, transition_data.cpp) ---
    LIM_KEYS = ("lum_arr", "pump_eff_arr", "loss_rate_arr", "pump_rate_arr", "emiss_coeff_arr", "exc_temp_arr")

    def lim_luminosity(self, pop, up, low, beaming_factor=1.0, lay_nb_hg=None, layer0_intensity=True):
        """Pump / loss rates and limiting luminosity of listed transitions.  layer0_intensity=True is the reference's
        literal behaviour (J of the loss terms from the populations of layer 0); lay_nb_hg (maser_postpass
        scalars[:, 4]) switches the saturation depth on."""
        pop = np.ascontiguousarray(pop, dtype=np.float64)
        up = np.ascontiguousarray(up, dtype=np.int32); low = np.ascontiguousarray(low, dtype=np.int32)
        T = up.shape[0]
        out = {k: np.zeros((T, self.L)) for k in self.LIM_KEYS}
        out.update(lum=np.zeros(T), tau_sat=np.zeros(T), temp_sat=np.zeros(T))
        hg = None if lay_nb_hg is None else np.ascontiguousarray(lay_nb_hg, dtype=np.int32)
        self._ck(self.L_.rt_lim_luminosity(self.h, abi.dptr(pop), T, abi.iptr(up), abi.iptr(low), int(layer0_intensity),
                                           *[abi.dptr(out[k]) for k in self.LIM_KEYS], abi.dptr(out["lum"]),
                                           abi.iptr(hg) if hg is not None else None, C.c_double(beaming_factor),
                                           abi.dptr(out["tau_sat"]), abi.dptr(out["temp_sat"])), "rt_lim_luminosity")
        return out

    # -- instrumentation --------------------------------------------------------------------------------------
    def profile_reset(self, timing: bool = False):
        self._ck(self.L_.rt_profile_reset(self.h, int(timing)), "rt_profile_reset")

    def profile_get(self):
        """{kernel class: dict(ms, launches, flops, bytes)}; flops / bytes are ALGORITHMIC."""
        cap = 64
        names = (C.c_char_p * cap)(); ms = (C.c_double * cap)(); ln = (C.c_longlong * cap)(); n = C.c_int()
        fl = (C.c_double * cap)(); by = (C.c_double * cap)()
        self._ck(self.L_.rt_profile_get(self.h, cap, names, ms, ln, C.byref(n)), "rt_profile_get")
        self._ck(self.L_.rt_profile_get_work(self.h, cap, fl, by), "rt_profile_get_work")
        return {names[i].decode(): dict(ms=ms[i], launches=ln[i], flops=fl[i], bytes=by[i]) for i in range(n.value)}

    def cloud_rebuild(self):
        self._ck(self.L_.rt_cloud_rebuild(self.h), "rt_cloud_rebuild")

    def timer_begin(self):
        self._ck(self.L_.rt_timer_begin(self.h), "rt_timer_begin")

    def timer_end(self) -> float:
        ms = C.c_double()
        self._ck(self.L_.rt_timer_end(self.h, C.byref(ms)), "rt_timer_end")
        return ms.value

    def iterations_enqueued(self) -> int:
        return int(self.L_.rt_iterations_enqueued(self.h))


def device_fp64_peak(device: int = 0, reps: int = 3):
    """(DFMA TFLOP/s, DMMA TFLOP/s) measured on `device` by the library's microbenchmark kernels."""
    a = C.c_double(); b = C.c_double()
    rc = lib().rt_device_fp64_peak(int(device), int(reps), C.byref(a), C.byref(b))
    if rc != abi.RT_OK:
        raise RtError(f"rt_device_fp64_peak: {lib().rt_status_string(rc).decode()}")
    return a.value, b.value


def launch_count() -> int:
    """Kernels launched by librt_b200.so in this process so far."""
    return int(lib().rt_launch_count())
