"""CPU: librt_b200.so loads and exports every symbol include/rt_b200.h declares; the ctypes mirror matches the header;
the library refuses to compute without a GPU (no CPU fallback)."""
import ctypes as C
import os
import re

import pytest

from radiative_transfer_b200 import abi, solver

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def header_functions():
    txt = open(os.path.join(ROOT, "include", "rt_b200.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z_0-9]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported():
    lib = solver.lib()
    names = header_functions()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/rt_b200.h but not exported"
    assert sorted(solver.ENTRY_POINTS) == names
    assert lib.rt_abi_version() == solver.ABI_VERSION


def test_params_default_matches_python_mirror():
    lib = solver.lib()
    p = abi.rt_params()
    lib.rt_params_default(C.byref(p))
    q = abi.default_params()
    for f, _ in abi.rt_params._fields_:
        assert getattr(p, f) == getattr(q, f), f


def test_status_strings():
    lib = solver.lib()
    assert lib.rt_status_string(0) == b"ok"
    assert b"dust" in lib.rt_status_string(4)  # iteration_lvg.cpp:82


def test_struct_sizes_match_header():
    """sizeof of every struct as the C compiler sees it (a tiny C program built with gcc) == ctypes mirror."""
    import subprocess, tempfile
    src = r'''
#include <stdio.h>
#include "rt_b200.h"
int main(void){ printf("%zu %zu %zu %zu %zu %zu %zu %zu\n", sizeof(rt_params), sizeof(rt_coll_table), sizeof(rt_dust_comp),
  sizeof(rt_lvg_table), sizeof(rt_overlap_table), sizeof(rt_model_desc), sizeof(rt_layer_soa), sizeof(rt_layer_status)); return 0; }
'''
    with tempfile.TemporaryDirectory() as d:
        f = os.path.join(d, "s.c")
        open(f, "w").write(src)
        exe = os.path.join(d, "s")
        subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), f, "-o", exe], check=True)
        got = [int(x) for x in subprocess.run([exe], check=True, capture_output=True, text=True).stdout.split()]
    want = [C.sizeof(t) for t in (abi.rt_params, abi.rt_coll_table, abi.rt_dust_comp, abi.rt_lvg_table, abi.rt_overlap_table,
                                  abi.rt_model_desc, abi.rt_layer_soa, abi.rt_layer_status)]
    assert got == want


def test_no_cpu_fallback(problems):
    """Without a CUDA device rt_model_create must fail with RT_ERR_CUDA -- it must not compute on the host."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(solver.RtError, match="CUDA|device"):
        solver.Model(problems("generic", 2))


def test_product_does_not_import_oracle():
    """Only tests/, bench.py and __graft_entry__.smoke() may touch oracle/."""
    pkg = os.path.join(ROOT, "radiative_transfer_b200")
    for dp, _, fs in os.walk(pkg):
        for f in fs:
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cpp", "Makefile")):
                txt = open(os.path.join(dp, f), errors="ignore").read()
                assert "oracle.loader" not in txt and "librt_oracle" not in txt and "libref_oracle" not in txt, f
                assert "from oracle" not in txt and "import oracle" not in txt, f
