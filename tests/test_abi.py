"""The C-ABI library loads and exports every symbol include/psgpu.h declares; host-only entry points behave; without a
CUDA device context creation fails loudly (there is no CPU fallback)."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

import parsip_b200 as ps
from parsip_b200 import api, blobtree as bt

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_symbols():
    txt = open(os.path.join(ROOT, "include", "psgpu.h")).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(psgpu_[a-z0-9_]+)\s*\(", txt)))


def test_every_declared_symbol_is_exported():
    lib = ps.load_library()
    syms = declared_symbols()
    assert len(syms) >= 19
    for s in syms:
        assert hasattr(lib, s), "libpsgpu.so does not export " + s
    assert sorted(api.EXPORTS) == syms, "api.EXPORTS out of date with include/psgpu.h"


def test_struct_sizes_match_header():
    """ctypes mirrors vs the C compiler's view of include/psgpu.h"""
    src = r'''
#include <stdio.h>
#include "psgpu.h"
int main(){ printf("%zu %zu %zu %zu\n", sizeof(psgpu_params), sizeof(psgpu_mesh), sizeof(psgpu_stats), sizeof(psgpu_tree)); return 0; }
'''
    exe = "/tmp/psgpu_sizes"
    subprocess.run(["gcc", "-x", "c", "-", "-I", os.path.join(ROOT, "include"), "-o", exe], input=src.encode(), check=True)
    out = subprocess.check_output([exe]).decode().split()
    assert [int(v) for v in out] == [C.sizeof(api.Params), C.sizeof(api._Mesh), C.sizeof(api._Stats), C.sizeof(api._Tree)]


def test_compat_header_layouts_compile():
    """include/PS_SimdPolyCompat.h carries static_asserts pinning the reference struct sizes (sizeof(MPU) == 21524 ...)."""
    src = '#include "PS_SimdPolyCompat.h"\nint main(){ PS::SIMDPOLY::SOABlobOps o; o.ctOps = 0; return (int)o.ctOps; }\n'
    subprocess.run(["g++", "-std=c++11", "-fsyntax-only", "-x", "c++", "-", "-I", os.path.join(ROOT, "include")], input=src.encode(), check=True)


def test_linearize_errors_and_info():
    flat = bt.flatten(bt.Union(bt.Blend(bt.Point([0, 0, 0]), bt.Point([0.5, 0, 0]), bt.Point([1, 0, 0])), bt.Line([0, 1, 0], [1, 1, 0])))
    p = ps.linearize(flat, ps.SEM_SCALAR)
    assert p.n_prims == 4 and p.max_value_stack == 1 and p.n_quads == 1 + 1 + 1 + 2 + 1  # 3 points, segment (2), END; union of prim folds in
    with pytest.raises(ps.PsGpuError) as e:
        ps.linearize(flat, ps.SEM_SIMD, 0)  # 3-child blend: the SIMD path needs binary operators (error -3 like the reference linearizer)
    assert e.value.code == -3
    bad = bt.flatten(bt.Blend(bt.Node(9), bt.Point([0, 0, 0])))
    with pytest.raises(ps.PsGpuError) as e:
        ps.linearize(bad, ps.SEM_SCALAR)
    assert e.value.code == -4


def test_no_cpu_fallback_without_device():
    lib = ps.load_library()
    if lib.psgpu_device_count() > 0:
        pytest.skip("a GPU is present")
    with pytest.raises(ps.PsGpuError) as e:
        ps.Context(0)
    assert e.value.code == -6


def test_count_mpus_matches_reference_formula():
    lo, hi = np.array([-1.2, -1.5, -2.0], np.float32), np.array([1.1, 1.0, 1.7], np.float32)
    from oracle.pyoracle import _load_oracle
    o = _load_oracle()
    for cs in (0.02, 0.05, 0.14, 0.2):
        assert ps.count_mpus(cs, lo, hi) == o.pso_count_mpus(C.c_float(cs), lo.ctypes.data, hi.ctypes.data)
        d = ps.lattice_dims(cs, lo, hi)
        assert d[0] * d[1] * d[2] == ps.count_mpus(cs, lo, hi)
