"""TEST INFRASTRUCTURE ONLY: ctypes binding of tests/hostemu/libhostemu.so (the product's per-thread device
functions compiled for the host). Lets the CPU suite check linearizer + interpreter + pruning walk without a GPU."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
_lib = None


def lib():
    global _lib
    if _lib is None:
        l = C.CDLL(os.path.join(HERE, "hostemu", "libhostemu.so"))
        vp = C.c_void_p
        l.hostemu_field.argtypes = [vp, vp, C.c_int64, vp, C.c_int]
        l.hostemu_field_color.argtypes = [vp, vp, vp, C.c_int64, vp, vp, C.c_int]
        l.hostemu_field_packed_columns.argtypes = [vp, vp, vp, C.c_int64, vp]
        l.hostemu_field_packed.argtypes = [vp, vp, C.c_int64, vp]
        l.hostemu_field_color_packed.argtypes = [vp, vp, vp, C.c_int64, C.c_float, vp, vp]
        l.hostemu_prune.restype = C.c_uint32
        l.hostemu_prune.argtypes = [vp, vp, vp, C.c_int, C.c_float, vp, C.c_uint32, C.c_int]
        _lib = l
    return _lib


def field(quads: np.ndarray, xyz, sem: int) -> np.ndarray:
    quads = np.ascontiguousarray(quads, np.uint32)
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    out = np.zeros(len(xyz), np.float32)
    lib().hostemu_field(quads.ctypes.data, xyz.ctypes.data, len(xyz), out.ctypes.data, sem)
    return out


def field_packed_columns(quads, xy, z) -> np.ndarray:
    """interp_packed.cuh, column form: xy (n, 2), z (n, 8) -> f (n, 8)"""
    quads = np.ascontiguousarray(quads, np.uint32)
    xy = np.ascontiguousarray(xy, np.float32).reshape(-1, 2)
    z = np.ascontiguousarray(z, np.float32).reshape(-1, 8)
    out = np.zeros((len(xy), 8), np.float32)
    lib().hostemu_field_packed_columns(quads.ctypes.data, xy.ctypes.data, z.ctypes.data, len(xy), out.ctypes.data)
    return out


def field_packed(quads, xyz) -> np.ndarray:
    """interp_packed.cuh, general form (4 points per call); len(xyz) must be a multiple of 4"""
    quads = np.ascontiguousarray(quads, np.uint32)
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    assert len(xyz) % 4 == 0
    out = np.zeros(len(xyz), np.float32)
    lib().hostemu_field_packed(quads.ctypes.data, xyz.ctypes.data, len(xyz), out.ctypes.data)
    return out


def field_color_packed(quads, colors, xyz, delta=1e-3):
    """packed PS_SimdPoly walk with colour: fields at (p, p+dx, p+dy, p+dz) and the colour at p"""
    quads = np.ascontiguousarray(quads, np.uint32)
    colors = np.ascontiguousarray(colors, np.float32)
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    f4, col = np.zeros((len(xyz), 4), np.float32), np.zeros((len(xyz), 3), np.float32)
    lib().hostemu_field_color_packed(quads.ctypes.data, colors.ctypes.data, xyz.ctypes.data, len(xyz), float(delta), f4.ctypes.data, col.ctypes.data)
    return f4, col


def field_color(quads, colors, xyz, sem):
    quads = np.ascontiguousarray(quads, np.uint32)
    colors = np.ascontiguousarray(colors, np.float32)
    xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
    f, col = np.zeros(len(xyz), np.float32), np.zeros((len(xyz), 3), np.float32)
    lib().hostemu_field_color(quads.ctypes.data, colors.ctypes.data, xyz.ctypes.data, len(xyz), f.ctypes.data, col.ctypes.data, sem)
    return f, col


def prune(quads, lo, hi, cull_mode: int, iso: float = 0.5, color_exact: bool = False):
    """Returns the pruned program (n,4) uint32, or None when the region needs no evaluation."""
    quads = np.ascontiguousarray(quads, np.uint32)
    lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
    cap = len(quads) + 64
    out = np.zeros((cap, 4), np.uint32)
    n = lib().hostemu_prune(quads.ctypes.data, lo.ctypes.data, hi.ctypes.data, cull_mode, iso, out.ctypes.data, cap, int(color_exact))
    if n == 0:
        return None
    assert n <= cap
    return out[:n].copy()
