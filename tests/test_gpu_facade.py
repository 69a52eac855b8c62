"""-m gpu: the drop-in C++ facade (include/PS_SimdPolyCompat.h: PrepareBBoxes / CountMPUNeeded / Polygonize / FieldValues and
the SimdPoly wrapper class) driven by a client written like a caller of the reference's PS_SimdPoly
(tests/facade/facade_probe.cpp), against the golden vectors of the reference's own unmodified build: the caller-owned
PolyMPUs must hold the reference's content byte for byte."""
import ctypes as C
import os

import numpy as np
import pytest

import gpu_helpers as gh
import scenes_small
from parsip_b200 import blobtree as bt
import parsip_b200 as ps
from parsip_b200 import api as ps_api
from test_gpu_parity import GOLD, RSQRT_SCENES, bits

pytestmark = pytest.mark.gpu
HERE = os.path.dirname(os.path.abspath(__file__))


class _Desc(C.Structure):
    _fields_ = [("n_prims", C.c_int32), ("n_ops", C.c_int32), ("n_mats", C.c_int32)] + [(n, C.c_void_p) for n in (
        "prim_type", "prim_pos", "prim_dir", "prim_res", "prim_color", "prim_mat", "op_type", "op_left", "op_right", "op_kind", "op_res",
        "mats_bwd", "mats_fwd")]


def _lib():
    l = C.CDLL(os.path.join(HERE, "facade", "libfacade_probe.so"))
    l.fac_scene_create.restype = C.c_void_p
    l.fac_scene_create.argtypes = [C.c_void_p]
    for f in ("fac_scene_destroy", "fac_get_bbox", "fac_get_boxes", "fac_mesh_concat"):
        getattr(l, f).restype = None
    l.fac_scene_destroy.argtypes = [C.c_void_p]
    l.fac_prepare_bboxes.argtypes = [C.c_void_p, C.c_float]
    l.fac_get_bbox.argtypes = [C.c_void_p] * 3
    l.fac_get_boxes.argtypes = [C.c_void_p] * 3
    l.fac_count_mpus.restype = C.c_uint32
    l.fac_count_mpus.argtypes = [C.c_void_p, C.c_float]
    l.fac_field.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_void_p]
    l.fac_newton.argtypes = [C.c_void_p, C.c_int] + [C.c_void_p] * 4 + [C.c_float, C.c_int] + [C.c_void_p] * 3
    l.fac_polygonize.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
    l.fac_mesh_concat.argtypes = [C.c_void_p] * 6
    l.fac_simdpoly_run.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
    return l


class Scene:
    def __init__(self, lib, soa):
        self.lib = lib
        self.k = {k: np.ascontiguousarray(v) for k, v in soa.items() if isinstance(v, np.ndarray)}
        k = self.k
        p = lambda a: a.ctypes.data
        self.d = _Desc(soa["n_prims"], soa["n_ops"], soa["n_mats"], p(k["prim_type"]), p(k["prim_pos"]), p(k["prim_dir"]), p(k["prim_res"]),
                       p(k["prim_color"]), p(k["prim_mat"]), p(k["op_type"]), p(k["op_left"]), p(k["op_right"]), p(k["op_kind"]), p(k["op_res"]),
                       p(k["mats_bwd"]), p(k["mats_fwd"]))
        self.h = lib.fac_scene_create(C.addressof(self.d))
        assert self.h, "facade scene rejected"
        self.n_prims, self.n_ops = soa["n_prims"], soa["n_ops"]

    def close(self):
        self.lib.fac_scene_destroy(self.h)


@pytest.fixture(scope="module")
def lib(ctx):  # ctx first: fails loudly without a device
    return _lib()


@pytest.mark.parametrize("name", sorted(scenes_small.SIMD_SCENES))
def test_facade_polygonize_fills_polympus_like_the_reference(lib, name):
    g = np.load(os.path.join(GOLD, "%s_padded.npz" % name))
    soa = bt.to_binary_soa(bt.binarize(scenes_small.SIMD_SCENES[name]()))
    sc = Scene(lib, soa)
    try:
        assert lib.fac_prepare_bboxes(sc.h, 0.05) == 1
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        lib.fac_get_bbox(sc.h, lo.ctypes.data, hi.ctypes.data)
        assert np.array_equal(bits(lo), bits(g["bbox_lo"])) and np.array_equal(bits(hi), bits(g["bbox_hi"]))
        pb, ob = np.zeros((sc.n_prims, 6), np.float32), np.zeros((max(sc.n_ops, 1), 6), np.float32)
        lib.fac_get_boxes(sc.h, pb.ctypes.data, ob.ctypes.data)
        assert np.array_equal(bits(pb), bits(g["prim_boxes"])), "PrepareBBoxes primitive boxes differ from the reference"
        assert np.allclose(ob[:sc.n_ops], np.load(os.path.join(GOLD, "%s_asis.npz" % name))["op_boxes"], rtol=0, atol=0)
        cs = float(g["cellsize"])
        assert lib.fac_count_mpus(sc.h, cs) == int(g["n_mpus"])
        # FieldComputer::fieldValue through the facade
        pts = np.ascontiguousarray(g["pts"], np.float32)
        f = np.zeros(len(pts), np.float32)
        assert lib.fac_field(sc.h, len(pts), pts.ctypes.data, f.ctypes.data) == 1
        if name not in RSQRT_SCENES:
            assert np.array_equal(bits(f), bits(g["field_bcast"]))
        # FieldComputer::computeRootNewtonRaphson through the facade (a20)
        gn = np.load(os.path.join(GOLD, "simd_newton.npz"))
        k = name + "_padded_"
        p1, p2, f1, f2 = (np.ascontiguousarray(gn[k + q], np.float32) for q in ("p1", "p2", "f1", "f2"))
        root, fr, ev = np.zeros((len(p1), 3), np.float32), np.zeros(len(p1), np.float32), np.zeros(len(p1), np.int32)
        assert lib.fac_newton(sc.h, len(p1), p1.ctypes.data, p2.ctypes.data, f1.ctypes.data, f2.ctypes.data, 0.5, 8, root.ctypes.data, fr.ctypes.data,
                              ev.ctypes.data) == 1
        if name not in RSQRT_SCENES:
            assert np.array_equal(bits(root), bits(gn[k + "root8"])) and np.array_equal(ev, gn[k + "evals8"])
        st = np.zeros(4, np.int64)
        assert lib.fac_polygonize(sc.h, cs, st.ctypes.data) == 1
        assert int(st[0]) == int(g["n_mpus"])
        nv, nt = int(st[2]), int(st[3])
        pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
        tri = np.zeros((nt, 3), np.uint32)
        mlo = np.zeros((int(st[0]), 3), np.float32)
        lib.fac_mesh_concat(sc.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data, mlo.ctypes.data)
        if name in RSQRT_SCENES:
            assert abs(nv - int(g["n_vertices"])) <= 0.01 * int(g["n_vertices"])
        else:
            assert int(st[1]) == int(g["n_crossed"]) and nv == int(g["n_vertices"]) and nt == int(g["n_triangles"])
            assert gh.sha(tri) == str(g["tri_sha"]) and gh.sha(pos) == str(g["pos_sha"]) and gh.sha(col) == str(g["col_sha"])
        assert np.isfinite(nrm).all()
        # the MPU lattice origins follow PS_Polygonizer.cpp:360-370 (x outer, z inner)
        side = np.float32(cs) * np.float32(7)
        assert np.array_equal(bits(mlo[1]), bits(np.array([lo[0], lo[1], lo[2] + np.float32(1) * side], np.float32)))
    finally:
        sc.close()


def test_simdpoly_wrapper_class_and_error_codes(lib):
    root = bt.binarize(scenes_small.SIMD_SCENES["s1"]())
    g = np.load(os.path.join(GOLD, "s1_padded.npz"))
    flat = bt.flatten(root)
    tree, keep = ps_api._tree_struct(flat)
    st = np.zeros(4, np.int64)
    assert lib.fac_simdpoly_run(C.addressof(tree), float(g["cellsize"]), st.ctypes.data) == 1
    assert int(st[2]) == int(g["n_vertices"]) and int(st[3]) == int(g["n_triangles"])
    # an n-ary operator is PS_ERROR_NON_BINARY_OP (-3), as in PS_HighPerformanceRender.cpp:81-86
    nary, keep2 = ps_api._tree_struct(bt.flatten(scenes_small.SIMD_SCENES["s1"]()))
    assert lib.fac_simdpoly_run(C.addressof(nary), 0.1, st.ctypes.data) == -3


def test_facade_large_lattice_uses_the_threaded_writer(lib, ctx):
    """~20 000 MPUs (just under the reference's MAX_MPU_COUNT of 24 000): the PolyMPUs writer splits the lattice over several
    host threads; its content must be the C ABI's PSGPU_OUT_MPU mesh of the same scene, MPU by MPU in lattice order."""
    from test_gpu_parity import compat_params, simd_setup
    from parsip_b200 import scenes
    name = "blob"
    soa = bt.to_binary_soa(bt.binarize(scenes_small.SIMD_SCENES[name]()))
    sc = Scene(lib, soa)
    try:
        assert lib.fac_prepare_bboxes(sc.h, 0.05) == 1
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        lib.fac_get_bbox(sc.h, lo.ctypes.data, hi.ctypes.data)
        cs = scenes.cellsize_for(lo, hi, 186)  # 27 MPUs on the longest axis
        n_mpus = lib.fac_count_mpus(sc.h, cs)
        assert 8192 < n_mpus < 24000
        st = np.zeros(4, np.int64)
        assert lib.fac_polygonize(sc.h, cs, st.ctypes.data) == 1
        nv, nt = int(st[2]), int(st[3])
        pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
        tri = np.zeros((nt, 3), np.uint32)
        mlo = np.zeros((int(st[0]), 3), np.float32)
        lib.fac_mesh_concat(sc.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data, mlo.ctypes.data)
    finally:
        sc.close()
    simd_setup(ctx, name)
    m = ctx.polygonize(compat_params(cs, lo, hi))
    p2, n2, c2, t2, ijk, counts = gh.lattice_ordered(m)
    assert int(st[0]) == n_mpus == ctx.stats()["mpus_total"] and int(st[1]) == m["n_mpus_out"]
    assert (nv, nt) == (m["n_vertices"], m["n_triangles"]) and nv > 100000
    for a, b in ((pos, p2), (nrm, n2), (col, c2)):
        assert np.array_equal(bits(a), bits(b))
    assert np.array_equal(tri, t2)
    # lattice origins: x outer, z inner, lo + idx * side (PS_Polygonizer.cpp:360-370)
    dims = ps.lattice_dims(cs, lo, hi)
    side = np.float32(cs) * np.float32(7)
    q = (3 * dims[1] + 5) * dims[2] + 7
    assert np.array_equal(bits(mlo[q]), bits(np.array([lo[0] + np.float32(3) * side, lo[1] + np.float32(5) * side, lo[2] + np.float32(7) * side], np.float32)))
