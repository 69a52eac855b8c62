"""TEST INFRASTRUCTURE ONLY: ctypes bindings for the oracles.

 * `RefScene`  - oracle/_ref/libpsref*.so, the reference's own SIMD polygonizer compiled from
                 /root/reference by oracle/Makefile (unmodified, or capacity-widened variants).
 * `COracle`   - oracle/libps_oracle.so, the plain-C restatement (oracle/ps_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may import this.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


class _RefDesc(C.Structure):
    _fields_ = [
        ("n_prims", C.c_int32), ("n_ops", C.c_int32), ("n_mats", C.c_int32),
        ("prim_type", C.POINTER(C.c_int32)),
        ("prim_pos", C.POINTER(C.c_float)), ("prim_dir", C.POINTER(C.c_float)),
        ("prim_res", C.POINTER(C.c_float)), ("prim_color", C.POINTER(C.c_float)),
        ("prim_mat", C.POINTER(C.c_int32)),
        ("op_type", C.POINTER(C.c_int32)), ("op_left", C.POINTER(C.c_int32)),
        ("op_right", C.POINTER(C.c_int32)), ("op_kind", C.POINTER(C.c_int32)),
        ("op_res", C.POINTER(C.c_float)),
        ("mats_bwd", C.POINTER(C.c_float)), ("mats_fwd", C.POINTER(C.c_float)),
    ]


def ref_lib_path(n_nodes: int = 128) -> str:
    if n_nodes <= 128:
        name = "libpsref.so"
    elif n_nodes <= 1024:
        name = "libpsref_w1k.so"
    else:
        name = "libpsref_w16k.so"
    return os.path.join(HERE, "_ref", name)


def ref_available(n_nodes: int = 128) -> bool:
    return os.path.exists(ref_lib_path(n_nodes))


_libs = {}


def _load_ref(path):
    if path in _libs:
        return _libs[path]
    lib = C.CDLL(path)
    lib.psref_scene_create.restype = C.c_void_p
    lib.psref_scene_create.argtypes = [C.POINTER(_RefDesc)]
    lib.psref_scene_destroy.argtypes = [C.c_void_p]
    lib.psref_prepare_bboxes.argtypes = [C.c_void_p, C.c_float]
    lib.psref_get_bbox.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.psref_set_bbox.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.psref_get_prim_boxes.argtypes = [C.c_void_p, C.c_void_p]
    lib.psref_get_op_boxes.argtypes = [C.c_void_p, C.c_void_p]
    lib.psref_pad_op_boxes.argtypes = [C.c_void_p, C.c_float]
    lib.psref_count_mpus.restype = C.c_uint32
    lib.psref_count_mpus.argtypes = [C.c_void_p, C.c_float]
    lib.psref_field.argtypes = [C.c_void_p, C.c_int64, C.c_void_p, C.c_void_p]
    lib.psref_field_color_normal.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 4
    if hasattr(lib, "psref_newton"):
        lib.psref_newton.argtypes = [C.c_void_p, C.c_int64] + [C.c_void_p] * 4 + [C.c_float, C.c_int] + [C.c_void_p] * 3
    lib.psref_polygonize.argtypes = [C.c_void_p, C.c_float, C.c_void_p]
    lib.psref_mpu_info.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
    lib.psref_mpu_copy.argtypes = [C.c_void_p, C.c_uint32] + [C.c_void_p] * 4
    lib.psref_mesh_concat.argtypes = [C.c_void_p] + [C.c_void_p] * 4
    lib.psref_polygonize_stream.argtypes = [C.c_void_p, C.c_float, C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
    lib.psref_stream_counts.restype = C.c_int64
    lib.psref_stream_counts.argtypes = [C.c_void_p, C.c_void_p]
    lib.psref_stream_copy.argtypes = [C.c_void_p] + [C.c_void_p] * 6
    lib.psref_set_threads.argtypes = [C.c_int]
    lib.psref_build_info.argtypes = [C.c_void_p]
    _libs[path] = lib
    return lib


class RefScene:
    """A scene held by the reference's own SOA structs (PS_Polygonizer.h:96-172)."""

    def __init__(self, soa: dict, lib_path: str | None = None):
        n = max(soa["n_prims"], soa["n_ops"])
        self.lib = _load_ref(lib_path or ref_lib_path(n))
        self._keep = {k: np.ascontiguousarray(v) for k, v in soa.items() if isinstance(v, np.ndarray)}
        k = self._keep
        d = _RefDesc(soa["n_prims"], soa["n_ops"], soa["n_mats"],
                     _p(k["prim_type"], C.c_int32), _p(k["prim_pos"], C.c_float), _p(k["prim_dir"], C.c_float),
                     _p(k["prim_res"], C.c_float), _p(k["prim_color"], C.c_float), _p(k["prim_mat"], C.c_int32),
                     _p(k["op_type"], C.c_int32), _p(k["op_left"], C.c_int32), _p(k["op_right"], C.c_int32),
                     _p(k["op_kind"], C.c_int32), _p(k["op_res"], C.c_float),
                     _p(k["mats_bwd"], C.c_float), _p(k["mats_fwd"], C.c_float))
        self.h = self.lib.psref_scene_create(C.byref(d))
        if not self.h:
            raise RuntimeError("psref_scene_create failed (capacity or node type)")
        self.n_prims, self.n_ops = soa["n_prims"], soa["n_ops"]

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.psref_scene_destroy(self.h)
            self.h = None

    def build_info(self):
        out = np.zeros(16, np.int64)
        self.lib.psref_build_info(out.ctypes.data)
        return dict(max_tree_nodes=int(out[0]), max_mpu_count=int(out[1]), sizeof_mpu=int(out[2]),
                    sizeof_polympus=int(out[3]), simd_flen=int(out[4]), grid_dim=int(out[5]),
                    sizeof_prims=int(out[6]), sizeof_ops=int(out[7]), sizeof_primmats=int(out[8]),
                    sizeof_boxmats=int(out[9]))

    def set_threads(self, n):
        self.lib.psref_set_threads(int(n))

    def prepare_bboxes(self, cellsize):
        return self.lib.psref_prepare_bboxes(self.h, float(cellsize))

    def bbox(self):
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        self.lib.psref_get_bbox(self.h, lo.ctypes.data, hi.ctypes.data)
        return lo, hi

    def set_bbox(self, lo, hi):
        lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
        self.lib.psref_set_bbox(self.h, lo.ctypes.data, hi.ctypes.data)

    def prim_boxes(self):
        out = np.zeros((self.n_prims, 6), np.float32)
        self.lib.psref_get_prim_boxes(self.h, out.ctypes.data)
        return out

    def op_boxes(self):
        out = np.zeros((max(self.n_ops, 1), 6), np.float32)
        self.lib.psref_get_op_boxes(self.h, out.ctypes.data)
        return out[:self.n_ops]

    def pad_op_boxes(self, pad):
        self.lib.psref_pad_op_boxes(self.h, float(pad))

    def count_mpus(self, cellsize):
        return int(self.lib.psref_count_mpus(self.h, float(cellsize)))

    def field(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self.lib.psref_field(self.h, len(xyz), xyz.ctypes.data, out.ctypes.data)
        return out

    def newton(self, p1, p2, f1, f2, target=0.5, iterations=8):
        """FieldComputer::computeRootNewtonRaphson per point pair: (root, field at the root, the member's return value)."""
        p1 = np.ascontiguousarray(p1, np.float32).reshape(-1, 3)
        p2 = np.ascontiguousarray(p2, np.float32).reshape(-1, 3)
        f1, f2 = np.ascontiguousarray(f1, np.float32), np.ascontiguousarray(f2, np.float32)
        n = len(p1)
        root, fr, ev = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
        self.lib.psref_newton(self.h, n, p1.ctypes.data, p2.ctypes.data, f1.ctypes.data, f2.ctypes.data, float(target), int(iterations),
                    root.ctypes.data, fr.ctypes.data, ev.ctypes.data)
        return root, fr, ev

    def field_color_normal(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        n = len(xyz)
        f, col, nrm = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        self.lib.psref_field_color_normal(self.h, n, xyz.ctypes.data, f.ctypes.data, col.ctypes.data, nrm.ctypes.data)
        return f, col, nrm

    def polygonize(self, cellsize):
        """Whole-grid reference Polygonize. Returns (ret, stats dict)."""
        st = np.zeros(8, np.int64)
        ret = self.lib.psref_polygonize(self.h, float(cellsize), st.ctypes.data)
        return ret, dict(mpus=int(st[0]), crossed=int(st[1]), vertices=int(st[2]), triangles=int(st[3]),
                         simd_evals=int(st[4]))

    def mpu(self, idx):
        info, lo = np.zeros(3, np.int32), np.zeros(3, np.float32)
        self.lib.psref_mpu_info(self.h, idx, info.ctypes.data, lo.ctypes.data)
        nv, nt = int(info[0]), int(info[1])
        pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
        tri = np.zeros((nt, 3), np.uint16)
        if nv or nt:
            self.lib.psref_mpu_copy(self.h, idx, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data)
        return dict(lo=lo, nv=nv, nt=nt, field_evals=int(info[2]), pos=pos, nrm=nrm, col=col, tri=tri)

    def mesh(self, stats):
        nv, nt = stats["vertices"], stats["triangles"]
        pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
        tri = np.zeros((nt, 3), np.uint32)
        self.lib.psref_mesh_concat(self.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data)
        return pos, nrm, col, tri

    def polygonize_stream(self, cellsize, b0=None, e0=None, keep_mesh=False):
        st = np.zeros(8, np.int64)
        b = np.ascontiguousarray(b0, np.int32) if b0 is not None else None
        e = np.ascontiguousarray(e0, np.int32) if e0 is not None else None
        ret = self.lib.psref_polygonize_stream(self.h, float(cellsize), b.ctypes.data if b is not None else None,
                                               e.ctypes.data if e is not None else None, int(keep_mesh), st.ctypes.data)
        stats = dict(mpus=int(st[0]), crossed=int(st[1]), vertices=int(st[2]), triangles=int(st[3]),
                     simd_evals=int(st[4]), lattice=(int(st[5]), int(st[6]), int(st[7])))
        if not keep_mesh:
            return ret, stats
        c = np.zeros(3, np.int64)
        self.lib.psref_stream_counts(self.h, c.ctypes.data)
        nv, nt, nm = int(c[0]), int(c[1]), int(c[2])
        pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
        tri = np.zeros((nt, 3), np.uint32)
        midx, mcnt = np.zeros((nm, 3), np.int32), np.zeros((nm, 2), np.uint32)
        self.lib.psref_stream_copy(self.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data,
                                   midx.ctypes.data, mcnt.ctypes.data)
        stats.update(pos=pos, nrm=nrm, col=col, tri=tri, mpu_idx=midx, mpu_counts=mcnt)
        return ret, stats


# ----------------------------------------------------------------------------------------------
# plain-C restatement (oracle/ps_oracle.c)
# ----------------------------------------------------------------------------------------------
class _SoaDesc(_RefDesc):
    pass


class _Tree(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int32), ("root", C.c_int32),
        ("type", C.POINTER(C.c_int32)), ("first_child", C.POINTER(C.c_int32)),
        ("n_children", C.POINTER(C.c_int32)), ("child", C.POINTER(C.c_int32)),
        ("pos", C.POINTER(C.c_float)), ("dir", C.POINTER(C.c_float)), ("res", C.POINTER(C.c_float)),
        ("color", C.POINTER(C.c_float)), ("xform", C.POINTER(C.c_float)), ("has_xform", C.POINTER(C.c_uint8)),
        ("srt", C.POINTER(C.c_float)), ("has_srt", C.POINTER(C.c_uint8)), ("node_id", C.POINTER(C.c_int32)),
        ("op_box_lo", C.POINTER(C.c_float)), ("op_box_hi", C.POINTER(C.c_float)),
    ]


def make_tree_struct(flat: dict):
    """psgpu_tree (include/psgpu_types.h) over the arrays of blobtree.flatten(); returns (struct, keepalive)."""
    k = {n: np.ascontiguousarray(v) for n, v in flat.items() if isinstance(v, np.ndarray)}
    t = _Tree(flat["n_nodes"], flat["root"], _p(k["type"], C.c_int32), _p(k["first_child"], C.c_int32),
              _p(k["n_children"], C.c_int32), _p(k["child"], C.c_int32), _p(k["pos"], C.c_float),
              _p(k["dir"], C.c_float), _p(k["res"], C.c_float), _p(k["color"], C.c_float),
              _p(k["xform"], C.c_float), _p(k["has_xform"], C.c_uint8))
    return t, k


def oracle_lib_path():
    return os.path.join(HERE, "libps_oracle.so")


_olib = None


def _load_oracle():
    global _olib
    if _olib is not None:
        return _olib
    lib = C.CDLL(oracle_lib_path())
    vp = C.c_void_p
    lib.pso_simd_create.restype = vp
    lib.pso_simd_create.argtypes = [C.POINTER(_RefDesc)]
    lib.pso_simd_destroy.argtypes = [vp]
    lib.pso_simd_prepare_bboxes.argtypes = [vp, C.c_float]
    lib.pso_simd_get_bbox.argtypes = [vp, vp, vp]
    lib.pso_simd_set_bbox.argtypes = [vp, vp, vp]
    lib.pso_simd_get_prim_boxes.argtypes = [vp, vp]
    lib.pso_simd_get_op_boxes.argtypes = [vp, vp]
    lib.pso_simd_pad_op_boxes.argtypes = [vp, C.c_float]
    lib.pso_count_mpus.restype = C.c_uint32
    lib.pso_count_mpus.argtypes = [C.c_float, vp, vp]
    lib.pso_simd_field.argtypes = [vp, C.c_int64, vp, vp]
    lib.pso_simd_field_color_normal.argtypes = [vp, C.c_int64, vp, vp, vp, vp]
    lib.pso_simd_newton.argtypes = [vp, C.c_int64, vp, vp, vp, vp, C.c_float, C.c_int, vp, vp, vp]
    lib.pso_simd_mpu_cache.argtypes = [vp, C.c_float, vp, vp, vp]
    lib.pso_simd_polygonize.argtypes = [vp, C.c_float, C.c_int, vp, vp, C.c_int, vp]
    lib.pso_simd_result_counts.argtypes = [vp, vp]
    lib.pso_simd_result_copy.argtypes = [vp] * 7
    lib.pso_scalar_create.restype = vp
    lib.pso_scalar_create.argtypes = [C.POINTER(_Tree)]
    lib.pso_scalar_destroy.argtypes = [vp]
    lib.pso_scalar_set_prim_boxes.argtypes = [vp, vp, vp]
    lib.pso_scalar_field.argtypes = [vp, C.c_int64, vp, vp]
    lib.pso_scalar_field_color_normal.argtypes = [vp, C.c_int64, vp, vp, vp, vp]
    lib.pso_scalar_newton.argtypes = [vp, C.c_int64, vp, vp, vp, vp, C.c_float, C.c_int, vp, vp, vp]
    lib.pso_scalar_mpu_cache.argtypes = [vp, vp, C.c_float, vp, vp, vp]
    lib.pso_scalar_polygonize.argtypes = [vp, vp, vp, C.c_float, C.c_int, vp, vp, C.c_int, vp]
    lib.pso_scalar_result_counts.argtypes = [vp, vp]
    lib.pso_scalar_result_copy.argtypes = [vp] * 7
    _olib = lib
    return lib


def _stats(st):
    return dict(mpus=int(st[0]), crossed=int(st[1]), vertices=int(st[2]), triangles=int(st[3]),
                simd_evals=int(st[4]), lattice=(int(st[5]), int(st[6]), int(st[7])), over_capacity=int(st[8]))


def _fetch_mesh(count_fn, copy_fn, h, stats):
    c = np.zeros(3, np.int64)
    count_fn(h, c.ctypes.data)
    nv, nt, nm = int(c[0]), int(c[1]), int(c[2])
    pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
    tri = np.zeros((nt, 3), np.uint32)
    midx, mcnt = np.zeros((nm, 3), np.int32), np.zeros((nm, 2), np.uint32)
    copy_fn(h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data, midx.ctypes.data, mcnt.ctypes.data)
    stats.update(pos=pos, nrm=nrm, col=col, tri=tri, mpu_idx=midx, mpu_counts=mcnt)


class SimdOracle:
    """PS_SimdPoly semantics restated in C (binary SOA input, like RefScene)."""

    def __init__(self, soa: dict):
        self.lib = _load_oracle()
        self._keep = {k: np.ascontiguousarray(v) for k, v in soa.items() if isinstance(v, np.ndarray)}
        k = self._keep
        d = _RefDesc(soa["n_prims"], soa["n_ops"], soa["n_mats"],
                     _p(k["prim_type"], C.c_int32), _p(k["prim_pos"], C.c_float), _p(k["prim_dir"], C.c_float),
                     _p(k["prim_res"], C.c_float), _p(k["prim_color"], C.c_float), _p(k["prim_mat"], C.c_int32),
                     _p(k["op_type"], C.c_int32), _p(k["op_left"], C.c_int32), _p(k["op_right"], C.c_int32),
                     _p(k["op_kind"], C.c_int32), _p(k["op_res"], C.c_float),
                     _p(k["mats_bwd"], C.c_float), _p(k["mats_fwd"], C.c_float))
        self.h = self.lib.pso_simd_create(C.byref(d))
        if not self.h:
            raise RuntimeError("pso_simd_create failed")
        self.n_prims, self.n_ops = soa["n_prims"], soa["n_ops"]

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.pso_simd_destroy(self.h)
            self.h = None

    def prepare_bboxes(self, cellsize):
        return self.lib.pso_simd_prepare_bboxes(self.h, float(cellsize))

    def bbox(self):
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        self.lib.pso_simd_get_bbox(self.h, lo.ctypes.data, hi.ctypes.data)
        return lo, hi

    def set_bbox(self, lo, hi):
        lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
        self.lib.pso_simd_set_bbox(self.h, lo.ctypes.data, hi.ctypes.data)

    def prim_boxes(self):
        out = np.zeros((self.n_prims, 6), np.float32)
        self.lib.pso_simd_get_prim_boxes(self.h, out.ctypes.data)
        return out

    def op_boxes(self):
        out = np.zeros((max(self.n_ops, 1), 6), np.float32)
        self.lib.pso_simd_get_op_boxes(self.h, out.ctypes.data)
        return out[:self.n_ops]

    def pad_op_boxes(self, pad):
        self.lib.pso_simd_pad_op_boxes(self.h, float(pad))

    def count_mpus(self, cellsize):
        lo, hi = self.bbox()
        return int(self.lib.pso_count_mpus(float(cellsize), lo.ctypes.data, hi.ctypes.data))

    def field(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self.lib.pso_simd_field(self.h, len(xyz), xyz.ctypes.data, out.ctypes.data)
        return out

    def newton(self, p1, p2, f1, f2, target=0.5, iterations=8):
        """FieldComputer::computeRootNewtonRaphson per point pair: (root, field at the root, the member's return value)."""
        p1 = np.ascontiguousarray(p1, np.float32).reshape(-1, 3)
        p2 = np.ascontiguousarray(p2, np.float32).reshape(-1, 3)
        f1, f2 = np.ascontiguousarray(f1, np.float32), np.ascontiguousarray(f2, np.float32)
        n = len(p1)
        root, fr, ev = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
        self.lib.pso_simd_newton(self.h, n, p1.ctypes.data, p2.ctypes.data, f1.ctypes.data, f2.ctypes.data, float(target), int(iterations),
                    root.ctypes.data, fr.ctypes.data, ev.ctypes.data)
        return root, fr, ev

    def field_color_normal(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        n = len(xyz)
        f, col, nrm = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        self.lib.pso_simd_field_color_normal(self.h, n, xyz.ctypes.data, f.ctypes.data, col.ctypes.data, nrm.ctypes.data)
        return f, col, nrm

    def mpu_cache(self, cellsize, ijk):
        ijk = np.ascontiguousarray(ijk, np.int32)
        fv, cs = np.zeros(512, np.float32), np.zeros(343, np.uint8)
        self.lib.pso_simd_mpu_cache(self.h, float(cellsize), ijk.ctypes.data, fv.ctypes.data, cs.ctypes.data)
        return fv, cs

    def polygonize(self, cellsize, nthreads=0, b0=None, e0=None, keep_mesh=True):
        st = np.zeros(16, np.int64)
        b = np.ascontiguousarray(b0, np.int32) if b0 is not None else None
        e = np.ascontiguousarray(e0, np.int32) if e0 is not None else None
        ret = self.lib.pso_simd_polygonize(self.h, float(cellsize), int(nthreads or os.cpu_count() or 1),
                                           b.ctypes.data if b is not None else None,
                                           e.ctypes.data if e is not None else None, int(keep_mesh), st.ctypes.data)
        stats = _stats(st)
        if keep_mesh:
            _fetch_mesh(self.lib.pso_simd_result_counts, self.lib.pso_simd_result_copy, self.h, stats)
        return ret, stats


class ScalarOracle:
    """Scalar-twin semantics (COMPACTBLOBTREE / CParsipOptimized) restated in C; n-ary flat tree input."""

    def __init__(self, flat: dict):
        self.lib = _load_oracle()
        t, self._keep = make_tree_struct(flat)
        self.h = self.lib.pso_scalar_create(C.byref(t))
        if not self.h:
            raise RuntimeError("pso_scalar_create failed")

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.pso_scalar_destroy(self.h)
            self.h = None

    def set_prim_boxes(self, lo, hi):
        """Per-node octrees for the scalar polygonizer's MPU filter (CPolyParsipOptimized.cpp:160-183); None switches it off."""
        if lo is None:
            self.lib.pso_scalar_set_prim_boxes(self.h, None, None)
            return
        lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
        self.lib.pso_scalar_set_prim_boxes(self.h, lo.ctypes.data, hi.ctypes.data)

    def field(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self.lib.pso_scalar_field(self.h, len(xyz), xyz.ctypes.data, out.ctypes.data)
        return out

    def field_color_normal(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        n = len(xyz)
        f, col, nrm = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        self.lib.pso_scalar_field_color_normal(self.h, n, xyz.ctypes.data, f.ctypes.data, col.ctypes.data, nrm.ctypes.data)
        return f, col, nrm

    def newton(self, p1, p2, f1, f2, iso=0.5, iterations=8):
        p1 = np.ascontiguousarray(p1, np.float32).reshape(-1, 3)
        p2 = np.ascontiguousarray(p2, np.float32).reshape(-1, 3)
        f1 = np.ascontiguousarray(f1, np.float32)
        f2 = np.ascontiguousarray(f2, np.float32)
        n = len(p1)
        root, fr, ev = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
        self.lib.pso_scalar_newton(self.h, n, p1.ctypes.data, p2.ctypes.data, f1.ctypes.data, f2.ctypes.data,
                                   float(iso), int(iterations), root.ctypes.data, fr.ctypes.data, ev.ctypes.data)
        return root, fr, ev

    def mpu_cache(self, lo, cellsize, ijk):
        lo = np.ascontiguousarray(lo, np.float32)
        ijk = np.ascontiguousarray(ijk, np.int32)
        fv, cs = np.zeros(512, np.float32), np.zeros(343, np.uint8)
        self.lib.pso_scalar_mpu_cache(self.h, lo.ctypes.data, float(cellsize), ijk.ctypes.data, fv.ctypes.data, cs.ctypes.data)
        return fv, cs

    def polygonize(self, lo, hi, cellsize, nthreads=0, b0=None, e0=None, keep_mesh=True):
        lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
        st = np.zeros(16, np.int64)
        b = np.ascontiguousarray(b0, np.int32) if b0 is not None else None
        e = np.ascontiguousarray(e0, np.int32) if e0 is not None else None
        ret = self.lib.pso_scalar_polygonize(self.h, lo.ctypes.data, hi.ctypes.data, float(cellsize),
                                             int(nthreads or os.cpu_count() or 1),
                                             b.ctypes.data if b is not None else None,
                                             e.ctypes.data if e is not None else None, int(keep_mesh), st.ctypes.data)
        stats = _stats(st)
        if keep_mesh:
            _fetch_mesh(self.lib.pso_scalar_result_counts, self.lib.pso_scalar_result_copy, self.h, stats)
        return ret, stats


# ----------------------------------------------------------------------------------------------
# the reference's own scalar twin (oracle/_ref/libcompactref.so, oracle/compact_wrap.cpp)
# ----------------------------------------------------------------------------------------------
class _CrefDesc(C.Structure):
    _fields_ = [
        ("n_nodes", C.c_int32),
        ("type", C.POINTER(C.c_int32)), ("first_child", C.POINTER(C.c_int32)),
        ("n_children", C.POINTER(C.c_int32)), ("child", C.POINTER(C.c_int32)),
        ("pos", C.POINTER(C.c_float)), ("dir", C.POINTER(C.c_float)), ("res", C.POINTER(C.c_float)),
        ("color", C.POINTER(C.c_float)), ("srt", C.POINTER(C.c_float)), ("has_srt", C.POINTER(C.c_uint8)),
    ]


def compactref_lib_path():
    return os.path.join(HERE, "_ref", "libcompactref.so")


def compactref_available():
    return os.path.exists(compactref_lib_path())


_clib = None


def _load_compactref():
    global _clib
    if _clib is not None:
        return _clib
    lib = C.CDLL(compactref_lib_path())
    vp = C.c_void_p
    lib.cref_create.restype = vp
    lib.cref_create.argtypes = [C.POINTER(_CrefDesc)]
    lib.cref_destroy.argtypes = [vp]
    lib.cref_convert_rc.argtypes = [vp]
    lib.cref_walk_rc.argtypes = [vp]
    lib.cref_counts.argtypes = [vp, vp]
    lib.cref_walk_copy.argtypes = [vp] * 13
    lib.cref_octree.argtypes = [vp, vp, vp]
    lib.cref_field.argtypes = [vp, C.c_int64, vp, vp]
    lib.cref_field_color_normal.argtypes = [vp, C.c_int64, vp, vp, vp, vp]
    lib.cref_newton.argtypes = [vp, C.c_int64, vp, vp, vp, vp, C.c_float, C.c_int, vp, vp, vp]
    lib.cref_scene_field.argtypes = [vp, C.c_int64, vp, vp]
    lib.cref_polygonize.argtypes = [vp, vp, vp, C.c_float, C.c_float, C.c_int]
    lib.cref_result_counts.argtypes = [vp, vp]
    lib.cref_result_copy.argtypes = [vp] * 7
    lib.cref_export_mesh.argtypes = [vp] * 6
    lib.cref_walk_srt.argtypes = [vp, vp, vp]
    lib.cref_flatten_transforms.argtypes = [vp]
    lib.cref_recompute_octrees.argtypes = [vp]
    lib.cref_save_scene.argtypes = [vp, C.c_char_p]
    _clib = lib
    return lib


class CompactRef:
    """The reference's OWN scalar path on a tree built with its scene-API classes: COMPACTBLOBTREE::convert / fieldvalue /
    normal / baseColor, ComputeRootNewtonRaphsonVEC4, CBlobNode::fieldValue, CParsipOptimized (all compiled unmodified from
    /root/reference). `flat` is blobtree.flatten() output; nodes may carry `srt` (scale, quaternion xyzw, translate), which the
    reference turns into its backward matrix itself. `walk()` returns the tree as the PRODUCT walker
    include/psgpu_blobtree_walk.h reads it back from the CBlobNode objects."""

    def __init__(self, flat: dict):
        self.lib = _load_compactref()
        k = {n: np.ascontiguousarray(v) for n, v in flat.items() if isinstance(v, np.ndarray)}
        self._keep = k
        if flat.get("root", 0) != 0:
            raise ValueError("CompactRef expects the root at index 0 (blobtree.flatten order)")
        srt = k.get("srt")
        has = k.get("has_srt")
        d = _CrefDesc(flat["n_nodes"], _p(k["type"], C.c_int32), _p(k["first_child"], C.c_int32), _p(k["n_children"], C.c_int32),
                      _p(k["child"], C.c_int32), _p(k["pos"], C.c_float), _p(k["dir"], C.c_float), _p(k["res"], C.c_float),
                      _p(k["color"], C.c_float), _p(srt, C.c_float) if srt is not None else None,
                      _p(has, C.c_uint8) if has is not None else None)
        self.h = self.lib.cref_create(C.byref(d))
        if not self.h:
            raise RuntimeError("cref_create: a node kind the reference builder does not know")
        c = np.zeros(4, np.int32)
        self.lib.cref_counts(self.h, c.ctypes.data)
        self.n_prims, self.n_ops, self.n_nodes, self.n_child = (int(v) for v in c)
        self.convert_rc = self.lib.cref_convert_rc(self.h)
        self.walk_rc = self.lib.cref_walk_rc(self.h)

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.cref_destroy(self.h)
            self.h = None

    def walk(self):
        N = self.n_nodes
        out = dict(n_nodes=N, root=0, type=np.zeros(N, np.int32), first_child=np.zeros(N, np.int32), n_children=np.zeros(N, np.int32),
                   child=np.zeros(max(1, self.n_child), np.int32), pos=np.zeros((N, 3), np.float32), dir=np.zeros((N, 3), np.float32),
                   res=np.zeros((N, 4), np.float32), color=np.zeros((N, 3), np.float32), xform=np.zeros((N, 12), np.float32),
                   has_xform=np.zeros(N, np.uint8), oct_lo=np.zeros((N, 3), np.float32), oct_hi=np.zeros((N, 3), np.float32))
        self.lib.cref_walk_copy(self.h, *(out[k].ctypes.data for k in ("type", "first_child", "n_children", "child", "pos", "dir", "res",
                                                                        "color", "xform", "has_xform", "oct_lo", "oct_hi")))
        out["srt"], out["node_id"] = np.zeros((N, 10), np.float32), np.zeros(N, np.int32)
        self.lib.cref_walk_srt(self.h, out["srt"].ctypes.data, out["node_id"].ctypes.data)
        out["has_srt"] = np.ones(N, np.uint8)
        return out

    def flatten_transforms(self):
        """CLayer::recursive_FlattenTransformations with the reference's CAffineTransformation::add, in place on the CBlobNode tree."""
        self.lib.cref_flatten_transforms(self.h)

    def recompute_octrees(self):
        """CLayer::recursive_RecomputeAllOctrees with the reference's CMatrix / COctree, in place."""
        self.lib.cref_recompute_octrees(self.h)

    def save_scene(self, path):
        """CBlobNode::saveScript of the root (the reference's own .scene writer)."""
        return self.lib.cref_save_scene(self.h, os.fsencode(path))

    def octree(self):
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        self.lib.cref_octree(self.h, lo.ctypes.data, hi.ctypes.data)
        return lo, hi

    def field(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self.lib.cref_field(self.h, len(xyz), xyz.ctypes.data, out.ctypes.data)
        return out

    def scene_field(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self.lib.cref_scene_field(self.h, len(xyz), xyz.ctypes.data, out.ctypes.data)
        return out

    def field_color_normal(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        n = len(xyz)
        f, col, nrm = np.zeros(n, np.float32), np.zeros((n, 4), np.float32), np.zeros((n, 3), np.float32)
        self.lib.cref_field_color_normal(self.h, n, xyz.ctypes.data, f.ctypes.data, col.ctypes.data, nrm.ctypes.data)
        return f, col, nrm

    def newton(self, p1, p2, f1, f2, iso=0.5, iterations=8):
        p1 = np.ascontiguousarray(p1, np.float32).reshape(-1, 3)
        p2 = np.ascontiguousarray(p2, np.float32).reshape(-1, 3)
        f1 = np.ascontiguousarray(f1, np.float32)
        f2 = np.ascontiguousarray(f2, np.float32)
        n = len(p1)
        root, fr, ev = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
        self.lib.cref_newton(self.h, n, p1.ctypes.data, p2.ctypes.data, f1.ctypes.data, f2.ctypes.data, float(iso), int(iterations),
                             root.ctypes.data, fr.ctypes.data, ev.ctypes.data)
        return root, fr, ev

    def polygonize(self, lo, hi, cellsize, iso=0.5, nthreads=1):
        """CParsipOptimized::setup(box) + run(). Returns the per-MPU meshes of the MPUs that kept surface, in list order."""
        lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
        self.lib.cref_polygonize(self.h, lo.ctypes.data, hi.ctypes.data, float(cellsize), float(iso), int(nthreads))
        c = np.zeros(4, np.int64)
        self.lib.cref_result_counts(self.h, c.ctypes.data)
        nm, nv, nt = int(c[0]), int(c[1]), int(c[2])
        pos, nrm, col = np.zeros((nv, 3), np.float32), np.zeros((nv, 3), np.float32), np.zeros((nv, 4), np.float32)
        tri = np.zeros((nt, 3), np.uint32)
        org, cnt = np.zeros((nm, 3), np.float32), np.zeros((nm, 2), np.int32)
        self.lib.cref_result_copy(self.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data, org.ctypes.data, cnt.ctypes.data)
        return dict(mpus=nm, vertices=nv, triangles=nt, field_evals=int(c[3]), pos=pos, nrm=nrm, col=col, tri=tri, mpu_origin=org, mpu_counts=cnt)

    def export_mesh(self):
        """CParsipOptimized::exportMesh of the last polygonize()."""
        c = np.zeros(2, np.int64)
        self.lib.cref_export_mesh(self.h, None, None, None, None, c.ctypes.data)
        nv, nt = int(c[0]), int(c[1])
        pos, nrm, col = np.zeros((nv, 3), np.float32), np.zeros((nv, 3), np.float32), np.zeros((nv, 4), np.float32)
        tri = np.zeros((nt, 3), np.uint32)
        ok = self.lib.cref_export_mesh(self.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data, c.ctypes.data)
        return ok, dict(vertices=nv, triangles=nt, pos=pos, nrm=nrm, col=col, tri=tri)
