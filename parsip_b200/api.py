"""ctypes binding of libpsgpu.so (include/psgpu.h). Thin: arrays in, arrays out, status codes raised."""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

SEM_SIMD, SEM_SCALAR = 0, 1
ROOT_FOURSAMPLE, ROOT_NEWTON_REF, ROOT_BISECT = 0, 1, 2
NOALLOC_MAX_POINTS = 65536  # PSGPU_NOALLOC_MAX_POINTS
COORD_MPU_LOCAL, COORD_GLOBAL = 0, 1
OUT_INDEXED, OUT_MPU = 0, 1
CULL_NONE, CULL_SUPPORT, CULL_INTERVAL = 0, 1, 2
LIN_REASSOCIATE = 1
LIN_FULL_AFFINE, LIN_AS_SHIPPED = 2, 4

SUCCESS = 1
ERRORS = {-1: "PARAM_ERROR", -2: "NOT_ENOUGH_MEM", -3: "INVALID_BVH", -4: "UNSUPPORTED_NODE", -5: "CUDA", -6: "NO_DEVICE"}

# every symbol include/psgpu.h declares (tests/test_abi.py checks the library exports them all)
EXPORTS = [
    "psgpu_create", "psgpu_destroy", "psgpu_last_error", "psgpu_device_count", "psgpu_linearize", "psgpu_program_free",
    "psgpu_program_info", "psgpu_program_copy", "psgpu_program_colors", "psgpu_prepare_bboxes", "psgpu_count_mpus", "psgpu_upload_program",
    "psgpu_polygonize", "psgpu_polygonize_host", "psgpu_mesh_download", "psgpu_mesh_map_host", "psgpu_field_eval", "psgpu_field_color_normal", "psgpu_newton_roots", "psgpu_field_eval_noalloc",
    "psgpu_tree_binarize", "psgpu_tree_view", "psgpu_tree_free", "psgpu_tree_flatten_transforms", "psgpu_tree_octrees", "psgpu_scene_load", "psgpu_scene_save", "psgpu_mpu_cache", "psgpu_mpu_caches_culled", "psgpu_prune_region", "psgpu_get_stats", "psgpu_measure_fp32_peak",
    "psgpu_comm_get_unique_id", "psgpu_comm_init", "psgpu_comm_destroy", "psgpu_slab_bounds", "psgpu_polygonize_sharded", "psgpu_polygonize_sharded_host", "psgpu_mesh_gather", "psgpu_layer_work",
]


class PsGpuError(RuntimeError):
    def __init__(self, code, what=""):
        self.code = code
        super().__init__("psgpu error %s (%d) %s" % (ERRORS.get(code, "?"), code, what))


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


class Params(C.Structure):
    _fields_ = [
        ("cellsize", C.c_float), ("bbox_lo", C.c_float * 3), ("bbox_hi", C.c_float * 3), ("iso", C.c_float),
        ("root_mode", C.c_int32), ("root_iterations", C.c_int32), ("coord_mode", C.c_int32), ("output", C.c_int32),
        ("cull", C.c_int32), ("z_begin", C.c_int32), ("z_end", C.c_int32), ("want_colors", C.c_int32),
        ("want_normals", C.c_int32), ("collect_stats", C.c_int32),
    ]

    @classmethod
    def make(cls, cellsize, lo, hi, root_mode=ROOT_BISECT, coord_mode=COORD_GLOBAL, output=OUT_INDEXED, cull=CULL_INTERVAL,
             z_begin=0, z_end=0, want_colors=1, want_normals=1, collect_stats=0, root_iterations=0, iso=0.0):
        p = cls()
        p.cellsize = float(cellsize)
        for a in range(3):
            p.bbox_lo[a] = float(lo[a])
            p.bbox_hi[a] = float(hi[a])
        p.iso = iso
        p.root_mode, p.root_iterations, p.coord_mode, p.output, p.cull = root_mode, root_iterations, coord_mode, output, cull
        p.z_begin, p.z_end, p.want_colors, p.want_normals, p.collect_stats = z_begin, z_end, want_colors, want_normals, collect_stats
        return p


class _Mesh(C.Structure):
    _fields_ = [
        ("n_vertices", C.c_int64), ("n_triangles", C.c_int64),
        ("pos", C.c_void_p), ("nrm", C.c_void_p), ("col", C.c_void_p), ("tri", C.c_void_p),
        ("n_mpus_out", C.c_int64), ("mpu_ijk", C.c_void_p), ("mpu_vbase", C.c_void_p), ("mpu_tbase", C.c_void_p),
        ("lattice", C.c_int32 * 3), ("over_capacity", C.c_int32),
    ]


class _Stats(C.Structure):
    _fields_ = [
        ("mpus_total", C.c_int64), ("mpus_evaluated", C.c_int64), ("mpus_crossed", C.c_int64), ("field_evals", C.c_int64),
        ("prim_evals", C.c_int64), ("algorithmic_flops", C.c_int64), ("program_quads", C.c_int64), ("pruned_quads", C.c_int64),
        ("ms_cull", C.c_float), ("ms_field", C.c_float), ("ms_scan", C.c_float), ("ms_emit", C.c_float), ("ms_total", C.c_float),
        ("gpu_launches", C.c_int64), ("ms_tri", C.c_float), ("ms_vertex", C.c_float), ("ms_color", C.c_float), ("ms_exchange", C.c_float),
        ("vertex_eval_flops", C.c_int64), ("unresolved_edges", C.c_int64),
    ]


class _Shard(C.Structure):
    _fields_ = [
        ("nranks", C.c_int32), ("rank", C.c_int32), ("z_begin", C.c_int32), ("z_end", C.c_int32),
        ("vertex_offset", C.c_int64), ("triangle_offset", C.c_int64), ("total_vertices", C.c_int64), ("total_triangles", C.c_int64),
        ("ms_exchange", C.c_float), ("stitched", C.c_int32),
    ]


def lib_path() -> str:
    # PSGPU_LIB: development aid for A/B runs of two builds of the same library (never a fallback: the file must exist)
    return os.environ.get("PSGPU_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libpsgpu.so")


_lib = None


def load_library():
    """Loads libpsgpu.so; raises (never falls back) if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    path = lib_path()
    if not os.path.exists(path):
        raise PsGpuError(-6, "libpsgpu.so is not built: run `python -c 'import __graft_entry__ as g; g.build()'` (there is no CPU fallback)")
    lib = C.CDLL(path)
    vp = C.c_void_p
    lib.psgpu_create.argtypes = [C.c_int, C.POINTER(vp)]
    lib.psgpu_destroy.argtypes = [vp]
    lib.psgpu_last_error.restype = C.c_char_p
    lib.psgpu_last_error.argtypes = [vp]
    lib.psgpu_linearize.argtypes = [C.POINTER(_Tree), C.c_int, C.c_int, C.POINTER(vp)]
    lib.psgpu_program_free.argtypes = [vp]
    lib.psgpu_program_info.argtypes = [vp, vp]
    lib.psgpu_program_copy.argtypes = [vp, vp]
    lib.psgpu_program_colors.argtypes = [vp, vp]
    lib.psgpu_prepare_bboxes.argtypes = [C.POINTER(_Tree), C.c_float, vp, vp]
    lib.psgpu_count_mpus.restype = C.c_uint32
    lib.psgpu_count_mpus.argtypes = [C.c_float, vp, vp]
    lib.psgpu_lattice_dims.argtypes = [C.c_float, vp, vp, vp]
    lib.psgpu_upload_program.argtypes = [vp, vp]
    lib.psgpu_polygonize.argtypes = [vp, C.POINTER(Params), C.POINTER(_Mesh)]
    lib.psgpu_polygonize_host.argtypes = [vp, C.POINTER(Params), C.POINTER(_Mesh), vp, vp, vp, vp, C.c_int64, C.c_int64]
    lib.psgpu_mesh_download.argtypes = [vp] * 8
    lib.psgpu_mesh_map_host.argtypes = [vp] * 8
    lib.psgpu_tree_binarize.argtypes = [vp, C.c_int, C.c_int, vp, vp]
    lib.psgpu_tree_view.restype = vp
    lib.psgpu_tree_view.argtypes = [vp]
    lib.psgpu_tree_free.argtypes = [vp]
    lib.psgpu_tree_flatten_transforms.argtypes = [vp, vp]
    lib.psgpu_tree_octrees.argtypes = [vp, C.c_int, vp, vp]
    lib.psgpu_scene_load.argtypes = [C.c_char_p, C.c_int, vp]
    lib.psgpu_scene_save.argtypes = [vp, C.c_char_p]
    lib.psgpu_field_eval.argtypes = [vp, vp, C.c_int64, vp]
    lib.psgpu_field_eval_noalloc.argtypes = [vp, vp, C.c_int64, vp]
    lib.psgpu_newton_roots.argtypes = [vp, vp, vp, vp, vp, C.c_int64, C.c_float, C.c_int, vp, vp, vp]
    lib.psgpu_field_color_normal.argtypes = [vp, vp, C.c_int64, vp, vp, vp]
    lib.psgpu_mpu_cache.argtypes = [vp, C.POINTER(Params), vp, vp, vp]
    lib.psgpu_mpu_caches_culled.argtypes = [vp, C.POINTER(Params), vp, vp, C.c_int64, vp]
    lib.psgpu_prune_region.argtypes = [vp, vp, vp, C.c_int, C.c_float, C.c_int, vp, C.c_uint32, vp]
    lib.psgpu_get_stats.argtypes = [vp, C.POINTER(_Stats)]
    lib.psgpu_comm_get_unique_id.argtypes = [vp]
    lib.psgpu_comm_init.argtypes = [vp, C.c_int, C.c_int, vp]
    lib.psgpu_comm_destroy.argtypes = [vp]
    lib.psgpu_slab_bounds.argtypes = [C.c_int, C.c_int, C.c_int, vp, vp, vp]
    lib.psgpu_polygonize_sharded.argtypes = [vp, C.POINTER(Params), C.POINTER(_Mesh), C.POINTER(_Shard)]
    lib.psgpu_mesh_gather.argtypes = [vp, C.c_int, vp, vp, vp, vp]
    lib.psgpu_layer_work.argtypes = [vp, vp, vp, C.c_int32]
    lib.psgpu_polygonize_sharded_host.argtypes = [vp, C.POINTER(Params), C.POINTER(_Mesh), C.POINTER(_Shard), vp, vp, vp, vp, C.c_int64, C.c_int64]
    lib.psgpu_measure_fp32_peak.argtypes = [vp, vp, vp]
    _lib = lib
    return lib


def _tree_struct(flat: dict):
    k = {n: np.ascontiguousarray(v) for n, v in flat.items() if isinstance(v, np.ndarray)}

    def p(a, t):
        return a.ctypes.data_as(C.POINTER(t))

    t = _Tree(flat["n_nodes"], flat["root"], p(k["type"], C.c_int32), p(k["first_child"], C.c_int32), p(k["n_children"], C.c_int32),
              p(k["child"], C.c_int32), p(k["pos"], C.c_float), p(k["dir"], C.c_float), p(k["res"], C.c_float),
              p(k["color"], C.c_float), p(k["xform"], C.c_float), p(k["has_xform"], C.c_uint8))
    if "srt" in k and "has_srt" in k:
        k["srt"], k["has_srt"] = np.ascontiguousarray(k["srt"], np.float32), np.ascontiguousarray(k["has_srt"], np.uint8)
        t.srt, t.has_srt = p(k["srt"], C.c_float), p(k["has_srt"], C.c_uint8)
    if "node_id" in k:
        k["node_id"] = np.ascontiguousarray(k["node_id"], np.int32)
        t.node_id = p(k["node_id"], C.c_int32)
    if "op_box_lo" in k and "op_box_hi" in k:  # PSGPU_LIN_AS_SHIPPED: the operator boxes of the reference's PrepareBBoxes
        k["op_box_lo"], k["op_box_hi"] = np.ascontiguousarray(k["op_box_lo"], np.float32), np.ascontiguousarray(k["op_box_hi"], np.float32)
        t.op_box_lo, t.op_box_hi = p(k["op_box_lo"], C.c_float), p(k["op_box_hi"], C.c_float)
    return t, k


class Program:
    """A linearized BlobTree (host side)."""

    def __init__(self, handle, lib):
        self.h, self.lib = handle, lib
        info = np.zeros(5, np.int64)
        lib.psgpu_program_info(self.h, info.ctypes.data)
        self.n_quads, self.n_prims, self.max_value_stack, self.max_point_stack, self.n_instr = (int(v) for v in info)

    def quads(self) -> np.ndarray:
        out = np.zeros((self.n_quads, 4), np.uint32)
        self.lib.psgpu_program_copy(self.h, out.ctypes.data)
        return out

    def colors(self) -> np.ndarray:
        out = np.zeros((max(self.n_prims, 1), 3), np.float32)
        self.lib.psgpu_program_colors(self.h, out.ctypes.data)
        return out[:self.n_prims]

    def __del__(self):
        if getattr(self, "h", None):
            self.lib.psgpu_program_free(self.h)
            self.h = None


def linearize(flat: dict, semantics: int = SEM_SCALAR, flags: int = LIN_REASSOCIATE) -> Program:
    lib = load_library()
    t, keep = _tree_struct(flat)
    h = C.c_void_p()
    rc = lib.psgpu_linearize(C.byref(t), semantics, flags, C.byref(h))
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_linearize")
    return Program(h, lib)


def prepare_bboxes(flat: dict, cellsize: float = 0.14):
    lib = load_library()
    t, keep = _tree_struct(flat)
    lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
    rc = lib.psgpu_prepare_bboxes(C.byref(t), float(cellsize), lo.ctypes.data, hi.ctypes.data)
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_prepare_bboxes")
    return lo, hi


def binarize_tree(flat: dict, split_always: bool = True, pad_with_null: bool = True):
    """psgpu_tree_binarize (ConvertToBinaryTree, CLayerManager.cpp:81-233) on a flat tree -> (flat tree dict, n_converted)."""
    lib = load_library()
    t, keep = _tree_struct(flat)
    h, nconv = C.c_void_p(), C.c_int32(0)
    rc = lib.psgpu_tree_binarize(C.byref(t), int(split_always), int(pad_with_null), C.byref(h), C.byref(nconv))
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_tree_binarize")
    try:
        v = C.cast(lib.psgpu_tree_view(h), C.POINTER(_Tree)).contents
        N = v.n_nodes
        arr = lambda ptr, n, dt: np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt).copy() if n else np.zeros(0, dt)
        nk = arr(v.n_children, N, np.int32)
        out = dict(n_nodes=N, root=v.root, type=arr(v.type, N, np.int32), first_child=arr(v.first_child, N, np.int32), n_children=nk,
                   child=arr(v.child, max(1, int(nk.sum())), np.int32), pos=arr(v.pos, N * 3, np.float32).reshape(N, 3),
                   dir=arr(v.dir, N * 3, np.float32).reshape(N, 3), res=arr(v.res, N * 4, np.float32).reshape(N, 4),
                   color=arr(v.color, N * 3, np.float32).reshape(N, 3), xform=arr(v.xform, N * 12, np.float32).reshape(N, 12),
                   has_xform=arr(v.has_xform, N, np.uint8))
    finally:
        lib.psgpu_tree_free(h)
    return out, int(nconv.value)


def comm_unique_id() -> bytes:
    """psgpu_comm_get_unique_id: 128 bytes rank 0 hands to the other ranks."""
    lib = load_library()
    buf = C.create_string_buffer(128)
    rc = lib.psgpu_comm_get_unique_id(buf)
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_comm_get_unique_id (libnccl.so.2 missing?)")
    return buf.raw


def slab_bounds(n_layers: int, nranks: int, rank: int, weights=None):
    lib = load_library()
    b, e = C.c_int32(), C.c_int32()
    w = None if weights is None else np.ascontiguousarray(weights, np.float32)
    rc = lib.psgpu_slab_bounds(int(n_layers), int(nranks), int(rank), None if w is None else w.ctypes.data, C.byref(b), C.byref(e))
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_slab_bounds")
    return int(b.value), int(e.value)


def _tree_dict(lib, h) -> dict:
    """numpy copy of a library-owned psgpu_tree_buf"""
    v = C.cast(lib.psgpu_tree_view(h), C.POINTER(_Tree)).contents
    N = v.n_nodes
    arr = lambda ptr, n, dt: np.ctypeslib.as_array(ptr, shape=(n,)).astype(dt).copy() if n else np.zeros(0, dt)
    nk = arr(v.n_children, N, np.int32)
    fc = arr(v.first_child, N, np.int32)
    nchild = max(1, int((fc + nk).max()) if N else 1)
    out = dict(n_nodes=N, root=v.root, type=arr(v.type, N, np.int32), first_child=fc, n_children=nk,
               child=arr(v.child, nchild, np.int32), pos=arr(v.pos, N * 3, np.float32).reshape(N, 3),
               dir=arr(v.dir, N * 3, np.float32).reshape(N, 3), res=arr(v.res, N * 4, np.float32).reshape(N, 4),
               color=arr(v.color, N * 3, np.float32).reshape(N, 3), xform=arr(v.xform, N * 12, np.float32).reshape(N, 12),
               has_xform=arr(v.has_xform, N, np.uint8))
    if v.srt:
        out["srt"] = arr(v.srt, N * 10, np.float32).reshape(N, 10)
        out["has_srt"] = arr(v.has_srt, N, np.uint8) if v.has_srt else np.ones(N, np.uint8)
    if v.node_id:
        out["node_id"] = arr(v.node_id, N, np.int32)
    return out


def flatten_transforms(flat: dict) -> dict:
    """psgpu_tree_flatten_transforms (CLayer::flattenTransformations); the tree must carry srt / has_srt."""
    lib = load_library()
    t, keep = _tree_struct(flat)
    h = C.c_void_p()
    rc = lib.psgpu_tree_flatten_transforms(C.byref(t), C.byref(h))
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_tree_flatten_transforms")
    try:
        return _tree_dict(lib, h)
    finally:
        lib.psgpu_tree_free(h)


def tree_octrees(flat: dict, recompute: bool = False):
    """psgpu_tree_octrees: (lo, hi) boxes of all nodes as the scene API computes them; [root] is the grid domain."""
    lib = load_library()
    t, keep = _tree_struct(flat)
    N = flat["n_nodes"]
    lo, hi = np.zeros((N, 3), np.float32), np.zeros((N, 3), np.float32)
    rc = lib.psgpu_tree_octrees(C.byref(t), int(recompute), lo.ctypes.data, hi.ctypes.data)
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_tree_octrees")
    return lo, hi


def scene_load(path: str, root_id: int = -1) -> dict:
    """psgpu_scene_load: a .scene file as a flat tree (with srt and node ids)."""
    lib = load_library()
    h = C.c_void_p()
    rc = lib.psgpu_scene_load(os.fsencode(path), int(root_id), C.byref(h))
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_scene_load " + path)
    try:
        return _tree_dict(lib, h)
    finally:
        lib.psgpu_tree_free(h)


def scene_save(flat: dict, path: str):
    lib = load_library()
    t, keep = _tree_struct(flat)
    rc = lib.psgpu_scene_save(C.byref(t), os.fsencode(path))
    if rc != SUCCESS:
        raise PsGpuError(rc, "psgpu_scene_save " + path)


def count_mpus(cellsize, lo, hi) -> int:
    lib = load_library()
    lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
    return int(lib.psgpu_count_mpus(float(cellsize), lo.ctypes.data, hi.ctypes.data))


def lattice_dims(cellsize, lo, hi):
    lib = load_library()
    lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
    nm = np.zeros(3, np.int32)
    lib.psgpu_lattice_dims(float(cellsize), lo.ctypes.data, hi.ctypes.data, nm.ctypes.data)
    return tuple(int(v) for v in nm)


class Context:
    """One GPU context (psgpu_ctx). Not thread-safe, like the reference's Polygonize."""

    def __init__(self, device: int = 0):
        self.lib = load_library()
        self.h = C.c_void_p()
        rc = self.lib.psgpu_create(int(device), C.byref(self.h))
        if rc != SUCCESS:
            self.h = None
            raise PsGpuError(rc, "psgpu_create: no usable CUDA device (there is no CPU fallback)")
        self.device = device

    def close(self):
        if getattr(self, "h", None):
            self.lib.psgpu_destroy(self.h)
            self.h = None

    def __del__(self):
        self.close()

    def _check(self, rc, what):
        if rc != SUCCESS:
            raise PsGpuError(rc, "%s: %s" % (what, self.lib.psgpu_last_error(self.h).decode()))

    def upload(self, program: Program):
        self._check(self.lib.psgpu_upload_program(self.h, program.h), "psgpu_upload_program")

    def field(self, xyz) -> np.ndarray:
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self._check(self.lib.psgpu_field_eval(self.h, xyz.ctypes.data, len(xyz), out.ctypes.data), "psgpu_field_eval")
        return out

    def field_noalloc(self, xyz) -> np.ndarray:
        """psgpu_field_eval_noalloc: the same kernel with its staging in the library's static device array (<= NOALLOC_MAX_POINTS points)."""
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        out = np.zeros(len(xyz), np.float32)
        self._check(self.lib.psgpu_field_eval_noalloc(self.h, xyz.ctypes.data, len(xyz), out.ctypes.data), "psgpu_field_eval_noalloc")
        return out

    def newton_roots(self, p1, p2, f1, f2, target=0.5, iterations=8):
        """Shrink-wrap Newton per point pair in the uploaded program's semantics (computeRootNewtonRaphson / ...VEC4):
        (root, field at the root, field evaluations as the reference returns them)."""
        p1 = np.ascontiguousarray(p1, np.float32).reshape(-1, 3)
        p2 = np.ascontiguousarray(p2, np.float32).reshape(-1, 3)
        f1, f2 = np.ascontiguousarray(f1, np.float32), np.ascontiguousarray(f2, np.float32)
        n = len(p1)
        root, fr, ev = np.zeros((n, 3), np.float32), np.zeros(n, np.float32), np.zeros(n, np.int32)
        self._check(self.lib.psgpu_newton_roots(self.h, p1.ctypes.data, p2.ctypes.data, f1.ctypes.data, f2.ctypes.data, n, float(target),
                                                int(iterations), root.ctypes.data, fr.ctypes.data, ev.ctypes.data), "psgpu_newton_roots")
        return root, fr, ev

    def field_color_normal(self, xyz):
        xyz = np.ascontiguousarray(xyz, np.float32).reshape(-1, 3)
        n = len(xyz)
        f, col, nrm = np.zeros(n, np.float32), np.zeros((n, 3), np.float32), np.zeros((n, 3), np.float32)
        self._check(self.lib.psgpu_field_color_normal(self.h, xyz.ctypes.data, n, f.ctypes.data, col.ctypes.data, nrm.ctypes.data),
                    "psgpu_field_color_normal")
        return f, col, nrm

    def mpu_cache(self, params: Params, ijk):
        ijk = np.ascontiguousarray(ijk, np.int32)
        fv, cs = np.zeros(512, np.float32), np.zeros(343, np.uint8)
        self._check(self.lib.psgpu_mpu_cache(self.h, C.byref(params), ijk.ctypes.data, fv.ctypes.data, cs.ctypes.data), "psgpu_mpu_cache")
        return fv, cs

    def mpu_caches_culled(self, params: Params):
        """(ijk (n,3), fv (n,512)) of every MPU the culled pipeline evaluates, each with its own pruned program."""
        n = C.c_int64()
        self._check(self.lib.psgpu_mpu_caches_culled(self.h, C.byref(params), None, None, 0, C.byref(n)), "psgpu_mpu_caches_culled")
        ijk, fv = np.zeros((max(n.value, 1), 3), np.int32), np.zeros((max(n.value, 1), 512), np.float32)
        self._check(self.lib.psgpu_mpu_caches_culled(self.h, C.byref(params), ijk.ctypes.data, fv.ctypes.data, n.value, C.byref(n)),
                    "psgpu_mpu_caches_culled")
        return ijk[:n.value], fv[:n.value]

    def prune_region(self, lo, hi, cull: int, iso: float = 0.5, color_exact: bool = False, cap: int = 1 << 16):
        """Device-side prune of the uploaded program against a box; returns the pruned quads or None (nothing to evaluate)."""
        lo, hi = np.ascontiguousarray(lo, np.float32), np.ascontiguousarray(hi, np.float32)
        out, n = np.zeros((cap, 4), np.uint32), C.c_uint32()
        self._check(self.lib.psgpu_prune_region(self.h, lo.ctypes.data, hi.ctypes.data, cull, iso, int(color_exact), out.ctypes.data, cap,
                                                C.byref(n)), "psgpu_prune_region")
        return None if n.value == 0 else out[:n.value].copy()

    def polygonize(self, params: Params, download: bool = True, allow_over_capacity: bool = False) -> dict:
        """cull -> field -> scan -> emit on the device; optionally copies the mesh to numpy arrays."""
        m = _Mesh()
        rc = self.lib.psgpu_polygonize(self.h, C.byref(params), C.byref(m))
        if rc != SUCCESS and not (allow_over_capacity and rc == -2 and m.over_capacity > 0):
            self._check(rc, "psgpu_polygonize")
        out = dict(n_vertices=int(m.n_vertices), n_triangles=int(m.n_triangles), n_mpus_out=int(m.n_mpus_out),
                   lattice=tuple(m.lattice), over_capacity=int(m.over_capacity))
        if download:
            nv, nt, nm = out["n_vertices"], out["n_triangles"], out["n_mpus_out"]
            pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
            tri = np.zeros((nt, 3), np.uint32)
            ijk, vb, tb = np.zeros((nm, 3), np.int32), np.zeros(nm + 1, np.uint32), np.zeros(nm + 1, np.uint32)
            self._check(self.lib.psgpu_mesh_download(self.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data,
                                                     ijk.ctypes.data, vb.ctypes.data, tb.ctypes.data), "psgpu_mesh_download")
            out.update(pos=pos, nrm=nrm, col=col, tri=tri, mpu_ijk=ijk, mpu_vbase=vb, mpu_tbase=tb)
        return out

    # ---- multi-GPU: one Context per rank --------------------------------------------------------------------------------
    def comm_init(self, nranks: int, rank: int, unique_id: bytes):
        """psgpu_comm_init: joins the NCCL communicator whose 128-byte id rank 0 made with comm_unique_id()."""
        buf = C.create_string_buffer(bytes(unique_id), 128)
        self._check(self.lib.psgpu_comm_init(self.h, int(nranks), int(rank), buf), "psgpu_comm_init")
        self.nranks, self.rank = int(nranks), int(rank)

    def polygonize_sharded(self, params: Params, download: bool = False) -> dict:
        """psgpu_polygonize_sharded (collective): this rank's stitched share of one mesh; tri[] holds GLOBAL vertex ids."""
        m, sh = _Mesh(), _Shard()
        self._check(self.lib.psgpu_polygonize_sharded(self.h, C.byref(params), C.byref(m), C.byref(sh)), "psgpu_polygonize_sharded")
        out = dict(n_vertices=int(m.n_vertices), n_triangles=int(m.n_triangles), lattice=tuple(m.lattice),
                   shard={f: getattr(sh, f) for f, _ in _Shard._fields_})
        if download:
            nv, nt = out["n_vertices"], out["n_triangles"]
            pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
            tri = np.zeros((nt, 3), np.uint32)
            self._check(self.lib.psgpu_mesh_download(self.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data, None, None, None),
                        "psgpu_mesh_download")
            out.update(pos=pos, nrm=nrm, col=col, tri=tri)
        return out

    def layer_work(self, n_layers: int):
        """psgpu_layer_work: (evaluated, crossed) MPU counts per layer of the global lattice in the last polygonize of this context."""
        ev, cr = np.zeros(n_layers, np.int32), np.zeros(n_layers, np.int32)
        self._check(self.lib.psgpu_layer_work(self.h, ev.ctypes.data, cr.ctypes.data, int(n_layers)), "psgpu_layer_work")
        return ev, cr

    def polygonize_sharded_host(self, params: Params, pos, nrm, col, tri) -> dict:
        """psgpu_polygonize_sharded_host (collective): the rank's share lands in HOST buffers, copies overlapping the emission kernels."""
        m, sh = _Mesh(), _Shard()
        self._check(self.lib.psgpu_polygonize_sharded_host(self.h, C.byref(params), C.byref(m), C.byref(sh), pos.ctypes.data, nrm.ctypes.data,
                                                           col.ctypes.data, tri.ctypes.data, pos.shape[0], tri.shape[0]), "psgpu_polygonize_sharded_host")
        return dict(n_vertices=int(m.n_vertices), n_triangles=int(m.n_triangles), shard={f: getattr(sh, f) for f, _ in _Shard._fields_})

    def mesh_gather(self, root: int, pos_ptr: int, nrm_ptr: int, col_ptr: int, tri_ptr: int):
        """psgpu_mesh_gather (collective): device pointers on the root (ints, e.g. torch tensors' data_ptr()); any non-zero value on
        the other ranks means "send this array"."""
        self._check(self.lib.psgpu_mesh_gather(self.h, int(root), C.c_void_p(pos_ptr or None), C.c_void_p(nrm_ptr or None),
                                               C.c_void_p(col_ptr or None), C.c_void_p(tri_ptr or None)), "psgpu_mesh_gather")

    def polygonize_host(self, params: Params, pos, nrm, col, tri) -> dict:
        """The reference-facing call with HOST buffers (device -> host copies inside)."""
        m = _Mesh()
        rc = self.lib.psgpu_polygonize_host(self.h, C.byref(params), C.byref(m), pos.ctypes.data, nrm.ctypes.data, col.ctypes.data,
                                            tri.ctypes.data, pos.shape[0], tri.shape[0])
        self._check(rc, "psgpu_polygonize_host")
        return dict(n_vertices=int(m.n_vertices), n_triangles=int(m.n_triangles), lattice=tuple(m.lattice))

    def stats(self) -> dict:
        s = _Stats()
        self.lib.psgpu_get_stats(self.h, C.byref(s))
        return {f: getattr(s, f) for f, _ in _Stats._fields_}

    def fp32_peak(self):
        t, mhz = C.c_float(), C.c_float()
        self._check(self.lib.psgpu_measure_fp32_peak(self.h, C.byref(t), C.byref(mhz)), "psgpu_measure_fp32_peak")
        return float(t.value), float(mhz.value)
