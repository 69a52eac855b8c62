"""-m gpu: the CUDA path in scalar-twin semantics (PSGPU_SEM_SCALAR) against fixtures produced by the reference's OWN scalar code
(tests/golden/scalar_*.npz: PS_BlobTree scene API + COMPACTBLOBTREE + CParsipOptimized compiled unmodified, tools/gen_golden.py).
The tree each test uploads is the one include/psgpu_blobtree_walk.h read back from the reference's CBlobNode objects.

North-star bars: field 1e-5 relative; vertices 1e-4 cell; symmetric Hausdorff 1e-3 cell. Where a wider bound is used the measured
reason is stated next to it."""
import os

import numpy as np
import pytest

import gpu_helpers as gh
from test_compactref import COLOR_UNPINNED, FIELD_FLOOR, GOLDEN_NAMES as NAMES, golden, golden_tree
import parsip_b200 as ps

pytestmark = pytest.mark.gpu


def upload(ctx, g):
    ctx.upload(ps.linearize(golden_tree(g), ps.SEM_SCALAR, ps.LIN_REASSOCIATE))


@pytest.mark.parametrize("name", NAMES)
def test_field_vs_reference_compact_tree(ctx, name):
    g = golden(name)
    upload(ctx, g)
    f, fr = ctx.field(g["pts"]), g["field"]
    rel = np.abs(f - fr) / np.maximum(np.abs(fr), FIELD_FLOOR.get(name, 1e-3))
    # north-star bar for every scene: primitives and combines use FMA (the reference's CPU build does not), point transforms
    # (matrices, warps) keep the reference's operation order and take sin / cos rounded once from double
    assert rel.max() < 1e-5, rel.max()
    # inside / outside agrees except within 1e-6 of the iso-value (north star: cube cases bit-exact outside that band)
    band = np.abs(fr - 0.5) < 1e-6
    assert np.array_equal((f > 0.5)[~band], (fr > 0.5)[~band])


@pytest.mark.parametrize("name", NAMES)
def test_color_and_normal_vs_reference_compact_tree(ctx, name):
    g = golden(name)
    upload(ctx, g)
    f, col, nrm = ctx.field_color_normal(g["pts"])
    fr, cr, nr = g["field"], g["color"][:, :3], g["normal"]
    on = fr > 1e-3
    cd = np.abs(col[on] - cr[on]).max(axis=1)
    # blend colours are f-weighted means: same 1e-5-relative inputs. A max / min / difference pick flips where two children tie
    # to within the field tolerance; those points are counted, not hidden.
    assert name in COLOR_UNPINNED or (np.percentile(cd, 99.5) < 2e-4 and (cd > 2e-4).mean() < 5e-3), (cd.max(), (cd > 2e-4).mean())
    flat_n = (nr == 1.0).all(axis=1)  # the reference's normalizeXYZ returns (1,1,1) for a zero gradient
    good = on & ~flat_n & (np.linalg.norm(nr, axis=1) > 0.5) & (fr > 0.05)
    ang = np.degrees(np.arccos(np.clip((nrm[good] * nr[good]).sum(1), -1, 1)))
    # forward differences with delta 1e-3 of fields that agree to ~1e-7 absolute: the quotient carries ~1e-4 absolute noise
    # against gradients of order 1, i.e. ~0.01 degrees; SURVEY Q7 asks for 0.1 degrees. The tail is plateaus with |grad| ~ 1e-3.
    # train: Ricci blends of power 16-20 make the field a near-step between their operands; its difference quotient is rounding noise
    # over a wider band (measured p99 0.20 degrees)
    bar = 0.5 if name == "train" else 0.1
    assert np.percentile(ang, 99) < bar and np.median(ang) < 0.02, (np.median(ang), np.percentile(ang, 99), ang.max())


def _by_mpu(g):
    """golden per-MPU meshes keyed by lattice index (origin = lo + side * ijk, CPolyParsipOptimized.cpp:371-384)"""
    cs = float(g["cellsize"])
    ijk = np.rint((g["mesh_origin"].astype(np.float64) - g["bbox_lo"].astype(np.float64)) / (7.0 * cs)).astype(np.int64)
    vb = np.concatenate([[0], np.cumsum(g["mesh_counts"][:, 0])])
    tb = np.concatenate([[0], np.cumsum(g["mesh_counts"][:, 1])])
    return {tuple(ijk[i]): (int(vb[i]), int(vb[i + 1]), int(tb[i]), int(tb[i + 1])) for i in range(len(ijk))}


@pytest.mark.parametrize("name", NAMES)
def test_newton_mode_mesh_vs_reference_polygonizer(ctx, name):
    """PSGPU_ROOT_NEWTON_REF + MPU-local coordinates + per-MPU output = CParsipOptimized on the GPU, against the meshes the
    reference's own CParsipOptimized::run produced. The reference skips MPUs that miss every primitive's heuristic octree and
    thereby drops surface (tests/test_compactref.py shows it on the CPU): every MPU the reference kept must be
    matched here vertex for vertex, and the GPU's extra surface MPUs are counted."""
    g = golden(name)
    upload(ctx, g)
    cs, lo, hi = float(g["cellsize"]), g["bbox_lo"], g["bbox_hi"]
    m = ctx.polygonize(ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_NEWTON_REF, coord_mode=ps.COORD_MPU_LOCAL, output=ps.OUT_MPU),
                       allow_over_capacity=True)
    ref = _by_mpu(g)
    vb, tb = m["mpu_vbase"].astype(np.int64), m["mpu_tbase"].astype(np.int64)
    mine = {tuple(int(v) for v in m["mpu_ijk"][r]): r for r in range(len(m["mpu_ijk"])) if tb[r + 1] > tb[r]}
    assert set(ref) <= set(mine), "reference MPUs missing on the GPU: %s" % sorted(set(ref) - set(mine))[:5]
    extra = len(set(mine) - set(ref))
    d_all, tri_equal, cd_all, ang_all = [], 0, [], []
    for key, (v0, v1, t0, t1) in ref.items():
        r = mine[key]
        assert vb[r + 1] - vb[r] == v1 - v0 and tb[r + 1] - tb[r] == t1 - t0, key
        tri_equal += int(np.array_equal(m["tri"][tb[r]:tb[r + 1]].astype(np.int64), g["mesh_tri"][t0:t1].astype(np.int64) - v0))
        p, q = m["pos"][vb[r]:vb[r + 1]].astype(np.float64), g["mesh_pos"][v0:v1].astype(np.float64)
        fin = np.isfinite(q).all(axis=1) & np.isfinite(g["mesh_col"][v0:v1]).all(axis=1)  # a Newton step over a zero gradient leaves inf / NaN vertices in the reference mesh (2 in the train model)
        d_all.append(np.linalg.norm(p[fin] - q[fin], axis=1) / cs)
        cd_all.append(np.abs(m["col"][vb[r]:vb[r + 1]][fin] - g["mesh_col"][v0:v1, :3][fin]).max(axis=1))
        n1, n2 = m["nrm"][vb[r]:vb[r + 1]], g["mesh_nrm"][v0:v1]
        ang_all.append(np.degrees(np.arccos(np.clip((n1[fin] * n2[fin]).sum(1), -1, 1))))
    assert tri_equal == len(ref)  # cube cases, table walk and per-MPU vertex numbering identical
    d, cd, ang = np.concatenate(d_all), np.concatenate(cd_all), np.concatenate(ang_all)
    # The reference's root finder is a 3-D shrink-wrap Newton with a |f - iso| < 1e-3 stopping rule (CompactBlobTree.cpp:1581-1622):
    # each step divides by |grad|^2 and the iteration count depends on which side of 1e-3 an iterate lands, so an FMA-level
    # field difference (1e-7) either vanishes (median) or costs one extra step (tail, bounded by the rule's own 1e-3 / |grad|
    # ~ 5e-4 units ~ 5e-3 cell ball). Vertices that collapse onto a corner with ~zero gradient (plateaus of clamped differences)
    # are the max.
    assert np.median(d) < 1e-4 and np.percentile(d, 99) < 5e-3 and d.max() < 5e-2, (np.median(d), np.percentile(d, 99), d.max())
    assert name in COLOR_UNPINNED or (np.percentile(cd, 99) < 1e-3 and (cd > 1e-3).mean() < 0.01), (cd.max(), (cd > 1e-3).mean())
    assert np.percentile(ang, 95) < 0.5, (np.median(ang), np.percentile(ang, 95))
    print("%s: %d reference MPUs matched, %d extra surface MPUs on the GPU (reference octree filter), vertex d/cell median %.2e p99 %.2e max %.2e"
          % (name, len(ref), extra, np.median(d), np.percentile(d, 99), d.max()))


@pytest.mark.parametrize("name", NAMES)
def test_native_mesh_vs_reference_mesh_hausdorff(ctx, name):
    """The product's native output (bracketed root search on the lattice edge, global vertex dedup) against the reference mesh.
    Reference vertices sit where its Newton stopped: |f - iso| < 1e-3, i.e. up to 1e-3 / |grad| off the iso-surface, and its
    shrink-wrap collapses all edges that start at one corner onto a single vertex. The reference mesh is therefore itself up to
    ~half a cell (measured below) away from the true iso-surface the native mesh tracks to 2e-4 in field value; the north-star
    1e-3-cell Hausdorff bar is met against the reference where the comparison is meaningful: native vertices vs the reference
    FIELD (residual), and reference vertices vs the native surface within the reference's own error ball."""
    g = golden(name)
    upload(ctx, g)
    cs, lo, hi = float(g["cellsize"]), g["bbox_lo"], g["bbox_hi"]
    m = ctx.polygonize(ps.Params.make(cs, lo, hi))
    assert m["n_triangles"] > 0
    # (a) native vertices against the REFERENCE field (the reference-built oracle restated bit-exactly on the CPU)
    from oracle.pyoracle import ScalarOracle
    fr = ScalarOracle(golden_tree(g)).field(m["pos"])
    res = np.abs(fr - 0.5)
    assert res.max() < 2e-4, res.max()
    # |grad f| near the surface is >= ~0.5 for these scenes: 2e-4 in f is <= 4e-4 units, i.e. a few 1e-3 of a cell
    # (b) every reference vertex lies within the reference's own Newton ball of the native surface
    ref_res = np.abs(g["newton_field"] - 0.5).max()
    ref_pos = g["mesh_pos"][np.isfinite(g["mesh_pos"]).all(axis=1)]
    h = gh.one_sided_hausdorff(ref_pos, m["pos"], m["tri"]) / cs
    # corner-collapsed vertices: up to ~sqrt(3)/2 cell off the edge crossing they stand for; on the train model the reference's Newton
    # overshoots across the near-steps of its power-20 Ricci blends and leaves a few vertices two cells out (its own residual is printed)
    assert h < (2.5 if name == "train" else 0.75), h
    print("%s: native residual vs reference field max %.2e; reference vertices to native surface: max %.3f cell (reference Newton residual %.1e)"
          % (name, res.max(), h, ref_res))


@pytest.mark.parametrize("name", NAMES)
def test_newton_roots_entry_vs_reference(ctx, name):
    """psgpu_newton_roots in the scalar semantics = ComputeRootNewtonRaphsonVEC4 (CompactBlobTree.cpp:1581-1622) on the edges the reference
    build solved for the goldens. The iteration is the reference's; the field under it differs at FMA level (1e-7), which the 1e-3
    stopping rule turns into one round more or less on a few edges."""
    g = golden(name)
    upload(ctx, g)
    root, fr, ev = ctx.newton_roots(g["newton_p1"], g["newton_p2"], g["newton_f1"], g["newton_f2"], 0.5, 8)
    same = ev == g["newton_evals"]
    assert same.mean() > 0.97, same.mean()
    d = np.linalg.norm(root[same].astype(np.float64) - g["newton_root"][same, :3], axis=1) / float(g["cellsize"])
    assert np.median(d) < 1e-4 and np.percentile(d, 99) < 5e-3, (np.median(d), d.max())  # the vertex bars of the mesh test above
    assert (np.abs(fr - 0.5) < 1e-3)[ev < 36].all()


def test_train_model_counts_vs_reference_run(ctx):
    """The reference's one real model at the cell size of its recorded runs (Distrib/ParsipHaptics_Release.csv:2: 0.13, 146 352 faces,
    82 370 vertices with the older 16^3-MPU build). tools/gen_golden.py repeated that run with the reference's own code as built here
    (8^3 MPUs, GUI domain from setOctreeFromBlobTree) and stored its counts; the GPU in the same mode (Newton roots, per-MPU vertices)
    must reproduce the triangle count of every MPU the reference kept, the native mesh the same faces with shared vertices."""
    g = golden("train")
    upload(ctx, g)
    lo, hi, cs = g["gui_lo"], g["gui_hi"], float(g["gui_cellsize"])
    m = ctx.polygonize(ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_NEWTON_REF, coord_mode=ps.COORD_MPU_LOCAL, output=ps.OUT_MPU), allow_over_capacity=True)
    ref_t, ref_v = int(g["gui_triangles"]), int(g["gui_vertices"])
    # the reference's octree filter may drop surface MPUs (Q12); never the other way round
    assert m["n_triangles"] >= ref_t and (m["n_triangles"] - ref_t) <= 0.02 * ref_t, (m["n_triangles"], ref_t)
    assert abs(m["n_vertices"] - ref_v) <= 0.02 * ref_v, (m["n_vertices"], ref_v)
    assert abs(m["n_triangles"] - 146352) <= 0.02 * 146352  # and the recorded run, loosely (other MPU size, same cells)
    nat = ctx.polygonize(ps.Params.make(cs, lo, hi))
    assert nat["n_triangles"] == m["n_triangles"] and nat["n_vertices"] < m["n_vertices"]
    print("train @0.13: reference build %d triangles / %d vertices, recorded 146352 / 82370; GPU Newton mode %d / %d, native %d / %d"
          % (ref_t, ref_v, m["n_triangles"], m["n_vertices"], nat["n_triangles"], nat["n_vertices"]))


def test_parsipcmd_cli_on_a_scene_file(ctx, tmp_path):
    """parsip_b200/parsipcmd (the ParsipCmd role of BASELINE configs[0]): .scene file -> flatten -> octree domain -> polygonize; its
    counts must equal the library called directly on the same tree and domain."""
    import subprocess
    g = golden("train")
    tree = golden_tree(g)
    tree["srt"], tree["has_srt"], tree["node_id"] = g["tree_srt"], np.ones(int(g["n_nodes"]), np.uint8), g["tree_node_id"]
    path = str(tmp_path / "train.scene")
    ps.scene_save(tree, path)
    exe = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "parsip_b200", "parsipcmd")
    r = subprocess.run([exe, path, "0.13", "--reps", "3", "--obj", str(tmp_path / "train.obj")], capture_output=True, text=True, timeout=300)
    assert r.returncode == 0, r.stdout + r.stderr
    row = r.stdout.strip().splitlines()[-1].split(",")
    faces, verts = int(row[10]), int(row[11])
    t2 = ps.flatten_transforms(ps.scene_load(path))
    lo, hi = ps.tree_octrees(t2, True)
    ctx.upload(ps.linearize(t2, ps.SEM_SCALAR, ps.LIN_REASSOCIATE))
    m = ctx.polygonize(ps.Params.make(0.13, lo[0], hi[0]))
    assert (faces, verts) == (m["n_triangles"], m["n_vertices"]) and int(row[1]) == 95
    assert abs(faces - 146352) <= 0.02 * 146352  # ParsipHaptics_Release.csv:2, loosely (%f scene precision, other MPU size)
    lines = open(tmp_path / "train.obj").read().splitlines()
    assert sum(l.startswith("v ") for l in lines) == verts and sum(l.startswith("f ") for l in lines) == faces


def test_export_mesh_into_the_reference_cmeshvv(ctx):
    """include/psgpu_meshvv_export.h compiled against the reference's CMeshVV declaration (oracle/_ref/libblobnode_probe.so,
    tests/facade/blobnode_probe.cpp): CBlobNode* -> walker -> GPU polygonize -> psgpu::export_mesh, against (1) the same mesh fetched
    through the C ABI and rebased the way CMeshVV::appendFrom does, and (2) CParsipOptimized::exportMesh of the reference build."""
    import ctypes as C
    from oracle import pyoracle
    from parsip_b200 import blobtree as bt
    import scenes_small
    probe = os.path.join(os.path.dirname(pyoracle.compactref_lib_path()), "libblobnode_probe.so")
    if not (os.path.exists(probe) and os.path.exists(pyoracle.compactref_lib_path())):
        pytest.skip("oracle/_ref reference builds not present")
    lib = C.CDLL(probe)
    lib.bnp_export_mesh.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_float, C.c_int] + [C.c_void_p] * 5
    cref = pyoracle._load_compactref()
    cref.cref_root.restype, cref.cref_root.argtypes = C.c_void_p, [C.c_void_p]
    ref = pyoracle.CompactRef(bt.flatten(scenes_small.COMPACTREF_SCENES["blob"]()))
    g = golden("blob")
    lo, hi, cs = np.ascontiguousarray(g["bbox_lo"], np.float32), np.ascontiguousarray(g["bbox_hi"], np.float32), float(g["cellsize"])
    rm = ref.polygonize(lo, hi, cs, nthreads=2)
    ok, rex = ref.export_mesh()
    assert ok and rex["triangles"] == rm["triangles"]
    upload(ctx, g)
    for per_mpu in (1, 0):
        cnt = np.zeros(2, np.int64)
        assert lib.bnp_export_mesh(cref.cref_root(ref.h), lo.ctypes.data, hi.ctypes.data, cs, per_mpu, cnt.ctypes.data, None, None, None, None) == 1
        nv, nt = int(cnt[0]), int(cnt[1])
        pos, nrm, col, tri = np.zeros((nv, 3), np.float32), np.zeros((nv, 3), np.float32), np.zeros((nv, 4), np.float32), np.zeros((nt, 3), np.uint32)
        assert lib.bnp_export_mesh(None, None, None, cs, per_mpu, cnt.ctypes.data, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data) == 1
        kw = dict(root_mode=ps.ROOT_NEWTON_REF, coord_mode=ps.COORD_MPU_LOCAL, output=ps.OUT_MPU) if per_mpu else {}
        m = ctx.polygonize(ps.Params.make(cs, lo, hi, **kw))
        assert (nv, nt) == (m["n_vertices"], m["n_triangles"])
        want = m["tri"].astype(np.int64)
        if per_mpu:
            vb, tb = m["mpu_vbase"].astype(np.int64), m["mpu_tbase"].astype(np.int64)
            want = want + np.repeat(vb[:-1], np.diff(tb))[:, None]
        assert np.array_equal(tri.astype(np.int64), want) and tri.max() < nv
        assert np.array_equal(pos, m["pos"]) and np.array_equal(nrm, m["nrm"]) and np.array_equal(col[:, :3], m["col"]) and (col[:, 3] == 1).all()
        if per_mpu:  # the reference's own export of the same scene: its primitive-octree filter drops some surface MPUs (Q12; the mesh test
            # above matches the kept ones one by one), so it never has more faces and here 2.3 % fewer
            assert nt >= rex["triangles"] and nt - rex["triangles"] <= 0.05 * nt, (nt, rex["triangles"])
