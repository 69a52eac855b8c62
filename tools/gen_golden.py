"""Generates tests/golden/*.npz from the reference's OWN code, compiled unmodified by oracle/Makefile:
  * <scene>_{asis,padded}.npz  - the SIMD polygonizer, oracle/_ref/libpsref.so (Parsip100/PS_SimdPoly/include/PS_Polygonizer.cpp)
  * scalar_<scene>.npz         - the scalar twin, oracle/_ref/libcompactref.so (PS_BlobTree scene API + COMPACTBLOBTREE +
                                 CParsipOptimized): the tree as include/psgpu_blobtree_walk.h reads it back from the reference's
                                 CBlobNode objects, fields / colours / normals / scene-API fields at random points, Newton roots
                                 on crossing lattice edges, and the per-MPU meshes of CParsipOptimized::run.
Run in the build container, where /root/reference exists; the fixtures are committed so the GPU box (no reference there) can
check against them.

    python tools/gen_golden.py [simd|scalar]
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.pyoracle import CompactRef, RefScene  # noqa: E402
from parsip_b200 import blobtree as bt, scenes  # noqa: E402
import scenes_small  # noqa: E402

CELLS = {"s1": 48, "blob": 40, "csg": 40, "discs_rings": 36, "xforms": 36}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    for name, mk in scenes_small.SIMD_SCENES.items():
        root = bt.binarize(mk())
        soa = bt.to_binary_soa(root)
        for padded in (False, True):
            ref = RefScene(soa)
            assert ref.build_info()["max_tree_nodes"] == 128, "golden vectors come from the UNMODIFIED build"
            ref.prepare_bboxes(0.05)
            lo, hi = ref.bbox()
            if padded:
                ref.pad_op_boxes(1.0e6)  # infinite lines reach beyond any finite box
            rs = np.random.RandomState(42)
            pts = rs.uniform(lo, hi, (3000, 3)).astype(np.float32)
            # field(): 4 consecutive points per SIMD call; field_color_normal(): one point broadcast to all lanes.
            # With the reference's own (unpadded) operator boxes the depth>3 prune decides per packet, so the two differ.
            fb, col, nrm = ref.field_color_normal(pts)
            f = ref.field(pts)
            assert padded is False or np.array_equal(f, fb)
            cs = scenes.cellsize_for(lo, hi, CELLS[name])
            ret, st = ref.polygonize(cs)
            assert ret == 1
            pos, mn, mc, tri = ref.mesh(st)
            np.savez_compressed(
                os.path.join(out, "%s_%s.npz" % (name, "padded" if padded else "asis")),
                bbox_lo=lo, bbox_hi=hi, prim_boxes=ref.prim_boxes(), op_boxes=ref.op_boxes(), pts=pts, field=f, field_bcast=fb, color=col, normal=nrm,
                cellsize=np.float32(cs), n_mpus=np.int64(st["mpus"]), n_crossed=np.int64(st["crossed"]),
                n_vertices=np.int64(st["vertices"]), n_triangles=np.int64(st["triangles"]), simd_evals=np.int64(st["simd_evals"]),
                pos_sha=sha(pos), tri_sha=sha(tri), col_sha=sha(mc), pos_head=pos[:256], tri_head=tri[:256],
                pos=pos.astype(np.float32) if len(pos) < 40000 else pos[:40000], tri=tri if len(tri) < 60000 else tri[:60000])
            print(name, "padded" if padded else "asis", st)


def main_simd_newton():
    """FieldComputer::computeRootNewtonRaphson (PS_Polygonizer.cpp:1625-1681; public, not called by Polygonize) on lattice-style crossing
    edges, with the reference's own and with padded operator boxes -> tests/golden/simd_newton.npz"""
    out, rec = os.path.join(ROOT, "tests", "golden"), {}
    for name, mk in scenes_small.SIMD_SCENES.items():
        soa = bt.to_binary_soa(bt.binarize(mk()))
        for padded in (False, True):
            ref = RefScene(soa)
            ref.prepare_bboxes(0.05)
            lo, hi = ref.bbox()
            if padded:
                ref.pad_op_boxes(1.0e6)
            rs = np.random.RandomState(44)
            cs = np.float32(scenes.cellsize_for(lo, hi, CELLS[name]))
            cand = rs.uniform(lo, hi, (40000, 3)).astype(np.float32)
            p2 = cand.copy()
            p2[np.arange(len(cand)), rs.randint(0, 3, len(cand))] += cs
            f1, f2 = ref.field_color_normal(cand)[0], ref.field_color_normal(p2)[0]  # broadcast evaluation, as the root finder's own
            cross = np.nonzero((f1 > 0.5) != (f2 > 0.5))[0][:300]
            k = "%s_%s_" % (name, "padded" if padded else "asis")
            rec[k + "p1"], rec[k + "p2"], rec[k + "f1"], rec[k + "f2"] = cand[cross], p2[cross], f1[cross], f2[cross]
            for iters in (8, 2):
                root, fr, ev = ref.newton(cand[cross], p2[cross], f1[cross], f2[cross], 0.5, iters)
                rec[k + "root%d" % iters], rec[k + "field%d" % iters], rec[k + "evals%d" % iters] = root, fr, ev
            print("simd newton", k, len(cross), "edges, evals histogram", np.bincount(rec[k + "evals8"] // 4))
    np.savez_compressed(os.path.join(out, "simd_newton.npz"), **rec)


SCALAR_CELLS = 36
WALK_KEYS = ("type", "first_child", "n_children", "child", "pos", "dir", "res", "color", "xform", "has_xform", "oct_lo", "oct_hi")


TRAIN_SCENE = "/root/reference/Parsip100/Distrib/train_corrected.scene"


def train_flat():
    """The reference's one real model (Distrib/train_corrected.scene, 95 primitives), read by the product's .scene reader and put
    through the GUI's flattenTransformations - the tree CLayer hands to its polygonizers."""
    import parsip_b200 as ps
    return ps.flatten_transforms(ps.scene_load(TRAIN_SCENE))


def main_scalar():
    out = os.path.join(ROOT, "tests", "golden")
    todo = [(name, (lambda mk=mk: bt.flatten(mk()))) for name, mk in scenes_small.COMPACTREF_SCENES.items()]
    if os.path.exists(TRAIN_SCENE):
        todo.append(("train", train_flat))
    for name, mk_flat in todo:
        ref = CompactRef(mk_flat())
        assert ref.convert_rc >= 0 and ref.walk_rc == 0, (name, ref.convert_rc, ref.walk_rc)
        w = ref.walk()
        lo, hi = ref.octree()
        rs = np.random.RandomState(43)
        pts = rs.uniform(lo - 0.2, hi + 0.2, (3000, 3)).astype(np.float32)
        f, col, nrm = ref.field_color_normal(pts)
        cs = np.float32(scenes.cellsize_for(lo, hi, SCALAR_CELLS))
        # Newton inputs: lattice-style edges (one cell long, axis-aligned) whose ends lie on different sides of the iso-surface
        cand = rs.uniform(lo, hi, (60000, 3)).astype(np.float32)
        axis = rs.randint(0, 3, len(cand))
        p2 = cand.copy()
        p2[np.arange(len(cand)), axis] += cs
        f1, f2 = ref.field(cand), ref.field(p2)
        cross = np.nonzero((f1 > 0.5) != (f2 > 0.5))[0][:400]
        root, fr, ev = ref.newton(cand[cross], p2[cross], f1[cross], f2[cross])
        m = ref.polygonize(lo, hi, float(cs), nthreads=1)
        ok, em = ref.export_mesh()
        assert ok and em["vertices"] == m["vertices"] and em["triangles"] == m["triangles"]
        extra = {}
        if name == "train":
            # the run the reference recorded (Distrib/ParsipHaptics_Release.csv:2: cell size 0.13, 146 352 faces / 82 370 vertices with its
            # older 16^3 MPUs) repeated with today's 8^3 build on the GUI's domain (setOctreeFromBlobTree): counts only
            ref.recompute_octrees()
            glo, ghi = ref.octree()
            big = ref.polygonize(glo, ghi, 0.13, nthreads=0)
            extra = dict(gui_lo=glo, gui_hi=ghi, gui_cellsize=np.float32(0.13), gui_mpus=np.int64(big["mpus"]), gui_vertices=np.int64(big["vertices"]),
                         gui_triangles=np.int64(big["triangles"]))
            print("train @0.13 on", glo, ghi, "->", big["mpus"], "MPUs", big["vertices"], "vertices", big["triangles"], "triangles")
        np.savez_compressed(
            os.path.join(out, "scalar_%s.npz" % name),
            n_nodes=np.int32(w["n_nodes"]), **{"tree_" + k: w[k] for k in WALK_KEYS},
            n_prims=np.int32(ref.n_prims), n_ops=np.int32(ref.n_ops), bbox_lo=lo, bbox_hi=hi, pts=pts, field=f, color=col, normal=nrm,
            scene_field=ref.scene_field(pts), newton_p1=cand[cross], newton_p2=p2[cross], newton_f1=f1[cross], newton_f2=f2[cross],
            newton_root=root, newton_field=fr, newton_evals=ev, cellsize=cs, mesh_mpus=np.int64(m["mpus"]),
            mesh_field_evals=np.int64(m["field_evals"]), mesh_origin=m["mpu_origin"], mesh_counts=m["mpu_counts"], mesh_pos=m["pos"],
            mesh_nrm=m["nrm"], mesh_col=m["col"], mesh_tri=m["tri"], export_pos_sha=sha(em["pos"]), export_tri_sha=sha(em["tri"]),
            tree_srt=w["srt"], tree_node_id=w["node_id"], **extra)
        print("scalar", name, "prims", ref.n_prims, "ops", ref.n_ops, "surface pts", int((f > 0.5).sum()), "newton edges", len(cross),
              "mesh", m["mpus"], m["vertices"], m["triangles"])


if __name__ == "__main__":
    what = sys.argv[1] if len(sys.argv) > 1 else "all"
    if what in ("all", "simd"):
        main()
    if what in ("all", "simd", "simd_newton"):
        main_simd_newton()
    if what in ("all", "scalar"):
        main_scalar()
