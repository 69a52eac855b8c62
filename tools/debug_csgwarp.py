"""Development aid: csgwarp scene, GPU (scalar semantics) vs the scalar oracle, per-MPU."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import numpy as np
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes
from oracle.pyoracle import ScalarOracle
import scenes_small, gpu_helpers as gh

name = sys.argv[1] if len(sys.argv) > 1 else "csgwarp"
cells = int(sys.argv[2]) if len(sys.argv) > 2 else 64
root = scenes_small.SCALAR_SCENES[name]()
flat = bt.flatten(root)
lo, hi = ps.prepare_bboxes(bt.flatten(bt.binarize(root)))
lo, hi = lo - 0.25, hi + 0.25
cs = scenes.cellsize_for(lo, hi, cells)
ctx = ps.Context(0)
ctx.upload(ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE))
orc = ScalarOracle(flat)
ret, ref = orc.polygonize(lo, hi, cs)
print("oracle", ref["vertices"], ref["triangles"], ref["crossed"], ref["lattice"])
refc = {tuple(i): tuple(c) for i, c in zip(ref["mpu_idx"], ref["mpu_counts"])}
for cull in (ps.CULL_NONE, ps.CULL_SUPPORT, ps.CULL_INTERVAL):
    m = ctx.polygonize(ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_NEWTON_REF, coord_mode=ps.COORD_MPU_LOCAL, output=ps.OUT_MPU, cull=cull),
                       allow_over_capacity=True)
    pos, nrm, col, tri, ijk, counts = gh.lattice_ordered(m)
    gc = {tuple(i): tuple(c) for i, c in zip(ijk, counts)}
    diff = [k for k in sorted(set(refc) | set(gc)) if refc.get(k) != gc.get(k)]
    print("cull", cull, "gpu", m["n_vertices"], m["n_triangles"], m["n_mpus_out"], "differing MPUs", len(diff), ctx.stats()["mpus_evaluated"])
    for k in diff[:6]:
        prm = ps.Params.make(cs, lo, hi, coord_mode=ps.COORD_MPU_LOCAL, output=ps.OUT_MPU)
        fv, cases = ctx.mpu_cache(prm, list(k))
        fvr, cr = orc.mpu_cache(lo, cs, list(k))
        print("   ", k, "ref", refc.get(k), "gpu", gc.get(k), "cache maxdiff", np.abs(fv - fvr).max(), "case mismatches", int((cases != cr).sum()),
              "near-iso", int((np.abs(fvr - 0.5) < 1e-6).sum()))
    if not diff:
        d = np.linalg.norm(pos.astype(np.float64) - ref["pos"].astype(np.float64), axis=1)
        print("    vertex dist / cell: p50 %.2e p99 %.2e max %.2e; tri equal %s; col maxdiff %.2e" % (
            np.percentile(d, 50) / cs, np.percentile(d, 99) / cs, d.max() / cs, np.array_equal(tri, ref["tri"]), np.abs(col - ref["col"]).max()))
for cull in (ps.CULL_NONE, ps.CULL_INTERVAL):
    n = ctx.polygonize(ps.Params.make(cs, lo, hi, cull=cull))
    bad = (n["tri"] >= n["n_vertices"]).any(axis=1)
    print("native cull", cull, n["n_vertices"], n["n_triangles"], "bad triangles", int(bad.sum()))
    if bad.any():
        t = n["tri"][bad][:5]
        print(t)
        for row in t:
            ok = row[row < n["n_vertices"]]
            print("   ", (n["pos"][ok] - lo) / cs / 7)
