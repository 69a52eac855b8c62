"""Quick GPU shake-out (development aid): S1 scene, GPU vs C oracle."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes
from oracle.pyoracle import SimdOracle, ScalarOracle

ctx = ps.Context(0)
print("fp32 peak (TFLOP/s, MHz):", ctx.fp32_peak())
root = bt.binarize(scenes.small())
soa = bt.to_binary_soa(root)
orc = SimdOracle(soa)
orc.prepare_bboxes(0.05)
orc.pad_op_boxes(1.0e6)  # neutralise the reference op-box prune
lo, hi = orc.bbox()
cs = scenes.cellsize_for(lo, hi, 128)
prog = ps.linearize(bt.flatten(root), ps.SEM_SIMD, 0)
ctx.upload(prog)
pts = np.random.RandomState(0).uniform(lo, hi, (100000, 3)).astype(np.float32)
fg, fr = ctx.field(pts), orc.field(pts)
rel = np.abs(fg - fr) / np.maximum(np.abs(fr), 1e-3)
print("field: max rel", rel.max(), "max abs", np.abs(fg - fr).max(), "bit-equal frac", (fg == fr).mean())
fg2, cg, ng = ctx.field_color_normal(pts[:2000])
fr2, cr, nr = orc.field_color_normal(pts[:2000])
ok = np.isfinite(nr).all(axis=1) & np.isfinite(ng).all(axis=1)
print("color max diff", np.abs(cg - cr).max(), "normal max diff", np.abs(ng[ok] - nr[ok]).max())

for cull in (ps.CULL_NONE, ps.CULL_SUPPORT, ps.CULL_INTERVAL):
    prm = ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE, coord_mode=ps.COORD_MPU_LOCAL, output=ps.OUT_MPU, cull=cull, collect_stats=1)
    t = time.time(); m = ctx.polygonize(prm); dt = time.time() - t
    st = ctx.stats()
    print("cull", cull, "nv", m["n_vertices"], "nt", m["n_triangles"], "mpus out", m["n_mpus_out"], "wall %.1f ms" % (dt * 1e3),
          {k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()})
_, ref = orc.polygonize(cs)
print("oracle nv", ref["vertices"], "nt", ref["triangles"], "crossed", ref["crossed"])
# per-MPU comparison
order = np.lexsort((m["mpu_ijk"][:, 2], m["mpu_ijk"][:, 1], m["mpu_ijk"][:, 0]))
ro = ref["mpu_idx"]
print("mpu sets equal:", np.array_equal(m["mpu_ijk"][order], ro))
vb_ref = np.concatenate([[0], np.cumsum(ref["mpu_counts"][:, 0].astype(np.int64))]).astype(np.int64); tb_ref = np.concatenate([[0], np.cumsum(ref["mpu_counts"][:, 1].astype(np.int64))]).astype(np.int64)
maxd = 0.0; badc = 0; tri_bad = 0
for q, r in enumerate(order):
    v0, v1, t0, t1 = int(m["mpu_vbase"][r]), int(m["mpu_vbase"][r + 1]), int(m["mpu_tbase"][r]), int(m["mpu_tbase"][r + 1])
    if v1 - v0 != ref["mpu_counts"][q, 0] or t1 - t0 != ref["mpu_counts"][q, 1]:
        badc += 1; continue
    pg, pr = m["pos"][v0:v1], ref["pos"][vb_ref[q]:vb_ref[q + 1]]
    fin = np.isfinite(pr).all(axis=1) & np.isfinite(pg).all(axis=1)
    if fin.any(): maxd = max(maxd, np.abs(pg[fin] - pr[fin]).max())
    tg, tr = m["tri"][t0:t1], ref["tri"][tb_ref[q]:tb_ref[q + 1]] - vb_ref[q]
    tri_bad += int(not np.array_equal(tg, tr))
print("MPUs with count mismatch", badc, "max vertex diff / cell", maxd / cs, "MPUs with triangle index mismatch", tri_bad)

# native indexed mesh
for cull in (ps.CULL_NONE, ps.CULL_INTERVAL):
    prm = ps.Params.make(cs, lo, hi, cull=cull)
    t = time.time(); n = ctx.polygonize(prm); dt = time.time() - t
    print("native cull", cull, "nv", n["n_vertices"], "nt", n["n_triangles"], "wall %.1f ms" % (dt * 1e3), "max idx", n["tri"].max() if n["n_triangles"] else -1)
    if cull == ps.CULL_NONE: base = n
print("native cull on/off identical:", np.array_equal(base["pos"], n["pos"]) and np.array_equal(base["tri"], n["tri"]))
