"""Development aid: field caches from per-MPU pruned programs vs the full program, on the GPU."""
import os, sys
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes
import scenes_small

root = scenes_small.SCALAR_SCENES["csgwarp"]()
flat = bt.flatten(root)
lo, hi = ps.prepare_bboxes(bt.flatten(bt.binarize(root)))
lo, hi = lo - 0.25, hi + 0.25
cs = scenes.cellsize_for(lo, hi, 64)
ctx = ps.Context(0)
ctx.upload(ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE))
for cull in (ps.CULL_NONE, ps.CULL_SUPPORT):
    prm = ps.Params.make(cs, lo, hi, coord_mode=ps.COORD_MPU_LOCAL, output=ps.OUT_MPU, cull=cull, root_mode=ps.ROOT_NEWTON_REF)
    ijk, fv = ctx.mpu_caches_culled(prm)
    nbad = 0
    for k, f in zip(ijk, fv):
        full, _ = ctx.mpu_cache(prm, k)
        if not np.array_equal(f.view(np.uint32), full.view(np.uint32)):
            nbad += 1
            if nbad <= 5:
                d = np.abs(f - full)
                w = np.argsort(d)[-3:]
                print("  MPU", tuple(k), "corners differing", int((d > 0).sum()), "max", d.max(), "at idx", w, "pruned", f[w], "full", full[w])
                a = (lo + np.float32(k) * np.float32(cs) * np.float32(7)).astype(np.float32) - np.float32(6e-5)
                b = (a + np.float32(cs) * np.float32(7)) + np.float32(1.2e-4)
                g = ctx.prune_region(a, b, cull)
                print("     program:", None if g is None else [(int(w & 31), int((w >> 5) & 15)) for w in g[:, 0]])
    print("cull", cull, "evaluated", len(ijk), "MPUs whose cache differs from the full program:", nbad)
