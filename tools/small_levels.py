"""Development aid: cull hierarchy variants on the small configs (configs[0] small@128^3 compat, one frame of configs[4] anim@256^3)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes
ctx = ps.Context(0)
def run(prm, reps=9):
    ms = []
    for _ in range(reps):
        ctx.polygonize(prm, download=False); ms.append(ctx.stats()["ms_total"])
    st = ctx.stats()
    return round(float(np.median(ms[3:])), 4), round(st["ms_cull"], 4), st["mpus_evaluated"], st["gpu_launches"]
flat = bt.flatten(bt.binarize(scenes.small()))
lo, hi = ps.prepare_bboxes(flat); cs = scenes.cellsize_for(lo, hi, 128)
ctx.upload(ps.linearize(flat, ps.SEM_SIMD, 0))
for lv in ("4,1", "1"):
    os.environ["PSGPU_CULL_LEVELS"] = lv
    print("small", lv, run(ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE)))
print("small none", run(ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE, cull=ps.CULL_NONE)))
f0 = bt.flatten(scenes.anim_frame(0)); lo, hi = ps.prepare_bboxes(f0); cs = scenes.cellsize_for(lo, hi, 256)
ctx.upload(ps.linearize(f0, ps.SEM_SCALAR, ps.LIN_REASSOCIATE))
for lv in ("4,1", "1"):
    os.environ["PSGPU_CULL_LEVELS"] = lv
    print("anim", lv, run(ps.Params.make(cs, lo, hi)))
print("anim none", run(ps.Params.make(cs, lo, hi, cull=ps.CULL_NONE)))
