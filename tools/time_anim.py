"""Development aid: where a frame of BASELINE configs[4] (anim@256^3) goes - linearize / upload / polygonize wall clock against the GPU's own
event time - and the same for configs[0]."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes

ctx = ps.Context(0)
f0 = bt.flatten(scenes.anim_frame(0))
lo, hi = ps.prepare_bboxes(f0)
for f in (30, 60, 90):
    l2, h2 = ps.prepare_bboxes(bt.flatten(scenes.anim_frame(f)))
    lo, hi = np.minimum(lo, l2), np.maximum(hi, h2)
cs = scenes.cellsize_for(lo, hi, 256)
prm = ps.Params.make(cs, lo, hi)
rows = []
for f in range(0, 120, 2):
    flat = bt.flatten(scenes.anim_frame(f))
    t0 = time.perf_counter(); prog = ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
    t1 = time.perf_counter(); ctx.upload(prog)
    t2 = time.perf_counter(); m = ctx.polygonize(prm, download=False)
    t3 = time.perf_counter(); st = ctx.stats()
    rows.append([(t1 - t0) * 1e3, (t2 - t1) * 1e3, (t3 - t2) * 1e3, st["ms_total"], st["ms_cull"], st["ms_field"], st["ms_emit"], st["gpu_launches"]])
r = np.median(np.array(rows[5:]), axis=0)
print("anim: linearize %.3f upload %.3f polygonize wall %.3f | events total %.3f cull %.3f field %.3f emit %.3f launches %d | quads %d mpus %s"
      % (*r[:7], int(r[7]), prog.n_quads, ps.lattice_dims(cs, lo, hi)))
root = bt.binarize(scenes.small()); flat = bt.flatten(root)
lo, hi = ps.prepare_bboxes(flat); cs = scenes.cellsize_for(lo, hi, 128)
ctx.upload(ps.linearize(flat, ps.SEM_SIMD, 0))
prm = ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE)
rows = []
for i in range(30):
    t2 = time.perf_counter(); m = ctx.polygonize(prm, download=False); t3 = time.perf_counter(); st = ctx.stats()
    rows.append([(t3 - t2) * 1e3, st["ms_total"], st["ms_cull"], st["ms_field"], st["ms_scan"], st["ms_emit"], st["gpu_launches"]])
r = np.median(np.array(rows[5:]), axis=0)
print("small: polygonize wall %.3f | events total %.3f cull %.3f field %.3f scan %.3f emit %.3f launches %d" % (*r[:6], int(r[6])))
