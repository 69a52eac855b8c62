"""Measures the BASELINE configs that are not the bench workload (development aid; results go to profiles/):
  configs[0]  small BlobTree (10 point/line primitives, union + blend) at 128^3   - PS_SimdPoly compat semantics
  configs[2]  CSG + warp-heavy tree (difference / intersection, twist / taper / bend) at 512^3 - native semantics (warps exist only there)
  configs[4]  animated BlobTree re-polygonized per frame at 256^3: re-linearize + upload + polygonize, ms per frame
One JSON line per config."""
import json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes

ctx = ps.Context(0)


def run(prog, prm, reps=7):
    ctx.upload(prog)
    ms, st, m = [], None, None
    for _ in range(reps):
        m = ctx.polygonize(prm, download=False)
        st = ctx.stats()
        ms.append(st["ms_total"])
    return float(np.median(ms[2:])), st, m


# configs[0]
root = bt.binarize(scenes.small())
flat = bt.flatten(root)
lo, hi = ps.prepare_bboxes(flat)
cs = scenes.cellsize_for(lo, hi, 128)
ms, st, m = run(ps.linearize(flat, ps.SEM_SIMD, 0), ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE))
dims = ps.lattice_dims(cs, lo, hi)
print(json.dumps({"config": "configs[0] small@128^3 compat", "ms": ms, "mpus": st["mpus_total"], "vertices": m["n_vertices"], "triangles": m["n_triangles"],
                  "cells_per_s": dims[0] * dims[1] * dims[2] * 343 / (ms * 1e-3), "stages": {k: st[k] for k in ("ms_cull", "ms_field", "ms_tri", "ms_vertex", "ms_color")}}))

# configs[2]
root = scenes.csgwarp()
flat = bt.flatten(root)
lo, hi = ps.prepare_bboxes(flat)
cs = scenes.cellsize_for(lo, hi, 512)
prog = ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
ms, st, m = run(prog, ps.Params.make(cs, lo, hi))
dims = ps.lattice_dims(cs, lo, hi)
info = prog.info() if hasattr(prog, "info") else None
print(json.dumps({"config": "configs[2] csgwarp@512^3 native", "ms": ms, "mpus": st["mpus_total"], "evaluated": st["mpus_evaluated"], "crossed": st["mpus_crossed"],
                  "vertices": m["n_vertices"], "triangles": m["n_triangles"], "cells_per_s": dims[0] * dims[1] * dims[2] * 343 / (ms * 1e-3),
                  "program_quads": st["program_quads"], "stages": {k: st[k] for k in ("ms_cull", "ms_field", "ms_tri", "ms_vertex", "ms_color")}}))

# configs[4]: per frame = build flat tree (host), linearize, upload, polygonize (mesh stays on the device for rendering)
f0 = bt.flatten(scenes.anim_frame(0))
lo, hi = ps.prepare_bboxes(f0)
for f in (30, 60, 90):  # the grid domain must hold every frame
    l2, h2 = ps.prepare_bboxes(bt.flatten(scenes.anim_frame(f)))
    lo, hi = np.minimum(lo, l2), np.maximum(hi, h2)
cs = scenes.cellsize_for(lo, hi, 256)
prm = ps.Params.make(cs, lo, hi)
frames, gpu_ms, lin_ms = [], [], []
for f in range(120):
    flat = bt.flatten(scenes.anim_frame(f))
    t0 = time.perf_counter()
    prog = ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
    t1 = time.perf_counter()
    ctx.upload(prog)
    m = ctx.polygonize(prm, download=False)
    t2 = time.perf_counter()
    frames.append((t2 - t0) * 1e3); lin_ms.append((t1 - t0) * 1e3); gpu_ms.append(ctx.stats()["ms_total"])
fr = np.array(frames[5:])
print(json.dumps({"config": "configs[4] anim@256^3 native, 120 frames", "ms_per_frame_median": float(np.median(fr)), "ms_per_frame_p99": float(np.percentile(fr, 99)),
                  "linearize_ms_median": float(np.median(lin_ms[5:])), "gpu_ms_median": float(np.median(gpu_ms[5:])),
                  "includes": "psgpu_linearize + psgpu_upload_program + psgpu_polygonize (wall clock); tree construction in Python excluded",
                  "vertices_last": m["n_vertices"], "triangles_last": m["n_triangles"]}))
