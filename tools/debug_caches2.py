import os, sys, ctypes as C
R = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes
import scenes_small, hostemu_binding as emu

root = scenes_small.SCALAR_SCENES["csgwarp"]()
flat = bt.flatten(root)
lo, hi = ps.prepare_bboxes(bt.flatten(bt.binarize(root)))
lo, hi = lo - 0.25, hi + 0.25
cs = scenes.cellsize_for(lo, hi, 64)
ctx = ps.Context(0)
prog = ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
quads = prog.quads()
ctx.upload(prog)
prm = ps.Params.make(cs, lo, hi, coord_mode=ps.COORD_MPU_LOCAL, output=ps.OUT_MPU, cull=ps.CULL_SUPPORT, root_mode=ps.ROOT_NEWTON_REF)
ijk, fv = ctx.mpu_caches_culled(prm)
lib = ctx.lib
lib.psgpu_debug_candidate_program.argtypes = [C.c_void_p, C.c_int, C.c_void_p, C.c_uint32, C.c_void_p]
shown = 0
ii = np.arange(8)
for idx, (k, f) in enumerate(zip(ijk, fv)):
    full, _ = ctx.mpu_cache(prm, k)
    if np.array_equal(f.view(np.uint32), full.view(np.uint32)):
        continue
    out = np.zeros((4096, 4), np.uint32); n = C.c_uint32()
    lib.psgpu_debug_candidate_program(ctx.h, idx, out.ctypes.data, 4096, C.byref(n))
    pipe = out[:n.value]
    side = np.float32(cs) * np.float32(7)
    a = (lo + np.float32(k) * side).astype(np.float32) - np.float32(6e-5)
    b = (a + side) + np.float32(1.2e-4)
    g = ctx.prune_region(a, b, ps.CULL_SUPPORT)
    h = emu.prune(quads, a, b, ps.CULL_SUPPORT)
    org = (lo + np.float32(k) * side).astype(np.float32)
    X = org[0] + ii.astype(np.float32) * np.float32(cs); Y = org[1] + ii.astype(np.float32) * np.float32(cs); Z = org[2] + ii.astype(np.float32) * np.float32(cs)
    # fv layout: i*8 + j*64 + k
    pts = np.zeros((512, 3), np.float32)
    for i in range(8):
        for j in range(8):
            for kk in range(8):
                pts[i * 8 + j * 64 + kk] = (X[i], Y[j], Z[kk])
    hf_full = emu.field(quads, pts, 1); hf_pipe = emu.field(pipe, pts, 1); hf_h = emu.field(h, pts, 1)
    gf_pts = ctx.field(pts)
    print("MPU", tuple(int(v) for v in k), "pipe len", len(pipe), "region len", None if g is None else len(g), "host len", None if h is None else len(h),
          "pipe==region", g is not None and np.array_equal(pipe, g), "| host eval: full vs pipeprog maxdiff", np.abs(hf_full - hf_pipe).max(),
          "full vs hostprog", np.abs(hf_full - hf_h).max(), "| gpu cache(pruned) vs host eval of pipe prog", np.abs(f - hf_pipe).max(),
          "| gpu full cache vs host full", np.abs(full - hf_full).max(), "gpu points-kernel vs host full", np.abs(gf_pts - hf_full).max())
    shown += 1
    if shown >= 4:
        break
    # (appended) where do they differ?
print("---- diff detail for last MPU")
d = np.where((pipe != g).any(axis=1))[0]
print("differing quads", d[:20], "of", len(pipe))
for q in d[:6]:
    print(q, "pipe", [hex(int(w)) for w in pipe[q]], pipe[q].view(np.float32), "region", [hex(int(w)) for w in g[q]], g[q].view(np.float32))
