"""Development aid: device prune_walk vs host-emulated prune_walk on every MPU box of the csgwarp lattice."""
import os, sys
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
prog = ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
quads = prog.quads()
ctx = ps.Context(0)
ctx.upload(prog)
side = np.float32(cs) * np.float32(7)
dims = ps.lattice_dims(cs, lo, hi)
ndiff = 0
for I in range(dims[0]):
    for J in range(dims[1]):
        for K in range(dims[2]):
            a = (lo + np.float32([I, J, K]) * side).astype(np.float32) - np.float32(6e-5)
            b = (a + side).astype(np.float32) + np.float32(1.2e-4)
            g = ctx.prune_region(a, b, ps.CULL_SUPPORT)
            h = emu.prune(quads, a, b, ps.CULL_SUPPORT)
            same = (g is None and h is None) or (g is not None and h is not None and np.array_equal(g, h))
            if not same:
                ndiff += 1
                if ndiff <= 4:
                    print("DIFF", (I, J, K), "gpu", None if g is None else len(g), "host", None if h is None else len(h))
                    for name, q in (("gpu", g), ("host", h)):
                        if q is not None:
                            print("  ", name, [(int(w & 31), int((w >> 5) & 15)) for w in q[:, 0] if (w & 31) not in (0,)][:40])
print("boxes", dims[0] * dims[1] * dims[2], "differences", ndiff)
