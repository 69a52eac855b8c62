"""Development aid: wall-clock cost of psgpu_upload_program + the first polygonize after it (slab program rebuilt) against a polygonize
on an unchanged program, on one rank's slab of blob10k@1024^3."""
import os, sys, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes, sharding

root = scenes.blob10k()
flat = bt.flatten(bt.binarize(root))
lo, hi = ps.prepare_bboxes(flat)
cs = scenes.cellsize_for(lo, hi, 1024)
ctx = ps.Context(0)
prog = ps.linearize(flat, ps.SEM_SIMD, 0)
ctx.upload(prog)
prm = ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE)
prm.z_begin, prm.z_end = sharding.slab_bounds(ps.lattice_dims(cs, lo, hi)[2], 8, 3)
for i in range(3):
    ctx.polygonize(prm, download=False)
for mode in ("same program", "upload each step"):
    ts = []
    for i in range(6):
        t = time.perf_counter()
        if mode != "same program":
            ctx.upload(prog)
        t1 = time.perf_counter()
        ctx.polygonize(prm, download=False)
        ts.append((round((t1 - t) * 1e3, 3), round((time.perf_counter() - t1) * 1e3, 3), round(ctx.stats()["ms_total"], 3)))
    print(mode, "(upload ms, polygonize wall ms, event ms):", ts[2:])
