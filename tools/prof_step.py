"""Profiling target (development aid): the N=1 bench workload (blob1k@512^3, BASELINE configs[1]), 1 warm-up + 2 steps,
no CPU legs - short enough to sit under ncu. `--native` runs the canonical semantics instead of the PS_SimdPoly compat ones."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes

native = "--native" in sys.argv
cells = 512
for a in sys.argv[1:]:
    if a.startswith("--cells="):
        cells = int(a.split("=")[1])
slab = None  # --slab=r/R : rank r's z-slab of R on blob10k@1024^3 (one rank of BASELINE configs[3])
for a in sys.argv[1:]:
    if a.startswith("--slab="):
        slab = tuple(int(v) for v in a.split("=")[1].split("/"))
        cells = 1024
zrange = None  # --zrange=a,b : an explicit slab of MPU layers of blob10k@1024^3
for a in sys.argv[1:]:
    if a.startswith("--zrange="):
        zrange = tuple(int(v) for v in a.split("=")[1].split(","))
        slab = (0, 1)
        cells = 1024
csgwarp = "--csgwarp" in sys.argv  # BASELINE configs[2], native semantics only (the SIMD path ignores warps)
root = scenes.csgwarp() if csgwarp else (scenes.blob10k() if slab else scenes.blob1k())
native = native or csgwarp
flat_bin = bt.flatten(bt.binarize(root))
lo, hi = ps.prepare_bboxes(flat_bin)
cs = scenes.cellsize_for(lo, hi, cells)
ctx = ps.Context(0)
if native:
    ctx.upload(ps.linearize(bt.flatten(root), ps.SEM_SCALAR, ps.LIN_REASSOCIATE))
    prm = ps.Params.make(cs, lo, hi)
else:
    ctx.upload(ps.linearize(flat_bin, ps.SEM_SIMD, 0))
    prm = ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE)
if slab:
    os.environ.setdefault("PSGPU_SLAB_PROGRAM_NOW", "1")  # the steady state of a sharded run: the slab program exists from the third frame on
    from parsip_b200 import sharding
    prm.z_begin, prm.z_end = zrange if zrange else sharding.slab_bounds(ps.lattice_dims(cs, lo, hi)[2], slab[1], slab[0])
for i in range(3):
    m = ctx.polygonize(prm, download=False)
    st = ctx.stats()
    print({k: (round(v, 3) if isinstance(v, float) else v) for k, v in st.items()}, m["n_vertices"], m["n_triangles"])
