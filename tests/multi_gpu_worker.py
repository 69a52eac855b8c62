"""Worker of tests/test_gpu_multi.py, one process per GPU (python -m torch.distributed.run --nproc-per-node N tests/multi_gpu_worker.py).
Every rank polygonizes its z-slab with psgpu_polygonize_sharded; the shares are gathered on rank 0 with psgpu_mesh_gather and
compared with the mesh ONE GPU produces for the whole lattice: same vertex count, same triangle soup, bit for bit."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import gpu_helpers as gh  # noqa: E402
import parsip_b200 as ps  # noqa: E402
from parsip_b200 import blobtree as bt, scenes  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sem = os.environ.get("PSGPU_TEST_SEM", "simd")
    prims = int(os.environ.get("PSGPU_TEST_PRIMS", "600"))
    cells = int(os.environ.get("PSGPU_TEST_CELLS", "160"))
    root = scenes.blobs(prims, 9.0, 11)
    if sem == "simd":
        flat = bt.flatten(bt.binarize(root))
        prog = ps.linearize(flat, ps.SEM_SIMD, 0)
        lo, hi = ps.prepare_bboxes(flat)
        mk = lambda **kw: ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE, **kw)
    else:
        prog = ps.linearize(bt.flatten(root), ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
        lo, hi = ps.prepare_bboxes(bt.flatten(bt.binarize(root)))
        mk = lambda **kw: ps.Params.make(cs, lo, hi, **kw)
    cs = scenes.cellsize_for(lo, hi, cells)
    ctx = ps.Context(local)
    ctx.upload(prog)
    # the 128-byte NCCL id travels over the application's own channel (here: a torch.distributed broadcast)
    idt = torch.zeros(128, dtype=torch.uint8, device="cuda")
    if rank == 0:
        idt.copy_(torch.frombuffer(bytearray(ps.comm_unique_id()), dtype=torch.uint8))
    dist.broadcast(idt, 0)
    ctx.comm_init(world, rank, bytes(idt.cpu().numpy().tobytes()))
    m = ctx.polygonize_sharded(mk())
    sh = m["shard"]
    assert sh["stitched"] == 1 and sh["nranks"] == world and sh["rank"] == rank
    st = ctx.stats()
    nv, nt = sh["total_vertices"], sh["total_triangles"]
    if rank == 0:
        pos = torch.empty((nv, 3), dtype=torch.float32, device="cuda")
        nrm, col = torch.empty_like(pos), torch.empty_like(pos)
        tri = torch.empty((nt, 3), dtype=torch.int32, device="cuda")
        ctx.mesh_gather(0, pos.data_ptr(), nrm.data_ptr(), col.data_ptr(), tri.data_ptr())
    else:
        ctx.mesh_gather(0, 1, 1, 1, 1)
    ok = True
    if rank == 0:
        P, T = pos.cpu().numpy(), tri.cpu().numpy().view(np.uint32)
        N, Cc = nrm.cpu().numpy(), col.cpu().numpy()
        ref = ps.Context(local)
        ref.upload(prog)
        one = ref.polygonize(mk())
        same_counts = (one["n_vertices"], one["n_triangles"]) == (nv, nt)
        soup = np.array_equal(gh.triangle_soup(P, T), gh.triangle_soup(one["pos"], one["tri"]))
        # normals and colours ride on the same vertices: compare the sorted (pos | nrm | col) rows
        rows = lambda p, n, c: (lambda a: a[np.lexsort(a.view(np.uint32).T[::-1])])(np.ascontiguousarray(np.concatenate([p, n, c], axis=1)))
        attrs = same_counts and np.array_equal(rows(P, N, Cc).view(np.uint32), rows(one["pos"], one["nrm"], one["col"]).view(np.uint32))
        in_range = int(T.max()) == nv - 1 if nt else True
        ok = same_counts and soup and in_range and attrs
        print("RESULT world=%d sem=%s vertices=%d (single GPU %d) triangles=%d (single GPU %d) soup_equal=%s attrs_equal=%s index_max_ok=%s exchange_ms=%.3f cull_ms=%.3f"
              % (world, sem, nv, one["n_vertices"], nt, one["n_triangles"], soup, attrs, in_range, sh["ms_exchange"], st["ms_cull"]), flush=True)
    flag = torch.tensor([1 if ok else 0], device="cuda")
    dist.broadcast(flag, 0)
    dist.destroy_process_group()
    sys.exit(0 if int(flag.item()) == 1 else 1)


if __name__ == "__main__":
    main()
