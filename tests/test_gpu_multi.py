"""-m gpu, needs >= 2 GPUs on the box (gpurun --gpus 2): the stitched z-slab mesh of psgpu_polygonize_sharded under real NCCL ranks
equals the single-GPU mesh. On a one-GPU box the collective cannot run; the single-rank path of the same entry point is checked."""
import os
import subprocess
import sys

import numpy as np
import pytest

import gpu_helpers as gh
import parsip_b200 as ps
from parsip_b200 import blobtree as bt, scenes

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _n_gpus():
    return ps.load_library().psgpu_device_count()


@pytest.mark.parametrize("sem", ["simd", "scalar"])
@pytest.mark.parametrize("world", [2, 4])
def test_sharded_mesh_equals_single_gpu_mesh(world, sem):
    if _n_gpus() < world:
        pytest.skip("needs %d GPUs on the box (run with gpurun --gpus %d)" % (world, world))
    env = dict(os.environ, PSGPU_TEST_SEM=sem)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", str(world), "--master-addr", "127.0.0.1",
           "--master-port", str(29500 + world + (7 if sem == "scalar" else 0)), os.path.join(ROOT, "tests", "multi_gpu_worker.py")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    print(r.stdout[-2000:])
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "soup_equal=True" in r.stdout and "index_max_ok=True" in r.stdout


def test_sharded_entry_point_single_rank(ctx):
    """Without a communicator psgpu_polygonize_sharded is psgpu_polygonize; explicit slabs still stitch by hand (vertex ids local)."""
    root = scenes.blobs(200, 6.0, 5)
    flat = bt.flatten(bt.binarize(root))
    ctx.upload(ps.linearize(flat, ps.SEM_SIMD, 0))
    lo, hi = ps.prepare_bboxes(flat)
    cs = scenes.cellsize_for(lo, hi, 96)
    prm = ps.Params.make(cs, lo, hi, root_mode=ps.ROOT_FOURSAMPLE)
    a = ctx.polygonize(prm)
    b = ctx.polygonize_sharded(prm, download=True)
    assert b["shard"]["nranks"] == 1 and b["shard"]["total_vertices"] == a["n_vertices"]
    assert np.array_equal(gh.triangle_soup(a["pos"], a["tri"]), gh.triangle_soup(b["pos"], b["tri"]))
    assert ps.slab_bounds(69, 8, 3) == (27, 36) and ps.slab_bounds(10, 2, 1, [0, 0, 0, 0, 1, 1, 1, 1, 1, 1]) == (7, 10)
