"""z-slab sharding helpers and the count/offset exchange (world_size-2 gloo on CPU; NCCL on the GPU box)."""
import os
import socket
import subprocess
import sys

import numpy as np
import pytest

from parsip_b200 import sharding

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("n_layers,world", [(69, 1), (69, 2), (69, 4), (69, 8), (147, 8), (3, 8), (8, 8)])
def test_slab_bounds_partition_the_lattice(n_layers, world):
    slabs = [sharding.slab_bounds(n_layers, world, r) for r in range(world)]
    assert slabs[0][0] == 0 and slabs[-1][1] == n_layers
    for (b0, e0), (b1, e1) in zip(slabs, slabs[1:]):
        assert e0 == b1 and b0 <= e0
    sizes = [e - b for b, e in slabs]
    assert max(sizes) - min(sizes) <= 1


def test_weighted_slab_bounds_balance_the_active_mpus():
    rs = np.random.RandomState(0)
    w = np.concatenate([np.zeros(10), rs.uniform(50, 100, 40), np.zeros(19)])  # empty layers at both ends
    slabs = [sharding.slab_bounds(len(w), 4, r, list(w)) for r in range(4)]
    assert slabs[0][0] == 0 and slabs[-1][1] == len(w)
    loads = [w[b:e].sum() for b, e in slabs]
    assert max(loads) < 1.15 * (w.sum() / 4)
    with pytest.raises(ValueError):
        sharding.slab_bounds(10, 2, 0, [1.0] * 9)
    with pytest.raises(ValueError):
        sharding.slab_bounds(10, 2, 2)


def test_offsets_from_counts():
    counts = [(10, 20), (0, 0), (7, 9)]
    assert sharding.offsets_from_counts(counts, 0) == ((0, 0), (17, 29))
    assert sharding.offsets_from_counts(counts, 2) == ((10, 20), (17, 29))
    tri = np.array([[0, 1, 2]], np.uint32)
    assert np.array_equal(sharding.rebase_triangles(tri, 10), [[10, 11, 12]])


_WORKER = r"""
import os, sys
sys.path.insert(0, %r)
import torch.distributed as dist
from parsip_b200 import sharding
dist.init_process_group("gloo")
r, w = dist.get_rank(), dist.get_world_size()
b, e = sharding.slab_bounds(69, w, r)
offs, tot = sharding.exchange_counts(100 * (r + 1) + (e - b), 1000 * (r + 1))
print("RESULT", r, b, e, offs[0], offs[1], tot[0], tot[1], flush=True)
dist.destroy_process_group()
"""


def test_exchange_counts_world2_gloo(tmp_path):
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    script = tmp_path / "worker.py"
    script.write_text(_WORKER % ROOT)
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    res = {}
    for p in procs:
        out, err = p.communicate(timeout=180)
        assert p.returncode == 0, err[-2000:]
        for line in out.splitlines():
            if line.startswith("RESULT"):
                v = [int(x) for x in line.split()[1:]]
                res[v[0]] = v[1:]
    # slabs 35 + 34 layers; counts (135, 1000) and (234, 2000)
    assert res[0] == [0, 35, 0, 0, 369, 3000]
    assert res[1] == [35, 69, 135, 1000, 369, 3000]
