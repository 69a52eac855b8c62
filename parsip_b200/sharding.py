"""z-slab sharding of the MPU lattice over ranks (one process per GPU) and the count/offset exchange.

The reference's only parallelism is a parallel_for over independent MPUs
(Parsip100/PS_SimdPoly/include/PS_Polygonizer.cpp:379-381); across GPUs the same independence lets the lattice be
cut into contiguous slabs of MPU layers along z (the reference's fastest MPU index, :360-370) with no data-path
collective: every rank holds the whole instruction stream, evaluates shared boundary corners from the same global
integer indices (bit-identical on both sides) and emits its own indexed mesh. The only exchange is an all-gather of
the per-rank {vertices, triangles} counts (16 bytes per rank; NCCL over NVLink when the backend is nccl), from which
every rank derives its global vertex / triangle offsets.
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def slab_bounds(n_layers: int, world: int, rank: int, weights: Sequence[float] | None = None) -> Tuple[int, int]:
    """[begin, end) MPU layers of `rank`. Without weights: contiguous slabs whose sizes differ by at most one layer.
    With per-layer weights (e.g. active-MPU counts from the culling pass, computed identically on every rank): the
    cut points that balance the weight prefix sum, every rank still getting a contiguous (possibly empty) slab."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    if weights is None:
        base, extra = divmod(n_layers, world)
        b = rank * base + min(rank, extra)
        return b, b + base + (1 if rank < extra else 0)
    if len(weights) != n_layers:
        raise ValueError("weights must have one entry per layer")
    total = float(sum(weights))
    if total <= 0.0:
        return slab_bounds(n_layers, world, rank)
    cuts = [0]
    acc, layer = 0.0, 0
    for r in range(1, world):
        target = total * r / world
        while layer < n_layers and acc + 0.5 * weights[layer] < target:
            acc += weights[layer]
            layer += 1
        cuts.append(layer)
    cuts.append(n_layers)
    return cuts[rank], cuts[rank + 1]


def offsets_from_counts(counts: Sequence[Tuple[int, int]], rank: int) -> Tuple[Tuple[int, int], Tuple[int, int]]:
    """Exclusive prefix of the per-rank (vertices, triangles) counts -> ((vertex offset, triangle offset), totals)."""
    ov = sum(c[0] for c in counts[:rank])
    ot = sum(c[1] for c in counts[:rank])
    return (ov, ot), (sum(c[0] for c in counts), sum(c[1] for c in counts))


def exchange_counts(n_vertices: int, n_triangles: int, device=None):
    """All-gathers {n_vertices, n_triangles} over the default process group (nccl on GPUs, gloo in the CPU tests) and
    returns ((vertex offset, triangle offset) of this rank, (total vertices, total triangles))."""
    import torch
    import torch.distributed as dist

    world, rank = dist.get_world_size(), dist.get_rank()
    mine = torch.tensor([int(n_vertices), int(n_triangles)], dtype=torch.int64, device=device)
    allc = torch.empty(world * 2, dtype=torch.int64, device=device)  # flat: gloo insists on a 1-D concatenation
    dist.all_gather_into_tensor(allc, mine)
    counts: List[Tuple[int, int]] = [(int(a), int(b)) for a, b in allc.view(world, 2).cpu().tolist()]
    return offsets_from_counts(counts, rank)


def rebase_triangles(tri, vertex_offset: int):
    """Triangle indices of a rank-local indexed mesh -> indices into the concatenated global vertex array."""
    return tri + type(tri.flat[0])(vertex_offset) if hasattr(tri, "flat") else tri + vertex_offset
