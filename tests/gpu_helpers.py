"""Helpers shared by the -m gpu parity tests."""
import hashlib

import numpy as np


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def lattice_ordered(m):
    """PSGPU_OUT_MPU result -> (pos, nrm, col, tri_global, mpu_ijk, counts) with MPUs in the reference's lattice order
    (x outer, z inner: PS_Polygonizer.cpp:360-370) and triangle indices rebased onto the concatenated vertex array, i.e.
    the byte layout of the reference's PolyMPUs content concatenated over MPUs with triangles."""
    ijk = m["mpu_ijk"]
    order = np.lexsort((ijk[:, 2], ijk[:, 1], ijk[:, 0]))
    vb, tb = m["mpu_vbase"].astype(np.int64), m["mpu_tbase"].astype(np.int64)
    pos, nrm, col, tri, counts = [], [], [], [], []
    base = 0
    for r in order:
        v0, v1, t0, t1 = vb[r], vb[r + 1], tb[r], tb[r + 1]
        pos.append(m["pos"][v0:v1]); nrm.append(m["nrm"][v0:v1]); col.append(m["col"][v0:v1])
        tri.append(m["tri"][t0:t1].astype(np.int64) + base)
        counts.append((v1 - v0, t1 - t0))
        base += v1 - v0
    cat = lambda xs, w, dt: (np.concatenate(xs).astype(dt) if xs else np.zeros((0, w), dt))
    return (cat(pos, 3, np.float32), cat(nrm, 3, np.float32), cat(col, 3, np.float32), cat(tri, 3, np.uint32), ijk[order],
            np.asarray(counts, np.int64).reshape(-1, 2))


def triangle_soup(pos, tri):
    """Order-independent form of an indexed mesh: each triangle's three positions, rotated so the smallest vertex comes
    first (orientation kept), triangles sorted. Two meshes with the same soup are the same surface."""
    p = pos[tri.astype(np.int64)]  # (T, 3, 3)
    key = p.view(np.uint32).astype(np.uint64)
    k = (key[..., 0] << np.uint64(42)) ^ (key[..., 1] << np.uint64(21)) ^ key[..., 2]  # (T, 3) per-vertex key
    first = np.argmin(k, axis=1)
    idx = (first[:, None] + np.arange(3)[None, :]) % 3
    p = np.take_along_axis(p, idx[:, :, None], axis=1).reshape(-1, 9)
    order = np.lexsort(p.T[::-1])
    return p[order]


def edge_manifold_stats(tri):
    """(number of undirected edges, edges used once, edges used more than twice) of an indexed mesh."""
    t = tri.astype(np.int64)
    e = np.concatenate([t[:, [0, 1]], t[:, [1, 2]], t[:, [2, 0]]])
    e.sort(axis=1)
    key = e[:, 0] * (int(t.max()) + 1 if len(t) else 1) + e[:, 1]
    _, cnt = np.unique(key, return_counts=True)
    return len(cnt), int((cnt == 1).sum()), int((cnt > 2).sum())


def point_triangle_distance(p, a, b, c):
    """Distances from points p (N,3) to triangles (a,b,c) (N,3) each, float64 (Ericson, Real-Time Collision Detection 5.1.5)."""
    ab, ac, ap = b - a, c - a, p - a
    d1, d2 = (ab * ap).sum(1), (ac * ap).sum(1)
    bp = p - b
    d3, d4 = (ab * bp).sum(1), (ac * bp).sum(1)
    cp = p - c
    d5, d6 = (ab * cp).sum(1), (ac * cp).sum(1)
    va, vb, vc = d3 * d6 - d5 * d4, d5 * d2 - d1 * d6, d1 * d4 - d3 * d2
    out = np.empty_like(p)
    done = np.zeros(len(p), bool)

    def put(mask, q):
        m = mask & ~done
        out[m] = q[m]
        done[m] = True

    with np.errstate(divide="ignore", invalid="ignore"):
        put((d1 <= 0) & (d2 <= 0), a)
        put((d3 >= 0) & (d4 <= d3), b)
        put((d6 >= 0) & (d5 <= d6), c)
        put((vc <= 0) & (d1 >= 0) & (d3 <= 0), a + ab * (d1 / (d1 - d3))[:, None])
        put((vb <= 0) & (d2 >= 0) & (d6 <= 0), a + ac * (d2 / (d2 - d6))[:, None])
        put((va <= 0) & (d4 - d3 >= 0) & (d5 - d6 >= 0), b + (c - b) * ((d4 - d3) / ((d4 - d3) + (d5 - d6)))[:, None])
        den = 1.0 / (va + vb + vc)
        put(np.ones(len(p), bool), a + ab * (vb * den)[:, None] + ac * (vc * den)[:, None])
    d = np.linalg.norm(p - out, axis=1)
    # degenerate triangles (the scalar twin's shrink-wrap Newton collapses edges that share a start corner): nearest vertex
    dv = np.minimum(np.minimum(np.linalg.norm(ap, axis=1), np.linalg.norm(bp, axis=1)), np.linalg.norm(cp, axis=1))
    return np.where(np.isfinite(d), np.minimum(d, dv), dv)


def one_sided_hausdorff(pts, pos, tri, k=24):
    """max over pts of the distance to the mesh (pos, tri): k nearest triangle centroids as candidates (meshes here are
    uniform marching-cubes meshes, so the closest triangle is among the nearest centroids)."""
    from scipy.spatial import cKDTree
    P = pos.astype(np.float64)
    T = tri.astype(np.int64)
    cen = P[T].mean(axis=1)
    tree = cKDTree(cen)
    k = min(k, len(T))
    _, nb = tree.query(pts.astype(np.float64), k=k)
    nb = nb.reshape(len(pts), k)
    best = np.full(len(pts), np.inf)
    for j in range(k):
        t = T[nb[:, j]]
        d = point_triangle_distance(pts.astype(np.float64), P[t[:, 0]], P[t[:, 1]], P[t[:, 2]])
        best = np.minimum(best, d)
    return float(best.max())
