"""Seeded synthetic BlobTree scenes for BASELINE.json's configs.

All generators use numpy's MT19937 (`np.random.RandomState(seed)`), seeds fixed here.
Wyvill support radius is 1, the iso radius 0.454 in local units.
"""
from __future__ import annotations

import math

import numpy as np

from . import blobtree as bt


def _unit(rs):
    v = rs.normal(size=3)
    return v / np.linalg.norm(v)


def _f32(v):
    return np.asarray(v, np.float32)


def small(seed: int = 1) -> bt.Node:
    """S1 / config 1: 10 primitives (6 points, 4 lines of length 0.4-0.7) on a helix of radius 0.6,
    union(blend(5), blend(5))."""
    rs = np.random.RandomState(seed)
    prims = []
    for i in range(10):
        t = i / 10.0
        c = np.array([0.6 * math.cos(4 * math.pi * t), 0.6 * math.sin(4 * math.pi * t), 2.4 * t - 1.2])
        col = rs.uniform(0, 1, 3)
        if i % 5 in (1, 3):  # 4 lines
            d = _unit(rs)
            ln = rs.uniform(0.4, 0.7)
            prims.append(bt.Line(_f32(c - 0.5 * ln * d), _f32(c + 0.5 * ln * d), color=col))
        else:
            prims.append(bt.Point(_f32(c), color=col))
    return bt.Union(bt.Blend(*prims[:5]), bt.Blend(*prims[5:]))


def _blob_prims(rs, n, extent, kinds=("point", "line", "cylinder"), probs=(0.5, 0.3, 0.2)):
    out = []
    ks = rs.choice(len(kinds), size=n, p=probs)
    for i in range(n):
        c = rs.uniform(0, extent, 3)
        col = rs.uniform(0, 1, 3)
        k = kinds[ks[i]]
        if k == "point":
            out.append(bt.Point(_f32(c), color=col))
        elif k == "line":
            d = _unit(rs)
            ln = rs.uniform(0.2, 1.0)
            out.append(bt.Line(_f32(c - 0.5 * ln * d), _f32(c + 0.5 * ln * d), color=col))
        elif k == "cylinder":
            out.append(bt.Cylinder(_f32(c), _unit(rs), rs.uniform(0.05, 0.3), rs.uniform(0.2, 1.0), color=col))
        elif k == "disc":
            out.append(bt.Disc(_f32(c), _unit(rs), rs.uniform(0.1, 0.4), color=col))
        elif k == "ring":
            out.append(bt.Ring(_f32(c), _unit(rs), rs.uniform(0.3, 0.6), color=col))
        elif k == "cube":
            out.append(bt.Cube(_f32(c), rs.uniform(0.1, 0.3), color=col))
    return out


def blobs(n_prims: int, extent: float, seed: int, group: int = 20) -> bt.Node:
    """S2 'blob1k' (n=1000, extent 12, seed 2) and S4 'blob10k' (n=10000, extent 26, seed 4):
    50 % point / 30 % line / 20 % cylinder, union of blends of `group`."""
    rs = np.random.RandomState(seed)
    prims = _blob_prims(rs, n_prims, extent)
    groups = [bt.Blend(*prims[i:i + group]) if len(prims[i:i + group]) > 1 else prims[i]
              for i in range(0, n_prims, group)]
    return bt.Union(*groups) if len(groups) > 1 else groups[0]


def blob1k() -> bt.Node:
    return blobs(1000, 12.0, 2)


def blob10k() -> bt.Node:
    return blobs(10000, 26.0, 4)


def csgwarp(n_groups: int = 32, extent: float = 10.0, seed: int = 3, survey_rates: bool = False) -> bt.Node:
    """S3 / config 3: 8 primitives per group, all six skeleton kinds,
    bend(taper(twist(dif(intersect(blend(a, b, c, d), e), blend(f, g)) ... h))) chains; identity
    affines under warps.
    Warp rates: the warps of the compact tree act about the WORLD origin (twist angle = k * coordinate, CompactBlobTree.cpp:1355-1385),
    and the groups sit up to `extent` away from it, so SURVEY 8(d)'s rates (twist k in [0.2, 1], taper [-0.3, 0.3], bend [0.1, 0.5])
    wind a group at distance 10 by up to 10 radians. The default divides them by `extent` (same angles per group as the survey's
    numbers would give a group of unit size at the origin); survey_rates=True takes them as written."""
    rs = np.random.RandomState(seed)
    kinds = ("point", "line", "cylinder", "disc", "ring", "cube")
    groups = []
    for g in range(n_groups):
        centre = rs.uniform(1.5, extent - 1.5, 3)
        ps = []
        for q in range(8):
            k = kinds[(g + q) % 6]
            c = centre + rs.uniform(-0.45, 0.45, 3)
            col = rs.uniform(0, 1, 3)
            if k == "point":
                ps.append(bt.Point(_f32(c), color=col))
            elif k == "line":
                d = _unit(rs); ln = rs.uniform(0.3, 0.8)
                ps.append(bt.Line(_f32(c - 0.5 * ln * d), _f32(c + 0.5 * ln * d), color=col))
            elif k == "cylinder":
                ps.append(bt.Cylinder(_f32(c), _unit(rs), rs.uniform(0.05, 0.25), rs.uniform(0.2, 0.7), color=col))
            elif k == "disc":
                ps.append(bt.Disc(_f32(c), _unit(rs), rs.uniform(0.1, 0.35), color=col))
            elif k == "ring":
                ps.append(bt.Ring(_f32(c), _unit(rs), rs.uniform(0.25, 0.5), color=col))
            else:
                ps.append(bt.Cube(_f32(c), rs.uniform(0.1, 0.25), color=col))
        body = bt.Blend(ps[0], ps[1], ps[2], ps[3])
        carved = bt.Dif(bt.Intersect(body, bt.Blend(ps[4], ps[0], ps[1])), bt.Blend(ps[5], ps[6]))
        solid = bt.SmoothDif(bt.Union(carved, ps[7]), bt.Point(_f32(centre + rs.uniform(-0.6, 0.6, 3)),
                                                                color=rs.uniform(0, 1, 3)))
        den = 1.0 if survey_rates else extent
        w = bt.Twist(rs.uniform(0.2, 1.0) / den, int(rs.randint(0, 3)), solid)
        w = bt.Taper(rs.uniform(-0.3, 0.3) / den, int(rs.randint(0, 2)), 2, w)
        y0 = float(centre[1])
        w = bt.Bend(rs.uniform(0.1, 0.5) / den, y0, y0 - 1.0, y0 + 1.0, w)
        groups.append(w)
    return bt.Union(*groups)


def anim_frame(frame: int, n_frames: int = 120, seed: int = 5) -> bt.Node:
    """S5 / config 5: 200 primitives, a 20-primitive subtree translated along a closed path
    (the calling pattern of PH/CBlobTreeAnimation.cpp:37-80: translate, bump, re-polygonize)."""
    rs = np.random.RandomState(seed)
    prims = _blob_prims(rs, 200, 7.0)
    t = frame / float(n_frames)
    s = 0.5 - 0.5 * math.cos(math.pi * t)  # sine ease
    off = np.array([2.0 * math.sin(2 * math.pi * s), 1.5 * math.sin(4 * math.pi * s), 1.0 * math.cos(2 * math.pi * s) - 1.0])
    moved = []
    for p in prims[:20]:
        q = bt.Node(p.type, (), _f32(np.asarray(p.pos) + off),
                    _f32(np.asarray(p.dir) + off) if p.type == bt.PRIM_LINE else p.dir, p.res, p.color)
        moved.append(q)
    groups = [bt.Blend(*moved)] + [bt.Blend(*prims[i:i + 20]) for i in range(20, 200, 20)]
    return bt.Union(*groups)


def cellsize_for(lo, hi, n_cells: int) -> float:
    """Cell size that puts n_cells cells on the longest bbox axis."""
    ext = float(np.max(np.asarray(hi, np.float64) - np.asarray(lo, np.float64)))
    return float(np.float32(ext / n_cells))
