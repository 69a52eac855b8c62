"""CPU checks of the packed-pair STRICT interpreter (parsip_b200/csrc/interp_packed.cuh compiled for the host): it must give
the same BITS as the scalar STRICT interpreter of interp.cuh, which tests/test_linearizer.py pins to the reference goldens.
The device build replaces the host's 1/b and 1/sqrt(x) by MUFU approximations + the same correction steps; the -m gpu
parity tests cover that side (field bits vs the goldens through the field kernel)."""
import numpy as np
import pytest

import hostemu_binding as emu
import scenes_small
from parsip_b200 import blobtree as bt
import parsip_b200 as ps


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def _program(name):
    root = bt.binarize(scenes_small.SIMD_SCENES[name]())
    flat = bt.flatten(root)
    return ps.linearize(flat, ps.SEM_SIMD, 0).quads(), ps.prepare_bboxes(flat)


@pytest.mark.parametrize("name", sorted(scenes_small.SIMD_SCENES))
def test_packed_general_form_equals_scalar_strict(name):
    quads, (lo, hi) = _program(name)
    rs = np.random.RandomState(3)
    pts = rs.uniform(lo - 0.3, hi + 0.3, (4000, 3)).astype(np.float32)
    ref = emu.field(quads, pts, ps.SEM_SIMD)
    got = emu.field_packed(quads, pts)
    assert np.array_equal(_bits(got), _bits(ref)), "max |d| = %g" % np.abs(got - ref).max()
    assert (ref > 0).sum() > 100


@pytest.mark.parametrize("name", sorted(scenes_small.SIMD_SCENES))
def test_packed_column_form_equals_scalar_strict(name):
    quads, (lo, hi) = _program(name)
    rs = np.random.RandomState(4)
    n = 600
    xy = rs.uniform(lo[:2], hi[:2], (n, 2)).astype(np.float32)
    z0 = rs.uniform(lo[2], hi[2], (n, 1)).astype(np.float32)
    z = (z0 + np.float32(0.02) * np.arange(8, dtype=np.float32)[None, :]).astype(np.float32)
    got = emu.field_packed_columns(quads, xy, z)
    pts = np.stack([np.repeat(xy[:, 0], 8), np.repeat(xy[:, 1], 8), z.reshape(-1)], axis=1)
    ref = emu.field(quads, pts, ps.SEM_SIMD).reshape(n, 8)
    assert np.array_equal(_bits(got), _bits(ref))


def test_packed_paths_on_degenerate_geometry():
    """Points ON a cylinder axis (sqrt of 0 / of a rounding-negative number: the reference's NaN), on a line, at a cube face,
    and primitives at the origin: the fast-path tests must hand these to the scalar code or agree with it."""
    col = (0.5, 0.5, 0.5)
    d = np.float32([0.3, -0.5, 0.81]); d /= np.linalg.norm(d)
    root = bt.Union(bt.Blend(bt.Cylinder([0, 0, 0], d, 0.2, 0.7, col), bt.Line([0, 0, 0], [0.5, 0.25, -0.125], col)),
                    bt.Blend(bt.Cube([0.5, 0.5, 0.5], 0.25, col), bt.Cylinder([1, 1, 1], [0, 0, 1], 0.1, 0.5, col)))
    quads = ps.linearize(bt.flatten(bt.binarize(root)), ps.SEM_SIMD, 0).quads()
    t = np.linspace(-1.5, 1.5, 401, dtype=np.float32)
    on_axis = (t[:, None] * d[None, :]).astype(np.float32)
    on_axis2 = np.stack([np.ones_like(t), np.ones_like(t), 1 + t], axis=1).astype(np.float32)
    on_line = (t[:, None] * np.float32([0.5, 0.25, -0.125])[None, :]).astype(np.float32)
    faces = np.float32([[0.25, 0.5, 0.5], [0.75, 0.75, 0.75], [0.5, 0.5, 0.5], [0, 0, 0], [0.25, 0.25, 0.25]])
    pts = np.concatenate([on_axis, on_axis2, on_line, faces])
    pts = np.concatenate([pts, pts[: (-len(pts)) % 4]])
    ref = emu.field(quads, pts, ps.SEM_SIMD)
    got = emu.field_packed(quads, pts)
    same = (_bits(got) == _bits(ref)) | (np.isnan(got) & np.isnan(ref))
    assert same.all(), np.nonzero(~same)[0][:10]


def test_packed_division_matches_ieee_on_random_lines():
    """The unclamped line divides by |dir|^2 per point; the packed form multiplies by a reciprocal and corrects with two FMAs."""
    rs = np.random.RandomState(8)
    col = (1, 1, 1)
    prims = []
    for _ in range(40):
        a = rs.uniform(-2, 2, 3)
        ln = 10.0 ** rs.uniform(-3, 1)
        v = rs.normal(size=3); v /= np.linalg.norm(v)
        prims.append(bt.Line(np.float32(a), np.float32(a + ln * v), col))
    root = bt.Blend(*prims)
    quads = ps.linearize(bt.flatten(bt.binarize(root)), ps.SEM_SIMD, 0).quads()
    pts = rs.uniform(-3, 3, (20000, 3)).astype(np.float32)
    assert np.array_equal(_bits(emu.field_packed(quads, pts)), _bits(emu.field(quads, pts, ps.SEM_SIMD)))


@pytest.mark.parametrize("name", sorted(scenes_small.SIMD_SCENES))
def test_packed_walk_with_colour_equals_scalar_strict(name):
    """eval_program_packed<.., COLOR> (the colour + normal pass of color_kernel<.., NRM>): fields at (p, p+dx, p+dy, p+dz) and the colour
    at p against the scalar STRICT walk - on the full program and on pruned ones, whose colour-scale markers (OPC_CSCALE, header counts,
    PRE-counted folded groups) only exist after culling."""
    root = bt.binarize(scenes_small.SIMD_SCENES[name]())
    flat = bt.flatten(root)
    prog = ps.linearize(flat, ps.SEM_SIMD, 0)
    quads, colors = prog.quads(), prog.colors()
    lo, hi = ps.prepare_bboxes(flat)
    rs = np.random.RandomState(6)
    dl = np.float32(1e-3)

    def check(q, pts):
        f4, col = emu.field_color_packed(q, colors, pts, dl)
        for k, off in enumerate(([0, 0, 0], [dl, 0, 0], [0, dl, 0], [0, 0, dl])):
            assert np.array_equal(_bits(f4[:, k]), _bits(emu.field(q, (pts + np.float32(off)).astype(np.float32), ps.SEM_SIMD)))
        fr, cr = emu.field_color(q, colors, pts, ps.SEM_SIMD)
        assert np.array_equal(_bits(f4[:, 0]), _bits(fr)) and np.array_equal(_bits(col), _bits(cr)), np.abs(col - cr).max()
        return fr

    f = check(quads, rs.uniform(lo, hi, (3000, 3)).astype(np.float32))
    assert (f > 0.3).sum() > 50
    pruned = 0
    for _ in range(60):  # pruned twice, as the hierarchy does (exact colours: markers ride on headers, fold into PRE counts, overflow into quads)
        a = rs.uniform(lo, hi - 0.6, 3).astype(np.float32)
        big = emu.prune(quads, a, a + np.float32(0.6), ps.CULL_INTERVAL, 0.5, True)
        if big is None:
            continue
        b = (a + rs.uniform(0, 0.45, 3)).astype(np.float32)
        small = emu.prune(big, b, b + np.float32(0.15), ps.CULL_INTERVAL, 0.5, True)
        if small is None:
            continue
        pruned += 1
        pts = rs.uniform(b, b + np.float32(0.149), (64, 3)).astype(np.float32)  # + delta stays inside the box's pad? compare like with like:
        f4, col = emu.field_color_packed(small, colors, pts, dl)
        fr, cr = emu.field_color(small, colors, pts, ps.SEM_SIMD)
        assert np.array_equal(_bits(f4[:, 0]), _bits(fr)) and np.array_equal(_bits(col), _bits(cr))
        fq, cq = emu.field_color(quads, colors, pts, ps.SEM_SIMD)  # and the pruned colour is the full program's colour
        assert np.array_equal(_bits(col), _bits(cq))
    assert pruned >= 5
