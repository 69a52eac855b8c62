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
