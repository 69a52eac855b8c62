"""CPU checks of the per-MPU pruning walk (parsip_b200/csrc/cull_walk.cuh compiled for the host): culling must be
EXACT - the pruned program evaluates to bit-identical floats inside its region, dropped regions hold no surface
("cull on/off must give bit-identical fields")."""
import numpy as np
import pytest

import hostemu_binding as emu
import scenes_small
from parsip_b200 import blobtree as bt
import parsip_b200 as ps

ISO = 0.5


def _regions(lo, hi, rs, n, side):
    c = rs.uniform(lo, hi, (n, 3)).astype(np.float32)
    return c, (c + np.float32(side)).astype(np.float32)


@pytest.mark.parametrize("name", sorted(set(scenes_small.SCALAR_SCENES) | {"quadric", "instances", "xforms_srt"}))
@pytest.mark.parametrize("side", [0.14, 0.56])
def test_pruned_program_is_bit_identical_inside_its_region(name, side):
    root = scenes_small.COMPACTREF_SCENES.get(name, scenes_small.SCALAR_SCENES.get(name))()
    prog = ps.linearize(bt.flatten(root), ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
    quads = prog.quads()
    lo, hi = ps.prepare_bboxes(bt.flatten(bt.binarize(root)))
    rs = np.random.RandomState(5)
    blo, bhi = _regions(lo - 0.5, hi + 0.2, rs, 150, side)
    kept = dropped_support = dropped_interval = shorter = 0
    for a, b in zip(blo, bhi):
        pts = rs.uniform(a, b, (64, 3)).astype(np.float32)
        # the 8 corners too: the lattice evaluates exactly there
        corners = np.array([[x, y, z] for x in (a[0], b[0]) for y in (a[1], b[1]) for z in (a[2], b[2])], np.float32)
        pts = np.concatenate([pts, corners])
        full = emu.field(quads, pts, ps.SEM_SCALAR)
        sup = emu.prune(quads, a, b, ps.CULL_SUPPORT, ISO)
        if sup is None:
            dropped_support += 1
            assert (full == 0.0).all(), "support cull dropped a region with a non-zero field"
        else:
            kept += 1
            shorter += int(len(sup) < len(quads))
            f = emu.field(sup, pts, ps.SEM_SCALAR)
            assert np.array_equal(f.view(np.uint32), full.view(np.uint32)), "pruned program changed the field"
        itv = emu.prune(quads, a, b, ps.CULL_INTERVAL, ISO)
        if itv is None:
            dropped_interval += 1
            inside = full > ISO
            assert inside.all() or (~inside).all(), "interval cull dropped a region the surface crosses"
        else:
            assert sup is not None and np.array_equal(itv, sup)
    assert kept > 0 and dropped_interval >= dropped_support
    if name in ("blob", "csgwarp"):
        assert shorter > 0  # the pruning does shorten programs


def test_simd_semantics_programs_prune_exactly_too():
    root = bt.binarize(scenes_small.blob())
    prog = ps.linearize(bt.flatten(root), ps.SEM_SIMD, 0)
    quads = prog.quads()
    lo, hi = ps.prepare_bboxes(bt.flatten(root))
    rs = np.random.RandomState(9)
    blo, bhi = _regions(lo, hi, rs, 100, 0.2)
    for a, b in zip(blo, bhi):
        pts = rs.uniform(a, b, (48, 3)).astype(np.float32)
        full = emu.field(quads, pts, ps.SEM_SIMD)
        pr = emu.prune(quads, a, b, ps.CULL_SUPPORT, ISO)
        if pr is None:
            assert (full == 0.0).all()
        else:
            assert np.array_equal(emu.field(pr, pts, ps.SEM_SIMD).view(np.uint32), full.view(np.uint32))


def test_zero_folding_rules_of_difference():
    """difference min(L, 1 - 0) must become min(L, 1), not L (blend sums exceed 1)."""
    L = bt.Blend(bt.Point([0, 0, 0]), bt.Point([0.05, 0, 0]), bt.Point([0, 0.05, 0]))
    tree = bt.Dif(L, bt.Point([5, 5, 5]))  # subtrahend is identically 0 near the origin
    quads = ps.linearize(bt.flatten(tree), ps.SEM_SCALAR, ps.LIN_REASSOCIATE).quads()
    a, b = np.float32([-0.1, -0.1, -0.1]), np.float32([0.1, 0.1, 0.1])
    pts = np.random.RandomState(1).uniform(a, b, (200, 3)).astype(np.float32)
    full = emu.field(quads, pts, ps.SEM_SCALAR)
    assert full.max() == 1.0  # the clamp is active here
    pr = emu.prune(quads, a, b, ps.CULL_SUPPORT, ISO)
    assert pr is not None and (pr[-2, 0] & 31) == 19  # the subtrahend is replaced by OPC_CLAMP1
    assert np.array_equal(emu.field(pr, pts, ps.SEM_SCALAR), full)
    # all-inside region: the interval cull may drop it (no surface), the support cull must not
    assert emu.prune(quads, a, b, ps.CULL_INTERVAL, ISO) is None


def _has_colour_scale(quads):
    """walks the instruction headers: OPC_CSCALE quad or a non-zero colour-scale count in header bits 27..31"""
    base = {2: 2, 3: 2, 5: 2, 6: 2, 7: 2, 26: 2, 4: 3, 28: 4}
    pc = 0
    while pc < len(quads):
        h = int(quads[pc, 0])
        opc = h & 31
        if opc == 0:
            return False
        if opc == 21 or (h >> 27):
            return True
        pc += base.get(opc, 1) + (3 if (opc < 16 and (h >> 9) & 1) else 0)
    return False


def test_simd_colour_rule_survives_pruning():
    """The PS_SimdPoly blend colour 2*f_L*c_L + 2*f_R*c_R is unnormalised (PS_Polygonizer.cpp:1475-1487): dropping an
    all-zero operand must keep the survivor's weight (OPC_CSCALE markers), or mesh colours would depend on the culling."""
    root = bt.binarize(scenes_small.blob())
    prog = ps.linearize(bt.flatten(root), ps.SEM_SIMD, 0)
    quads, colors = prog.quads(), prog.colors()
    lo, hi = ps.prepare_bboxes(bt.flatten(root))
    rs = np.random.RandomState(21)
    blo, bhi = _regions(lo, hi, rs, 200, 0.15)
    checked = marked = 0
    for a, b in zip(blo, bhi):
        pts = rs.uniform(a, b, (32, 3)).astype(np.float32)
        full_f, full_c = emu.field_color(quads, colors, pts, ps.SEM_SIMD)
        pr = emu.prune(quads, a, b, ps.CULL_SUPPORT, ISO, color_exact=True)
        if pr is None:
            continue
        marked += int(_has_colour_scale(pr))
        f, c = emu.field_color(pr, colors, pts, ps.SEM_SIMD)
        assert np.array_equal(f.view(np.uint32), full_f.view(np.uint32))
        on = full_f > 0  # where the field is exactly 0 the reference's tie rule adds up colours of zero operands: never a vertex
        assert np.array_equal(c[on].view(np.uint32), full_c[on].view(np.uint32))
        checked += int(on.sum())
    assert checked > 1000 and marked > 0


def test_simd_colour_rule_survives_two_pruning_levels():
    """The device prunes hierarchically (region of 4x4x4 MPUs, then each MPU): markers left by the coarse level must survive
    when the fine level drops the instruction that carried them."""
    root = bt.binarize(scenes_small.blob())
    prog = ps.linearize(bt.flatten(root), ps.SEM_SIMD, 0)
    quads, colors = prog.quads(), prog.colors()
    lo, hi = ps.prepare_bboxes(bt.flatten(root))
    rs = np.random.RandomState(22)
    checked = 0
    for _ in range(60):
        a = rs.uniform(lo, hi - 0.6, 3).astype(np.float32)
        coarse = emu.prune(quads, a, a + np.float32(0.6), ps.CULL_SUPPORT, ISO, color_exact=True)
        if coarse is None:
            continue
        for _ in range(6):
            c = (a + rs.uniform(0, 0.45, 3)).astype(np.float32)
            d = (c + np.float32(0.15)).astype(np.float32)
            fine = emu.prune(coarse, c, d, ps.CULL_SUPPORT, ISO, color_exact=True)
            pts = rs.uniform(c, d, (24, 3)).astype(np.float32)
            full_f, full_c = emu.field_color(quads, colors, pts, ps.SEM_SIMD)
            if fine is None:
                assert (full_f == 0).all()
                continue
            f, col = emu.field_color(fine, colors, pts, ps.SEM_SIMD)
            on = full_f > 0
            assert np.array_equal(f.view(np.uint32), full_f.view(np.uint32))
            assert np.array_equal(col[on].view(np.uint32), full_c[on].view(np.uint32))
            checked += int(on.sum())
    assert checked > 1000


@pytest.mark.parametrize("sem", ["simd", "scalar"])
def test_slab_prune_with_tiled_bounds_is_exact_and_tighter(sem):
    """The one-off prune of a program against a rank's z-slab (context.cu: ensure_slab_program) covers the elongated slab box with a grid
    of balls and tests every primitive's exact support box (cull_walk.cuh: prim_bound_tiled, prim_outside_box) - one ball around a
    146 x 146 x 18-MPU slab reaches across the whole model and drops nothing. The slab program must give bit-identical fields at every
    point of the slab, and be much shorter than the single-ball prune."""
    from parsip_b200 import scenes
    root = scenes.blobs(1200, 12.0, 9)
    simd = sem == "simd"
    flat = bt.flatten(bt.binarize(root)) if simd else bt.flatten(root)
    quads = (ps.linearize(flat, ps.SEM_SIMD, 0) if simd else ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE)).quads()
    lo, hi = ps.prepare_bboxes(bt.flatten(bt.binarize(root)))
    rs = np.random.RandomState(4)
    semc = ps.SEM_SIMD if simd else ps.SEM_SCALAR
    for z0, z1 in ((lo[2], lo[2] + 1.5), (5.0, 6.2), (hi[2] - 2.0, hi[2])):
        a, b = np.array([lo[0], lo[1], z0], np.float32), np.array([hi[0], hi[1], z1], np.float32)
        ball = emu.prune(quads, a, b, ps.CULL_SUPPORT, ISO, 1 if simd else 0)
        tiled = emu.prune(quads, a, b, ps.CULL_SUPPORT, ISO, (1 if simd else 0) | 2)
        assert tiled is not None and len(tiled) < 0.6 * len(ball), (len(tiled), len(ball))
        pts = rs.uniform(a, b, (20000, 3)).astype(np.float32)
        pts[:4000, 2] = np.where(rs.rand(4000) < 0.5, a[2], b[2])  # the slab's own boundary planes
        assert np.array_equal(emu.field(quads, pts, semc).view(np.uint32), emu.field(tiled, pts, semc).view(np.uint32))
