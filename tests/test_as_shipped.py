"""PSGPU_LIN_AS_SHIPPED (include/psgpu.h): the reference's prune of operators deeper than 3 levels, reproduced as OPC_BOXMASK
instructions. Host-compiled interpreters (tests/hostemu) against the golden vectors of the UNMODIFIED reference build with its OWN
operator boxes (tests/golden/*_asis.npz: `field` = 4 consecutive points per SIMD call, `field_bcast` = one point broadcast to all lanes),
which the default linearization - every contribution kept - does not match."""
import os

import numpy as np
import pytest

import hostemu_binding as emu
import scenes_small
from parsip_b200 import blobtree as bt
import parsip_b200 as ps

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
RSQRT_SCENES = ("discs_rings",)


def _bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


def shipped_program(name):
    b, soa, orc = scenes_small.simd_oracle(scenes_small.SIMD_SCENES[name](), pad_op_boxes=False)
    flat = scenes_small.with_op_boxes(bt.flatten(b), orc.op_boxes())
    return ps.linearize(flat, ps.SEM_SIMD, ps.LIN_AS_SHIPPED), ps.linearize(bt.flatten(b), ps.SEM_SIMD, 0), orc


@pytest.mark.parametrize("name", sorted(scenes_small.SIMD_SCENES))
def test_as_shipped_field_equals_the_unmodified_reference(name):
    g = np.load(os.path.join(GOLD, "%s_asis.npz" % name))
    prog, plain, orc = shipped_program(name)
    assert np.array_equal(orc.op_boxes(), g["op_boxes"])
    pts = g["pts"]
    one = emu.field(prog.quads(), pts, ps.SEM_SIMD)       # a point per evaluation
    four = emu.field_packed(prog.quads(), pts)            # 4 consecutive points = one SIMD packet
    if name in RSQRT_SCENES:  # 12-bit rsqrt in the reference's disc / ring distances (Q7): tolerance pin, and bits against the exact-rsqrt oracle
        assert np.abs(one - g["field_bcast"]).max() < 5e-4 and np.abs(four[:len(pts) // 4 * 4] - g["field"][:len(pts) // 4 * 4]).max() < 5e-4
        assert np.array_equal(_bits(one), _bits(orc.field_color_normal(pts)[0]))
        assert np.array_equal(_bits(four[:len(pts) // 4 * 4]), _bits(orc.field(pts)[:len(pts) // 4 * 4]))
        return
    assert np.array_equal(_bits(one), _bits(g["field_bcast"])), np.abs(one - g["field_bcast"]).max()
    n4 = len(pts) // 4 * 4
    assert np.array_equal(_bits(four[:n4]), _bits(g["field"][:n4])), np.abs(four[:n4] - g["field"][:n4]).max()
    if name == "blob":  # the prune does fire there: the default program (every contribution kept) is NOT the shipped reference
        keep = emu.field(plain.quads(), pts, ps.SEM_SIMD)
        assert prog.n_quads > plain.n_quads and (_bits(keep) != _bits(one)).any()


@pytest.mark.parametrize("name", ["blob", "csg"])
def test_as_shipped_program_survives_pruning(name):
    """The culling walk copies OPC_BOXMASK with the entry it masks and drops it with an all-zero one: pruned programs evaluate to the
    same bits inside their region, for single points and for packets."""
    prog, _, orc = shipped_program(name)
    quads = prog.quads()
    lo, hi = orc.bbox()
    rs = np.random.RandomState(8)
    kept = 0
    for _ in range(80):
        a = rs.uniform(lo, hi - 0.5, 3).astype(np.float32)
        b = (a + np.float32(0.5)).astype(np.float32)
        sub = emu.prune(quads, a, b, ps.CULL_INTERVAL, 0.5, True)
        pts = rs.uniform(a, b, (64, 3)).astype(np.float32)
        full = emu.field(quads, pts, ps.SEM_SIMD)
        if sub is None:
            assert (full < 0.5).all() or (full >= 0.5).all()
            continue
        kept += 1
        assert np.array_equal(_bits(emu.field(sub, pts, ps.SEM_SIMD)), _bits(full))
        assert np.array_equal(_bits(emu.field_packed(sub, pts)), _bits(emu.field_packed(quads, pts)))
    assert kept > 10


def test_as_shipped_needs_boxes_and_binary_trees():
    b = bt.binarize(scenes_small.SIMD_SCENES["s1"]())
    with pytest.raises(ps.PsGpuError):
        ps.linearize(bt.flatten(b), ps.SEM_SIMD, ps.LIN_AS_SHIPPED)  # no op_box_lo / op_box_hi
