"""CPU checks of the host linearizer + the device interpreter (compiled for the host by tests/hostemu):
the postfix quad stream evaluated by eval_program must reproduce the oracle's field."""
import numpy as np
import pytest

import hostemu_binding as emu
import scenes_small
from oracle.pyoracle import ScalarOracle
from parsip_b200 import blobtree as bt
import parsip_b200 as ps

REL = 1e-5  # north star: field values within 1e-5 relative (FP32); denominator floored at 1e-3


def rel_err(a, b):
    return np.abs(a - b) / np.maximum(np.abs(b), 1e-3)


def sample_points(lo, hi, n, seed=0):
    return np.random.RandomState(seed).uniform(lo, hi, (n, 3)).astype(np.float32)


@pytest.mark.parametrize("name", sorted(scenes_small.SIMD_SCENES))
def test_simd_semantics_program_matches_simd_oracle(name):
    root = scenes_small.SIMD_SCENES[name]()
    b, soa, orc = scenes_small.simd_oracle(root)
    lo, hi = orc.bbox()
    prog = ps.linearize(bt.flatten(b), ps.SEM_SIMD, 0)
    pts = sample_points(lo, hi, 20000)
    f = emu.field(prog.quads(), pts, ps.SEM_SIMD)
    fr = orc.field(pts)
    # PS_SimdPoly semantics run STRICT arithmetic (reference operation order, no FMA): bit-identical, not merely 1e-5
    assert np.array_equal(f.view(np.uint32), fr.view(np.uint32)), "max abs diff %g" % np.abs(f - fr).max()
    if any(n > 2 for n in bt.flatten(root)["n_children"]):
        # like SimdPoly::linearizeBlobTree (PS_HighPerformanceRender.cpp:81-86) the SIMD semantics insist on binary operators
        with pytest.raises(ps.PsGpuError) as e:
            ps.linearize(bt.flatten(root), ps.SEM_SIMD, 0)
        assert e.value.code == -3


@pytest.mark.parametrize("name", sorted(scenes_small.SCALAR_SCENES))
def test_scalar_semantics_program_matches_scalar_oracle(name):
    root = scenes_small.SCALAR_SCENES[name]()
    flat = bt.flatten(root)
    orc = ScalarOracle(flat)
    lo, hi = ps.prepare_bboxes(bt.flatten(bt.binarize(root)))
    for flags in (0, ps.LIN_REASSOCIATE):
        prog = ps.linearize(flat, ps.SEM_SCALAR, flags)
        pts = sample_points(lo - 0.3, hi + 0.3, 20000, seed=3)
        f = emu.field(prog.quads(), pts, ps.SEM_SCALAR)
        fr = orc.field(pts)
        ok = np.isfinite(fr)
        # native semantics use fused multiply-adds; the scalar oracle (like the reference's CPU build) does not.
        # Tolerance: 1e-5 relative + 5e-8 absolute - (1-d^2)^3 amplifies the 1-ulp rounding of d^2 by 3/(1-d^2)
        # relative as f -> 0, so a purely relative bound is ill-posed there; warps (sincos) and
        # Ricci (powf) add libm differences.
        tol = (3e-5 if name in ("ricci", "csgwarp") else REL) * np.abs(fr[ok]) + 5e-8
        assert (np.abs(f[ok] - fr[ok]) <= tol).all(), (name, np.abs(f[ok] - fr[ok]).max())


def test_program_structure_and_stack_depths():
    root = scenes_small.csgwarp_small()
    prog = ps.linearize(bt.flatten(root), ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
    q = prog.quads()
    assert q.shape == (prog.n_quads, 4)
    assert (q[-1, 0] & 31) == 0  # OPC_END terminates the stream
    assert prog.max_point_stack == 3  # bend(taper(twist(...)))
    assert 2 <= prog.max_value_stack <= 8
    assert prog.n_prims == sum(1 for t in bt.flatten(root)["type"] if t < 16)


def test_deep_right_leaning_chain_reassociates_to_constant_stack():
    # ConvertToBinaryTree produces op(c0, op(c1, ...)) chains: 500 deep here
    prims = [bt.Point([0.1 * i, 0.0, 0.0]) for i in range(500)]
    chain = bt.binarize(bt.Blend(*prims))
    flat = bt.flatten(chain)
    deep = ps.linearize(flat, ps.SEM_SCALAR, ps.LIN_REASSOCIATE)
    assert deep.max_value_stack == 1
    with pytest.raises(ps.PsGpuError) as e:
        ps.linearize(flat, ps.SEM_SCALAR, 0)  # literal order needs a 500-deep value stack
    assert e.value.code == -2
