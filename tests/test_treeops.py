"""CPU checks of the pre-linearization pass psgpu_tree_binarize: the restated ConvertToBinaryTree::run
(Parsip100/ParsipHaptics/include/CLayerManager.cpp:81-233) - chain shape, clone-order node ids, the "all children are
primitives" exemption, NULL padding - and that the binarized tree is what the PS_SimdPoly linearizer accepts and evaluates
bit for bit like the harness' own binarization."""
import numpy as np
import pytest

import hostemu_binding as emu
import scenes_small
from parsip_b200 import blobtree as bt, scenes
import parsip_b200 as ps


def nested(flat, n=None):
    """Order-preserving structural form of a flat tree: (type, params, children...)."""
    n = flat["root"] if n is None else n
    kids = flat["child"][flat["first_child"][n]:flat["first_child"][n] + flat["n_children"][n]]
    key = (int(flat["type"][n]), flat["pos"][n].tobytes(), flat["dir"][n].tobytes(), flat["res"][n].tobytes(), flat["color"][n].tobytes(),
           int(flat["has_xform"][n]), flat["xform"][n].tobytes() if flat["has_xform"][n] else b"")
    return (key,) + tuple(nested(flat, int(c)) for c in kids)


@pytest.mark.parametrize("name", sorted(scenes_small.SCALAR_SCENES))
def test_binarize_equals_the_harness_binarization(name):
    import sys
    sys.setrecursionlimit(20000)
    root = scenes_small.SCALAR_SCENES[name]()
    flat = bt.flatten(root)
    out, nconv = ps.binarize_tree(flat, split_always=True, pad_with_null=False)
    ref = bt.flatten(bt.binarize(root))
    assert nested(out) == nested(ref)
    assert out["n_nodes"] == ref["n_nodes"]
    assert nconv == int(sum(1 for n in range(flat["n_nodes"]) if flat["n_children"][n] > 2))
    ops = out["type"] >= 16
    assert (out["n_children"][ops & (out["type"] < 32)] == 2).all() and (out["n_children"][~ops] == 0).all()


def test_clone_order_ids_chain_shape_and_exemptions():
    P = lambda x: bt.Point((x, 0, 0))
    # union(blend(p0,p1,p2,p3), p4, blend(p5,p6))
    root = bt.Union(bt.Blend(P(0), P(1), P(2), P(3)), P(4), bt.Blend(P(5), P(6)))
    flat = bt.flatten(root)
    out, nconv = ps.binarize_tree(flat, True, True)
    U, B, PT = 16, 20, 0  # psgpu_node_type
    # clone order (CLayerManager.cpp:42-52, :140-176): root U0; its child 0 = blend B1 (split: p, link, p, link, p, p);
    # then the second union link, p4, and the last child blend with its two points
    assert out["type"].tolist() == [U, B, PT, B, PT, B, PT, PT, U, PT, B, PT, PT]
    kids = lambda n: out["child"][out["first_child"][n]:out["first_child"][n] + out["n_children"][n]].tolist()
    assert kids(0) == [1, 8] and kids(8) == [9, 10] and kids(10) == [11, 12]
    assert kids(1) == [2, 3] and kids(3) == [4, 5] and kids(5) == [6, 7]
    assert out["pos"][[2, 4, 6, 7, 9, 11, 12], 0].tolist() == [0, 1, 2, 3, 4, 5, 6]
    assert nconv == 2
    # split_always = 0: an operator whose children are all primitives stays n-ary (:136-139, :178-186); the union still splits
    out2, nconv2 = ps.binarize_tree(flat, False, True)
    assert nconv2 == 1 and sorted(out2["n_children"][out2["type"] == B].tolist()) == [2, 4]
    assert out2["n_children"][out2["type"] == U].tolist() == [2, 2]
    # a non-unary operator with one child gets a NULL primitive (:196-213); a warp does not
    lone = bt.flatten(bt.Union(bt.Twist(0.5, 2, bt.Blend(P(0)))))
    out3, nconv3 = ps.binarize_tree(lone, True, True)
    assert out3["type"].tolist() == [U, 32, B, PT, 6, 6] and nconv3 == 2
    out4, nconv4 = ps.binarize_tree(lone, True, False)
    assert out4["type"].tolist() == [U, 32, B, PT] and nconv4 == 0


def test_binarized_tree_feeds_the_simd_linearizer_bit_exactly():
    root = scenes_small.SIMD_SCENES["blob"]()
    flat = bt.flatten(root)
    with pytest.raises(ps.PsGpuError):
        ps.linearize(flat, ps.SEM_SIMD, 0)  # n-ary: PS_ERROR_NON_BINARY_OP
    out, _ = ps.binarize_tree(flat)
    q1 = ps.linearize(out, ps.SEM_SIMD, 0).quads()
    q2 = ps.linearize(bt.flatten(bt.binarize(root)), ps.SEM_SIMD, 0).quads()
    assert np.array_equal(q1, q2)
    lo, hi = ps.prepare_bboxes(out)
    pts = np.random.RandomState(3).uniform(lo, hi, (5000, 3)).astype(np.float32)
    assert np.array_equal(emu.field(q1, pts, ps.SEM_SIMD).view(np.uint32), emu.field(q2, pts, ps.SEM_SIMD).view(np.uint32))


def test_deep_chain_and_bad_input():
    root = scenes.blob1k()
    flat = bt.flatten(root)
    out, nconv = ps.binarize_tree(flat)
    assert nconv == 51 and out["n_nodes"] == 1000 + 999  # a binary tree over 1000 leaves
    again, n2 = ps.binarize_tree(out)
    assert n2 == 0 and np.array_equal(again["type"], out["type"]) and np.array_equal(again["child"], out["child"])
    bad = dict(flat)
    bad["child"] = flat["child"].copy()
    bad["child"][0] = flat["n_nodes"] + 5
    with pytest.raises(ps.PsGpuError) as e:
        ps.binarize_tree(bad)
    assert e.value.code == -1
