"""Small seeded scenes used by both the CPU and GPU tests (sized so the oracles finish in seconds)."""
import numpy as np

from parsip_b200 import blobtree as bt, scenes


def s1():
    """BASELINE configs[0]: 10 point/line primitives, union of two blends (runs on the UNMODIFIED reference)."""
    return scenes.small()


def blob(n=60, extent=3.5, seed=7, group=6):
    return scenes.blobs(n, extent, seed, group=group)


def csg_simd(seed=11):
    """point / cylinder / cube / disc / ring under every operator the SIMD evaluator implements; no lines (Q1), no Ricci (Q6)."""
    rs = np.random.RandomState(seed)
    P = lambda: rs.uniform(0.4, 2.6, 3).astype(np.float32)
    col = lambda: rs.uniform(0, 1, 3)
    a = bt.Blend(bt.Point(P(), col()), bt.Cylinder(P(), rs.normal(size=3), 0.2, 0.6, col()))
    b = bt.Union(bt.Cube(P(), 0.25, col()), bt.Point(P(), col()))
    c = bt.Intersect(bt.Blend(bt.Point([1.5, 1.5, 1.5], col()), bt.Point([1.8, 1.5, 1.5], col())), bt.Cube([1.6, 1.5, 1.5], 0.5, col()))
    d = bt.Dif(bt.Blend(bt.Point([0.8, 2.2, 0.9], col()), bt.Point([1.1, 2.2, 0.9], col())), bt.Point([0.95, 2.3, 1.0], col()))
    e = bt.SmoothDif(bt.Cylinder([2.2, 0.8, 2.0], [0, 1, 0], 0.25, 0.8, col()), bt.Point([2.2, 1.2, 2.1], col()))
    return bt.Union(bt.Union(a, b), bt.Union(c, bt.Union(d, e)))


def discs_rings(seed=13):
    rs = np.random.RandomState(seed)
    P = lambda: rs.uniform(0.5, 2.5, 3).astype(np.float32)
    col = lambda: rs.uniform(0, 1, 3)
    return bt.Union(bt.Blend(bt.Disc(P(), rs.normal(size=3), 0.3, col()), bt.Ring(P(), rs.normal(size=3), 0.45, col())),
                    bt.Blend(bt.Ring(P(), rs.normal(size=3), 0.35, col()), bt.Point(P(), col())))


def with_xforms(seed=17):
    """primitives carrying a backward 3x4 matrix (rotation + non-uniform scale + translation)."""
    rs = np.random.RandomState(seed)

    def M():
        q, _ = np.linalg.qr(rs.normal(size=(3, 3)))
        s = np.diag(rs.uniform(0.7, 1.4, 3))
        m = np.zeros((3, 4), np.float32)
        m[:, :3] = (q @ s).astype(np.float32)
        m[:, 3] = rs.uniform(-0.3, 0.3, 3)
        return m
    col = lambda: rs.uniform(0, 1, 3)
    P = lambda: rs.uniform(0.6, 2.0, 3).astype(np.float32)
    return bt.Union(bt.Blend(bt.Point(P(), col(), M()), bt.Cylinder(P(), rs.normal(size=3), 0.15, 0.5, col(), M())),
                    bt.Blend(bt.Cube(P(), 0.2, col(), M()), bt.Point(P(), col())))


def csgwarp_small():
    return scenes.csgwarp(n_groups=3, extent=5.0, seed=3)


def ricci():
    return bt.Ricci(2.5, bt.Point([1.0, 1.0, 1.0], (1, 0, 0)), bt.Point([1.5, 1.0, 1.0], (0, 1, 0)), bt.Line([1.0, 1.4, 1.0], [1.6, 1.5, 1.1], (0, 0, 1)))


def _quat(rs):
    q = rs.normal(size=4)
    return tuple(float(v) for v in (q / np.linalg.norm(q)).astype(np.float32))


def xforms_srt(seed=19):
    """primitives AND operators carrying the scene API's scale / rotation / translation (CAffineTransformation); the compact
    tree applies operator matrices too (CompactBlobTree.cpp:685-694). Only the reference-built oracle and the
    walker know the float32 backward matrices the reference derives from them."""
    rs = np.random.RandomState(seed)
    col = lambda: rs.uniform(0, 1, 3)
    P = lambda: rs.uniform(0.8, 2.2, 3).astype(np.float32)
    S = lambda: tuple(float(v) for v in rs.uniform(0.7, 1.3, 3).astype(np.float32))
    T = lambda: tuple(float(v) for v in rs.uniform(-0.3, 0.3, 3).astype(np.float32))
    a = bt.with_srt(bt.Point(P(), col()), S(), _quat(rs), T())
    b = bt.with_srt(bt.Cylinder(P(), rs.normal(size=3), 0.15, 0.5, col()), S(), _quat(rs), T())
    c = bt.with_srt(bt.Cube(P(), 0.2, col()), S(), _quat(rs), T())
    d = bt.with_srt(bt.Line(P(), P(), col()), (1, 1, 1), _quat(rs), T())
    e = bt.with_srt(bt.Disc(P(), rs.normal(size=3), 0.3, col()), S(), (0, 0, 0, 1), T())
    grp = bt.with_srt(bt.Blend(a, b, bt.Point(P(), col())), (1, 1, 1), _quat(rs), T())
    tw = bt.with_srt(bt.Twist(0.4, 1, bt.Blend(c, d)), S(), (0, 0, 0, 1), T())
    return bt.Union(grp, tw, e)


def quadric(seed=23):
    """CQuadricPoint primitives (Medusa's building block, PS_BlobTree/include/CQuadricPoint.h) blended with skeletal ones."""
    rs = np.random.RandomState(seed)
    col = lambda: rs.uniform(0, 1, 3)
    P = lambda: rs.uniform(0.8, 2.4, 3).astype(np.float32)
    qs = [bt.QuadricPoint(P(), float(np.float32(rs.uniform(0.5, 0.9))), float(np.float32(rs.uniform(0.6, 1.2))), col()) for _ in range(9)]
    return bt.Union(bt.Blend(*qs[:5]), bt.Blend(qs[5], qs[6], bt.Point(P(), col())), bt.Dif(bt.Blend(qs[7], qs[8]), bt.Point(P(), col())))


def instances(seed=29):
    """CInstance: a sub-tree (and a single primitive) re-used under the instance's own transform (CompactBlobTree.cpp:1069-1078)."""
    rs = np.random.RandomState(seed)
    col = lambda: rs.uniform(0, 1, 3)
    arm = bt.Blend(bt.Point([1.0, 1.0, 1.0], col()), bt.Line([1.0, 1.0, 1.0], [1.5, 1.2, 1.1], col()), bt.Cylinder([1.5, 1.2, 1.1], [0, 1, 0], 0.12, 0.4, col()))
    ball = bt.Point([2.2, 2.0, 1.0], col())
    i1 = bt.with_srt(bt.Instance(arm), (1, 1, 1), _quat(rs), (0.9, 0.4, 0.3))
    i2 = bt.with_srt(bt.Instance(arm), (0.8, 0.8, 0.8), (0, 0, 0, 1), (-0.2, 1.1, 0.6))
    i3 = bt.with_srt(bt.Instance(ball), (1, 1, 1), (0, 0, 0, 1), (0.0, -0.9, 0.5))
    return bt.Union(arm, ball, bt.Blend(i1, i3), i2)


SIMD_SCENES = {"s1": s1, "blob": blob, "csg": csg_simd, "discs_rings": discs_rings, "xforms": with_xforms}
SCALAR_SCENES = {"s1": s1, "blob": blob, "csg": csg_simd, "discs_rings": discs_rings, "xforms": with_xforms, "csgwarp": csgwarp_small, "ricci": ricci}
# scenes only the reference-built scalar twin can judge (tests/golden/scalar_*.npz carry the tree the walker read back)
COMPACTREF_SCENES = dict(SCALAR_SCENES, xforms_srt=xforms_srt, quadric=quadric, instances=instances)
del COMPACTREF_SCENES["xforms"]  # raw backward matrices have no scale / rotation / translation form


def simd_oracle(root, pad_op_boxes=True):
    """(binary tree, SOA arrays, SimdOracle with boxes prepared). pad_op_boxes neutralises the reference's depth>3
    operator-box prune (boxes padded by 0.504 although the Wyvill support is 1.0)."""
    from oracle.pyoracle import SimdOracle
    b = bt.binarize(root)
    soa = bt.to_binary_soa(b)
    o = SimdOracle(soa)
    o.prepare_bboxes(0.05)
    if pad_op_boxes:
        o.pad_op_boxes(1.0e6)  # infinite lines (Q1) reach beyond any finite box
    return b, soa, o


def with_op_boxes(flat, op_boxes):
    """flat tree + the operator boxes of the reference's PrepareBBoxes (oracle / RefScene .op_boxes(): one [lo xyz, hi xyz] row per
    operator in pre-order, the order blobtree.flatten numbers nodes in) as psgpu_tree.op_box_lo / op_box_hi for PSGPU_LIN_AS_SHIPPED."""
    import numpy as np
    is_op = flat["type"] >= 16
    assert int(is_op.sum()) == len(op_boxes)
    out = dict(flat)
    out["op_box_lo"] = np.zeros((flat["n_nodes"], 3), np.float32)
    out["op_box_hi"] = np.zeros((flat["n_nodes"], 3), np.float32)
    out["op_box_lo"][is_op] = np.asarray(op_boxes, np.float32)[:, :3]
    out["op_box_hi"][is_op] = np.asarray(op_boxes, np.float32)[:, 3:]
    return out
