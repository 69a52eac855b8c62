"""Host-side BlobTree description used by the Python tests and bench (a thin mirror of the
reference scene API's node kinds: Parsip100/PS_BlobTree/include/CBlobTree.h:29-394,
CSkeleton{Point,Line,Cylinder,Disc,Ring,Cube}.h, CBlend.h, CUnion.h, CIntersection.h,
CDifference.h, CSmoothDifference.h, CRicciBlend.h, CWarp{Twist,Taper,Bend,Shear}.h).

Nodes are plain records; `flatten()` produces the arrays of `psgpu_tree` (include/psgpu_types.h).
Nothing here computes a field.
"""
from __future__ import annotations

import numpy as np

# include/psgpu_types.h
PRIM_POINT, PRIM_LINE, PRIM_CYLINDER, PRIM_DISC, PRIM_RING, PRIM_CUBE, PRIM_NULL, PRIM_QUADRICPOINT, PRIM_INSTANCE = range(9)
OP_UNION, OP_INTERSECT, OP_DIF, OP_SMOOTHDIF, OP_BLEND, OP_RICCI = range(16, 22)
OP_WARP_TWIST, OP_WARP_TAPER, OP_WARP_BEND, OP_WARP_SHEAR = range(32, 36)

TYPE_NAMES = {
    PRIM_POINT: "point", PRIM_LINE: "line", PRIM_CYLINDER: "cylinder", PRIM_DISC: "disc",
    PRIM_RING: "ring", PRIM_CUBE: "cube", PRIM_NULL: "null", PRIM_QUADRICPOINT: "quadricpoint", PRIM_INSTANCE: "instance",
    OP_UNION: "union", OP_INTERSECT: "intersect", OP_DIF: "dif", OP_SMOOTHDIF: "smoothdif",
    OP_BLEND: "blend", OP_RICCI: "ricci",
    OP_WARP_TWIST: "twist", OP_WARP_TAPER: "taper", OP_WARP_BEND: "bend", OP_WARP_SHEAR: "shear",
}


def is_prim(t: int) -> bool:
    return t < 16


def is_warp(t: int) -> bool:
    return t >= 32


class Node:
    __slots__ = ("type", "children", "pos", "dir", "res", "color", "xform", "srt", "origin")

    def __init__(self, type, children=(), pos=(0, 0, 0), dir=(0, 0, 0), res=(0, 0, 0, 0),
                 color=(1, 1, 1), xform=None, srt=None, origin=None):
        self.type = int(type)
        self.children = list(children)
        self.pos = tuple(float(v) for v in pos)
        self.dir = tuple(float(v) for v in dir)
        r = tuple(float(v) for v in res)
        self.res = r + (0.0,) * (4 - len(r))
        self.color = tuple(float(v) for v in color)
        # backward (world -> local) 3x4 row-major, or None for identity
        self.xform = None if xform is None else np.asarray(xform, np.float32).reshape(3, 4)
        # the scene API's own parametrisation (CAffineTransformation: scale xyz, quaternion xyzw, translate xyz). A node built
        # with srt and no xform gets its backward matrix from srt_backward(); the reference-built oracle derives its own.
        self.srt = None if srt is None else np.asarray(srt, np.float32).reshape(10)
        if self.srt is not None and self.xform is None:
            self.xform = srt_backward(self.srt)
        self.origin = origin  # PRIM_INSTANCE: the instanced node (must be part of the same tree)

    @property
    def is_op(self):
        return not is_prim(self.type)

    def count(self):
        return 1 + sum(c.count() for c in self.children)


# -- constructors named after the reference's node classes ------------------------------------
def Point(pos, color=(1, 1, 1), xform=None):
    return Node(PRIM_POINT, pos=pos, color=color, xform=xform)


def Line(start, end, color=(1, 1, 1), xform=None):
    return Node(PRIM_LINE, pos=start, dir=end, color=color, xform=xform)


def Cylinder(pos, direction, radius, height, color=(1, 1, 1), xform=None):
    d = np.asarray(direction, np.float64)
    d = d / np.linalg.norm(d)
    return Node(PRIM_CYLINDER, pos=pos, dir=np.asarray(d, np.float32), res=(radius, height, 0),
                color=color, xform=xform)


def Disc(pos, normal, radius, color=(1, 1, 1), xform=None):
    d = np.asarray(normal, np.float64)
    d = d / np.linalg.norm(d)
    return Node(PRIM_DISC, pos=pos, dir=np.asarray(d, np.float32),
                res=(radius, np.float32(radius) * np.float32(radius), 0), color=color, xform=xform)


def Ring(pos, normal, radius, color=(1, 1, 1), xform=None):
    d = np.asarray(normal, np.float64)
    d = d / np.linalg.norm(d)
    return Node(PRIM_RING, pos=pos, dir=np.asarray(d, np.float32),
                res=(radius, np.float32(radius) * np.float32(radius), 0), color=color, xform=xform)


def Cube(pos, side, color=(1, 1, 1), xform=None):
    return Node(PRIM_CUBE, pos=pos, res=(side, 0, 0), color=color, xform=xform)


def Null():
    return Node(PRIM_NULL)


def QuadricPoint(pos, radius, scale, color=(1, 1, 1), xform=None, srt=None):
    """CQuadricPoint (PS_BlobTree/include/CQuadricPoint.h:43-52): f = scale * (1 - d^2/R^2)^2 inside R."""
    return Node(PRIM_QUADRICPOINT, pos=pos, res=(radius, scale, 0), color=color, xform=xform, srt=srt)


def Instance(origin, xform=None, srt=None):
    """CInstance (PS_BlobTree/include/CInstance.h): the origin's field at the instance's backward-transformed point."""
    return Node(PRIM_INSTANCE, origin=origin, xform=xform, srt=srt)


def with_srt(node, scale=(1, 1, 1), quat=(0, 0, 0, 1), translate=(0, 0, 0)):
    """Gives a node the scene API's transform (CAffineTransformation::set(scale, rotation, translate))."""
    node.srt = np.asarray(tuple(scale) + tuple(quat) + tuple(translate), np.float32)
    node.xform = srt_backward(node.srt)
    return node


def srt_backward(srt):
    """Backward 3x4 of CAffineTransformation::setReverseTransformations (PS_FrameWork/include/PS_AffineTransformation.h:160-181):
    forward = T(t) . R(q) . S(s) about the origin, backward = its inverse. Float64 here; the reference inverts in float32, so a
    tree evaluated against the reference-built oracle takes the matrix the product walker reads back instead."""
    s, q, t = np.asarray(srt[0:3], np.float64), np.asarray(srt[3:7], np.float64), np.asarray(srt[7:10], np.float64)
    x, y, z, w = q
    R = np.array([[1 - 2 * (y * y + z * z), 2 * (x * y - z * w), 2 * (x * z + y * w)],
                  [2 * (x * y + z * w), 1 - 2 * (x * x + z * z), 2 * (y * z - x * w)],
                  [2 * (x * z - y * w), 2 * (y * z + x * w), 1 - 2 * (x * x + y * y)]])
    F = np.eye(4)
    F[:3, :3] = R @ np.diag(s)
    F[:3, 3] = t
    return np.linalg.inv(F)[:3, :].astype(np.float32)


def Union(*kids):
    return Node(OP_UNION, kids)


def Intersect(*kids):
    return Node(OP_INTERSECT, kids)


def Dif(*kids):
    return Node(OP_DIF, kids)


def SmoothDif(*kids):
    return Node(OP_SMOOTHDIF, kids)


def Blend(*kids):
    return Node(OP_BLEND, kids)


def Ricci(n, *kids):
    return Node(OP_RICCI, kids, res=(n, 1.0 / n if n != 0 else 0.0, 0, 0))


def Twist(factor, axis, kid):
    return Node(OP_WARP_TWIST, [kid], res=(factor, axis, 0, 0))


def Taper(factor, axis_along, axis_taper, kid):
    return Node(OP_WARP_TAPER, [kid], res=(factor, axis_along, axis_taper, 0))


def Bend(rate, center, left, right, kid):
    return Node(OP_WARP_BEND, [kid], res=(rate, center, left, right))


def Shear(factor, axis_along, axis_dep, kid):
    return Node(OP_WARP_SHEAR, [kid], res=(factor, axis_along, axis_dep, 0))


# -- flat n-ary form (psgpu_tree) ---------------------------------------------------------------
def flatten(root: Node) -> dict:
    """Pre-order numbering, children ranges into one index array."""
    nodes = []
    stack = [root]
    order = {}
    # iterative pre-order (10k-deep chains must not hit the recursion limit)
    while stack:
        n = stack.pop()
        order[id(n)] = len(nodes)
        nodes.append(n)
        for c in reversed(n.children):
            stack.append(c)
    N = len(nodes)
    out = dict(
        n_nodes=N, root=0,
        type=np.zeros(N, np.int32), first_child=np.zeros(N, np.int32), n_children=np.zeros(N, np.int32),
        pos=np.zeros((N, 3), np.float32), dir=np.zeros((N, 3), np.float32), res=np.zeros((N, 4), np.float32),
        color=np.zeros((N, 3), np.float32), xform=np.zeros((N, 12), np.float32), has_xform=np.zeros(N, np.uint8),
        srt=np.zeros((N, 10), np.float32), has_srt=np.zeros(N, np.uint8),
    )
    child = []
    for i, n in enumerate(nodes):
        out["type"][i] = n.type
        out["first_child"][i] = len(child)
        out["n_children"][i] = len(n.children)
        child.extend(order[id(c)] for c in n.children)
        out["pos"][i] = n.pos
        out["dir"][i] = n.dir
        out["res"][i] = n.res
        out["color"][i] = n.color
        if n.xform is not None:
            out["xform"][i] = n.xform.reshape(12)
            out["has_xform"][i] = 1
        else:
            out["xform"][i] = np.eye(3, 4, dtype=np.float32).reshape(12)
        if n.srt is not None:
            out["srt"][i] = n.srt
            out["has_srt"][i] = 1
        if n.type == PRIM_INSTANCE:
            out["res"][i, 0] = order[id(n.origin)]
    out["child"] = np.asarray(child if child else [0], np.int32)
    return out


def binarize(root: Node) -> Node:
    """Right-leaning binary split of every operator with more than two children, the shape the
    reference's ConvertToBinaryTree produces (Parsip100/ParsipHaptics/include/CLayerManager.cpp:
    131-176: op(c0, op(c1, ... op(c_{n-2}, c_{n-1})))). Iterative (deep chains)."""
    clones = []

    def clone(n, kids):
        m = Node(n.type, kids, n.pos, n.dir, n.res, n.color, n.xform, origin=n.origin)
        m.srt = n.srt
        clones.append(m)
        return m

    # post-order rebuild without recursion
    done = {}
    stack = [(root, False)]
    while stack:
        n, visited = stack.pop()
        if not visited:
            stack.append((n, True))
            for c in n.children:
                stack.append((c, False))
            continue
        kids = [done[id(c)] for c in n.children]
        if n.is_op and len(kids) > 2:
            acc = clone(n, [kids[-2], kids[-1]])
            for c in reversed(kids[:-2]):
                acc = clone(n, [c, acc])
            done[id(n)] = acc
        else:
            done[id(n)] = clone(n, kids)
    for m in clones:  # instances point at the rebuilt copy of their origin
        if m.origin is not None:
            m.origin = done[id(m.origin)]
    return done[id(root)]


def to_binary_soa(root: Node) -> dict:
    """The SOA arrays SimdPoly::linearizeBlobTree fills (PS_HighPerformanceRender.cpp:42-364):
    pre-order ids per kind, left/right child ids, childKind = isOp(left)*2 + isOp(right),
    matrix slot 0 = identity. Input must already be binary; warps (unary) are padded with a NULL
    right child the way ConvertToBinaryTree pads (CLayerManager.cpp:196-213 only pads non-unary ops,
    so a unary warp reaching the SIMD linearizer is an error -3 there; the SIMD oracle has no
    warp semantics anyway)."""
    prims, ops = [], []
    mats = [np.eye(3, 4, dtype=np.float32)]
    rec = {}
    # explicit-stack pre-order assigning ids in visit order
    stack = [root]
    while stack:
        n = stack.pop()
        if n.is_op:
            if len(n.children) != 2:
                raise ValueError("to_binary_soa: operator %s has %d children" % (TYPE_NAMES[n.type], len(n.children)))
            rec[id(n)] = len(ops)
            ops.append(n)
            stack.append(n.children[1])
            stack.append(n.children[0])
        else:
            rec[id(n)] = len(prims)
            prims.append(n)
    P, O = len(prims), len(ops)
    d = dict(
        n_prims=P, n_ops=O,
        prim_type=np.array([p.type for p in prims], np.int32),
        prim_pos=np.array([p.pos for p in prims], np.float32).reshape(P, 3),
        prim_dir=np.array([p.dir for p in prims], np.float32).reshape(P, 3),
        prim_res=np.array([p.res[:3] for p in prims], np.float32).reshape(P, 3),
        prim_color=np.array([p.color for p in prims], np.float32).reshape(P, 3),
        prim_mat=np.zeros(P, np.int32),
        op_type=np.array([o.type for o in ops], np.int32).reshape(O),
        op_left=np.array([rec[id(o.children[0])] for o in ops], np.int32).reshape(O),
        op_right=np.array([rec[id(o.children[1])] for o in ops], np.int32).reshape(O),
        op_kind=np.array([int(o.children[0].is_op) * 2 + int(o.children[1].is_op) for o in ops], np.int32).reshape(O),
        op_res=np.array([o.res for o in ops], np.float32).reshape(O, 4),
    )
    fwd = [np.eye(4, dtype=np.float32)]
    for i, p in enumerate(prims):
        if p.xform is not None:
            d["prim_mat"][i] = len(mats)
            mats.append(p.xform)
            m4 = np.eye(4, dtype=np.float64)
            m4[:3, :] = p.xform
            # forward 4x4 in the memory order mat4Transform reads (e[col][row], translation in e[3][0..2], PS_MATRIX4.h:199-209)
            fwd.append(np.ascontiguousarray(np.linalg.inv(m4).T).astype(np.float32))
    d["n_mats"] = len(mats)
    d["mats_bwd"] = np.stack(mats).reshape(-1, 12).astype(np.float32)
    d["mats_fwd"] = np.stack(fwd).reshape(-1, 16).astype(np.float32)
    return d
