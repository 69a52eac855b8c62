"""Generates tests/golden/*.npz from the reference's OWN polygonizer (oracle/_ref/libpsref.so = unmodified
Parsip100/PS_SimdPoly/include/PS_Polygonizer.cpp compiled by oracle/Makefile). Run in the build container, where
/root/reference exists; the fixtures are committed so the GPU box (no reference there) can check against them.

    python tools/gen_golden.py
"""
import hashlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

from oracle.pyoracle import RefScene  # noqa: E402
from parsip_b200 import blobtree as bt, scenes  # noqa: E402
import scenes_small  # noqa: E402

CELLS = {"s1": 48, "blob": 40, "csg": 40, "discs_rings": 36, "xforms": 36}


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def main():
    out = os.path.join(ROOT, "tests", "golden")
    os.makedirs(out, exist_ok=True)
    for name, mk in scenes_small.SIMD_SCENES.items():
        root = bt.binarize(mk())
        soa = bt.to_binary_soa(root)
        for padded in (False, True):
            ref = RefScene(soa)
            assert ref.build_info()["max_tree_nodes"] == 128, "golden vectors come from the UNMODIFIED build"
            ref.prepare_bboxes(0.05)
            lo, hi = ref.bbox()
            if padded:
                ref.pad_op_boxes(1.0e6)  # infinite lines reach beyond any finite box
            rs = np.random.RandomState(42)
            pts = rs.uniform(lo, hi, (3000, 3)).astype(np.float32)
            # field(): 4 consecutive points per SIMD call; field_color_normal(): one point broadcast to all lanes.
            # With the reference's own (unpadded) operator boxes the depth>3 prune decides per packet, so the two differ.
            fb, col, nrm = ref.field_color_normal(pts)
            f = ref.field(pts)
            assert padded is False or np.array_equal(f, fb)
            cs = scenes.cellsize_for(lo, hi, CELLS[name])
            ret, st = ref.polygonize(cs)
            assert ret == 1
            pos, mn, mc, tri = ref.mesh(st)
            np.savez_compressed(
                os.path.join(out, "%s_%s.npz" % (name, "padded" if padded else "asis")),
                bbox_lo=lo, bbox_hi=hi, prim_boxes=ref.prim_boxes(), op_boxes=ref.op_boxes(), pts=pts, field=f, field_bcast=fb, color=col, normal=nrm,
                cellsize=np.float32(cs), n_mpus=np.int64(st["mpus"]), n_crossed=np.int64(st["crossed"]),
                n_vertices=np.int64(st["vertices"]), n_triangles=np.int64(st["triangles"]), simd_evals=np.int64(st["simd_evals"]),
                pos_sha=sha(pos), tri_sha=sha(tri), col_sha=sha(mc), pos_head=pos[:256], tri_head=tri[:256],
                pos=pos.astype(np.float32) if len(pos) < 40000 else pos[:40000], tri=tri if len(tri) < 60000 else tri[:60000])
            print(name, "padded" if padded else "asis", st)


if __name__ == "__main__":
    main()
