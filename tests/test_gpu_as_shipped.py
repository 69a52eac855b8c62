"""-m gpu: PSGPU_LIN_AS_SHIPPED through the C ABI and the facade against the goldens of the UNMODIFIED reference build run with its OWN
operator boxes (tests/golden/*_asis.npz): field bits, per-MPU counts, and the hashes of its concatenated PolyMPUs positions and indices.
Colours are not compared: where its prune fires the shipped reference reads uninitialised stack arrays (PS_Polygonizer.cpp:1384-1385) and
its own colour hash changes from run to run."""
import os

import numpy as np
import pytest

import gpu_helpers as gh
import scenes_small
from parsip_b200 import blobtree as bt
import parsip_b200 as ps
from test_gpu_parity import GOLD, RSQRT_SCENES, bits, compat_params

pytestmark = pytest.mark.gpu
NAMES = sorted(scenes_small.SIMD_SCENES)


def same_floats(a, b):
    """bit-identical, except that any NaN equals any NaN: where the prune zeroes all four root samples of an edge the reference divides
    0.5 by 0 and multiplies the infinity by a zero step - x86 SSE produces the NaN 0xFFC00000 there, the GPU 0x7FFFFFFF"""
    a, b = np.ascontiguousarray(a, np.float32), np.ascontiguousarray(b, np.float32)
    return a.shape == b.shape and bool(((bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))).all())


def shipped_setup(ctx, name):
    b, soa, orc = scenes_small.simd_oracle(scenes_small.SIMD_SCENES[name](), pad_op_boxes=False)
    ctx.upload(ps.linearize(scenes_small.with_op_boxes(bt.flatten(b), orc.op_boxes()), ps.SEM_SIMD, ps.LIN_AS_SHIPPED))
    return soa, orc


@pytest.mark.parametrize("name", NAMES)
def test_as_shipped_field_bits(ctx, name):
    g = np.load(os.path.join(GOLD, "%s_asis.npz" % name))
    _, orc = shipped_setup(ctx, name)
    f = ctx.field(g["pts"])  # one point per evaluation = the reference's broadcast call
    if name in RSQRT_SCENES:
        assert np.abs(f - g["field_bcast"]).max() < 5e-4
        assert np.array_equal(bits(f), bits(orc.field_color_normal(g["pts"])[0]))
        return
    assert np.array_equal(bits(f), bits(g["field_bcast"])), np.abs(f - g["field_bcast"]).max()


@pytest.mark.parametrize("name", NAMES)
def test_as_shipped_mesh_equals_the_unmodified_reference(ctx, name):
    g = np.load(os.path.join(GOLD, "%s_asis.npz" % name))
    _, orc = shipped_setup(ctx, name)
    lo, hi = orc.bbox()
    cs = float(g["cellsize"])
    for cull in (ps.CULL_INTERVAL, ps.CULL_NONE):
        m = ctx.polygonize(compat_params(cs, lo, hi, cull=cull))
        pos, nrm, col, tri, ijk, counts = gh.lattice_ordered(m)
        if name in RSQRT_SCENES:
            assert abs(m["n_vertices"] - int(g["n_vertices"])) <= 0.01 * int(g["n_vertices"])
        else:
            assert m["n_mpus_out"] == int(g["n_crossed"])
            assert m["n_vertices"] == int(g["n_vertices"]) and m["n_triangles"] == int(g["n_triangles"])
            assert gh.sha(tri) == str(g["tri_sha"]), "triangle indices differ from the shipped reference"
            n = len(g["pos"])  # the golden keeps the first 40 000 vertices and the hash of all of them
            assert same_floats(pos[:n], g["pos"]), "vertex positions differ from the shipped reference"
            if not np.isnan(pos).any():
                assert gh.sha(pos) == str(g["pos_sha"])
        ret, ref = orc.polygonize(cs)  # the restatement with the same boxes (pinned to the same goldens and hashes on the CPU)
        assert ret == 1 and np.array_equal(ijk, ref["mpu_idx"]) and np.array_equal(counts, ref["mpu_counts"].astype(np.int64))
        assert np.array_equal(tri, ref["tri"]) and same_floats(pos, ref["pos"])


def test_as_shipped_through_the_facade(ctx, monkeypatch):
    """PSGPU_FACADE_AS_SHIPPED=1: PrepareBBoxes + Polygonize of include/PS_SimdPolyCompat.h with the caller's own operator boxes."""
    import test_gpu_facade as tf
    lib = tf._lib()
    monkeypatch.setenv("PSGPU_FACADE_AS_SHIPPED", "1")
    for name in ("blob", "csg", "s1"):
        g = np.load(os.path.join(GOLD, "%s_asis.npz" % name))
        soa = bt.to_binary_soa(bt.binarize(scenes_small.SIMD_SCENES[name]()))
        sc = tf.Scene(lib, soa)
        try:
            assert lib.fac_prepare_bboxes(sc.h, 0.05) == 1
            st = np.zeros(4, np.int64)
            assert lib.fac_polygonize(sc.h, float(g["cellsize"]), st.ctypes.data) == 1
            nv, nt = int(st[2]), int(st[3])
            assert (nv, nt) == (int(g["n_vertices"]), int(g["n_triangles"]))
            pos, nrm, col = (np.zeros((nv, 3), np.float32) for _ in range(3))
            tri = np.zeros((nt, 3), np.uint32)
            mlo = np.zeros((int(st[0]), 3), np.float32)
            lib.fac_mesh_concat(sc.h, pos.ctypes.data, nrm.ctypes.data, col.ctypes.data, tri.ctypes.data, mlo.ctypes.data)
            assert gh.sha(tri) == str(g["tri_sha"]) and same_floats(pos[:len(g["pos"])], g["pos"])
        finally:
            sc.close()
