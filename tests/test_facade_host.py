"""CPU: the host-side entry points of the C++ facade (PrepareBBoxes, CountMPUNeeded of include/PS_SimdPolyCompat.h) against the
golden vectors of the reference's own build - no GPU needed for these."""
import os

import numpy as np
import pytest

import scenes_small
from parsip_b200 import blobtree as bt
import test_gpu_facade as tf

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


@pytest.mark.parametrize("name", sorted(scenes_small.SIMD_SCENES))
def test_facade_prepare_bboxes_and_count_mpus_equal_the_reference(name):
    lib = tf._lib()
    g, ga = np.load(os.path.join(GOLD, "%s_padded.npz" % name)), np.load(os.path.join(GOLD, "%s_asis.npz" % name))
    sc = tf.Scene(lib, bt.to_binary_soa(bt.binarize(scenes_small.SIMD_SCENES[name]())))
    try:
        assert lib.fac_prepare_bboxes(sc.h, 0.05) == 1
        lo, hi = np.zeros(3, np.float32), np.zeros(3, np.float32)
        lib.fac_get_bbox(sc.h, lo.ctypes.data, hi.ctypes.data)
        pb, ob = np.zeros((sc.n_prims, 6), np.float32), np.zeros((max(sc.n_ops, 1), 6), np.float32)
        lib.fac_get_boxes(sc.h, pb.ctypes.data, ob.ctypes.data)
        assert np.array_equal(lo, g["bbox_lo"]) and np.array_equal(hi, g["bbox_hi"])
        assert np.array_equal(pb, g["prim_boxes"]) and np.array_equal(ob[:sc.n_ops], ga["op_boxes"])
        assert lib.fac_count_mpus(sc.h, float(g["cellsize"])) == int(g["n_mpus"])
    finally:
        sc.close()
