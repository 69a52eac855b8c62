"""-m gpu: psgpu_field_eval_noalloc (include/psgpu.h) - the allocation-free variant of psgpu_field_eval, device staging in a static device
array of the library instead of growable context buffers. No counterpart in the reference.
Called through the C ABI and compared with psgpu_field_eval on the same points: the same kernel on the same program, so the bar is the one
the field tests use for two runs of one kernel - BIT-IDENTICAL - on two calls in a row (the second, smaller and different, also shows that
what the first call left in the shared array does not leak into the next). Named test_zz_* so that pytest collects it last.
Passes alone in a fresh process and after the other tests."""
import ctypes as C

import numpy as np
import pytest

import scenes_small
from parsip_b200 import api, blobtree as bt
import parsip_b200 as ps

pytestmark = pytest.mark.gpu


def bits(a):
    return np.ascontiguousarray(a, np.float32).view(np.uint32)


@pytest.fixture()
def own_ctx():
    c = ps.Context(0)  # a context of its own: nothing here depends on what other tests left in the shared one
    yield c
    c.close()


@pytest.mark.parametrize("sem", ["simd", "scalar"])
def test_noalloc_field_equals_field_eval_on_two_calls_in_a_row(own_ctx, sem):
    ctx = own_ctx
    if sem == "simd":
        root = bt.binarize(scenes_small.SIMD_SCENES["blob"]())
        ctx.upload(ps.linearize(bt.flatten(root), ps.SEM_SIMD, 0))
        lo, hi = ps.prepare_bboxes(bt.flatten(root))
    else:
        root = scenes_small.COMPACTREF_SCENES["csgwarp"]()
        ctx.upload(ps.linearize(bt.flatten(root), ps.SEM_SCALAR, ps.LIN_REASSOCIATE))
        lo, hi = ps.prepare_bboxes(bt.flatten(bt.binarize(root)))
    rs = np.random.RandomState(11)
    first = rs.uniform(lo, hi, (api.NOALLOC_MAX_POINTS, 3)).astype(np.float32)  # the maximum: fills the whole array
    second = rs.uniform(lo, hi, (12345, 3)).astype(np.float32)
    for pts in (first, second):
        want = ctx.field(pts)
        got = ctx.field_noalloc(pts)
        assert (want > 0).any(), "test points miss the model"
        assert np.array_equal(bits(got), bits(want)), "max abs diff %g" % np.abs(got - want).max()
    # and once more the first batch: the array now holds the second call's data in its head
    assert np.array_equal(bits(ctx.field_noalloc(first)), bits(ctx.field(first)))


def test_noalloc_rejects_oversize_and_bad_arguments(own_ctx):
    ctx = own_ctx
    root = bt.binarize(scenes_small.SIMD_SCENES["s1"]())
    pts = np.zeros((api.NOALLOC_MAX_POINTS + 1, 3), np.float32)
    out = np.zeros(len(pts), np.float32)
    lib = ctx.lib
    # no program uploaded yet
    assert lib.psgpu_field_eval_noalloc(ctx.h, pts.ctypes.data, 4, out.ctypes.data) == -1
    ctx.upload(ps.linearize(bt.flatten(root), ps.SEM_SIMD, 0))
    rc = lib.psgpu_field_eval_noalloc(ctx.h, pts.ctypes.data, len(pts), out.ctypes.data)
    assert rc == -2  # PSGPU_NOT_ENOUGH_MEM
    msg = lib.psgpu_last_error(ctx.h).decode()
    assert "PSGPU_NOALLOC_MAX_POINTS" in msg and str(api.NOALLOC_MAX_POINTS) in msg
    with pytest.raises(ps.PsGpuError):
        ctx.field_noalloc(pts)
    assert lib.psgpu_field_eval_noalloc(ctx.h, None, 4, out.ctypes.data) == -1
    assert lib.psgpu_field_eval_noalloc(ctx.h, pts.ctypes.data, 0, out.ctypes.data) == 1  # empty batch: nothing to do
    # still usable after the rejections
    p = np.random.RandomState(2).uniform(-1, 1, (1000, 3)).astype(np.float32)
    assert np.array_equal(bits(ctx.field_noalloc(p)), bits(ctx.field(p)))


def test_noalloc_call_allocates_no_device_memory(own_ctx):
    """cudaMemGetInfo before and after: the no-alloc call leaves the device's free memory where it was (the context buffers of
    psgpu_field_eval grow by the batch on their first use, which is what this entry point exists to avoid)."""
    ctx = own_ctx
    root = bt.binarize(scenes_small.SIMD_SCENES["blob"]())
    ctx.upload(ps.linearize(bt.flatten(root), ps.SEM_SIMD, 0))
    rt = None
    for name in ("libcudart.so.12", "libcudart.so"):
        try:
            rt = C.CDLL(name)
            break
        except OSError:
            continue
    if rt is None:
        pytest.skip("libcudart not loadable by name from Python")
    rt.cudaMemGetInfo.argtypes = [C.POINTER(C.c_size_t), C.POINTER(C.c_size_t)]

    def free_bytes():
        f, t = C.c_size_t(), C.c_size_t()
        assert rt.cudaMemGetInfo(C.byref(f), C.byref(t)) == 0
        return f.value

    pts = np.random.RandomState(3).uniform(-2, 2, (api.NOALLOC_MAX_POINTS, 3)).astype(np.float32)
    ctx.field_noalloc(pts[:16])  # first call on this device: the symbol look-up
    before = free_bytes()
    for _ in range(3):
        ctx.field_noalloc(pts)
    assert free_bytes() == before
