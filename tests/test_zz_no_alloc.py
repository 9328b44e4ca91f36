"""GPU: hgm_map_device_static, the entry point for callers that may not allocate device memory
(a map slot inside an HBM budget owned by the TaskTracker or a pool allocator).  Its scratch is
a static device array shared by every call, so this checks it against the allocating entry
point (hgm_set_centroids + hgm_map_device) on two calls in a row with different jobs -- the
second call finds the first call's index and sums in the scratch.

Collected last on purpose (file name), so an -x run has already run everything else; it must
also pass alone in a fresh process: it depends on no state left by other tests."""
import numpy as np
import pytest

import cases
from conftest import check_partials
from hgpumap import capi

pytestmark = pytest.mark.gpu


def reference_result(slot, d_pts, n, cent, d_asg):
    """The existing, allocating path on the same device buffers."""
    slot.set_centroids(cent)
    slot.reset_partials()
    slot.map_device(d_pts, n, d_asg)
    sums, counts = slot.read_partials()
    assign = np.empty(n, np.int32)
    slot.memcpy_d2h(assign, d_asg)
    assert slot.last_kernel() == capi.KERNEL_GRID2D
    return sums, counts, assign


def test_static_scratch_entry_point_matches_allocating_path_twice(oracle):
    jobs = [cases.blobs(300001, 2, 128, seed=41), cases.blobs(123457, 2, 37, seed=42)]
    nmax = max(p.shape[0] for p, _ in jobs)
    with capi.GpuMapSlot(0) as slot:
        # the caller owns inputs and outputs; these buffers stand for the arena it was given
        d_pts = slot.device_alloc(nmax * 8)
        d_asg = slot.device_alloc(nmax * 4)
        d_asg_ref = slot.device_alloc(nmax * 4)
        try:
            for pts, cent in jobs:  # two calls in a row, second one over a dirty scratch
                n = pts.shape[0]
                slot.memcpy_h2d(d_pts, pts)
                ref_sums, ref_counts, ref_assign = reference_result(slot, d_pts, n, cent, d_asg_ref)
                slot.sync()
                sums, counts = capi.map_device_static(0, d_pts, n, cent, d_assign=d_asg)
                assign = np.empty(n, np.int32)
                slot.memcpy_d2h(assign, d_asg)
                np.testing.assert_array_equal(assign, ref_assign)
                np.testing.assert_array_equal(counts, ref_counts)
                o_sums, o_counts, o_abs, o_assign = oracle.map_combine(pts, cent)
                np.testing.assert_array_equal(assign, o_assign)
                check_partials(sums, counts, o_sums, o_counts, o_abs)
                check_partials(sums, counts, ref_sums, ref_counts, o_abs)
            # without the assignment output, and an empty split
            pts, cent = jobs[0]
            slot.memcpy_h2d(d_pts, pts)
            sums, counts = capi.map_device_static(0, d_pts, pts.shape[0], cent)
            o = oracle.map_combine(pts, cent)
            check_partials(sums, counts, o[0], o[1], o[2])
            sums, counts = capi.map_device_static(0, d_pts, 0, cent)
            assert counts.sum() == 0 and not sums.any()
        finally:
            slot.device_free(d_pts)
            slot.device_free(d_asg)
            slot.device_free(d_asg_ref)


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    cent3 = np.zeros((4, 3), np.float32)
    with pytest.raises(capi.HgmError) as err:
        capi.map_device_static(0, 0, 0, cent3)
    assert err.value.code == -22 and "d == 2" in str(err.value)
    big = np.zeros((1025, 2), np.float32)
    with pytest.raises(capi.HgmError) as err:
        capi.map_device_static(0, 0, 0, big)
    assert err.value.code == -22 and "1..1024" in str(err.value)
    bad = np.array([[0.0, 1.0], [np.nan, 2.0]], np.float32)
    with pytest.raises(capi.HgmError) as err:
        capi.map_device_static(0, 0, 0, bad)
    assert "cannot be indexed" in str(err.value)
