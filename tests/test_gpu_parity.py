"""GPU: the CUDA path, called through the C ABI, against the oracle on the same inputs.
Bar: assignments and counts bit-exact; sums within
1e-5 * max(|s_ref|, 1e-5 * sum|x|)."""
import os

import numpy as np
import pytest

import cases
from conftest import GOLDEN, check_partials
from hgpumap import capi, synth

pytestmark = pytest.mark.gpu

KERNELS_2D = [capi.KERNEL_GRID2D, capi.KERNEL_GRID2D_SORT, capi.KERNEL_EXACT]


@pytest.fixture(scope="module")
def slot():
    s = capi.GpuMapSlot(0)
    yield s
    s.close()


@pytest.fixture
def tensor_form(slot):
    """Force one form of the tcgen05 score kernel for a test; AUTO again afterwards."""
    def force(name):
        slot.set_tensor_form({"single": capi.TENSOR_FORM_SINGLE, "pair": capi.TENSOR_FORM_PAIR,
                              "pair_fused": capi.TENSOR_FORM_PAIR_FUSED,
                              "pair_segmented": capi.TENSOR_FORM_PAIR_SEGMENTED}[name])
    yield force
    slot.set_tensor_form(capi.TENSOR_FORM_AUTO)


def run_case(slot, oracle, pts, cent, kernel, fmt=capi.FMT_F32LE, want_kernel=None):
    slot.set_kernel(kernel)
    slot.set_centroids(cent)
    data = pts.astype("<f4") if fmt == capi.FMT_F32LE else pts.astype(">f4")
    sums, counts, assign = slot.map_split(data, fmt=fmt, want_assign=True)
    assert slot.last_points == pts.shape[0]
    if pts.shape[0]:
        assert slot.last_kernel() == (want_kernel or kernel)
    r_sums, r_counts, r_abs, r_assign = oracle.map_combine(pts, cent)
    np.testing.assert_array_equal(assign, r_assign)
    check_partials(sums, counts, r_sums, r_counts, r_abs)
    # without the assignment output the kernel takes another template path
    sums2, counts2, _ = slot.map_split(data, fmt=fmt, want_assign=False)
    check_partials(sums2, counts2, r_sums, r_counts, r_abs)
    slot.set_kernel(capi.KERNEL_AUTO)


@pytest.mark.parametrize("fname", ["kmeans_blobs_d2_k128.npz", "kmeans_adversarial_2d.npz"])
@pytest.mark.parametrize("kernel", KERNELS_2D)
def test_golden_2d(slot, fname, kernel):
    """Against the committed golden vectors (made by the reference runtime / C restatement)."""
    g = np.load(os.path.join(GOLDEN, fname))
    slot.set_kernel(kernel)
    slot.set_centroids(g["centroids"])
    sums, counts, assign = slot.map_split(g["points"], want_assign=True)
    np.testing.assert_array_equal(assign, g["assign"].astype(np.int32))
    check_partials(sums, counts, g["sums"], g["counts"], g["abs_sums"])
    slot.set_kernel(capi.KERNEL_AUTO)


def test_golden_text_case_as_binary_rows(slot):
    g = np.load(os.path.join(GOLDEN, "kmeans_text_d5_k7.npz"))
    slot.set_centroids(g["centroids"])
    sums, counts, assign = slot.map_split(g["points"], want_assign=True)
    np.testing.assert_array_equal(assign, g["assign"].astype(np.int32))
    check_partials(sums, counts, g["sums"], g["counts"], g["abs_sums"])
    assert slot.last_kernel() == capi.KERNEL_EXACT


@pytest.mark.parametrize("kernel", KERNELS_2D)
def test_config1_one_million_points(slot, oracle, kernel):
    """BASELINE config 1 shape: 2^20 2-D points, k = 128."""
    pts, cent = cases.blobs(1 << 20, 2, 128)
    run_case(slot, oracle, pts, cent, kernel)


@pytest.mark.parametrize("n", [0, 1, 2, 3, 511, 512, 513, 8191, 8192, 8193, 16385, 100001])
@pytest.mark.parametrize("kernel", KERNELS_2D)
def test_ragged_sizes(slot, oracle, n, kernel):
    pts, cent = cases.blobs(n, 2, 128, seed=n + 1)
    run_case(slot, oracle, pts.reshape(n, 2), cent, kernel)


@pytest.mark.parametrize("k", [1, 2, 7, 31, 33, 100, 128, 129, 200, 254, 255, 257, 1000, 1024])
@pytest.mark.parametrize("kernel", KERNELS_2D)
def test_cluster_counts(slot, oracle, k, kernel):
    pts, cent = cases.blobs(30011, 2, k, seed=k)
    run_case(slot, oracle, pts, cent, kernel)


def test_combiner_form_follows_the_shape(slot):
    """Fixed-point combiner for k <= 254 with ordinary magnitudes; counting sort when forced, for larger k and
    for coordinates whose scale a float cannot carry."""
    pts, cent = cases.blobs(1000, 2, 128)
    slot.set_centroids(cent)
    assert slot.grid2d_form()[0] == capi.GRID2D_FORM_FIXED_POINT
    slot.set_kernel(capi.KERNEL_GRID2D_SORT)  # rebuilds the installed index for the other form
    assert slot.grid2d_form()[0] == capi.GRID2D_FORM_SORT
    slot.map_split(pts)
    assert slot.last_kernel() == capi.KERNEL_GRID2D_SORT
    slot.set_kernel(capi.KERNEL_AUTO)
    assert slot.grid2d_form()[0] == capi.GRID2D_FORM_FIXED_POINT
    slot.set_centroids(cases.blobs(10, 2, 300)[1])
    assert slot.grid2d_form()[0] == capi.GRID2D_FORM_SORT
    slot.set_centroids((cent.astype(np.float64) * 1e-36).astype(np.float32))
    assert slot.grid2d_form()[0] == capi.GRID2D_FORM_SORT
    slot.set_centroids(cases.general_d(10, 3, 4, 1)[1])
    assert slot.grid2d_form()[0] == capi.GRID2D_FORM_NONE


@pytest.mark.parametrize("share,k", [(1.0, 1), (1.0, 128), (0.9, 128), (0.6, 200)])
@pytest.mark.parametrize("kernel", [capi.KERNEL_GRID2D, capi.KERNEL_GRID2D_SORT])
def test_one_cluster_holds_most_points(slot, oracle, share, k, kernel):
    """Skew: 6 M points, most of them in one cluster -- tens of thousands of adds per (cluster, lane) word of the
    fixed-point combiner (it must flush long before a word can wrap), thousands per tile for the counting sort."""
    pts, cent = cases.skewed_2d(6_000_000, k=k, share=share)
    run_case(slot, oracle, pts, cent, kernel)


@pytest.mark.parametrize("maker", [cases.adversarial_2d, cases.tiny_scale_2d, cases.close_centroids_2d,
                                   cases.degenerate_axis_2d, cases.tiny_values_2d])
@pytest.mark.parametrize("kernel", KERNELS_2D)
def test_edge_cases_2d(slot, oracle, maker, kernel):
    pts, cent = maker()
    run_case(slot, oracle, pts, cent, kernel)


def test_uniform_points_hit_many_edge_cells(slot, oracle):
    """Uniform points: ~13 % land in multi-candidate cells, some outside the grid."""
    rng = np.random.default_rng(5)
    cent = rng.uniform(-100, 100, size=(128, 2)).astype(np.float32)
    pts = rng.uniform(-140, 140, size=(300000, 2)).astype(np.float32)
    run_case(slot, oracle, pts, cent, capi.KERNEL_GRID2D)


def test_unindexable_centroids_fall_to_exact_scan(slot, oracle):
    """Non-finite or astronomically large centroids cannot be indexed; AUTO must pick the
    scan, and forcing the grid kernel must be refused (no silent substitution)."""
    pts, cent = cases.blobs(5000, 2, 16, seed=3)
    cent = cent.copy()
    cent[4, 1] = np.nan
    cent[7, 0] = np.inf
    run_case(slot, oracle, pts, cent, capi.KERNEL_AUTO, want_kernel=capi.KERNEL_EXACT)
    slot.set_kernel(capi.KERNEL_GRID2D)
    slot.set_centroids(cent)
    with pytest.raises(capi.HgmError):
        slot.map_split(pts)
    slot.set_kernel(capi.KERNEL_AUTO)


@pytest.mark.parametrize("d,k,n", [(1, 5, 4001), (3, 17, 5003), (4, 64, 9001), (8, 33, 7001), (9, 12, 3001),
                                   (16, 50, 6007), (64, 100, 4099), (128, 257, 2053), (300, 9, 777),
                                   (32, 8200, 1501)])
def test_general_dimensions_exact_scan(slot, oracle, d, k, n):
    pts, cent = cases.general_d(n, d, k, seed=d * 1000 + k)
    tensor = d % 4 == 0 and 8 <= d <= 128 and k >= 8 and k * d >= (256 if d > 8 else 2048)  # hgm_api.cu:pick_kernel
    run_case(slot, oracle, pts, cent, capi.KERNEL_AUTO, want_kernel=capi.KERNEL_TENSOR if tensor else capi.KERNEL_EXACT)
    if tensor:
        run_case(slot, oracle, pts, cent, capi.KERNEL_EXACT)


def test_big_endian_rows_are_decoded_on_device(slot, oracle):
    pts, cent = cases.blobs(50001, 2, 128, seed=8)
    run_case(slot, oracle, pts, cent, capi.KERNEL_AUTO, fmt=capi.FMT_F32BE, want_kernel=capi.KERNEL_GRID2D)
    pts, cent = cases.general_d(3000, 5, 7, seed=2)
    run_case(slot, oracle, pts, cent, capi.KERNEL_AUTO, fmt=capi.FMT_F32BE, want_kernel=capi.KERNEL_EXACT)


def test_malformed_split_is_rejected(slot):
    slot.set_centroids(np.zeros((4, 2), np.float32))
    with pytest.raises(capi.HgmError) as err:
        slot.map_split(b"\x00" * 12)  # 1.5 rows
    assert err.value.code == -1002
    with pytest.raises(capi.HgmError):
        slot.map_split(b"\x00" * 16, fmt=99)


def test_map_file_and_offsets(slot, oracle, tmp_path):
    pts, cent = cases.blobs(70000, 2, 128, seed=21)
    path = str(tmp_path / "split.bin")
    with open(path, "wb") as f:
        f.write(b"\xff" * 24)  # foreign bytes before the split
        f.write(pts.astype("<f4").tobytes())
    slot.set_centroids(cent)
    sums, counts, assign = slot.map_file(path, offset=24, want_assign=True)
    r_sums, r_counts, r_abs, r_assign = oracle.map_combine(pts, cent)
    np.testing.assert_array_equal(assign, r_assign)
    check_partials(sums, counts, r_sums, r_counts, r_abs)
    # FileSplit-style sub-range (mapred/FileSplit.java:81-85): rows 1000..50999
    sums, counts, _ = slot.map_file(path, offset=24 + 1000 * 8, length=50000 * 8)
    r = oracle.map_combine(pts[1000:51000], cent)
    check_partials(sums, counts, r[0], r[1], r[2])
    with pytest.raises(capi.HgmError):
        slot.map_file(str(tmp_path / "missing.bin"))


def test_multi_chunk_split_streams_through_the_pipeline(slot, oracle):
    """A split larger than one 64 MiB staging chunk (d=2: 8,388,608 rows), pageable host memory."""
    n = 2 * 8388608 + 12345
    centres = synth.centres(synth.SEED_POINTS, 128, 2)
    cent = synth.perturb(synth.SEED_CENTROIDS, centres)
    pts = capi.synth_points_host(synth.SEED_POINTS, 0, n, centres)
    slot.set_centroids(cent)
    sums, counts, assign = slot.map_split(pts, want_assign=True)
    assert counts.sum() == n
    # oracle on three windows (start, chunk boundary, end) + global linearity of the sums
    for lo in (0, 8388608 - 5000, n - 10000):
        np.testing.assert_array_equal(assign[lo:lo + 10000], oracle.assign(pts[lo:lo + 10000], cent))
    np.testing.assert_array_equal(np.bincount(assign, minlength=128), counts)
    np.testing.assert_allclose(sums.sum(axis=0), pts.astype(np.float64).sum(axis=0), rtol=1e-9)
    ref = np.zeros((128, 2))
    np.add.at(ref, assign, pts.astype(np.float64))
    np.testing.assert_allclose(sums, ref, rtol=1e-9, atol=1e-6)


def test_device_resident_path_accumulates_across_calls(slot, oracle):
    """hgm_map_device ADDS into the partials: two halves of a split == the whole split."""
    pts, cent = cases.blobs(200001, 2, 128, seed=77)
    slot.set_centroids(cent)
    nbytes = pts.nbytes
    d_pts = slot.device_alloc(nbytes)
    d_asg = slot.device_alloc(pts.shape[0] * 4)
    try:
        slot.memcpy_h2d(d_pts, pts)
        slot.reset_partials()
        half = 100000  # even -> second half stays 16-byte aligned
        slot.map_device(d_pts, half, d_asg)
        slot.map_device(d_pts + half * 8, pts.shape[0] - half, d_asg + half * 4)
        sums, counts = slot.read_partials()
        assign = np.empty(pts.shape[0], np.int32)
        slot.memcpy_d2h(assign, d_asg)
        r_sums, r_counts, r_abs, r_assign = oracle.map_combine(pts, cent)
        np.testing.assert_array_equal(assign, r_assign)
        check_partials(sums, counts, r_sums, r_counts, r_abs)
        # misaligned start (odd point index -> 8-byte alignment only) takes the float2 path
        slot.reset_partials()
        slot.map_device(d_pts + 8, pts.shape[0] - 1, d_asg + 4)
        sums, counts = slot.read_partials()
        slot.memcpy_d2h(assign, d_asg)
        r = oracle.map_combine(pts[1:], cent)
        np.testing.assert_array_equal(assign[1:], r[3])
        check_partials(sums, counts, r[0], r[1], r[2])
        ms, launches = slot.last_map_ms()
        assert ms > 0 and launches == 1
    finally:
        slot.device_free(d_pts)
        slot.device_free(d_asg)


def test_device_synth_matches_host_bits(slot):
    centres = synth.centres(1, 128, 2)
    n = 100003
    d_out = slot.device_alloc(n * 8)
    try:
        slot.synth_points_device(9, 5, n, centres, d_out)
        got = np.empty((n, 2), np.float32)
        slot.memcpy_d2h(got, d_out)
        assert np.array_equal(got.view(np.uint32), synth.points(9, 5, n, centres).view(np.uint32))
    finally:
        slot.device_free(d_out)


def test_idempotent_and_centroid_update_roundtrip(slot, oracle):
    """Property: mapping twice gives identical counts; one Lloyd step from the partials
    matches the oracle's step (the reducer's centroid update)."""
    pts, cent = cases.blobs(150000, 2, 64, seed=31)
    slot.set_centroids(cent)
    s1, c1, a1 = slot.map_split(pts, want_assign=True)
    s2, c2, a2 = slot.map_split(pts, want_assign=True)
    np.testing.assert_array_equal(c1, c2)
    np.testing.assert_array_equal(a1, a2)
    new = (s1 / np.maximum(c1, 1)[:, None]).astype(np.float32)
    r = oracle.map_combine(pts, cent)
    ref_new = (r[0] / np.maximum(r[1], 1)[:, None]).astype(np.float32)
    np.testing.assert_allclose(new, ref_new, rtol=1e-5, atol=1e-6)
    slot.set_centroids(new)
    s3, c3, a3 = slot.map_split(pts, want_assign=True)
    np.testing.assert_array_equal(a3, oracle.assign(pts, new))


# ---- the tcgen05 path (K2 scores + K3 exact re-check + K4 accumulate) ---------------------------

@pytest.mark.parametrize("d,k,n", [(128, 4096, 20011), (64, 1024, 200003),   # BASELINE configs 5 and 4 at their real d and k
                                   (32, 128, 5000), (64, 1024, 20011), (128, 512, 6001), (96, 300, 4097),
                                   (64, 129, 255), (64, 256, 256), (64, 256, 1), (64, 100, 3001), (32, 9, 2000),
                                   (128, 8, 1000), (16, 200, 3000), (48, 64, 3001), (100, 300, 2500),
                                   (8, 1024, 4000), (20, 128, 2999), (124, 40, 1000), (36, 70, 777)])
@pytest.mark.parametrize("form", ["single", "pair", "pair_fused", "pair_segmented"])
def test_tensor_path_matches_oracle(slot, oracle, tensor_form, form, d, k, n):
    """BASELINE config 4/5 shapes at test size: assignments bit-exact through the TF32 GEMM +
    FP32 re-check, counts exact, sums within tolerance -- in both forms of K2 (one CTA per tile,
    and the cta_group::2 pair; the library picks by shape, hgm_set_tensor_form forces one) and, for the
    pair, with the per-cluster sums made inside K2 and by K4's second pass ("pair" picks by k)."""
    tensor_form(form)
    pts, cent = cases.general_d(n, d, k, seed=d + k)
    run_case(slot, oracle, pts, cent, capi.KERNEL_TENSOR)


@pytest.mark.parametrize("form", ["single", "pair"])
def test_tensor_path_decides_almost_everything_on_the_tensor_cores(slot, oracle, tensor_form, form):
    """If the GEMM scores were wrong every point would be flagged and the exact re-check would
    hide it: on well-separated data the re-check must stay rare, on duplicated centroids it
    must take (at least) the points of the duplicated clusters."""
    tensor_form(form)
    pts, cent = cases.general_d(50000, 64, 1024, seed=77)
    slot.set_kernel(capi.KERNEL_TENSOR)
    slot.set_centroids(cent)
    sums, counts, assign = slot.map_split(pts, want_assign=True)
    rechecks = slot.last_rechecks()
    np.testing.assert_array_equal(assign, oracle.assign(pts, cent))
    assert rechecks < 0.02 * len(pts), f"{rechecks} of {len(pts)} points re-checked"
    dup = cent.copy()
    dup[1] = dup[0]
    slot.set_centroids(dup)
    _, counts, assign = slot.map_split(pts, want_assign=True)
    np.testing.assert_array_equal(assign, oracle.assign(pts, dup))
    assert slot.last_rechecks() >= counts[0] > 0
    slot.set_kernel(capi.KERNEL_AUTO)


@pytest.mark.parametrize("d,k,n,hot", [(64, 1024, 30011, 3), (64, 512, 20001, 1), (128, 600, 9001, 2), (100, 300, 7001, 5),
                                       (36, 20, 5003, 2)])
@pytest.mark.parametrize("form", ["pair_fused", "pair_segmented"])
def test_tensor_path_sums_on_skewed_clusters(slot, oracle, tensor_form, form, d, k, n, hot):
    """A few clusters hold most of the points, so that every 64-row window of a tile has many rows of one
    cluster: the fused form adds those up in registers (groups within a half, merged across the halves)
    before they leave as one reduction per cluster -- the sums must not notice.  One cluster alone gets
    every point in the last case."""
    tensor_form(form)
    rng = np.random.default_rng(d * 1000 + k)
    _, cent = cases.general_d(16, d, k, seed=d + k + 1)
    which = np.where(rng.random(n) < 0.8, rng.integers(0, hot, n), rng.integers(0, k, n))
    pts = (cent[which] + rng.normal(0, 0.5, size=(n, d))).astype(np.float32)
    run_case(slot, oracle, pts, cent, capi.KERNEL_TENSOR)
    one = (cent[np.full(3000, k - 1)] + rng.normal(0, 0.5, size=(3000, d))).astype(np.float32)
    run_case(slot, oracle, one, cent, capi.KERNEL_TENSOR)


def test_tensor_path_is_the_default_for_gemm_shapes(slot, oracle):
    pts, cent = cases.general_d(3000, 64, 1024, seed=3)
    run_case(slot, oracle, pts, cent, capi.KERNEL_AUTO, want_kernel=capi.KERNEL_TENSOR)
    pts, cent = cases.general_d(3000, 64, 100, seed=4)  # a quarter-filled centroid tile still beats the scan
    run_case(slot, oracle, pts, cent, capi.KERNEL_AUTO, want_kernel=capi.KERNEL_TENSOR)
    pts, cent = cases.general_d(3000, 8, 100, seed=5)   # d = 8, k*d < 2048: the register scan is cheaper than a padded tile
    run_case(slot, oracle, pts, cent, capi.KERNEL_AUTO, want_kernel=capi.KERNEL_EXACT)


def test_tensor_path_near_ties_are_rechecked_exactly(slot, oracle):
    """Duplicated and nearly-duplicated centroids, points on centroids and on bisectors: the
    TF32 scores cannot separate them, the exact re-check must (lowest index wins ties)."""
    rng = np.random.default_rng(9)
    d, k = 64, 256
    cent = rng.normal(0, 4, size=(k, d)).astype(np.float32)
    cent[17] = cent[5]                                   # exact duplicate
    cent[200] = cent[5]
    cent[33] = cent[32] + np.float32(1e-4)               # closer than TF32 can see
    cent[64] = np.nextafter(cent[63], np.float32(np.inf))  # one ulp apart
    pts = [cent[rng.integers(0, k, 2000)] + rng.normal(0, 1.0, size=(2000, d)).astype(np.float32)]
    pts.append(cent.copy())                              # points exactly on centroids
    pts.append(((cent[:-1].astype(np.float64) + cent[1:]) / 2).astype(np.float32))  # bisector points
    nan_row = np.zeros((3, d), np.float32)
    nan_row[0, 3] = np.nan
    nan_row[1, 0] = np.inf
    nan_row[2] = 1e30
    pts.append(nan_row)
    pts = np.concatenate(pts).astype(np.float32)
    run_case(slot, oracle, pts, cent, capi.KERNEL_TENSOR)
    a = oracle.assign(pts, cent)
    assert not np.isin(a, [17, 200]).any()


@pytest.mark.parametrize("form", ["single", "pair"])
def test_tensor_path_near_ties_across_centroid_tiles_at_k4096(slot, oracle, tensor_form, form):
    """Config 5's k: 32 centroid tiles (16 pair tiles) per point tile.  Duplicates and one-ulp neighbours are
    planted in DIFFERENT tiles (and at tile edges), so the running (best, second, tile) of a row has to be merged
    across tiles correctly and the lowest index has to win through the exact re-check."""
    tensor_form(form)
    rng = np.random.default_rng(4096)
    d, k = 128, 4096
    cent = rng.normal(0, 4, size=(k, d)).astype(np.float32)
    pairs = [(5, 4000), (127, 128), (255, 256), (128, 3967), (1000, 3000), (2047, 2048), (4095, 0), (300, 2348)]
    for lo, hi in pairs[:5]:
        cent[max(lo, hi)] = cent[min(lo, hi)]                       # exact duplicates, far apart in index
    for lo, hi in pairs[5:]:
        cent[hi] = np.nextafter(cent[lo], np.float32(np.inf))       # one ulp apart
    owners = np.array([min(p) for p in pairs] + [max(p) for p in pairs])
    which = np.concatenate([rng.choice(owners, 6000), rng.integers(0, k, 6000)])
    pts = (cent[which] + rng.normal(0, 0.5, size=(len(which), d))).astype(np.float32)
    pts = np.concatenate([pts, cent[owners]])                        # points exactly on the planted centroids
    run_case(slot, oracle, pts, cent, capi.KERNEL_TENSOR)
    assert slot.last_rechecks() >= 6000
    a = oracle.assign(pts, cent)
    assert not np.isin(a, [max(p) for p in pairs[:5]]).any()


@pytest.mark.parametrize("form", ["single", "pair"])
def test_tensor_path_with_non_finite_centroids(slot, oracle, tensor_form, form):
    """A NaN or infinite centroid poisons every score bound: nothing may be decided on the tensor
    cores then, and the exact re-check must reproduce the scan (NaN and +inf distances never win)."""
    tensor_form(form)
    pts, cent = cases.general_d(1500, 64, 256, seed=12)
    cent = cent.copy()
    cent[3, 5] = np.nan
    cent[100] = np.inf
    cent[200, 0] = -np.inf
    run_case(slot, oracle, pts, cent, capi.KERNEL_TENSOR)
    assert slot.last_rechecks() == len(pts)
