"""GPU: the HBM split cache and the iteration driver (SURVEY.md section 8(f)4, config 3's shape:
iteration 1 staged from the host, later iterations from HBM)."""
import numpy as np
import pytest

import cases
from conftest import check_partials
from hgpumap import capi, driver, shard

pytestmark = pytest.mark.gpu


def test_cache_put_map_drop(oracle, tmp_path):
    pts, cent = cases.blobs(120001, 2, 128, seed=51)
    parts = np.array_split(pts, 3)
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        slot.cache_put(0, parts[0])
        slot.cache_put(1, parts[1].astype(">f4"), fmt=capi.FMT_F32BE)       # decoded once, on the way in
        path = str(tmp_path / "part2.bin")
        parts[2].astype("<f4").tofile(path)
        slot.cache_put_file(2, path)
        assert slot.cache_info() == (3, pts.nbytes)
        for _ in range(2):  # mapping from the cache is repeatable and adds into the partials
            slot.reset_partials()
            n = sum(slot.cache_map(i) for i in range(3))
            assert n == len(pts)
            sums, counts = slot.read_partials()
            o = oracle.map_combine(pts, cent)
            check_partials(sums, counts, o[0], o[1], o[2])
        slot.cache_put(1, parts[0])                                          # replace an entry
        assert slot.cache_info()[0] == 3
        slot.cache_drop(1)
        assert slot.cache_info()[0] == 2
        with pytest.raises(capi.HgmError):
            slot.cache_map(1)
        slot.cache_drop()
        assert slot.cache_info() == (0, 0)
        with pytest.raises(capi.HgmError) as err:
            slot.cache_put(5, b"\x00" * 12)
        assert err.value.code == -1002


def test_map_all_merges_neighbours_into_one_launch(oracle):
    pts, cent = cases.blobs(6 * 40000, 2, 128, seed=53)
    parts = np.array_split(pts, 6)  # 320000 bytes each: whole 16-byte groups, so they sit back to back
    with capi.GpuMapSlot(0) as slot:
        slot.set_centroids(cent)
        for i, part in enumerate(parts):
            slot.cache_put(10 + i, part)
        slot.reset_partials()
        assert slot.cache_map_all() == (len(pts), 1)
        sums, counts = slot.read_partials()
        o = oracle.map_combine(pts, cent)
        check_partials(sums, counts, o[0], o[1], o[2])
        # a task's list of splits: neighbours share a launch, a gap starts a new one, a repeated id is mapped twice
        slot.reset_partials()
        assert slot.cache_map_many([14, 10, 11, 13]) == (4 * 40000, 2)
        sums, counts = slot.read_partials()
        o = oracle.map_combine(np.concatenate([parts[i] for i in (0, 1, 3, 4)]), cent)
        check_partials(sums, counts, o[0], o[1], o[2])
        ms, launches = slot.last_map_ms()
        assert ms > 0 and launches == 2  # one timed region over both launches
        slot.reset_partials()
        assert slot.cache_map_many([12, 12]) == (2 * 40000, 2)
        assert slot.read_partials()[1].sum() == 2 * 40000
        with pytest.raises(capi.HgmError):
            slot.cache_map_many([10, 77])
        slot.cache_drop(12)  # a hole: two runs now
        slot.reset_partials()
        n, groups = slot.cache_map_all()
        assert (n, groups) == (len(pts) - 40000, 2)
        keep = np.concatenate([parts[i] for i in (0, 1, 3, 4, 5)])
        sums, counts = slot.read_partials()
        o = oracle.map_combine(keep, cent)
        check_partials(sums, counts, o[0], o[1], o[2])
        slot.cache_put(12, parts[2][:1001])  # appended right behind the last split; odd size: padded at ITS end
        slot.cache_put(99, parts[2][:10])    # ... so this one does not touch its neighbour
        slot.reset_partials()
        assert slot.cache_map_all() == (len(pts) - 40000 + 1011, 3)
        slot.set_centroids(np.zeros((3, 5), np.float32))  # another d: the rows mean nothing now
        assert slot.cache_info() == (0, 0)


def test_ten_iterations_from_the_cache_follow_the_oracle(oracle):
    """Each iteration's assignments are checked against the oracle run on the SAME centroids the
    GPU path used (sums differ in their last bits, so free-running trajectories may part by an ulp)."""
    pts, cent0 = cases.blobs(90000, 2, 64, seed=61)
    rng = np.random.default_rng(1)
    cent0 = (cent0 + rng.normal(0, 3, cent0.shape)).astype(np.float32)  # a poor start, so centroids move
    splits = np.array_split(pts, 4)
    seen = []

    def check(it, new_cent, sums, counts):
        used = seen[-1] if seen else cent0
        o_sums, o_counts, o_abs, _ = oracle.map_combine(pts, used)
        check_partials(sums, counts, o_sums, o_counts, o_abs)
        np.testing.assert_array_equal(new_cent, shard.update_centroids(sums, counts, used))
        seen.append(new_cent.copy())

    with capi.GpuMapSlot(0) as slot:
        final, history = driver.run_kmeans(slot, splits, cent0, iterations=10, on_iteration=check)
        assert slot.cache_info() == (4, pts.nbytes)
    assert len(history) == 10 and all(h[1].sum() == len(pts) for h in history)
    # Lloyd's objective never increases
    def cost(c):
        a = oracle.assign(pts, c)
        return float(((pts.astype(np.float64) - c[a]) ** 2).sum())
    costs = [cost(cent0)] + [cost(c) for c in seen]
    assert all(b <= a * (1 + 1e-9) for a, b in zip(costs, costs[1:]))
    assert costs[-1] < 0.5 * costs[0]
