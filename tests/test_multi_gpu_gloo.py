"""CPU: the N > 1 host logic with world_size 2 over gloo -- split sharding and the allreduce of
partials give every rank the single-process result.  Per-rank partials come from the oracle
(there is no GPU here); on the GPU box the same flow runs through hgm_allreduce_partials."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.multiprocessing as mp

import cases
from hgpumap import shard


def test_split_ranges_never_cut_a_row():
    assert shard.split_ranges(100, 32, 8) == [(0, 32), (32, 32), (64, 32)]  # trailing 4 bytes are no row
    r = shard.split_ranges(8 * 1000, 8 * 300 + 5, 8)
    assert r == [(0, 2400), (2400, 2400), (4800, 2400), (7200, 800)]
    assert sum(l for _, l in r) == 8000 and all(o % 8 == 0 and l % 8 == 0 for o, l in r)
    with pytest.raises(ValueError):
        shard.split_ranges(100, 4, 8)


def test_splits_for_rank_partitions_everything_once():
    for n in (0, 1, 7, 8, 298):
        for world in (1, 2, 3, 8):
            got = [i for r in range(world) for i in shard.splits_for_rank(n, world, r)]
            assert got == list(range(n))
            sizes = [len(shard.splits_for_rank(n, world, r)) for r in range(world)]
            assert max(sizes) - min(sizes) <= 1
    with pytest.raises(ValueError):
        shard.splits_for_rank(4, 2, 2)


def _worker(rank, world, port, tmp):
    import torch.distributed as dist

    sys.path[:0] = [os.path.join(os.path.dirname(__file__)), os.path.join(os.path.dirname(__file__), "..", "hadoop-1.0.3-gpu_b200")]
    from conftest import Oracle, ORACLE_DIR

    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    oracle = Oracle(os.path.join(ORACLE_DIR, "libkmeans_oracle.so"))
    pts, cent = cases.blobs(40000, 2, 128, seed=5)
    ranges = shard.split_ranges(pts.nbytes, 8 * 3000, 8)
    sums = np.zeros((128, 2))
    counts = np.zeros(128, np.int64)
    for i in shard.splits_for_rank(len(ranges), world, rank):
        off, length = ranges[i]
        s, c, _, _ = oracle.map_combine(pts[off // 8:(off + length) // 8], cent)
        sums += s
        counts += c
    sums, counts = shard.allreduce_partials(sums, counts)
    np.savez(os.path.join(tmp, f"rank{rank}.npz"), sums=sums, counts=counts)
    dist.destroy_process_group()


def test_world_size_two_allreduce_equals_single_process(oracle, tmp_path):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    mp.spawn(_worker, args=(2, port, str(tmp_path)), nprocs=2, join=True)
    pts, cent = cases.blobs(40000, 2, 128, seed=5)
    ref_sums, ref_counts, _, _ = oracle.map_combine(pts, cent)
    for rank in range(2):
        got = np.load(tmp_path / f"rank{rank}.npz")
        np.testing.assert_array_equal(got["counts"], ref_counts)
        np.testing.assert_allclose(got["sums"], ref_sums, rtol=1e-12, atol=1e-9)


def test_centroid_update_keeps_empty_clusters():
    old = np.array([[1, 1], [5, 5], [9, 9]], np.float32)
    sums = np.array([[4.0, 2.0], [0.0, 0.0], [3.0, 30.0]])
    counts = np.array([2, 0, 3])
    new = shard.update_centroids(sums, counts, old)
    np.testing.assert_array_equal(new, np.array([[2, 1], [5, 5], [1, 10]], np.float32))
