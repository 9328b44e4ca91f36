"""GPU, two or more devices: split-sharded map tasks + hgm_allreduce_partials (NCCL over NVLink, csrc/hgm_comm.cu)
against the oracle on the concatenated input.  Skipped on a one-GPU box; run with `gpurun --gpus 2`.

One process per GPU, as the library is used (one GPU map slot = one child process, slot i <-> device i,
mapred/TaskTrackerStatus.java:536-551).  Counts must be bit-exact on every rank, sums within the stated tolerance."""
import os
import sys

import numpy as np
import pytest

import cases
from conftest import check_partials

pytestmark = pytest.mark.gpu


def _n_devices():
    try:
        import torch

        return torch.cuda.device_count()
    except Exception:
        return 0


def _exchange_handles(tmp, rank, world, mine, timeout=60.0):
    """All-gather of the 64-byte inbox handles through files (the host's job: JobConf / a file / torch.distributed)."""
    import time

    with open(os.path.join(tmp, f"handle{rank}.tmp"), "wb") as f:
        f.write(mine)
    os.replace(os.path.join(tmp, f"handle{rank}.tmp"), os.path.join(tmp, f"handle{rank}"))
    t0, out = time.time(), []
    for r in range(world):
        path = os.path.join(tmp, f"handle{r}")
        while not os.path.exists(path):
            if time.time() - t0 > timeout:
                raise TimeoutError(path)
            time.sleep(0.01)
        out.append(open(path, "rb").read())
    return out


def _worker(rank, world, comm_id, tmp, d, k, n, iterations, use_peer):
    here = os.path.dirname(os.path.abspath(__file__))
    sys.path[:0] = [here, os.path.join(here, "..", "hadoop-1.0.3-gpu_b200")]
    from hgpumap import capi, shard

    pts, cent = (cases.blobs(n, 2, k, seed=21) if d == 2 else cases.general_d(n, d, k, seed=d + k))
    ranges = shard.split_ranges(pts.nbytes, 4 * d * 4099, 4 * d)
    mine = shard.splits_for_rank(len(ranges), world, rank)
    with capi.GpuMapSlot(rank) as slot:
        slot.set_centroids(cent)
        slot.comm_init(world, rank, comm_id)
        if use_peer:  # small partials go through peer memory, larger ones still through NCCL
            slot.peer_open(rank, _exchange_handles(tmp, rank, world, slot.peer_handle(world)))
        row = 4 * d
        for it in range(iterations):  # the second iteration checks that the counts came back as integers
            slot.reset_partials()
            for i in mine:
                off, length = ranges[i]
                chunk = np.ascontiguousarray(pts[off // row:(off + length) // row])
                d_pts = slot.device_alloc(max(chunk.nbytes, 16))
                slot.memcpy_h2d(d_pts, chunk)
                slot.map_device(d_pts, chunk.shape[0])
                slot.sync()
                slot.device_free(d_pts)
            local_sums, local_counts = slot.read_partials()
            slot.allreduce_partials()
            sums, counts = slot.read_partials()
        np.savez(os.path.join(tmp, f"rank{rank}.npz"), sums=sums, counts=counts, local_counts=local_counts,
                 kernel=slot.last_kernel())


@pytest.mark.skipif(_n_devices() < 2, reason="needs at least two GPUs")
@pytest.mark.parametrize("use_peer", [False, True], ids=["nccl", "peer-memory"])
@pytest.mark.parametrize("d,k,n", [(2, 128, 300001), (2, 1000, 50001), (64, 1024, 30011), (5, 7, 20001)])
def test_allreduce_of_partials_matches_oracle_on_the_whole_input(oracle, tmp_path, d, k, n, use_peer):
    import torch.multiprocessing as mp

    from hgpumap import capi

    world = min(_n_devices(), 8)
    comm_id = capi.GpuMapSlot.comm_id()
    mp.spawn(_worker, args=(world, comm_id, str(tmp_path), d, k, n, 3, use_peer), nprocs=world, join=True)
    pts, cent = (cases.blobs(n, 2, k, seed=21) if d == 2 else cases.general_d(n, d, k, seed=d + k))
    r_sums, r_counts, r_abs, _ = oracle.map_combine(pts, cent)
    local_total = np.zeros(k, np.int64)
    first = np.load(tmp_path / "rank0.npz")
    for rank in range(world):
        got = np.load(tmp_path / f"rank{rank}.npz")
        if use_peer and (k * d + k) * 8 <= 64 * 1024:  # summed in rank order on every rank: bit-identical everywhere
            np.testing.assert_array_equal(got["sums"].view(np.uint64), first["sums"].view(np.uint64))
        check_partials(got["sums"], got["counts"], r_sums, r_counts, r_abs)
        local_total += got["local_counts"]
    np.testing.assert_array_equal(local_total, r_counts)  # the shards covered every point exactly once
