"""CPU: run the reference's own Pipes runtime (oracle/_ref) live, when it is there.
These re-derive what tests/golden/ froze, so a drift of either side shows up here."""
import os

import numpy as np
import pytest

import cases
import pipes_harness as ph
from conftest import GOLDEN
from hgpumap import wire


@pytest.fixture(scope="module")
def live(ref_bins, tmp_path_factory):
    tmp = tmp_path_factory.mktemp("ref")
    pts, cent = cases.blobs(6000, 2, 128, seed=99)
    cpath = str(tmp / "cent.bin")
    cent.astype("<f4").tofile(cpath)
    jobs = [
        (ref_bins["wordcount-part"], ph.wordcount_downlink(2), None),
        (ref_bins["wordcount-simple"], ph.wordcount_downlink(0), None),
        # one point per record, tiny io.sort.mb -> many combiner spills (HadoopPipes.cc:683-690)
        (ref_bins["kmeans-pipes-cpu"], ph.kmeans_downlink(pts, cpath, 128, 2, "binary", 1, 1, io_sort_mb=0), None),
        # 64 points per record, no reduces -> no combiner, raw map output
        (ref_bins["kmeans-pipes-cpu"], ph.kmeans_downlink(pts[:1000], cpath, 128, 2, "binary", 0, 64), None),
        # abort: child must fail without DONE (HadoopPipes.cc:855-857)
        (ref_bins["wordcount-part"], wire.Downlink().authenticate().start().abort(), None),
    ]
    return pts, cent, ph.run_children(jobs)


def test_live_wordcount_equals_golden(live):
    _, _, res = live
    assert res[0][1] == open(os.path.join(GOLDEN, "wordcount_part.uplink"), "rb").read()
    assert res[1][1] == open(os.path.join(GOLDEN, "wordcount_simple.uplink"), "rb").read()
    assert res[0][0].returncode == 1  # `return runTask(...)`: true -> exit status 1 (wordcount-simple.cc:63-66)


def test_live_kmeans_with_spills_matches_oracle(live, oracle):
    pts, cent, res = live
    up = wire.parse_uplink(res[2][1])
    assert up.done
    counts, sums = wire.kmeans_partials([(k, v) for _, k, v in up.partitioned], 128, 2)
    assert len(up.partitioned) == 6000  # io.sort.mb=0: a spill per emit, one record each
    o_sums, o_counts, _, _ = oracle.map_combine(pts, cent)
    np.testing.assert_array_equal(counts, o_counts)
    np.testing.assert_allclose(sums, o_sums, rtol=1e-12, atol=1e-9)
    assert up.counter("KMEANS", "POINTS") == 6000


def test_live_kmeans_without_reduces_emits_raw_map_output(live, oracle):
    pts, cent, res = live
    up = wire.parse_uplink(res[3][1])
    assert up.done and not up.partitioned and len(up.outputs) == 1000
    assign = oracle.assign(pts[:1000], cent)
    assert [int(k) for k, _ in up.outputs] == assign.tolist()
    first = up.outputs[0][1].split()
    assert first[0] == b"1" and np.float32(float(first[1])) == pts[0, 0]


def test_live_abort_fails_the_task(live):
    _, _, res = live
    proc, data, _ = res[4]
    assert b"Aborted by driver" in proc.stderr
    assert proc.returncode == 0  # false -> exit status 0
    assert not wire.parse_uplink(data).done
