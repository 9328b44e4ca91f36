"""CPU: the C-ABI library loads and exports exactly what include/hgpumap.h declares.
No compute is called here -- there is no GPU in the build container."""
import ctypes as C
import os
import re

import numpy as np
import pytest

from conftest import ROOT, _have_gpu
from hgpumap import capi

HEADER = os.path.join(ROOT, "include", "hgpumap.h")


def declared_symbols():
    text = open(HEADER).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(hgm_[a-z0-9_]+)\s*\(", text)))


def test_header_and_binding_list_agree():
    assert declared_symbols() == sorted(capi.SIGNATURES)


def test_library_exports_every_declared_symbol():
    lib = capi.load_library()
    for name in declared_symbols():
        assert hasattr(lib, name), f"{name} declared in include/hgpumap.h but not exported"
    assert lib.hgm_version() == 1001  # major 1, minor 1: entry points were added in round 2, none changed


def test_every_entry_point_cites_the_reference():
    """Each declaration's comment names the reference interface it stands under."""
    text = open(HEADER).read()
    for name in ("hgm_open", "hgm_set_centroids", "hgm_map_split", "hgm_map_file", "hgm_last_error",
                 "hgm_device_count", "hgm_comm_id"):
        decl = re.search(rf"^(?:int|void|const char \*)\s*{name}\(", text, flags=re.M)
        head = text[:decl.start()]
        comment = head[head.rindex("/*"):]
        assert re.search(r"\.(java|cc|hh):[\s*]*\d+", comment), f"{name} has no reference file:line citation"


def test_library_is_sm100a_only():
    """The shipped code object must be sm_100a SASS (no PTX-JIT, no other arch)."""
    import subprocess
    cuobjdump = "/usr/local/cuda/bin/cuobjdump"
    if not os.path.exists(cuobjdump):
        pytest.skip("no cuobjdump")
    out = subprocess.run([cuobjdump, "-lelf", capi.library_path()], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_\d+a?", out))
    assert archs == {"sm_100a"}, archs


@pytest.mark.skipif(_have_gpu(), reason="negative test for GPU-less hosts")
def test_open_fails_loudly_without_a_gpu():
    with pytest.raises(capi.HgmError) as err:
        capi.GpuMapSlot(0)
    assert err.value.code == -19 and "no CPU fallback" in str(err.value)


def test_host_side_synth_matches_numpy_restatement():
    from hgpumap import synth
    c = capi.synth_centres(5, 9, 3)
    assert np.array_equal(c.view(np.uint32), synth.centres(5, 9, 3).view(np.uint32))
    x = capi.synth_points_host(5, 77, 1000, c)
    assert np.array_equal(x.view(np.uint32), synth.points(5, 77, 1000, c).view(np.uint32))
    p = capi.synth_perturb(6, c)
    assert np.array_equal(p.view(np.uint32), synth.perturb(6, c).view(np.uint32))


def test_bad_arguments_are_rejected_without_a_device():
    lib = capi.load_library()
    assert lib.hgm_device_count(None) == -22
    assert lib.hgm_set_centroids(None, None, 1, 1) == -22
    assert b"NULL" in lib.hgm_last_error(None)
    assert lib.hgm_synth_centres(1, 0, 2, None) == -22


def test_tensor_form_constants_match_the_header():
    """hgpumap/capi.py mirrors include/hgpumap.h: a form added on one side only would be rejected as HGM_EINVAL."""
    import re

    from hgpumap import capi

    text = open(os.path.join(os.path.dirname(__file__), "..", "include", "hgpumap.h")).read()
    header = {m.group(1): int(m.group(2)) for m in re.finditer(r"#define HGM_(TENSOR_FORM_\w+) (\d+)", text)}
    assert header == {name: getattr(capi, name) for name in header}
    assert sorted(header.values()) == list(range(len(header)))
