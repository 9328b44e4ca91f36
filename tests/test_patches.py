"""CPU: the Java patches under patches/ are real unified diffs that apply to the reference tree (they cannot be
compiled here: no JDK).  SURVEY.md section 8(f)1 asks for them "as patches"."""
import glob
import os
import shutil
import subprocess

import pytest

from conftest import ROOT

REF = "/root/reference/hadoop-1.0.3"
PATCHES = sorted(glob.glob(os.path.join(ROOT, "patches", "*.patch")))

pytestmark = pytest.mark.skipif(not os.path.isdir(REF) or shutil.which("patch") is None,
                                reason="needs the reference tree and patch(1)")


def test_there_are_patches():
    assert len(PATCHES) >= 4


@pytest.mark.parametrize("path", PATCHES, ids=[os.path.basename(p) for p in PATCHES])
def test_patch_applies_to_the_reference_tree(path):
    """Dry run against the read-only reference: every hunk finds its place without fuzz or offset."""
    proc = subprocess.run(["patch", "--dry-run", "-p1", "-d", REF, "-i", path], capture_output=True, text=True)
    assert proc.returncode == 0, proc.stdout + proc.stderr
    assert "fuzz" not in proc.stdout and "offset" not in proc.stdout, proc.stdout


def test_patches_apply_in_sequence(tmp_path):
    """All of them, one after the other, on a copy of the files they touch."""
    touched = set()
    for path in PATCHES:
        for line in open(path):
            if line.startswith("+++ b/"):
                touched.add(line[6:].split("\t")[0].strip())
    for rel in touched:
        dst = tmp_path / rel
        dst.parent.mkdir(parents=True, exist_ok=True)
        shutil.copy(os.path.join(REF, rel), dst)
    for path in PATCHES:
        proc = subprocess.run(["patch", "-p1", "-d", str(tmp_path), "-i", path], capture_output=True, text=True)
        assert proc.returncode == 0, os.path.basename(path) + ": " + proc.stdout + proc.stderr
    runner = (tmp_path / "src/mapred/org/apache/hadoop/mapred/pipes/PipesGPUMapRunner.java").read_text(errors="replace")
    assert 'job.getInt("mapred.task.gpu.device.id", 0)' in runner
    conf = (tmp_path / "src/mapred/org/apache/hadoop/mapred/JobConf.java").read_text(errors="replace")
    assert "mapred.map.runnner.gpu.class" not in conf
