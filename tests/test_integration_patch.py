"""integration/SConstruct.sm100a.patch (north_star: "SConstruct gains an nvcc sm_100a target") must
apply to the reference's SConstruct; runs only where /root/reference exists (this container)."""
import ast
import os
import shutil
import subprocess

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATCH = os.path.join(ROOT, "integration", "SConstruct.sm100a.patch")
REF_SCONSTRUCT = "/root/reference/SConstruct"


@pytest.mark.skipif(not os.path.exists(REF_SCONSTRUCT) or shutil.which("patch") is None,
                    reason="needs the reference checkout and patch(1)")
def test_sconstruct_patch_applies(tmp_path):
    shutil.copy(REF_SCONSTRUCT, tmp_path / "SConstruct")
    dry = subprocess.run(["patch", "-p1", "--dry-run", "-d", str(tmp_path), "-i", PATCH],
                         capture_output=True, text=True)
    assert dry.returncode == 0, dry.stdout + dry.stderr
    subprocess.run(["patch", "-p1", "-d", str(tmp_path), "-i", PATCH], check=True, capture_output=True)
    text = (tmp_path / "SConstruct").read_text()
    ast.parse(text)  # SConstruct is Python
    assert "compute_100a" in text and "linearham_b200_build" in text and "--devices" in text
    # the run step keeps the reference's command line when --b200 is not given
    assert '"_build/linearham/linearham") + " --pipeline"' in text
