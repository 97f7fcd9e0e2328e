"""The reference's own Catch TEST_CASE("PhyloHMM") (test/test.cpp:750-1399: ~200 assertions on the state space, the
transition matrices, the xMSA, three known-answer log-likelihoods and a sampled annotation) compiled from where it
lies in the reference tree against the product's classes, through the source-compatibility adapter
integration/compat/linearham_compat.hpp, and run on the GPU. The binary is built by __graft_entry__.build() (or
integration/compat/build_catch.sh) where /root/reference exists and travels to the GPU box under build/compat/."""
import os
import re
import subprocess
import tempfile

import numpy as np
import pytest

from conftest import GOLDEN_DIR

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
BINARY = os.path.join(ROOT, "build", "compat", "catch_phylohmm")
HAVE_REF = os.path.isfile("/root/reference/test/test.cpp")


def _data_dir():
    z = np.load(os.path.join(GOLDEN_DIR, "fixture_catch_files.npz"), allow_pickle=False)
    d = tempfile.mkdtemp(prefix="lh_catch_")
    for name, text in zip(z["names"], z["texts"]):
        path = os.path.join(d, "data", str(name))
        os.makedirs(os.path.dirname(path), exist_ok=True)
        open(path, "w").write(str(text))
    return d


@pytest.mark.skipif(not HAVE_REF, reason="needs the reference tree (test/test.cpp, lib/eigen)")
def test_reference_catch_case_compiles_against_the_adapter_and_refuses_to_run_without_a_gpu():
    import torch
    out = subprocess.run([os.path.join(ROOT, "integration", "compat", "build_catch.sh")], capture_output=True, text=True)
    assert out.returncode == 0, out.stderr[-2000:]
    assert os.path.isfile(BINARY)
    if torch.cuda.is_available():
        return
    run = subprocess.run([BINARY], cwd=_data_dir(), capture_output=True, text=True)
    assert run.returncode != 0 and "no CPU fallback" in run.stdout   # the product path fails loudly


@pytest.mark.gpu
def test_reference_catch_case_passes_on_the_gpu():
    if not os.path.isfile(BINARY):
        pytest.skip("build/compat/catch_phylohmm was not built (no reference tree where build() ran)")
    run = subprocess.run([BINARY], cwd=_data_dir(), capture_output=True, text=True, timeout=600)
    assert run.returncode == 0, run.stdout[-3000:]
    m = re.search(r"All tests passed \((\d+) assertions? in 1 test case\)", run.stdout)
    assert m and int(m.group(1)) >= 190, run.stdout[-500:]
