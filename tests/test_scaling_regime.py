"""The underflow regime (libpll test/out/scaling.out): deep ladders, tiny alpha, 4 and 8 rate
categories, 1000-3000 tips, scored by the reference's own C++ (tests/golden/scaling_*.npz). libpll rescales a
site when ALL 4R entries of a node are below 2^-256 (core_partials.c:648-659); these families are where a
per-category rule would keep different counts. The CPU restatement and the CUDA path must both give the
reference's doubles (1e-9 where they are finite and non-zero, the same class where not) and its scaler counts."""
import numpy as np
import pytest

from conftest import SCALING_NAMES, compare_in_underflow_regime, load_golden
from oracle import pyoracle as po


def test_there_are_scaling_goldens():
    assert len(SCALING_NAMES) >= 6


@pytest.mark.parametrize("name", SCALING_NAMES)
def test_port_matches_reference_in_underflow_regime(name):
    fam, tb, exp, _ = load_golden(name)
    one = {k: v[:1] for k, v in tb.items()}   # (a 3000-tip tree takes the scalar port a few seconds)
    ll, out = po.port_score_batch(fam, one)
    compare_in_underflow_regime(ll, out, {k: v[:1] for k, v in exp.items()})


@pytest.mark.gpu
@pytest.mark.parametrize("patterns", [1, 0])
@pytest.mark.parametrize("name", SCALING_NAMES)
def test_cuda_matches_reference_in_underflow_regime(name, patterns):
    from linearham_b200 import device
    fam, tb, exp, _ = load_golden(name)
    f = device.Family(fam)
    f.set_option("site_patterns", patterns)
    ll, out = f.score(tb, outputs=True)
    f.close()
    worst = compare_in_underflow_regime(ll, out, exp)
    print(f"{name} patterns={patterns}: worst relative deviation of a finite emission {worst:.2e}")
