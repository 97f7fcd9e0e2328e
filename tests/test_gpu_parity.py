"""GPU parity tests proper: the CUDA path (through the C ABI) against the reference's outputs
in the golden bundles and against the oracle port on the same inputs.

Tolerances: log-likelihood 1e-9 relative (BASELINE.json north_star); per-site emissions and
forward tables 1e-9 relative; scaler counts exact."""
import numpy as np
import pytest

from conftest import load_golden, rel_err, GOLDEN_NAMES
from linearham_b200 import device
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _check(fam, tb, exp_ll, exp, tag, **opts):
    f = device.Family(fam)
    for k, v in opts.items():
        f.set_option(k, v)
    ll, out = f.score(tb, outputs=True)
    tm = f.timings()
    f.close()
    assert tm["kernel_launches"] >= 4
    assert rel_err(out["xmsa_emission"], exp["xmsa_emission"]) <= TOL, tag
    assert np.array_equal(out["germ_scalers"], exp["germ_scalers"]), tag
    assert np.array_equal(out["vd_scalers"], exp["vd_scalers"]), tag
    keys = ["vgerm_forward", "vd_forward", "jgerm_forward"]
    if fam["has_d"]:
        assert np.array_equal(out["dj_scalers"], exp["dj_scalers"]), tag
        keys += ["dgerm_forward", "dj_forward"]
    for k in keys:
        assert rel_err(out[k], exp[k]) <= TOL, (tag, k)
    assert rel_err(ll, exp_ll) <= TOL, tag
    return ll


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_matches_reference_golden(name):
    fam, tb, exp, _ = load_golden(name)
    _check(fam, tb, exp["loglik"], exp, name)


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_matches_oracle_port(name):
    fam, tb, _, _ = load_golden(name)
    pll, pout = po.port_score_batch(fam, tb)
    _check(fam, tb, pll, pout, name)


@pytest.mark.parametrize("ns,slots", [(1, 0), (2, 1), (3, -1), (4, 0), (4, 2), (3, 1)])
def test_tiling_and_stack_placement_do_not_change_results(ns, slots):
    """Results are independent of the site tiling (xMSA columns per thread) and of where the CLV
    stack lives (shared memory / L2 scratch)."""
    fam, tb, exp, _ = load_golden("synth_n40")
    a = _check(fam, tb, exp["loglik"], exp, "synth_n40", sites_per_thread=ns, smem_slots=slots)
    f = device.Family(fam)
    b = f.score(tb)
    assert np.array_equal(a, b)  # bit-identical


def test_batch_replication_and_chunking_are_bit_identical():
    fam, tb, exp, _ = load_golden("synth_n12")
    rep = 50
    big = {k: np.concatenate([v] * rep) for k, v in tb.items()}
    f = device.Family(fam)
    f.set_option("chunk_bytes", 1 << 16)  # force many chunks
    ll = f.score(big)
    ll1 = f.score(tb)
    assert np.array_equal(ll, np.tile(ll1, rep))
    assert rel_err(ll1, exp["loglik"]) <= TOL


def test_empty_batch_and_errors():
    fam, tb, _, _ = load_golden("catch_a")
    f = device.Family(fam)
    empty = {k: v[:0] for k, v in tb.items()}
    empty["rates"] = np.zeros((0, 4))
    assert f.score(empty).shape == (0,)
    bad = {k: v.copy() for k, v in tb.items()}
    bad["children"][0, 0, 0] = 77
    with pytest.raises(device.DeviceError):
        f.score(bad)
    bad = {k: v.copy() for k, v in tb.items()}
    bad["pi"][0, 0] = 0.0
    with pytest.raises(device.DeviceError):
        f.score(bad)
    badfam = dict(fam)
    badfam["tip_codes"] = fam["tip_codes"].copy()
    badfam["tip_codes"][0, 0] = 9
    with pytest.raises(device.DeviceError):
        device.Family(badfam)
