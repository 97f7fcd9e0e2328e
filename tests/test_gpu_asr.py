"""lh_asr_sample (linearham_b200/csrc/asr.cu) - ancestral-state sampling, the second half of the reference's R step
(scripts/run_bootstrap_asr_ess.R:81-89, phylomd asr.sim) - through the C ABI:
  * its draws equal the oracle's (oracle/asr_oracle.py: same Philox counters, same order of operations) exactly, its
    per-category site log-likelihoods within 1e-9 - of the oracle, of k_edge's read-out of the main path, and through
    their category mean of the compiled reference's xmsa_emission;
  * its draws follow the exact joint posterior of the reference's 4-tip Catch fixture (chi-square, fixed seed);
  * the result depends on the arguments only (seed, not chunking), and on a 1000-tip tree whose partials underflow.
The reference's own draws (R's generator on a worker cluster) cannot be reproduced: that parity is unpinned."""
import numpy as np
import pytest

from conftest import load_golden
from linearham_b200 import device
from oracle import asr_oracle as ao
from test_asr_oracle import xmsa_as_alignment, posterior_case, check_against_exact_posterior, _logmeanexp

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["catch_a", "synth_n12", "synth_n40", "synth_igk_n30"])
def test_draws_equal_the_oracle(name):
    fam, tb, exp, _ = load_golden(name)
    B = min(4, len(tb["root_child"]))
    tb = {k: v[:B] for k, v in tb.items()}
    msa, naive = xmsa_as_alignment(fam, B)
    rng = np.random.default_rng(5)
    cases = [(msa, naive, 0)]
    # the alignment of the R step: a naive sequence per sample (N where the sampler left one), N / gaps in the tips
    msa2 = msa.copy()
    msa2[rng.random(msa.shape) < 0.05] = 4
    cases.append((msa2, rng.integers(0, 5, size=naive.shape).astype(np.uint8), 0))
    cases.append((msa2, naive, -1))          # no row replaced
    f = device.Family(fam)
    for m, nv, row in cases:
        got = f.asr_sample(tb, m, nv, naive_row=row, seed=77, loglik=True)
        want = ao.asr_sample(fam["n_tips"], tb, m, nv, naive_row=row, seed=77)
        assert np.array_equal(got["site_rate"], want["site_rate"])
        assert np.array_equal(got["node_states"], want["node_states"])
        fin = np.isfinite(want["site_rate_loglik"])
        assert np.array_equal(fin, np.isfinite(got["site_rate_loglik"]))
        assert np.max(np.abs(got["site_rate_loglik"][fin] - want["site_rate_loglik"][fin])) <= 1e-9
    # against the main path: k_edge's per-category read-out and the reference's emission
    got = f.asr_sample(tb, msa, naive, naive_row=0, seed=1, loglik=True)
    ll, out = f.score(tb, outputs=["site_rate_loglik"])
    f.close()
    assert np.max(np.abs(got["site_rate_loglik"] - out["site_rate_loglik"])) <= 1e-9
    assert np.max(np.abs(_logmeanexp(got["site_rate_loglik"], 1) - np.log(exp["xmsa_emission"][:B]))) <= 1e-9


def test_draws_follow_the_exact_posterior():
    T, tb, msa, naive, _ = posterior_case(n_rep=4000, n_cols=8)
    fam = load_golden("catch_a")[0]
    f = device.Family(fam)
    out = f.asr_sample(tb, msa, naive, naive_row=0, seed=31337)
    assert check_against_exact_posterior(T, tb, msa, out) > 1e-3
    # the arguments decide the result: the seed does, the chunking does not
    assert np.array_equal(out["node_states"], f.asr_sample(tb, msa, naive, naive_row=0, seed=31337)["node_states"])
    assert not np.array_equal(out["node_states"], f.asr_sample(tb, msa, naive, naive_row=0, seed=31338)["node_states"])
    f.set_option("asr_scratch_mb", 1)
    small = f.asr_sample(tb, msa, naive, naive_row=0, seed=31337)
    f.close()
    assert np.array_equal(out["node_states"], small["node_states"]) and np.array_equal(out["site_rate"], small["site_rate"])


def test_a_thousand_tips_whose_partials_underflow():
    fam, tb, exp, _ = load_golden("scaling_rand2000_a003_r4")
    tb = {k: v[:1] for k, v in tb.items()}
    msa, naive = xmsa_as_alignment(fam, 1)
    msa, naive = msa[:, :48], naive[:, :48]
    f = device.Family(fam)
    got = f.asr_sample(tb, msa, naive, naive_row=0, seed=9, loglik=True)
    f.close()
    want = ao.asr_sample(fam["n_tips"], tb, msa, naive, naive_row=0, seed=9)
    fin = np.isfinite(want["site_rate_loglik"])
    assert fin.any() and np.array_equal(fin, np.isfinite(got["site_rate_loglik"]))
    g, w = got["site_rate_loglik"][fin], want["site_rate_loglik"][fin]
    assert np.max(np.abs(g - w) / np.maximum(1.0, np.abs(w))) <= 1e-9
    assert (w < -1000).any()                       # e^-1000 is not a double: the rescaling counts are what is compared
    assert np.array_equal(got["site_rate"], want["site_rate"])
    assert np.mean(got["node_states"] != want["node_states"]) <= 1e-4   # a draw within 1e-13 of a boundary may differ
    assert (got["node_states"] <= 4).all()


def test_arguments_are_checked():
    fam, tb, exp, _ = load_golden("synth_n12")
    tb = {k: v[:2].copy() for k, v in tb.items()}
    msa, naive = xmsa_as_alignment(fam, 2)
    f = device.Family(fam)
    bad = dict(tb)
    bad["children"] = tb["children"].copy()
    bad["children"][1, 0, 0] = fam["n_tips"] + 5
    with pytest.raises(device.DeviceError, match="tree 1: operation 0 refers to node"):
        f.asr_sample(bad, msa, naive)
    bad = dict(tb)
    bad["blen"] = -tb["blen"]
    with pytest.raises(device.DeviceError, match="negative branch length"):
        f.asr_sample(bad, msa, naive)
    with pytest.raises(device.DeviceError, match="naive_row"):
        f.asr_sample(tb, msa, naive, naive_row=fam["n_tips"])
    f.close()
