"""The oracle of the ancestral-state sampling step (oracle/asr_oracle.py; scripts/run_bootstrap_asr_ess.R:81-89,
lib/phylomd/src/simulation_functions.cpp:57-117) before it is trusted as the checker of the CUDA path:
  * its random stream against the published Philox4x32-10 known-answer vectors (Random123 kat_vectors);
  * its per-category site likelihoods against the compiled reference's goldens (their category mean is
    xmsa_emission) and against the port's per-category read-out;
  * its draws against the exact joint posterior of (rate category, inner states) on the reference's 4-tip Catch
    fixture, by enumeration - the distribution the R script samples from. The reference's own draws (R's generator)
    cannot be reproduced here: parity with them is unpinned and stays so."""
import numpy as np
import pytest

from conftest import load_golden
from oracle import asr_oracle as ao


def test_philox_known_answers():
    kat = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
           ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
           ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
            (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kat:
        got = ao.philox4x32_10(*[np.uint32(c) for c in ctr], *key)
        assert tuple(int(g) for g in got) == want


def _logmeanexp(a, axis):
    m = a.max(axis=axis, keepdims=True)
    return np.squeeze(m, axis) + np.log(np.exp(a - m).mean(axis=axis))


def xmsa_as_alignment(fam, B):
    """The xMSA as the alignment: every row from the family, the naive row's own codes per tree."""
    codes = np.asarray(fam["tip_codes"], dtype=np.uint8).reshape(fam["n_tips"], fam["n_sites"])
    return codes, np.repeat(codes[0][None, :], B, axis=0)


@pytest.mark.parametrize("name", ["catch_a", "synth_n12", "synth_igk_n30"])
def test_site_rate_logliks_against_the_reference_goldens(name):
    from oracle import pyoracle as po
    fam, tb, exp, _ = load_golden(name)
    B = min(3, len(tb["root_child"]))
    tb = {k: v[:B] for k, v in tb.items()}
    msa, naive = xmsa_as_alignment(fam, B)
    out = ao.asr_sample(fam["n_tips"], tb, msa, naive, naive_row=0, seed=1)
    srl = out["site_rate_loglik"]
    assert np.max(np.abs(_logmeanexp(srl, 1) - np.log(exp["xmsa_emission"][:B]))) <= 1e-9
    _, pout = po.port_score_batch(fam, tb, fields=["site_rate_loglik"])
    assert np.max(np.abs(srl - pout["site_rate_loglik"])) <= 1e-9


def chi_square_p(counts, probs):
    """Pearson chi-square of observed counts against cell probabilities (cells with small expectation pooled)."""
    from scipy.stats import chi2
    n = counts.sum()
    e = probs * n
    order = np.argsort(e)
    o_s, e_s = counts[order], e[order]
    pooled_o, pooled_e, acc_o, acc_e = [], [], 0.0, 0.0
    for o, ee in zip(o_s, e_s):
        acc_o += o
        acc_e += ee
        if acc_e >= 8:
            pooled_o.append(acc_o)
            pooled_e.append(acc_e)
            acc_o = acc_e = 0.0
    if acc_e > 0:
        pooled_o[-1] += acc_o
        pooled_e[-1] += acc_e
    o, ee = np.array(pooled_o), np.array(pooled_e)
    stat = ((o - ee) ** 2 / ee).sum()
    return float(chi2.sf(stat, len(o) - 1)), len(o)


def posterior_case(n_rep=1500, n_cols=6):
    """The 4-tip Catch fixture, sample 0 repeated n_rep times, on n_cols of its xMSA columns (with an N among the
    tips where the fixture has one) -> (T, tb, msa, naive, columns)."""
    fam, tb, exp, _ = load_golden("catch_a")
    T = fam["n_tips"]
    codes = np.asarray(fam["tip_codes"], dtype=np.uint8).reshape(T, fam["n_sites"])
    cols = np.linspace(0, fam["n_sites"] - 1, n_cols).astype(int)
    tb1 = {k: np.repeat(v[:1], n_rep, axis=0) for k, v in tb.items()}
    tb1["blen"] = tb1["blen"] * 6.0          # longer branches: a posterior with mass on many cells
    tb1["root_blen"] = tb1["root_blen"] * 6.0
    msa = codes[:, cols]
    return T, tb1, msa, np.repeat(msa[0][None, :], n_rep, axis=0), cols


def check_against_exact_posterior(T, tb, msa, out):
    n_rep, S = out["site_rate"].shape
    worst = 1.0
    for j in range(S):
        post = ao.exact_posterior(T, tb, 0, msa[:, j])
        keys = sorted(post)
        index = {k: i for i, k in enumerate(keys)}
        counts = np.zeros(len(keys))
        for b in range(n_rep):
            counts[index[(int(out["site_rate"][b, j]),) + tuple(int(s) for s in out["node_states"][b, :, j])]] += 1
        p, cells = chi_square_p(counts, np.array([post[k] for k in keys]))
        assert cells >= 4, "the case is too concentrated to test anything"
        worst = min(worst, p)
    return worst


def test_draws_follow_the_exact_posterior():
    T, tb, msa, naive, _ = posterior_case()
    out = ao.asr_sample(T, tb, msa, naive, naive_row=0, seed=2024)
    assert check_against_exact_posterior(T, tb, msa, out) > 1e-3   # six chi-square tests, fixed seed
    again = ao.asr_sample(T, tb, msa, naive, naive_row=0, seed=2024)
    other = ao.asr_sample(T, tb, msa, naive, naive_row=0, seed=2025)
    assert np.array_equal(out["node_states"], again["node_states"])
    assert not np.array_equal(out["node_states"], other["node_states"])
