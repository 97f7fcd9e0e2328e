"""Per-category site log-likelihoods (lh_outputs_t.site_rate_loglik): what the reference's ancestral-sequence
step recomputes per kept tree, site and rate category with phylomd (scripts/run_bootstrap_asr_ess.R:84-87:
phylo.likelihood(rate-scaled tree, GTR(scale = TRUE), column) / pi[naive base]).

Pinned three ways: (1) the mean over categories is xmsa_emission, which the compiled reference fixes;
(2) an independent restatement here (scipy expm of the rate-scaled GTR generator + plain pruning);
(3) GPU against the port."""
import numpy as np
import pytest
from scipy.linalg import expm

from conftest import load_golden, rel_err

FIELDS = ["xmsa_emission", "site_rate_loglik"]


def _logmeanexp(a, axis):
    m = a.max(axis=axis, keepdims=True)
    m = np.where(np.isfinite(m), m, 0.0)
    return np.squeeze(m, axis) + np.log(np.exp(a - m).mean(axis=axis))


def _independent(fam, tb, b):
    """log( L_k(column) / pi[naive] ) for sample b by brute force -> [R][S]."""
    T, S = fam["n_tips"], fam["n_sites"]
    er, pi, rates = tb["er"][b], tb["pi"][b], tb["rates"][b]
    Q = np.zeros((4, 4))
    iu = np.triu_indices(4, 1)
    Q[iu] = er
    Q = (Q + Q.T) * pi[None, :]
    np.fill_diagonal(Q, -Q.sum(axis=1))
    Q /= -(pi * np.diag(Q)).sum()                      # mean rate 1 (phylomd scale = TRUE; models.c:240-244)
    codes = fam["tip_codes"]
    tip = np.concatenate([np.eye(4), np.ones((1, 4))])  # A C G T N
    out = np.zeros((len(rates), S))
    for k, r in enumerate(rates):
        clv = {t: tip[codes[t]] for t in range(T)}      # [S][4]
        for op, (c1, c2) in enumerate(tb["children"][b]):
            pl, pr = expm(Q * r * tb["blen"][b][op, 0]), expm(Q * r * tb["blen"][b][op, 1])
            clv[T + op] = (clv[int(c1)] @ pl.T) * (clv[int(c2)] @ pr.T)
        pm = expm(Q * r * tb["root_blen"][b])
        lk = ((clv[2 * T - 3] * pi[None, :]) * (clv[int(tb["root_child"][b])] @ pm.T)).sum(axis=1)
        naive_p = np.where(codes[0] == 4, 1.0, pi[np.minimum(codes[0], 3)])
        out[k] = np.log(lk / naive_p)
    return out


@pytest.mark.parametrize("name", ["catch_a", "catch_c", "synth_n12", "synth_n2", "synth_igk_n30"])
def test_port_rate_logliks_are_pinned(name):
    from oracle import pyoracle as po
    fam, tb, exp, _ = load_golden(name)
    ll, out = po.port_score_batch(fam, tb, fields=FIELDS)
    srl = out["site_rate_loglik"]
    assert srl.shape == (len(ll), tb["rates"].shape[1], fam["n_sites"])
    # (1) category mean = the emission the reference's own output fixes
    assert np.max(np.abs(_logmeanexp(srl, 1) - np.log(exp["xmsa_emission"]))) <= 1e-9
    # (2) independent restatement
    for b in range(min(2, len(ll))):
        assert np.max(np.abs(srl[b] - _independent(fam, tb, b))) <= 1e-9, (name, b)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["catch_a", "catch_b", "synth_n12", "synth_n40", "synth_n24_r1", "synth_igk_n30", "synth_n2"])
def test_gpu_rate_logliks_match_the_port(name):
    from linearham_b200 import device
    from oracle import pyoracle as po
    fam, tb, exp, _ = load_golden(name)
    pll, pout = po.port_score_batch(fam, tb, fields=FIELDS)
    res = []
    for patterns in (1, 0):
        f = device.Family(fam)
        f.set_option("site_patterns", patterns)
        ll, out = f.score(tb, outputs=FIELDS)
        plain = f.score(tb)                      # a later call without the read-out is unaffected
        f.close()
        assert np.array_equal(ll, plain, equal_nan=True)
        g, p = out["site_rate_loglik"], pout["site_rate_loglik"]
        assert np.array_equal(np.isfinite(g), np.isfinite(p))
        ok = np.isfinite(p)
        assert np.max(np.abs(g[ok] - p[ok])) <= 1e-9 * max(1.0, np.max(np.abs(p[ok])))
        assert rel_err(out["xmsa_emission"], exp["xmsa_emission"]) <= 1e-9
        res.append(g)
    assert np.array_equal(res[0], res[1])        # pattern columns or all columns: the same doubles


@pytest.mark.gpu
def test_gpu_rate_logliks_on_liao():
    from conftest import load_liao
    from linearham_b200 import device
    fam, tb, exp, _ = load_liao()
    sub = {k: v[:64] for k, v in tb.items()}
    f = device.Family(fam)
    ll, out = f.score(sub, outputs=FIELDS)
    f.close()
    assert np.max(np.abs(_logmeanexp(out["site_rate_loglik"], 1)[:4] - np.log(exp["xmsa_emission"]))) <= 1e-9
    assert np.max(np.abs(_logmeanexp(out["site_rate_loglik"], 1) - np.log(out["xmsa_emission"]))) <= 1e-9
    assert np.max(np.abs(out["site_rate_loglik"][0] - _independent(fam, sub, 0))) <= 1e-8
