"""BASELINE config 2: the Liao clonal family (data/liao_dataset.fasta, 312 sequences x 363 nt; real partis
germline/HMM parameters; stand-in annotation and tree samples, scripts/make_liao_bundle.py) scored over a
batch of 1024 tree samples. Expected values come from the compiled reference."""
import numpy as np
import pytest

from conftest import load_liao, rel_err

TOL = 1e-9


def test_bundle_shape():
    fam, tb, exp, z = load_liao()
    assert fam["n_tips"] == 313 and fam["has_d"] == 1 and (fam["n_v"], fam["n_d"], fam["n_j"]) == (3, 5, 2)
    assert tb["children"].shape == (1024, 311, 2) and tb["rates"].shape == (1024, 4)
    assert len(str(z["naive"])) == 363 and np.isfinite(exp["loglik"]).all()
    # every sample is its own (topology, noise vector) pair
    assert len({(int(a), int(b)) for a, b in zip(z["topo_of"], z["noise_of"])}) == 1024


def test_port_matches_the_reference_on_liao():
    from oracle import pyoracle as po
    fam, tb, exp, _ = load_liao()
    sub = {k: v[:6] for k, v in tb.items()}
    ll, out = po.port_score_batch(fam, sub, want_outputs=True)
    assert rel_err(ll, exp["loglik"][:6]) <= TOL
    assert rel_err(out["xmsa_emission"][:4], exp["xmsa_emission"]) <= TOL
    assert np.array_equal(out["germ_scalers"][:4], exp["germ_scalers"])


@pytest.mark.gpu
def test_gpu_scores_all_1024_liao_samples():
    from linearham_b200 import device
    fam, tb, exp, _ = load_liao()
    f = device.Family(fam)
    ll = f.score(tb)
    tm = f.timings()
    assert tm["kernel_launches"] >= 4 and tm["pattern_trees"] == 1024  # naive under the trifurcation
    assert rel_err(ll, exp["loglik"]) <= TOL
    sub = {k: v[:4] for k, v in tb.items()}
    l4, out = f.score(sub, outputs=True)
    f.close()
    assert np.array_equal(l4, ll[:4])
    for k in ("xmsa_emission", "vgerm_forward", "vd_forward", "dgerm_forward", "dj_forward", "jgerm_forward"):
        assert rel_err(out[k], exp[k]) <= TOL, k
    for k in ("germ_scalers", "vd_scalers", "dj_scalers"):
        assert np.array_equal(out[k], exp[k]), k


@pytest.mark.gpu
def test_liao_on_all_columns_and_in_a_repertoire():
    from linearham_b200 import device
    from conftest import load_golden
    fam, tb, exp, _ = load_liao()
    sub = {k: v[:96] for k, v in tb.items()}
    f = device.Family(fam)
    a = f.score(sub)
    f.set_option("site_patterns", 0)
    b = f.score(sub)
    f.close()
    assert np.array_equal(a, b) and rel_err(a, exp["loglik"][:96]) <= TOL
    other = load_golden("synth_n40")
    rep = device.Repertoire(n_slots=2)
    lls, _ = rep.score([(other[0], other[1]), (fam, sub)])
    rep.close()
    assert np.array_equal(lls[1], a)
