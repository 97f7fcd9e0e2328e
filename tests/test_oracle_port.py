"""Pins the oracle: the CPU restatement (oracle/port) must reproduce the outputs the
reference's own C++ produced for the golden bundles (tests/golden/make_golden.py), and the
published Catch constants (test/test.cpp:968,1291,1398)."""
import numpy as np
import pytest

from conftest import load_golden, rel_err
from oracle import pyoracle as po

TOL = 1e-12  # port vs compiled reference, relative (both FP64, same operation order)


def test_port_matches_reference_outputs(golden):
    name, fam, tb, exp, meta = golden
    ll, out = po.port_score_batch(fam, tb)
    assert rel_err(ll, exp["loglik"]) <= TOL
    assert rel_err(out["xmsa_emission"], exp["xmsa_emission"]) <= TOL
    assert np.array_equal(out["germ_scalers"], exp["germ_scalers"])
    assert np.array_equal(out["vd_scalers"], exp["vd_scalers"])
    for k in ("vgerm_forward", "vd_forward", "jgerm_forward"):
        assert rel_err(out[k], exp[k]) <= 1e-11, k
    if fam["has_d"]:
        assert np.array_equal(out["dj_scalers"], exp["dj_scalers"])
        for k in ("dgerm_forward", "dj_forward"):
            assert rel_err(out[k], exp[k]) <= 1e-11, k


@pytest.mark.parametrize("name,value", [("catch_a", -75.8136), ("catch_b", -75.1122515055), ("catch_c", -55.73483)])
def test_catch_known_answers(name, value):
    fam, tb, exp, meta = load_golden(name)
    ll, _ = po.port_score_batch(fam, {k: v[:1] for k, v in tb.items()}, want_outputs=False)
    assert ll[0] == pytest.approx(value, rel=1.2e-5)          # Catch Approx default epsilon
    assert exp["loglik"][0] == pytest.approx(value, rel=1.2e-5)


def test_catch_a_xmsa_exact():
    """xMSA of fixture A, test/test.cpp:997-1002 (rows: naive, 0, 1, 2; A,C,G,T,N = 0..4)."""
    fam, _, _, meta = load_golden("catch_a")
    assert fam["tip_codes"].shape == (4, 36)
    seqs = ["".join("ACGTN"[c] for c in row) for row in fam["tip_codes"]]
    assert seqs == list(meta["xmsa_seqs"])
    assert list(meta["labels"]) == ["naive", "0", "1", "2"]


def test_gamma_rates_match_reference_values():
    # alpha = 1, 4 categories: values printed by the reference build
    r = po.port_gamma_cats(1.0, 4)
    ref = [0.13695378267140107, 0.47675185617665189, 0.99999999997958422, 2.3862943611723626]
    assert np.allclose(r, ref, rtol=1e-14, atol=0)
    assert po.port_gamma_cats(0.5, 1)[0] == 1.0
    with pytest.raises(ValueError):
        po.port_gamma_cats(0.01, 4)  # gamma.c:238 rejects alpha < 0.02


@pytest.mark.skipif(not po.have_ref(), reason="oracle/_ref not built")
def test_gamma_rates_bitexact_vs_reference_sweep():
    ref = po.RefPhyloHMM.__new__(po.RefPhyloHMM)
    import ctypes as C
    lib = C.CDLL(po.REF_LIB)
    lib.ref_gamma_cats.argtypes = [C.c_double, C.c_int, C.POINTER(C.c_double)]
    rng = np.random.default_rng(5)
    alphas = np.concatenate([[0.02, 0.05, 0.1, 0.16, 0.32, 0.5, 1.0, 2.0, 10.0, 50.0, 200.0],
                             np.maximum(0.02, rng.exponential(1.0, 300))])
    for a in alphas:
        for n in (2, 4, 5, 8):
            out = np.zeros(n)
            assert lib.ref_gamma_cats(float(a), n, out.ctypes.data_as(C.POINTER(C.c_double))) == 0
            assert np.array_equal(out, po.port_gamma_cats(a, n)), (a, n)
