import glob
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.dirname(os.path.abspath(__file__)))

GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")
# sample bundles (family + tree batch + reference outputs); family_*.npz hold a family description only
GOLDEN_NAMES = sorted(n for n in (os.path.splitext(os.path.basename(p))[0]
                                  for p in glob.glob(os.path.join(GOLDEN_DIR, "*.npz")))
                      if not n.startswith(("family_", "pipeline_", "liao_", "scaling_", "fixture_")))
# underflow-regime families scored by the reference (scripts/make_scaling_goldens.py)
SCALING_NAMES = sorted(os.path.splitext(os.path.basename(p))[0] for p in glob.glob(os.path.join(GOLDEN_DIR, "scaling_*.npz")))


def compare_in_underflow_regime(got_ll, got, exp, tol=1e-9):
    """xmsa_emission / log-likelihood: within tol where the reference's value is finite and non-zero, the same
    class (zero, +-inf, nan) where it is not; every scaler count exact for the samples whose reference
    log-likelihood is finite (past a NaN the reference's own counts are the residue of comparisons with NaN)."""
    xe, rxe = np.asarray(got["xmsa_emission"]), np.asarray(exp["xmsa_emission"])
    ok = np.isfinite(rxe) & (np.abs(rxe) >= np.finfo(np.float64).tiny)   # (a subnormal has no 1e-9 to offer)
    sub = np.isfinite(rxe) & (rxe != 0) & ~ok
    assert np.array_equal(np.isnan(xe), np.isnan(rxe)) and np.array_equal(xe == 0, rxe == 0)
    assert np.all(np.abs(xe[sub]) < np.finfo(np.float64).tiny)
    assert np.array_equal(np.isinf(xe), np.isinf(rxe))
    worst = float(np.max(np.abs(xe[ok] - rxe[ok]) / np.abs(rxe[ok]))) if ok.any() else 0.0
    assert worst <= tol, worst
    rll = np.asarray(exp["loglik"])
    fin = np.isfinite(rll)
    assert np.array_equal(np.isnan(got_ll), np.isnan(rll)) and np.array_equal(np.isinf(got_ll), np.isinf(rll))
    if fin.any():
        assert float(np.max(np.abs(got_ll[fin] - rll[fin]) / np.abs(rll[fin]))) <= tol
        for k in ("germ_scalers", "vd_scalers", "dj_scalers"):
            assert np.array_equal(np.asarray(got[k])[fin], np.asarray(exp[k])[fin]), k
    return worst


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"), allow_pickle=False)
    fam = {k[4:]: (int(z[k]) if z[k].ndim == 0 else z[k]) for k in z.files if k.startswith("fam_")}
    tb = {k[3:]: z[k] for k in z.files if k.startswith("tb_")}
    exp = {k[4:]: z[k] for k in z.files if k.startswith("exp_")}
    meta = {k: z[k] for k in ("alpha", "newicks", "labels", "xmsa_seqs")}
    return fam, tb, exp, meta


def load_family_bundle(name):
    """tests/golden/family_<name>.npz -> (family dict, labels, sha of the generating files)."""
    z = np.load(os.path.join(GOLDEN_DIR, f"family_{name}.npz"), allow_pickle=False)
    fam = {k[4:]: (int(z[k]) if z[k].ndim == 0 else z[k]) for k in z.files if k.startswith("fam_")}
    return fam, [str(s) for s in z["labels"]], str(z["sha"])


def load_liao():
    """BASELINE config 2 bundle -> (family, tree batch of 1024 samples, reference outputs, bundle)."""
    from linearham_b200 import workload
    z = np.load(os.path.join(GOLDEN_DIR, "liao_b1024.npz"), allow_pickle=False)
    fam = {k[4:]: (int(z[k]) if z[k].ndim == 0 else z[k]) for k in z.files if k.startswith("fam_")}
    exp = {k[4:]: z[k] for k in z.files if k.startswith("exp_")}
    return fam, workload.liao_batch(z), exp, z


@pytest.fixture(params=GOLDEN_NAMES)
def golden(request):
    return (request.param,) + load_golden(request.param)


def rel_err(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    denom = np.maximum(np.abs(b), 1e-300)
    return float(np.max(np.abs(a - b) / denom)) if a.size else 0.0
