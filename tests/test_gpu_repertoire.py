"""BASELINE config 5 (repertoire sweep): many independent clonal families on one device.

lh_family_rebind and lh_repertoire_* against (i) the reference's golden outputs, (ii) the oracle port
on the same inputs, (iii) a fresh handle per family - the last bit for bit: the result of a family must
not depend on what the handle scored before, on the slot, on the order or on the number of slots."""
import tempfile

import numpy as np
import pytest

from conftest import load_golden, rel_err, GOLDEN_NAMES
from linearham_b200 import device, workload
from oracle import pyoracle as po

pytestmark = pytest.mark.gpu
TOL = 1e-9


def _same(a, b):
    return np.array_equal(a, b, equal_nan=True)


def test_rebind_gives_the_results_of_a_fresh_handle():
    # heavy (7 regions) -> light chain (5 regions) -> a larger family -> the 4-tip Catch fixture -> back
    names = ["synth_n40", "synth_igk_n30", "synth_n200", "catch_c", "synth_n24_r1", "synth_n40"]
    fresh = {}
    for n in set(names):
        fam, tb, exp, _ = load_golden(n)
        f = device.Family(fam)
        fresh[n] = f.score(tb, outputs=True)
        f.close()
    f = None
    for n in names:
        fam, tb, exp, _ = load_golden(n)
        if f is None:
            f = device.Family(fam)
        else:
            f.rebind(fam)
        ll, out = f.score(tb, outputs=True)
        assert _same(ll, fresh[n][0]), n
        for k in out:
            assert _same(out[k], fresh[n][1][k]), (n, k)
        assert rel_err(ll, exp["loglik"]) <= TOL, n
        assert rel_err(out["xmsa_emission"], exp["xmsa_emission"]) <= TOL, n
        assert np.array_equal(out["germ_scalers"], exp["germ_scalers"]), n
    f.close()


def test_failed_rebind_leaves_no_family_bound():
    fam, tb, exp, _ = load_golden("synth_n12")
    f = device.Family(fam)
    bad = dict(fam)
    bad["tip_codes"] = fam["tip_codes"].copy()
    bad["tip_codes"][1, 3] = 9
    with pytest.raises(device.DeviceError, match="tip_codes"):
        f.rebind(bad)
    with pytest.raises(device.DeviceError, match="no family bound"):
        f.score(tb)
    f.rebind(fam)
    assert rel_err(f.score(tb), exp["loglik"]) <= TOL
    f.close()


@pytest.mark.parametrize("n_slots", [1, 3])
def test_repertoire_of_golden_families_matches_the_reference(n_slots):
    fams = [load_golden(n) for n in GOLDEN_NAMES]
    rep = device.Repertoire(n_slots=n_slots)
    lls, st = rep.score([(fam, tb) for fam, tb, _, _ in fams])
    rep.close()
    assert st["n_failed"] == 0 and st["n_families"] == len(fams) and all(s == 0 for s in st["status"])
    assert st["n_trees"] == sum(tb["root_child"].shape[0] for _, tb, _, _ in fams)
    assert st["kernel_launches"] >= 4 * len(fams)
    for name, ll, (fam, tb, exp, _) in zip(GOLDEN_NAMES, lls, fams):
        assert rel_err(ll, exp["loglik"]) <= TOL, name


def test_c5_families_match_the_port_and_do_not_depend_on_slots():
    cfg = workload.CONFIGS["c5"]
    with tempfile.TemporaryDirectory() as d:
        fams = workload.repertoire(cfg, d, 10, B=24, n_topo=8)
    assert min(f["n_leaves"] for f in fams) < 40 and max(f["n_leaves"] for f in fams) > 400
    pairs = [(f["fam"], f["tb"]) for f in fams]
    r1 = device.Repertoire(n_slots=1)
    a, st1 = r1.score(pairs)
    a2, _ = r1.score(list(reversed(pairs)))  # the handles now hold grown buffers and another family
    r1.close()
    r3 = device.Repertoire(n_slots=3, host_threads=2)
    b, st3 = r3.score(pairs)
    r3.close()
    assert st3["n_slots"] == 3 and st3["host_threads_per_slot"] == 2
    for i, (fam, tb) in enumerate(pairs):
        assert _same(a[i], b[i]) and _same(a[i], a2[len(pairs) - 1 - i]), i
        h = device.Family(fam)
        assert _same(h.score(tb), a[i]), i
        h.close()
        port, _ = po.port_score_batch(fam, tb, want_outputs=False)
        ok = np.isfinite(port)
        assert np.array_equal(np.isfinite(a[i]), ok), i
        assert rel_err(a[i][ok], port[ok]) <= TOL, i


def test_a_bad_family_fails_alone():
    good = load_golden("synth_n12")
    other = load_golden("synth_n40")
    bad_tb = {k: v.copy() for k, v in good[1].items()}
    bad_tb["blen"][0, 0, 0] = -1.0
    rep = device.Repertoire(n_slots=2)
    with pytest.raises(device.DeviceError, match="family 1: .*negative branch length") as ei:
        rep.score([(good[0], good[1]), (good[0], bad_tb), (other[0], other[1])])
    assert ei.value.stats["status"] == [0, 1, 0] and ei.value.stats["n_failed"] == 1
    # the repertoire stays usable
    lls, st = rep.score([(good[0], good[1]), (other[0], other[1])])
    assert rel_err(lls[0], good[2]["loglik"]) <= TOL and rel_err(lls[1], other[2]["loglik"]) <= TOL
    lls, st = rep.score([])
    assert lls == [] and st["n_families"] == 0
    rep.close()


def test_many_slots_stress_is_bit_identical_to_fresh_handles():
    """Six slots rebinding 48 families twice over, the copy engine kept busy by all of them: every family must
    come out as from a fresh handle. (Regression: the re-tiled tip codes of a rebound handle once went up with a
    synchronous cudaMemcpy from pageable memory, whose DMA is not ordered with the handle's non-blocking
    streams - under load k_prune could start on the previous family's codes.)"""
    cfg = workload.CONFIGS["c5"]
    with tempfile.TemporaryDirectory() as d:
        fams = workload.repertoire(cfg, d, 48, B=400, n_topo=8)
    pairs = [(f["fam"], f["tb"]) for f in fams]
    want = []
    for fam, tb in pairs:
        h = device.Family(fam)
        want.append(h.score(tb))
        h.close()
    rep = device.Repertoire(n_slots=6)
    for _ in range(2):
        got, st = rep.score(pairs)
        assert st["n_failed"] == 0
        for i, (a, b) in enumerate(zip(got, want)):
            assert _same(a, b), (i, fams[i]["n_leaves"])
    rep.close()
    port, _ = po.port_score_batch(pairs[0][0], {k: v[:8] for k, v in pairs[0][1].items()}, want_outputs=False)
    assert rel_err(want[0][:8], port) <= TOL
