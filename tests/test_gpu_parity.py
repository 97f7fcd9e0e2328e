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
    f.set_option("chunk_trees", 37)  # force many chunks
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


def _score_with(fam, tb, **opts):
    f = device.Family(fam)
    for k, v in opts.items():
        f.set_option(k, v)
    ll, out = f.score(tb, outputs=True)
    tm = f.timings()
    f.close()
    return ll, out, tm


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_site_patterns_are_bit_identical_to_all_columns(name):
    """Pruning the distinct site patterns of rows 1.. and applying the naive row in k_edge gives
    the same doubles as pruning every xMSA column."""
    fam, tb, _, _ = load_golden(name)
    a, oa, ta = _score_with(fam, tb, site_patterns=1)
    b, ob, tb_ = _score_with(fam, tb, site_patterns=0)
    assert tb_["pattern_trees"] == 0 and tb_["prune_columns"] == fam["n_sites"]
    if name.startswith("synth"):
        assert ta["pattern_trees"] > 0 and ta["prune_columns"] < fam["n_sites"]
    assert np.array_equal(a, b)
    assert np.array_equal(oa["xmsa_emission"], ob["xmsa_emission"])


def _rotate_root_to_naive(tb):
    """Moves the evaluation edge onto the naive tip's branch (same unrooted tree): the naive tip,
    a child of the last operation, becomes root_child."""
    out = {k: v.copy() for k, v in tb.items()}
    for t in range(len(out["root_child"])):
        last, lb = out["children"][t, -1], out["blen"][t, -1]
        j = int(np.flatnonzero(last == 0)[0])
        x, lx, ln = last[1 - j], lb[1 - j], lb[j]
        c, lc = out["root_child"][t], out["root_blen"][t]
        out["children"][t, -1] = (c, x)
        out["blen"][t, -1] = (lc, lx)
        out["root_child"][t], out["root_blen"][t] = 0, ln
    return out


@pytest.mark.parametrize("name", ["synth_n12", "synth_n40"])
def test_naive_tip_across_the_root_edge(name):
    fam, tb, _, _ = load_golden(name)
    keep = [t for t in range(len(tb["root_child"])) if 0 in tb["children"][t, -1]]
    tb = {k: v[keep] for k, v in tb.items()}
    rot = _rotate_root_to_naive(tb)
    pll, pout = po.port_score_batch(fam, rot)
    a, oa, ta = _score_with(fam, rot, site_patterns=1)
    b, ob, _ = _score_with(fam, rot, site_patterns=0)
    assert ta["pattern_trees"] == len(keep)
    assert np.array_equal(a, b) and np.array_equal(oa["xmsa_emission"], ob["xmsa_emission"])
    assert rel_err(a, pll) <= TOL and rel_err(oa["xmsa_emission"], pout["xmsa_emission"]) <= TOL


def test_naive_tip_deep_in_the_tree_falls_back_to_all_columns():
    """With the naive row swapped onto a tip that is not next to the virtual root, no column can be
    shared: the batch runs on all xMSA columns and still matches the oracle."""
    fam, tb, _, _ = load_golden("synth_n40")
    near_root = set(tb["children"][:, -1].ravel()) | set(tb["root_child"])
    r = next(i for i in range(1, fam["n_tips"]) if i not in near_root)
    fam2 = dict(fam)
    codes = fam["tip_codes"].copy()
    codes[[0, r]] = codes[[r, 0]]
    fam2["tip_codes"] = codes
    relabel = {k: v.copy() for k, v in tb.items()}
    for key in ("children", "root_child"):
        a = relabel[key]
        is0, isr = a == 0, a == r
        a[is0], a[isr] = r, 0
    pll, pout = po.port_score_batch(fam2, relabel)
    ll, out, tm = _score_with(fam2, relabel)
    assert tm["pattern_trees"] == 0 and tm["prune_columns"] == fam["n_sites"]
    assert rel_err(ll, pll) <= TOL and rel_err(out["xmsa_emission"], pout["xmsa_emission"]) <= TOL


def test_pipelined_call_matches_resident_batch():
    """lh_score_batch (segments planned and copied while earlier chunks run) and the split
    upload / run / fetch form give the same doubles, across short opening chunks and a ragged tail."""
    fam, tb, _, _ = load_golden("synth_n12")
    rep = 397
    big = {k: np.concatenate([v] * rep) for k, v in tb.items()}
    f = device.Family(fam)
    f.set_option("chunk_trees", 100)
    a = f.score(big)
    n_launch = f.timings()["prune_launches"]
    f.upload(big)
    f.run()
    b = f.fetch()
    f.close()
    assert n_launch >= 3
    assert np.array_equal(a, b)
    assert np.array_equal(a, np.tile(a[:len(tb["root_child"])], rep))


def test_mixed_pattern_batch_and_chunks():
    """Trees with the naive tip deep in the tree run on all columns next to trees pruned on site
    patterns; many chunks; pipelined and resident forms agree."""
    fam, tb, _, _ = load_golden("synth_n12")
    rep = 41
    big = {k: np.concatenate([v] * rep) for k, v in tb.items()}
    f = device.Family(fam)
    f.set_option("chunk_trees", 29)
    a = f.score(big)
    tm = f.timings()
    f.upload(big)
    f.run()
    b = f.fetch()
    f.set_option("site_patterns", 0)
    c = f.score(big)
    f.close()
    assert 0 < tm["pattern_trees"] < len(a)
    assert np.array_equal(a, b) and np.array_equal(a, c)


@pytest.mark.parametrize("slots,lo", [(1, 0), (1, 1), (1, 2), (2, 1), (2, 3)])
def test_shared_memory_slot_window_does_not_change_results(slots, lo):
    """Which CLV stack levels sit in shared memory (the rest in the L2 scratch) is a placement choice."""
    fam, tb, exp, _ = load_golden("synth_n40")
    a, oa, ta = _score_with(fam, tb, smem_slots=slots, smem_lo=lo, site_patterns=0)
    b, ob, _ = _score_with(fam, tb, site_patterns=0)
    assert np.array_equal(a, b) and np.array_equal(oa["xmsa_emission"], ob["xmsa_emission"])
    assert rel_err(a, exp["loglik"]) <= TOL


@pytest.mark.parametrize("n_rates", [1, 2, 3, 8])
def test_rate_category_counts_match_the_port(n_rates):
    """1..LH_MAX_RATES discrete-Gamma categories (the goldens hold R = 1 and 4)."""
    from linearham_b200 import host
    fam, tb, _, meta = load_golden("synth_n40")
    tb = dict(tb)
    tb["rates"] = host.gamma_cats(meta["alpha"], n_rates)
    pll, pout = po.port_score_batch(fam, tb)
    for sp in (1, 0):
        ll, out, _ = _score_with(fam, tb, site_patterns=sp)
        assert rel_err(ll, pll) <= TOL and rel_err(out["xmsa_emission"], pout["xmsa_emission"]) <= TOL
    with pytest.raises(device.DeviceError):
        bad = dict(tb)
        bad["rates"] = np.ones((len(tb["root_child"]), 9))
        device.Family(fam).score(bad)
