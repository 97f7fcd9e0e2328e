"""k_plan (device traversal planner) against the host planner (plan.cpp): bit-identical programs and
branch-length order, the same errors, and the same log-likelihoods with either planner."""
import tempfile

import numpy as np
import pytest

from conftest import load_golden, load_liao, rel_err, GOLDEN_NAMES
from linearham_b200 import device, workload
from test_plan import plan as host_plan

pytestmark = pytest.mark.gpu


def _device_plan(fam, tb):
    f = device.Family(fam)
    f.set_option("device_plan", 1)
    f.upload(tb)
    steps, pblen = f.fetch_plan()
    tm = f.timings()
    f.close()
    assert tm["device_plan"] == 1
    return steps, pblen


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_device_programs_equal_host_programs(name):
    fam, tb, _, _ = load_golden(name)
    hs, hp, _ = host_plan(fam["n_tips"], tb)
    ds, dp = _device_plan(fam, tb)
    assert np.array_equal(ds, hs) and np.array_equal(dp, hp)


def test_device_programs_on_liao_and_random_shapes():
    fam, tb, _, _ = load_liao()
    sub = {k: v[:256] for k, v in tb.items()}
    hs, hp, _ = host_plan(fam["n_tips"], sub)
    ds, dp = _device_plan(fam, sub)
    assert np.array_equal(ds, hs) and np.array_equal(dp, hp)
    cfg = workload.CONFIGS["c5"]
    with tempfile.TemporaryDirectory() as d:
        fams = workload.repertoire(cfg, d, 6, B=40, n_topo=40)
    for f in fams:   # 40 distinct topologies each, 20-500 leaves
        hs, hp, _ = host_plan(f["fam"]["n_tips"], f["tb"])
        ds, dp = _device_plan(f["fam"], f["tb"])
        assert np.array_equal(ds, hs) and np.array_equal(dp, hp), f["n_leaves"]


@pytest.mark.parametrize("name", ["catch_a", "synth_n40", "synth_n200", "synth_igk_n30"])
def test_results_do_not_depend_on_the_planner(name):
    fam, tb, exp, _ = load_golden(name)
    big = {k: np.concatenate([v] * 50) for k, v in tb.items()}
    out = []
    for dp in (1, 0):
        f = device.Family(fam)
        f.set_option("device_plan", dp)
        a = f.score(big)
        tm = f.timings()
        assert tm["device_plan"] == dp
        f.upload(big)
        f.run()
        b = f.fetch()
        f.close()
        assert np.array_equal(a, b, equal_nan=True)
        out.append(a)
    assert np.array_equal(out[0], out[1], equal_nan=True)
    assert rel_err(out[0][:len(exp["loglik"])], exp["loglik"]) <= 1e-9


def test_device_planner_reports_bad_trees():
    fam, tb, exp, _ = load_golden("synth_n40")
    big = {k: np.concatenate([v] * 20) for k, v in tb.items()}
    f = device.Family(fam)
    f.set_option("device_plan", 1)
    bad = {k: v.copy() for k, v in big.items()}
    bad["children"][70, 5, 0] = 10 ** 6               # out of range (and past the host-planned probe trees)
    with pytest.raises(device.DeviceError, match="tree 70 is not a valid post-order"):
        f.score(bad)
    bad = {k: v.copy() for k, v in big.items()}
    T = fam["n_tips"]
    op = int(np.argmax((bad["children"][71, :, 0] >= T) & (bad["children"][71, :, 1] < T)))
    assert bad["children"][71, op, 0] >= T
    bad["children"][71, op, 1] = bad["children"][71, op, 0]   # an inner node consumed twice
    with pytest.raises(device.DeviceError, match="tree 71 is not a valid post-order"):
        f.score(bad)
    for dp, tree in ((1, 72), (0, 73)):  # a tip referenced twice: device and host planner
        bad = {k: v.copy() for k, v in big.items()}
        tips = np.argwhere(bad["children"][tree] < T)
        (o1, c1), (o2, c2) = tips[0], tips[-1]
        bad["children"][tree, o2, c2] = bad["children"][tree, o1, c1]
        f.set_option("device_plan", dp)
        with pytest.raises(device.DeviceError, match=f"tree {tree} is not a valid post-order"):
            f.score(bad)
    f.set_option("device_plan", 1)
    bad = {k: v.copy() for k, v in big.items()}
    bad["blen"][75, 3, 1] = -0.5
    bad["blen"][79, 3, 1] = np.nan
    with pytest.raises(device.DeviceError, match="tree 75: negative branch length"):
        f.score(bad)
    bad = {k: v.copy() for k, v in big.items()}
    bad["root_child"][77] = 2 * fam["n_tips"] - 3      # the virtual root itself
    with pytest.raises(device.DeviceError, match="tree 77 is not a valid"):
        f.score(bad)
    # the handle is fine afterwards
    assert rel_err(f.score(tb), exp["loglik"]) <= 1e-9
    f.close()


def test_large_family_3000_leaves_matches_the_port():
    """A tree far beyond BASELINE's largest (1000 leaves): CLV stack 12 deep, most levels in the L2 scratch,
    36 KB of planner tables per tree. Emissions and scaler counts must match the port; at this size the
    reference's own forward pass no longer yields a finite log-likelihood, and the class must agree."""
    from linearham_b200 import host
    from oracle import pyoracle as po
    cfg = dict(n_leaves=3000, B=3, R=4, fam_seed=1700, tree_seed=2700, outgroup="naive")
    with tempfile.TemporaryDirectory() as d:
        ff = workload.write_family_files(cfg, d)
        ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
        fam = ph.family()
        ph.close()
        tb, _, _ = workload.tree_batch(cfg, ff, 3)
    assert fam["n_tips"] == 3001
    pll, pout = po.port_score_batch(fam, tb, want_outputs=True)
    hs, hp, _ = host_plan(fam["n_tips"], tb)
    for dp in (1, 0):
        f = device.Family(fam)
        f.set_option("device_plan", dp)
        ll, out = f.score(tb, outputs=True)
        if dp:
            f.upload(tb)
            ds, dpb = f.fetch_plan()
            assert np.array_equal(ds, hs) and np.array_equal(dpb, hp)
        f.close()
        assert np.array_equal(np.isfinite(ll), np.isfinite(pll))
        ok = np.isfinite(pll)
        assert rel_err(ll[ok], pll[ok]) <= 1e-9
        assert rel_err(out["xmsa_emission"], pout["xmsa_emission"]) <= 1e-9
        for k in ("germ_scalers", "vd_scalers", "dj_scalers"):
            assert np.array_equal(out[k], pout[k]), k


def test_schedule_options_do_not_change_results():
    """Opening unit chunk, growing chunks and either planner: a multi-chunk pipelined call gives the doubles of
    the resident batch."""
    from linearham_b200 import host
    cfg = dict(workload.CONFIGS["c3"])
    with tempfile.TemporaryDirectory() as d:
        ff = workload.write_family_files(cfg, d)
        ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
        fam = ph.family()
        ph.close()
        tb, _, _ = workload.tree_batch(cfg, ff, 6000, n_topo=64)
    f = device.Family(fam)
    f.upload(tb)
    f.run()
    want = f.fetch()
    f.close()
    assert np.isfinite(want).mean() > 0.9
    seen = set()
    for open_unit in (0, 1):
        for growth in (0, 1):
            for dp in (0, 1):
                f = device.Family(fam)
                f.set_option("open_unit", open_unit)
                f.set_option("chunk_growth", growth)
                f.set_option("device_plan", dp)
                got = f.score(tb)
                seen.add(f.timings()["kernel_launches"])
                f.close()
                assert np.array_equal(got, want, equal_nan=True), (open_unit, growth, dp)
    assert len(seen) > 1   # the options did change the schedule
