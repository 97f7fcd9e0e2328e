"""lh_group_t: one family on several GPUs inside one process. The doubles must not depend on
the number of devices: every result of a group is np.array_equal to the single-device handle's. The two-device
cases need a box with two GPUs (gpurun --gpus 2); on one GPU they are skipped and the one-device group runs."""
import os
import subprocess

import numpy as np
import pytest

from conftest import load_golden, rel_err
from linearham_b200 import device, host

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_parse_devices():
    assert device.parse_devices("0-3") == [0, 1, 2, 3]
    assert device.parse_devices("0,2,5") == [0, 2, 5]
    assert device.parse_devices("1") == [1]
    assert device.parse_devices("0-1,4-5") == [0, 1, 4, 5]
    for bad in ("", "a", "3-1", "0,0", "1-", "0;1"):
        with pytest.raises(device.DeviceError):
            device.parse_devices(bad)


def _devices(n):
    if device.device_count() < n:
        pytest.skip(f"needs {n} GPUs")
    return list(range(n))


@pytest.mark.gpu
@pytest.mark.parametrize("n_dev", [1, 2])
def test_group_equals_single_device(n_dev):
    fam, tb, exp, meta = load_golden("synth_n40")
    rep = 150
    big = {k: np.concatenate([v] * rep) for k, v in tb.items()}
    big = {k: v[:-7] for k, v in big.items()}   # a ragged last shard
    f = device.Family(fam)
    want, want_out = f.score(big, outputs=True)
    f.close()
    g = device.Group(fam, _devices(n_dev))
    assert g.size() == n_dev
    got, out = g.score(big, outputs=True)
    tm = g.timings()
    assert tm["n_trees"] == len(want)
    assert np.array_equal(got, want, equal_nan=True)
    for k in want_out:
        assert np.array_equal(out[k], want_out[k], equal_nan=True), k
    # Newick text in
    nw, labels = [str(s) for s in meta["newicks"]] * rep, [str(s) for s in meta["labels"]]
    nw = nw[:-7]
    g.set_labels(labels)
    got_text = g.score_newick(nw, big["er"], big["pi"], big["rates"].shape[1], rates=big["rates"])
    assert np.array_equal(got_text, want, equal_nan=True)
    # fewer trees than devices, and none
    small = {k: v[:1] for k, v in tb.items()}
    assert np.array_equal(g.score(small), want[:1], equal_nan=True)
    assert len(g.score({k: v[:0] for k, v in tb.items()})) == 0
    # an error names the tree of the caller's batch
    bad = {k: v.copy() for k, v in big.items()}
    t_bad = len(want) - 3
    bad["blen"][t_bad, 0, 0] = -1.0
    with pytest.raises(device.DeviceError, match=f"tree {t_bad}: negative branch length"):
        g.score(bad)
    g.close()
    assert rel_err(want[:len(exp["loglik"])], exp["loglik"]) <= 1e-9


@pytest.mark.gpu
@pytest.mark.parametrize("n_dev", [1, 2])
def test_phylohmm_and_cli_over_devices(n_dev, tmp_path):
    """PhyloHMM::set_devices / `linearham --pipeline --devices`: the output TSV does not depend on the devices."""
    devs = _devices(n_dev)
    from linearham_b200 import synth
    d = str(tmp_path)
    genes = synth.make_germline_set(os.path.join(d, "hmm"), seed=7)
    ff = synth.make_family(d, genes, 24, seed=1500, name="family")
    samples = synth.tree_samples(np.random.default_rng(3), ff["tree"], 37, outgroup="naive")
    tsv = os.path.join(d, "in.tsv")
    synth.write_tsv(tsv, samples)
    outs = []
    for spec in (None, ",".join(str(x) for x in devs)):
        out = os.path.join(d, f"out_{spec}.tsv")
        cmd = [os.path.join(ROOT, "linearham_b200", "linearham"), "--pipeline", "--yaml-path", ff["yaml_path"], "--cluster-ind", "0",
               "--hmm-param-dir", os.path.join(d, "hmm"), "--seed", "5", "--num-rates", "4", "--input-path", tsv,
               "--output-path", out, "--batch-size", "16"]
        env = dict(os.environ)
        if spec is not None and n_dev == 2:
            env["LINEARHAM_DEVICES"] = spec      # the variable ...
        elif spec is not None:
            cmd += ["--devices", spec]           # ... or the flag
        subprocess.run(cmd, check=True, env=env, capture_output=True)
        outs.append(open(out).read())
    assert outs[0] == outs[1] and outs[0].count("\n") == 38
    ph = host.PhyloHMM(ff["yaml_path"], 0, os.path.join(d, "hmm"), 0)
    ph.set_devices(devs)
    er, pi, al = [s["er"] for s in samples], [s["pi"] for s in samples], [s["alpha"] for s in samples]
    ll = ph.ScoreSamples([s["newick"] for s in samples], er, pi, al, 4)
    ph.close()
    ph1 = host.PhyloHMM(ff["yaml_path"], 0, os.path.join(d, "hmm"), 0)
    assert np.array_equal(ll, ph1.ScoreSamples([s["newick"] for s in samples], er, pi, al, 4), equal_nan=True)
    ph1.close()


@pytest.mark.gpu
@pytest.mark.parametrize("n_dev", [1, 2])
def test_repertoire_over_devices(n_dev):
    devs = _devices(n_dev)
    fams = []
    for name in ("synth_n12", "synth_n40", "synth_n24_r1", "synth_igk_n30", "synth_n12", "synth_n40"):
        fam, tb, exp, _ = load_golden(name)
        fams.append((fam, {k: np.concatenate([v] * 20) for k, v in tb.items()}, np.tile(exp["loglik"], 20)))
    rep = device.Repertoire(n_slots=2, devices=devs)
    lls, st = rep.score([(f, t) for f, t, _ in fams])
    rep.close()
    assert st["n_slots"] == 2 * n_dev and st["n_failed"] == 0
    for ll, (_, _, want) in zip(lls, fams):
        assert rel_err(ll, want) <= 1e-9
    one = device.Repertoire(n_slots=1)
    lls1, _ = one.score([(f, t) for f, t, _ in fams])
    one.close()
    for a, b in zip(lls, lls1):
        assert np.array_equal(a, b, equal_nan=True)
