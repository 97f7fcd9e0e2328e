"""Host-side product code (C++ behind include/linearham_b200_host.h) against the checkers:
Gamma categories bit-exact vs the oracle port (itself bit-exact vs libpll, test_oracle_port.py),
Newick ingest identical to the independent pure-Python flattener in oracle/pytrees.py and to the
flattened batches stored in the golden bundles (made from the same Newick strings)."""
import numpy as np
import pytest

from conftest import load_golden, GOLDEN_NAMES
from linearham_b200 import host, synth
from oracle import pyoracle as po, pytrees as pt


def test_gamma_bitexact_vs_port():
    rng = np.random.default_rng(3)
    alphas = np.concatenate([[0.02, 0.05, 0.1, 0.16, 0.32, 0.5, 1.0, 2.0, 10.0, 50.0, 200.0],
                             np.maximum(0.02, rng.exponential(1.0, 500))])
    for n in (1, 2, 4, 5, 8):
        got = host.gamma_cats(alphas, n)
        for a, g in zip(alphas, got):
            assert np.array_equal(g, po.port_gamma_cats(a, n)), (a, n)
    with pytest.raises(host.HostError):
        host.gamma_cats(0.01, 4)  # gamma.c:238


@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_newick_ingest_matches_golden_batches(name):
    fam, tb, exp, meta = load_golden(name)
    got = host.newick_to_ops(list(meta["newicks"]), list(meta["labels"]))
    for k in ("children", "blen", "root_child", "root_blen"):
        assert np.array_equal(got[k], tb[k]), k
    full = host.tree_batch(list(meta["newicks"]), list(meta["labels"]), tb["er"], tb["pi"], meta["alpha"],
                           tb["rates"].shape[1])
    assert np.array_equal(full["rates"], tb["rates"])  # bundle rates came from libpll itself


def test_newick_ingest_random_trees_and_threads():
    rng = np.random.default_rng(9)
    tree = synth.random_tree(rng, 57)
    labels = ["naive"] + [str(i) for i in range(57)]
    samples = synth.tree_samples(rng, tree, 300)
    nw = [s["newick"] for s in samples]
    a = host.newick_to_ops(nw, labels, n_threads=1)
    b = host.newick_to_ops(nw, labels, n_threads=4)
    ref = pt.flatten_batch(nw, labels, np.ones((300, 6)), np.ones((300, 4)), np.ones((300, 1)))
    for k in ("children", "blen", "root_child", "root_blen"):
        assert np.array_equal(a[k], ref[k]), k
        assert np.array_equal(a[k], b[k]), k


def test_newick_edge_cases():
    labels = ["naive", "0", "1", "2"]
    # missing and zero lengths -> EPS (src/PhyloHMM.cpp:422); inner labels and comments ignored
    got = host.newick_to_ops(["((0:0.03,1)x[&index=5]:0.0 , naive:0.3,2[&index=1]);"], labels)
    assert got["children"].tolist() == [[[1, 2], [0, 3]]]
    assert got["blen"].tolist() == [[[0.03, 1e-6], [0.3, 1e-6]]]
    assert got["root_child"].tolist() == [4] and got["root_blen"].tolist() == [1e-6]
    for bad in ["(0,1,2);",                 # missing tip
                "((0,1),naive,2,3);",       # unknown label / 4 children
                "((0,1),(naive,2));",       # not a trifurcation
                "((0,1,2),naive,2);",       # non-binary inner node
                "((0,1),naive,1);",         # duplicate tip
                "((0:x,1),naive,2);",       # malformed length
                "((0,1),naive,2)"]:         # no ';'
        with pytest.raises(host.HostError):
            host.newick_to_ops([bad], labels)
