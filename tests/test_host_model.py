"""The product's host-side family construction (C++ linearham::PhyloHMM of this repo: YAML -> state
space -> transition matrices -> xMSA) must be BIT-EXACT against the reference's own C++:
  * against the committed family bundles the compiled reference produced for the synthetic
    bench families (tests/golden/family_*.npz) - runs anywhere;
  * member by member against the compiled reference (oracle/_ref) on the reference's Catch
    fixtures and on synthetic families - runs where /root/reference exists.
Also pins the literal xMSA of the Catch fixture (test/test.cpp:997-1002)."""
import os
import tempfile

import numpy as np
import pytest

from conftest import load_family_bundle, load_golden
from linearham_b200 import host, workload
from oracle import pyoracle as po

REFDATA = "/root/reference/data/"
have_ref_data = os.path.isdir(REFDATA) and po.have_ref()

I32_FIELDS = ["msa", "xmsa", "xmsa_naive_ind", "vd_junction_xmsa_inds", "dj_junction_xmsa_inds"] + \
    [f"{r}_{f}" for r in ("vpadding", "vgerm", "dgerm", "jgerm", "jpadding") for f in ("xmsa_inds", "naive_bases", "site_inds", "ggene_ranges")] + \
    [f"{r}_{f}" for r in ("vgerm", "dgerm", "jgerm") for f in ("germ_inds", "left_del", "right_del")] + \
    [f"{r}_{f}" for r in ("vd_junction", "dj_junction") for f in ("naive_bases", "germ_inds", "site_inds", "del", "ggene_ranges")]
F64_FIELDS = ["vpadding_transition", "vgerm_vd_junction_transition", "vd_junction_transition", "vd_junction_dgerm_transition",
              "dgerm_dj_junction_transition", "dj_junction_transition", "dj_junction_jgerm_transition", "jpadding_transition",
              "vgerm_gene_prob", "vgerm_trans_prod"]
STR_FIELDS = ["xmsa_seqs", "xmsa_labels", "locus", "alphabet", "vgerm_state_strs", "dgerm_state_strs", "jgerm_state_strs",
              "vd_junction_state_strs", "dj_junction_state_strs"]


def same(a, b):
    """exact equality; Eigen column vectors (n x 1) and plain vectors (1 x n) count as the same shape"""
    a, b = np.asarray(a), np.asarray(b)
    if 1 in a.shape or 1 in b.shape or a.size == 0:
        return a.size == b.size and np.array_equal(a.reshape(-1), b.reshape(-1))
    return np.array_equal(a, b)


def assert_family_equal(got, exp):
    for k, v in exp.items():
        g = got[k]
        if isinstance(v, (int, np.integer)):
            assert int(g) == int(v), k
        else:
            v = np.asarray(v)
            assert np.array_equal(np.asarray(g).reshape(-1), v.reshape(-1)), k   # exact, doubles included


@pytest.mark.parametrize("name", ["tiny", "c3", "c4"])
def test_family_bundles_bitexact(name):
    exp, labels, sha = load_family_bundle(name)
    with tempfile.TemporaryDirectory() as d:
        ff = workload.write_family_files(workload.CONFIGS[name], d)
        assert ff["sha"] == sha
        h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
        assert h.strs("xmsa_labels") == labels
        assert_family_equal(h.family(), exp)
        h.close()


@pytest.mark.skipif(not have_ref_data, reason="needs /root/reference and oracle/_ref")
@pytest.mark.parametrize("yaml,params", [("phylo_hmm_input.yaml", "hmm_params"), ("phylo_hmm_input_extra.yaml", "hmm_params"),
                                         ("phylo_likelihood_hmm_input.yaml", "phylo_likelihood_hmm_params")])
def test_catch_fixtures_member_by_member(yaml, params):
    ref = po.RefPhyloHMM(REFDATA + yaml, 0, REFDATA + params, 0)
    h = host.PhyloHMM(REFDATA + yaml, 0, REFDATA + params, 0)
    for f in I32_FIELDS:
        assert same(h.i32(f), ref.i32(f)), f
    for f in F64_FIELDS:
        assert same(h.f64(f), ref.f64(f)), f
    for f in STR_FIELDS:
        assert h.strs(f) == ref.strs(f), f
    assert_family_equal(h.family(), ref.family())


@pytest.mark.skipif(not have_ref_data, reason="needs /root/reference and oracle/_ref")
@pytest.mark.parametrize("n_leaves,seed", [(5, 31), (17, 32), (60, 33)])
def test_synthetic_families_member_by_member(n_leaves, seed):
    cfg = dict(workload.CONFIGS["tiny"], n_leaves=n_leaves, fam_seed=seed)
    with tempfile.TemporaryDirectory() as d:
        ff = workload.write_family_files(cfg, d)
        ref = po.RefPhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
        h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
        for f in I32_FIELDS:
            assert same(h.i32(f), ref.i32(f)), f
        for f in F64_FIELDS:
            assert same(h.f64(f), ref.f64(f)), f
        for f in STR_FIELDS:
            assert h.strs(f) == ref.strs(f), f


@pytest.mark.skipif(not os.path.isdir(REFDATA), reason="needs /root/reference/data")
def test_catch_xmsa_literal():
    """xMSA of fixture A as written in the reference's test (test/test.cpp:997-1002)."""
    h = host.PhyloHMM(REFDATA + "phylo_hmm_input.yaml", 0, REFDATA + "hmm_params", 0)
    fam, _, _, meta = load_golden("catch_a")
    assert np.array_equal(h.i32("xmsa"), fam["tip_codes"].astype(np.int32))
    assert h.strs("xmsa_seqs") == list(meta["xmsa_seqs"])
    assert h.i32("xmsa").shape == (4, 36)


def test_constructor_errors():
    with tempfile.TemporaryDirectory() as d:
        ff = workload.write_family_files(workload.CONFIGS["tiny"], d)
        with pytest.raises(host.HostError, match="does not exist"):
            host.PhyloHMM(ff["yaml_path"], 0, os.path.join(d, "nope"), 0)          # src/VDJGermline.cpp:51-52
        bad = os.path.join(d, "bad.yaml")
        open(bad, "w").write('{"germline-info": {"locus": "igh"}, "events": [{"unique_ids": ["0"]}]}')
        with pytest.raises(host.HostError, match="flexbounds"):
            host.PhyloHMM(bad, 0, ff["hmm_dir"], 0)                                # src/HMM.cpp:37-44
        open(bad, "w").write('{"events": []}')
        with pytest.raises(host.HostError, match="locus"):
            host.PhyloHMM(bad, 0, ff["hmm_dir"], 0)                                # src/HMM.cpp:31-36


def test_export_newick_format():
    # pll_utree_export_newick: %f lengths, EPS for missing/zero, ":0.0;" at the root (utree.c:148-270)
    assert host.export_newick("((0:0.03,1)x[&index=5]:0.0,naive:0.3,2[&index=1]);") == \
        "((0:0.030000,1:0.000001)x:0.000001,naive:0.300000,2:0.000001):0.0;"


@pytest.mark.skipif(not have_ref_data, reason="needs /root/reference and oracle/_ref")
@pytest.mark.parametrize("locus,n_leaves,seed", [("igk", 9, 41), ("igl", 30, 42)])
def test_light_chain_families_member_by_member(locus, n_leaves, seed):
    """V-J loci have one junction and no D region (src/HMM.cpp:125-170, src/PhyloHMM.cpp:76,106); the
    reference's own tests hold no light-chain fixture, so the compiled reference is the oracle here."""
    from linearham_b200 import synth
    with tempfile.TemporaryDirectory() as d:
        genes = synth.make_germline_set(os.path.join(d, "hmm"), seed=9, locus=locus)
        ff = synth.make_light_family(d, genes, n_leaves, seed=seed, locus=locus)
        ref = po.RefPhyloHMM(ff["yaml_path"], 0, os.path.join(d, "hmm"), 0)
        h = host.PhyloHMM(ff["yaml_path"], 0, os.path.join(d, "hmm"), 0)
        assert h.strs("locus") == [locus]
        skip = ("dgerm", "dj_junction")
        for f in I32_FIELDS:
            if not f.startswith(skip):
                assert same(h.i32(f), ref.i32(f)), f
        for f in F64_FIELDS:
            if not f.startswith(skip) and "dgerm" not in f:
                assert same(h.f64(f), ref.f64(f)), f
        for f in STR_FIELDS:
            if not f.startswith(skip):
                assert h.strs(f) == ref.strs(f), f
        fam, rfam = h.family(), ref.family()
        assert fam["has_d"] == 0 and rfam["has_d"] == 0
        for k in ("tip_codes", "vpad_inds", "vgerm_inds", "jgerm_inds", "jpad_inds", "vd_inds", "v_gene_prob", "v_trans_prod",
                  "vpad_trans", "jpad_trans", "t_v_vd", "t_vd", "t_vd_d"):
            assert np.array_equal(np.asarray(fam[k]).reshape(-1), np.asarray(rfam[k]).reshape(-1)), k
