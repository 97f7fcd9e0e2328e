"""Generates tests/golden/*.npz by running the REFERENCE's own C++ (compiled in place into
oracle/_ref by oracle/refbuild/Makefile) on its Catch fixtures and on small synthetic families.

    python tests/golden/make_golden.py          # needs /root/reference; run in the build container

Each bundle holds: the flat family description (device-ABI layout), the flattened tree batch, the
Newick strings and model parameters, and the reference's outputs for every sample
(log-likelihood, xmsa_emission, forward tables, scaler counts). The bundles travel to the GPU box,
where /root/reference does not exist.
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)

from linearham_b200 import synth  # noqa: E402
from oracle import pyoracle as po, pytrees as pt  # noqa: E402

REFDATA = "/root/reference/data/"
OUT_FIELDS_F64 = ["xmsa_emission", "vgerm_forward", "vd_junction_forward", "dgerm_forward",
                  "dj_junction_forward", "jgerm_forward"]


def ref_outputs(ref, has_d):
    o = {"xmsa_emission": ref.f64("xmsa_emission").reshape(-1),
         "vgerm_forward": ref.f64("vgerm_forward").reshape(-1),
         "vd_forward": np.asfortranarray(ref.f64("vd_junction_forward")).reshape(-1, order="F"),
         "jgerm_forward": ref.f64("jgerm_forward").reshape(-1),
         "vd_scalers": ref.i32("vd_junction_scaler_counts").reshape(-1)}
    gs = [int(ref.i32("vgerm_scaler_count")[0, 0]), 0, int(ref.i32("jgerm_scaler_count")[0, 0])]
    if has_d:
        gs[1] = int(ref.i32("dgerm_scaler_count")[0, 0])
        o["dgerm_forward"] = ref.f64("dgerm_forward").reshape(-1)
        o["dj_forward"] = np.asfortranarray(ref.f64("dj_junction_forward")).reshape(-1, order="F")
        o["dj_scalers"] = ref.i32("dj_junction_scaler_counts").reshape(-1)
    o["germ_scalers"] = np.array(gs, dtype=np.int32)
    return o


def bundle(name, yaml_path, hmm_dir, samples, num_rates):
    ref = po.RefPhyloHMM(yaml_path, 0, hmm_dir, 0)
    fam = ref.family()
    labels = ref.strs("xmsa_labels")
    newicks = [s["newick"] for s in samples]
    er = np.array([s["er"] for s in samples], dtype=np.float64)
    pi = np.array([s["pi"] for s in samples], dtype=np.float64)
    alpha = np.array([s["alpha"] for s in samples], dtype=np.float64)
    rates = np.array([ref.gamma_cats(a, num_rates) for a in alpha])
    tb = pt.flatten_batch(newicks, labels, er, pi, rates)
    exp = {}
    ll = []
    for i, s in enumerate(samples):
        ll.append(ref.score_newick(s["newick"], er[i], pi[i], alpha[i], num_rates))
        for k, v in ref_outputs(ref, fam["has_d"]).items():
            exp.setdefault(k, []).append(v)
    data = {"fam_" + k: np.asarray(v) for k, v in fam.items()}
    data.update({"tb_" + k: v for k, v in tb.items()})
    data.update({"exp_" + k: np.stack(v) for k, v in exp.items()})
    data["exp_loglik"] = np.array(ll)
    data["alpha"] = alpha
    data["newicks"] = np.array(newicks)
    data["labels"] = np.array(labels)
    data["xmsa_seqs"] = np.array(ref.strs("xmsa_seqs"))
    path = os.path.join(HERE, name + ".npz")
    np.savez_compressed(path, **data)
    print(f"{name}: T={fam['n_tips']} S_x={fam['n_sites']} B={len(samples)} R={num_rates} "
          f"loglik[0]={ll[0]!r} scalers[0]={exp['germ_scalers'][0]} -> {os.path.getsize(path)} bytes")


def main():
    if not po.have_ref():
        po.build_ref()
    newton = open(REFDATA + "newton.tree").read().strip()
    base = {"newick": newton, "er": [1.0] * 6, "pi": [0.17, 0.19, 0.25, 0.39], "alpha": 1.0}
    # the three Catch fixtures (test/test.cpp:750-1399), plus perturbed model parameters on fixture A
    rng = np.random.default_rng(11)
    extra = []
    for _ in range(5):
        pi = np.maximum(rng.dirichlet(np.ones(4)), 0.01)
        extra.append({"newick": "((0:0.03,1:0.4):0.0,naive:0.3,2);" if _ == 0 else newton,
                      "er": rng.dirichlet(np.ones(6)), "pi": pi / pi.sum(), "alpha": max(0.05, rng.exponential(1.0))})
    bundle("catch_a", REFDATA + "phylo_hmm_input.yaml", REFDATA + "hmm_params", [base] + extra, 4)
    bundle("catch_b", REFDATA + "phylo_hmm_input_extra.yaml", REFDATA + "hmm_params", [base], 4)
    bundle("catch_c", REFDATA + "phylo_likelihood_hmm_input.yaml", REFDATA + "phylo_likelihood_hmm_params", [base], 1)
    # synthetic families (SURVEY.md 8(d) recipe, synthetic germline parameters)
    with tempfile.TemporaryDirectory() as d:
        genes = synth.make_germline_set(os.path.join(d, "hmm"), seed=7)
        for name, n, B, R, seed in [("synth_n12", 12, 6, 4, 1301), ("synth_n40", 40, 4, 4, 1302),
                                    ("synth_n24_r1", 24, 3, 1, 1303), ("synth_n200", 200, 2, 4, 1304)]:
            fam = synth.make_family(d, genes, n, seed=seed, name=name)
            samples = synth.tree_samples(np.random.default_rng(seed + 1000), fam["tree"], B, walk="cumulative")
            bundle(name, fam["yaml_path"], os.path.join(d, "hmm"), samples, R)
        # the smallest tree the path accepts: two observed sequences + naive = one operation (3 tips)
        fam = synth.make_family(d, genes, 2, seed=1310, name="synth_n2")
        samples = synth.tree_samples(np.random.default_rng(2310), fam["tree"], 5, outgroup="naive")
        bundle("synth_n2", fam["yaml_path"], os.path.join(d, "hmm"), samples, 4)
    make_light()


def make_light():
    """Light-chain family (igk: V-J, no D region; src/HMM.cpp:125-170, src/PhyloHMM.cpp:76,106) -
    the reference's own tests hold no light-chain fixture."""
    with tempfile.TemporaryDirectory() as d:
        genes = synth.make_germline_set(os.path.join(d, "hmmk"), seed=9, locus="igk")
        fam = synth.make_light_family(d, genes, 30, seed=1305, name="synth_igk_n30")
        samples = synth.tree_samples(np.random.default_rng(2305), fam["tree"], 4, outgroup="naive")
        bundle("synth_igk_n30", fam["yaml_path"], os.path.join(d, "hmmk"), samples, 4)


if __name__ == "__main__":
    main()
