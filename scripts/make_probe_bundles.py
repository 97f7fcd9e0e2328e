"""Build-container helper: writes build/probe_*.npz (family + flattened tree batches) using the
compiled reference for the family description. Not committed; travels with gpurun snapshots."""
import os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from linearham_b200 import synth
from oracle import pyoracle as po, pytrees as pt

def make(name, n_leaves, B, R, fam_seed, tree_seed, n_ref=4):
    with tempfile.TemporaryDirectory() as d:
        genes = synth.make_germline_set(os.path.join(d, "hmm"), seed=7)
        fam = synth.make_family(d, genes, n_leaves, seed=fam_seed, name=name)
        ref = po.RefPhyloHMM(fam["yaml_path"], 0, os.path.join(d, "hmm"), 0)
        F = ref.family()
        samples = synth.tree_samples(np.random.default_rng(tree_seed), fam["tree"], B)
        alpha = np.array([s["alpha"] for s in samples])
        rates = np.array([po.port_gamma_cats(a, R) for a in alpha])
        t0 = time.time()
        tb = pt.flatten_batch([s["newick"] for s in samples], fam["labels"], [s["er"] for s in samples],
                              [s["pi"] for s in samples], rates)
        ll = [ref.score_newick(samples[i]["newick"], samples[i]["er"], samples[i]["pi"], alpha[i], R) for i in range(n_ref)]
        data = {"fam_" + k: np.asarray(v) for k, v in F.items()}
        data.update({"tb_" + k: v for k, v in tb.items()})
        data["ref_loglik"] = np.array(ll)
        np.savez(os.path.join(ROOT, "build", name + ".npz"), **data)
        print(name, "T", F["n_tips"], "S_x", F["n_sites"], "B", B, "flatten s", round(time.time() - t0, 1), "ref ll", ll[:2])

if __name__ == "__main__":
    make("probe_c3", 128, 4096, 4, 1300, 2003)
    make("probe_c4", 1000, 256, 4, 1400, 2004, n_ref=2)
