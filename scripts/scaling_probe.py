"""Experiment (CPU): do libpll's all-categories scaling rule and a per-category rule give different doubles?
Scores candidate underflow-regime families with the scalar port in both modes (LH_PORT_RATE_SCALERS) and, where
oracle/_ref exists, with the reference itself.   python scripts/scaling_probe.py <case> <mode: site|rate|ref> <out.npz>"""
import os, sys, tempfile
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
from linearham_b200 import synth, host
from oracle import pyoracle as po, pytrees as pt

BASES = "ACGT"


def two_clade_tree(n_const, n_var, blen_const, blen_var):
    def fn(rng, n_leaves):
        t = synth.Tree()
        ids = list(range(n_leaves))
        def ladder(parent, labels, bl):
            node = t.add(parent, bl)
            if len(labels) == 1:
                t.label[node] = str(labels[0]); return
            ladder(node, labels[:-1], bl)
            tip = t.add(node, bl); t.label[tip] = str(labels[-1])
        import sys as _s; _s.setrecursionlimit(20000)
        ladder(0, ids[:n_const], blen_const)
        ladder(0, ids[n_const:], blen_var)
        t.add(0, 0.002, "naive")
        return t
    return fn


def const_and_random(n_const):
    def fn(rng, tree, naive):
        seqs = {}
        for i, lab in enumerate(tree.label):
            if lab is None or lab == "naive":
                continue
            if int(lab) < n_const:
                seqs[lab] = naive
            else:
                seqs[lab] = "".join(c if c == "N" else BASES[int(rng.integers(4))] for c in naive)
        return seqs
    return fn


CASES = {
    # name: (n_leaves, tree_fn, evolve_fn, alpha, R, branch scale)
    "cat1000_a002_r4": (1000, lambda r, n: synth.caterpillar_tree(r, n, 0.1), None, 0.02, 4),
    "cat3000_a005_r8": (3000, lambda r, n: synth.caterpillar_tree(r, n, 0.05), None, 0.05, 8),
    "cat1500_a002_r4_long": (1500, lambda r, n: synth.caterpillar_tree(r, n, 0.5), None, 0.02, 4),
    "rand2000_a003_r4": (2000, lambda r, n: synth.random_tree(r, n, 0.3), None, 0.03, 4),
    "two_clade_a002_r4": (1200, two_clade_tree(1150, 50, 0.2, 0.2), const_and_random(1150), 0.02, 4),
    "two_clade_a005_r8": (3000, two_clade_tree(2900, 100, 0.1, 0.3), const_and_random(2900), 0.05, 8),
    "two_clade_a05_r4": (2000, two_clade_tree(1900, 100, 0.3, 0.3), const_and_random(1900), 0.5, 4),
}


def build(case, d):
    n, tree_fn, evolve_fn, alpha, R = CASES[case]
    genes = synth.make_germline_set(os.path.join(d, "hmm"), seed=7)
    fam = synth.make_family(d, genes, n, seed=4242, name=case, tree_fn=tree_fn, evolve_fn=evolve_fn)
    rng = np.random.default_rng(99)
    samples = []
    for i in range(2):
        pi = np.maximum(rng.dirichlet(np.ones(4)), 0.05); pi /= pi.sum()
        samples.append({"newick": fam["tree"].newick(with_index=True), "er": rng.dirichlet(np.ones(6)), "pi": pi,
                        "alpha": alpha if i == 0 else alpha * 1.5})
    return fam, samples, R, os.path.join(d, "hmm")


if __name__ == "__main__":
    case, mode, out = sys.argv[1:4]
    with tempfile.TemporaryDirectory() as d:
        fam, samples, R, hmm = build(case, d)
        if mode == "ref":
            ref = po.RefPhyloHMM(fam["yaml_path"], 0, hmm, 0)
            ll, xe = [], []
            for s in samples:
                ll.append(ref.score_newick(s["newick"], s["er"], s["pi"], s["alpha"], R))
                xe.append(ref.f64("xmsa_emission").reshape(-1))
            np.savez(out, ll=np.array(ll), xe=np.stack(xe))
        else:
            ph = host.PhyloHMM(fam["yaml_path"], 0, hmm, 0)
            f = ph.family(); labels = ph.strs("xmsa_labels")
            tb = host.tree_batch([s["newick"] for s in samples], labels, [s["er"] for s in samples], [s["pi"] for s in samples],
                                 [s["alpha"] for s in samples], R)
            ll, o = po.port_score_batch(f, tb, fields=["xmsa_emission"])
            np.savez(out, ll=ll, xe=o["xmsa_emission"])
