"""tests/golden/scaling_*.npz: families in the underflow regime (libpll test/out/scaling.out is the
model: deep caterpillars whose per-site likelihoods leave the double range), scored by the REFERENCE's own C++
(oracle/_ref). They pin the scaling semantics where a per-category rescaling rule and libpll's all-categories rule
(core_partials.c:648-659) could part: alpha 0.02-0.5 with 4 and 8 categories, 1000-3000 tips, ladders, random trees
and a constant clade joined to a hypermutated one.    python scripts/make_scaling_goldens.py   (needs /root/reference)"""
import os, sys, tempfile
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
sys.path.insert(0, os.path.join(ROOT, "scripts"))
import numpy as np
import make_golden
import scaling_probe as sp
from linearham_b200 import synth

CASES = {
    "scaling_cat1000_a002_r4": (1000, lambda r, n: synth.caterpillar_tree(r, n, 0.1), None, 0.02, 4),
    "scaling_cat3000_a005_r8": (3000, lambda r, n: synth.caterpillar_tree(r, n, 0.01), None, 0.05, 8),
    "scaling_cat1500_a002_r4": (1500, lambda r, n: synth.caterpillar_tree(r, n, 0.02), None, 0.02, 4),
    "scaling_rand2000_a003_r4": (2000, lambda r, n: synth.random_tree(r, n, 0.02), None, 0.03, 4),
    "scaling_rand3000_a005_r8": (3000, lambda r, n: synth.random_tree(r, n, 0.01), None, 0.05, 8),
    # milder: every log-likelihood finite, the high-rate categories far below 2^-256 while the low-rate ones are not
    "scaling_cat3000_mild_r8": (3000, lambda r, n: synth.caterpillar_tree(r, n, 0.002), None, 0.05, 8),
    "scaling_rand3000_mild_r4": (3000, lambda r, n: synth.random_tree(r, n, 0.003), None, 0.1, 4),
    "scaling_clades_a002_r4": (1200, sp.two_clade_tree(1150, 50, 0.2, 0.2), sp.const_and_random(1150), 0.02, 4),
    "scaling_clades_a005_r8": (3000, sp.two_clade_tree(2900, 100, 0.1, 0.3), sp.const_and_random(2900), 0.05, 8),
    "scaling_clades_a05_r4": (2000, sp.two_clade_tree(1900, 100, 0.3, 0.3), sp.const_and_random(1900), 0.5, 4),
}

if __name__ == "__main__":
    only = sys.argv[1:]
    with tempfile.TemporaryDirectory() as d:
        genes = synth.make_germline_set(os.path.join(d, "hmm"), seed=7)
        for name, (n, tree_fn, evolve_fn, alpha, R) in CASES.items():
            if only and name not in only:
                continue
            fam = synth.make_family(d, genes, n, seed=4242, name=name, tree_fn=tree_fn, evolve_fn=evolve_fn)
            rng = np.random.default_rng(99)
            samples = []
            for i in range(2):
                pi = np.maximum(rng.dirichlet(np.ones(4)), 0.05)
                samples.append({"newick": fam["tree"].newick(with_index=True), "er": rng.dirichlet(np.ones(6)),
                                "pi": pi / pi.sum(), "alpha": alpha if i == 0 else alpha * 1.5})
            make_golden.bundle(name, fam["yaml_path"], os.path.join(d, "hmm"), samples, R)
