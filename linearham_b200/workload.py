"""Synthetic workloads of BASELINE.json's configs (SURVEY.md section 8(d) recipe): the files the
reference reads (HMM parameter dir, cluster YAML, Newick samples) plus the flat tree batch the
device C ABI takes. No likelihood arithmetic here. Used by bench.py, tests and the golden-bundle
scripts so that every arm sees byte-identical inputs."""
import hashlib
import os

import numpy as np

from . import host, synth

# name -> leaves, trees per batch (whole job), Gamma categories, family seed, tree seed
CONFIGS = {
    "c3": dict(n_leaves=128, B=4096, R=4, fam_seed=1300, tree_seed=2003, outgroup="naive",
               desc="synthetic 128-leaf IGH family, 400 nt naive, Gamma-4 GTR, batch 4096 trees"),
    "c4": dict(n_leaves=1000, B=16384, R=4, fam_seed=1400, tree_seed=2004, outgroup="naive",
               desc="synthetic 1000-leaf IGH family, 400 nt, Gamma-4, batch 16384 trees"),
    # repertoire sweep: independent families, leaves log-uniform in [20, 500] (SURVEY.md 8(d) "C5");
    # BASELINE names 10 000 families x 1000 samples - bench and tests state the subset they run
    "c5": dict(n_families=10000, leaves=(20, 500), B=1000, R=4, fam_seed=1500000, tree_seed=2500000, outgroup="naive",
               desc="repertoire sweep: synthetic clonal families of 20-500 leaves scored independently"),
    "tiny": dict(n_leaves=24, B=256, R=4, fam_seed=1500, tree_seed=2005,
                 desc="synthetic 24-leaf IGH family (tests)"),
}
GERMLINE_SEED = 7


def write_family_files(cfg, outdir):
    """-> dict(yaml_path, hmm_dir, labels, tree, sha) ; files are deterministic in cfg."""
    hmm_dir = os.path.join(outdir, "hmm")
    genes = synth.make_germline_set(hmm_dir, seed=GERMLINE_SEED)
    fam = synth.make_family(outdir, genes, cfg["n_leaves"], seed=cfg["fam_seed"], name="family")
    h = hashlib.sha256()
    h.update(open(fam["yaml_path"], "rb").read())
    for fn in sorted(os.listdir(hmm_dir)):
        h.update(open(os.path.join(hmm_dir, fn), "rb").read())
    fam["hmm_dir"] = hmm_dir
    fam["sha"] = h.hexdigest()
    return fam


def tree_batch(cfg, fam_files, B, n_topo=None, n_threads=0):
    """B tree samples for the family. The first n_topo samples are a Markov chain of topologies
    (Poisson(3) NNIs per step - with cfg["outgroup"] the naive tip stays under the trifurcation, as
    RevBayes' outgroup constraint keeps it - LogNormal(0,0.3) branch noise anchored on the true tree) written
    as RevBayes-style Newick strings; sample b >= n_topo reuses topology b % n_topo with fresh
    LogNormal(0,0.1) noise on every branch. er ~ Dirichlet(1), pi ~ Dirichlet(1) floored at
    0.01, alpha = max(0.05, Exp(1)) are fresh for every sample.
    -> (tb dict for the C ABI, newicks[n_topo], alpha[B])"""
    rng = np.random.default_rng(cfg["tree_seed"])
    n_topo = min(B, n_topo or B)
    samples = synth.tree_samples(rng, fam_files["tree"], n_topo, outgroup=cfg.get("outgroup"))
    newicks = [s["newick"] for s in samples]
    base = host.newick_to_ops(newicks, fam_files["labels"], n_threads)
    reps = (B + n_topo - 1) // n_topo
    tb = {k: np.concatenate([v] * reps)[:B].copy() for k, v in base.items()}
    if B > n_topo:
        tb["blen"][n_topo:] *= rng.lognormal(0.0, 0.1, tb["blen"][n_topo:].shape)
        tb["root_blen"][n_topo:] *= rng.lognormal(0.0, 0.1, B - n_topo)
    er = np.array([s["er"] for s in samples])
    pi = np.array([s["pi"] for s in samples])
    alpha = np.array([s["alpha"] for s in samples])
    if B > n_topo:
        m = B - n_topo
        er = np.concatenate([er, rng.dirichlet(np.ones(6), m)])
        p2 = np.maximum(rng.dirichlet(np.ones(4), m), 0.01)
        pi = np.concatenate([pi, p2 / p2.sum(axis=1, keepdims=True)])
        alpha = np.concatenate([alpha, np.maximum(0.05, rng.exponential(1.0, m))])
    tb["er"], tb["pi"] = np.ascontiguousarray(er), np.ascontiguousarray(pi)
    tb["rates"] = host.gamma_cats(alpha, cfg["R"])
    return tb, newicks, alpha


def repertoire_sizes(cfg, n_families):
    """Leaf counts of the first n_families families of a repertoire config: round(exp(U(log lo, log hi)))."""
    lo, hi = cfg["leaves"]
    rng = np.random.default_rng(cfg["fam_seed"])
    return np.rint(np.exp(rng.uniform(np.log(lo), np.log(hi), n_families))).astype(int)


def repertoire(cfg, outdir, n_families, B=None, n_topo=16, device=0, n_threads=0):
    """The first n_families families of a repertoire config, each with B tree samples:
    -> list of dict(fam=family description built by the product's C++ PhyloHMM host class, tb, newicks,
    alpha, files, n_leaves). The HMM parameter directory is shared (one germline set per repertoire)."""
    B = B or cfg["B"]
    hmm_dir = os.path.join(outdir, "hmm")
    genes = synth.make_germline_set(hmm_dir, seed=GERMLINE_SEED)
    out = []
    for i, n_leaves in enumerate(repertoire_sizes(cfg, n_families)):
        ff = synth.make_family(outdir, genes, int(n_leaves), seed=cfg["fam_seed"] + 1 + i, name=f"family_{i}")
        ff["hmm_dir"] = hmm_dir
        ph = host.PhyloHMM(ff["yaml_path"], 0, hmm_dir, 0, device=device)
        fam, labels = ph.family(), ph.strs("xmsa_labels")
        ph.close()
        if labels != ff["labels"]:
            raise RuntimeError("xMSA labels do not match the generated family")
        sub = dict(cfg, tree_seed=cfg["tree_seed"] + i)
        tb, newicks, alpha = tree_batch(sub, ff, B, n_topo=n_topo, n_threads=n_threads)
        out.append(dict(fam=fam, tb=tb, newicks=newicks, alpha=alpha, files=ff, n_leaves=int(n_leaves)))
    return out


def liao_batch(z):
    """BASELINE config 2: rebuilds the 1024-sample tree batch (device-ABI layout) of tests/golden/liao_b1024.npz
    from its arrays - 32 topologies x 128 noise vectors (int8 indices into a stored table of doubles), exact
    IEEE products, so every machine gets the doubles the reference scored (scripts/make_liao_bundle.py)."""
    topo, noise, table = z["topo_of"], z["noise_of"], z["noise_table"]
    return {"children": np.ascontiguousarray(z["base_children"][topo]),
            "blen": np.ascontiguousarray(z["base_blen"][topo] * table[z["noise_k"][noise].astype(np.int64) + 128]),
            "root_child": np.ascontiguousarray(z["base_root_child"][topo]),
            "root_blen": np.ascontiguousarray(z["base_root_blen"][topo]
                                              * table[z["noise_root_k"][noise].astype(np.int64) + 128]),
            "er": np.ascontiguousarray(z["er"]), "pi": np.ascontiguousarray(z["pi"]),
            "rates": np.ascontiguousarray(z["rates"])}


def bytes_per_tree(T, S, R=4):
    """ALGORITHMIC bytes of one tree evaluation under the streaming model:
    every inner CLV written once and read once with its per-site uint32 scaler, tip codes read once,
    P-matrices written and read once, one FP64 emission per column out."""
    return (T - 2) * S * (2 * 32 * R + 2 * 4) + T * S + 2 * (2 * T - 3) * R * 128 + S * 8


def flops_per_tree(T, S, R=4):
    """FP64 flop of pruning, inner x inner accounting."""
    return 60 * R * (T - 2) * S
