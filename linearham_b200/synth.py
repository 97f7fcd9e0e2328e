"""Deterministic synthetic inputs for the phylo-HMM likelihood path (bench + tests).

Writes files in exactly the formats the reference reads:
  * a partis HMM parameter directory, one YAML per germline gene
    (schema parsed by /root/reference/src/{Germline,NTInsertion,NPadding}.cpp)
  * a partis cluster YAML with `linearham-info` flexbounds/relpos (src/HMM.cpp:30-44,71-83)
  * RevBayes-format tree-sample TSVs (src/PhyloHMM.cpp:396-400)
following the recipe of SURVEY.md section 8(d). The germline parameter values themselves are
synthetic (gene lengths / landing structure modelled on human IGH), because the real partis
parameter files live in the reference checkout and do not travel to the GPU box.
No numerics of the hot path live here.
"""
import os

import numpy as np

BASES = "ACGT"


# ------------------------------------------------------------------ germline parameter files ---

def _rand_seq(rng, n):
    return "".join(BASES[i] for i in rng.integers(0, 4, n))


def _mutate_seq(rng, s, k):
    s = list(s)
    for p in rng.choice(len(s), size=min(k, len(s)), replace=False):
        s[p] = BASES[(BASES.index(s[p]) + int(rng.integers(1, 4))) % 4]
    return "".join(s)


def _fmt(x):
    return repr(float(x))


def _emission_block(probs, germ, indent="  "):
    lines = [f"{indent}emissions:", f"{indent}  probs:"]
    for b, p in zip(BASES, probs):
        lines.append(f"{indent}    {b}: {_fmt(p)}")
    lines += [f"{indent}  track: nukes", f"{indent}extras:", f"{indent}  ambiguous_char: N",
              f"{indent}  ambiguous_emission_prob: 0.25", f"{indent}  germline: {germ}"]
    return lines


def _state(name, transitions, emis=None, germ=None):
    lines = ["- !!python/object:python.hmmwriter.State"]
    if emis is None:
        lines += ["  emissions: null", "  extras: {}"]
    else:
        lines += _emission_block(emis, germ)
    lines.append(f"  name: {name}")
    lines.append("  transitions:")
    for k, v in transitions.items():
        lines.append(f"    {k}: {_fmt(v)}")
    return lines


def _germ_emission(rng, base, mute):
    p = np.full(4, mute / 3.0)
    p[BASES.index(base)] = 1.0 - mute
    return p


def write_gene_yaml(path, gname, seq, gtype, gene_prob, rng):
    """gname uses partis file-name escaping (IGHV1-2_star_01). gtype in 'VDJ'."""
    n = len(seq)
    lines = ["!!python/object:python.hmmwriter.HMM", "extras:", f"  gene_prob: {_fmt(gene_prob)}",
             f"name: {gname}", "states:"]
    mute = rng.uniform(0.02, 0.15, n)
    if gtype == "V":
        pn = float(rng.uniform(0.2, 0.5))
        init = {f"{gname}_0": 1.0 - pn, "insert_left_N": pn}
        lines += _state("init", init)
        lines += _state("insert_left_N", init, [0.25] * 4, "N")
    else:
        # landing-in over the first positions + the four NTI states
        k = min(n, 12 if gtype == "D" else 20)
        w = rng.dirichlet(np.ones(k + 4))
        init = {f"{gname}_{i}": w[i] for i in range(k)}
        init.update({f"insert_left_{b}": w[k + j] for j, b in enumerate(BASES)})
        lines += _state("init", init)
        for b in BASES:
            w = rng.dirichlet(np.ones(k + 4))
            tr = {f"{gname}_{i}": w[i] for i in range(k)}
            tr.update({f"insert_left_{c}": w[k + j] for j, c in enumerate(BASES)})
            em = _germ_emission(rng, b, float(rng.uniform(0.05, 0.3)))
            lines += _state(f"insert_left_{b}", tr, em, b)
    # germline states
    erosion = 14 if gtype != "J" else 0  # positions near the 3' end that may exit early
    for i, base in enumerate(seq):
        if i == n - 1:
            if gtype == "J":
                tr = {"end": 0.04, "insert_right_N": 0.96}
            else:
                tr = {"end": 1.0}
        else:
            if gtype != "J" and i >= n - 1 - erosion:
                pend = float(rng.uniform(0.02, 0.3))
                tr = {f"{gname}_{i + 1}": 1.0 - pend, "end": pend}
            else:
                tr = {f"{gname}_{i + 1}": 1.0}
        lines += _state(f"{gname}_{i}", tr, _germ_emission(rng, base, mute[i]), base)
    if gtype == "J":
        lines += _state("insert_right_N", {"end": 0.04, "insert_right_N": 0.96}, [0.25] * 4, "N")
    lines += ["tracks:", "  nukes:", "  - A", "  - C", "  - G", "  - T", ""]
    with open(path, "w") as f:
        f.write("\n".join(lines))


def make_germline_set(outdir, seed=7, n_v=3, n_d=5, n_j=2, v_len=296, j_len=63, locus="igh"):
    """Writes the parameter dir; returns {display name: (type, sequence)} with '*' names.
    locus "igk" / "igl": light chain, no D genes (src/HMM.cpp:125-170)."""
    pre = {"igh": "IGH", "igk": "IGK", "igl": "IGL"}[locus]
    if locus != "igh":
        n_d = 0
    rng = np.random.default_rng(seed)
    os.makedirs(outdir, exist_ok=True)
    genes = {}
    v0 = _rand_seq(rng, v_len)
    j0 = _rand_seq(rng, j_len)
    gp = {"V": rng.dirichlet(np.ones(n_v) * 5), "D": rng.dirichlet(np.ones(max(n_d, 1)) * 5),
          "J": rng.dirichlet(np.ones(n_j) * 5)}
    for i in range(n_v):
        seq = v0 if i == 0 else _mutate_seq(rng, v0, 6)[int(rng.integers(0, 4)):]
        genes[f"{pre}V{i + 1}-1*01"] = ("V", seq, gp["V"][i])
    for i in range(n_d):
        genes[f"{pre}D{i + 1}-1*01"] = ("D", _rand_seq(rng, int(rng.integers(17, 32))), gp["D"][i])
    for i in range(n_j):
        seq = j0 if i == 0 else _mutate_seq(rng, j0, 4)[int(rng.integers(0, 6)):]
        genes[f"{pre}J{i + 1}*01"] = ("J", seq, gp["J"][i])
    for name, (t, seq, p) in genes.items():
        fname = name.replace("*", "_star_").replace("/", "_slash_")
        write_gene_yaml(os.path.join(outdir, fname + ".yaml"), fname, seq, t, p, rng)
    return {k: (v[0], v[1]) for k, v in genes.items()}


# ------------------------------------------------------------------------------- trees ---

class Tree:
    """Unrooted tree stored rooted at its trifurcation. Node 0 is the root (3 children);
    every other inner node has 2 children. Leaves carry labels."""

    def __init__(self):
        self.children = [[]]
        self.parent = [-1]
        self.blen = [0.0]
        self.label = [None]

    def add(self, parent, blen, label=None):
        self.children.append([])
        self.parent.append(parent)
        self.blen.append(float(blen))
        self.label.append(label)
        i = len(self.parent) - 1
        self.children[parent].append(i)
        return i

    def copy(self):
        t = Tree()
        t.children = [list(c) for c in self.children]
        t.parent = list(self.parent)
        t.blen = list(self.blen)
        t.label = list(self.label)
        return t

    def newick(self, with_index=False):
        import sys
        sys.setrecursionlimit(max(sys.getrecursionlimit(), 4 * len(self.parent) + 1000))
        out = []
        counter = [0]

        def rec(i):
            if not self.children[i]:
                out.append(self.label[i])
            else:
                out.append("(")
                for k, c in enumerate(self.children[i]):
                    if k:
                        out.append(",")
                    rec(c)
                out.append(")")
            if with_index:
                counter[0] += 1
                out.append(f"[&index={counter[0]}]")
            if i != 0:
                out.append(":" + repr(self.blen[i]))

        rec(0)
        out.append(";")
        return "".join(out)

    def inner_nonroot(self):
        return [i for i in range(1, len(self.parent)) if self.children[i]]

    def nni(self, rng, fixed=None):
        """One nearest-neighbour interchange across a random inner edge. `fixed`: a node that never
        moves (the outgroup tip under the root)."""
        cand = self.inner_nonroot()
        if not cand:
            return
        v = cand[int(rng.integers(len(cand)))]
        p = self.parent[v]
        sibs = [c for c in self.children[p] if c != v and c != fixed]
        s = sibs[int(rng.integers(len(sibs)))]
        c = self.children[v][int(rng.integers(2))]
        # swap subtree c (child of v) with subtree s (sibling of v)
        self.children[v][self.children[v].index(c)] = s
        self.children[p][self.children[p].index(s)] = c
        self.parent[s], self.parent[c] = v, p


def random_tree(rng, n_leaves, blen_mean=0.01):
    """Trifurcating root: two random subtrees + the 'naive' tip (branch 0.002); SURVEY 8(d)."""
    t = Tree()
    leaf_ids = list(range(n_leaves))

    def bl():
        return 1e-4 + rng.exponential(blen_mean)

    def grow(parent, labels):
        node = t.add(parent, bl())
        if len(labels) == 1:
            t.label[node] = str(labels[0])
            return
        k = int(rng.integers(1, len(labels)))
        grow(node, labels[:k])
        grow(node, labels[k:])

    import sys
    sys.setrecursionlimit(max(10000, 4 * n_leaves))
    k = int(rng.integers(1, n_leaves)) if n_leaves > 1 else 1
    grow(0, leaf_ids[:k])
    if n_leaves > 1:
        grow(0, leaf_ids[k:])
    t.add(0, 0.002, "naive")
    if n_leaves == 1:
        raise ValueError("need at least 2 observed sequences")
    return t


def caterpillar_tree(rng, n_leaves, blen=0.05):
    """A ladder: tip 0 and tip 1 deepest, every further tip one node up; the naive tip and the last tip hang under
    the trifurcation. Depth n_leaves - 2: the shape of libpll's own scaling test (test/out/scaling.out)."""
    t = Tree()
    chain = [t.add(0, blen) for _ in range(1)]
    for _ in range(n_leaves - 3):
        chain.append(t.add(chain[-1], blen))
    # chain[k] is an inner node; its second child is tip (n_leaves - 2 - k); the deepest holds tips 0 and 1
    for k, node in enumerate(chain):
        tip = t.add(node, blen)
        t.label[tip] = str(n_leaves - 2 - k)
    last = t.add(chain[-1], blen)
    t.label[last] = "0"
    top = t.add(0, blen)
    t.label[top] = str(n_leaves - 1)
    t.add(0, 0.002, "naive")
    return t


def evolve(rng, tree, naive_seq):
    """Mutate non-N sites with probability = branch length, uniformly to another base."""
    root_seq = np.array([BASES.index(c) if c in BASES else 4 for c in naive_seq], dtype=np.int8)
    seqs = {}
    stack = [(0, root_seq)]
    while stack:
        node, s = stack.pop()
        for c in tree.children[node]:
            if tree.label[c] == "naive":
                continue
            cs = s.copy()
            hit = (rng.random(len(cs)) < tree.blen[c]) & (cs != 4)
            cs[hit] = (cs[hit] + rng.integers(1, 4, int(hit.sum()))) % 4
            if tree.children[c]:
                stack.append((c, cs))
            else:
                seqs[tree.label[c]] = "".join("ACGTN"[x] for x in cs)
    return seqs


def tree_samples(rng, tree, B, walk="anchored", outgroup=None):
    """Sample 0 = `tree`; sample i = sample i-1 after Poisson(3) NNIs, with LogNormal(0,0.3) branch noise.

    walk="anchored" (default): the noise multiplies the branch lengths of `tree` (node ids are stable
    under NNI), so lengths stay on the scale of the family however long the chain is.
    walk="cumulative": the noise compounds from sample to sample (SURVEY.md 8(d) as written). Over
    thousands of samples the lengths random-walk over many orders of magnitude and the reference
    itself returns inf/NaN; the committed golden bundles (<= 6 samples) were made this way.
    outgroup: label of a tip that stays under the trifurcation, as the naive sequence does in the
    trees linearham scores (templates/revbayes_template.rev:55, dnUniformTopology(outgroup=...))."""
    cur = tree.copy()
    fixed = None
    if outgroup is not None:
        fixed = next(i for i, lab in enumerate(tree.label) if lab == outgroup)
    base = list(tree.blen)
    out = []
    for i in range(B):
        if i:
            cur = cur.copy()
            for _ in range(int(rng.poisson(3))):
                cur.nni(rng, fixed)
            for j in range(1, len(cur.blen)):
                noise = float(rng.lognormal(0.0, 0.3))
                cur.blen[j] = (cur.blen[j] if walk == "cumulative" else base[j]) * noise
        er = rng.dirichlet(np.ones(6))
        pi = np.maximum(rng.dirichlet(np.ones(4)), 0.01)
        pi /= pi.sum()
        alpha = max(0.05, float(rng.exponential(1.0)))
        out.append({"newick": cur.newick(with_index=True), "er": er, "pi": pi, "alpha": alpha})
    return out


def write_tsv(path, samples):
    """RevBayes-style trees file (src/PhyloHMM.cpp:396-400)."""
    cols = ["Iteration", "Posterior", "Likelihood", "Prior", "alpha"] + [f"er[{i}]" for i in range(1, 7)] + \
           [f"pi[{i}]" for i in range(1, 5)] + ["tree"]
    with open(path, "w") as f:
        f.write("\t".join(cols) + "\n")
        for i, s in enumerate(samples):
            row = [str(i * 10), "-1000.5", repr(-1000.0 - i), "-50.25", repr(s["alpha"])] + \
                  [repr(float(x)) for x in s["er"]] + [repr(float(x)) for x in s["pi"]] + [s["newick"]]
            f.write("\t".join(row) + "\n")


# ------------------------------------------------------------------------------ families ---

def make_light_family(outdir, genes, n_leaves, seed, L=400, name="family", locus="igk"):
    """Light-chain family (V-J, one junction): naive = lead Ns + V + 8 inserted bases + J + Ns."""
    rng = np.random.default_rng(seed)
    vs = [g for g, (t, _) in genes.items() if t == "V"]
    js = [g for g, (t, _) in genes.items() if t == "J"]
    V0, J0 = genes[vs[0]][1], genes[js[0]][1]
    lead = 15
    naive = "N" * lead + V0[:len(V0) - 6]
    vend = len(naive)
    naive += _rand_seq(rng, 8)
    jstart = len(naive)
    naive += J0[4:]
    jend = len(naive)
    assert jend <= L, (jend, L)
    naive += "N" * (L - jend)
    flex = {"v_l": [0, lead], "v_r": [vend - 20, vend - 10], "j_l": [jstart + 4, jstart + 14], "j_r": [jend, L]}
    relpos = {vs[0]: lead, js[0]: jstart - 4}
    for v in vs[1:]:
        relpos[v] = lead + len(V0) - len(genes[v][1])
    for j in js[1:]:
        relpos[j] = (jstart - 4) + len(J0) - len(genes[j][1])
    tree = random_tree(rng, n_leaves)
    seqs = evolve(rng, tree, naive)
    ids = [str(i) for i in range(n_leaves)]
    import json
    doc = {"germline-info": {"locus": locus},
           "events": [{"unique_ids": ids, "naive_seq": naive, "input_seqs": [seqs[i] for i in ids],
                       "has_shm_indels": [False] * n_leaves,
                       "linearham-info": {"flexbounds": flex, "relpos": relpos}}]}
    os.makedirs(outdir, exist_ok=True)
    path = os.path.join(outdir, name + ".yaml")
    with open(path, "w") as f:
        json.dump(doc, f)
    return {"yaml_path": path, "tree": tree, "labels": ["naive"] + ids, "naive": naive,
            "flexbounds": flex, "relpos": relpos}


def make_family(outdir, genes, n_leaves, seed, L=400, name="family", tree_fn=None, evolve_fn=None):
    """Writes <outdir>/<name>.yaml; returns dict(yaml_path, tree, labels, naive).
    tree_fn(rng, n_leaves) / evolve_fn(rng, tree, naive) replace random_tree / evolve (test families with a
    chosen shape and chosen sequences)."""
    rng = np.random.default_rng(seed)
    vs = [g for g, (t, _) in genes.items() if t == "V"]
    ds = [g for g, (t, _) in genes.items() if t == "D"]
    js = [g for g, (t, _) in genes.items() if t == "J"]
    V0, D0, J0 = genes[vs[0]][1], genes[ds[0]][1], genes[js[0]][1]
    lead = 15
    naive = "N" * lead + V0[:len(V0) - 6]
    vend = len(naive)
    naive += _rand_seq(rng, 6)
    dstart = len(naive)
    naive += D0[3:15]
    naive += _rand_seq(rng, 7)
    jstart = len(naive)
    naive += J0[4:]
    jend = len(naive)
    assert jend <= L, (jend, L)
    naive += "N" * (L - jend)
    flex = {"v_l": [0, lead], "v_r": [vend - 20, vend - 10], "d_l": [dstart, dstart + 4],
            "d_r": [dstart + 5, dstart + 9], "j_l": [jstart + 4, jstart + 14], "j_r": [jend, L]}
    relpos = {vs[0]: lead, ds[0]: dstart - 3, js[0]: jstart - 4}
    for v in vs[1:]:
        relpos[v] = lead + len(V0) - len(genes[v][1])
    for j in js[1:]:
        relpos[j] = (jstart - 4) + len(J0) - len(genes[j][1])
    dl2 = flex["d_l"][1]
    for d in ds[1:]:
        n = len(genes[d][1])
        # cover site d_l.second .. d_r.first and reach into the V-D junction
        lo = max(flex["d_r"][0] - n + 1, flex["v_r"][0] + 1)
        hi = min(dl2 - 1, dl2)
        relpos[d] = int(rng.integers(min(lo, hi), hi + 1))
    tree = (tree_fn or random_tree)(rng, n_leaves)
    seqs = (evolve_fn or evolve)(rng, tree, naive)
    ids = [str(i) for i in range(n_leaves)]
    import json
    doc = {"germline-info": {"locus": "igh"},
           "events": [{"unique_ids": ids, "naive_seq": naive, "input_seqs": [seqs[i] for i in ids],
                       "has_shm_indels": [False] * n_leaves,
                       "linearham-info": {"flexbounds": flex, "relpos": relpos}}]}
    os.makedirs(outdir, exist_ok=True)
    path = os.path.join(outdir, name + ".yaml")
    with open(path, "w") as f:
        json.dump(doc, f)
    return {"yaml_path": path, "tree": tree, "labels": ["naive"] + ids, "naive": naive,
            "flexbounds": flex, "relpos": relpos}
