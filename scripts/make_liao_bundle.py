"""BASELINE config 2 (Liao clonal family scored over a batch of 1024 tree samples): builds
tests/golden/liao_b1024.npz with the REFERENCE's own C++ (oracle/_ref). Build container only
(reads /root/reference); the bundle travels to the GPU box.

    python scripts/make_liao_bundle.py

The real inputs of this config are a partis annotation of data/liao_dataset.fasta (312 sequences x 363 nt)
and RevBayes tree samples. Neither partis nor RevBayes runs here, so this script makes
the STAND-IN SURVEY.md 8(c) prescribes, from real data wherever it exists:
  * germline genes and HMM parameters: the reference's partis parameter directory
    lib/partis/test/ref-results-slow/test/parameters/data/hmm/hmms (read in place by the reference's
    HMM constructor and by this repo's; nothing is copied);
  * annotation: ungapped alignment of the family consensus against every V / D / J allele of that
    directory: 3 V, 5 D and 2 J candidates, relpos from the alignments, flexbounds by the scheme of
    lib/partis/python/utils.py:3137-3310 as SURVEY.md 8(d) restates it (V/J bounds 10 wide, the D
    "germline" region one site); naive sequence = consensus with the V/D/J stretches set to germline;
  * trees: neighbour joining on Hamming distances (naive included, rooted at the naive tip's
    neighbour so that the naive tip hangs under the trifurcation as RevBayes' outgroup constraint keeps
    it), then a chain of 32 topologies (Poisson(3) NNIs per step, LogNormal(0, 0.3) branch noise)
    crossed with 128 exact power-of-two-table noise vectors -> 1024 distinct samples; er, pi ~
    Dirichlet(1), alpha = max(0.05, Exp(1)) fresh per sample (templates/revbayes_template.rev:23-66).
The reference scores all 1024 samples; its log-likelihoods (and full outputs of the first 4) are the
expected values. The bundle stores the 32 topologies, the noise vectors (int8 indices into a stored
table of doubles) and the per-sample parameters; tests rebuild the batch with exact IEEE products.
"""
import json
import os
import sys
import tempfile

import numpy as np
import yaml

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from linearham_b200 import host, synth, workload  # noqa: E402
from oracle import pyoracle as po, pytrees as pt  # noqa: E402

REF = "/root/reference/"
HMM_DIR = REF + "lib/partis/test/ref-results-slow/test/parameters/data/hmm/hmms"
FASTA = REF + "data/liao_dataset.fasta"
N_TOPO, N_NOISE, B, R = 32, 128, 1024, 4


class _Loader(yaml.SafeLoader):
    pass


_Loader.add_multi_constructor("tag:yaml.org,2002:python/object",
                              lambda loader, suffix, node: loader.construct_mapping(node, deep=True))


def germline_genes():
    """{gene: (type, sequence)} from the HMM YAMLs (states[].extras.germline of the <gene>_<pos> states)."""
    genes = {}
    for fn in sorted(os.listdir(HMM_DIR)):
        if not fn.endswith(".yaml"):
            continue
        doc = yaml.load(open(os.path.join(HMM_DIR, fn)), Loader=_Loader)
        name = doc["name"]
        seq = {}
        for st in doc["states"]:
            if st["name"].startswith(name + "_") and st.get("extras") and "germline" in st["extras"]:
                seq[int(st["name"][len(name) + 1:])] = st["extras"]["germline"]
        genes[name] = (name[3], "".join(seq[i] for i in range(len(seq))))
    return genes


def read_fasta(path):
    ids, seqs = [], []
    for line in open(path):
        line = line.strip()
        if line.startswith(">"):
            ids.append(line[1:].split()[0])
            seqs.append("")
        elif line:
            seqs[-1] += line.upper()
    return ids, seqs


def run_score(g, cons, relpos, lo, hi, reverse=False):
    """Best-scoring ungapped run of g placed at relpos, anchored at the 5' end (or 3' end when
    reverse) inside sites [lo, hi): +1 match, -3 mismatch. -> (score, first site, end site)."""
    idx = [s for s in range(max(lo, relpos), min(hi, relpos + len(g)))]
    if reverse:
        idx = idx[::-1]
    best, cur, stop = 0, 0, None
    for s in idx:
        cur += 1 if g[s - relpos] == cons[s] else -3
        if cur > best:
            best, stop = cur, s
    if stop is None:
        return 0, None, None
    return (best, stop, idx[0] + 1) if reverse else (best, idx[0], stop + 1)


def local_score(g, cons, relpos, lo, hi):
    """Best local ungapped segment (Kadane) of g at relpos inside [lo, hi). -> (score, start, end)."""
    best, cur, start, seg = 0, 0, None, (None, None)
    for s in range(max(lo, relpos), min(hi, relpos + len(g))):
        v = 1 if g[s - relpos] == cons[s] else -3
        if cur <= 0:
            cur, start = v, s
        else:
            cur += v
        if cur > best:
            best, seg = cur, (start, s + 1)
    return best, seg[0], seg[1]


def annotate(seqs, genes):
    """-> dict(naive, flexbounds, relpos, picks) for the family; see the module docstring."""
    L = len(seqs[0])
    arr = np.array([[("ACGT".find(c) if c in "ACGT" else 4) for c in s] for s in seqs])
    cons = "".join("ACGTN"[np.bincount(arr[:, j], minlength=5)[:4].argmax()] for j in range(L))
    by = {t: {g: s for g, (tt, s) in genes.items() if tt == t} for t in "VDJ"}
    # V: anchored at the 5' end of the read
    vs = []
    for g, s in by["V"].items():
        vs.append(max((run_score(s, cons, r, 0, L) + (r, g) for r in range(-40, 6)), key=lambda x: x[0]))
    vs.sort(key=lambda x: (-x[0], x[4]))
    vs = vs[:3]
    vend = vs[0][2]
    # J: anchored at the 3' end
    js = []
    for g, s in by["J"].items():
        js.append(max((run_score(s, cons, r, 0, L, reverse=True) + (r, g)
                       for r in range(L - len(s) - 5, L - len(s) + 12)), key=lambda x: x[0]))
    js.sort(key=lambda x: (-x[0], x[4]))
    js = js[:2]
    jstart = js[0][1]
    # D: best local segment between the V and J stretches
    ds = []
    for g, s in by["D"].items():
        ds.append(max((local_score(s, cons, r, vend - 2, jstart + 2) + (r, g)
                       for r in range(vend - len(s), jstart)), key=lambda x: x[0]))
    ds.sort(key=lambda x: (-x[0], x[4]))
    d0 = ds[0]
    dstart, dend = d0[1], d0[2]
    c = (dstart + dend) // 2
    flex = {"v_l": [0, max(0, vs[0][3])], "v_r": [vend - 20, vend - 10], "d_l": [c - 4, c], "d_r": [c + 1, c + 5],
            "j_l": [jstart + 4, jstart + 14], "j_r": [L, L]}
    assert flex["v_r"][1] < flex["d_l"][0] + 5 and flex["d_r"][1] <= flex["j_l"][1], flex
    relpos = {v[4]: int(v[3]) for v in vs}
    relpos.update({j[4]: int(j[3]) for j in js})
    relpos[d0[4]] = int(d0[3])
    assert d0[3] < c and d0[3] + len(by["D"][d0[4]]) - 1 >= c + 1
    # the other D candidates must start before d_l.second, reach d_r.first and start after v_r.first
    others = []
    for g, s in by["D"].items():
        if g == d0[4]:
            continue
        lo, hi = max(c + 1 - len(s) + 1, flex["v_r"][0] + 1), c - 1
        if lo > hi:
            continue
        others.append(max((local_score(s, cons, r, vend - 2, jstart + 2) + (r, g) for r in range(lo, hi + 1)),
                          key=lambda x: x[0]))
    others.sort(key=lambda x: (-x[0], x[4]))
    for o in others[:4]:
        relpos[o[4]] = int(o[3])
    naive = list(cons)
    v0, j0 = by["V"][vs[0][4]], by["J"][js[0][4]]
    for s in range(max(0, vs[0][3]), vend):
        naive[s] = v0[s - vs[0][3]]
    for s in range(dstart, dend):
        naive[s] = by["D"][d0[4]][s - d0[3]]
    for s in range(jstart, min(L, js[0][3] + len(j0))):
        naive[s] = j0[s - js[0][3]]
    picks = {"V": [(v[4], v[0], v[3]) for v in vs], "J": [(j[4], j[0], j[3]) for j in js],
             "D": [(d0[4], d0[0], d0[3])] + [(o[4], o[0], o[3]) for o in others[:4]],
             "vend": vend, "dseg": (dstart, dend), "jstart": jstart}
    # relpos is keyed by the partis gene names (src/VDJGermline.cpp:56-59)
    relpos = {g.replace("_star_", "*").replace("_slash_", "/"): r for g, r in relpos.items()}
    return {"naive": "".join(naive), "flexbounds": flex, "relpos": relpos, "picks": picks}


def nj_tree(labels, seqs, floor=1e-4):
    """Neighbour joining on Hamming distances; rooted at the neighbour of labels[0] ('naive'), whose tip
    is the last child of the trifurcation. -> synth.Tree"""
    arr = np.array([[("ACGT".find(c) if c in "ACGT" else 4) for c in s] for s in seqs], dtype=np.int8)
    n = len(labels)
    D = np.zeros((n, n))
    for i in range(n):
        D[i] = (arr != arr[i]).mean(axis=1)
    node = list(range(n))          # matrix row -> node id
    adj = {i: [] for i in range(n)}
    nxt = n
    while len(node) > 3:
        m = len(node)
        r = D.sum(axis=1)
        Q = (m - 2) * D - r[:, None] - r[None, :]
        np.fill_diagonal(Q, np.inf)
        i, j = np.unravel_index(np.argmin(Q), Q.shape)
        i, j = min(i, j), max(i, j)
        di = 0.5 * D[i, j] + (r[i] - r[j]) / (2 * (m - 2))
        dj = D[i, j] - di
        u = nxt
        nxt += 1
        adj[u] = [(node[i], di), (node[j], dj)]
        adj[node[i]].append((u, di))
        adj[node[j]].append((u, dj))
        du = 0.5 * (D[i] + D[j] - D[i, j])
        D[i] = du
        D[:, i] = du
        D[i, i] = 0.0
        D = np.delete(np.delete(D, j, 0), j, 1)
        node[i] = u
        del node[j]
    a, b, c = node
    u = nxt
    adj[u] = []
    for x, d in ((a, (D[0, 1] + D[0, 2] - D[1, 2]) / 2), (b, (D[0, 1] + D[1, 2] - D[0, 2]) / 2),
                 (c, (D[0, 2] + D[1, 2] - D[0, 1]) / 2)):
        adj[u].append((x, d))
        adj[x].append((u, d))
    root = adj[0][0][0]  # the naive tip's neighbour
    t = synth.Tree()
    order = [x for x in adj[root] if x[0] != 0] + [x for x in adj[root] if x[0] == 0]
    stack = [(x, d, 0, root) for x, d in reversed(order)]
    while stack:
        x, d, parent, came = stack.pop()
        me = t.add(parent, max(float(d), floor), labels[x] if x < n else None)
        for y, dy in reversed([e for e in adj[x] if e[0] != came]):
            stack.append((y, dy, me, x))
    return t


def ops_newick(children, blen, root_child, root_blen, labels):
    """Newick string of one flattened tree (the doubles survive repr() exactly)."""
    T = len(labels)
    text = {}
    for i in range(T - 2):
        parts = []
        for k in range(2):
            ch = int(children[i, k])
            parts.append((labels[ch] if ch < T else text.pop(ch)) + ":" + repr(float(blen[i, k])))
        text[T + i] = "(" + ",".join(parts) + ")"
    top = text.pop(2 * T - 3)
    rc = int(root_child)
    third = (labels[rc] if rc < T else text.pop(rc)) + ":" + repr(float(root_blen))
    assert not text
    return top[:-1] + "," + third + ");"


def main():
    sys.path.insert(0, os.path.join(ROOT, "tests", "golden"))
    from make_golden import ref_outputs
    if not po.have_ref():
        po.build_ref()
    genes = germline_genes()
    ids, seqs = read_fasta(FASTA)
    ann = annotate(seqs, genes)
    print("stand-in annotation:", json.dumps(ann["picks"]), json.dumps(ann["flexbounds"]))
    doc = {"germline-info": {"locus": "igh"},
           "events": [{"unique_ids": ids, "naive_seq": ann["naive"], "input_seqs": seqs,
                       "has_shm_indels": [False] * len(ids),
                       "linearham-info": {"flexbounds": ann["flexbounds"], "relpos": ann["relpos"]}}]}
    rng = np.random.default_rng(2002)
    with tempfile.TemporaryDirectory() as d:
        ypath = os.path.join(d, "liao.yaml")
        json.dump(doc, open(ypath, "w"))
        ref = po.RefPhyloHMM(ypath, 0, HMM_DIR, 0)
        fam = ref.family()
        labels = ref.strs("xmsa_labels")
        assert labels == ["naive"] + ids, labels[:3]
        # this repo's host class builds the same family from the same files (bit-exact)
        mine = host.PhyloHMM(ypath, 0, HMM_DIR, 0)
        fm = mine.family()
        for k, v in fam.items():
            assert np.array_equal(np.asarray(v), np.asarray(fm[k])), k
        mine.close()
        tree = nj_tree(labels, [ann["naive"]] + seqs)
        samples = synth.tree_samples(rng, tree, N_TOPO, outgroup="naive")
        T = fam["n_tips"]
        er = rng.dirichlet(np.ones(6), B)
        pi = np.maximum(rng.dirichlet(np.ones(4), B), 0.01)
        pi /= pi.sum(axis=1, keepdims=True)
        alpha = np.maximum(0.05, rng.exponential(1.0, B))
        rates = np.array([ref.gamma_cats(a, R) for a in alpha])
        base = pt.flatten_batch([s["newick"] for s in samples], labels, er[:N_TOPO], pi[:N_TOPO], rates[:N_TOPO])
        z = {"base_children": base["children"], "base_blen": base["blen"], "base_root_child": base["root_child"],
             "base_root_blen": base["root_blen"],
             "noise_table": np.exp2(np.arange(-128, 128) / 64.0),
             "noise_k": np.clip(np.rint(rng.normal(0, 0.1 * 64 / np.log(2), (N_NOISE, T - 2, 2))), -128, 127).astype(np.int8),
             "noise_root_k": np.clip(np.rint(rng.normal(0, 0.1 * 64 / np.log(2), N_NOISE)), -128, 127).astype(np.int8),
             "topo_of": (np.arange(B) % N_TOPO).astype(np.int32),
             "noise_of": (((np.arange(B) // N_TOPO) * 4 + np.arange(B) % 4) % N_NOISE).astype(np.int32),
             "er": er, "pi": pi, "rates": rates, "alpha": alpha}
        assert len({(int(a), int(b)) for a, b in zip(z["topo_of"], z["noise_of"])}) == B
        tb = workload.liao_batch(z)
        ll, exp = [], {}
        for b in range(B):
            nw = ops_newick(tb["children"][b], tb["blen"][b], tb["root_child"][b], tb["root_blen"][b], labels)
            ll.append(ref.score_newick(nw, er[b], pi[b], alpha[b], R))
            if b < 4:
                for k, v in ref_outputs(ref, fam["has_d"]).items():
                    exp.setdefault(k, []).append(v)
            if b % 128 == 0:
                print("reference scored", b, ll[-1], flush=True)
        data = {"fam_" + k: np.asarray(v) for k, v in fam.items()}
        data.update(z)
        data.update({"exp_" + k: np.stack(v) for k, v in exp.items()})
        data["exp_loglik"] = np.array(ll)
        data["labels"] = np.array(labels)
        data["naive"] = np.array(ann["naive"])
        data["annotation"] = np.array(json.dumps({"flexbounds": ann["flexbounds"], "relpos": ann["relpos"],
                                                   "picks": ann["picks"]}))
        path = os.path.join(ROOT, "tests", "golden", "liao_b1024.npz")
        np.savez_compressed(path, **data)
        print(f"liao_b1024: T={T} S_x={fam['n_sites']} vd={fam['vd_sites']}x{fam['vd_states']} "
              f"dj={fam['dj_sites']}x{fam['dj_states']} B={B} loglik[0]={ll[0]!r} finite={np.isfinite(ll).mean():.3f} "
              f"-> {os.path.getsize(path)} bytes")


if __name__ == "__main__":
    main()
