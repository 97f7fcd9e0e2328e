"""TEST INFRASTRUCTURE ONLY - nothing under linearham_b200/ imports this.

CPU restatement (numpy) of the ancestral-state sampling step of the reference's R script
(scripts/run_bootstrap_asr_ess.R:81-89): per tree and alignment site, the rate category drawn in proportion to the
site's likelihood on the rate-scaled tree (phylomd::phylo.likelihood, R:84-87), then phylomd::asr.sim on that tree
(lib/phylomd/src/simulation_functions.cpp:57-117: post-order partial likelihoods, root state from pi * L_root, every
other inner node top-down from P(parent state -> s) * L_node[s]).

Parity status. R and phylomd are absent from this image and the script draws with R's generator on a worker cluster,
so the DRAWS of the reference cannot be reproduced by anything outside R: **parity with the reference's draws is
unpinned**. What is pinned: (1) the per-category site likelihoods against the compiled reference's goldens (their
category mean is xmsa_emission; tests/test_asr_oracle.py), (2) the sampling distribution against the exact posterior of
small trees by enumeration (exact_posterior below), (3) the random stream against the published Philox4x32-10
known-answer vectors. The CUDA path (linearham_b200/csrc/asr.cu) uses the same counters and the same order of
floating-point operations, so its draws equal this file's exactly (tests/test_gpu_asr.py).

The transition matrices here come from scipy.linalg.expm of the rate-scaled GTR generator - deliberately not the
eigen-decomposition the device uses - built as libpll and phylomd::GTR(scale = TRUE) build it (models.c:198-244)."""
import itertools

import numpy as np
from scipy.linalg import expm

LN2_256 = 177.445678223345993274


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    """Philox4x32-10 on arrays of uint32 counters (Salmon et al., SC'11) -> four uint32 arrays."""
    M0, M1, W0, W1 = 0xD2511F53, 0xCD9E8D57, 0x9E3779B9, 0xBB67AE85
    c0, c1, c2, c3 = (np.asarray(c, dtype=np.uint64) & 0xFFFFFFFF for c in np.broadcast_arrays(c0, c1, c2, c3))
    k0, k1 = int(k0) & 0xFFFFFFFF, int(k1) & 0xFFFFFFFF
    for _ in range(10):
        p0, p1 = M0 * c0, M1 * c2
        hi0, lo0, hi1, lo1 = p0 >> 32, p0 & 0xFFFFFFFF, p1 >> 32, p1 & 0xFFFFFFFF
        c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
        k0, k1 = (k0 + W0) & 0xFFFFFFFF, (k1 + W1) & 0xFFFFFFFF
    return tuple(c.astype(np.uint32) for c in (c0, c1, c2, c3))


def uniform(tree, site, draw, seed):
    x = philox4x32_10(np.uint32(tree), np.asarray(site, dtype=np.uint32), np.uint32(draw), np.uint32(0),
                      seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)[0]
    return (x.astype(np.float64) + 0.5) * (1.0 / 4294967296.0)


def gtr_generator(er, pi):
    """Q with Q[i][j] = r_ij * pi_j, rows summing to zero, scaled to one expected substitution per unit time."""
    r = np.asarray(er, dtype=np.float64)
    if r[5] > 0:
        r = r / r[5]
    pi = np.asarray(pi, dtype=np.float64)
    Q = np.zeros((4, 4))
    k = 0
    for i in range(4):
        for j in range(i + 1, 4):
            Q[i, j] = r[k] * pi[j]
            Q[j, i] = r[k] * pi[i]
            k += 1
    Q -= np.diag(Q.sum(axis=1))
    return Q / -(pi * np.diag(Q)).sum()


def _branches(tb, b):
    """[(parent op, child id, branch length)] in the device's order: 2*op + c, then the root edge."""
    ch, bl = np.asarray(tb["children"][b]), np.asarray(tb["blen"][b])
    n_ops = ch.shape[0]
    out = [(op, int(ch[op, c]), float(bl[op, c])) for op in range(n_ops) for c in range(2)]
    out.append((n_ops - 1, int(tb["root_child"][b]), float(tb["root_blen"][b])))
    return out


def _pmats(tb, b):
    Q = gtr_generator(tb["er"][b], tb["pi"][b])
    rates = np.asarray(tb["rates"][b], dtype=np.float64)
    return [[np.eye(4) if t == 0.0 else expm(Q * (r * t)) for r in rates] for (_, _, t) in _branches(tb, b)]


def _sum4(a, b, c, d):
    return ((a + b) + c) + d


def _tip_message(P, codes):
    m = np.empty((4, codes.shape[0]))
    rows = _sum4(P[:, 0], P[:, 1], P[:, 2], P[:, 3])
    known = codes < 4
    m[:, known] = P[:, codes[known]]
    m[:, ~known] = rows[:, None]
    return m


def _message(P, v):
    return np.stack([_sum4(P[i, 0] * v[0], P[i, 1] * v[1], P[i, 2] * v[2], P[i, 3] * v[3]) for i in range(4)])


def _draw4(w, u):
    c1 = w[0] + w[1]
    c2 = c1 + w[2]
    total = c2 + w[3]
    thr = u * total
    s = np.where(thr < w[0], 0, np.where(thr < c1, 1, np.where(thr < c2, 2, 3)))
    return np.where((total > 0) & np.isfinite(total), s, 4).astype(np.uint8)


def asr_sample(n_tips, tb, msa_codes, naive_codes, naive_row=0, seed=0):
    """-> dict(node_states [B][T-2][S] uint8, site_rate [B][S] uint8, site_rate_loglik [B][R][S])."""
    T = int(n_tips)
    msa = np.asarray(msa_codes, dtype=np.uint8)
    naive = np.asarray(naive_codes, dtype=np.uint8)
    B, S = naive.shape
    R = np.asarray(tb["rates"]).shape[1]
    n_ops = T - 2
    states = np.zeros((B, n_ops, S), dtype=np.uint8)
    site_rate = np.zeros((B, S), dtype=np.uint8)
    srl = np.zeros((B, R, S))
    sites = np.arange(S)
    for b in range(B):
        pm = _pmats(tb, b)
        br = _branches(tb, b)
        pi = np.asarray(tb["pi"][b], dtype=np.float64)
        codes = msa.copy()
        if naive_row >= 0:
            codes[naive_row] = naive[b]

        def upward(k, cols=slice(None)):
            clv = [None] * n_ops
            count = np.zeros(codes[0, cols].shape[0])
            for op in range(n_ops):
                node = None
                js = [2 * op, 2 * op + 1] + ([2 * n_ops] if op == n_ops - 1 else [])
                for jb in js:
                    cid = br[jb][1]
                    m = _tip_message(pm[jb][k], codes[cid, cols]) if cid < T else _message(pm[jb][k], clv[cid - T])
                    node = m if node is None else node * m
                mx = node.max(axis=0)
                while True:
                    small = (mx < 2.0 ** -256) & (mx > 0)
                    if not small.any():
                        break
                    node[:, small] *= 2.0 ** 256
                    mx[small] *= 2.0 ** 256
                    count[small] += 1
                clv[op] = node
            root = clv[n_ops - 1]
            L = _sum4(pi[0] * root[0], pi[1] * root[1], pi[2] * root[2], pi[3] * root[3])
            with np.errstate(divide="ignore"):
                return np.log(L) - count * LN2_256, clv

        logl = np.stack([upward(k)[0] for k in range(R)])                     # [R][S]
        with np.errstate(divide="ignore"):
            lp = np.where(naive[b] < 4, np.log(pi)[np.minimum(naive[b], 3)], 0.0)
        srl[b] = logl - lp
        best = logl.max(axis=0)
        ok = np.isfinite(best)
        with np.errstate(invalid="ignore"):
            w = np.exp(logl - best)
        total = np.zeros(S)
        for k in range(R):
            total = total + w[k]
        thr = uniform(b, sites, 0, seed) * total
        k_samp = np.full(S, R - 1)
        run = np.zeros(S)
        found = np.zeros(S, dtype=bool)
        for k in range(R):
            run = run + w[k]
            hit = (thr < run) & ~found
            k_samp[hit] = k
            found |= hit
        k_samp = np.where(ok, k_samp, R)
        site_rate[b] = k_samp
        states[b][:, ~ok] = 4
        for k in range(R):
            cols = np.flatnonzero(k_samp == k)
            if cols.size == 0:
                continue
            _, clv = upward(k, cols)
            st = np.zeros((n_ops, cols.size), dtype=np.uint8)
            root = clv[n_ops - 1]
            st[n_ops - 1] = _draw4([pi[i] * root[i] for i in range(4)], uniform(b, cols, n_ops, seed))
            for op in range(n_ops - 1, -1, -1):
                ps = st[op]
                js = ([2 * n_ops] if op == n_ops - 1 else []) + [2 * op, 2 * op + 1]
                for jb in js:
                    cid = br[jb][1]
                    if cid < T:
                        continue
                    node = cid - T
                    P = pm[jb][k]
                    prow = P[np.minimum(ps, 3)]                                # [n][4]
                    s = _draw4([prow[:, i] * clv[node][i] for i in range(4)], uniform(b, cols, 1 + node, seed))
                    st[node] = np.where(ps < 4, s, 4)
            states[b][:, cols] = st
    return {"node_states": states, "site_rate": site_rate, "site_rate_loglik": srl}


def exact_posterior(n_tips, tb, b, codes):
    """The joint distribution of (rate category, inner states) for one site by enumeration: P(k) ~ L_k, then the
    states given k from pi[root] * prod over branches - what sampling the category (R:84-87) and asr.sim define.
    `codes` are the T tip codes of the site. -> dict {(k, s_0, ..., s_{T-3}): probability}. Small trees only."""
    T = int(n_tips)
    n_ops = T - 2
    assert n_ops <= 5
    pm, br = _pmats(tb, b), _branches(tb, b)
    pi = np.asarray(tb["pi"][b], dtype=np.float64)
    R = np.asarray(tb["rates"]).shape[1]
    joint = {}
    for k in range(R):
        for st in itertools.product(range(4), repeat=n_ops):
            p = pi[st[n_ops - 1]]
            for jb, (op, cid, _) in enumerate(br):
                P = pm[jb][k]
                if cid >= T:
                    p *= P[st[op], st[cid - T]]
                else:
                    c = int(codes[cid])
                    p *= P[st[op], c] if c < 4 else P[st[op]].sum()
            joint[(k,) + st] = p
    z = sum(joint.values())
    return {key: v / z for key, v in joint.items()}
