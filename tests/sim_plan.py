"""Numpy emulation of the pruning program the CUDA kernel executes (REG + slot stack),
used on the CPU to check the traversal planner (linearham_b200/csrc/plan.cpp) against the
oracle port without a GPU. Mirrors k_prune in linearham_b200/csrc/kernels.cu step by step,
vectorised over xMSA columns. Test infrastructure only."""
import numpy as np

FACTOR = 2.0 ** 256
THRESH = 2.0 ** -256


def gtr_pmatrices(er, pi, rates, blens):
    """-> P[branch, k, 4, 4] via a symmetric eigendecomposition (independent of the product code)."""
    er = np.asarray(er, float)
    pi = np.asarray(pi, float)
    if er[5] > 0:
        er = er / er[5]
    q = np.zeros((4, 4))
    k = 0
    for i in range(4):
        for j in range(i + 1, 4):
            q[i, j] = q[j, i] = er[k] * np.sqrt(pi[i] * pi[j])
            q[i, i] -= er[k] * pi[j]
            q[j, j] -= er[k] * pi[i]
            k += 1
    q /= np.sum(pi * -np.diag(q))
    w, v = np.linalg.eigh(q)
    sq = np.sqrt(pi)
    out = np.zeros((len(blens), len(rates), 4, 4))
    for b, t in enumerate(blens):
        for k, r in enumerate(rates):
            if t == 0:
                out[b, k] = np.eye(4)
            else:
                out[b, k] = np.eye(4) + ((v / sq[:, None]) * np.expm1(w * r * t)) @ (v.T * sq[None, :])
    return out


def run_program(steps, pblen, tip_codes, er, pi, rates, n_slots):
    """steps [T-1,4], pblen [2T-3], tip_codes [T,S] -> xmsa_emission [S], max slot used."""
    T, S = tip_codes.shape
    R = len(rates)
    P = gtr_pmatrices(er, pi, rates, pblen)
    reg = np.zeros((S, R, 4))
    reg_cnt = np.zeros(S, dtype=np.int64)
    slots = {}

    def message(src, branch):
        if src >= 0:
            codes = tip_codes[src]
            m = np.zeros((S, R, 4))
            for k in range(R):
                pk = P[branch, k]
                table = np.concatenate([pk.T, pk.sum(axis=1)[None, :]], axis=0)  # [5, 4]: code -> P[:, code]
                m[:, k, :] = table[codes]
            return m, np.zeros(S, dtype=np.int64)
        if src == -1:
            clv, cnt = reg, reg_cnt
        else:
            clv, cnt = slots[-2 - src]
        return np.einsum("kij,skj->ski", P[branch], clv), cnt

    n_steps = T - 2
    for s in range(n_steps):
        a, b, dst, _ = steps[s]
        ma, ca = message(a, 2 * s)
        mb, cb = message(b, 2 * s + 1)
        new = ma * mb
        cnt = ca + cb
        small = np.all(new.reshape(S, -1) < THRESH, axis=1)
        new[small] *= FACTOR
        cnt = cnt + small
        reg, reg_cnt = new, cnt
        if dst >= 0:
            assert dst < n_slots
            slots[dst] = (reg.copy(), reg_cnt.copy())
    a, b, _, _ = steps[n_steps]
    mc, cc = message(b, 2 * n_steps)
    clvp, cp = (reg, reg_cnt) if a == -1 else slots[-2 - a]
    pi = np.asarray(pi, float)
    terma_r = np.sum(pi[None, None, :] * mc * clvp, axis=2)  # [S, R]
    terma = np.sum(np.where(terma_r > 0, terma_r, 0.0) / R, axis=1)
    with np.errstate(divide="ignore"):
        lk = np.log(terma) + (cp + cc) * np.log(THRESH)
    nb = tip_codes[0]
    corr = np.where(nb != 4, np.log(pi[np.minimum(nb, 3)]), 0.0)
    return np.exp(lk - corr)
