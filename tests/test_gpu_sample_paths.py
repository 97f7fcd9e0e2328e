"""lh_sample_paths (k_sample): the index draws of HMM::SampleNaiveSequence (src/HMM.cpp:323-341, 1222-1353) on the
device, against a restatement here of what libstdc++'s std::discrete_distribution / generate_canonical compute
(bits/random.tcc: weights summed left to right, divided by the sum, running sum with the last entry forced to 1.0,
u = (x0 + x1 * 2^32) / 2^64, lower_bound) over the DENSE transition columns and the forward tables the same run gave.
The end-to-end check - the reference's own RunPipeline output, byte for byte - is tests/test_gpu_phylohmm.py."""
import numpy as np
import pytest

from conftest import load_golden, SCALING_NAMES
from linearham_b200 import device

pytestmark = pytest.mark.gpu


def _draw(weights, x0, x1):
    n = len(weights)
    if n < 2:
        return 0, 0
    s = 0.0
    for w in weights:
        s += float(w)
    u = (float(x0) + float(x1) * 4294967296.0) / 18446744073709551616.0
    if u >= 1.0:
        u = np.nextafter(1.0, 0.0)
    cp, out = 0.0, []
    for w in weights:
        cp += float(w) / s
        out.append(cp)
    out[-1] = 1.0
    lo, length = 0, n                       # std::lower_bound
    while length > 0:
        half = length >> 1
        if out[lo + half] < u:
            lo += half + 1
            length -= half + 1
        else:
            length = half
    return lo, 2


def _paths(fam, tables, words):
    """tables: dict of the run's forward read-outs; -> idx like lh_sample_paths lays it out."""
    has_d = int(fam["has_d"])
    n_v, n_d, n_j = int(fam["n_v"]), int(fam["n_d"]), int(fam["n_j"])
    Lvd, Svd, Ldj, Sdj = int(fam["vd_sites"]), int(fam["vd_states"]), int(fam["dj_sites"]), int(fam["dj_states"])
    mat = lambda name, r, c: np.asarray(fam[name]).reshape(c, r).T          # column-major -> [r][c]
    B = words.shape[0]
    out = []
    for b in range(B):
        w, at, idx = words[b], 0, []

        def draw(weights):
            nonlocal at
            k, used = _draw(weights, int(w[at]) if at < len(w) else 0, int(w[at + 1]) if at + 1 < len(w) else 0)
            at += used
            return k
        germ = draw(tables["jgerm_forward"][b])
        idx.append(germ)
        for dj in ([True, False] if has_d else [False]):
            L, S = (Ldj, Sdj) if dj else (Lvd, Svd)
            fw = np.asarray(tables["dj_forward" if dj else "vd_forward"][b]).reshape(S, L).T        # [site][state]
            t_out = mat("t_dj_j", Sdj, n_j) if dj else mat("t_vd_d", Svd, n_d if has_d else n_j)
            t_jj = mat("t_dj", Sdj, Sdj) if dj else mat("t_vd", Svd, Svd)
            states, prev = [0] * L, germ
            for i in range(L - 1, -1, -1):
                col = t_out[:, prev] if i == L - 1 else t_jj[:, prev]
                prev = draw(col * fw[i])
                states[i] = prev
            idx += states
            if dj:
                germ = draw(mat("t_d_dj", n_d, Sdj)[:, prev] * tables["dgerm_forward"][b])
            else:
                germ = draw(mat("t_v_vd", n_v, Svd)[:, prev] * tables["vgerm_forward"][b])
            idx.append(germ)
        assert at == len(w)
        out.append(idx)
    return np.array(out, dtype=np.int32)


def _n_words(fam):
    d = lambda n, calls: 2 * calls if n >= 2 else 0
    n = d(int(fam["n_j"]), 1) + d(int(fam["vd_states"]), int(fam["vd_sites"])) + d(int(fam["n_v"]), 1)
    if int(fam["has_d"]):
        n += d(int(fam["dj_states"]), int(fam["dj_sites"])) + d(int(fam["n_d"]), 1)
    return n


@pytest.mark.parametrize("name", ["catch_a", "synth_n12", "synth_igk_n30", "synth_n2"])
def test_draws_equal_the_restated_standard_library(name):
    fam, tb, exp, _ = load_golden(name)
    B = len(tb["root_child"])
    f = device.Family(fam)
    ll, tables = f.score(tb, outputs=True)          # the read-outs keep the tables on the device as well
    rng = np.random.default_rng(4)
    words = rng.integers(0, 2 ** 32, size=(B, _n_words(fam)), dtype=np.uint64).astype(np.uint32)
    words[0, :] = 0                                   # u = 0: the first state, whatever its weight
    if B > 1:
        words[1, :] = 0xFFFFFFFF                      # u rounds to 1: nextafter(1, 0)
    idx, flags = f.sample_paths(words)
    again, _ = f.sample_paths(words)
    f.close()
    assert np.array_equal(idx, again)
    # (the all-zero row picks state 0 whatever its weight; if nothing leads on from there the weights of the next draw
    # sum to zero and the tree is handed to the host - every other row must be drawn here)
    assert not flags[2:].any()
    ok = flags == 0
    want = _paths(fam, {k: np.asarray(v)[ok] for k, v in tables.items()}, words[ok])
    assert np.array_equal(idx[ok], want)


def test_trees_with_non_finite_tables_are_left_to_the_host():
    name = [n for n in SCALING_NAMES if "cat3000_a005" in n or "cat1000" in n][0]
    fam, tb, exp, _ = load_golden(name)
    B = len(tb["root_child"])
    f = device.Family(fam)
    ll, tables = f.score(tb, outputs=True)
    words = np.random.default_rng(5).integers(0, 2 ** 32, size=(B, _n_words(fam)), dtype=np.uint64).astype(np.uint32)
    idx, flags = f.sample_paths(words)
    f.close()
    finite = np.array([all(np.isfinite(np.asarray(tables[k][b])).all() for k in tables if k.endswith("_forward")) for b in range(B)])
    assert (flags[~finite] == 1).all()
    ok = flags == 0
    if ok.any():
        want = _paths(fam, {k: np.asarray(v)[ok] for k, v in tables.items()}, words[ok])
        assert np.array_equal(idx[ok], want)


def test_arguments_are_checked():
    fam, tb, exp, _ = load_golden("synth_n12")
    B = len(tb["root_child"])
    f = device.Family(fam)
    f.score(tb)                                                       # no tables kept
    words = np.random.default_rng(6).integers(0, 2 ** 32, size=(B, _n_words(fam)), dtype=np.uint64).astype(np.uint32)
    with pytest.raises(device.DeviceError, match="keep_tables"):
        f.sample_paths(words)
    f.set_option("keep_tables", 1)
    f.score(tb)
    idx, flags = f.sample_paths(words)
    assert idx.shape[0] == B and not flags.any()
    with pytest.raises(device.DeviceError, match="generator outputs"):
        f.sample_paths(words[:, :-2])
    with pytest.raises(device.DeviceError, match="keep"):
        f.sample_paths(words[:-1])
    f.close()
