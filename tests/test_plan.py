"""CPU tests of the traversal planner: the register/stack program must (1) be well formed and
(2) reproduce the oracle's xmsa_emission when executed by the numpy emulation of the kernel."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_golden, rel_err
from linearham_b200 import _abi, synth
from oracle import pytrees as pt
import sim_plan


def plan(n_tips, tb):
    lib = C.CDLL(_abi.LIB_PATH)
    lib.lhh_last_error.restype = C.c_char_p
    B = tb["root_child"].shape[0]
    ch = np.ascontiguousarray(tb["children"], dtype=np.int32)
    bl = np.ascontiguousarray(tb["blen"], dtype=np.float64)
    rc = np.ascontiguousarray(tb["root_child"], dtype=np.int32)
    rb = np.ascontiguousarray(tb["root_blen"], dtype=np.float64)
    steps = np.zeros((B, n_tips - 1, 4), dtype=np.int32)
    pblen = np.zeros((B, 2 * n_tips - 3))
    mx = C.c_int32()
    p = lambda a, t: a.ctypes.data_as(C.POINTER(t))
    err = lib.lhh_plan_trees(n_tips, B, p(ch, C.c_int32), p(bl, C.c_double), p(rc, C.c_int32), p(rb, C.c_double),
                             p(steps, C.c_int32), p(pblen, C.c_double), C.byref(mx))
    if err:
        raise ValueError(lib.lhh_last_error().decode())
    return steps, pblen, mx.value


def check_wellformed(steps, n_tips, n_slots):
    T = n_tips
    live = {}
    reg_valid = False
    tips_seen = []
    for s, (a, b, dst, _) in enumerate(steps[:-1]):
        for src in (a, b):
            if src >= 0:
                tips_seen.append(src)
            elif src == -1:
                assert reg_valid, f"step {s} reads REG before it holds a result"
            else:
                assert live.pop(-2 - src, None) is not None, f"step {s} reads an empty slot"
        assert not (a == -1 and b == -1)
        reg_valid = True
        if dst >= 0:
            assert 0 <= dst < n_slots and dst not in live
            live[dst] = s
    a, b, _, _ = steps[-1]
    for src in (a, b):
        if src >= 0:
            tips_seen.append(src)
        elif src <= -2:
            assert live.pop(-2 - src, None) is not None
    assert not live, "parked results never consumed"
    assert sorted(tips_seen) == list(range(T)), "every tip must be consumed exactly once"


def test_plan_reproduces_golden_emissions(golden):
    name, fam, tb, exp, meta = golden
    steps, pblen, n_slots = plan(fam["n_tips"], tb)
    for t in range(steps.shape[0]):
        check_wellformed(steps[t], fam["n_tips"], n_slots)
        xe = sim_plan.run_program(steps[t], pblen[t], fam["tip_codes"], tb["er"][t], tb["pi"][t], tb["rates"][t], n_slots)
        assert rel_err(xe, exp["xmsa_emission"][t]) <= 1e-10, (name, t)


def _caterpillar(n, bad_order):
    # ((((t0,t1),t2),t3)...) or its mirror image; worst cases for a naive post-order
    s = "0:0.1"
    for i in range(1, n):
        s = f"({s},{i}:0.1):0.05" if not bad_order else f"({i}:0.1,{s}):0.05"
    return f"({s},naive:0.01,{n}:0.2);"


def _comb_of_cherries(n_pairs):
    s = "(0:0.1,1:0.1):0.05"
    for i in range(1, n_pairs):
        s = f"(({2 * i}:0.1,{2 * i + 1}:0.1):0.05,{s}):0.05"
    return f"({s},naive:0.01,{2 * n_pairs}:0.2);"


@pytest.mark.parametrize("newick,n_leaves,max_slots", [
    (_caterpillar(300, False), 301, 1), (_caterpillar(300, True), 301, 1),
    (_comb_of_cherries(200), 401, 2),
], ids=["caterpillar", "caterpillar_mirrored", "comb_of_cherries"])
def test_stack_depth_is_bounded(newick, n_leaves, max_slots):
    labels = ["naive"] + [str(i) for i in range(n_leaves)]
    tb = pt.flatten_batch([newick], labels, [[1] * 6], [[.25] * 4], [[1.0]])
    steps, pblen, n_slots = plan(len(labels), tb)
    check_wellformed(steps[0], len(labels), n_slots)
    assert n_slots <= max_slots


def test_random_trees_depth_and_wellformed():
    rng = np.random.default_rng(3)
    for n in (2, 3, 5, 17, 128, 1000):
        tree = synth.random_tree(rng, n)
        labels = ["naive"] + [str(i) for i in range(n)]
        samples = synth.tree_samples(rng, tree, 3)
        tb = pt.flatten_batch([s["newick"] for s in samples], labels, [s["er"] for s in samples],
                              [s["pi"] for s in samples], [[1.0]] * 3)
        steps, pblen, n_slots = plan(n + 1, tb)
        assert n_slots <= int(np.floor(np.log2(n + 1)))
        for t in range(3):
            check_wellformed(steps[t], n + 1, n_slots)
            assert np.isclose(pblen[t].sum(), tb["blen"][t].sum() + tb["root_blen"][t])


def test_malformed_tree_is_rejected():
    fam, tb, _, _ = load_golden("catch_a")
    bad = {k: v.copy() for k, v in tb.items()}
    bad["children"][0, 0, 0] = 99
    with pytest.raises(ValueError):
        plan(fam["n_tips"], bad)
    bad = {k: v.copy() for k, v in tb.items()}
    bad["children"][0, 1, 0] = bad["children"][0, 1, 1] = 4  # inner node consumed twice
    with pytest.raises(ValueError):
        plan(fam["n_tips"], bad)
    # a tip referenced twice (and so another one never): the reference refuses a tip-set mismatch in
    # SetTipStates (lib/libptpll/src/pll_partition.cpp:153-194)
    bad = {k: v.copy() for k, v in tb.items()}
    T = fam["n_tips"]
    tips = np.argwhere(bad["children"][0] < T)
    (o1, c1), (o2, c2) = tips[0], tips[1]
    bad["children"][0, o2, c2] = bad["children"][0, o1, c1]
    with pytest.raises(ValueError):
        plan(T, bad)
