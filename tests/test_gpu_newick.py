"""The device Newick parser (k_newick, csrc/newick_dev.cu) and lh_score_newick against the host parser
(host/newick.cpp - itself pinned to the compiled reference's flattened trees, test_host_ingest.py): the
operation arrays must be bit-identical, trees the device leaves to the host must come out the same, and a
malformed tree must fail the call with the host parser's message."""
import re

import numpy as np
import pytest

from conftest import load_golden, load_liao, GOLDEN_NAMES, rel_err
from linearham_b200 import device, host, synth

pytestmark = pytest.mark.gpu
OPS = ("children", "blen", "root_child", "root_blen")


KERNELS = [1, 2, 3]   # k_newick (one thread per tree), k_newick_cta (one CTA per tree; the default where it fits),
                      # k_newick_warp (one warp per tree)


def _family(fam, labels, kernel=0):
    f = device.Family(fam)
    f.set_labels(labels)
    f.set_option("newick_kernel", kernel)
    return f


def _on_patterns(ops):
    return ((ops["root_child"] == 0) | (ops["children"][:, -1, 0] == 0) | (ops["children"][:, -1, 1] == 0)).astype(np.uint8)


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("name", GOLDEN_NAMES)
def test_ops_bit_identical_on_goldens(name, kernel):
    fam, tb, exp, meta = load_golden(name)
    nw, labels = [str(s) for s in meta["newicks"]], [str(s) for s in meta["labels"]]
    f = _family(fam, labels, kernel)
    got = f.parse_newick(nw)
    assert not got["left_to_host"].any()
    for k in OPS:
        assert np.array_equal(got[k], tb[k]), k
    assert np.array_equal(got["on_patterns"], _on_patterns(tb))
    # and the scores: text in == arrays in, bit for bit; both match the reference
    ll_text, out_text = f.score_newick(nw, tb["er"], tb["pi"], tb["rates"].shape[1], alpha=meta["alpha"], outputs=True)
    tm = f.timings()
    ll_arr, out_arr = f.score(tb, outputs=True)
    f.close()
    assert tm["newick_host_trees"] == 0 and tm["ms_newick"] > 0
    assert np.array_equal(ll_text, ll_arr, equal_nan=True)
    for k in out_arr:
        assert np.array_equal(out_text[k], out_arr[k], equal_nan=True), k
    assert rel_err(ll_text, exp["loglik"]) <= 1e-9


def test_liao_family_real_labels():
    fam, tb, exp, z = load_liao()
    labels = [str(s) for s in z["labels"]]
    text = host.NewickBuffer(tb, labels)   # the bundle stores arrays; write them out as RevBayes would
    nw = [text[i] for i in range(len(text))]
    ref = host.newick_to_ops(nw, labels)
    for k in OPS:
        assert np.array_equal(ref[k], tb[k]), k   # the strings say what the bundle holds
    f = _family(fam, labels)
    got = f.parse_newick(nw)
    assert not got["left_to_host"].any()
    for k in OPS:
        assert np.array_equal(got[k], tb[k]), k
    ll = f.score_newick(nw, tb["er"], tb["pi"], tb["rates"].shape[1], rates=tb["rates"], lengths=True)
    f.close()
    assert rel_err(ll, exp["loglik"]) <= 1e-9


def _decorate(rng, nwk):
    """The same tree spelled differently: [&index=N] comments, blanks, exponent notation, missing and zero
    lengths (-> EPS), inner labels, a length on the trifurcation."""
    out, i, n = [], 0, len(nwk)
    while i < n:
        c = nwk[i]
        if c == ":":
            j = i + 1
            while j < n and nwk[j] not in ",);":
                j += 1
            v = float(nwk[i + 1:j])
            r = rng.random()
            if r < 0.08:
                s = ""                                  # missing
            elif r < 0.12:
                s = ":0" if rng.random() < 0.5 else ":0.0"
            elif r < 0.35:
                s = f":{v:.{int(rng.integers(1, 17))}e}"
            elif r < 0.5:
                s = f":{v:.{int(rng.integers(3, 18))}f}"
            else:
                s = ":" + repr(v)
            if rng.random() < 0.2:
                s = f"[&index={int(rng.integers(0, 999))}]" + s
            elif rng.random() < 0.04:   # a long comment holding everything a lexer could trip over, '[' included
                s = "[&note=((a:1,b:2)x:3,[c);" + " pad" * int(rng.integers(0, 30)) + "]" + s
            if rng.random() < 0.05 and s:
                s = s[0] + " " + s[1:]
            out.append(s)
            i = j
            continue
        out.append(c)
        if c == ")" and rng.random() < 0.15:
            out.append(f"n{int(rng.integers(0, 99))}")
        if c in ",(" and rng.random() < 0.05:
            out.append(" ")
        i += 1
    s = "".join(out)
    if rng.random() < 0.3:
        s = s[:-1] + ":0.5;"
    return s


@pytest.mark.parametrize("kernel", KERNELS)
@pytest.mark.parametrize("n_leaves,n_trees", [(3, 400), (7, 1500), (57, 4000), (200, 3000), (1000, 1200), (3000, 64)])
def test_ops_bit_identical_on_random_trees(n_leaves, n_trees, kernel):
    rng = np.random.default_rng(100 + n_leaves)
    labels = ["naive"] + [f"seq_{i:04d}_long" if i % 3 == 0 else str(i) for i in range(n_leaves)]
    nw = []
    while len(nw) < n_trees:
        tree = synth.random_tree(rng, n_leaves)
        for s in synth.tree_samples(rng, tree, 50, outgroup=None if len(nw) % 100 else "naive"):
            nw.append(s["newick"])
    nw = nw[:n_trees]
    # rename the tips (synth writes them as their index) to the labels above
    import re
    ren = {str(i): labels[i + 1] for i in range(n_leaves)}
    nw = [re.sub(r"(?<=[(,])(\d+)(?=[:,)\[])", lambda m: ren[m.group(1)], s) for s in nw]
    nw = [_decorate(rng, s) if i % 2 else s for i, s in enumerate(nw)]
    fam, _, _, _ = load_golden("synth_n12")  # any bound family will do for parsing only ...
    ref = host.newick_to_ops(nw, labels)
    fam = dict(fam)
    # ... as long as it has the right number of tips: parse_newick only reads n_tips
    fam["n_tips"] = n_leaves + 1
    fam["tip_codes"] = np.zeros((n_leaves + 1, fam["n_sites"]), dtype=np.uint8)
    f = _family(fam, labels, kernel)
    got = f.parse_newick(nw)
    f.close()
    assert not got["left_to_host"].any()
    for k in OPS:
        assert np.array_equal(got[k], ref[k]), k
    assert np.array_equal(got["on_patterns"], _on_patterns(ref))


def test_deep_caterpillar_goes_through_the_second_device_pass():
    """A caterpillar of 600 tips nests 598 deep - beyond the per-level tables of k_newick_cta (256): the tree is
    handed to k_newick on the device, not to the host, and comes out the same."""
    T = 601
    labels = ["naive"] + [f"t{i}" for i in range(1, T)]
    n_ops = T - 2
    ch = np.zeros((3, n_ops, 2), dtype=np.int32)
    ch[:, 0] = (1, 2)
    for op in range(1, n_ops - 1):
        ch[:, op] = (T + op - 1, op + 2)
    ch[:, n_ops - 1] = (T + n_ops - 2, T - 1)
    rng = np.random.default_rng(5)
    tb = {"children": ch, "blen": rng.uniform(1e-4, 0.1, (3, n_ops, 2)), "root_child": np.zeros(3, dtype=np.int32),
          "root_blen": rng.uniform(1e-4, 0.1, 3)}
    text = host.NewickBuffer(tb, labels)
    nw = [text[i] for i in range(3)]
    ref = host.newick_to_ops(nw, labels)
    fam = dict(load_golden("synth_n12")[0])
    fam["n_tips"] = T
    fam["tip_codes"] = np.zeros((T, fam["n_sites"]), dtype=np.uint8)
    f = _family(fam, labels, 2)
    got = f.parse_newick(nw)
    f.close()
    assert not got["left_to_host"].any()
    for k in OPS:
        assert np.array_equal(got[k], ref[k]), k


def test_lenient_spellings_are_left_to_the_host_by_the_cta_parser():
    """What the lenient host parser takes beyond the strict grammar - a missing comma, text after ';' - the
    one-CTA-per-tree parser declines (the one-thread-per-tree parser reads them as the host does); the call's
    results are the host parser's either way."""
    fam, tb, exp, meta = load_golden("synth_n12")
    nw, labels = [str(s) for s in meta["newicks"]], [str(s) for s in meta["labels"]]
    odd = list(nw)
    odd[0] = nw[0] + " trailing text (ignored"
    i = nw[1].index(",")
    odd[1] = nw[1][:i] + " " + nw[1][i + 1:]
    odd[2] = "  \t" + nw[2].replace(",", " ,\n", 3)          # blanks are fine everywhere
    ref = host.newick_to_ops(odd, labels)
    R = tb["rates"].shape[1]
    for kernel, left in ((1, []), (2, [0, 1]), (3, [0, 1])):
        f = _family(fam, labels, kernel)
        got = f.parse_newick(odd)
        assert np.flatnonzero(got["left_to_host"]).tolist() == left
        keep = got["left_to_host"] == 0
        for k in OPS:
            assert np.array_equal(got[k][keep], ref[k][keep]), k
        ll = f.score_newick(odd, tb["er"], tb["pi"], R, alpha=meta["alpha"])
        assert f.timings()["newick_host_trees"] == len(left)
        ll_arr = f.score(dict(tb, **ref))
        f.close()
        assert np.array_equal(ll, ll_arr, equal_nan=True)


@pytest.mark.parametrize("kernel", KERNELS)
def test_hard_spellings_go_to_the_host_parser_and_score_the_same(kernel):
    fam, tb, exp, meta = load_golden("synth_n12")
    nw, labels = [str(s) for s in meta["newicks"]], [str(s) for s in meta["labels"]]
    B = len(nw)
    hard = list(nw)
    # the same values spelled so that the device parser declines: 25 significant digits, '+', hex float
    def respell(s, how):
        i = s.index(":")
        j = i + 1
        while s[j] not in ",);":
            j += 1
        v = float(s[i + 1:j])
        return s[:i + 1] + {"long": f"{v:.30f}", "plus": "+" + repr(v), "hex": v.hex()}[how] + s[j:]
    hard[1], hard[3], hard[B - 1] = respell(nw[1], "long"), respell(nw[3], "plus"), respell(nw[B - 1], "hex")
    ref = host.newick_to_ops(hard, labels)
    f = _family(fam, labels, kernel)
    got = f.parse_newick(hard)
    assert got["left_to_host"].tolist() == [1 if i in (1, 3, B - 1) else 0 for i in range(B)]
    keep = got["left_to_host"] == 0
    for k in OPS:
        assert np.array_equal(got[k][keep], ref[k][keep]), k
    R = tb["rates"].shape[1]
    ll, out = f.score_newick(hard, tb["er"], tb["pi"], R, alpha=meta["alpha"], outputs=True)
    tm = f.timings()
    tb2 = dict(tb, **ref)
    ll_arr, out_arr = f.score(tb2, outputs=True)
    f.close()
    assert tm["newick_host_trees"] == 3
    assert np.array_equal(ll, ll_arr, equal_nan=True)
    for k in out_arr:
        assert np.array_equal(out[k], out_arr[k], equal_nan=True), k


@pytest.mark.parametrize("kernel", KERNELS)
def test_malformed_trees_fail_with_the_host_parsers_message(kernel):
    fam, tb, exp, meta = load_golden("synth_n12")
    nw, labels = [str(s) for s in meta["newicks"]], [str(s) for s in meta["labels"]]
    R = tb["rates"].shape[1]
    f = _family(fam, labels, kernel)
    cases = {
        "unknown": nw[2].replace("," + labels[3] + "[", ",nobody[", 1).replace("(" + labels[3] + "[", "(nobody[", 1),
        "twice": nw[2].replace("," + labels[3] + "[", "," + labels[4] + "[", 1).replace("(" + labels[3] + "[", "(" + labels[4] + "[", 1),
        "nosemi": nw[2].rstrip(";"),
        "badlen": nw[2].replace(":", ":x", 1),
        "unbalanced": nw[2].replace("(", "", 1),
        # the right number of ')' but a tip short (a unary node): the item list is one short of what S5 / S6 expect
        "unary": re.sub(r",[^,():;\[\]\s]+(\[[^\]]*\])?(:[0-9.eE+-]+)?\)", ")", nw[2], count=1),
        "two_words": nw[2].replace("naive[", "na[x]ive[", 1),
        "empty": "",
        "negative": None,
    }
    for what, bad in cases.items():
        batch = list(nw)
        if what == "negative":
            i = nw[5].index(":")
            batch[5] = nw[5][:i + 1] + "-" + nw[5][i + 1:]
            with pytest.raises(device.DeviceError, match="tree 5: negative branch length"):
                f.score_newick(batch, tb["er"], tb["pi"], R, alpha=meta["alpha"])
            continue
        batch[2] = bad
        assert bad != nw[2], what
        with pytest.raises(host.HostError) as host_err:
            host.newick_to_ops([bad], labels)
        want = str(host_err.value).replace("tree 0: ", "tree 2: ")
        with pytest.raises(device.DeviceError) as dev_err:
            f.score_newick(batch, tb["er"], tb["pi"], R, alpha=meta["alpha"])
        assert str(dev_err.value) == want, what
    # the handle is fine afterwards
    assert rel_err(f.score_newick(nw, tb["er"], tb["pi"], R, alpha=meta["alpha"]), exp["loglik"]) <= 1e-9
    with pytest.raises(device.DeviceError, match="alpha"):
        f.score_newick(nw, tb["er"], tb["pi"], R, alpha=np.full(len(nw), 0.001))
    f.close()


def test_many_chunks_and_segments_match_the_array_path():
    """A batch long enough for several chunks (and k_split lists on the device) with a mix of trees that run
    on site patterns and trees that do not."""
    fam, tb, exp, meta = load_golden("synth_n40")
    nw, labels = [str(s) for s in meta["newicks"]], [str(s) for s in meta["labels"]]
    rep = 60
    big_nw = nw * rep
    big = {k: np.concatenate([v] * rep) for k, v in tb.items()}
    f = _family(fam, labels)
    f.set_option("chunk_trees", 97)
    a = f.score_newick(big_nw, big["er"], big["pi"], big["rates"].shape[1], rates=big["rates"])
    tm = f.timings()
    b = f.score(big)
    tm_b = f.timings()
    f.close()
    assert tm["prune_launches"] >= 6 and tm["pattern_trees"] == tm_b["pattern_trees"]
    assert np.array_equal(a, b, equal_nan=True)
    assert rel_err(a[:len(nw)], exp["loglik"]) <= 1e-9


def test_pinned_text_is_copied_in_place_and_scores_the_same():
    """Strings the caller holds in pinned memory at a constant stride, with their lengths (host.NewickBuffer(pinned=
    True)), go to the device with one 2-D copy per chunk instead of through the pinned ring: the same doubles, through
    lh_score_newick and through PhyloHMM::ScoreSamples, and lh_timings_t.text_direct says which way was taken."""
    fam, tb, exp, meta = load_golden("synth_n40")
    labels = [str(s) for s in meta["labels"]]
    rep = 40
    big = {k: np.concatenate([v] * rep) for k, v in tb.items()}
    R = tb["rates"].shape[1]
    f = _family(fam, labels)
    ll_arr = f.score(big)
    for pinned in (False, True):
        text = host.NewickBuffer(big, labels, pinned=pinned)
        assert text.pinned == pinned
        ll = f.score_newick(text.ptrs, big["er"], big["pi"], R, rates=big["rates"], lengths=text.lengths)
        tm = f.timings()
        assert np.array_equal(ll, ll_arr, equal_nan=True)
        assert (tm["text_direct"] > 0) == pinned and tm["newick_host_trees"] == 0
        if pinned:
            f.set_option("text_direct", 0)
            ll = f.score_newick(text.ptrs, big["er"], big["pi"], R, rates=big["rates"], lengths=text.lengths)
            assert np.array_equal(ll, ll_arr, equal_nan=True) and f.timings()["text_direct"] == 0
            f.set_option("text_direct", 1)
            # without the lengths the strings are staged (the library would have to read them to find their ends)
            ll = f.score_newick(text.ptrs, big["er"], big["pi"], R, rates=big["rates"])
            assert np.array_equal(ll, ll_arr, equal_nan=True) and f.timings()["text_direct"] == 0
            # pinned, but not at a constant stride: staged
            order = np.arange(len(text))
            order[[3, 7]] = order[[7, 3]]
            ptrs = (type(text.ptrs))()
            import ctypes as C
            src, dst = C.cast(text.ptrs, C.POINTER(C.c_void_p)), C.cast(ptrs, C.POINTER(C.c_void_p))
            for i, j in enumerate(order):
                dst[i] = src[int(j)]
            shuffled = {k: v[order] for k, v in big.items()}
            ll = f.score_newick(ptrs, shuffled["er"], shuffled["pi"], R, rates=shuffled["rates"], lengths=text.lengths[order])
            assert np.array_equal(ll, ll_arr[order], equal_nan=True)
        del text
    f.close()
