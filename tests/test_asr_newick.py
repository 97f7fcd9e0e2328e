"""linearham_b200.asr.annotated_newick: the annotated tree of the R step (scripts/run_bootstrap_asr_ess.R:91-104).
The string must read back (host parser, which skips the comments) as the same operation arrays, and every node must
carry its sequence: the observed one on a tip, the sampled one on an inner node, in the order the arrays list them."""
import re

import numpy as np

from conftest import load_golden
from linearham_b200 import asr, host
from oracle import asr_oracle as ao


def test_annotated_tree_reads_back_and_carries_every_sequence():
    fam, tb, exp, meta = load_golden("synth_n12")
    labels = [str(s) for s in meta["labels"]]
    T, S = fam["n_tips"], 24
    codes = np.asarray(fam["tip_codes"], dtype=np.uint8).reshape(T, fam["n_sites"])[:, :S]
    naive = np.repeat(codes[0][None, :], 2, axis=0)
    sub = {k: v[:2] for k, v in tb.items()}
    out = ao.asr_sample(T, sub, codes, naive, naive_row=0, seed=3)     # (the oracle: this test needs no GPU)
    for b in range(2):
        s = asr.annotated_newick(T, sub["children"][b], sub["blen"][b], sub["root_child"][b], sub["root_blen"][b], labels,
                                 codes, out["node_states"][b])
        back = host.newick_to_ops([s], labels)
        for k in ("children", "root_child"):
            assert np.array_equal(back[k][0], sub[k][b]), k
        for k in ("blen", "root_blen"):
            assert np.array_equal(back[k][0], sub[k][b]), k               # shortest round-trip text: the same doubles
        seqs = re.findall(r'\[&ancestral="([ACGTN]+)"\]', s)
        assert len(seqs) == 2 * T - 2 and all(len(q) == S for q in seqs)
        # inner nodes close in operation order (post-order), the trifurcation last
        inner = re.findall(r'\)\[&ancestral="([ACGTN]+)"\]', s)
        assert inner == [asr._seq(out["node_states"][b][op]) for op in range(T - 2)]
        for t in range(T):
            assert f'{labels[t]}[&ancestral="{asr._seq(codes[t])}"]' in s
