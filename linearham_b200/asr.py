"""The text artifact of the reference's ancestral-sequence step (scripts/run_bootstrap_asr_ess.R:91-104): every
kept tree in Newick with an `[&ancestral="..."]` comment on every node - the observed sequence on a tip, the sampled
one (lh_asr_sample, device.Family.asr_sample) on an inner node. Host-side string assembly only.

The R script re-roots the tree on the naive outgroup (ape::root(resolve.root = TRUE)) before it annotates, which adds a
root node next to the naive tip; here the tree is written as the batch arrays hold it - unrooted, the trifurcation
outermost, its third neighbour `root_child` first - and the trifurcation's comment is that node's sampled sequence
(the script's root when the naive tip is `root_child`)."""
import numpy as np

ALPHABET = "ACGTN"


def _seq(codes):
    return "".join(ALPHABET[int(c)] for c in codes)


def annotated_newick(n_tips, children, blen, root_child, root_blen, labels, tip_codes, node_states):
    """children [T-2][2], blen [T-2][2], root_child, root_blen: one tree as in lh_tree_batch_t; tip_codes [T][S] the
    alignment the states were sampled on (the sample's naive sequence in the naive row); node_states [T-2][S] from
    lh_asr_sample. -> the annotated Newick string (branch lengths written so that they read back exactly)."""
    T = int(n_tips)
    children, blen = np.asarray(children), np.asarray(blen)

    def fmt(x):
        return repr(float(x))    # the shortest text that reads back as the same double

    def node(i):
        """text of node i without its branch length (iterative: trees are a thousand levels deep at most, but
        Python's recursion limit is not something to depend on)"""
        out, stack = [], [(i, 0)]
        while stack:
            n, state = stack.pop()
            if n < T:
                out.append(f'{labels[n]}[&ancestral="{_seq(tip_codes[n])}"]')
                continue
            op = n - T
            if state == 0:
                out.append("(")
                stack.append((n, 1))
                stack.append((int(children[op, 0]), 0))
            elif state == 1:
                out.append(f":{fmt(blen[op, 0])},")
                stack.append((n, 2))
                stack.append((int(children[op, 1]), 0))
            else:
                out.append(f':{fmt(blen[op, 1])})[&ancestral="{_seq(node_states[op])}"]')
        return "".join(out)

    top = T - 3                      # the last operation: the trifurcation
    a, b = int(children[top, 0]), int(children[top, 1])
    return (f"({node(int(root_child))}:{fmt(root_blen)},{node(a)}:{fmt(blen[top, 0])},{node(b)}:{fmt(blen[top, 1])})"
            f'[&ancestral="{_seq(node_states[top])}"];')
