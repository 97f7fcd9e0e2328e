"""TEST INFRASTRUCTURE ONLY. Pure-Python Newick -> flat operation arrays, independent of the
product's C++ ingest (linearham_b200/csrc/host/newick.cpp), so the two can be checked against
each other. Follows the reference's conventions:

  * grammar + ring orientation: libpll parse_utree.y:179-246
  * virtual root = the top-level trifurcation; likelihood edge = (root, first listed subtree)
    (lib/libptpll/src/pll_util.cpp:368-371, pll_partition.cpp:309-315)
  * operations in libpll post-order: first subtree, then second, third, then the root itself
    (utree.c:379-442, 284-329)
  * missing or zero branch length -> EPS = 1e-6 (src/PhyloHMM.cpp:355,422; pll_util.cpp:46-62)
"""
import re

import numpy as np

EPS = 1e-6
_INDEX_RE = re.compile(r"\[\&index=[0-9]+\]")
_TOK = re.compile(r"\s*([(),:;]|[^\s(),:;\[\]]+)")


def parse_newick(s):
    """-> nested: leaf = (label, length); inner = ([child, child(, child)], label, length)."""
    s = _INDEX_RE.sub("", s)
    toks = _TOK.findall(s)
    pos = [0]

    def peek():
        return toks[pos[0]] if pos[0] < len(toks) else None

    def take():
        t = toks[pos[0]]
        pos[0] += 1
        return t

    def length():
        if peek() == ":":
            take()
            return float(take())
        return 0.0

    def subtree():
        if peek() == "(":
            take()
            kids = [subtree()]
            while peek() == ",":
                take()
                kids.append(subtree())
            assert take() == ")"
            label = None
            if peek() not in (":", ",", ")", ";", None):
                label = take()
            return (kids, label, length())
        label = take()
        return (label, length())

    t = subtree()
    assert peek() == ";", "newick must end with ';'"
    return t


def flatten(newick, label_to_row):
    """-> children[T-2,2] int32, blen[T-2,2] f64, root_child int, root_blen float."""
    top = parse_newick(newick)
    kids = top[0]
    assert isinstance(kids, list) and len(kids) == 3, "top level must be a trifurcation"
    T = len(label_to_row)
    children, blen = [], []

    def fix(x):
        return x if x else EPS

    def visit(node):
        """returns node id"""
        if not isinstance(node[0], list):
            return label_to_row[node[0]]
        a, b = node[0]
        assert len(node[0]) == 2, "inner nodes must be binary"
        ia = visit(a)
        ib = visit(b)
        children.append((ia, ib))
        blen.append((fix(a[-1]), fix(b[-1])))
        return T + len(children) - 1

    s0 = visit(kids[0])
    s1 = visit(kids[1])
    s2 = visit(kids[2])
    children.append((s1, s2))
    blen.append((fix(kids[1][-1]), fix(kids[2][-1])))
    assert len(children) == T - 2, (len(children), T)
    return (np.array(children, dtype=np.int32), np.array(blen, dtype=np.float64), int(s0), float(fix(kids[0][-1])))


def flatten_batch(newicks, labels, er, pi, rates):
    """labels: xMSA row labels (row 0 = 'naive'). -> tree-batch dict for linearham_b200._abi.tree_batch."""
    m = {lab: i for i, lab in enumerate(labels)}
    ch, bl, rc, rb = [], [], [], []
    for s in newicks:
        c, b, r, l = flatten(s, m)
        ch.append(c)
        bl.append(b)
        rc.append(r)
        rb.append(l)
    return {"children": np.stack(ch), "blen": np.stack(bl), "root_child": np.array(rc, dtype=np.int32),
            "root_blen": np.array(rb), "er": np.asarray(er, dtype=np.float64).reshape(len(newicks), 6),
            "pi": np.asarray(pi, dtype=np.float64).reshape(len(newicks), 4),
            "rates": np.asarray(rates, dtype=np.float64).reshape(len(newicks), -1)}
