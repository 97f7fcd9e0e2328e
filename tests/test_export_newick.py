"""ExportNewick's branch lengths (host/phylo_hmm.cpp: pll_utree_export_newick prints every length with "%f",
utree.c:148-270): the fast path rounds plain decimals on their digits, everything near a tie or spelled otherwise goes
through the parse-and-print path - both must give printf's digits, i.e. Python's '%f' % float(text)."""
import numpy as np

from linearham_b200 import host


def _export_lengths(texts):
    out = []
    for i in range(0, len(texts), 3):
        a, b, c = texts[i:i + 3]
        s = host.export_newick(f"(x:{a},y:{b},z:{c});")
        assert s.startswith("(x:") and s.endswith("):0.0;")
        body = s[1:-6].split(",")
        out += [t.split(":")[1] for t in body]
    return out


def _expect(t):
    v = float(t)
    return "%f" % (v if v != 0.0 else 1e-6)


def test_lengths_print_as_printf_does():
    rng = np.random.default_rng(3)
    texts = []
    for _ in range(20000):                       # what RevBayes writes: 0.xxxxx with 1..18 digits
        nd = int(rng.integers(1, 19))
        texts.append("0." + "".join(str(d) for d in rng.integers(0, 10, nd)))
    for _ in range(6000):                        # larger values, carries
        texts.append(str(int(rng.integers(0, 1000))) + "." + "".join(str(d) for d in rng.integers(0, 10, int(rng.integers(1, 12)))))
    for _ in range(6000):                        # at and around the ties of the sixth place
        head = "".join(str(d) for d in rng.integers(0, 10, 6))
        tail = rng.choice(["5", "50", "500000", "5000000000", "50000000001", "4999999999", "49999999999999", "499999",
                           "5000001", "4999998", "500000000000000000000001", "45", "55", "9999999", "5" + "0" * 25 + "1"])
        texts.append(f"{int(rng.integers(0, 3))}.{head}{tail}")
    texts += ["0.9999995", "0.9999994999", "9.9999995", "99.9999999", "999.9999996", "0.0", "0.000000", "0.0000004", "0.0000005",
              "0.00000051", "1e-3", "1.5E-7", "12345.678", "007.5", ".5", "5.", "5", "0.1", "1.0000000", "0.1234565", "0.1234575",
              "0.00000049999999999999999", "2.5e0", "0.3"]
    while len(texts) % 3:
        texts.append("0.25")
    got = _export_lengths(texts)
    want = [_expect(t) for t in texts]
    bad = [(t, g, w) for t, g, w in zip(texts, got, want) if g != w]
    assert not bad, bad[:10]
