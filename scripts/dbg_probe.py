import os, sys, numpy as np
sys.path.insert(0, "/root/repo"); sys.path.insert(0, "/root/repo/scripts")
from gpu_probe import load
from linearham_b200 import device
from oracle import pyoracle as po
fam, tb, ref_ll = load("probe_c3")
f = device.Family(fam)
ll = f.score(tb); ll2 = f.score(tb)
print("nan", np.isnan(ll).sum(), np.isnan(ll2).sum(), "inf", np.isinf(ll).sum(), "ndiff", (ll != ll2).sum())
d = np.nonzero(ll != ll2)[0]
print(d[:20], ll[d[:10]], ll2[d[:10]])
print("ref", ref_ll, ll[:4])
print(f.timings())
n=8
sub = {k: v[:n] for k, v in tb.items()}
pll, _ = po.port_score_batch(fam, sub, want_outputs=False)
print("port", pll, ll[:n])
if len(d):
    i = d[0]
    sub = {k: v[i:i+1] for k, v in tb.items()}
    pll, _ = po.port_score_batch(fam, sub, want_outputs=False)
    print("port at", i, pll, ll[i], ll2[i], f.score(sub))
