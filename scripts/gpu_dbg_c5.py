import os, sys, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from linearham_b200 import device, workload
from oracle import pyoracle as po
cfg = dict(workload.CONFIGS["c5"])
d = tempfile.mkdtemp()
fams = workload.repertoire(cfg, d, 64, B=1000, n_topo=16)
ports = []
for f in fams:
    tb = {k: v[:4] for k, v in f["tb"].items()}
    ports.append(po.port_score_batch(f["fam"], tb, want_outputs=False)[0])
def check(tag, lls):
    bad = []
    for i, (ll, p) in enumerate(zip(lls, ports)):
        ok = np.isfinite(p)
        e = np.max(np.abs(ll[:4][ok] - p[ok]) / np.abs(p[ok])) if ok.any() else 0
        if e > 1e-9: bad.append((i, fams[i]["n_leaves"], float(e)))
    print(tag, "bad families:", bad[:10], len(bad), flush=True)
pairs = [(f["fam"], f["tb"]) for f in fams]
for slots in (1, 6):
    for opts in ([], [("open_unit", 0)], [("chunk_growth", 0)], [("device_plan", 1)], [("open_unit", 0), ("chunk_growth", 0)]):
        rep = device.Repertoire(n_slots=slots)
        for k, v in opts: rep.set_option(k, v)
        lls, st = rep.score(pairs)
        check(f"slots={slots} {opts}", lls)
        rep.close()
# single handles
bad = []
for i, (fam, tb) in enumerate(pairs[:64]):
    h = device.Family(fam); ll = h.score(tb); h.close()
    p = ports[i]; ok = np.isfinite(p)
    e = np.max(np.abs(ll[:4][ok] - p[ok]) / np.abs(p[ok])) if ok.any() else 0
    if e > 1e-9: bad.append((i, fams[i]["n_leaves"], float(e)))
print("fresh handles bad:", bad)
