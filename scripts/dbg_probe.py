"""Debug: c4 sample of trees, GPU (site patterns on/off, pipelined/resident) vs the oracle port."""
import os, sys, tempfile, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from linearham_b200 import device, host, workload
from oracle import pyoracle as po
name = sys.argv[1] if len(sys.argv) > 1 else "c4"
n = int(sys.argv[2]) if len(sys.argv) > 2 else 256
cfg = workload.CONFIGS[name]
d = tempfile.mkdtemp()
ff = workload.write_family_files(cfg, d)
phmm = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
fam = phmm.family()
tb, _, _ = workload.tree_batch(cfg, ff, n, n_topo=min(n, 256))
pll, pout = po.port_score_batch(fam, tb)
res = {}
for sp in (1, 0):
    f = device.Family(fam)
    f.set_option("site_patterns", sp)
    ll, out = f.score(tb, outputs=True)
    f.upload(tb); f.run(); ll2 = f.fetch()
    res[sp] = (ll, out["xmsa_emission"], ll2)
    print("sp", sp, "nonfinite gpu", (~np.isfinite(ll)).sum(), "port", (~np.isfinite(pll)).sum(),
          "class equal", np.array_equal(np.isfinite(ll), np.isfinite(pll)), "pipelined==resident", np.array_equal(ll, ll2, equal_nan=True),
          f.timings()["pattern_trees"])
    bad = np.nonzero(np.isfinite(ll) != np.isfinite(pll))[0]
    for i in bad[:5]:
        xe, pxe = out["xmsa_emission"][i], pout["xmsa_emission"][i]
        dd = np.nonzero(~(np.abs(xe - pxe) <= 1e-9 * np.abs(pxe)))[0]
        print("  tree", i, "gpu", ll[i], "port", pll[i], "xe diff cols", dd[:10], xe[dd[:5]], pxe[dd[:5]])
    f.close()
print("patterns vs all identical:", np.array_equal(res[1][0], res[0][0], equal_nan=True), np.array_equal(res[1][1], res[0][1], equal_nan=True))
