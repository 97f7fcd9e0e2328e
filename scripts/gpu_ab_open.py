"""A/B of the opening chunk on C2, C3 and a C4 shard of 2048 trees (what a rank of 8 scores)."""
import os, sys, time, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_liao
from linearham_b200 import device, host, workload
cases = {}
fam, tb, exp, _ = load_liao(); cases["c2"] = (fam, tb)
for name, B in (("c3", 4096), ("c4", 2048), ("c4", 16384)):
    cfg = dict(workload.CONFIGS[name]); d = tempfile.mkdtemp(); ff = workload.write_family_files(cfg, d)
    ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0); fam3 = ph.family()
    tb3, _, _ = workload.tree_batch(cfg, ff, B, n_topo=64); cases[f"{name}x{B}"] = (fam3, tb3)
for name, (fam, tb) in cases.items():
    hs = {ou: device.Family(fam) for ou in (0, 1)}
    for ou, f in hs.items():
        f.set_option("chunk_growth", ou)
        if name == "c4x2048": f.set_option("host_threads", 4)
        for _ in range(3): f.score(tb)
    ts = {0: [], 1: []}
    for rep in range(12 if "16384" in name else 30):
        for ou, f in hs.items():
            t0 = time.perf_counter(); f.score(tb); ts[ou].append(time.perf_counter() - t0)
    for ou in (0, 1):
        tm = hs[ou].timings()
        print(name, "chunk_growth", ou, "median %.3f ms  p10 %.3f" % (1e3 * np.median(ts[ou]), 1e3 * np.percentile(ts[ou], 10)),
              {k: round(tm[k], 3) for k in ("ms_plan", "ms_h2d", "ms_kernels", "kernel_launches", "device_plan")})
    for f in hs.values(): f.close()
