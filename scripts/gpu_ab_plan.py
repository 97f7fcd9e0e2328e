"""A/B of the planners on small batches (C2 Liao, C3) in one process + pinned H2D bandwidth."""
import os, sys, time, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
from conftest import load_liao
from linearham_b200 import device, host, workload
for mb in (1, 8, 64, 512):
    a = torch.empty(mb << 20, dtype=torch.uint8).pin_memory(); b = torch.empty(mb << 20, dtype=torch.uint8, device="cuda")
    for _ in range(3): b.copy_(a, non_blocking=True)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(10): b.copy_(a, non_blocking=True)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 10
    print("H2D pinned %4d MB: %.3f ms  %.1f GB/s" % (mb, 1e3 * dt, (mb << 20) / dt / 1e9))
cases = {}
fam, tb, exp, _ = load_liao(); cases["c2"] = (fam, tb)
cfg = dict(workload.CONFIGS["c3"]); d = tempfile.mkdtemp(); ff = workload.write_family_files(cfg, d)
ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0); fam3 = ph.family()
tb3, _, _ = workload.tree_batch(cfg, ff, 4096, n_topo=256); cases["c3"] = (fam3, tb3)
for name, (fam, tb) in cases.items():
    hs = {dp: device.Family(fam) for dp in (0, 1)}
    for dp, f in hs.items():
        f.set_option("device_plan", dp)
        for _ in range(5): f.score(tb)
    ts = {0: [], 1: []}
    for rep in range(40):
        for dp, f in hs.items():
            t0 = time.perf_counter(); f.score(tb); ts[dp].append(time.perf_counter() - t0)
    for dp in (0, 1):
        tm = hs[dp].timings()
        print(name, "device_plan", dp, "median %.3f ms  p10 %.3f" % (1e3 * np.median(ts[dp]), 1e3 * np.percentile(ts[dp], 10)),
              {k: round(tm[k], 3) for k in ("ms_plan", "ms_h2d", "ms_kernels", "kernel_launches")}, tm["h2d_bytes"])
