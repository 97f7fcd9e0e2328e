"""One C5-sized family (150 leaves, 1000 trees) through lh_score_batch: the host timeline of a call (LH_TRACE=1)."""
import sys, tempfile, time
sys.path.insert(0, ".")
import numpy as np
from linearham_b200 import device, host, workload
cfg = dict(workload.CONFIGS["c3"]); cfg["n_leaves"] = int(sys.argv[1]) if len(sys.argv) > 1 else 150
ff = workload.write_family_files(cfg, tempfile.mkdtemp())
ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
fam = ph.family()
tb, _, _ = workload.tree_batch(cfg, ff, 1000, n_topo=16)
f = device.Family(fam)
if len(sys.argv) > 2: f.set_option("host_threads", int(sys.argv[2]))
for _ in range(3): f.score(tb)
print("=== traced call ===", file=sys.stderr, flush=True)
t0 = time.perf_counter(); f.score(tb); dt = time.perf_counter() - t0
tm = f.timings()
print("call ms", dt * 1e3, {k: round(tm[k], 3) for k in ("ms_call", "ms_setup", "ms_plan", "ms_h2d", "ms_kernels", "ms_d2h", "ms_plan_dev", "device_plan")})
