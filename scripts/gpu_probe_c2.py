"""Per-call breakdown of lh_score_batch on the Liao batch (small-batch latency)."""
import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_liao
from linearham_b200 import device
fam, tb, exp, _ = load_liao()
for opts in ([], [("device_plan", 1)], [("device_plan", 0)], [("chunk_trees", 1024)], [("chunk_trees", 512)], [("chunk_trees", 256)],
             [("host_threads", 4)], [("host_threads", 8)], [("timing", 0)]):
    f = device.Family(fam)
    for k, v in opts:
        f.set_option(k, v)
    for _ in range(5):
        f.score(tb)
    ts = []
    for _ in range(20):
        t0 = time.perf_counter(); f.score(tb); ts.append(time.perf_counter() - t0)
    tm = f.timings()
    print(opts, "median ms %.3f min %.3f" % (1e3 * np.median(ts), 1e3 * min(ts)),
          {k: round(tm[k], 3) for k in ("ms_call", "ms_setup", "ms_plan", "ms_h2d", "ms_kernels", "ms_d2h", "ms_plan_dev", "device_plan", "kernel_launches")})
    f.close()
