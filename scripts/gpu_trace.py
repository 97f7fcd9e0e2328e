import os, sys, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_liao
from linearham_b200 import device
fam, tb, exp, _ = load_liao()
for dp in (1, 0):
    f = device.Family(fam)
    f.set_option("device_plan", dp)
    for i in range(4):
        print("--- dp", dp, "call", i, file=sys.stderr, flush=True)
        t0 = time.perf_counter(); f.score(tb); print("wall %.3f" % (1e3 * (time.perf_counter() - t0)), file=sys.stderr, flush=True)
    f.close()
