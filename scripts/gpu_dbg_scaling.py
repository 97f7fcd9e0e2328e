import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from conftest import load_golden
from linearham_b200 import device
from oracle import pyoracle as po
name = sys.argv[1]
fam, tb, exp, _ = load_golden(name)
f = device.Family(fam)
ll, out = f.score(tb, outputs=True)
pll, pout = po.port_score_batch(fam, tb)
print("gpu ll", ll, "port ll", pll, "ref ll", exp["loglik"])
for k in out:
    a, b = np.asarray(out[k]), np.asarray(pout[k])
    same = np.array_equal(a, b, equal_nan=True)
    cls = np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(np.isinf(a), np.isinf(b))
    print(k, a.shape, "identical" if same else "DIFF", "class-equal" if cls else "CLASS-DIFF")
    if not cls:
        d = np.argwhere((np.isnan(a) != np.isnan(b)) | (np.isinf(a) != np.isinf(b)))
        print("   first class diffs", d[:5].tolist(), [a[tuple(i)] for i in d[:5]], [b[tuple(i)] for i in d[:5]])
for k in ("vd_scalers", "dj_scalers", "germ_scalers"):
    print(k, out[k][0][:40], pout[k][0][:40])
