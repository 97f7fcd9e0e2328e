import os, sys, numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import load_golden
from linearham_b200 import device
name = sys.argv[1] if len(sys.argv) > 1 else "synth_n12"
fam, tb, exp, _ = load_golden(name)
sub = {k: v[:1] for k, v in tb.items()}
f = device.Family(fam); f.set_option("site_patterns", 0)
ll0, o0 = f.score(sub, outputs=True); f.close()
for arena in (0, 256, 128, 96, 64):
    f = device.Family(fam); f.set_option("site_repeats", 1); f.set_option("sr_arena", arena)
    ll, o = f.score(sub, outputs=True); tm = f.timings(); f.close()
    d = np.nonzero(o["xmsa_emission"][0] != o0["xmsa_emission"][0])[0]
    print("arena", arena, tm["sr_arena"], tm["sr_ctas_per_sm"], "ll", ll[0], ll0[0], "ndiff cols", len(d), d[:12])
