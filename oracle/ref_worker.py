"""TEST INFRASTRUCTURE ONLY. One CPU worker of the reference arm / cpu_baseline leg of bench.py.

Scores tree samples with the reference's own C++ (oracle/_ref, kind "reference") or, where that
library is absent, with the scalar restatement (oracle/port, kind "port"). Protocol on
stdin/stdout (one line each way): after start-up the worker prints "ready"; for every line
"go <first> <count>" it scores samples [first, first+count) and prints a JSON line with its own
elapsed seconds and the log-likelihoods; "quit" ends it. The parent starts one worker per host core
and releases them together, so the reported time is concurrent work on all cores.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    kind, path = sys.argv[1], sys.argv[2]
    z = np.load(path, allow_pickle=False)
    from oracle import pyoracle as po
    if kind in ("reference", "reference_shipped"):  # the tuned build, or the -O0 build the reference's SConstruct makes
        ref = po.RefPhyloHMM(str(z["yaml_path"]), 0, str(z["hmm_dir"]), 0,
                             lib_path=po.REF_LIB_SHIPPED if kind == "reference_shipped" else None)
        newicks = [str(s) for s in z["newicks"]]

        def score(a, n):
            ll, st = ref.score_batch_timed(newicks[a:a + n], z["er"][a:a + n], z["pi"][a:a + n],
                                           z["alpha"][a:a + n], int(z["R"]))
            return ll, st
    else:
        fam = {k[4:]: (int(z[k]) if z[k].ndim == 0 else z[k]) for k in z.files if k.startswith("fam_")}
        tb = {k[3:]: z[k] for k in z.files if k.startswith("tb_")}
        po.port_lib()

        def score(a, n):
            ll, _ = po.port_score_batch(fam, {k: v[a:a + n] for k, v in tb.items()}, want_outputs=False)
            return ll, {}
    print("ready", flush=True)
    for line in sys.stdin:
        w = line.split()
        if not w or w[0] == "quit":
            break
        a, n = int(w[1]), int(w[2])
        t0 = time.perf_counter()
        ll, st = score(a, n) if n > 0 else (np.zeros(0), {})
        dt = time.perf_counter() - t0
        print(json.dumps({"elapsed_s": dt, "first": a, "loglik": [float(x) for x in ll], "stages": st}), flush=True)


if __name__ == "__main__":
    main()
