"""GPU-box probe: parity spot-check + stage timings on the probe bundles."""
import os, sys, time, json
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from linearham_b200 import device
from oracle import pyoracle as po

def load(name):
    z = np.load(os.path.join(ROOT, "build", name + ".npz"))
    fam = {k[4:]: (int(z[k]) if z[k].ndim == 0 else z[k]) for k in z.files if k.startswith("fam_")}
    tb = {k[3:]: z[k] for k in z.files if k.startswith("tb_")}
    return fam, tb, z["ref_loglik"]

def bytes_tree(T, S, R=4):
    return (T - 2) * S * (2 * 32 * R + 8) + T * S + 2 * (2 * T - 3) * R * 128 + S * 8

def probe(name, opts_list, n_port=2, reps=3, rep=1):
    fam, tb, ref_ll = load(name)
    if rep > 1:
        tb = {k: np.concatenate([v] * rep) for k, v in tb.items()}
    B = tb["root_child"].shape[0]
    T, S = fam["n_tips"], fam["n_sites"]
    sub = {k: v[:n_port] for k, v in tb.items()}
    pll, _ = po.port_score_batch(fam, sub, want_outputs=False)
    for opts in opts_list:
        f = device.Family(fam)
        for k, v in opts.items():
            f.set_option(k, v)
        t0 = time.time(); ll = f.score(tb); t_e2e = time.time() - t0
        err_ref = np.max(np.abs(ll[:len(ref_ll)] - ref_ll) / np.abs(ref_ll))
        err_port = np.max(np.abs(ll[:n_port] - pll) / np.abs(pll))
        best = None
        for r in range(reps):
            t0 = time.time(); ll2 = f.score(tb); t = time.time() - t0
            tm = f.timings()
            if best is None or tm["ms_prune"] < best["ms_prune"]:
                best = dict(tm); best["wall_e2e_s"] = t
        assert np.array_equal(ll, ll2, equal_nan=True)
        tps = B / (best["ms_prune"] + best["ms_hmm"]) * 1e3
        print(json.dumps({"bundle": name, "opts": opts, "B": B, "T": T, "S": S, "err_vs_ref": float(err_ref),
                          "err_vs_port": float(err_port), "ms_prune_stage": best["ms_prune"], "ms_hmm": best["ms_hmm"],
                          "ms_h2d": best["ms_h2d"], "ms_d2h": best["ms_d2h"], "wall_e2e_s": best["wall_e2e_s"],
                          "trees_per_s_kernels": tps, "algo_GBps": tps * bytes_tree(T, S) / 1e9,
                          "slots": best["prune_slots"], "smem_slots": best["prune_smem_slots"], "tile": best["prune_tile_sites"], "ns": best["prune_sites_per_thread"], "occ": best["prune_ctas_per_sm"], "grid": best["prune_grid"],
                          "launches": best["kernel_launches"], "finite": bool(np.all(np.isfinite(ll)))}), flush=True)
        f.close()

if __name__ == "__main__":
    print(json.dumps({"fp64_peak_tflops": device.fp64_peak_tflops(0)}), flush=True)
    probe("probe_c3", [{}, {"sites_per_thread": 3}, {"sites_per_thread": 2}, {"smem_slots": 0}, {"smem_slots": 2}])
    probe("probe_c4", [{}, {"sites_per_thread": 3}, {"smem_slots": 0}, {"smem_slots": 2}, {"smem_slots": 4}], n_port=1, rep=8)
