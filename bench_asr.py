"""Throughput of lh_asr_sample (ancestral-state sampling) on the BASELINE families, through the C ABI with
host buffers (H2D of the trees and D2H of the sampled states inside the timed region). Not the north_star metric:
a side record for the R-step row of SURVEY 8(f).  python bench_asr.py [--config c3|c4] [--trees N] [--steps K]
One JSON line: site-draws (tree x alignment site) per second, the algorithmic scratch traffic of k_asr against the
measured HBM peak, and the oracle (numpy, one core) on a bounded sample as the CPU figure."""
import argparse, json, os, sys, tempfile, time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--config", default="c3")
    ap.add_argument("--trees", type=int, default=0)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=2)
    ap.add_argument("--no-cpu", action="store_true")
    args = ap.parse_args()
    from linearham_b200 import device, host, workload
    cfg = dict(workload.CONFIGS[args.config])
    B = args.trees or (256 if args.config == "c4" else 1024)
    ff = workload.write_family_files(cfg, tempfile.mkdtemp(prefix="lh_asr_"))
    ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    fam = ph.family()
    tb, _, _ = workload.tree_batch(cfg, ff, B, n_topo=min(B, 256))
    T, Sx = int(fam["n_tips"]), int(fam["n_sites"])
    codes = np.asarray(fam["tip_codes"], dtype=np.uint8).reshape(T, Sx)
    # the alignment of the R step: one column per site of the MSA - the xMSA's distinct (site, naive base) columns
    # collapse to the sites; the first column of every site stands in for it here
    rng = np.random.default_rng(11)
    msa = codes
    naive = np.repeat(codes[0][None, :], B, axis=0)
    flip = rng.random(naive.shape) < 0.1
    naive[flip] = rng.integers(0, 4, size=int(flip.sum()))
    S, R = msa.shape[1], int(tb["rates"].shape[1])
    f = device.Family(fam)
    for _ in range(args.warmup):
        out = f.asr_sample(tb, msa, naive, naive_row=0, seed=1)
    t0 = time.perf_counter()
    for s in range(args.steps):
        out = f.asr_sample(tb, msa, naive, naive_row=0, seed=2 + s)
    dt = (time.perf_counter() - t0) / args.steps
    f.close()
    draws = B * S
    # k_asr's scratch traffic per (tree, site): every upward pass writes 32 B per inner node and reads 32 B per inner
    # child (T-3 of them); R passes + (R-1)/R expected repeats; the downward pass reads 32 B per inner node
    passes = R + (R - 1) / R
    bytes_per_draw = passes * ((T - 2) * 32 + (T - 3) * 32) + (T - 2) * 32
    peak = None
    try:
        peak = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))).get("hbm_gbps_sustained")
    except Exception:
        pass
    line = {"metric": "asr site-draws/s (tree x alignment site: rate category + every inner node)", "value": draws / dt,
            "unit": "site-draws/s", "trees_per_s": B / dt, "ms_per_call": dt * 1e3, "n_gpus": 1, "steps": args.steps,
            "dtype": "f64", "data": "synthetic", "config": {"workload": args.config, "trees": B, "tips": T, "sites": S, "rates": R},
            "e2e": {"value": draws / dt, "unit": "site-draws/s",
                    "h2d_bytes_per_step": int(sum(np.asarray(v).nbytes for v in tb.values()) + msa.nbytes + naive.nbytes),
                    "d2h_bytes_per_step": int(out["node_states"].nbytes + out["site_rate"].nbytes)},
            "scratch_traffic": {"bytes_per_site_draw": bytes_per_draw, "achieved_GBps_whole_call": draws * bytes_per_draw / dt / 1e9,
                                "hbm_peak_GBps": peak}}
    if not args.no_cpu:
        from oracle import asr_oracle as ao
        nb = 2 if T > 500 else 8
        sub = {k: np.asarray(v)[:nb] for k, v in tb.items()}
        t0 = time.perf_counter()
        want = ao.asr_sample(T, sub, msa, naive[:nb], naive_row=0, seed=2 + args.steps - 1)
        cdt = time.perf_counter() - t0
        line["cpu_baseline"] = {"value": nb * S / cdt, "unit": "site-draws/s", "cores": 1, "kind": "port",
                                "sample": f"{nb} trees of the same batch (numpy restatement, scipy expm per branch)"}
        line["parity"] = {"states_equal": bool(np.array_equal(want["node_states"], out["node_states"][:nb])),
                          "site_rate_equal": bool(np.array_equal(want["site_rate"], out["site_rate"][:nb]))}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
