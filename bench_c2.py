"""bench.py --config c2: BASELINE config 2, the Liao clonal family (312 sequences x 363 nt; real partis
germline / HMM parameters; stand-in annotation and tree samples, see scripts/make_liao_bundle.py) scored
over a batch of 1024 tree samples on one B200. Inputs come from tests/golden/liao_b1024.npz, which also
holds the compiled reference's log-likelihood of every sample: `parity` compares all 1024.
The headline bench line stays config c4 (bench.py)."""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))


def run(env, args):
    """-> the record (rank 0) or None; bench.py prints it."""
    import bench as B
    from linearham_b200 import device, host, sharding, workload
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from conftest import load_liao
    torch, dist, rank, world, dev = env.torch, env.dist, env.rank, env.world, env.dev
    fam, tb, exp, z = load_liao()
    labels = [str(x) for x in z["labels"]]
    T, S, R = fam["n_tips"], fam["n_sites"], int(tb["rates"].shape[1])
    Bt = int(tb["root_child"].shape[0])
    shard, lo, hi = sharding.shard_batch(tb, rank, world)
    Bs = hi - lo
    f = device.Family(fam, device=dev)
    f.set_labels(labels)
    if world > 1:
        f.set_option("host_threads", env.host_threads)
    text = host.NewickBuffer(shard, labels)   # the shard's trees as text (the bundle stores arrays)

    def sync_all():
        f.sync()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    out = None
    f.upload(shard)
    for _ in range(max(3, args.warmup)):
        f.run()
    sync_all()
    sampler = B.ClockSampler(dev)
    sampler.start()
    acc = {"ms_prune": 0.0, "ms_pmatrix": 0.0, "ms_edge": 0.0, "ms_hmm": 0.0, "prune_launches": 0, "kernel_launches": 0}
    f.mark(0)
    for _ in range(args.steps):
        f.run()
        tm = f.timings()
        for k in acc:
            acc[k] += tm[k]
    f.mark(1)
    sync_all()
    ms_kernels = f.elapsed_ms(0, 1)
    ll = f.fetch()
    for _ in range(2):
        f.score(shard)
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ll_e2e = f.score(shard)
        t2 = f.timings()
    sync_all()
    e2e_s = time.perf_counter() - t0
    assert np.array_equal(ll, ll_e2e, equal_nan=True)
    # Newick text in (lh_score_newick: parsed on the device)
    for _ in range(2):
        f.score_newick(text.ptrs, shard["er"], shard["pi"], R, rates=shard["rates"])
    sync_all()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        ll_text = f.score_newick(text.ptrs, shard["er"], shard["pi"], R, rates=shard["rates"])
        t3 = f.timings()
    sync_all()
    text_s = time.perf_counter() - t0
    clocks = sampler.stop()
    assert np.array_equal(ll, ll_text, equal_nan=True)
    if dist is not None:
        ms_kernels, e2e_s, text_s = env.max_over_ranks([ms_kernels, e2e_s, text_s])
        full = sharding.gather_loglik(dist, torch.from_numpy(ll).to(f"cuda:{dev}"), Bt).cpu().numpy()
    else:
        full = ll
    if rank == 0:
        ref = exp["loglik"]
        parity = {"n": int(Bt), "max_rel_err_vs_reference": float(np.max(np.abs(full - ref) / np.abs(ref))),
                  "tolerance": 1e-9, "against": "log-likelihoods of the compiled reference stored in the bundle (all samples)"}
        cpu_baseline = None
        if not args.no_cpu_baseline:
            from oracle import pyoracle as po
            po.build_port()
            n = 16
            sub = {k: v[:n] for k, v in tb.items()}
            ta = time.perf_counter()
            port, _ = po.port_score_batch(fam, sub, want_outputs=False)
            dt = time.perf_counter() - ta
            cpu_baseline = {"value": n / dt, "unit": B.UNIT, "cores": 1, "kind": "port",
                            "sample": f"first {n} tree samples, scalar port on one core, {dt:.2f} s",
                            "max_rel_err_port_vs_reference": float(np.max(np.abs(port - ref[:n]) / np.abs(ref[:n])))}
        peak, peak_src = B.measured_peaks()
        bytes_tree = workload.bytes_per_tree(T, S, R)
        prune_s = acc["ms_prune"] * 1e-3
        achieved = bytes_tree * Bs * args.steps / prune_s / 1e9
        cols = int(tm["prune_columns"])
        fp64_peak = device.fp64_peak_tflops(dev)
        fl = workload.flops_per_tree(T, cols, R) * Bs * args.steps / prune_s / 1e12
        out = {
            "metric": B.METRIC, "value": Bt * args.steps / (ms_kernels * 1e-3), "unit": B.UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": max(3, args.warmup), "ms_per_step": ms_kernels / args.steps,
            "higher_is_better": True, "scaling": "strong", "vs_baseline": None, "dtype": "f64",
            "data": "Liao family sequences + partis HMM parameters (real); stand-in annotation and tree samples",
            "impl": "b200",
            "config": {"workload": "c2: Liao clonal family (312 sequences x 363 nt) scored over a batch of 1024 tree samples",
                       "n_tips": T, "xmsa_columns": S, "prune_columns": cols, "gamma_categories": R, "global_batch": Bt,
                       "hmm": {"vd": [fam["vd_sites"], fam["vd_states"]], "dj": [fam["dj_sites"], fam["dj_states"]],
                               "genes": [fam["n_v"], fam["n_d"], fam["n_j"]]},
                       "l2": "L2 flush not needed: a step reads what the step itself wrote (branch records, "
                             f"{f.timings()['prune_grid']} CTAs); inputs are {t2['h2d_bytes'] >> 20} MB"},
            "roofline": {"bound": "fp64", "achieved": fl, "peak": fp64_peak, "unit": "TFLOP/s", "frac": fl / fp64_peak,
                         "traffic": None, "kernel": "k_prune", "peak_source": "lh_measure_fp64_peak (this run)",
                         "avg_launch_ms": acc["ms_prune"] / max(1, acc["prune_launches"]),
                         "hbm_streaming_model": {"achieved": achieved, "peak": peak, "unit": "GB/s", "ratio": achieved / peak,
                                                 "peak_source": peak_src, "algorithmic_bytes_per_tree": bytes_tree}},
            "stage_ms_per_step": {k[3:]: acc[k] / args.steps for k in ("ms_pmatrix", "ms_prune", "ms_edge", "ms_hmm")},
            "e2e": {"value": Bt * args.steps / text_s, "unit": B.UNIT, "h2d_bytes_per_step": int(t3["h2d_bytes"]) * world,
                    "d2h_bytes_per_step": int(t3["d2h_bytes"]) * world, "ms_newick_rank0": t3["ms_newick"],
                    "api": "lh_score_newick (Newick strings in, host log-likelihoods out)"},
            "e2e_arrays": {"value": Bt * args.steps / e2e_s, "unit": B.UNIT, "h2d_bytes_per_step": int(t2["h2d_bytes"]) * world,
                           "d2h_bytes_per_step": int(t2["d2h_bytes"]) * world,
                           "api": "lh_score_batch (host arrays in, host log-likelihoods out)"},
            "cpu_baseline": cpu_baseline, "parity": parity, "gpu_launches": int(acc["kernel_launches"]),
            "clocks": clocks}
    f.close()
    return out if rank == 0 else None
