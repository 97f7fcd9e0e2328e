"""bench.py --config c5 (and the `c5` record of the default run): BASELINE config 5, the repertoire sweep -
independent clonal families of 20-500 leaves, whole families assigned to the GPUs longest-first
(sharding.assign_families), no data-path collective; one all-gather of the per-tree log-likelihoods per pass.
The reference runs one linearham process per cluster (SConstruct:344-352, 428-447; src/linearham.cpp:377-386).

A step = one pass over the rank's families through lh_repertoire_score: HOST family descriptions and tree
batches in, host log-likelihoods out - bind (one blob copy per family), traversal planning, H2D, kernels, D2H
all inside the timed region, so `value` IS the end-to-end number (a repertoire has no HBM-resident form: its
inputs are new families). BASELINE names 10 000 families x 1000 samples; a run scores the first --families of
them and says so in `config`. Every rank generates only its own families (process pool) and checks the first
trees of EVERY one of them against the scalar port; rank 0 also scores a 64-family sample with the
reference's own C++ (oracle/_ref)."""
import concurrent.futures as cf
import os
import tempfile
import time

import numpy as np


def _make_family(job):
    """Worker: one family of the repertoire (files + description + tree batch)."""
    cfg, outdir, i, n_leaves, B, n_topo, genes = job
    from linearham_b200 import host, synth, workload
    ff = synth.make_family(outdir, genes, int(n_leaves), seed=cfg["fam_seed"] + 1 + i, name=f"family_{i}")
    ff["hmm_dir"] = os.path.join(outdir, "hmm")
    ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    fam, labels = ph.family(), ph.strs("xmsa_labels")
    ph.close()
    if labels != ff["labels"]:
        raise RuntimeError("xMSA labels do not match the generated family")
    sub = dict(cfg, tree_seed=cfg["tree_seed"] + i)
    tb, newicks, alpha = workload.tree_batch(sub, ff, B, n_topo=n_topo, n_threads=1)
    files = {k: ff[k] for k in ("yaml_path", "hmm_dir")}
    return i, dict(fam=fam, tb=tb, newicks=newicks[:4], alpha=alpha[:4], files=files, n_leaves=int(n_leaves))


def generate(cfg, outdir, ids, sizes, B, n_topo, workers):
    from linearham_b200 import synth, workload
    genes = synth.make_germline_set(os.path.join(outdir, "hmm"), seed=workload.GERMLINE_SEED)
    jobs = [(cfg, outdir, int(i), int(sizes[i]), B, n_topo, genes) for i in ids]
    out = {}
    if workers <= 1 or len(jobs) < 4:
        for j in jobs:
            i, fam = _make_family(j)
            out[i] = fam
        return out
    import multiprocessing as mp
    # (spawn: the parent holds a CUDA context and NCCL threads, which a forked child must not inherit)
    with cf.ProcessPoolExecutor(max_workers=workers, mp_context=mp.get_context("spawn")) as ex:
        for i, fam in ex.map(_make_family, jobs, chunksize=max(1, len(jobs) // (workers * 8))):
            out[i] = fam
    return out


def run(env, args):
    """-> the record (rank 0) or None; bench.py prints it."""
    import bench as B
    from linearham_b200 import device, sharding, workload
    torch, dist, rank, world, dev = env.torch, env.dist, env.rank, env.world, env.dev
    cfg = dict(workload.CONFIGS["c5"])
    n_fam = args.families or 256
    n_trees_f = args.batch or cfg["B"]
    cores = env.cores
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world))

    # whole families to ranks, longest first, on a cost known before anything is generated (leaves x trees;
    # the xMSA width is ~600 for every family of this repertoire)
    sizes = workload.repertoire_sizes(cfg, n_fam)
    costs = [sharding.family_cost(int(n) + 1, 600, n_trees_f, cfg["R"]) for n in sizes]
    parts = sharding.assign_families(costs, world)
    mine = parts[rank]
    workdir = tempfile.mkdtemp(prefix=f"lhbench_c5_r{rank}_")
    t0 = time.perf_counter()
    fams = generate(cfg, workdir, mine, sizes, n_trees_f, args.topologies_c5, max(1, cores // local_world))
    gen_s = time.perf_counter() - t0
    n_trees = [n_trees_f] * n_fam
    slots = args.slots
    host_threads = max(1, cores // local_world // slots)
    rep = device.Repertoire(device=dev, n_slots=slots, host_threads=host_threads)
    for kv in args.opt:
        k, v = kv.split("=")
        rep.set_option(k, int(v))
    prep = rep.prepare([(fams[i]["fam"], fams[i]["tb"]) for i in mine])

    gathered = None

    def gather(lls):
        nonlocal gathered
        if dist is None:
            return
        local = {f: torch.from_numpy(ll).to(f"cuda:{dev}") for f, ll in zip(mine, lls)}
        gathered = sharding.gather_family_loglik(dist, local, parts, n_trees, device=f"cuda:{dev}")

    for _ in range(max(args.warmup, 3)):
        lls, st = rep.score_prepared(prep)
        gather(lls)
    env.sync_all()
    sampler = B.ClockSampler(dev)
    sampler.start()
    acc = {"kernel_launches": 0, "h2d_bytes": 0, "d2h_bytes": 0, "ms_bind": 0.0, "ms_score": 0.0}
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lls, st = rep.score_prepared(prep)
        gather(lls)
        for k in acc:
            acc[k] += st[k]
    env.sync_all()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    wall, = env.max_over_ranks([wall])
    launches, h2d, d2h = (int(x) for x in env.sum_over_ranks([acc["kernel_launches"], acc["h2d_bytes"], acc["d2h_bytes"]]))

    # ---- parity: every rank checks the first trees of EVERY family it scored against the scalar port --------
    parity, cpu_baseline = None, None
    if not args.no_cpu_baseline:
        from oracle import pyoracle as po
        po.build_port()
        k_trees, worst, n_cmp, cpu_s, classes_equal = 2, 0.0, 0, 0.0, 1.0
        mine_ll = {f: ll for f, ll in zip(mine, lls)}
        for f in mine:
            tb = {k: v[:k_trees] for k, v in fams[f]["tb"].items()}
            ta = time.perf_counter()
            port, _ = po.port_score_batch(fams[f]["fam"], tb, want_outputs=False)
            cpu_s += time.perf_counter() - ta
            got = mine_ll[f][:k_trees]
            if gathered is not None:   # what the job returns: the gathered copy
                got = gathered[f][:k_trees].cpu().numpy()
            e, cls, _ = B.rel_err_finite(got, port)
            worst = max(worst, e)
            classes_equal = min(classes_equal, 1.0 if cls else 0.0)
            n_cmp += len(port)
        worst, neg_cls = env.max_over_ranks([worst, -classes_equal])
        n_cmp_all, cpu_s_all = env.sum_over_ranks([n_cmp, cpu_s])
        parity = {"n": int(n_cmp_all), "families_checked": n_fam, "max_rel_err_vs_cpu": worst, "tolerance": 1e-9,
                  "non_finite_class_equal": bool(-neg_cls >= 1.0),
                  "checker": f"scalar port (oracle/port), first {k_trees} tree samples of every family, on the rank that scored it"}
        cpu_baseline = {"value": n_cmp_all / max(cpu_s_all, 1e-9), "unit": B.UNIT, "cores": 1, "kind": "port",
                        "sample": f"first {k_trees} tree samples of each of the {n_fam} families, scalar port, one core per rank, "
                                  f"{cpu_s_all:.2f} core-seconds in all"}
        if rank == 0 and po.have_ref():
            # a 64-family sample of rank 0's families through the reference's own C++ (Newick strings in)
            sample = mine[:: max(1, len(mine) // 64)][:64]
            worst_ref, cls_ref, n_ref = 0.0, True, 0
            for f in sample:
                ref = po.RefPhyloHMM(fams[f]["files"]["yaml_path"], 0, fams[f]["files"]["hmm_dir"], 0)
                nw = fams[f]["newicks"][:2]
                want = np.array([ref.score_newick(s, fams[f]["tb"]["er"][j], fams[f]["tb"]["pi"][j], fams[f]["alpha"][j], cfg["R"])
                                 for j, s in enumerate(nw)])
                ref.close()
                e, cls, _ = B.rel_err_finite(mine_ll[f][:len(nw)], want)
                worst_ref, cls_ref, n_ref = max(worst_ref, e), cls_ref and cls, n_ref + len(nw)
            parity["vs_reference"] = {"families": len(sample), "n": n_ref, "max_rel_err": worst_ref, "non_finite_class_equal": cls_ref,
                                      "checker": "oracle/_ref (the reference's own C++), first 2 tree samples of a 64-family "
                                                 "sample of rank 0's families"}
    out = None
    if rank == 0:
        total_trees = sum(n_trees)
        v = total_trees * args.steps / wall
        out = {
            "metric": B.METRIC, "value": v, "unit": B.UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "b200",
            "config": {"workload": f"c5: {cfg['desc']}; the first {n_fam} of BASELINE's 10 000 families, "
                                   f"{n_trees_f} tree samples each (BASELINE: 1000)",
                       "families": n_fam, "trees_per_family": n_trees_f, "leaves_min_median_max":
                           [int(sizes.min()), int(np.median(sizes)), int(sizes.max())],
                       "gamma_categories": cfg["R"], "distinct_topologies_per_family": args.topologies_c5,
                       "slots_per_gpu": slots, "options": list(args.opt), "host_threads_per_slot": host_threads,
                       "families_rank0": len(mine),
                       "parallelism": f"whole families assigned to {world} GPU(s) longest-first (sharding.assign_families); "
                                      "no data-path collective; one all-gather of the per-tree log-likelihoods per pass",
                       "l2": "every family is new to the device each pass (bind + H2D inside the timed region)"},
            "families_per_s": n_fam * args.steps / wall,
            "value_is_end_to_end": True,
            "e2e": {"value": v, "unit": B.UNIT, "h2d_bytes_per_step": h2d // args.steps, "d2h_bytes_per_step": d2h // args.steps,
                    "api": "lh_repertoire_score (host family descriptions + tree batches in, host log-likelihoods out)",
                    "slot_ms_per_step_rank0": {"bind": acc["ms_bind"] / args.steps, "score": acc["ms_score"] / args.steps}},
            "roofline": None, "cpu_baseline": cpu_baseline, "parity": parity,
            "gpu_launches": launches, "clocks": clocks, "host_generation_s_rank0": gen_s,
        }
    rep.close()
    return out
