"""bench.py --config c5: BASELINE config 5, the repertoire sweep (independent clonal families of 20-500
leaves, whole families sharded over the GPUs longest-first, no data-path collective; one all-gather of
the per-tree log-likelihoods per pass).

A step = one pass over the rank's families through lh_repertoire_score: HOST family descriptions and
tree batches in, host log-likelihoods out - bind (one blob copy per family), traversal planning, H2D,
kernels, D2H all inside the timed region, so `value` IS the end-to-end number (a repertoire has no
HBM-resident form: its inputs are new families). BASELINE names 10 000 families x 1000 samples; the
default run scores the first --families of them with --batch samples each and says so in `config`.
The headline bench line stays config c4 (bench.py)."""
import json
import os
import tempfile
import time

import numpy as np


def run(args, rank, world, local_rank):
    import torch
    import bench as B  # ClockSampler, METRIC, UNIT
    from linearham_b200 import device, sharding, workload
    cfg = dict(workload.CONFIGS["c5"])
    n_fam = args.families
    n_trees_f = args.batch or cfg["B"]
    if not torch.cuda.is_available() or device.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: linearham_b200 has no CPU fallback")
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = local_rank if world > 1 else 0
    cores = os.cpu_count() or 1

    workdir = tempfile.mkdtemp(prefix="lhbench_c5_")
    t0 = time.perf_counter()
    fams = workload.repertoire(cfg, workdir, n_fam, B=n_trees_f, n_topo=args.topologies_c5, device=dev)
    gen_s = time.perf_counter() - t0
    n_trees = [int(f["tb"]["root_child"].shape[0]) for f in fams]
    costs = [sharding.family_cost(f["fam"]["n_tips"], f["fam"]["n_sites"], n, cfg["R"]) for f, n in zip(fams, n_trees)]
    parts = sharding.assign_families(costs, world)
    mine = parts[rank]
    local_world = int(os.environ.get("LOCAL_WORLD_SIZE", world))
    slots = args.slots
    host_threads = max(1, cores // local_world // slots)
    rep = device.Repertoire(device=dev, n_slots=slots, host_threads=host_threads)
    for kv in args.opt:
        k, v = kv.split("=")
        rep.set_option(k, int(v))
    prep = rep.prepare([(fams[i]["fam"], fams[i]["tb"]) for i in mine])

    def sync_all():
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

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
    sync_all()
    sampler = B.ClockSampler(dev)
    sampler.start()
    acc = {"kernel_launches": 0, "h2d_bytes": 0, "d2h_bytes": 0, "ms_bind": 0.0, "ms_score": 0.0}
    t0 = time.perf_counter()
    for _ in range(args.steps):
        lls, st = rep.score_prepared(prep)
        gather(lls)
        for k in acc:
            acc[k] += st[k]
    sync_all()
    wall = time.perf_counter() - t0
    clocks = sampler.stop()
    tot = torch.tensor([wall], dtype=torch.float64)
    sums = torch.tensor([acc["kernel_launches"], acc["h2d_bytes"], acc["d2h_bytes"]], dtype=torch.float64)
    if dist is not None:
        tot, sums = tot.to(f"cuda:{dev}"), sums.to(f"cuda:{dev}")
        dist.all_reduce(tot, op=dist.ReduceOp.MAX)
        dist.all_reduce(sums, op=dist.ReduceOp.SUM)
    wall = float(tot.item())
    launches, h2d, d2h = (int(x) for x in sums.tolist())

    if rank == 0:
        # parity and the CPU baseline: the scalar port (one core) on the first trees of every family of a sample
        cpu_baseline, parity = None, None
        if not args.no_cpu_baseline:
            from oracle import pyoracle as po
            po.build_port()
            k_trees, worst, n_cmp, cpu_s, classes_equal = 4, 0.0, 0, 0.0, True
            full = {f: ll for f, ll in zip(mine, lls)}
            if gathered is not None:
                full = {f: g.cpu().numpy() for f, g in enumerate(gathered)}
            sample = list(range(n_fam))   # every family: a wrong family is a per-family event
            for f in sample:
                tb = {k: v[:k_trees] for k, v in fams[f]["tb"].items()}
                ta = time.perf_counter()
                port, _ = po.port_score_batch(fams[f]["fam"], tb, want_outputs=False)
                cpu_s += time.perf_counter() - ta
                ok = np.isfinite(port)
                classes_equal &= bool(np.array_equal(np.isfinite(full[f][:k_trees]), ok))
                if ok.any():
                    worst = max(worst, float(np.max(np.abs(full[f][:k_trees][ok] - port[ok]) / np.abs(port[ok]))))
                n_cmp += int(len(port))
            cpu_baseline = {"value": n_cmp / cpu_s, "unit": B.UNIT, "cores": 1, "kind": "port",
                            "sample": f"first {k_trees} tree samples of each of the {len(sample)} families, "
                                      f"scalar port on one core, {cpu_s:.2f} s"}
            parity = {"n": n_cmp, "max_rel_err_vs_cpu": worst, "tolerance": 1e-9, "non_finite_class_equal": classes_equal}
        total_trees = sum(n_trees)
        leaves = [f["n_leaves"] for f in fams]
        v = total_trees * args.steps / wall
        out = {
            "metric": B.METRIC, "value": v, "unit": B.UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": 1e3 * wall / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "impl": "b200",
            "config": {"workload": f"c5: {cfg['desc']}; the first {n_fam} of BASELINE's 10 000 families, "
                                   f"{n_trees_f} tree samples each (BASELINE: 1000)",
                       "families": n_fam, "trees_per_family": n_trees_f, "leaves_min_median_max":
                           [int(min(leaves)), int(np.median(leaves)), int(max(leaves))],
                       "gamma_categories": cfg["R"], "distinct_topologies_per_family": args.topologies_c5,
                       "slots_per_gpu": slots, "options": list(args.opt), "host_threads_per_slot": host_threads,
                       "parallelism": f"whole families assigned to {world} GPU(s) longest-first; no data-path "
                                      "collective; one all-gather of the per-tree log-likelihoods per pass",
                       "l2": "every family is new to the device each pass (bind + H2D inside the timed region)"},
            "families_per_s": n_fam * args.steps / wall,
            "value_is_end_to_end": True,
            "e2e": {"value": v, "unit": B.UNIT, "h2d_bytes_per_step": h2d // args.steps, "d2h_bytes_per_step": d2h // args.steps,
                    "api": "lh_repertoire_score (host family descriptions + tree batches in, host log-likelihoods out)",
                    "slot_ms_per_step_rank0": {"bind": acc["ms_bind"] / args.steps, "score": acc["ms_score"] / args.steps}},
            "roofline": None, "cpu_baseline": cpu_baseline, "parity": parity,
            "gpu_launches": launches, "clocks": clocks, "host_generation_s": gen_s,
        }
        print(json.dumps(out))
    rep.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0
