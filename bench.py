#!/usr/bin/env python
"""bench.py - PhyloHMM tree-likelihood evaluations per second (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config c4|c3|c2|c5]

A step is one pass of the hot path over one batch of synthetic tree samples of a fixed clonal
family: for every (tree, er, pi, alpha) sample, Felsenstein pruning over the xMSA + the scaled HMM
forward pass = one `InitializePhyloEmission() + LogLikelihood()` of the reference
(src/PhyloHMM.cpp:429-430). Default workload = BASELINE.json configs[3] ("c4": synthetic 1000-leaf
IGH family, 400 nt, Gamma-4, batch 16384 trees sharded over the N GPUs).

  value       trees/s, kernels only, batch already resident in HBM (lh_batch_upload once, then
              lh_batch_run per step); CUDA events on the library's stream, max over ranks
  e2e         trees/s through the class API a linearham user calls, PhyloHMM::ScoreSamples: NEWICK STRINGS
              + er/pi/alpha on the host in, log-likelihoods of the whole batch on rank 0's host out. The
              text lives in PINNED host memory (the contract's "inputs from pinned host memory"; --pageable-text
              for ordinary memory), goes to the GPU and is parsed there (k_newick); H2D, Gamma rates (host threads), parse,
              planning, kernels, D2H and - at N > 1 - the NCCL gather are all inside the timed region. This is
              what the reference arm times (src/PhyloHMM.cpp:419-430: parse + Gamma + emission + forward).
  e2e_arrays  the same through lh_score_batch with pre-flattened operation arrays and pre-computed rates
              (round 1's `e2e`)
  roofline    k_prune against the FP64 pipe (the binding roofline: inner CLVs never leave the chip), with the
              streaming-model HBM figure and the measured DRAM traffic beside it
  cpu_baseline  the reference's own C++ (oracle/_ref) on all host cores, bounded sample (N=1 only), plus a
              single-thread figure and the reference "as shipped" (-O0, SConstruct:266)
  e2e_pipeline  (c3) rows/s of the product's real entry point, PhyloHMM::RunPipeline: trace TSV in, linearham TSV
              out (scoring from the Newick text + the annotation's draws on the GPU), beside the reference's own
  c3, c5      the other single-box configurations, each with its own parity check (c5 = families sharded
              over the ranks through lh_repertoire_score)
  parity_vs_n1  (N > 1) rank 0 re-scores a slice of every rank's shard on its own GPU: bit-identical

--impl reference times the reference's CPU implementation alone (rank 0 only) and loads nothing of the product.
"""
import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "PhyloHMM tree-likelihood evals/sec"
UNIT = "trees/s"


def measured_peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        try:
            return float(json.load(open(p))["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.path = tempfile.mktemp(suffix=".csv")
        self.proc = None

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.device)],
                                         stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except OSError:
            self.proc = None

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        self.proc.wait()
        sm, mx, pw, reasons = [], [], [], set()
        for line in open(self.path):
            f = [x.strip() for x in line.split(",")]
            if len(f) < 8:
                continue
            try:
                sm.append(float(f[1])); mx.append(float(f[2])); pw.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[4:8]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        os.unlink(self.path)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)), "power_w_max": float(max(pw)),
                "samples": len(sm), "reasons": sorted(reasons)}


# ------------------------------------------------------------------ CPU arm (oracle) ---

class CpuPool:
    """One oracle/ref_worker.py process per host core, released together."""

    def __init__(self, kind, data_path, cores):
        self.cores = cores
        env = dict(os.environ, OMP_NUM_THREADS="1")
        self.procs = [subprocess.Popen([sys.executable, os.path.join(ROOT, "oracle", "ref_worker.py"), kind, data_path],
                                       stdin=subprocess.PIPE, stdout=subprocess.PIPE, text=True, env=env)
                      for _ in range(cores)]
        for p in self.procs:
            line = p.stdout.readline().strip()
            if line != "ready":
                raise RuntimeError("CPU worker failed to start: " + line)

    def score(self, first, count):
        """-> (wall seconds = slowest worker, loglik[count], summed stage seconds)"""
        per = (count + self.cores - 1) // self.cores
        jobs = []
        for i, p in enumerate(self.procs):
            a = first + i * per
            n = max(0, min(per, first + count - a))
            jobs.append((p, a, n))
        for p, a, n in jobs:
            p.stdin.write(f"go {a} {n}\n")
            p.stdin.flush()
        ll = np.zeros(count)
        wall, stages = 0.0, {}
        for p, a, n in jobs:
            r = json.loads(p.stdout.readline())
            wall = max(wall, r["elapsed_s"])
            ll[a - first:a - first + n] = r["loglik"]
            for k, v in r["stages"].items():
                stages[k] = stages.get(k, 0.0) + v
        return wall, ll, stages

    def close(self):
        for p in self.procs:
            try:
                p.stdin.write("quit\n")
                p.stdin.flush()
                p.wait(timeout=10)
            except Exception:
                p.kill()


def cpu_sample_file(ff, R, newicks, er, pi, alpha, workdir, name="cpu_sample.npz"):
    """The worker data file of the reference kind: the family's files + the samples as the RevBayes trace holds them."""
    path = os.path.join(workdir, name)
    np.savez(path, yaml_path=ff["yaml_path"], hmm_dir=ff["hmm_dir"], newicks=np.array(newicks), er=np.asarray(er),
             pi=np.asarray(pi), alpha=np.asarray(alpha), R=R)
    return path


def cpu_sample_size(cfg_name, cores):
    # c4: ~0.14 s/tree/core, c3: ~10 ms/tree/core -> about 5-10 s of wall time
    per_core = {"c4": 32, "c3": 256}.get(cfg_name, 64)
    return per_core * cores


def reference_samples(cfg, ff, n_topo):
    """The first n_topo samples of the configuration exactly as workload.tree_batch draws them (same generator,
    same seed) - Newick strings, er, pi, alpha - without touching the product library."""
    from linearham_b200 import synth  # pure Python
    rng = np.random.default_rng(cfg["tree_seed"])
    samples = synth.tree_samples(rng, ff["tree"], n_topo, outgroup=cfg.get("outgroup"))
    return ([s["newick"] for s in samples], np.array([s["er"] for s in samples]), np.array([s["pi"] for s in samples]),
            np.array([s["alpha"] for s in samples]))


def rel_err_finite(a, b):
    ok = np.isfinite(b)
    worst = float(np.max(np.abs(a[ok] - b[ok]) / np.abs(b[ok]))) if ok.any() else 0.0
    return worst, bool(np.array_equal(np.isfinite(a), ok)), int(len(b) - ok.sum())


# ---------------------------------------------------------------------- reference arm ---

def reference_arm(args):
    from linearham_b200 import workload  # configuration table + synthetic family files (Python only)
    from oracle import pyoracle as po
    cfg = dict(workload.CONFIGS[args.config])
    cores = os.cpu_count() or 1
    workdir = tempfile.mkdtemp(prefix="lhbench_ref_")
    ff = workload.write_family_files(cfg, workdir)
    B_total = args.batch or cfg["B"]
    n_topo = min(args.topologies, B_total)
    newicks, er, pi, alpha = reference_samples(cfg, ff, n_topo)
    if not po.have_ref():
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libref_linearham.so is missing (built by "
                          "__graft_entry__.build() where /root/reference exists)"}))
        return 0
    ref = po.RefPhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    fam = ref.family()
    ref.close()
    config = workload_config(args, cfg, fam, B_total)
    n_step = max(cores, min(len(newicks), {"c4": 4, "c3": 64}.get(args.config, 16) * cores))
    path = cpu_sample_file(ff, cfg["R"], newicks[:n_step], er[:n_step], pi[:n_step], alpha[:n_step], workdir)
    pool = CpuPool("reference", path, cores)
    for _ in range(args.warmup):
        pool.score(0, n_step)
    total = 0.0
    for _ in range(args.steps):
        w, _, stages = pool.score(0, n_step)
        total += w
    pool.close()
    v = n_step * args.steps / total
    sample = (f"{n_step} tree samples per step (the first {n_step} of the batch), {cores} worker processes, each the "
              "reference's loop body src/PhyloHMM.cpp:419-430 (Newick parse + Gamma rates + emission + forward); "
              "the std::regex of :420 is built once per worker, not once per row as the reference does")
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
        "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "reference", "sample": sample},
        "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0}))
    return 0


def workload_config(args, cfg, fam, B_total):
    return {"workload": f"{args.config}: {cfg['desc']}", "n_tips": int(fam["n_tips"]), "xmsa_columns": int(fam["n_sites"]),
            "gamma_categories": cfg["R"], "global_batch": B_total, "distinct_topologies": min(args.topologies, B_total),
            "hmm": {"vd": [int(fam["vd_sites"]), int(fam["vd_states"])], "dj": [int(fam["dj_sites"]), int(fam["dj_states"])],
                    "genes": [int(fam["n_v"]), int(fam["n_d"]), int(fam["n_j"])]},
            "parallelism": f"trees sharded over {args.gpus} GPU(s), no data-path collective; "
                           "NCCL all-gather of the per-tree log-likelihoods",
            "l2": "inputs larger than L2"}


# -------------------------------------------------------------------------- B200 arm ---

class Env:
    """Process-wide state of the B200 arm: ranks, device, the torch stream everything is ordered on."""

    def __init__(self, args):
        import torch
        from linearham_b200 import device
        self.torch = torch
        self.rank = int(os.environ.get("RANK", "0"))
        self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local_rank = int(os.environ.get("LOCAL_RANK", "0"))
        if not torch.cuda.is_available() or device.device_count() < 1:
            raise RuntimeError("bench.py needs a CUDA device: linearham_b200 has no CPU fallback")
        self.dist = None
        self.dev = self.local_rank if self.world > 1 else 0
        torch.cuda.set_device(self.dev)
        if self.world > 1:
            import torch.distributed as dist
            dist.init_process_group("nccl", device_id=torch.device("cuda", self.dev))
            self.dist = dist
        self.cores = os.cpu_count() or 1
        local_world = int(os.environ.get("LOCAL_WORLD_SIZE", self.world))
        # ranks of one node share its host cores: each takes its share for staging and planning
        self.host_threads = max(2, self.cores // local_world) if self.world > 1 else self.cores
        # a non-default torch stream: the library runs its kernels on it (option "stream") and the NCCL gather
        # is queued behind them on the same stream
        self.stream = torch.cuda.Stream(device=self.dev)

    def sync_all(self):
        self.torch.cuda.synchronize()
        if self.dist is not None:
            self.dist.barrier()
            self.torch.cuda.synchronize()

    def max_over_ranks(self, values):
        if self.dist is None:
            return list(values)
        t = self.torch.tensor(list(values), dtype=self.torch.float64, device=f"cuda:{self.dev}")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return t.tolist()

    def sum_over_ranks(self, values):
        if self.dist is None:
            return list(values)
        t = self.torch.tensor(list(values), dtype=self.torch.float64, device=f"cuda:{self.dev}")
        self.dist.all_reduce(t, op=self.dist.ReduceOp.SUM)
        return t.tolist()


class _DevArr:
    def __init__(self, ptr, n):
        self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}


def bench_family(env, args, cfg_name, steps, warmup, full_record):
    """One tree-batch configuration (c3 / c4) on the env's ranks -> the record of rank 0 (None elsewhere)."""
    torch, dist = env.torch, env.dist
    from linearham_b200 import device, host, sharding, workload
    rank, world, dev = env.rank, env.world, env.dev
    cfg = dict(workload.CONFIGS[cfg_name])
    B_total = (args.batch if cfg_name == args.config else 0) or cfg["B"]
    workdir = tempfile.mkdtemp(prefix=f"lhbench_{cfg_name}_")
    ff = workload.write_family_files(cfg, workdir)
    # per-family host work of the product: partis YAML + HMM parameter dir -> state space, transition
    # matrices, xMSA (C++ linearham::PhyloHMM of this repo; bit-exact vs the reference, tests/test_host_model.py)
    phmm = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0, device=dev)
    fam, labels = phmm.family(), phmm.strs("xmsa_labels")
    if labels != ff["labels"]:
        raise RuntimeError("xMSA labels do not match the generated family")
    T, S, R = fam["n_tips"], fam["n_sites"], cfg["R"]
    tb, newicks, alpha = workload.tree_batch(cfg, ff, B_total, n_topo=args.topologies)
    cores = env.cores
    sub = argparse.Namespace(**vars(args))
    sub.config = cfg_name
    config = workload_config(sub, cfg, fam, B_total)
    config["tree_samples"] = ("every tree keeps the naive tip under the virtual root (RevBayes' outgroup=naive template), so "
                              "the whole batch is pruned on the distinct site patterns of xMSA rows 1.. (prune_columns)")

    # CPU baseline first (before this process touches the GPU heavily); rank 0, N=1 only
    cpu_baseline, parity, cpu_ll = None, None, None
    want_cpu = rank == 0 and not args.no_cpu_baseline
    if want_cpu:
        from oracle import pyoracle as po
        if po.have_ref():
            n_cpu = min(len(newicks), cpu_sample_size(cfg_name, cores) if (world == 1 and full_record) else 256)
            path = cpu_sample_file(ff, R, newicks[:n_cpu], tb["er"][:n_cpu], tb["pi"][:n_cpu], alpha[:n_cpu], workdir)
            pool = CpuPool("reference", path, cores)
            pool.score(0, cores)  # warm the workers (page in the library, first allocation)
            w, cpu_ll, stages = pool.score(0, n_cpu)
            pool.close()
            cpu_baseline = {"value": n_cpu / w, "unit": UNIT, "cores": cores, "kind": "reference",
                            "sample": f"first {n_cpu} tree samples of the batch, one worker process per core, {w:.2f} s wall",
                            "stage_core_seconds": stages}
            if world == 1 and full_record:
                # one worker alone, and the reference as its own build makes it (-O0 linearham + libptpll,
                # SConstruct:266, lib/libptpll/CMakeLists.txt:14) on all cores
                n1 = {"c4": 6, "c3": 64}.get(cfg_name, 8)
                p1 = CpuPool("reference", path, 1)
                p1.score(0, 1)
                w1, _, _ = p1.score(0, n1)
                p1.close()
                cpu_baseline["single_thread"] = {"value": n1 / w1, "unit": UNIT, "cores": 1, "sample": f"first {n1} tree samples, {w1:.2f} s"}
                if os.path.exists(po.REF_LIB_SHIPPED):
                    ns = {"c4": 1, "c3": 16}.get(cfg_name, 2) * cores
                    ps = CpuPool("reference_shipped", path, cores)
                    ws, _, _ = ps.score(0, min(ns, n_cpu))
                    ps.close()
                    cpu_baseline["as_shipped"] = {"value": min(ns, n_cpu) / ws, "unit": UNIT, "cores": cores,
                                                  "build": "make -C oracle/refbuild FLAVOR=shipped (-O0, as SConstruct:266 builds it)",
                                                  "sample": f"first {min(ns, n_cpu)} tree samples, one worker per core, {ws:.2f} s wall"}
                else:
                    cpu_baseline["as_shipped"] = None

    # shard the batch: contiguous, equal chunks (trees of one family cost the same)
    per = sharding.per_rank(B_total, world)
    shard, lo, hi = sharding.shard_batch(tb, rank, world)
    Bs = hi - lo
    alpha_s = alpha[lo:hi]
    # the shard's trees as the RevBayes trace would hold them (text), written once outside the timed regions
    # (in pinned host memory - the contract's "inputs from pinned host memory": the library then copies the strings to
    # the device where they lie; --pageable-text puts them in ordinary memory, staged through the library's pinned ring)
    try:
        text = host.NewickBuffer(shard, labels, pinned=not args.pageable_text)
    except host.HostError as e:   # (a box that refuses to page-lock that much: the text is staged instead)
        print(f"[bench] pinning the text failed ({e}); using pageable memory", file=sys.stderr)
        text = host.NewickBuffer(shard, labels, pinned=False)

    f = device.Family(fam, device=dev)
    f.set_labels(labels)
    for kv in args.opt:
        k, v = kv.split("=")
        f.set_option(k, int(v))
        phmm.set_option(k, int(v))
    if args.opt:
        config["options"] = list(args.opt)
    f.set_option("stream", env.stream.cuda_stream)
    if world > 1:
        f.set_option("host_threads", env.host_threads)
        phmm.set_host_threads(env.host_threads)
    gathered = None

    side = torch.cuda.Stream(device=dev) if dist is not None else None
    stage_v = torch.zeros(per, dtype=torch.float64, device=f"cuda:{dev}") if dist is not None else None
    ev_ready, ev_done = (torch.cuda.Event(), torch.cuda.Event()) if dist is not None else (None, None)
    gathers = 0

    def gather_dev(ptr, n):
        """NCCL all-gather of this rank's device-resident log-likelihoods. The rank's vector is copied (16 KB, on the
        kernels' stream, behind k_hmm) into a staging tensor and gathered from there on a side stream, so that the
        next step's kernels do not wait for the slowest rank; the staging tensor is not rewritten before the gather
        that reads it is done. Everything is inside the timed region and complete at its closing synchronize."""
        nonlocal gathered, gathers
        if dist is None:
            return None
        local = torch.as_tensor(_DevArr(ptr, n), device=f"cuda:{dev}")
        if gathered is None:
            gathered = torch.empty(world * per, dtype=torch.float64, device=f"cuda:{dev}")
        if gathers:
            env.stream.wait_event(ev_done)
        stage_v[:n].copy_(local)
        ev_ready.record(env.stream)
        with torch.cuda.stream(side):
            side.wait_event(ev_ready)
            sharding.gather_loglik(dist, stage_v, B_total, out=gathered)
            ev_done.record(side)
        gathers += 1
        return gathered[:B_total]

    sampler = ClockSampler(dev)
    rec = {}
    with torch.cuda.stream(env.stream):
        # -------- value: batch resident in HBM -------------------------------------
        f.upload(shard)
        for _ in range(warmup):
            f.run()
            gather_dev(*f.loglik_dev())
        env.sync_all()
        if full_record:
            sampler.start()
        f.mark(0)
        acc = {"ms_prune": 0.0, "ms_pmatrix": 0.0, "ms_edge": 0.0, "ms_hmm": 0.0, "prune_launches": 0, "kernel_launches": 0}
        for _ in range(steps):
            f.run()
            gather_dev(*f.loglik_dev())   # the collective is part of the step
            tm = f.timings()
            for k in acc:
                acc[k] += tm[k]
        f.mark(1)
        env.sync_all()
        ms_kernels = f.elapsed_ms(0, 1)
        ll_dev = f.fetch()
        full_value = gathered[:B_total].cpu().numpy() if gathered is not None else ll_dev

        # -------- e2e_arrays: flattened host arrays in, this rank's host log-likelihoods out ----------
        arr_s, arr_detail, h2d_a, d2h_a = None, None, 0, 0
        if not args.no_e2e:
            for _ in range(2):
                f.score(shard)
            env.sync_all()
            t0 = time.perf_counter()
            for _ in range(steps):
                ll_arr = f.score(shard)
                t2 = f.timings()
            env.sync_all()
            arr_s = time.perf_counter() - t0
            h2d_a, d2h_a = t2["h2d_bytes"], t2["d2h_bytes"]
            arr_detail = {k: t2[k] for k in ("ms_call", "ms_setup", "ms_plan", "ms_h2d", "ms_kernels", "ms_d2h", "ms_plan_dev", "device_plan")}
            assert np.array_equal(ll_arr, ll_dev, equal_nan=True)

        # -------- e2e: Newick text in through the class API, the whole batch's log-likelihoods on rank 0's host out
        e2e_s, e2e_detail, h2d_e, d2h_e, e2e_launches, ll_host = None, None, 0, 0, 0, None
        if not args.no_e2e:
            stage = torch.empty(per, dtype=torch.float64, device=f"cuda:{dev}") if dist is not None else None

            def e2e_step():
                ll = phmm.ScoreSamples(text, shard["er"], shard["pi"], alpha_s, R)
                if dist is None:
                    return ll
                stage[:Bs].copy_(torch.from_numpy(ll), non_blocking=True)
                allv = sharding.gather_loglik(dist, stage, B_total, out=gathered)
                out = allv.cpu().numpy() if rank == 0 else None   # the job's result lands on rank 0's host
                return out

            for _ in range(2):
                e2e_step()
            env.sync_all()
            t0 = time.perf_counter()
            for _ in range(steps):
                ll_host = e2e_step()
                t3 = phmm.timings()
                e2e_launches += t3["kernel_launches"]
            env.sync_all()
            e2e_s = time.perf_counter() - t0
            h2d_e, d2h_e = t3["h2d_bytes"], t3["d2h_bytes"]
            e2e_detail = {k: t3[k] for k in ("ms_call", "ms_setup", "ms_plan", "ms_h2d", "ms_kernels", "ms_d2h", "ms_plan_dev",
                                             "ms_newick", "newick_host_trees", "pattern_trees", "text_direct")}
            if rank == 0:
                assert np.array_equal(ll_host, full_value, equal_nan=True), "text in and arrays in must give the same doubles"
        # -------- e2e_pipeline: the product's real entry point, RunPipeline (src/PhyloHMM.cpp:393-446): RevBayes trace TSV in,
        # linearham TSV out (scoring + one SampleNaiveSequence per row), on c3; beside it the reference's own RunPipeline
        pipe = None
        if cfg_name == "c3" and world == 1 and not args.no_e2e:
            from linearham_b200 import synth
            rows = [{"newick": text[i], "er": shard["er"][i], "pi": shard["pi"][i], "alpha": float(alpha_s[i])} for i in range(Bs)]
            tsv_all, tsv_small = os.path.join(workdir, "trace.tsv"), os.path.join(workdir, "trace64.tsv")
            synth.write_tsv(tsv_all, rows)
            synth.write_tsv(tsv_small, rows[:64])
            out_tsv = os.path.join(workdir, "lh_out.tsv")
            ours = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0, device=dev)
            PIPE_BATCH = 256   # RunPipeline's default: rows per device call
            ours.RunPipeline(tsv_small, out_tsv, R, PIPE_BATCH)      # fresh object, seed 0: comparable with the reference's run
            ours_small = open(out_tsv).read()
            ours.RunPipeline(tsv_all, out_tsv, R, PIPE_BATCH)         # warm-up
            t0 = time.perf_counter()
            for _ in range(steps):
                ours.RunPipeline(tsv_all, out_tsv, R, PIPE_BATCH)
            pipe_s = time.perf_counter() - t0
            ours.close()
            pipe = {"value": Bs * steps / pipe_s, "unit": "rows/s", "rows_per_run": Bs, "batch_size": PIPE_BATCH,
                    "tsv_bytes": os.path.getsize(tsv_all),
                    "api": "PhyloHMM::RunPipeline (TSV in -> TSV out; reader | device scoring from Newick text + the "
                           "annotation's draws on the device (k_sample) | output lines in file order; the stages overlap)"}
            if want_cpu and po.have_ref():
                ref = po.RefPhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
                ref_out = os.path.join(workdir, "ref_out.tsv")
                t0 = time.perf_counter()
                ref.run_pipeline(tsv_small, ref_out, R)
                ref_s = time.perf_counter() - t0
                ref.close()
                pipe["reference"] = {"value": 64 / ref_s, "unit": "rows/s", "cores": 1, "sample": f"first 64 rows, {ref_s:.2f} s, "
                                     "the reference's own RunPipeline (oracle/_ref), one thread as the reference runs it"}
                pipe["output_identical_to_reference_on_sample"] = open(ref_out).read() == ours_small
        clocks = sampler.stop() if full_record else None

        # -------- N > 1: what the ranks computed equals what one GPU computes (bit for bit) ----------
        parity_vs_n1 = None
        if dist is not None:
            k = max(1, 512 // world)
            idx = np.concatenate([np.arange(r * per, min(B_total, r * per + k)) for r in range(world)])
            if rank == 0:
                sub_tb = {key: np.ascontiguousarray(v[idx]) for key, v in tb.items()}
                again = f.score(sub_tb)
                parity_vs_n1 = {"equal": bool(np.array_equal(again, full_value[idx], equal_nan=True)), "n": int(len(idx)),
                                "how": f"rank 0 re-scored the first {k} trees of every rank's shard on its own GPU; "
                                       "np.array_equal with the NCCL-gathered vector"}
            env.sync_all()

    # -------- max over ranks ------------------------------------------------------
    ms_kernels, arr_max, e2e_max = env.max_over_ranks([ms_kernels, arr_s or 0.0, e2e_s or 0.0])
    launches_total, h2d_a, d2h_a, h2d_e, d2h_e, e2e_launches = (int(x) for x in env.sum_over_ranks(
        [acc["kernel_launches"], h2d_a, d2h_a, h2d_e, d2h_e, e2e_launches]))
    out = None
    if rank == 0:
        value = B_total * steps / (ms_kernels * 1e-3)
        peak, peak_src = measured_peaks()
        bytes_tree = workload.bytes_per_tree(T, S, R)
        # the kernel walks the distinct site patterns when it can; utilisation counts executed flop
        cols = int(tm["prune_columns"])
        flops_tree = workload.flops_per_tree(T, cols, R)
        config["prune_columns"] = cols
        config["pattern_trees_rank0"] = int(tm["pattern_trees"])
        prune_s = acc["ms_prune"] * 1e-3
        launches_per_step = max(1, acc["prune_launches"] // steps)
        trees_per_launch = Bs / launches_per_step
        avg_launch_ms = acc["ms_prune"] / max(1, acc["prune_launches"])
        streaming = bytes_tree * Bs * steps / prune_s / 1e9
        fp64_peak = device.fp64_peak_tflops(dev)
        fp64_ach = flops_tree * Bs * steps / prune_s / 1e12
        traffic, traffic_gbs, traffic_note = None, None, None
        tp = os.path.join(ROOT, "profiles", f"k_prune_traffic_{cfg_name}.json")
        if os.path.exists(tp):
            try:
                tj = json.load(open(tp))
                traffic = tj["dram_bytes_per_launch"] / tj["trees_in_launch"] * trees_per_launch
                traffic_gbs = traffic / (avg_launch_ms * 1e-3) / 1e9
                traffic_note = (f"ncu --set full, dram__bytes_read.sum + dram__bytes_write.sum of one k_prune launch of "
                                f"{tj['trees_in_launch']} trees (profiles/k_prune_traffic_{cfg_name}.json, tag {tj.get('tag')}), "
                                "per tree x trees_per_launch")
            except Exception:
                traffic = None
        if cpu_ll is not None:
            n = len(cpu_ll)
            worst, cls, nonfin = rel_err_finite(full_value[:n], cpu_ll)
            parity = {"n": int(n), "max_rel_err_vs_cpu": worst, "tolerance": 1e-9, "non_finite_class_equal": cls,
                      "non_finite": nonfin, "checker": "oracle/_ref (the reference's own C++)"}
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": steps, "warmup": warmup,
            "ms_per_step": ms_kernels / steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config, "impl": "b200",
            "roofline": {"bound": "fp64", "achieved": fp64_ach, "peak": fp64_peak, "unit": "TFLOP/s", "frac": fp64_ach / fp64_peak,
                         "kernel": "k_prune", "peak_source": "lh_measure_fp64_peak (register-resident DFMA loop, this run)",
                         "flop_model": "60 flop per inner node, pruning column and rate category (SURVEY 8(d): two 4x4 "
                                       "mat-vecs + the product), on the prune_columns the kernel walks",
                         "traffic": traffic, "traffic_source": traffic_note, "dram_gbs": traffic_gbs,
                         "dram_frac_of_hbm_peak": None if traffic_gbs is None else traffic_gbs / peak,
                         "trees_per_launch": trees_per_launch, "avg_launch_ms": avg_launch_ms,
                         "hbm_streaming_model": {"achieved": streaming, "peak": peak, "unit": "GB/s", "ratio": streaming / peak,
                                                 "peak_source": peak_src, "algorithmic_bytes_per_tree": bytes_tree,
                                                 "note": "SURVEY 8(d)'s streaming-model bytes (every inner CLV written and read once) "
                                                         "over the launch time: not a bandwidth fraction - inner CLVs live in "
                                                         "registers / shared memory, so the ratio exceeds 1 and HBM does not bind"}},
            "stage_ms_per_step": {"k_pmatrix": acc["ms_pmatrix"] / steps, "k_prune": acc["ms_prune"] / steps,
                                  "k_edge": acc["ms_edge"] / steps, "k_hmm": acc["ms_hmm"] / steps, "rank": 0},
            "cpu_baseline": cpu_baseline,
            "e2e": None if e2e_s is None else {
                "value": B_total * steps / e2e_max, "unit": UNIT, "h2d_bytes_per_step": h2d_e, "d2h_bytes_per_step": d2h_e,
                "newick_bytes_per_step_rank0": text.total_bytes(), "host_threads_per_rank": env.host_threads,
                "text_memory": "pinned (copied in place, 2-D copy per chunk)" if text.pinned else "pageable (staged through the pinned ring)",
                "last_call_ms_rank0": e2e_detail, "gather_timed": world > 1,
                "api": "PhyloHMM::ScoreSamples (Newick strings + er/pi/alpha on the host in, log-likelihoods out): text to the "
                       "GPU, k_newick, k_plan, k_eigen, k_pmatrix, k_prune, k_edge, k_hmm; Gamma rates on the host threads"
                       + ("; then the ranks' results gathered over NCCL onto rank 0's host" if world > 1 else "")},
            "e2e_arrays": None if arr_s is None else {
                "value": B_total * steps / arr_max, "unit": UNIT, "h2d_bytes_per_step": h2d_a, "d2h_bytes_per_step": d2h_a,
                "last_call_ms_rank0": arr_detail,
                "api": "lh_score_batch (pre-flattened operation arrays + pre-computed rates in, each rank's log-likelihoods on its host out)"},
            "e2e_pipeline": pipe,
            "gpu_launches": launches_total + e2e_launches, "clocks": clocks, "parity": parity, "parity_vs_n1": parity_vs_n1,
            "finite_fraction": float(np.mean(np.isfinite(full_value))),
        }
    f.close()
    phmm.close()
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--config", default="c4")
    ap.add_argument("--batch", type=int, default=0, help="override the whole-job batch size")
    ap.add_argument("--topologies", type=int, default=256, help="distinct tree topologies in the batch")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--pageable-text", dest="pageable_text", action="store_true",
                    help="e2e: hold the Newick text in pageable instead of pinned host memory")
    ap.add_argument("--no-extras", action="store_true", help="skip the c3 / c5 records of the default run")
    ap.add_argument("--opt", action="append", default=[], metavar="KEY=VALUE",
                    help="lh_set_option on the family (e.g. pipe_waves=4); recorded in config")
    ap.add_argument("--families", type=int, default=0, help="config c5: families scored per pass (whole job; default 1024)")
    ap.add_argument("--slots", type=int, default=6, help="config c5: device handles per GPU")
    ap.add_argument("--topologies-c5", dest="topologies_c5", type=int, default=16,
                    help="config c5: distinct topologies per family")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus and world > 1:
        args.gpus = world

    if args.impl == "reference":
        if rank != 0:
            return 0
        if args.config in ("c5", "c2"):
            print(json.dumps({"impl": "reference", "unavailable": "the reference arm is timed on config c4 (bench.py default)"}))
            return 0
        return reference_arm(args)

    # stdout carries the ONE JSON line and nothing else: whatever the libraries print there (NCCL's version banner,
    # ...) goes to stderr
    sys.stdout.flush()
    json_fd = os.dup(1)
    os.dup2(2, 1)
    env = Env(args)
    if args.config in ("c5", "c2"):
        import bench_c2
        import bench_c5
        out = (bench_c5.run(env, args) if args.config == "c5" else bench_c2.run(env, args))
    else:
        out = bench_family(env, args, args.config, args.steps, args.warmup, True)
        if args.config == "c4" and not args.no_extras and not args.batch:
            # the other single-box configurations, so that one driver-run line carries them all
            c3 = bench_family(env, args, "c3", 3, 3, False)
            import bench_c5
            c5_args = argparse.Namespace(**vars(args))
            c5_args.families = args.families or 1024
            c5_args.steps, c5_args.warmup, c5_args.batch = 2, 3, 0
            c5 = bench_c5.run(env, c5_args)
            if env.rank == 0:
                keep = ("value", "unit", "ms_per_step", "steps", "warmup", "config", "roofline", "e2e", "e2e_arrays", "e2e_pipeline", "parity", "cpu_baseline",
                        "parity_vs_n1", "gpu_launches", "families_per_s", "value_is_end_to_end")
                out["c3"] = {k: c3[k] for k in keep if k in c3}
                out["c5"] = {k: c5[k] for k in keep if k in c5}
                out["gpu_launches"] += c3["gpu_launches"] + c5["gpu_launches"]
    if env.rank == 0:
        sys.stdout.flush()
        os.write(json_fd, (json.dumps(out) + "\n").encode())
    if env.dist is not None:
        env.dist.barrier()
        env.dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
