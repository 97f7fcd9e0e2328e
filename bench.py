#!/usr/bin/env python
"""bench.py - PhyloHMM tree-likelihood evaluations per second (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl b200|reference] [--config c4|c3]

A step is one pass of the hot path over one batch of synthetic tree samples of a fixed clonal
family: for every (tree, er, pi, alpha) sample, Felsenstein pruning over the xMSA + the scaled HMM
forward pass = one `InitializePhyloEmission() + LogLikelihood()` of the reference
(src/PhyloHMM.cpp:429-430). Default workload = BASELINE.json configs[3] ("c4": synthetic 1000-leaf
IGH family, 400 nt, Gamma-4, batch 16384 trees sharded over the N GPUs).

  value     trees/s, kernels only, batch already resident in HBM (lh_batch_upload once, then
            lh_batch_run per step); CUDA events on the library's stream, max over ranks
  e2e       trees/s through lh_score_batch with HOST buffers: host traversal planning, H2D from
            pinned staging, kernels, D2H of the log-likelihoods, all inside the timed region
  roofline  k_prune: algorithmic bytes / measured launch time vs measured HBM peak
  cpu_baseline  the reference's own C++ (oracle/_ref) on all host cores, bounded sample (N=1 only)

--impl reference times the reference's CPU implementation alone (rank 0 only).
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


def cpu_arm_setup(cfg_name, cfg, ff, fam, tb, newicks, alpha, n_sample, workdir):
    """Writes the worker data file; returns (kind, path)."""
    from oracle import pyoracle as po
    path = os.path.join(workdir, "cpu_sample.npz")
    if po.have_ref():
        np.savez(path, yaml_path=ff["yaml_path"], hmm_dir=ff["hmm_dir"], newicks=np.array(newicks[:n_sample]),
                 er=tb["er"][:n_sample], pi=tb["pi"][:n_sample], alpha=alpha[:n_sample], R=cfg["R"])
        return "reference", path
    po.build_port()
    data = {"fam_" + k: np.asarray(v) for k, v in fam.items()}
    data.update({"tb_" + k: v[:n_sample] for k, v in tb.items()})
    np.savez(path, **data)
    return "port", path


def cpu_sample_size(cfg_name, cores):
    # c4: ~0.14 s/tree/core, c3: ~10 ms/tree/core -> about 5-10 s of wall time
    per_core = {"c4": 32, "c3": 256}.get(cfg_name, 64)
    return per_core * cores


# ------------------------------------------------------------------------------ main ---

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
    ap.add_argument("--no-newick", action="store_true", help="skip the Newick-string class-API measurement")
    ap.add_argument("--newick-trees", type=int, default=4096, help="tree samples per step of e2e_newick")
    ap.add_argument("--opt", action="append", default=[], metavar="KEY=VALUE",
                    help="lh_set_option on the family (e.g. pipe_waves=4); recorded in config")
    ap.add_argument("--families", type=int, default=256, help="config c5: families scored per pass (whole job)")
    ap.add_argument("--slots", type=int, default=6, help="config c5: device handles per GPU")
    ap.add_argument("--topologies-c5", dest="topologies_c5", type=int, default=16,
                    help="config c5: distinct topologies per family")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 3) if args.impl == "b200" else args.warmup

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if world != args.gpus and world > 1:
        args.gpus = world
    if args.impl == "reference" and rank != 0:
        return 0

    if args.config in ("c5", "c2"):
        if args.impl == "reference":
            print(json.dumps({"impl": "reference", "unavailable": "the reference arm is timed on config c4 (bench.py default)"}))
            return 0
        import bench_c2
        import bench_c5
        return (bench_c5 if args.config == "c5" else bench_c2).run(args, rank, world, local_rank)

    from linearham_b200 import host, sharding, workload
    cfg = dict(workload.CONFIGS[args.config])
    B_total = args.batch or cfg["B"]
    workdir = tempfile.mkdtemp(prefix="lhbench_")
    ff = workload.write_family_files(cfg, workdir)
    # per-family host work of the product: partis YAML + HMM parameter dir -> state space, transition
    # matrices, xMSA (C++ linearham::PhyloHMM of this repo; bit-exact vs the reference, tests/test_host_model.py)
    phmm = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0, device=local_rank if world > 1 else 0)
    fam, labels = phmm.family(), phmm.strs("xmsa_labels")
    if labels != ff["labels"]:
        raise RuntimeError("xMSA labels do not match the generated family")
    T, S, R = fam["n_tips"], fam["n_sites"], cfg["R"]
    tb, newicks, alpha = workload.tree_batch(cfg, ff, B_total, n_topo=args.topologies)
    cores = os.cpu_count() or 1
    config = {"workload": f"{args.config}: {cfg['desc']}", "n_tips": T, "xmsa_columns": S, "gamma_categories": R,
              "global_batch": B_total, "distinct_topologies": min(args.topologies, B_total),
              "hmm": {"vd": [fam["vd_sites"], fam["vd_states"]], "dj": [fam["dj_sites"], fam["dj_states"]],
                      "genes": [fam["n_v"], fam["n_d"], fam["n_j"]]},
              "parallelism": f"trees sharded over {args.gpus} GPU(s), no data-path collective; "
                             "NCCL all-gather of the per-tree log-likelihoods",
              "l2": "inputs larger than L2"}

    # ---------------------------------------------------------------- reference arm ---
    if args.impl == "reference":
        n_step = max(cores, min(len(newicks), {"c4": 4, "c3": 64}.get(args.config, 16) * cores))
        kind, path = cpu_arm_setup(args.config, cfg, ff, fam, tb, newicks, alpha, n_step, workdir)
        pool = CpuPool(kind, path, cores)
        for _ in range(args.warmup):
            pool.score(0, n_step)
        total = 0.0
        for _ in range(args.steps):
            w, _, stages = pool.score(0, n_step)
            total += w
        pool.close()
        v = n_step * args.steps / total
        sample = f"{n_step} tree samples per step (the first {n_step} of the batch), {cores} worker processes"
        print(json.dumps({
            "impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * total / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": config,
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": kind, "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}))
        return 0

    # ------------------------------------------------------------------- B200 arm ---
    import torch
    from linearham_b200 import device
    if not torch.cuda.is_available() or device.device_count() < 1:
        raise RuntimeError("bench.py needs a CUDA device: linearham_b200 has no CPU fallback")
    dist = None
    if world > 1:
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("nccl", device_id=torch.device("cuda", local_rank))
    dev = local_rank if world > 1 else 0

    # CPU baseline first (before this process touches the GPU heavily); rank 0, N=1 only
    cpu_baseline, parity = None, None
    cpu_ll = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        n_cpu = min(len(newicks), cpu_sample_size(args.config, cores))
        kind, path = cpu_arm_setup(args.config, cfg, ff, fam, tb, newicks, alpha, n_cpu, workdir)
        pool = CpuPool(kind, path, cores)
        pool.score(0, cores)  # warm the workers (page in the library, first allocation)
        w, cpu_ll, stages = pool.score(0, n_cpu)
        pool.close()
        cpu_baseline = {"value": n_cpu / w, "unit": UNIT, "cores": cores, "kind": kind,
                        "sample": f"first {n_cpu} tree samples of the batch, one worker process per core, "
                                  f"{w:.2f} s wall",
                        "stage_core_seconds": stages}

    # shard the batch: contiguous, equal chunks (trees of one family cost the same)
    per = sharding.per_rank(B_total, world)
    shard, lo, hi = sharding.shard_batch(tb, rank, world)
    Bs = hi - lo

    f = device.Family(fam, device=dev)
    for kv in args.opt:
        k, v = kv.split("=")
        f.set_option(k, int(v))
    if args.opt:
        config["options"] = list(args.opt)
    gathered = None
    # ranks of one node share its host cores: each takes its share for traversal planning and Newick ingest
    host_threads = max(2, cores // int(os.environ.get("LOCAL_WORLD_SIZE", world))) if world > 1 else cores
    if world > 1:
        f.set_option("stream", torch.cuda.current_stream().cuda_stream)
        f.set_option("host_threads", host_threads)
        phmm.set_host_threads(host_threads)

    class _DevArr:
        def __init__(self, ptr, n):
            self.__cuda_array_interface__ = {"shape": (n,), "typestr": "<f8", "data": (ptr, False), "version": 2}

    def sync_all():
        f.sync()
        torch.cuda.synchronize()
        if dist is not None:
            dist.barrier()
            torch.cuda.synchronize()

    def gather_step():
        nonlocal gathered
        if dist is None:
            return
        ptr, n = f.loglik_dev()
        local = torch.as_tensor(_DevArr(ptr, n), device=f"cuda:{dev}")
        if gathered is None:
            gathered = torch.empty(world * per, dtype=torch.float64, device=f"cuda:{dev}")
        sharding.gather_loglik(dist, local, B_total, out=gathered)

    sampler = ClockSampler(dev)
    # -------- value: batch resident in HBM -------------------------------------
    f.upload(shard)
    for _ in range(args.warmup):
        f.run()
        gather_step()
    sync_all()
    sampler.start()
    f.mark(0)
    acc = {"ms_prune": 0.0, "ms_pmatrix": 0.0, "ms_repeats": 0.0, "ms_edge": 0.0, "ms_hmm": 0.0, "prune_launches": 0, "kernel_launches": 0}
    for _ in range(args.steps):
        f.run()
        gather_step()
        tm = f.timings()
        for k in acc:
            acc[k] += tm[k]
    f.mark(1)
    sync_all()
    ms_kernels = f.elapsed_ms(0, 1)
    ll_dev = f.fetch()

    # -------- e2e: host buffers in, host log-likelihoods out --------------------
    e2e_s, h2d, d2h, ms_plan, e2e_detail = None, 0, 0, 0.0, None
    if not args.no_e2e:
        for _ in range(max(1, min(args.warmup, 2))):
            f.score(shard)
        sync_all()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ll_e2e = f.score(shard)
            t2 = f.timings()
            h2d, d2h, ms_plan = t2["h2d_bytes"], t2["d2h_bytes"], t2["ms_plan"]
            e2e_detail = {k: t2[k] for k in ("ms_call", "ms_setup", "ms_plan", "ms_h2d", "ms_kernels", "ms_d2h",
                                             "ms_plan_dev", "device_plan")}
        sync_all()
        e2e_s = time.perf_counter() - t0
        assert np.array_equal(ll_e2e, ll_dev, equal_nan=True)
    # -------- e2e from Newick strings: the class API a linearham user calls -------------
    # PhyloHMM::ScoreSamples(newick strings, er, pi, alpha) -> log-likelihoods: Newick ingest on the host
    # threads overlapped with lh_score_batch of the previous chunk. The strings are marshalled to a C array
    # once, outside the timed region (a C++ caller holds them as char* already).
    nw_s, nw_n = None, 0
    if not args.no_e2e and not args.no_newick:
        nw_n = min(Bs, args.newick_trees)
        nt = len(newicks)
        arr = host.encode_newicks([newicks[(lo + i) % nt] for i in range(nw_n)])
        er_n, pi_n, al_n = shard["er"][:nw_n], shard["pi"][:nw_n], alpha[lo:lo + nw_n]
        ll_nw = phmm.ScoreSamples(arr, er_n, pi_n, al_n, R)
        sync_all()
        t0 = time.perf_counter()
        for _ in range(args.steps):
            ll_nw = phmm.ScoreSamples(arr, er_n, pi_n, al_n, R)
        sync_all()
        nw_s = time.perf_counter() - t0
        k = max(0, min(nw_n, nt - lo))  # samples whose branch lengths are the strings' own
        assert np.array_equal(ll_nw[:k], ll_dev[:k], equal_nan=True)
    clocks = sampler.stop()

    # -------- max over ranks ------------------------------------------------------
    if dist is not None:
        t = torch.tensor([ms_kernels, e2e_s or 0.0, acc["ms_prune"], nw_s or 0.0], dtype=torch.float64, device=f"cuda:{dev}")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms_kernels, e2e_max, _, nw_max = t.tolist()
        e2e_s = e2e_max if e2e_s is not None else None
        nw_s = nw_max if nw_s is not None else None
        full = gathered[:B_total].cpu().numpy() if gathered is not None else ll_dev
        lt = torch.tensor([acc["kernel_launches"]], dtype=torch.float64, device=f"cuda:{dev}")
        dist.all_reduce(lt, op=dist.ReduceOp.SUM)
        launches_total = int(lt.item())
    else:
        full = ll_dev
        launches_total = acc["kernel_launches"]

    if rank == 0:
        value = B_total * args.steps / (ms_kernels * 1e-3)
        peak, peak_src = measured_peaks()
        bytes_tree = workload.bytes_per_tree(T, S, R)
        # the kernel walks the distinct site patterns when it can; utilisation counts executed flop
        cols = int(tm["prune_columns"])
        flops_tree = workload.flops_per_tree(T, cols, R)
        config["prune_columns"] = cols
        config["pattern_trees_rank0"] = int(tm["pattern_trees"])
        config["site_repeats"] = bool(tm["site_repeats"])
        if tm["site_repeats"]:
            config["sr_arena_entries"], config["sr_ctas_per_sm"] = int(tm["sr_arena"]), int(tm["sr_ctas_per_sm"])
        prune_s = acc["ms_prune"] * 1e-3
        achieved = bytes_tree * Bs * args.steps / prune_s / 1e9
        fp64_peak = device.fp64_peak_tflops(dev)
        traffic = None
        tp = os.path.join(ROOT, "profiles", f"k_prune_traffic_{args.config}.json")
        if os.path.exists(tp):
            try:
                # the ncu capture ran one launch of `trees_in_launch` trees: scale to this run's launch size
                tj = json.load(open(tp))
                traffic = tj["dram_bytes_per_launch"] / tj["trees_in_launch"] * (Bs / max(1, acc["prune_launches"] // args.steps))
            except Exception:
                traffic = None
        if cpu_ll is not None:
            n = len(cpu_ll)
            ok = np.isfinite(cpu_ll)
            parity = {"n": int(n), "max_rel_err_vs_cpu": float(np.max(np.abs(full[:n][ok] - cpu_ll[ok]) / np.abs(cpu_ll[ok]))),
                      "tolerance": 1e-9, "non_finite_class_equal": bool(np.array_equal(np.isfinite(full[:n]), ok)),
                      "non_finite": int(n - ok.sum())}
        out = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
            "ms_per_step": ms_kernels / args.steps, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
            "dtype": "f64", "data": "synthetic", "config": config, "impl": "b200",
            "roofline": {"bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                         "traffic": traffic, "traffic_source": "ncu --set full dram__bytes_read.sum + dram__bytes_write.sum of one k_prune launch "
                         "(profiles/k_prune_traffic_<config>.json), per tree x trees_per_launch",
                         "kernel": "k_prune", "peak_source": peak_src,
                         "algorithmic_bytes_per_tree": bytes_tree, "trees_per_launch": Bs / max(1, acc["prune_launches"] // args.steps),
                         "avg_launch_ms": acc["ms_prune"] / max(1, acc["prune_launches"]),
                         "note": "inner CLVs live in registers/shared memory (on-chip stack), so DRAM traffic is far "
                                 "below the streaming-model bytes and frac may exceed 1; the binding limits are the "
                                 "FP64 pipe and the L1 data pipe. achieved counts the streaming-model bytes of all "
                                 "xmsa_columns; fp64 counts the flop executed on prune_columns",
                         "fp64": {"achieved_tflops": flops_tree * Bs * args.steps / prune_s / 1e12, "peak_tflops": fp64_peak,
                                  "frac": flops_tree * Bs * args.steps / prune_s / 1e12 / fp64_peak,
                                  "peak_source": "lh_measure_fp64_peak (DFMA loop, this run)"}},
            "stage_ms_per_step": {"k_pmatrix": acc["ms_pmatrix"] / args.steps, "k_repeats": acc["ms_repeats"] / args.steps, "k_prune": acc["ms_prune"] / args.steps,
                                  "k_edge": acc["ms_edge"] / args.steps, "k_hmm": acc["ms_hmm"] / args.steps, "rank": 0},
            "cpu_baseline": cpu_baseline,
            "e2e": None if e2e_s is None else {"value": B_total * args.steps / e2e_s, "unit": UNIT,
                                               "h2d_bytes_per_step": int(h2d) * world, "d2h_bytes_per_step": int(d2h) * world,
                                               "host_plan_ms_per_step_rank0": ms_plan, "host_threads_per_rank": host_threads, "last_call_ms_rank0": e2e_detail,
                                               "api": "lh_score_batch (host arrays in, host log-likelihoods out)"},
            "e2e_newick": None if nw_s is None else {
                "value": nw_n * world * args.steps / nw_s, "unit": UNIT, "trees_per_step": nw_n * world,
                "host_threads": host_threads,
                "api": "PhyloHMM::ScoreSamples (Newick strings + er/pi/alpha in, log-likelihoods out; "
                       "Newick ingest on host threads overlapped with lh_score_batch)"},
            "gpu_launches": launches_total, "clocks": clocks, "parity": parity,
            "finite_fraction": float(np.mean(np.isfinite(full))),
        }
        print(json.dumps(out))
    f.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return 0


if __name__ == "__main__":
    sys.exit(main())
