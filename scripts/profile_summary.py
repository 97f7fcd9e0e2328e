"""Summarises one GPU round's ncu artefacts (gpurun_out/) into profiles/<tag>_*.txt|json.

    python scripts/profile_summary.py <tag> [config]

  launches_<tag>.csv      -> profiles/<tag>_launches_summary.txt  (per-kernel share of device time)
  prof_prune_<tag>.ncu-rep -> profiles/<tag>_k_prune_metrics.txt  (selected raw metrics)
                              profiles/k_prune_traffic_<config>.json (DRAM bytes per launch, read by bench.py)
"""
import collections, csv, json, os, subprocess, sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
tag = sys.argv[1]
config = sys.argv[2] if len(sys.argv) > 2 else "c4"
G, P = os.path.join(ROOT, "gpurun_out"), os.path.join(ROOT, "profiles")
os.makedirs(P, exist_ok=True)

lc = os.path.join(G, f"launches_{tag}.csv")
if os.path.exists(lc):
    rows = [r for r in csv.reader(open(lc)) if len(r) > 5]
    hdr, agg = None, collections.defaultdict(lambda: [0, 0.0])
    for r in rows:
        if r[0] == "ID":
            hdr = r
            continue
        if hdr is None:
            continue
        d = dict(zip(hdr, r))
        v = float(d["Metric Value"].replace(",", ""))
        v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}[d["Metric Unit"]]
        name = d["Kernel Name"].split("(")[0].replace("void ", "")
        agg[name][0] += 1
        agg[name][1] += v
    tot = sum(v[1] for v in agg.values())
    with open(os.path.join(P, f"{tag}_launches_summary.txt"), "w") as f:
        f.write(f"# ncu --metrics gpu__time_duration.sum --clock-control none, command: python bench.py --no-cpu-baseline\n")
        f.write(f"# source: gpurun_out/launches_{tag}.csv (cold-cache, serialised launches: compare shares)\n")
        f.write(f"{'kernel':40s} {'launches':>8s} {'total_ms':>12s} {'share':>8s}\n")
        for k, v in sorted(agg.items(), key=lambda x: -x[1][1]):
            f.write(f"{k:40s} {v[0]:8d} {v[1]:12.3f} {v[1] / tot:8.4f}\n")
    print(open(os.path.join(P, f"{tag}_launches_summary.txt")).read())

rep = os.path.join(G, f"prof_prune_{tag}.ncu-rep")
if os.path.exists(rep):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    want = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
            "launch__registers_per_thread", "launch__grid_size", "launch__block_size", "launch__occupancy_limit_registers",
            "launch__occupancy_limit_shared_mem", "launch__shared_mem_per_block_dynamic", "sm__warps_active.avg.pct_of_peak_sustained_active",
            "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
            "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
            "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
            "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
            "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
            "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum", "sm__cycles_elapsed.max"]
    want += [h for h in hdr if h.startswith("smsp__average_warps_issue_stalled") and h.endswith("_per_issue_active.ratio") and "not_issued" not in h]
    out = {}
    with open(os.path.join(P, f"{tag}_k_prune_metrics.txt"), "w") as f:
        f.write(f"# ncu --set full --clock-control none --import-source on -k regex:k_prune (one launch), source gpurun_out/prof_prune_{tag}.ncu-rep\n")
        f.write(f"# kernel: {vals[hdr.index('Kernel Name')]}  grid {vals[hdr.index('launch__grid_size')]} x block {vals[hdr.index('launch__block_size')]}\n")
        for h, u, v in zip(hdr, units, vals):
            if h in want:
                f.write(f"{h:95s} {u:16s} {v}\n")
                out[h] = (v, u)
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}
    rd = float(out["dram__bytes_read.sum"][0]) * scale[out["dram__bytes_read.sum"][1]]
    wr = float(out["dram__bytes_write.sum"][0]) * scale[out["dram__bytes_write.sum"][1]]
    meta = {"dram_bytes_per_launch": rd + wr, "dram_read": rd, "dram_write": wr, "tag": tag,
            "note": "one k_prune launch of `python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-e2e --batch 4096` "
                    "(4096 trees in the launch)", "trees_in_launch": 4096}
    json.dump(meta, open(os.path.join(P, f"k_prune_traffic_{config}.json"), "w"), indent=1)
    print(open(os.path.join(P, f"{tag}_k_prune_metrics.txt")).read())
