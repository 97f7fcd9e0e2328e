"""RunPipeline at the C3 shape: the device sampler's output file against the host sampler's, 4096 rows."""
import os, sys, tempfile, time
sys.path.insert(0, ".")
from linearham_b200 import host, workload, synth
cfg = dict(workload.CONFIGS["c3"]); n_rows = 4096
d = tempfile.mkdtemp(); ff = workload.write_family_files(cfg, d)
tb, newicks, alpha = workload.tree_batch(cfg, ff, n_rows, n_topo=256)
text = host.NewickBuffer(tb, ff["labels"])
rows = [{"newick": text[i], "er": tb["er"][i], "pi": tb["pi"][i], "alpha": float(alpha[i])} for i in range(n_rows)]
tsv = os.path.join(d, "trace.tsv"); synth.write_tsv(tsv, rows)
res = {}
for dev in (False, True):
    ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0); ph.set_device_sampler(dev)
    out = os.path.join(d, f"out_{dev}.tsv"); t0 = time.perf_counter(); ph.RunPipeline(tsv, out, 4); dt = time.perf_counter() - t0
    nxt = [ph.SampleNaiveSequence() for _ in range(2)]; ph.close()
    res[dev] = (open(out).read(), nxt); print("device sampler" if dev else "host sampler", f"{dt*1e3:.0f} ms (first run, cold)")
print("same file:", res[True][0] == res[False][0], " same next draws:", res[True][1] == res[False][1], " rows:", res[True][0].count("\n") - 1)
