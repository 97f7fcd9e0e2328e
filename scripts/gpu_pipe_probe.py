"""RunPipeline on the c3 trace with different batch sizes (rows/s, stage busy times with LH_TRACE=1)."""
import os, sys, tempfile, time
sys.path.insert(0, ".")
import numpy as np
from linearham_b200 import host, workload, synth
cfg = dict(workload.CONFIGS["c3"])
n_rows = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
d = tempfile.mkdtemp()
ff = workload.write_family_files(cfg, d)
tb, newicks, alpha = workload.tree_batch(cfg, ff, n_rows, n_topo=256)
text = host.NewickBuffer(tb, ff["labels"])
rows = [{"newick": text[i], "er": tb["er"][i], "pi": tb["pi"][i], "alpha": float(alpha[i])} for i in range(n_rows)]
tsv = os.path.join(d, "trace.tsv"); synth.write_tsv(tsv, rows)
out = os.path.join(d, "out.tsv")
ref_text = None
for bs in (256, 1024):
    ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    ph.RunPipeline(tsv, out, 4, bs)
    first = open(out).read()
    if ref_text is None: ref_text = first
    t0 = time.perf_counter()
    for _ in range(3): ph.RunPipeline(tsv, out, 4, bs)
    dt = (time.perf_counter() - t0) / 3
    ph.close()
    print(f"batch {bs}: {n_rows / dt:.0f} rows/s  ({dt*1e3:.1f} ms per run)  same output as batch 1024: {first == ref_text}", flush=True)
