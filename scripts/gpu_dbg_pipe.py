import os, sys, tempfile
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
from linearham_b200 import host, synth, workload
d = tempfile.mkdtemp()
ff = workload.write_family_files(workload.CONFIGS["tiny"], d)
samples = synth.tree_samples(np.random.default_rng(77), ff["tree"], 300, outgroup="naive")
tin = os.path.join(d, "in300.trees"); synth.write_tsv(tin, samples)
outs = []
for threads in (1, 0):
    tout = os.path.join(d, f"o{threads}.trees")
    h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0); h.set_host_threads(threads)
    h.RunPipeline(tin, tout, 4, batch_size=128); h.close()
    outs.append(open(tout).read().splitlines())
hdr = outs[0][0].split("\t")
n = 0
for li, (a, b) in enumerate(zip(outs[0], outs[1])):
    if a != b:
        fa, fb = a.split("\t"), b.split("\t")
        diff = [(hdr[k] if k < len(hdr) else k, fa[k][:60], fb[k][:60]) for k in range(min(len(fa), len(fb))) if fa[k] != fb[k]]
        print("line", li, "fields", len(fa), len(fb), diff[:6])
        n += 1
        if n > 5: break
print("lines", len(outs[0]), len(outs[1]), "different", sum(a != b for a, b in zip(outs[0], outs[1])))
