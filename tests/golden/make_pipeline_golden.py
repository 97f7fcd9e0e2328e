"""Generates tests/golden/pipeline_tiny.npz with the REFERENCE's own C++ (oracle/_ref): for the synthetic
"tiny" family of linearham_b200.workload, a RevBayes-style samples TSV, the output TSV the reference's
PhyloHMM::RunPipeline (src/PhyloHMM.cpp:393-446) writes for it (seed 0), and for the first sample the
--compute-logl / --sample call sequence (src/linearham.cpp:322-389): LogLikelihood() and five
SampleNaiveSequence() draws (seed 0).

    python tests/golden/make_pipeline_golden.py     # needs /root/reference
"""
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from linearham_b200 import synth, workload  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

N_SAMPLES, R = 12, 4
cfg = workload.CONFIGS["tiny"]
with tempfile.TemporaryDirectory() as d:
    ff = workload.write_family_files(cfg, d)
    samples = synth.tree_samples(np.random.default_rng(cfg["tree_seed"]), ff["tree"], N_SAMPLES)
    tsv_in, tsv_out = os.path.join(d, "in.trees"), os.path.join(d, "out.trees")
    synth.write_tsv(tsv_in, samples)
    ref = po.RefPhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    ref.run_pipeline(tsv_in, tsv_out, R)
    ref.close()
    # single-sample API on a fresh object (fresh RNG)
    nwk_path = os.path.join(d, "first.tree")
    import re
    first_newick = re.sub(r"\[&index=[0-9]+\]", "", samples[0]["newick"])  # a plain Newick file, as --newick-path expects
    open(nwk_path, "w").write(first_newick)
    ref = po.RefPhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    ref.init_params_file(nwk_path, samples[0]["er"], samples[0]["pi"], samples[0]["alpha"], R)
    ll = ref.loglikelihood()
    draws = [ref.sample_naive() for _ in range(5)]
    np.savez_compressed(os.path.join(HERE, "pipeline_tiny.npz"), sha=ff["sha"], tsv_in=open(tsv_in).read(),
                        tsv_out=open(tsv_out).read(), first_newick=first_newick, first_er=samples[0]["er"],
                        first_pi=samples[0]["pi"], first_alpha=samples[0]["alpha"], first_loglik=ll, first_draws=np.array(draws),
                        num_rates=R)
    print("loglik", ll, "draw0", draws[0][:60], "...")
    print(open(tsv_out).read()[:600])
