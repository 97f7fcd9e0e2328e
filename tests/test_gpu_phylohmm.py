"""The product's PhyloHMM class (C++, reference API) on the GPU against outputs of the reference's own
C++ stored in tests/golden/pipeline_tiny.npz (made by tests/golden/make_pipeline_golden.py):
the --compute-logl / --sample call sequence and the batched RunPipeline.
Tolerances: log-likelihood 1e-9 relative; printed TSV numbers 2e-6 relative (6 significant digits
are printed); every sampled string (naive sequence, genes, insertions, deletions) identical."""
import os
import tempfile

import numpy as np
import pytest

from conftest import GOLDEN_DIR
from linearham_b200 import host, workload

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def tiny():
    z = np.load(os.path.join(GOLDEN_DIR, "pipeline_tiny.npz"), allow_pickle=False)
    d = tempfile.mkdtemp(prefix="lh_tiny_")
    ff = workload.write_family_files(workload.CONFIGS["tiny"], d)
    assert ff["sha"] == str(z["sha"])
    return z, ff, d


def test_compute_logl_and_sample_like_the_reference_cli(tiny):
    """src/linearham.cpp:322-389: ctor, InitializePhyloParameters, InitializePhyloEmission, LogLikelihood,
    N x SampleNaiveSequence; as the reference's Catch test does on its fixture (test/test.cpp:755-758, 968-974)."""
    z, ff, d = tiny
    path = os.path.join(d, "first.tree")
    open(path, "w").write(str(z["first_newick"]))
    h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    h.InitializePhyloParameters(path, z["first_er"], z["first_pi"], float(z["first_alpha"]), int(z["num_rates"]))
    h.InitializePhyloEmission()
    ll = h.LogLikelihood()
    assert abs(ll - float(z["first_loglik"])) <= 1e-9 * abs(float(z["first_loglik"]))
    draws = [h.SampleNaiveSequence() for _ in range(5)]
    assert draws == [str(s) for s in z["first_draws"]]
    assert h.LogLikelihood() == ll   # cached; sampling does not disturb it
    h.close()


@pytest.mark.parametrize("batch_size", [1, 5, 1024])
def test_run_pipeline_matches_reference_output(tiny, batch_size):
    z, ff, d = tiny
    tin, tout = os.path.join(d, "in.trees"), os.path.join(d, f"out_{batch_size}.trees")
    open(tin, "w").write(str(z["tsv_in"]))
    h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    h.RunPipeline(tin, tout, int(z["num_rates"]), batch_size=batch_size)
    h.close()
    got = [l.split("\t") for l in open(tout).read().splitlines()]
    exp = [l.split("\t") for l in str(z["tsv_out"]).splitlines()]
    assert got[0] == exp[0]                      # header
    assert len(got) == len(exp) == 13
    for g, e in zip(got[1:], exp[1:]):
        assert len(g) == len(e)
        for name, a, b in zip(exp[0], g, e):
            try:
                fb = float(b)
            except ValueError:
                assert a == b, name              # tree, naive sequence, genes, insertions
                continue
            assert abs(float(a) - fb) <= 2e-6 * max(1.0, abs(fb)), (name, a, b)


def test_score_samples_equals_pipeline_loglik(tiny):
    z, ff, d = tiny
    rows = [l.split("\t") for l in str(z["tsv_in"]).splitlines()]
    hdr, rows = rows[0], rows[1:]
    c = {n: i for i, n in enumerate(hdr)}
    newicks = [r[c["tree"]] for r in rows]
    er = np.array([[float(r[c[f"er[{i}]"]]) for i in range(1, 7)] for r in rows])
    pi = np.array([[float(r[c[f"pi[{i}]"]]) for i in range(1, 5)] for r in rows])
    alpha = np.array([float(r[c["alpha"]]) for r in rows])
    h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    ll = h.ScoreSamples(newicks, er, pi, alpha, int(z["num_rates"]))
    h.close()
    out = [l.split("\t") for l in str(z["tsv_out"]).splitlines()]
    k = out[0].index("LHLogLikelihood")
    exp = np.array([float(r[k]) for r in out[1:]])
    assert np.allclose(ll, exp, rtol=6e-6, atol=0)   # 6 significant digits are printed
    assert abs(ll[0] - float(z["first_loglik"])) <= 1e-9 * abs(float(z["first_loglik"]))


def test_score_samples_chunks_overlap_ingest_and_match_one_by_one(tiny):
    """More samples than one ingest chunk (256 first, then 1024): chunk c+1 is parsed while chunk c is scored; the result
    is the per-sample result, and a malformed string in a LATER chunk still raises."""
    z, ff, d = tiny
    rows = [l.split("\t") for l in str(z["tsv_in"]).splitlines()]
    hdr, rows = rows[0], rows[1:]
    c = {n: i for i, n in enumerate(hdr)}
    base = [r[c["tree"]] for r in rows]
    n = 5000
    rng = np.random.default_rng(3)
    newicks = [base[i % len(base)] for i in range(n)]
    er = rng.dirichlet(np.ones(6), n)
    pi = np.maximum(rng.dirichlet(np.ones(4), n), 0.05)
    pi /= pi.sum(axis=1, keepdims=True)
    alpha = np.maximum(0.1, rng.exponential(1.0, n))
    R = int(z["num_rates"])
    h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    ll = h.ScoreSamples(host.encode_newicks(newicks), er, pi, alpha, R)
    for lo, hi in ((0, 7), (250, 260), (1270, 1290), (4990, 5000)):
        part = h.ScoreSamples(newicks[lo:hi], er[lo:hi], pi[lo:hi], alpha[lo:hi], R)
        assert np.array_equal(part, ll[lo:hi])
    bad = list(newicks)
    bad[4500] = "((0:0.1,1:0.1):0.1,naive:0.1,nosuchtip:0.1);"
    with pytest.raises(host.HostError):
        h.ScoreSamples(bad, er, pi, alpha, R)
    again = h.ScoreSamples(newicks[:9], er[:9], pi[:9], alpha[:9], R)   # the handle stays usable
    assert np.array_equal(again, ll[:9])
    h.close()


def test_errors_surface_as_exceptions(tiny):
    z, ff, d = tiny
    h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
    with pytest.raises(host.HostError):
        h.InitializePhyloEmission()                           # no tree yet
    bad = os.path.join(d, "bad.tree")
    open(bad, "w").write("((0:0.1,1:0.1):0.1,naive:0.1,nosuchtip:0.1);")
    h.InitializePhyloParameters(bad, [1] * 6, [0.25] * 4, 1.0, 4)
    with pytest.raises(host.HostError):                       # label mismatch (pll_partition.cpp:157-189)
        h.InitializePhyloEmission()
    with pytest.raises(host.HostError):                       # gamma.c:238
        h.InitializePhyloParameters(bad, [1] * 6, [0.25] * 4, 0.001, 4)
    h.close()


# ---- the command line (linearham_b200/linearham; flags of src/linearham.cpp:268-455) ------------------
CLI = os.path.join(os.path.dirname(os.path.abspath(host.__file__)), "linearham")


def _cli(*args):
    import subprocess
    return subprocess.run([CLI, *[str(a) for a in args]], capture_output=True, text=True, timeout=300)


def _tree_flags(z, path):
    flags = ["--newick-path", path, "--alpha", repr(float(z["first_alpha"])), "--num-rates", int(z["num_rates"])]
    for x in z["first_er"]:
        flags += ["--er", repr(float(x))]
    for x in z["first_pi"]:
        flags += ["--pi", repr(float(x))]
    return flags


def test_cli_compute_logl_sample_and_pipeline(tiny):
    z, ff, d = tiny
    path = os.path.join(d, "cli_first.tree")
    open(path, "w").write(str(z["first_newick"]))
    common = ["--yaml-path", ff["yaml_path"], "--cluster-ind", 0, "--hmm-param-dir", ff["hmm_dir"], "--seed", 0]
    r = _cli("--compute-logl", *common, *_tree_flags(z, path))
    assert r.returncode == 0, r.stderr
    # std::cout << double prints 6 significant digits (src/linearham.cpp:344)
    assert r.stdout.strip() == "%g" % float(z["first_loglik"])
    r = _cli("--sample", *common, *_tree_flags(z, path), "--N", 5)
    assert r.returncode == 0, r.stderr
    assert r.stdout.split() == [str(s) for s in z["first_draws"]]
    tin, tout = os.path.join(d, "cli_in.trees"), os.path.join(d, "cli_out.trees")
    open(tin, "w").write(str(z["tsv_in"]))
    r = _cli("--pipeline", *common, "--num-rates", int(z["num_rates"]), "--input-path", tin, "--output-path", tout,
             "--batch-size", 7)
    assert r.returncode == 0, r.stderr
    got = [l.split("\t") for l in open(tout).read().splitlines()]
    exp = [l.split("\t") for l in str(z["tsv_out"]).splitlines()]
    assert got[0] == exp[0] and len(got) == len(exp)
    for g, e in zip(got[1:], exp[1:]):
        for name, a, b in zip(exp[0], g, e):
            try:
                fb = float(b)
            except ValueError:
                assert a == b, name
                continue
            assert abs(float(a) - fb) <= 2e-6 * max(1.0, abs(fb)), (name, a, b)


def test_cli_reports_errors_like_the_reference(tiny):
    z, ff, d = tiny
    common = ["--yaml-path", ff["yaml_path"], "--cluster-ind", 0, "--hmm-param-dir", ff["hmm_dir"]]
    bad = os.path.join(d, "cli_bad.tree")
    open(bad, "w").write("((0:0.1,1:0.1):0.1,naive:0.1,nosuchtip:0.1);")
    r = _cli("--compute-logl", *common, "--newick-path", bad, *sum((["--er", 1] for _ in range(6)), []),
             *sum((["--pi", 0.25] for _ in range(4)), []))
    assert r.returncode == 1 and r.stderr.startswith("ERROR: ") and r.stdout == ""


def test_run_pipeline_parallel_sampler_equals_the_serial_walk(tiny):
    """RunPipeline samples the rows of a batch on several threads, each starting from the one generator's state
    advanced to its first row (std::mt19937::discard): the output must be what the serial, file-order walk of the
    reference gives (src/HMM.cpp:325-329, 1252-1257) - compared here with the same code on one host thread, which
    takes the serial path - and the object must end in the same state (next draws equal)."""
    import numpy as np
    from linearham_b200 import synth
    z, ff, d = tiny
    samples = synth.tree_samples(np.random.default_rng(77), ff["tree"], 300, outgroup="naive")
    tin = os.path.join(d, "in300.trees")
    synth.write_tsv(tin, samples)
    outs, draws = [], []
    for threads, batch in ((1, 128), (0, 128), (0, 1024)):
        tout = os.path.join(d, f"out300_{threads}_{batch}.trees")
        h = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
        h.set_host_threads(threads)
        h.set_device_sampler(False)
        h.RunPipeline(tin, tout, int(z["num_rates"]), batch_size=batch)
        draws.append([h.SampleNaiveSequence() for _ in range(3)])   # continues the generator where the run left it
        h.close()
        outs.append(open(tout).read())
    assert outs[0].count("\n") == 301
    assert outs[1] == outs[0] and outs[2] == outs[0]
    assert draws[1] == draws[0] and draws[2] == draws[0]


def _pipeline_runs(yaml_path, hmm_dir, tin, d, num_rates, configs):
    outs, draws, lls = [], [], []
    for k, (device_sampler, threads, batch) in enumerate(configs):
        tout = os.path.join(d, f"out_ds_{k}.trees")
        h = host.PhyloHMM(yaml_path, 0, hmm_dir, 0)
        h.set_host_threads(threads)
        h.set_device_sampler(device_sampler)
        h.RunPipeline(tin, tout, num_rates, batch_size=batch)
        lls.append(h.LogLikelihood())                               # the object holds the last row
        draws.append([h.SampleNaiveSequence() for _ in range(3)])   # and the generator stands where the run left it
        h.close()
        outs.append(open(tout).read())
    return outs, draws, lls


def test_run_pipeline_device_sampler_equals_the_host_sampler(tiny):
    """The states of every row's annotation drawn on the device (k_sample through lh_sample_paths: libstdc++'s
    discrete_distribution and generate_canonical restated operation by operation, fed with the raw outputs of the
    run's std::mt19937) against the host's SampleNaiveSequence: the same output file, the same object afterwards -
    last row's likelihood, and the next draws, which need both the generator's state and the last row's forward
    tables. Heavy chain (two junctions, D gene) and light chain (one junction)."""
    import numpy as np
    from linearham_b200 import synth
    z, ff, d = tiny
    samples = synth.tree_samples(np.random.default_rng(78), ff["tree"], 700, outgroup="naive")
    tin = os.path.join(d, "in700.trees")
    synth.write_tsv(tin, samples)
    configs = ((False, 1, 128), (True, 0, 256), (True, 1, 64), (True, 0, 1024), (True, 0, 1))
    outs, draws, lls = _pipeline_runs(ff["yaml_path"], ff["hmm_dir"], tin, d, int(z["num_rates"]), configs)
    assert outs[0].count("\n") == 701
    for k in range(1, len(configs)):
        assert outs[k] == outs[0], configs[k]
        assert draws[k] == draws[0] and lls[k] == lls[0], configs[k]
    # light chain
    dk = os.path.join(d, "igk")
    os.makedirs(dk, exist_ok=True)
    genes = synth.make_germline_set(os.path.join(dk, "hmmk"), seed=9, locus="igk")
    fam = synth.make_light_family(dk, genes, 30, seed=1305, name="fam_igk")
    samples = synth.tree_samples(np.random.default_rng(79), fam["tree"], 300, outgroup="naive")
    tin = os.path.join(dk, "in300.trees")
    synth.write_tsv(tin, samples)
    outs, draws, lls = _pipeline_runs(fam["yaml_path"], os.path.join(dk, "hmmk"), tin, dk, 4, ((False, 1, 128), (True, 0, 128)))
    assert outs[0].count("\n") == 301 and outs[1] == outs[0] and draws[1] == draws[0] and lls[1] == lls[0]
