"""The data files of the reference's Catch fixture for TEST_CASE("PhyloHMM") (test/test.cpp:752-754, 1056, 1372-1373)
-> tests/golden/fixture_catch_files.npz, so that the Catch binary built by integration/compat/build_catch.sh can run on a box
without /root/reference. Data only (partis-style YAML, one Newick line); run here:  python tests/golden/make_catch_data.py"""
import os
import numpy as np

REF = "/root/reference/data"
FILES = ["newton.tree", "phylo_hmm_input.yaml", "phylo_hmm_input_extra.yaml", "phylo_likelihood_hmm_input.yaml"] + \
    ["hmm_params/" + f for f in sorted(os.listdir(os.path.join(REF, "hmm_params")))] + \
    ["phylo_likelihood_hmm_params/" + f for f in sorted(os.listdir(os.path.join(REF, "phylo_likelihood_hmm_params")))]
out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "fixture_catch_files.npz")
np.savez_compressed(out, names=np.array(FILES), texts=np.array([open(os.path.join(REF, f)).read() for f in FILES]))
print(out, len(FILES), "files", os.path.getsize(out), "bytes")
