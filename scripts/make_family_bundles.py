"""Build-container helper: runs the REFERENCE's host code (oracle/_ref) on the synthetic family files
of the bench configs and stores the flat family description under tests/golden/family_<cfg>.npz.
They pin the product's host-side family construction (bit-exact xMSA / index maps / transition
matrices) and let bench.py run on the GPU box, where /root/reference does not exist.

    python scripts/make_family_bundles.py
"""
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from linearham_b200 import workload  # noqa: E402
from oracle import pyoracle as po  # noqa: E402

for name in ("tiny", "c3", "c4"):
    cfg = workload.CONFIGS[name]
    with tempfile.TemporaryDirectory() as d:
        ff = workload.write_family_files(cfg, d)
        ref = po.RefPhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
        fam = ref.family()
        data = {"fam_" + k: np.asarray(v) for k, v in fam.items()}
        data["labels"] = np.array(ref.strs("xmsa_labels"))
        data["xmsa_seqs"] = np.array(ref.strs("xmsa_seqs"))
        data["sha"] = np.array(ff["sha"])
        path = os.path.join(ROOT, "tests", "golden", f"family_{name}.npz")
        np.savez_compressed(path, **data)
        print(name, "T", fam["n_tips"], "S_x", fam["n_sites"], "vd", fam["vd_sites"], fam["vd_states"],
              "dj", fam["dj_sites"], fam["dj_states"], os.path.getsize(path), "bytes")
