"""Standalone timing of the device Newick parser on a bench configuration (gpurun)."""
import sys, time, tempfile, json
sys.path.insert(0, ".")
import numpy as np
from linearham_b200 import device, host, workload
cfg_name, B = sys.argv[1], int(sys.argv[2])
cfg = dict(workload.CONFIGS[cfg_name])
ff = workload.write_family_files(cfg, tempfile.mkdtemp())
ph = host.PhyloHMM(ff["yaml_path"], 0, ff["hmm_dir"], 0)
fam, labels = ph.family(), ph.strs("xmsa_labels")
tb, newicks, alpha = workload.tree_batch(cfg, ff, B, n_topo=256)
text = host.NewickBuffer(tb, labels)
f = device.Family(fam)
f.set_labels(labels)
for kernel in (3, 2, 1):
    f.set_option("newick_kernel", kernel)
    for rep in range(3):
        t0 = time.perf_counter()
        nb_out = None
        B_, T = len(text), int(fam["n_tips"])
        import ctypes as C
        from linearham_b200 import _abi
        nb, keep = _abi.newick_batch(text.ptrs, np.zeros((B_, 6)), np.ones((B_, 4)), 1, rates=np.zeros((B_, 1)), lengths=text.lengths)
        out = {"children": np.zeros((B_, T - 2, 2), dtype=np.int32), "blen": np.zeros((B_, T - 2, 2)),
               "root_child": np.zeros(B_, dtype=np.int32), "root_blen": np.zeros(B_), "on": np.zeros(B_, dtype=np.uint8), "st": np.zeros(B_, dtype=np.uint8)}
        ms = C.c_float()
        i32p, u8p, f64p = C.POINTER(C.c_int32), C.POINTER(C.c_uint8), C.POINTER(C.c_double)
        rc = device.lib().lh_debug_parse_newick(f._h, C.byref(nb), out["children"].ctypes.data_as(i32p), out["blen"].ctypes.data_as(f64p),
            out["root_child"].ctypes.data_as(i32p), out["root_blen"].ctypes.data_as(f64p), out["on"].ctypes.data_as(u8p), out["st"].ctypes.data_as(u8p), C.byref(ms))
        wall = time.perf_counter() - t0
        ok = all(np.array_equal(out[k], tb[k]) for k in ("children", "blen", "root_child", "root_blen"))
        print(json.dumps({"cfg": cfg_name, "B": B, "kernel": kernel, "rc": rc, "ms_kernel_sum": ms.value, "wall_s": wall, "bytes": text.total_bytes(), "identical": ok, "left": int(out["st"].sum())}), flush=True)
