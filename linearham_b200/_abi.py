"""ctypes mirror of include/linearham_b200.h (the device C ABI).

Only struct layouts, argument marshalling and the library loader live here; no
arithmetic. The tests hand the same structs to their CPU checker, so that the CUDA path and
the checker see byte-identical inputs.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "liblinearham_b200.so")

LH_MAX_RATES = 8

_i32p = C.POINTER(C.c_int32)
_f64p = C.POINTER(C.c_double)
_u8p = C.POINTER(C.c_uint8)


class FamilyDesc(C.Structure):
    _fields_ = [
        ("n_tips", C.c_int32), ("n_sites", C.c_int32), ("tip_codes", _u8p),
        ("has_d", C.c_int32), ("n_v", C.c_int32), ("n_d", C.c_int32), ("n_j", C.c_int32),
        ("vd_sites", C.c_int32), ("vd_states", C.c_int32),
        ("dj_sites", C.c_int32), ("dj_states", C.c_int32),
        ("vpad_off", _i32p), ("vpad_inds", _i32p),
        ("vgerm_off", _i32p), ("vgerm_inds", _i32p),
        ("dgerm_off", _i32p), ("dgerm_inds", _i32p),
        ("jgerm_off", _i32p), ("jgerm_inds", _i32p),
        ("jpad_off", _i32p), ("jpad_inds", _i32p),
        ("vd_inds", _i32p), ("dj_inds", _i32p),
        ("v_gene_prob", _f64p), ("v_trans_prod", _f64p),
        ("vpad_trans", _f64p), ("jpad_trans", _f64p),
        ("t_v_vd", _f64p), ("t_vd", _f64p), ("t_vd_d", _f64p),
        ("t_d_dj", _f64p), ("t_dj", _f64p), ("t_dj_j", _f64p),
    ]


class TreeBatch(C.Structure):
    _fields_ = [
        ("n_trees", C.c_int32), ("n_rates", C.c_int32),
        ("children", _i32p), ("blen", _f64p),
        ("root_child", _i32p), ("root_blen", _f64p),
        ("er", _f64p), ("pi", _f64p), ("rates", _f64p),
    ]


class AsrBatch(C.Structure):
    _fields_ = [("trees", C.POINTER(TreeBatch)), ("n_sites", C.c_int32), ("naive_row", C.c_int32),
                ("msa_codes", C.POINTER(C.c_uint8)), ("naive_codes", C.POINTER(C.c_uint8)), ("seed", C.c_uint64)]


class NewickBatch(C.Structure):
    _fields_ = [
        ("n_trees", C.c_int32), ("n_rates", C.c_int32),
        ("newick", C.POINTER(C.c_char_p)), ("length", C.POINTER(C.c_int64)),
        ("er", _f64p), ("pi", _f64p), ("rates", _f64p), ("alpha", _f64p),
    ]


class Outputs(C.Structure):
    _fields_ = [
        ("xmsa_emission", _f64p), ("vgerm_forward", _f64p), ("vd_forward", _f64p),
        ("dgerm_forward", _f64p), ("dj_forward", _f64p), ("jgerm_forward", _f64p),
        ("germ_scalers", _i32p), ("vd_scalers", _i32p), ("dj_scalers", _i32p),
        ("site_rate_loglik", _f64p),
    ]


class Timings(C.Structure):
    _fields_ = [
        ("ms_h2d", C.c_float), ("ms_plan", C.c_float), ("ms_pmatrix", C.c_float),
        ("ms_prune", C.c_float), ("ms_edge", C.c_float), ("ms_hmm", C.c_float), ("ms_d2h", C.c_float),
        ("ms_total", C.c_float), ("ms_kernels", C.c_float),
        ("n_trees", C.c_int32), ("kernel_launches", C.c_int32), ("prune_launches", C.c_int32),
        ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64),
        ("prune_slots", C.c_int32), ("prune_smem_slots", C.c_int32), ("prune_smem_lo", C.c_int32), ("prune_tile_sites", C.c_int32),
        ("prune_sites_per_thread", C.c_int32), ("prune_grid", C.c_int32), ("prune_ctas_per_sm", C.c_int32),
        ("prune_columns", C.c_int32), ("pattern_trees", C.c_int32),
        ("ms_call", C.c_float), ("ms_setup", C.c_float),
        ("ms_plan_dev", C.c_float), ("device_plan", C.c_int32),
        ("ms_newick", C.c_float), ("newick_host_trees", C.c_int32), ("text_direct", C.c_int32),
    ]


_FAMILY_ARRAYS = {  # field -> dtype
    "tip_codes": np.uint8,
    "vpad_off": np.int32, "vpad_inds": np.int32, "vgerm_off": np.int32, "vgerm_inds": np.int32,
    "dgerm_off": np.int32, "dgerm_inds": np.int32, "jgerm_off": np.int32, "jgerm_inds": np.int32,
    "jpad_off": np.int32, "jpad_inds": np.int32, "vd_inds": np.int32, "dj_inds": np.int32,
    "v_gene_prob": np.float64, "v_trans_prod": np.float64, "vpad_trans": np.float64,
    "jpad_trans": np.float64, "t_v_vd": np.float64, "t_vd": np.float64, "t_vd_d": np.float64,
    "t_d_dj": np.float64, "t_dj": np.float64, "t_dj_j": np.float64,
}
_FAMILY_SCALARS = ["n_tips", "n_sites", "has_d", "n_v", "n_d", "n_j",
                   "vd_sites", "vd_states", "dj_sites", "dj_states"]
_TREE_ARRAYS = {"children": np.int32, "blen": np.float64, "root_child": np.int32,
                "root_blen": np.float64, "er": np.float64, "pi": np.float64, "rates": np.float64}


def _ptr(a, ctype):
    return a.ctypes.data_as(C.POINTER(ctype))


_CT = {np.uint8: C.c_uint8, np.int32: C.c_int32, np.float64: C.c_double}


def family_desc(fam):
    """dict of numpy arrays / ints (see FamilyDesc) -> (FamilyDesc, keepalive list)."""
    d = FamilyDesc()
    keep = []
    for k in _FAMILY_SCALARS:
        setattr(d, k, int(fam[k]))
    for k, dt in _FAMILY_ARRAYS.items():
        a = np.ascontiguousarray(np.asarray(fam[k]).reshape(-1), dtype=dt)
        if a.size == 0:
            a = np.zeros(1, dtype=dt)
        keep.append(a)
        setattr(d, k, _ptr(a, _CT[dt]))
    return d, keep


def tree_batch(tb):
    """dict with children[B,T-2,2], blen[B,T-2,2], root_child[B], root_blen[B],
    er[B,6], pi[B,4], rates[B,R] -> (TreeBatch, keepalive)."""
    t = TreeBatch()
    keep = []
    rates = np.asarray(tb["rates"], dtype=np.float64)
    t.n_trees = int(np.asarray(tb["root_child"]).shape[0])
    t.n_rates = int(rates.shape[1])
    for k, dt in _TREE_ARRAYS.items():
        a = np.ascontiguousarray(np.asarray(tb[k]).reshape(-1), dtype=dt)
        keep.append(a)
        setattr(t, k, _ptr(a, _CT[dt]))
    return t, keep


def alloc_outputs(fam, B, fields=None, n_rates=None):
    """Allocate numpy buffers for lh_outputs_t; returns (Outputs, dict of arrays)."""
    shapes = {
        "xmsa_emission": ((B, fam["n_sites"]), np.float64),
        "vgerm_forward": ((B, fam["n_v"]), np.float64),
        "vd_forward": ((B, fam["vd_sites"] * fam["vd_states"]), np.float64),
        "dgerm_forward": ((B, max(fam["n_d"], 1)), np.float64),
        "dj_forward": ((B, max(fam["dj_sites"] * fam["dj_states"], 1)), np.float64),
        "jgerm_forward": ((B, fam["n_j"]), np.float64),
        "germ_scalers": ((B, 3), np.int32),
        "vd_scalers": ((B, fam["vd_sites"]), np.int32),
        "dj_scalers": ((B, max(fam["dj_sites"], 1)), np.int32),
    }
    if fields is not None and "site_rate_loglik" in fields:   # only on request: B x R x S_x doubles
        shapes["site_rate_loglik"] = ((B, int(n_rates), fam["n_sites"]), np.float64)
    o = Outputs()
    arrs = {}
    for k, (shape, dt) in shapes.items():
        if fields is not None and k not in fields:
            continue
        a = np.zeros(shape, dtype=dt)
        arrs[k] = a
        setattr(o, k, _ptr(a, _CT[dt]))
    return o, arrs


def newick_batch(newicks, er, pi, n_rates, alpha=None, rates=None, lengths=False):
    """Newick strings (str or bytes) + er[B,6], pi[B,4] + alpha[B] or rates[B,R] -> (NewickBatch, keepalive)."""
    nb = NewickBatch()
    if isinstance(newicks, C.Array):   # a (c_char_p * B) array marshalled once by the caller (host.NewickBuffer.ptrs)
        arr, raw = newicks, None
        B = int(np.asarray(er).reshape(-1, 6).shape[0])
    else:
        raw = [s if isinstance(s, bytes) else str(s).encode() for s in newicks]
        B = len(raw)
        arr = (C.c_char_p * max(B, 1))(*raw)
    keep = [raw, arr]
    nb.n_trees, nb.n_rates = B, int(n_rates)
    nb.newick = C.cast(arr, C.POINTER(C.c_char_p))
    if lengths is not None and lengths is not False:
        ln = (np.array([len(s) for s in raw], dtype=np.int64) if lengths is True
              else np.ascontiguousarray(lengths, dtype=np.int64))
        keep.append(ln)
        nb.length = ln.ctypes.data_as(C.POINTER(C.c_int64))
    for name, val in (("er", er), ("pi", pi), ("alpha", alpha), ("rates", rates)):
        if val is None:
            continue
        a = np.ascontiguousarray(np.asarray(val, dtype=np.float64).reshape(-1))
        keep.append(a)
        setattr(nb, name, _ptr(a, C.c_double))
    return nb, keep


DEVICE_SYMBOLS = [
    "lh_last_error", "lh_version", "lh_device_count", "lh_family_create", "lh_family_destroy",
    "lh_score_batch", "lh_batch_upload", "lh_batch_run", "lh_batch_fetch", "lh_sync",
    "lh_loglik_dev", "lh_get_timings", "lh_set_option", "lh_measure_fp64_peak",
    "lh_timer_mark", "lh_timer_elapsed_ms", "lh_family_rebind", "lh_debug_fetch_plan",
    "lh_family_set_labels", "lh_score_newick", "lh_debug_parse_newick", "lh_newick_parse_noalloc",
    "lh_newick_parse_noalloc_max_trees", "lh_asr_sample", "lh_pin_host", "lh_unpin_host", "lh_sample_paths",
    "lh_group_create", "lh_group_destroy", "lh_group_size", "lh_group_member", "lh_group_set_labels", "lh_group_set_option",
    "lh_group_score_batch", "lh_group_score_newick", "lh_group_get_timings", "lh_group_last_error", "lh_parse_devices",
    "lh_repertoire_create_multi", "lh_repertoire_create", "lh_repertoire_score", "lh_repertoire_destroy", "lh_repertoire_set_option", "lh_repertoire_last_error",
]


class RepertoireItem(C.Structure):
    _fields_ = [("desc", C.POINTER(FamilyDesc)), ("trees", C.POINTER(TreeBatch)), ("loglik", _f64p),
                ("status", C.c_int32)]


class RepertoireStats(C.Structure):
    _fields_ = [("ms_wall", C.c_float), ("ms_bind", C.c_float), ("ms_score", C.c_float),
                ("n_families", C.c_int32), ("n_failed", C.c_int32), ("n_slots", C.c_int32),
                ("host_threads_per_slot", C.c_int32), ("n_trees", C.c_int64),
                ("kernel_launches", C.c_int64), ("h2d_bytes", C.c_int64), ("d2h_bytes", C.c_int64)]


def load_library(path=None):
    """Load the CUDA library. Fails loudly: there is no fallback implementation."""
    path = path or LIB_PATH
    if not os.path.exists(path):
        raise RuntimeError(
            f"{path} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). linearham_b200 has no CPU fallback.")
    lib = C.CDLL(path)
    lib.lh_last_error.restype = C.c_char_p
    lib.lh_version.restype = C.c_char_p
    lib.lh_device_count.argtypes = [C.POINTER(C.c_int)]
    lib.lh_family_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(FamilyDesc), C.c_int]
    lib.lh_family_destroy.argtypes = [C.c_void_p]
    lib.lh_family_destroy.restype = None
    lib.lh_score_batch.argtypes = [C.c_void_p, C.POINTER(TreeBatch), _f64p, C.POINTER(Outputs)]
    lib.lh_batch_upload.argtypes = [C.c_void_p, C.POINTER(TreeBatch)]
    lib.lh_batch_run.argtypes = [C.c_void_p]
    lib.lh_batch_fetch.argtypes = [C.c_void_p, _f64p, C.POINTER(Outputs)]
    lib.lh_sync.argtypes = [C.c_void_p]
    lib.lh_loglik_dev.argtypes = [C.c_void_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int32)]
    lib.lh_get_timings.argtypes = [C.c_void_p, C.POINTER(Timings)]
    lib.lh_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
    lib.lh_measure_fp64_peak.argtypes = [C.c_int, _f64p]
    lib.lh_timer_mark.argtypes = [C.c_void_p, C.c_int32]
    lib.lh_timer_elapsed_ms.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.POINTER(C.c_float)]
    lib.lh_debug_fetch_plan.argtypes = [C.c_void_p, _i32p, _f64p]
    lib.lh_family_rebind.argtypes = [C.c_void_p, C.POINTER(FamilyDesc)]
    lib.lh_family_set_labels.argtypes = [C.c_void_p, C.POINTER(C.c_char_p)]
    lib.lh_sample_paths.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_uint32), C.c_int32, C.POINTER(C.c_int32), C.c_int32,
                                    C.POINTER(C.c_uint8)]
    lib.lh_pin_host.argtypes = [C.c_void_p, C.c_size_t]
    lib.lh_unpin_host.argtypes = [C.c_void_p]
    lib.lh_asr_sample.argtypes = [C.c_void_p, C.POINTER(AsrBatch), C.POINTER(C.c_uint8), C.POINTER(C.c_uint8), _f64p]
    lib.lh_newick_parse_noalloc.argtypes = [C.c_void_p, C.c_int32] + [C.c_void_p] * 10
    lib.lh_newick_parse_noalloc_max_trees.argtypes = [C.c_int32]
    lib.lh_newick_parse_noalloc_max_trees.restype = C.c_int32
    lib.lh_score_newick.argtypes = [C.c_void_p, C.POINTER(NewickBatch), _f64p, C.POINTER(Outputs)]
    lib.lh_debug_parse_newick.argtypes = [C.c_void_p, C.POINTER(NewickBatch), _i32p, _f64p, _i32p, _f64p, _u8p, _u8p,
                                          C.POINTER(C.c_float)]
    lib.lh_group_create.argtypes = [C.POINTER(C.c_void_p), C.POINTER(FamilyDesc), C.POINTER(C.c_int), C.c_int]
    lib.lh_group_destroy.argtypes = [C.c_void_p]
    lib.lh_group_destroy.restype = None
    lib.lh_group_size.argtypes = [C.c_void_p]
    lib.lh_group_member.argtypes = [C.c_void_p, C.c_int]
    lib.lh_group_member.restype = C.c_void_p
    lib.lh_group_set_labels.argtypes = [C.c_void_p, C.POINTER(C.c_char_p)]
    lib.lh_group_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
    lib.lh_group_score_batch.argtypes = [C.c_void_p, C.POINTER(TreeBatch), _f64p, C.POINTER(Outputs)]
    lib.lh_group_score_newick.argtypes = [C.c_void_p, C.POINTER(NewickBatch), _f64p, C.POINTER(Outputs)]
    lib.lh_group_get_timings.argtypes = [C.c_void_p, C.c_int, C.POINTER(Timings)]
    lib.lh_group_last_error.restype = C.c_char_p
    lib.lh_parse_devices.argtypes = [C.c_char_p, C.POINTER(C.c_int), C.c_int, C.POINTER(C.c_int)]
    lib.lh_repertoire_create_multi.argtypes = [C.POINTER(C.c_void_p), C.POINTER(C.c_int), C.c_int, C.c_int, C.c_int]
    lib.lh_repertoire_create.argtypes = [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int]
    lib.lh_repertoire_score.argtypes = [C.c_void_p, C.POINTER(RepertoireItem), C.c_int32,
                                        C.POINTER(RepertoireStats)]
    lib.lh_repertoire_set_option.argtypes = [C.c_void_p, C.c_char_p, C.c_int64]
    lib.lh_repertoire_destroy.argtypes = [C.c_void_p]
    lib.lh_repertoire_destroy.restype = None
    lib.lh_repertoire_last_error.restype = C.c_char_p
    return lib


# ---- host-side C ABI (include/linearham_b200_host.h) ------------------------------------

HOST_SYMBOLS = ["lhh_last_error", "lhh_plan_trees", "lhh_gamma_cats", "lhh_gamma_cats_batch",
                "lhh_newick_to_ops", "lhh_parse_decimal_buf", "lhh_ops_to_newick"]


def bind_host(lib):
    lib.lhh_last_error.restype = C.c_char_p
    lib.lhh_gamma_cats.argtypes = [C.c_double, C.c_int32, _f64p]
    lib.lhh_gamma_cats_batch.argtypes = [C.c_int32, _f64p, C.c_int32, _f64p]
    lib.lhh_newick_to_ops.argtypes = [C.c_int32, C.POINTER(C.c_char_p), C.c_int32, C.POINTER(C.c_char_p),
                                      C.c_int32, _i32p, _f64p, _i32p, _f64p]
    lib.lhh_plan_trees.argtypes = [C.c_int32, C.c_int32, _i32p, _f64p, _i32p, _f64p, _i32p, _f64p, _i32p]
    lib.lhh_parse_decimal_buf.argtypes = [C.c_char_p, C.c_int32, _f64p, _i32p, _i32p]
    lib.lhh_ops_to_newick.argtypes = [C.c_int32, C.POINTER(C.c_char_p), C.c_int32, _i32p, _f64p, _i32p, _f64p, C.c_int32,
                                      C.c_void_p, C.c_int64]
    return lib


HOST_SYMBOLS += ["lhh_phylohmm_new", "lhh_phylohmm_free", "lhh_phylohmm_set_device", "lhh_phylohmm_set_devices", "lhh_phylohmm_set_host_threads", "lhh_phylohmm_set_device_sampler", "lhh_phylohmm_family_desc",
                 "lhh_phylohmm_init_params", "lhh_phylohmm_init_emission", "lhh_phylohmm_loglikelihood",
                 "lhh_phylohmm_sample_naive", "lhh_phylohmm_run_pipeline", "lhh_phylohmm_score_samples", "lhh_phylohmm_score_samples_len", "lhh_phylohmm_device_family",
                 "lhh_export_newick", "lhh_phylohmm_get_f64", "lhh_phylohmm_get_i32", "lhh_phylohmm_get_str"]


def bind_phylohmm(lib):
    lib.lhh_phylohmm_new.restype = C.c_void_p
    lib.lhh_phylohmm_new.argtypes = [C.c_char_p, C.c_int32, C.c_char_p, C.c_int32]
    lib.lhh_phylohmm_free.argtypes = [C.c_void_p]
    lib.lhh_phylohmm_free.restype = None
    lib.lhh_phylohmm_set_device.argtypes = [C.c_void_p, C.c_int32]
    lib.lhh_phylohmm_set_devices.argtypes = [C.c_void_p, _i32p, C.c_int32]
    lib.lhh_phylohmm_device_family.argtypes = [C.c_void_p, C.POINTER(C.c_void_p)]
    lib.lhh_phylohmm_set_host_threads.argtypes = [C.c_void_p, C.c_int32]
    lib.lhh_phylohmm_set_device_sampler.argtypes = [C.c_void_p, C.c_int32]
    lib.lhh_phylohmm_family_desc.argtypes = [C.c_void_p, C.POINTER(FamilyDesc)]
    lib.lhh_phylohmm_init_params.argtypes = [C.c_void_p, C.c_char_p, _f64p, _f64p, C.c_double, C.c_int32]
    lib.lhh_phylohmm_init_emission.argtypes = [C.c_void_p]
    lib.lhh_phylohmm_loglikelihood.argtypes = [C.c_void_p, _f64p]
    lib.lhh_phylohmm_sample_naive.argtypes = [C.c_void_p, C.c_char_p, C.c_int32]
    lib.lhh_phylohmm_run_pipeline.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32, C.c_int32]
    lib.lhh_phylohmm_score_samples.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), _f64p, _f64p, _f64p,
                                               C.c_int32, _f64p]
    lib.lhh_phylohmm_score_samples_len.argtypes = [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), C.POINTER(C.c_int64), _f64p,
                                                   _f64p, _f64p, C.c_int32, _f64p]
    lib.lhh_export_newick.argtypes = [C.c_char_p, C.c_char_p, C.c_int32]
    lib.lhh_phylohmm_get_f64.argtypes = [C.c_void_p, C.c_char_p, _f64p, _i32p, _i32p]
    lib.lhh_phylohmm_get_i32.argtypes = [C.c_void_p, C.c_char_p, _i32p, _i32p, _i32p]
    lib.lhh_phylohmm_get_str.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int32]
    return lib
