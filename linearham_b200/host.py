"""Thin Python handle over the host-side C ABI (include/linearham_b200_host.h): tree ingest and
Gamma categories. Marshalling only; the work is done by the C++ in linearham_b200/csrc/host."""
import ctypes as C

import numpy as np

from . import _abi

_lib = None


class HostError(ValueError):
    pass


def lib():
    global _lib
    if _lib is None:
        _lib = _abi.bind_host(_abi.load_library())
    return _lib


def _ck(rc):
    if rc:
        raise HostError(lib().lhh_last_error().decode())


def gamma_cats(alpha, n_rates):
    """pll_compute_gamma_cats(alpha, n, MEAN) (gamma.c:220-292). alpha scalar -> [n]; array -> [B, n]."""
    a = np.ascontiguousarray(np.atleast_1d(alpha), dtype=np.float64)
    out = np.zeros((a.shape[0], int(n_rates)))
    _ck(lib().lhh_gamma_cats_batch(a.shape[0], a.ctypes.data_as(_abi._f64p), int(n_rates),
                                   out.ctypes.data_as(_abi._f64p)))
    return out[0] if np.ndim(alpha) == 0 else out


def newick_to_ops(newicks, labels, n_threads=0):
    """Newick strings + xMSA row labels -> dict(children[B,T-2,2], blen[B,T-2,2], root_child[B], root_blen[B])."""
    T, B = len(labels), len(newicks)
    lab = (C.c_char_p * T)(*[str(s).encode() for s in labels])
    nwk = (C.c_char_p * B)(*[str(s).encode() for s in newicks])
    children = np.zeros((B, T - 2, 2), dtype=np.int32)
    blen = np.zeros((B, T - 2, 2))
    root_child = np.zeros(B, dtype=np.int32)
    root_blen = np.zeros(B)
    _ck(lib().lhh_newick_to_ops(T, lab, B, nwk, int(n_threads), children.ctypes.data_as(_abi._i32p),
                                blen.ctypes.data_as(_abi._f64p), root_child.ctypes.data_as(_abi._i32p),
                                root_blen.ctypes.data_as(_abi._f64p)))
    return {"children": children, "blen": blen, "root_child": root_child, "root_blen": root_blen}


def tree_batch(newicks, labels, er, pi, alpha, n_rates, n_threads=0):
    """Everything lh_score_batch needs for B samples given as the RevBayes TSV gives them
    (src/PhyloHMM.cpp:414-426): Newick, er[6], pi[4], alpha."""
    tb = newick_to_ops(newicks, labels, n_threads)
    B = len(newicks)
    tb["er"] = np.ascontiguousarray(er, dtype=np.float64).reshape(B, 6)
    tb["pi"] = np.ascontiguousarray(pi, dtype=np.float64).reshape(B, 4)
    tb["rates"] = gamma_cats(np.asarray(alpha, dtype=np.float64).reshape(B), n_rates)
    return tb
