"""TEST INFRASTRUCTURE ONLY. ctypes loaders for the two checkers:

  * RefPhyloHMM  - the reference's own C++ (oracle/_ref/libref_linearham.so, built from
                   /root/reference by oracle/refbuild/Makefile; prebuilt file travels to the GPU box)
  * port_*       - the scalar CPU restatement (oracle/port/lh_port.cpp), buildable anywhere with g++

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs import this.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
REF_LIB = os.path.join(_HERE, "_ref", "libref_linearham.so")
REF_LIB_SHIPPED = os.path.join(_HERE, "_ref", "libref_linearham_shipped.so")
PORT_SRC = os.path.join(_HERE, "port", "lh_port.cpp")
PORT_LIB = os.path.join(_HERE, "port", "libport.so")
REFERENCE_ROOT = "/root/reference"

_f64p = C.POINTER(C.c_double)
_i32p = C.POINTER(C.c_int32)


def build_port(force=False):
    if force or not os.path.exists(PORT_LIB) or os.path.getmtime(PORT_LIB) < os.path.getmtime(PORT_SRC):
        subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-fPIC", "-shared",
                               "-o", PORT_LIB, PORT_SRC])
    return PORT_LIB


def build_ref(flavor="tuned"):
    """Compile the reference in place (only possible where /root/reference exists)."""
    if not os.path.isdir(REFERENCE_ROOT):
        return None
    subprocess.check_call(["make", "-C", os.path.join(_HERE, "refbuild"), "-j8", f"FLAVOR={flavor}", "all"],
                          stdout=subprocess.DEVNULL)
    return REF_LIB if flavor == "tuned" else REF_LIB_SHIPPED


def have_ref():
    return os.path.exists(REF_LIB)


_port = None


def port_lib():
    global _port
    if _port is None:
        from linearham_b200 import _abi
        lib = C.CDLL(build_port())
        lib.port_gamma_cats.argtypes = [C.c_double, C.c_int, _f64p]
        lib.port_score_batch.argtypes = [C.POINTER(_abi.FamilyDesc), C.POINTER(_abi.TreeBatch), _f64p,
                                         C.POINTER(_abi.Outputs)]
        _port = lib
    return _port


def port_gamma_cats(alpha, n):
    out = np.zeros(n)
    rc = port_lib().port_gamma_cats(float(alpha), int(n), out.ctypes.data_as(_f64p))
    if rc:
        raise ValueError(f"invalid alpha {alpha}")
    return out


def port_score_batch(fam, tb, want_outputs=True, fields=None):
    """fam / tb: dicts as in linearham_b200._abi. Returns (loglik[B], outputs dict)."""
    from linearham_b200 import _abi
    d, k1 = _abi.family_desc(fam)
    t, k2 = _abi.tree_batch(tb)
    B = t.n_trees
    ll = np.zeros(B)
    if want_outputs:
        o, arrs = _abi.alloc_outputs(fam, B, fields, n_rates=t.n_rates)
        port_lib().port_score_batch(C.byref(d), C.byref(t), ll.ctypes.data_as(_f64p), C.byref(o))
    else:
        arrs = {}
        port_lib().port_score_batch(C.byref(d), C.byref(t), ll.ctypes.data_as(_f64p), None)
    return ll, arrs


class RefPhyloHMM:
    """The reference linearham::PhyloHMM, driven through oracle/refbuild/ref_capi.cpp."""

    def __init__(self, yaml_path, cluster_ind, hmm_param_dir, seed=0, lib_path=None):
        self.lib = C.CDLL(lib_path or REF_LIB)
        L = self.lib
        L.ref_last_error.restype = C.c_char_p
        L.ref_phylohmm_new.restype = C.c_void_p
        L.ref_phylohmm_new.argtypes = [C.c_char_p, C.c_int, C.c_char_p, C.c_int]
        L.ref_phylohmm_free.argtypes = [C.c_void_p]
        L.ref_score_newick.argtypes = [C.c_void_p, C.c_char_p, _f64p, _f64p, C.c_double, C.c_int, _f64p]
        L.ref_score_batch_timed.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_char_p), _f64p, _f64p, _f64p,
                                            C.c_int, _f64p, _f64p, _f64p, _f64p]
        L.ref_init_params_file.argtypes = [C.c_void_p, C.c_char_p, _f64p, _f64p, C.c_double, C.c_int]
        L.ref_loglikelihood.argtypes = [C.c_void_p, _f64p]
        L.ref_sample_naive.argtypes = [C.c_void_p, C.c_char_p, C.c_int]
        L.ref_run_pipeline.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int]
        L.ref_gamma_cats.argtypes = [C.c_double, C.c_int, _f64p]
        L.ref_get_f64.argtypes = [C.c_void_p, C.c_char_p, _f64p, _i32p, _i32p]
        L.ref_get_i32.argtypes = [C.c_void_p, C.c_char_p, _i32p, _i32p, _i32p]
        L.ref_get_str.argtypes = [C.c_void_p, C.c_char_p, C.c_char_p, C.c_int]
        self.h = L.ref_phylohmm_new(str(yaml_path).encode(), int(cluster_ind), str(hmm_param_dir).encode(), int(seed))
        if not self.h:
            raise RuntimeError("reference PhyloHMM ctor failed: " + L.ref_last_error().decode())

    def close(self):
        if self.h:
            self.lib.ref_phylohmm_free(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _check(self, rc):
        if rc:
            raise RuntimeError(self.lib.ref_last_error().decode())

    def f64(self, name):
        r, c = C.c_int32(), C.c_int32()
        self._check(self.lib.ref_get_f64(self.h, name.encode(), None, C.byref(r), C.byref(c)))
        out = np.zeros((c.value, r.value))  # column-major payload
        self._check(self.lib.ref_get_f64(self.h, name.encode(), out.ctypes.data_as(_f64p), C.byref(r), C.byref(c)))
        return out.T.copy()  # (rows, cols)

    def i32(self, name):
        r, c = C.c_int32(), C.c_int32()
        self._check(self.lib.ref_get_i32(self.h, name.encode(), None, C.byref(r), C.byref(c)))
        out = np.zeros((c.value, r.value), dtype=np.int32)
        self._check(self.lib.ref_get_i32(self.h, name.encode(), out.ctypes.data_as(_i32p), C.byref(r), C.byref(c)))
        return out.T.copy()

    def strs(self, name):
        n = self.lib.ref_get_str(self.h, name.encode(), None, 0)
        if n < 0:
            raise KeyError(name)
        buf = C.create_string_buffer(n)
        self.lib.ref_get_str(self.h, name.encode(), buf, n)
        s = buf.value.decode()
        return s.split("\n") if s else []

    def score_newick(self, newick, er, pi, alpha, num_rates):
        er = np.ascontiguousarray(er, dtype=np.float64)
        pi = np.ascontiguousarray(pi, dtype=np.float64)
        out = C.c_double()
        self._check(self.lib.ref_score_newick(self.h, newick.encode(), er.ctypes.data_as(_f64p),
                                              pi.ctypes.data_as(_f64p), float(alpha), int(num_rates), C.byref(out)))
        return out.value

    def score_batch_timed(self, newicks, er, pi, alpha, num_rates):
        n = len(newicks)
        arr = (C.c_char_p * n)(*[s.encode() for s in newicks])
        er = np.ascontiguousarray(er, dtype=np.float64)
        pi = np.ascontiguousarray(pi, dtype=np.float64)
        alpha = np.ascontiguousarray(alpha, dtype=np.float64)
        ll = np.zeros(n)
        tp, te, tf = C.c_double(), C.c_double(), C.c_double()
        self._check(self.lib.ref_score_batch_timed(self.h, n, arr, er.ctypes.data_as(_f64p), pi.ctypes.data_as(_f64p),
                                                   alpha.ctypes.data_as(_f64p), int(num_rates),
                                                   ll.ctypes.data_as(_f64p), C.byref(tp), C.byref(te), C.byref(tf)))
        return ll, {"parse_s": tp.value, "emission_s": te.value, "forward_s": tf.value}

    def init_params_file(self, newick_path, er, pi, alpha, num_rates):
        er = np.ascontiguousarray(er, dtype=np.float64)
        pi = np.ascontiguousarray(pi, dtype=np.float64)
        self._check(self.lib.ref_init_params_file(self.h, str(newick_path).encode(), er.ctypes.data_as(_f64p),
                                                  pi.ctypes.data_as(_f64p), float(alpha), int(num_rates)))

    def loglikelihood(self):
        out = C.c_double()
        self._check(self.lib.ref_loglikelihood(self.h, C.byref(out)))
        return out.value

    def sample_naive(self):
        buf = C.create_string_buffer(1 << 16)
        self._check(self.lib.ref_sample_naive(self.h, buf, len(buf)))
        return buf.value.decode()

    def run_pipeline(self, in_path, out_path, num_rates):
        self._check(self.lib.ref_run_pipeline(self.h, str(in_path).encode(), str(out_path).encode(), int(num_rates)))

    def gamma_cats(self, alpha, n):
        out = np.zeros(n)
        if self.lib.ref_gamma_cats(float(alpha), int(n), out.ctypes.data_as(_f64p)):
            raise ValueError(f"invalid alpha {alpha}")
        return out

    # ---- flat family description in the device-ABI layout -------------------
    def family(self):
        locus = self.strs("locus")[0]
        has_d = 1 if locus == "igh" else 0
        xmsa = self.i32("xmsa")  # (T, S_x)
        fam = {"n_tips": xmsa.shape[0], "n_sites": xmsa.shape[1], "tip_codes": xmsa.astype(np.uint8),
               "has_d": has_d}

        def csr(region):
            rng = self.i32(region + "_ggene_ranges")  # (2, n): start,end per gene in map order
            inds = self.i32(region + "_xmsa_inds").reshape(-1)
            n = rng.shape[1]
            off = np.zeros(n + 1, dtype=np.int32)
            for i in range(n):
                assert rng[0, i] == (rng[1, i - 1] if i else rng[0, 0])
                off[i], off[i + 1] = rng[0, i], rng[1, i]
            off -= off[0]
            return n, off, inds.astype(np.int32)

        fam["n_v"], fam["vpad_off"], fam["vpad_inds"] = csr("vpadding")
        _, fam["vgerm_off"], fam["vgerm_inds"] = csr("vgerm")
        if has_d:
            fam["n_d"], fam["dgerm_off"], fam["dgerm_inds"] = csr("dgerm")
        else:
            fam["n_d"], fam["dgerm_off"], fam["dgerm_inds"] = 0, np.zeros(1, np.int32), np.zeros(0, np.int32)
        fam["n_j"], fam["jgerm_off"], fam["jgerm_inds"] = csr("jgerm")
        _, fam["jpad_off"], fam["jpad_inds"] = csr("jpadding")
        vd = self.i32("vd_junction_xmsa_inds")  # (L, S)
        fam["vd_sites"], fam["vd_states"] = vd.shape
        fam["vd_inds"] = np.asfortranarray(vd).reshape(-1, order="F")
        if has_d:
            dj = self.i32("dj_junction_xmsa_inds")
            fam["dj_sites"], fam["dj_states"] = dj.shape
            fam["dj_inds"] = np.asfortranarray(dj).reshape(-1, order="F")
        else:
            fam["dj_sites"] = fam["dj_states"] = 0
            fam["dj_inds"] = np.zeros(0, np.int32)
        fam["v_gene_prob"] = self.f64("vgerm_gene_prob").reshape(-1)
        fam["v_trans_prod"] = self.f64("vgerm_trans_prod").reshape(-1)
        fam["vpad_trans"] = self.f64("vpadding_transition").reshape(-1)
        fam["jpad_trans"] = self.f64("jpadding_transition").reshape(-1)

        def colmajor(name):
            return np.asfortranarray(self.f64(name)).reshape(-1, order="F")

        fam["t_v_vd"] = colmajor("vgerm_vd_junction_transition")
        fam["t_vd"] = colmajor("vd_junction_transition")
        fam["t_vd_d"] = colmajor("vd_junction_dgerm_transition")
        if has_d:
            fam["t_d_dj"] = colmajor("dgerm_dj_junction_transition")
            fam["t_dj"] = colmajor("dj_junction_transition")
            fam["t_dj_j"] = colmajor("dj_junction_jgerm_transition")
        else:
            fam["t_d_dj"] = fam["t_dj"] = fam["t_dj_j"] = np.zeros(0)
        return fam
