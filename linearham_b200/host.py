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


def encode_newicks(newicks):
    """list of str -> (c_char_p * n) array for the C ABI (keeps the encoded bytes alive)"""
    return (C.c_char_p * len(newicks))(*[s if isinstance(s, bytes) else str(s).encode() for s in newicks])


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


class NewickBuffer:
    """The trees of a flat batch as Newick text in one buffer (lhh_ops_to_newick): .ptrs is the
    (c_char_p * B) array the text entry points take, .lengths their strlen, [i] the i-th string."""

    def __init__(self, tb, labels, n_threads=0, pinned=False):
        """pinned: page-lock the buffer (lh_pin_host) - the text calls then copy it to the device where it lies."""
        T, B = len(labels), int(np.asarray(tb["root_child"]).shape[0])
        lab = (C.c_char_p * T)(*[str(s).encode() for s in labels])
        stride = (max(len(str(s)) for s in labels) + 48) * 2 * T + 128
        stride = (stride + 63) // 64 * 64
        self.buf = np.zeros(B * stride, dtype=np.uint8)
        arrs = [np.ascontiguousarray(tb["children"], dtype=np.int32), np.ascontiguousarray(tb["blen"], dtype=np.float64),
                np.ascontiguousarray(tb["root_child"], dtype=np.int32), np.ascontiguousarray(tb["root_blen"], dtype=np.float64)]
        _ck(lib().lhh_ops_to_newick(T, lab, B, arrs[0].ctypes.data_as(_abi._i32p), arrs[1].ctypes.data_as(_abi._f64p),
                                    arrs[2].ctypes.data_as(_abi._i32p), arrs[3].ctypes.data_as(_abi._f64p), int(n_threads),
                                    self.buf.ctypes.data, stride))
        base = self.buf.ctypes.data
        self.n, self.stride = B, stride
        self.ptrs = (C.c_char_p * max(B, 1))()
        addr = C.cast(self.ptrs, C.POINTER(C.c_void_p))
        for i in range(B):
            addr[i] = base + i * stride
        self.lengths = (self.buf.reshape(B, stride) == 0).argmax(axis=1).astype(np.int64) if B else np.zeros(0, dtype=np.int64)
        self.pinned = False
        if pinned and B:
            from . import device
            if device.lib().lh_pin_host(self.buf.ctypes.data, self.buf.nbytes):
                raise HostError(device.lib().lh_last_error().decode())
            self.pinned = True

    def __del__(self):
        if getattr(self, "pinned", False):
            self.pinned = False
            try:
                from . import device
                device.lib().lh_unpin_host(self.buf.ctypes.data)
            except Exception:   # (interpreter shutdown: the library may be gone; the process's pages go with it)
                pass

    def __len__(self):
        return self.n

    def __getitem__(self, i):
        return self.ptrs[i].decode()

    def total_bytes(self):
        return int(self.lengths.sum())


def tree_batch(newicks, labels, er, pi, alpha, n_rates, n_threads=0):
    """Everything lh_score_batch needs for B samples given as the RevBayes TSV gives them
    (src/PhyloHMM.cpp:414-426): Newick, er[6], pi[4], alpha."""
    tb = newick_to_ops(newicks, labels, n_threads)
    B = len(newicks)
    tb["er"] = np.ascontiguousarray(er, dtype=np.float64).reshape(B, 6)
    tb["pi"] = np.ascontiguousarray(pi, dtype=np.float64).reshape(B, 4)
    tb["rates"] = gamma_cats(np.asarray(alpha, dtype=np.float64).reshape(B), n_rates)
    return tb


def export_newick(newick):
    """pll_utree_export_newick of a Newick string (the `tree` column of the output TSV)."""
    _abi.bind_phylohmm(lib())
    buf = C.create_string_buffer(len(newick) * 3 + 64)
    _ck(lib().lhh_export_newick(newick.encode(), buf, len(buf)))
    return buf.value.decode()


class PhyloHMM:
    """Handle over the C++ linearham::PhyloHMM of this repo (csrc/host/phylo_hmm.hpp), which mirrors the
    reference class (src/PhyloHMM.hpp:24-127). Construction is host-only; scoring needs the GPU."""

    def __init__(self, yaml_path, cluster_ind, hmm_param_dir, seed=0, device=0):
        self._lib = _abi.bind_phylohmm(lib())
        self._h = self._lib.lhh_phylohmm_new(str(yaml_path).encode(), int(cluster_ind), str(hmm_param_dir).encode(), int(seed))
        if not self._h:
            raise HostError(self._lib.lhh_last_error().decode())
        _ck(self._lib.lhh_phylohmm_set_device(self._h, int(device)))

    def set_devices(self, devices):
        """several GPUs of the box for the batched calls (before the first scoring call)"""
        arr = np.ascontiguousarray(devices, dtype=np.int32)
        _ck(self._lib.lhh_phylohmm_set_devices(self._h, arr.ctypes.data_as(_abi._i32p), len(arr)))

    def set_device_sampler(self, on):
        """RunPipeline: draw the annotations' states on the device (default) or with the host sampler."""
        _ck(self._lib.lhh_phylohmm_set_device_sampler(self._h, int(bool(on))))

    def set_host_threads(self, n):
        """host threads for Newick ingest and traversal planning (0 = all hardware threads)"""
        _ck(self._lib.lhh_phylohmm_set_host_threads(self._h, int(n)))

    def close(self):
        if getattr(self, "_h", None):
            self._lib.lhh_phylohmm_free(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    # ---- read-outs ---------------------------------------------------------------------------
    def f64(self, name):
        r, c = C.c_int32(), C.c_int32()
        _ck(self._lib.lhh_phylohmm_get_f64(self._h, name.encode(), None, C.byref(r), C.byref(c)))
        out = np.zeros((c.value, r.value))
        _ck(self._lib.lhh_phylohmm_get_f64(self._h, name.encode(), out.ctypes.data_as(_abi._f64p), C.byref(r), C.byref(c)))
        return out.T.copy()

    def i32(self, name):
        r, c = C.c_int32(), C.c_int32()
        _ck(self._lib.lhh_phylohmm_get_i32(self._h, name.encode(), None, C.byref(r), C.byref(c)))
        out = np.zeros((c.value, r.value), dtype=np.int32)
        _ck(self._lib.lhh_phylohmm_get_i32(self._h, name.encode(), out.ctypes.data_as(_abi._i32p), C.byref(r), C.byref(c)))
        return out.T.copy()

    def strs(self, name):
        n = self._lib.lhh_phylohmm_get_str(self._h, name.encode(), None, 0)
        if n < 0:
            raise KeyError(name)
        buf = C.create_string_buffer(n)
        self._lib.lhh_phylohmm_get_str(self._h, name.encode(), buf, n)
        s = buf.value.decode()
        return s.split("\n") if s else []

    def family(self):
        """The flat family description as a dict of numpy arrays (what lh_family_create takes)."""
        d = _abi.FamilyDesc()
        _ck(self._lib.lhh_phylohmm_family_desc(self._h, C.byref(d)))
        T, S = d.n_tips, d.n_sites
        fam = {k: int(getattr(d, k)) for k in _abi._FAMILY_SCALARS}

        def arr(ptr, n, dt):
            return np.ctypeslib.as_array(ptr, shape=(max(n, 1),))[:n].astype(dt, copy=True)

        fam["tip_codes"] = arr(d.tip_codes, T * S, np.uint8).reshape(T, S)
        for reg, n in (("vpad", d.n_v), ("vgerm", d.n_v), ("dgerm", d.n_d), ("jgerm", d.n_j), ("jpad", d.n_j)):
            off = arr(getattr(d, reg + "_off"), n + 1, np.int32)
            fam[reg + "_off"] = off
            fam[reg + "_inds"] = arr(getattr(d, reg + "_inds"), int(off[-1]), np.int32)
        fam["vd_inds"] = arr(d.vd_inds, d.vd_sites * d.vd_states, np.int32)
        fam["dj_inds"] = arr(d.dj_inds, d.dj_sites * d.dj_states, np.int32)
        fam["v_gene_prob"] = arr(d.v_gene_prob, d.n_v, np.float64)
        fam["v_trans_prod"] = arr(d.v_trans_prod, d.n_v, np.float64)
        fam["vpad_trans"] = arr(d.vpad_trans, d.n_v, np.float64)
        fam["jpad_trans"] = arr(d.jpad_trans, d.n_j, np.float64)
        n_mid = d.n_d if d.has_d else d.n_j
        fam["t_v_vd"] = arr(d.t_v_vd, d.n_v * d.vd_states, np.float64)
        fam["t_vd"] = arr(d.t_vd, d.vd_states * d.vd_states, np.float64)
        fam["t_vd_d"] = arr(d.t_vd_d, d.vd_states * n_mid, np.float64)
        fam["t_d_dj"] = arr(d.t_d_dj, d.n_d * d.dj_states, np.float64) if d.has_d else np.zeros(0)
        fam["t_dj"] = arr(d.t_dj, d.dj_states * d.dj_states, np.float64) if d.has_d else np.zeros(0)
        fam["t_dj_j"] = arr(d.t_dj_j, d.dj_states * d.n_j, np.float64) if d.has_d else np.zeros(0)
        return fam

    # ---- the reference API -------------------------------------------------------------------
    def InitializePhyloParameters(self, newick_path, er, pi, alpha, num_rates):
        er = np.ascontiguousarray(er, dtype=np.float64)
        pi = np.ascontiguousarray(pi, dtype=np.float64)
        _ck(self._lib.lhh_phylohmm_init_params(self._h, str(newick_path).encode(), er.ctypes.data_as(_abi._f64p),
                                               pi.ctypes.data_as(_abi._f64p), float(alpha), int(num_rates)))

    def InitializePhyloEmission(self):
        _ck(self._lib.lhh_phylohmm_init_emission(self._h))

    def LogLikelihood(self):
        v = C.c_double()
        _ck(self._lib.lhh_phylohmm_loglikelihood(self._h, C.byref(v)))
        return v.value

    def SampleNaiveSequence(self):
        buf = C.create_string_buffer(1 << 16)
        _ck(self._lib.lhh_phylohmm_sample_naive(self._h, buf, len(buf)))
        return buf.value.decode()

    def RunPipeline(self, input_path, output_path, num_rates, batch_size=256):
        _ck(self._lib.lhh_phylohmm_run_pipeline(self._h, str(input_path).encode(), str(output_path).encode(),
                                                int(num_rates), int(batch_size)))

    def timings(self):
        """lh_get_timings of the device handle behind the object (the last scoring call)."""
        h, t = C.c_void_p(), _abi.Timings()
        _ck(self._lib.lhh_phylohmm_device_family(self._h, C.byref(h)))
        if self._lib.lh_get_timings(h, C.byref(t)):
            raise HostError(self._lib.lh_last_error().decode())
        return {k: getattr(t, k) for k, _ in _abi.Timings._fields_}

    def set_option(self, key, value):
        h = C.c_void_p()
        _ck(self._lib.lhh_phylohmm_device_family(self._h, C.byref(h)))
        if self._lib.lh_set_option(h, key.encode(), int(value)):
            raise HostError(self._lib.lh_last_error().decode())

    def ScoreSamples(self, newicks, er, pi, alpha, num_rates):
        """newicks: list of str, or a ctypes array made by encode_newicks (marshalled once, reused)."""
        n = len(newicks)
        er = np.ascontiguousarray(er, dtype=np.float64).reshape(n, 6)
        pi = np.ascontiguousarray(pi, dtype=np.float64).reshape(n, 4)
        alpha = np.ascontiguousarray(alpha, dtype=np.float64).reshape(n)
        ll = np.zeros(n)
        if isinstance(newicks, NewickBuffer):   # the strings where they lie, with their lengths
            _ck(self._lib.lhh_phylohmm_score_samples_len(
                self._h, n, newicks.ptrs, newicks.lengths.ctypes.data_as(C.POINTER(C.c_int64)), er.ctypes.data_as(_abi._f64p),
                pi.ctypes.data_as(_abi._f64p), alpha.ctypes.data_as(_abi._f64p), int(num_rates), ll.ctypes.data_as(_abi._f64p)))
            return ll
        arr = newicks if isinstance(newicks, C.Array) else encode_newicks(newicks)
        _ck(self._lib.lhh_phylohmm_score_samples(self._h, n, arr, er.ctypes.data_as(_abi._f64p), pi.ctypes.data_as(_abi._f64p),
                                                 alpha.ctypes.data_as(_abi._f64p), int(num_rates), ll.ctypes.data_as(_abi._f64p)))
        return ll
