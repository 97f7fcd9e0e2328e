"""Thin Python handle over the device C ABI (include/linearham_b200.h) for tests and bench.py.
No arithmetic here; every number comes out of liblinearham_b200.so (CUDA, sm_100a).
There is no fallback: if the library or a GPU is missing the calls raise."""
import ctypes as C

import numpy as np

from . import _abi

_f64p = C.POINTER(C.c_double)
_lib = None


def lib():
    global _lib
    if _lib is None:
        _lib = _abi.load_library()
    return _lib


class DeviceError(RuntimeError):
    pass


def device_count():
    n = C.c_int()
    if lib().lh_device_count(C.byref(n)):
        return 0
    return n.value


def fp64_peak_tflops(device=0):
    """Measured DFMA rate of this GPU (diagnostic kernel in the library)."""
    v = C.c_double()
    if lib().lh_measure_fp64_peak(int(device), C.byref(v)):
        raise DeviceError(lib().lh_last_error().decode())
    return v.value


class Family:
    """Device-resident clonal family (lh_family_t)."""

    def __init__(self, fam, device=0):
        self.fam = fam
        self._h = C.c_void_p()
        d, keep = _abi.family_desc(fam)
        if lib().lh_family_create(C.byref(self._h), C.byref(d), int(device)):
            raise DeviceError(lib().lh_last_error().decode())
        self._tb_keep = None

    def rebind(self, fam):
        """The handle takes on another family (lh_family_rebind); buffers, streams and threads stay."""
        d, keep = _abi.family_desc(fam)
        self._ck(lib().lh_family_rebind(self._h, C.byref(d)))
        self.fam = fam
        self._tb_keep = None

    def close(self):
        if self._h:
            lib().lh_family_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise DeviceError(lib().lh_last_error().decode())

    def set_option(self, key, value):
        self._ck(lib().lh_set_option(self._h, key.encode(), int(value)))

    def score(self, tb, outputs=False):
        """Host buffers in, host results out (lh_score_batch). -> loglik[B] (, dict of read-outs)."""
        t, keep = _abi.tree_batch(tb)
        ll = np.zeros(t.n_trees)
        if outputs:
            o, arrs = _abi.alloc_outputs(self.fam, t.n_trees, None if outputs is True else outputs, n_rates=t.n_rates)
            self._ck(lib().lh_score_batch(self._h, C.byref(t), ll.ctypes.data_as(_f64p), C.byref(o)))
            return ll, arrs
        self._ck(lib().lh_score_batch(self._h, C.byref(t), ll.ctypes.data_as(_f64p), None))
        return ll

    def asr_sample(self, tb, msa_codes, naive_codes, naive_row=0, seed=0, loglik=False):
        """Ancestral-state sampling (lh_asr_sample): per tree and alignment site the rate category, then the base of
        every inner node. msa_codes [T][S] uint8 (row naive_row ignored), naive_codes [B][S] uint8.
        -> dict(node_states [B][T-2][S], site_rate [B][S] (, site_rate_loglik [B][R][S]))."""
        t, keep = _abi.tree_batch(tb)
        msa = np.ascontiguousarray(msa_codes, dtype=np.uint8)
        naive = np.ascontiguousarray(naive_codes, dtype=np.uint8)
        B, T, S = t.n_trees, int(self.fam["n_tips"]), msa.shape[1]
        assert msa.shape == (T, S) and naive.shape == (B, S)
        u8p = C.POINTER(C.c_uint8)
        a = _abi.AsrBatch(C.pointer(t), S, int(naive_row), msa.ctypes.data_as(u8p), naive.ctypes.data_as(u8p), int(seed))
        out = {"node_states": np.zeros((B, T - 2, S), dtype=np.uint8), "site_rate": np.zeros((B, S), dtype=np.uint8)}
        srl = None
        if loglik:
            out["site_rate_loglik"] = np.zeros((B, t.n_rates, S))
            srl = out["site_rate_loglik"].ctypes.data_as(_f64p)
        self._ck(lib().lh_asr_sample(self._h, C.byref(a), out["node_states"].ctypes.data_as(u8p),
                                     out["site_rate"].ctypes.data_as(u8p), srl))
        return out

    def sample_paths(self, words):
        """lh_sample_paths on the batch scored last (option keep_tables = 1, or read-outs requested):
        words [B][n_words] uint32 raw generator outputs -> (idx [B][n_idx] int32, flags [B] uint8)."""
        w = np.ascontiguousarray(words, dtype=np.uint32)
        B, fam = w.shape[0], self.fam
        n_idx = 2 + int(fam["vd_sites"]) + ((1 + int(fam["dj_sites"])) if int(fam["has_d"]) else 0)
        idx, flags = np.zeros((B, n_idx), dtype=np.int32), np.zeros(B, dtype=np.uint8)
        self._ck(lib().lh_sample_paths(self._h, B, w.ctypes.data_as(C.POINTER(C.c_uint32)), w.shape[1],
                                       idx.ctypes.data_as(C.POINTER(C.c_int32)), n_idx, flags.ctypes.data_as(C.POINTER(C.c_uint8))))
        return idx, flags

    # Newick text in ------------------------------------------------------------
    def set_labels(self, labels):
        """xMSA row labels (row 0 = "naive") for the device Newick parser (lh_family_set_labels)."""
        arr = (C.c_char_p * len(labels))(*[str(s).encode() for s in labels])
        self._ck(lib().lh_family_set_labels(self._h, arr))

    def score_newick(self, newicks, er, pi, n_rates, alpha=None, rates=None, outputs=False, lengths=False):
        """Newick strings in, host results out (lh_score_newick)."""
        nb, keep = _abi.newick_batch(newicks, er, pi, n_rates, alpha=alpha, rates=rates, lengths=lengths)
        ll = np.zeros(nb.n_trees)
        if outputs:
            o, arrs = _abi.alloc_outputs(self.fam, nb.n_trees, None if outputs is True else outputs, n_rates=n_rates)
            self._ck(lib().lh_score_newick(self._h, C.byref(nb), ll.ctypes.data_as(_f64p), C.byref(o)))
            return ll, arrs
        self._ck(lib().lh_score_newick(self._h, C.byref(nb), ll.ctypes.data_as(_f64p), None))
        return ll

    def parse_newick(self, newicks):
        """Only the device parser (lh_debug_parse_newick) -> dict of the operation arrays + flags + kernel ms."""
        B, T = len(newicks), int(self.fam["n_tips"])
        z = np.zeros((B, 1))
        nb, keep = _abi.newick_batch(newicks, np.zeros((B, 6)), np.ones((B, 4)), 1, rates=z)
        out = {"children": np.zeros((B, T - 2, 2), dtype=np.int32), "blen": np.zeros((B, T - 2, 2)),
               "root_child": np.zeros(B, dtype=np.int32), "root_blen": np.zeros(B),
               "on_patterns": np.zeros(B, dtype=np.uint8), "left_to_host": np.zeros(B, dtype=np.uint8)}
        ms = C.c_float()
        i32p, u8p = C.POINTER(C.c_int32), C.POINTER(C.c_uint8)
        self._ck(lib().lh_debug_parse_newick(
            self._h, C.byref(nb), out["children"].ctypes.data_as(i32p), out["blen"].ctypes.data_as(_f64p),
            out["root_child"].ctypes.data_as(i32p), out["root_blen"].ctypes.data_as(_f64p),
            out["on_patterns"].ctypes.data_as(u8p), out["left_to_host"].ctypes.data_as(u8p), C.byref(ms)))
        out["ms_kernel"] = ms.value
        return out

    # split form --------------------------------------------------------------
    def upload(self, tb):
        t, keep = _abi.tree_batch(tb)
        self._tb_keep = (t, keep)
        self._ck(lib().lh_batch_upload(self._h, C.byref(t)))
        self._n = t.n_trees

    def run(self):
        self._ck(lib().lh_batch_run(self._h))

    def sync(self):
        self._ck(lib().lh_sync(self._h))

    def fetch(self):
        ll = np.zeros(self._n)
        self._ck(lib().lh_batch_fetch(self._h, ll.ctypes.data_as(_f64p), None))
        return ll

    def mark(self, slot):
        self._ck(lib().lh_timer_mark(self._h, int(slot)))

    def elapsed_ms(self, a, b):
        ms = C.c_float()
        self._ck(lib().lh_timer_elapsed_ms(self._h, int(a), int(b), C.byref(ms)))
        return ms.value

    def loglik_dev(self):
        """(device pointer, n) of the log-likelihoods of the last run."""
        p, n = C.c_void_p(), C.c_int32()
        self._ck(lib().lh_loglik_dev(self._h, C.byref(p), C.byref(n)))
        return p.value, n.value

    def fetch_plan(self):
        """(steps [B][T-1][4] int32, pblen [B][2T-3]) of the current batch as the device holds them."""
        T = int(self.fam["n_tips"])
        steps = np.zeros((self._n, T - 1, 4), dtype=np.int32)
        pblen = np.zeros((self._n, 2 * T - 3))
        self._ck(lib().lh_debug_fetch_plan(self._h, steps.ctypes.data_as(C.POINTER(C.c_int32)),
                                           pblen.ctypes.data_as(_f64p)))
        return steps, pblen

    def timings(self):
        t = _abi.Timings()
        self._ck(lib().lh_get_timings(self._h, C.byref(t)))
        return {k: getattr(t, k) for k, _ in _abi.Timings._fields_}


def parse_devices(spec):
    """"0-3", "0,2,5" -> [0, 1, 2, 3] (lh_parse_devices: the --devices flag / LINEARHAM_DEVICES)."""
    out, n = (C.c_int * 64)(), C.c_int()
    if lib().lh_parse_devices(str(spec).encode(), out, 64, C.byref(n)):
        raise DeviceError(lib().lh_group_last_error().decode())
    return [out[i] for i in range(n.value)]


class Group:
    """One clonal family on several GPUs of the box, inside this process (lh_group_t): contiguous shards of the
    trees of a call, one per device; results in one host array."""

    def __init__(self, fam, devices):
        self.fam = fam
        self._h = C.c_void_p()
        d, keep = _abi.family_desc(fam)
        dev = (C.c_int * len(devices))(*[int(x) for x in devices])
        if lib().lh_group_create(C.byref(self._h), C.byref(d), dev, len(devices)):
            raise DeviceError(lib().lh_group_last_error().decode())

    def close(self):
        if self._h:
            lib().lh_group_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def _ck(self, rc):
        if rc:
            raise DeviceError(lib().lh_group_last_error().decode())

    def size(self):
        return lib().lh_group_size(self._h)

    def set_option(self, key, value):
        self._ck(lib().lh_group_set_option(self._h, key.encode(), int(value)))

    def set_labels(self, labels):
        arr = (C.c_char_p * len(labels))(*[str(s).encode() for s in labels])
        self._ck(lib().lh_group_set_labels(self._h, arr))

    def score(self, tb, outputs=False):
        t, keep = _abi.tree_batch(tb)
        ll = np.zeros(t.n_trees)
        if outputs:
            o, arrs = _abi.alloc_outputs(self.fam, t.n_trees, None if outputs is True else outputs, n_rates=t.n_rates)
            self._ck(lib().lh_group_score_batch(self._h, C.byref(t), ll.ctypes.data_as(_f64p), C.byref(o)))
            return ll, arrs
        self._ck(lib().lh_group_score_batch(self._h, C.byref(t), ll.ctypes.data_as(_f64p), None))
        return ll

    def score_newick(self, newicks, er, pi, n_rates, alpha=None, rates=None, outputs=False, lengths=False):
        nb, keep = _abi.newick_batch(newicks, er, pi, n_rates, alpha=alpha, rates=rates, lengths=lengths)
        ll = np.zeros(nb.n_trees)
        if outputs:
            o, arrs = _abi.alloc_outputs(self.fam, nb.n_trees, None if outputs is True else outputs, n_rates=n_rates)
            self._ck(lib().lh_group_score_newick(self._h, C.byref(nb), ll.ctypes.data_as(_f64p), C.byref(o)))
            return ll, arrs
        self._ck(lib().lh_group_score_newick(self._h, C.byref(nb), ll.ctypes.data_as(_f64p), None))
        return ll

    def timings(self, member=-1):
        t = _abi.Timings()
        self._ck(lib().lh_group_get_timings(self._h, int(member), C.byref(t)))
        return {k: getattr(t, k) for k, _ in _abi.Timings._fields_}


class Repertoire:
    """Many independent clonal families on one device (lh_repertoire_t): n_slots device handles, one
    host thread per slot."""

    def __init__(self, device=0, n_slots=3, host_threads=0, devices=None):
        """devices: several GPUs (n_slots handles on each, one queue); else the one `device`."""
        self._h = C.c_void_p()
        devs = [int(device)] if devices is None else [int(x) for x in devices]
        arr = (C.c_int * len(devs))(*devs)
        if lib().lh_repertoire_create_multi(C.byref(self._h), arr, len(devs), int(n_slots), int(host_threads)):
            raise DeviceError(lib().lh_repertoire_last_error().decode())

    def close(self):
        if self._h:
            lib().lh_repertoire_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_option(self, key, value):
        if lib().lh_repertoire_set_option(self._h, key.encode(), int(value)):
            raise DeviceError(lib().lh_repertoire_last_error().decode())

    @staticmethod
    def prepare(families):
        """[(family dict, tree-batch dict), ...] -> marshalled items (reusable across score calls)."""
        n = len(families)
        items = (_abi.RepertoireItem * max(n, 1))()
        keep, outs = [], []
        for i, (fam, tb) in enumerate(families):
            d, kd = _abi.family_desc(fam)
            t, kt = _abi.tree_batch(tb)
            ll = np.full(t.n_trees, np.nan)
            items[i].desc = C.pointer(d)
            items[i].trees = C.pointer(t)
            items[i].loglik = ll.ctypes.data_as(_f64p)
            keep.append((d, kd, t, kt))
            outs.append(ll)
        return dict(items=items, n=n, keep=keep, loglik=outs)

    def score_prepared(self, prep):
        """-> (list of loglik arrays (views, overwritten by the next call), stats dict)."""
        st = _abi.RepertoireStats()
        rc = lib().lh_repertoire_score(self._h, prep["items"], prep["n"], C.byref(st))
        stats = {k: getattr(st, k) for k, _ in _abi.RepertoireStats._fields_}
        stats["status"] = [prep["items"][i].status for i in range(prep["n"])]
        if rc:
            err = DeviceError(lib().lh_repertoire_last_error().decode())
            err.stats = stats
            raise err
        return prep["loglik"], stats

    def score(self, families):
        ll, stats = self.score_prepared(self.prepare(families))
        return [a.copy() for a in ll], stats
