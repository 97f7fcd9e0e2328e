"""lh_newick_parse_noalloc - the device Newick parser with its scratch in a static device array, for callers that
cannot call cudaMalloc (no counterpart in the reference) - against
the existing entry point over the same kernels, lh_debug_parse_newick: the operation arrays must be bit-identical
(integer work; branch lengths are correctly rounded doubles, compared exactly as in tests/test_gpu_newick.py), on
two calls in a row that share the scratch. Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from conftest import load_golden
from linearham_b200 import device, host

pytestmark = pytest.mark.gpu
OPS = ("children", "blen", "root_child", "root_blen")


def _pack(newicks):
    """strings -> (text bytes with every string NUL-terminated on a 16-byte boundary, offs uint64, lens uint32)"""
    raw = [s.encode() for s in newicks]
    lens = np.array([len(r) for r in raw], dtype=np.uint32)
    padded = (lens.astype(np.int64) + 1 + 15) // 16 * 16
    offs = np.concatenate([[0], np.cumsum(padded)[:-1]]).astype(np.uint64)
    text = np.zeros(int(padded.sum()) + 16, dtype=np.uint8)
    for r, o in zip(raw, offs):
        text[int(o):int(o) + len(r)] = np.frombuffer(r, dtype=np.uint8)
    return text, offs, lens


def _parse_noalloc(torch, f, newicks, T):
    n = len(newicks)
    text, offs, lens = _pack(newicks)
    dev = "cuda:0"
    d_text = torch.from_numpy(text).to(dev)
    d_offs = torch.from_numpy(offs.view(np.int64)).to(dev)
    d_lens = torch.from_numpy(lens.view(np.int32)).to(dev)
    out = {"children": torch.full((n, T - 2, 2), -7, dtype=torch.int32, device=dev),
           "blen": torch.full((n, T - 2, 2), -7.0, dtype=torch.float64, device=dev),
           "root_child": torch.full((n,), -7, dtype=torch.int32, device=dev),
           "root_blen": torch.full((n,), -7.0, dtype=torch.float64, device=dev),
           "on_patterns": torch.full((n,), 9, dtype=torch.uint8, device=dev),
           "left_to_host": torch.full((n,), 9, dtype=torch.uint8, device=dev)}
    torch.cuda.synchronize()
    free_before = torch.cuda.mem_get_info()[0]
    rc = device.lib().lh_newick_parse_noalloc(
        f._h, n, d_text.data_ptr(), d_offs.data_ptr(), d_lens.data_ptr(), out["children"].data_ptr(), out["blen"].data_ptr(),
        out["root_child"].data_ptr(), out["root_blen"].data_ptr(), out["on_patterns"].data_ptr(), out["left_to_host"].data_ptr(),
        C.c_void_p(torch.cuda.current_stream().cuda_stream))
    assert rc == 0, device.lib().lh_last_error().decode()
    torch.cuda.synchronize()
    free_after = torch.cuda.mem_get_info()[0]
    return {k: v.cpu().numpy() for k, v in out.items()}, free_before, free_after


def test_noalloc_parser_matches_the_existing_entry_point_twice_in_a_row():
    torch = pytest.importorskip("torch")
    if not torch.cuda.is_available():
        pytest.fail("needs a CUDA device: linearham_b200 has no CPU fallback")
    batches = []
    for name, rep in (("synth_n200", 40), ("synth_n200", 1)):
        fam, tb, exp, meta = load_golden(name)
        nw = [str(s) for s in meta["newicks"]] * rep
        batches.append(nw)
    # the second call parses other trees than the first: the same family's trees in reverse order with one spelled
    # so that the device declines it (status 1) - what the first call left in the scratch must not show
    hard = batches[1][0]
    i = hard.index(":")
    batches[1] = list(reversed(batches[0][:31])) + [hard[:i + 1] + "+" + hard[i + 1:]]
    labels = [str(s) for s in meta["labels"]]
    T = int(fam["n_tips"])
    f = device.Family(fam)
    f.set_labels(labels)
    assert device.lib().lh_newick_parse_noalloc_max_trees(T) >= len(batches[0])
    frees = []
    for call, nw in enumerate(batches):
        want = f.parse_newick(nw)                       # the existing entry point (allocating; same kernels)
        got, free_before, free_after = _parse_noalloc(torch, f, nw, T)
        frees.append((free_before, free_after))
        assert np.array_equal(got["left_to_host"], want["left_to_host"]), call
        keep = want["left_to_host"] == 0
        assert keep.sum() >= len(nw) - 1
        for k in OPS:
            assert np.array_equal(got[k][keep], want[k][keep]), (call, k)
        assert np.array_equal(got["on_patterns"][keep], want["on_patterns"][keep]), call
        ref = host.newick_to_ops([s for s, kp in zip(nw, keep) if kp], labels)   # and the host parser
        for k in OPS:
            assert np.array_equal(got[k][keep], ref[k]), (call, k)
    # the second call (kernels loaded, nothing lazy left) must not have taken device memory
    assert frees[1][1] >= frees[1][0], frees
    # sizes beyond the static scratch are refused with a clear error, and nothing is launched
    too_many = device.lib().lh_newick_parse_noalloc_max_trees(T) + 1
    rc = device.lib().lh_newick_parse_noalloc(f._h, too_many, *([C.c_void_p(16)] * 9), None)
    assert rc != 0 and "static scratch" in device.lib().lh_last_error().decode()
    assert device.lib().lh_newick_parse_noalloc_max_trees(40000) == 0
    f.close()
