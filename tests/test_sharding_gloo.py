"""N>1 path on CPU: two gloo ranks shard a tree batch, score their chunks (with the CPU checker, since no
GPU is present), all-gather the log-likelihoods and must reproduce the single-process result
bit for bit; plus the shard arithmetic and the family scheduler."""
import os
import socket
import sys

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from conftest import ROOT, load_golden
from linearham_b200 import sharding


def test_shard_bounds_cover_everything_once():
    for n in (0, 1, 5, 16, 17, 1000, 16384):
        for world in (1, 2, 3, 4, 8):
            spans = [sharding.shard_bounds(n, r, world) for r in range(world)]
            assert spans[0][0] == 0 and spans[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(spans, spans[1:]))
            assert max(hi - lo for lo, hi in spans) == (sharding.per_rank(n, world) if n else 0)


def test_family_assignment_balances_load():
    rng = np.random.default_rng(0)
    costs = list(rng.lognormal(0, 1, 200))
    parts = sharding.assign_families(costs, 8)
    assert sorted(f for p in parts for f in p) == list(range(200))
    loads = [sum(costs[f] for f in p) for p in parts]
    assert max(loads) <= 1.1 * (sum(costs) / 8) + max(costs) * 0.5


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _worker(rank, world, port, name, rep, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import pyoracle as po
    fam, tb, exp, _ = load_golden(name)
    tb = {k: np.concatenate([v] * rep) for k, v in tb.items()}     # an uneven batch over 2 ranks
    n = tb["root_child"].shape[0]
    shard, lo, hi = sharding.shard_batch(tb, rank, world)
    ll, _ = po.port_score_batch(fam, shard, want_outputs=False)
    full = sharding.gather_loglik(dist, torch.from_numpy(ll), n)
    q.put((rank, full.numpy().copy(), lo, hi))
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("name,rep", [("synth_n12", 1), ("catch_a", 3), ("synth_n40", 1)])
def test_two_gloo_ranks_reproduce_the_single_process_result(name, rep):
    from oracle import pyoracle as po
    fam, tb, exp, _ = load_golden(name)
    big = {k: np.concatenate([v] * rep) for k, v in tb.items()}
    single, _ = po.port_score_batch(fam, big, want_outputs=False)
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, name, rep, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    spans = sorted((lo, hi) for _, _, lo, hi in got)
    assert spans[0][0] == 0 and spans[0][1] == spans[1][0] and spans[1][1] == len(single)
    for _, full, _, _ in got:
        assert np.array_equal(full, single)          # bit-identical on every rank


# ---- BASELINE config 5: whole families sharded over ranks --------------------------------------

REPERTOIRE = ["synth_n200", "catch_a", "synth_n40", "synth_igk_n30", "synth_n12", "catch_b", "synth_n24_r1"]


def _rep_worker(rank, world, port, q):
    sys.path.insert(0, ROOT)
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from oracle import pyoracle as po
    fams = [load_golden(n) for n in REPERTOIRE]
    n_trees = [int(tb["root_child"].shape[0]) for _, tb, _, _ in fams]
    costs = [sharding.family_cost(f["n_tips"], f["n_sites"], n, tb["rates"].shape[1])
             for (f, tb, _, _), n in zip(fams, n_trees)]
    parts = sharding.assign_families(costs, world)
    local = {i: torch.from_numpy(po.port_score_batch(fams[i][0], fams[i][1], want_outputs=False)[0])
             for i in parts[rank]}
    res = sharding.gather_family_loglik(dist, local, parts, n_trees)
    q.put((rank, [r.numpy().copy() for r in res], parts))
    dist.barrier()
    dist.destroy_process_group()


def test_two_gloo_ranks_score_a_repertoire_of_families():
    from oracle import pyoracle as po
    single = [po.port_score_batch(f, tb, want_outputs=False)[0] for f, tb, _, _ in map(load_golden, REPERTOIRE)]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rep_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    parts = got[0][2]
    assert got[1][2] == parts and all(parts) and sorted(f for p in parts for f in p) == list(range(len(REPERTOIRE)))
    for _, res, _ in got:
        for a, b in zip(res, single):
            assert np.array_equal(a, b, equal_nan=True)   # bit-identical on every rank
