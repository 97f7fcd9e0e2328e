"""Sharding of a tree batch over ranks and the gather of the per-tree log-likelihoods.
Trees of one family cost the same, so a batch is cut into contiguous, equal chunks; the path has no
data-path exchange, the only collective is one all-gather of `per_rank` doubles per step. Works with
any torch.distributed backend (NCCL on the GPUs, gloo in the CPU tests)."""
import numpy as np


def per_rank(n_items, world):
    return (n_items + world - 1) // world


def shard_bounds(n_items, rank, world):
    """[lo, hi) of the rank's chunk; ranks past the end get an empty chunk."""
    per = per_rank(n_items, world)
    lo = min(n_items, rank * per)
    return lo, min(n_items, lo + per)


def shard_batch(tb, rank, world):
    """tb: dict of arrays with the tree axis first -> the rank's slice (contiguous copies)."""
    n = int(np.asarray(tb["root_child"]).shape[0])
    lo, hi = shard_bounds(n, rank, world)
    return {k: np.ascontiguousarray(v[lo:hi]) for k, v in tb.items()}, lo, hi


def gather_loglik(dist, local, n_items, out=None):
    """All-gather the ranks' log-likelihood chunks into one [n_items] tensor on every rank.
    local: 1-D float64 torch tensor holding this rank's chunk (device or CPU)."""
    import torch
    world, per = dist.get_world_size(), per_rank(n_items, dist.get_world_size())
    if local.numel() < per:  # the last chunk may be short: pad to the common length
        pad = torch.zeros(per, dtype=local.dtype, device=local.device)
        pad[:local.numel()] = local
        local = pad
    if out is None:
        out = torch.empty(world * per, dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, local.contiguous())
    return out[:n_items]


def assign_families(costs, world):
    """Longest-processing-time-first assignment of whole families to ranks (BASELINE config 5:
    families sharded across GPUs). costs[f] ~ B_f * (T_f - 2) * S_f. -> list of family ids per rank."""
    order = sorted(range(len(costs)), key=lambda f: -costs[f])
    load, out = [0.0] * world, [[] for _ in range(world)]
    for f in order:
        r = min(range(world), key=lambda i: load[i])
        out[r].append(f)
        load[r] += costs[f]
    return out


def family_cost(n_tips, n_sites, n_trees, n_rates=4):
    """Pruning cost model of one family: trees x inner nodes x xMSA columns x categories."""
    return float(n_trees) * max(1, n_tips - 2) * n_sites * n_rates


def gather_family_loglik(dist, local, assignment, n_trees, device=None):
    """Repertoire gather: every rank scored the families assignment[rank] (lists of family ids, the same
    on every rank - assign_families is deterministic); local = {family id: 1-D float64 torch tensor}.
    One all-gather of the ranks' concatenated log-likelihoods, padded to the longest rank.
    -> list over all families of 1-D tensors, on every rank."""
    import torch
    world, rank = dist.get_world_size(), dist.get_rank()
    per = max(1, max(sum(n_trees[f] for f in fams) for fams in assignment))
    mine = [local[f] for f in assignment[rank]]
    if device is None:
        device = mine[0].device if mine else "cpu"
    buf = torch.zeros(per, dtype=torch.float64, device=device)
    if mine:
        cat = torch.cat(mine)
        buf[:cat.numel()] = cat
    out = torch.empty(world * per, dtype=torch.float64, device=buf.device)
    dist.all_gather_into_tensor(out, buf)
    res = [None] * len(n_trees)
    for r, fams in enumerate(assignment):
        o = r * per
        for f in fams:
            res[f] = out[o:o + n_trees[f]]
            o += n_trees[f]
    return res
