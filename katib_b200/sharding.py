"""Host-side sharding rules of the multi-GPU TPE path (one process per GPU).

The candidate set is split into contiguous index ranges; every rank scores its range and reports, per parameter, its
local winner as (FP64 score, global index, value).  The global winner is the best score, lowest global index on ties --
the rule tpe.Sampler.compare applies to the unsharded set (first maximum wins).  The CUDA side does the same merge in
tpe_merge_kernel after one ncclAllGather; this module is the rule written once for bench.py and the CPU (gloo) tests.
"""
import numpy as np


def shard_range(n_total, world, rank):
    """[begin, end) of rank's contiguous slice; the first n_total % world ranks take one extra candidate."""
    base, extra = divmod(n_total, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def merge_winners(scores, indices, values):
    """scores/indices/values: [world, D] arrays of local winners (index < 0 = rank had no candidate).
    Returns ([D] score, [D] index, [D] value) of the global winners."""
    scores, indices, values = np.asarray(scores, float), np.asarray(indices, np.int64), np.asarray(values, float)
    world, D = scores.shape
    out_s, out_i, out_v = np.full(D, -np.inf), np.full(D, -1, np.int64), np.zeros(D)
    for d in range(D):
        for r in range(world):
            if indices[r, d] < 0:
                continue
            better = out_i[d] < 0 or scores[r, d] > out_s[d] or (scores[r, d] == out_s[d] and indices[r, d] < out_i[d])
            if better:
                out_s[d], out_i[d], out_v[d] = scores[r, d], indices[r, d], values[r, d]
    return out_s, out_i, out_v


def merge_status(status_rows):
    """The status row that travels with the winners through the all-gather (tpe_pick_kernel / tpe_merge_kernel):
    status_rows[r] = 1 + the highest parameter whose near-tie table overflowed on rank r, 0 = none.  The call's status is
    the maximum -- an overflow anywhere is an overflow everywhere, so every rank repeats the pass (or fails) together
    and no rank trusts winners merged from another rank's incomplete table."""
    return int(max(int(s) for s in status_rows)) if len(status_rows) else 0


def suggest_shard(n_ei_candidates, world, rank, shard_min=65536):
    """Index range of the Philox candidate stream rank draws and scores in kb_tpe_suggest: sharded (shard_range) from
    shard_min candidates on, otherwise the whole set with no collective."""
    if world > 1 and n_ei_candidates >= shard_min:
        return shard_range(n_ei_candidates, world, rank)
    return 0, n_ei_candidates


def population_row_weights(f_all, weights, begin, end):
    """CMA-ES with the population sharded by rows (cma_tell_sharded in kb_cmaes.cu): the recombination weight of each of
    this rank's rows [begin, end).  The ranking is global -- ascending objective value, ties by global row index (a stable
    sort, as Optimizer.Tell sorts its solutions) -- and weights[p] belongs to the row ranked p."""
    f_all = np.asarray(f_all, float)
    order = np.argsort(f_all, kind="stable")
    w = np.empty(len(f_all))
    w[order] = np.asarray(weights, float)
    return w[begin:end]
