"""N > 1 path on CPU: two gloo ranks each score their candidate shard with the CPU oracle, all-gather their local
winners, and merge by the library's rule (best score, lowest global index).  The result must equal the oracle's
winners on the unsharded set -- including when the best value is duplicated across shards."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import _oracle as O
from katib_b200.sharding import merge_status, merge_winners, shard_range, suggest_shard


def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _problem():
    rng = np.random.Generator(np.random.PCG64(42))
    n, nc = 200, 1501
    vals = np.stack([rng.uniform(-1, 1, n), np.floor(rng.uniform(0, 8, n))], axis=1)
    loss = vals[:, 0] ** 2 + 0.1 * vals[:, 1] + rng.normal(0, 0.1, n)
    cand = np.stack([rng.uniform(-1, 1, nc), np.floor(rng.uniform(0, 8, nc))], axis=1)
    cand[1200] = cand[3]  # duplicates across the shard boundary
    return vals, loss, cand


def _score(vals, loss, cand):
    s0, b0 = O.tpe_score_numeric(cand[:, 0], vals[:, 0], loss, -1.0, 1.0)
    s1, b1 = O.tpe_score_numeric(cand[:, 1], vals[:, 1], loss, -0.5, 7.5, 1.0)
    return np.stack([s0, s1], axis=1), [b0, b1]


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    vals, loss, cand = _problem()
    lo, hi = shard_range(len(cand), world, rank)
    sc, best = _score(vals, loss, cand[lo:hi])
    local = torch.tensor([[sc[best[d], d], float(lo + best[d]), cand[lo + best[d], d]] for d in range(2)], dtype=torch.float64)
    gathered = [torch.zeros_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    g = torch.stack(gathered).numpy()  # [world, D, 3]
    s, i, v = merge_winners(g[:, :, 0], g[:, :, 1].astype(np.int64), g[:, :, 2])
    # the status row rides in the same gather: only rank 1 "overflowed" (parameter 1), both ranks must see it
    st = torch.tensor([0.0 if rank == 0 else 2.0], dtype=torch.float64)
    sts = [torch.zeros_like(st) for _ in range(world)]
    dist.all_gather(sts, st)
    status = merge_status([float(x[0]) for x in sts])
    statuses = [torch.zeros(1, dtype=torch.float64) for _ in range(world)]
    dist.all_gather(statuses, torch.tensor([float(status)], dtype=torch.float64))
    assert all(float(x[0]) == 2.0 for x in statuses)  # every rank took the same decision
    if rank == 0:
        q.put((s.tolist(), i.tolist(), v.tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_range_covers_everything():
    for n, w in ((10, 3), (7, 8), (1 << 26, 8), (0, 2)):
        spans = [shard_range(n, w, r) for r in range(w)]
        assert spans[0][0] == 0 and spans[-1][1] == n
        assert all(spans[r][1] == spans[r + 1][0] for r in range(w - 1))
        assert max(e - b for b, e in spans) - min(e - b for b, e in spans) <= 1


def test_suggest_sharding_rule():
    assert suggest_shard(24, 8, 3) == (0, 24)                      # small requests: no collective, any rank alone
    spans = [suggest_shard(1 << 26, 8, r) for r in range(8)]
    assert spans[0][0] == 0 and spans[-1][1] == 1 << 26 and all(spans[r][1] == spans[r + 1][0] for r in range(7))
    assert suggest_shard(65536, 2, 1) == (32768, 65536) and suggest_shard(65535, 2, 1) == (0, 65535)
    assert merge_status([0, 0, 3, 1]) == 3 and merge_status([0, 0]) == 0


def test_two_rank_merge_equals_unsharded_oracle():
    O.lib()  # build the oracle before forking
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    s, i, v = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    vals, loss, cand = _problem()
    sc, best = _score(vals, loss, cand)
    assert i == best
    assert v == [cand[best[0], 0], cand[best[1], 1]]
    np.testing.assert_allclose(s, [sc[best[0], 0], sc[best[1], 1]], rtol=0, atol=0)


def test_merge_rule_ties_and_empty_ranks():
    s, i, v = merge_winners([[1.0, 2.0], [1.0, 3.0], [0.0, 0.0]], [[7, 5], [4, 9], [-1, -1]], [[.1, .2], [.3, .4], [0, 0]])
    assert list(i) == [4, 9] and list(v) == [.3, .4] and list(s) == [1.0, 3.0]


def _cma_problem():
    rng = np.random.Generator(np.random.PCG64(5))
    lam, n = 24, 7
    Y = rng.normal(size=(lam, n))
    f = np.sum(Y ** 2, axis=1)
    f[5] = f[20]  # a tie across the shard boundary: the global row index decides
    wts = np.log((lam + 1) / 2) - np.log(np.arange(1, lam + 1))  # positive for the better half, negative after
    return Y, f, wts


def _cma_worker(rank, world, port, q):
    from katib_b200.sharding import population_row_weights
    os.environ["MASTER_ADDR"], os.environ["MASTER_PORT"] = "127.0.0.1", str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    Y, f, wts = _cma_problem()
    lo, hi = shard_range(len(f), world, rank)
    mine = torch.tensor(f[lo:hi])
    gathered = [torch.zeros_like(mine) for _ in range(world)]
    dist.all_gather(gathered, mine)                      # the ranking needs every objective value
    w = population_row_weights(torch.cat(gathered).numpy(), wts, lo, hi)
    yw = torch.tensor((np.maximum(w, 0.0)[:, None] * Y[lo:hi]).sum(axis=0))
    P = torch.tensor((Y[lo:hi].T * w) @ Y[lo:hi])       # this rank's share of sum_k w_k y_k y_k^T
    dist.all_reduce(yw)
    dist.all_reduce(P)
    if rank == 0:
        q.put((w.tolist(), yw.numpy().tolist(), P.numpy().tolist()))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_cmaes_partials_equal_the_unsharded_sums():
    """Host-side rule of the sharded CMA-ES tell: global stable ranking -> per-row weights; the all-reduced y_w and
    rank-mu partials equal the single-process sums."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_cma_worker, args=(r, 2, port, q)) for r in range(2)]
    [p.start() for p in procs]
    w0, yw, P = q.get(timeout=120)
    [p.join(timeout=60) for p in procs]
    assert all(p.exitcode == 0 for p in procs)
    Y, f, wts = _cma_problem()
    order = sorted(range(len(f)), key=lambda k: (f[k], k))
    w = np.empty(len(f)); w[order] = wts
    assert w[5] > w[20]                                   # the tie went to the lower global index
    np.testing.assert_array_equal(w0, w[:12])
    np.testing.assert_allclose(yw, (np.maximum(w, 0)[:, None] * Y).sum(axis=0), rtol=1e-14, atol=1e-15)
    np.testing.assert_allclose(P, (Y.T * w) @ Y, rtol=1e-13, atol=1e-14)
