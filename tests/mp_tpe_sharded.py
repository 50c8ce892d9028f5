"""Two (or more) ranks, one GPU each: TPE winners of a candidate set sharded by index range over NCCL ranks must equal
the winners of the whole set on one GPU and the CPU oracle's (first maximum wins: the rule of tpe.Sampler.Sample's
argmax, reached from sample.go:45-90).

Launched by tests/test_gpu_tpe_sharded.py (needs >= 2 GPUs) as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/mp_tpe_sharded.py
Every rank holds the WHOLE candidate set (same seed), scores its own slice through kb_tpe_score_select with the
communicator set, and scores everything again on a second, communicator-less context.

Cases: (1) mixed space, the best candidate of every parameter copied to a higher index inside the other shards and one
row duplicated across each shard boundary (the lower global index must win on every rank), host-buffer and
device-pointer entry points, merged + pruned options, kb_tpe_suggest sharded against unsharded; (2) empty history = every score equal: all ranks overflow their
near-tie tables, repeat the pass together, index 0 wins; (3) a flat landscape where only the LAST rank's slice has
many distinct values: that rank alone overflows, and every rank must still repeat the pass (the status row travels
with the winners) and agree."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
sys.path.insert(0, ROOT)
sys.path.insert(0, HERE)


def _uid(dist, kb, rank):
    uid = [kb.Context.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    return uid[0]


def main():
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    import torch
    import torch.distributed as dist
    import katib_b200 as kb
    from katib_b200.sharding import shard_range
    from test_gpu_tpe import MIXED, make_history, oracle_scores, sample_candidates

    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    sharded, full = kb.Context(local), kb.Context(local)
    sharded.comm_init(world, rank, _uid(dist, kb, rank))
    failures = []

    def check(name, cond, detail=""):
        if not cond:
            failures.append(f"rank {rank} {name}: {detail}")

    # ---- case 1: mixed space, duplicates of the winners planted across the shard boundaries
    sp = MIXED
    D = len(sp["kinds"])
    rng = np.random.Generator(np.random.PCG64(2024))  # the same stream on every rank
    n_trials, n_cand = 400, 6000 * world + 37
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = (vals[:, 0] - 1.0) ** 2 + 0.05 * (vals[:, 1] - 3) ** 2 + (vals[:, 2] != 1) + rng.normal(0, 0.5, n_trials)
    cand = sample_candidates(rng, n_cand, sp, vals)
    _, best0 = oracle_scores(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"], vals, loss, cand)
    for d in range(D):  # a copy of each parameter's winner at a higher index inside every shard
        for r in range(world):
            lo_r, hi_r = shard_range(n_cand, world, r)
            j = lo_r + (7 * d + 3) % max(1, hi_r - lo_r)
            if j > best0[d]:
                cand[j, d] = cand[best0[d], d]
    for r in range(1, world):  # ... and the row before each boundary duplicated onto the next rank's first row
        lo_r, _ = shard_range(n_cand, world, r)
        cand[lo_r] = cand[lo_r - 1]
    ref, best = oracle_scores(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"], vals, loss, cand)
    for c in (sharded, full):
        c.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
        c.tpe_set_history(vals, loss)
    off, end = shard_range(n_cand, world, rank)
    mine = np.ascontiguousarray(cand[off:end])
    idx_f, val_f, sc_f = full.tpe_score_select(cand)
    idx_s, val_s, sc_s = sharded.tpe_score_select(mine, index_offset=off)
    check("case1 host: sharded == oracle", list(idx_s) == best, f"{list(idx_s)} vs {best}")
    check("case1 host: single == oracle", list(idx_f) == best, f"{list(idx_f)} vs {best}")
    check("case1 host: values/scores", np.array_equal(val_s, val_f) and np.array_equal(sc_s, sc_f))
    t = torch.from_numpy(mine).to(dev)
    torch.cuda.synchronize()
    idx_d, val_d, _ = sharded.tpe_score_select_device(end - off, t.data_ptr(), off)
    check("case1 device: sharded == oracle", list(idx_d) == best and np.array_equal(val_d, val_f), f"{list(idx_d)} vs {best}")
    for opt in ((1, 1), (2, 1)):  # merged duplicates and the pruned variant go through the same merge
        sharded.tpe_set_option(*opt)
    idx_m = sharded.tpe_score_select(mine, index_offset=off)[0]
    check("case1 merged+pruned: sharded == oracle", list(idx_m) == best, f"{list(idx_m)} vs {best}")
    for opt in ((1, 0), (2, 0)):
        sharded.tpe_set_option(*opt)
    # kb_tpe_suggest with a communicator: every rank draws and scores its index range of ONE Philox stream and all of
    # them return what the single-GPU call returns (200 000 candidates: above the sharding threshold)
    v_full = full.tpe_suggest(200_000, seed=77)
    v_sh = sharded.tpe_suggest(200_000, seed=77)
    check("case1 suggest: sharded == single", np.array_equal(v_sh, v_full), f"{v_sh} vs {v_full}")
    # the same with the bucketed kernel (KB_OPT_LOCAL_ORIGIN = 2: whatever the history length) on the sharded side only
    sharded.tpe_set_option(3, 2)
    v_loc = sharded.tpe_suggest(200_000, seed=77)
    idx_l = sharded.tpe_score_select(mine, index_offset=off)[0]
    sharded.tpe_set_option(3, 1)
    check("case1 suggest, bucketed kernel: sharded == single", np.array_equal(v_loc, v_full), f"{v_loc} vs {v_full}")
    check("case1 host, bucketed kernel: sharded == oracle", list(idx_l) == best, f"{list(idx_l)} vs {best}")
    if rank == 0:  # below the threshold there is no collective: one rank may call alone
        check("case1 suggest small: lone rank", np.array_equal(sharded.tpe_suggest(24, seed=5), full.tpe_suggest(24, seed=5)))

    # ---- case 2: empty history, every score is exactly 0: all ranks overflow, repeat together, index 0 wins
    kinds2, lows2, highs2 = [kb.KB_UNIFORM, kb.KB_INT], [0.0, 0], [1.0, 9]
    for c in (sharded, full):
        c.space_set(kinds2, lows2, highs2)
        c.tpe_set_history(np.zeros((0, 2)), np.zeros(0))
    n2 = 30000 * world
    cand2 = np.stack([rng.uniform(0, 1, n2), np.floor(rng.uniform(0, 10, n2))], axis=1)
    off, end = shard_range(n2, world, rank)
    idx_s, val_s, sc_s = sharded.tpe_score_select(np.ascontiguousarray(cand2[off:end]), index_offset=off)
    idx_f, val_f, _ = full.tpe_score_select(cand2)
    check("case2 flat: index 0 wins", list(idx_s) == [0, 0] and list(idx_f) == [0, 0], f"{list(idx_s)} {list(idx_f)}")
    check("case2 flat: values", list(val_s) == [cand2[0, 0], cand2[0, 1]] and abs(sc_s[0]) < 1e-12)

    # ---- case 3: only the LAST rank's slice has many distinct values (it alone overflows its table)
    fresh = kb.Context(local)  # a context whose table is still at its initial size
    fresh.comm_init(world, rank, _uid(dist, kb, rank))
    fresh.space_set(kinds2, lows2, highs2)
    fresh.tpe_set_history(np.zeros((0, 2)), np.zeros(0))
    cand3 = cand2.copy()
    lo_last, _ = shard_range(n2, world, world - 1)
    cand3[:lo_last, 0] = 0.25  # one distinct value on every rank but the last
    cand3[:lo_last, 1] = 4.0
    idx_s, val_s, _ = fresh.tpe_score_select(np.ascontiguousarray(cand3[off:end]), index_offset=off)
    check("case3 one-rank overflow: index 0 wins on every rank", list(idx_s) == [0, 0], f"{list(idx_s)}")
    check("case3 values", list(val_s) == [0.25, 4.0], f"{list(val_s)}")

    t = torch.tensor([len(failures)], device=dev, dtype=torch.int64)
    dist.all_reduce(t)
    ok = int(t[0]) == 0
    dist.barrier()
    for c in (sharded, full, fresh):
        c.close()
    dist.barrier()
    dist.destroy_process_group()
    for f in failures:
        print("FAIL", f, flush=True)
    if rank == 0:
        print(f"TPE SHARDED {'OK' if ok else 'MISMATCH'} world {world}: winners of every rank == single GPU == oracle "
              f"(mixed space with planted duplicates, flat landscape, one-rank overflow)", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
