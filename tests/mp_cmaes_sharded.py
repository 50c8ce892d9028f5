"""Two (or more) ranks, one GPU each: CMA-ES with the population sharded by rows must track the single-GPU optimiser.

Launched by tests/test_gpu_cmaes.py::test_sharded_population_tracks_single_gpu (needs >= 2 GPUs) as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tests/mp_cmaes_sharded.py
Every rank also runs the whole population on a second, communicator-less context and compares states after each tell."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    os.environ.setdefault("NCCL_DEBUG", "NONE")
    import torch
    import torch.distributed as dist
    import katib_b200 as kb
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    local = int(os.environ.get("LOCAL_RANK", rank))
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    sharded, full = kb.Context(local), kb.Context(local)
    uid = [kb.Context.unique_id() if rank == 0 else None]
    dist.broadcast_object_list(uid, src=0)
    sharded.comm_init(world, rank, uid[0])
    n, lam = 48, 16 * world
    L, r0 = lam // world, rank * (lam // world)
    rng = np.random.Generator(np.random.PCG64(11))  # same stream on every rank
    Q, _ = np.linalg.qr(rng.normal(size=(n, n)))
    scale = np.logspace(0, 3, n)
    fobj = lambda X: np.sum((X @ Q) ** 2 * scale, axis=1)
    m0, lo, hi = np.full(n, 1.5), np.full(n, -50.0), np.full(n, 50.0)
    for c in (sharded, full):
        c.cmaes_create(m0, 1.0, lam, lo, hi)
    worst = 0.0
    for g in range(8):
        if g % 2 == 0:  # z from the caller
            Z = rng.normal(size=(lam, n))
            Xf = full.cmaes_ask(lam, Z)
            Xs = sharded.cmaes_ask(L, Z[r0:r0 + L])
        else:           # z drawn on the device: one global Philox stream, each rank its rows
            Xf = full.cmaes_ask(lam, None, seed=100 + g)
            Xs = sharded.cmaes_ask(L, None, seed=100 + g)
        # x = m + sigma B D z: while eigenvalues are (nearly) degenerate the eigenbasis B is arbitrary, and two
        # contexts whose C differ in the last bit may pick different ones -- so compare what does not depend on the
        # basis, the Mahalanobis norm (x-m)^T C^-1 (x-m) / sigma^2 = |z|^2, row by row (bounds are wide enough here
        # that no coordinate is clipped).  It pins the z stream, including the per-rank row offset of the Philox draw.
        st = full.cmaes_state()
        Ci = np.linalg.inv(st["C"])
        maha = lambda X: np.einsum("ki,ij,kj->k", X - st["mean"], Ci, X - st["mean"]) / st["sigma"] ** 2
        da = float(np.max(np.abs(maha(Xs) - maha(Xf[r0:r0 + L])) / maha(Xf[r0:r0 + L])))
        worst = max(worst, da)
        if os.environ.get("KB_MP_VERBOSE"):
            print(f"rank {rank} gen {g} ask |z|^2 diff {da:.3e}", flush=True)
        f = fobj(Xf)
        f[3] = f[lam - 2]  # a tie across ranks: the global index decides
        full.cmaes_tell(Xf, f)
        sharded.cmaes_tell(Xf[r0:r0 + L].copy(), f[r0:r0 + L].copy())
        a, b = sharded.cmaes_state(), full.cmaes_state()
        for key in ("mean", "C", "p_sigma", "p_c"):
            den = max(1e-300, float(np.max(np.abs(b[key]))))
            worst = max(worst, float(np.max(np.abs(a[key] - b[key]))) / den)
            if os.environ.get("KB_MP_VERBOSE"):
                print(f"rank {rank} gen {g} {key} diff {float(np.max(np.abs(a[key] - b[key]))) / den:.3e}", flush=True)
        worst = max(worst, abs(a["sigma"] - b["sigma"]) / b["sigma"])
        assert a["generation"] == b["generation"] == g + 1
    t = torch.tensor([worst], device="cuda", dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ok = float(t[0]) < 1e-10
    dist.barrier()
    sharded.close(); full.close()
    dist.barrier()
    dist.destroy_process_group()
    if rank == 0:
        print(f"SHARDED {'OK' if ok else 'MISMATCH'} worst relative difference {float(t[0]):.3e} world {world}", flush=True)
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
