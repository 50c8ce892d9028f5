"""kb_tpe_scores_noalloc: the scoring entry point for callers that may not allocate GPU memory (a plugin inside an
application that owns the device memory, a loop that must avoid cudaMalloc's synchronisation).  Its scratch is a fixed
array inside the library, shared by every call, so the checks are: same scores as the allocating entry point
(kb_tpe_scores) AND as the CPU oracle (oracle/tpe.c) within the tolerance the TPE parity tests use, on calls in a row
whose second call has a smaller problem than the first (stale table contents must not leak), a finite answer for a
mixture with a zero sigma and for candidates far from every component, and a clear refusal beyond the stated size limit.
Collected last on purpose (file name), and self-contained so that it passes alone in a fresh process as well."""
import numpy as np
import pytest

import katib_b200 as kb
import _oracle as O

pytestmark = pytest.mark.gpu
TOL = 1e-5  # same bound as tests/test_gpu_tpe.py: |a - b| <= TOL * max(1, |b|)


def _case(ctx, torch, seed, n_trials, n_cand, low, high):
    rng = np.random.Generator(np.random.PCG64(seed))
    ctx.space_set([kb.KB_UNIFORM], [low], [high])
    vals = rng.uniform(low, high, (n_trials, 1))
    loss = (vals[:, 0] - 0.3 * (low + high)) ** 2 + rng.normal(0, 0.2 * (high - low), n_trials)
    ctx.tpe_set_history(vals, loss)
    cand = rng.uniform(low, high, (n_cand, 1))
    ref = ctx.tpe_scores(cand)[:, 0]                       # existing, allocating entry point
    dev = torch.device("cuda", 0)
    sides = []
    keep = []
    for side in (0, 1):
        w, mu, sg, _ = ctx.tpe_get_parzen(0, side)
        t = [torch.from_numpy(np.ascontiguousarray(a)).to(dev) for a in (w, mu, sg)]
        keep.append(t)
        sides.append((len(w), t[0].data_ptr(), t[1].data_ptr(), t[2].data_ptr()))
    x = torch.from_numpy(np.ascontiguousarray(cand[:, 0])).to(dev)
    out = torch.full((n_cand,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    kb.tpe_scores_noalloc(0, n_cand, x.data_ptr(), sides[0], sides[1], low, high, out.data_ptr(),
                          torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    got = out.cpu().numpy()
    err = np.max(np.abs(got - ref) / np.maximum(1.0, np.abs(ref)))
    assert np.all(np.isfinite(got)) and err < TOL, err
    want, _ = O.tpe_score_numeric(cand[:, 0], vals[:, 0], loss, low, high)   # the CPU oracle, same history
    err_o = np.max(np.abs(got - want) / np.maximum(1.0, np.abs(want)))
    assert err_o < TOL, err_o
    return max(err, err_o)


def test_noalloc_matches_allocating_entry_point_twice():
    import torch
    ctx = kb.Context(0)
    try:
        e1 = _case(ctx, torch, 1, 3000, 5000, -2.0, 6.0)     # K_above = 2976: fills most of the static table
        e2 = _case(ctx, torch, 2, 120, 777, 10.0, 10.5)      # then a much smaller problem on the same scratch
        e3 = _case(ctx, torch, 1, 3000, 5000, -2.0, 6.0)     # and the first again
        print("noalloc max rel diff vs kb_tpe_scores: %.2e %.2e %.2e" % (e1, e2, e3))
    finally:
        ctx.close()


def test_noalloc_stays_finite_on_degenerate_mixtures():
    """A caller may hand in any mixture, not only kb_tpe_get_parzen's: a zero sigma is floored like the oracle floors
    it, and a candidate that is many sigmas from every component gives a finite score, not NaN."""
    import torch
    dev = torch.device("cuda", 0)
    w = torch.tensor([0.5, 0.5], dtype=torch.float64, device=dev)
    mu = torch.tensor([0.25, 0.75], dtype=torch.float64, device=dev)
    sg0 = torch.tensor([0.0, 0.01], dtype=torch.float64, device=dev)     # one sigma exactly zero
    sg1 = torch.tensor([0.001, 0.001], dtype=torch.float64, device=dev)  # narrow: x = 0.5 is 250 sigma away
    x = torch.tensor([0.5, 0.25, 0.75, 0.1], dtype=torch.float64, device=dev)
    out = torch.full((4,), float("nan"), dtype=torch.float64, device=dev)
    torch.cuda.synchronize()
    kb.tpe_scores_noalloc(0, 4, x.data_ptr(), (2, w.data_ptr(), mu.data_ptr(), sg0.data_ptr()),
                          (2, w.data_ptr(), mu.data_ptr(), sg1.data_ptr()), 0.0, 1.0, out.data_ptr(),
                          torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    assert np.all(np.isfinite(out.cpu().numpy())), out


def test_noalloc_rejects_sizes_beyond_the_static_scratch():
    import torch
    dev = torch.device("cuda", 0)
    z = torch.ones(8, dtype=torch.float64, device=dev)
    side_ok = (4, z.data_ptr(), z.data_ptr(), z.data_ptr())
    side_big = (4097, z.data_ptr(), z.data_ptr(), z.data_ptr())
    with pytest.raises(kb.KbError) as e:
        kb.tpe_scores_noalloc(0, 8, z.data_ptr(), side_ok, side_big, 0.0, 1.0, z.data_ptr())
    assert "4096" in str(e.value) and e.value.code == -5
    with pytest.raises(kb.KbError):
        kb.tpe_scores_noalloc(0, 8, z.data_ptr(), side_ok, side_ok, 1.0, 1.0, z.data_ptr())  # empty range
