"""KB_OPT_LOCAL_ORIGIN (tpe_score_cont_local_kernel): candidates bucketed by value, every thread block measuring x - mu
from its own origin.  Same sums as the kernel that keeps the candidates' order: scores within 1e-5 relative of the FP64
oracle (north_star), winners identical -- first index among equal values included, although the blocks now meet the
candidates in another order (the rule of tpe.Sampler.Sample's argmax, reached from sample.go:45-90)."""
import numpy as np
import pytest

import _oracle as O
import katib_b200 as kb
from test_gpu_tpe import MIXED, TOL, make_history, oracle_scores, rel_err, sample_candidates

pytestmark = pytest.mark.gpu
KB_OPT_LOCAL_ORIGIN = 3


@pytest.fixture()
def ctx():
    c = kb.Context(0)
    yield c
    c.close()


def clustered(rng, n, lo, hi, centres, width):
    """most candidates in tight clusters (blocks that use their own origin), the rest spread over the range"""
    k = rng.integers(0, len(centres), n)
    x = np.asarray(centres)[k] + rng.normal(0, width * (hi - lo), n)
    spread = rng.uniform(lo, hi, n)
    x = np.where(rng.uniform(0, 1, n) < 0.9, x, spread)
    return np.clip(x, lo, np.nextafter(hi, lo))


def test_local_origin_scores_and_winners_vs_oracle(ctx):
    rng = np.random.Generator(np.random.PCG64(31))
    D, n, nc = 6, 1500, 40000
    lows = np.array([-5.0, 0.0, 1e-3, -1e4, 2.0, -0.5])
    highs = np.array([5.0, 1.0, 3e-3, 1e4, 2.5, 0.5])
    ctx.space_set([kb.KB_UNIFORM] * D, lows, highs)
    vals = rng.uniform(lows, highs, (n, D))
    loss = np.sum(((vals - lows) / (highs - lows) - 0.4) ** 2, axis=1) + rng.normal(0, 0.05, n)
    ctx.tpe_set_history(vals, loss)
    cand = np.stack([clustered(rng, nc, lows[d], highs[d], lows[d] + (highs[d] - lows[d]) * np.array([0.12, 0.4, 0.77]), 1e-3)
                     for d in range(D)], axis=1)
    ctx.tpe_set_option(KB_OPT_LOCAL_ORIGIN, 2)
    got = ctx.tpe_scores(cand)
    idx, val, sc = ctx.tpe_score_select(cand)
    own, wide = ctx.tpe_local_stats()
    assert own > 0 and wide > 0, (own, wide)  # both forms ran
    worst = 0.0
    for d in range(D):
        ref, b = O.tpe_score_numeric(cand[:, d], vals[:, d], loss, lows[d], highs[d])
        worst = max(worst, rel_err(got[:, d], ref))
        assert idx[d] == b and val[d] == cand[b, d], d
        assert abs(sc[d] - ref[b]) <= 1e-9 * max(1.0, abs(ref[b]))
    assert worst < TOL, worst
    # against the kernel that keeps the candidates' order: the same sums, a different FP32 rounding pattern
    ctx.tpe_set_option(KB_OPT_LOCAL_ORIGIN, 0)
    plain = ctx.tpe_scores(cand)
    idx0, val0, _ = ctx.tpe_score_select(cand)
    assert list(idx0) == list(idx) and np.array_equal(val0, val)
    diff = np.max(np.abs(plain - got) / np.maximum(1.0, np.abs(plain)))
    print("local origin vs plain: max rel diff %.2e, vs oracle %.2e, blocks own/wide %d/%d" % (diff, worst, own, wide))
    assert diff < 2.5e-6, diff


@pytest.mark.parametrize("n_cand", [1, 24, 2049, 5000])
def test_local_origin_small_and_ragged_sets_mixed_space(ctx, n_cand):
    rng = np.random.Generator(np.random.PCG64(100 + n_cand))
    sp = MIXED
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    vals = make_history(rng, 400, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = (vals[:, 0] - 1.0) ** 2 + 0.05 * (vals[:, 1] - 3) ** 2 + (vals[:, 2] != 1) + rng.normal(0, 0.5, 400)
    ctx.tpe_set_history(vals, loss)
    cand = sample_candidates(rng, n_cand, sp, vals)
    ref, best = oracle_scores(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"], vals, loss, cand)
    ctx.tpe_set_option(KB_OPT_LOCAL_ORIGIN, 2)
    got = ctx.tpe_scores(cand)
    for d in range(len(sp["kinds"])):
        assert rel_err(got[:, d], ref[:, d]) < TOL, (d, rel_err(got[:, d], ref[:, d]))
    idx, val, sc = ctx.tpe_score_select(cand, index_offset=77)
    assert list(idx - 77) == best


def test_local_origin_first_index_wins_among_equal_values(ctx):
    """every value three times, in an order the bucketing undoes: the lowest index must still win"""
    rng = np.random.Generator(np.random.PCG64(12))
    ctx.space_set([kb.KB_UNIFORM, kb.KB_UNIFORM], [0.0, -3.0], [1.0, 3.0])
    vals = np.stack([rng.uniform(0, 1, 300), rng.uniform(-3, 3, 300)], axis=1)
    loss = (vals[:, 0] - 0.3) ** 2 + 0.1 * (vals[:, 1] - 1.0) ** 2
    ctx.tpe_set_history(vals, loss)
    base = np.stack([clustered(rng, 15000, 0.0, 1.0, [0.3, 0.31], 2e-3), clustered(rng, 15000, -3.0, 3.0, [1.0], 1e-3)], axis=1)
    cand = np.concatenate([base, base[::-1], base])
    ctx.tpe_set_option(KB_OPT_LOCAL_ORIGIN, 2)
    idx, val, _ = ctx.tpe_score_select(cand)
    for d, (lo, hi) in enumerate([(0.0, 1.0), (-3.0, 3.0)]):
        _, b = O.tpe_score_numeric(cand[:, d], vals[:, d], loss, lo, hi)
        assert idx[d] == b, (d, idx[d], b)


def test_local_origin_suggest_same_suggestion(ctx):
    """kb_tpe_suggest draws into the staging buffers: the bucketed path must pick what the ordered path picks"""
    rng = np.random.Generator(np.random.PCG64(5))
    sp = MIXED
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    vals = make_history(rng, 600, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = (vals[:, 0] - 1.0) ** 2 + 0.05 * (vals[:, 1] - 3) ** 2 + (vals[:, 2] != 1) + rng.normal(0, 0.5, 600)
    ctx.tpe_set_history(vals, loss)
    out = {}
    for opt in (0, 1, 2):
        ctx.tpe_set_option(KB_OPT_LOCAL_ORIGIN, opt)
        out[opt] = (ctx.tpe_suggest(300_000, seed=9), ctx.tpe_suggest(5000, seed=9))
    for opt in (1, 2):
        assert np.array_equal(out[opt][0], out[0][0]) and np.array_equal(out[opt][1], out[0][1]), opt
    own, wide = ctx.tpe_local_stats()
    assert own > 0


def test_local_origin_option_range(ctx):
    with pytest.raises(kb.KbError):
        ctx.tpe_set_option(KB_OPT_LOCAL_ORIGIN, 3)
