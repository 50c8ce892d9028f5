"""On-device candidate generation (sampleFromGMM / sampleFromCategoricalDist / RandomSampler) and the full
kb_tpe_suggest step.  goptuna's math/rand stream cannot be reproduced, so draws are compared in distribution with
the oracle's sampler (two-sample Kolmogorov-Smirnov per parameter), and suggest is checked for self-consistency:
it must equal score_select over the candidates the same seed generates."""
import numpy as np
import pytest
from scipy import stats

import _oracle as O
import katib_b200 as kb
from test_gpu_tpe import MIXED, make_history, working_domain

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = kb.Context(0)
    yield c
    c.close()


def _setup(ctx, n_trials, seed):
    rng = np.random.Generator(np.random.PCG64(seed))
    sp = MIXED
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = (vals[:, 0] - 1.0) ** 2 + 0.05 * (vals[:, 1] - 3) ** 2 + (vals[:, 2] != 1) + rng.normal(0, 0.5, n_trials)
    ctx.tpe_set_history(vals, loss)
    return sp, vals, loss


def test_samples_follow_the_below_mixture(ctx):
    sp, vals, loss = _setup(ctx, 400, 21)
    n = 40000
    cand = ctx.tpe_sample_candidates(n, seed=1234)
    assert cand.shape == (n, len(sp["kinds"]))
    for d, (k, lo, hi, st, c) in enumerate(zip(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])):
        x = cand[:, d]
        if k == kb.KB_CATEGORICAL:
            pl, _ = O.categorical_tables(vals[:, d], loss, c)
            freq = np.bincount(x.astype(int), minlength=c) / n
            assert np.all(np.abs(freq - pl) < 5 * np.sqrt(pl * (1 - pl) / n) + 1e-3), (d, freq, pl)
            continue
        wl, wh, q, sh = working_domain(k, lo, hi, st)
        below, _ = O.tpe_split(vals[:, d] - sh, loss)
        w, mu, sg = O.parzen_build(below, wl, wh)
        ref = np.zeros(n)
        O.lib().kbo_gmm_sample(99, n, len(w), w, mu, sg, wl, wh, q, ref)
        if q > 0:
            assert np.allclose(x / q, np.round(x / q)) and x.min() >= wl and x.max() <= wh
            lv = np.unique(np.concatenate([x, ref]))
            fa = np.array([(x == v).mean() for v in lv]); fb = np.array([(ref == v).mean() for v in lv])
            assert np.max(np.abs(np.cumsum(fa) - np.cumsum(fb))) < 0.02, d
        else:
            assert x.min() >= wl and x.max() < wh
            ks = stats.ks_2samp(x, ref).statistic
            assert ks < 0.02, (d, ks)


def test_samples_are_deterministic_and_seeded(ctx):
    _setup(ctx, 120, 22)
    a = ctx.tpe_sample_candidates(500, seed=7)
    b = ctx.tpe_sample_candidates(500, seed=7)
    c = ctx.tpe_sample_candidates(500, seed=8)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert np.array_equal(ctx.tpe_sample_candidates(200, seed=7), a[:200])  # counter-based: prefix-stable


@pytest.mark.parametrize("n_ei", [1, 24, 1000])
def test_suggest_equals_select_over_its_own_candidates(ctx, n_ei):
    sp, vals, loss = _setup(ctx, 300, 23)
    out = ctx.tpe_suggest(n_ei, seed=42)
    cand = ctx.tpe_sample_candidates(n_ei, seed=42)
    idx, val, sc = ctx.tpe_score_select(cand)
    assert np.array_equal(out, val)
    for d, k in enumerate(sp["kinds"]):  # the suggestion lies inside the distribution
        if k == kb.KB_CATEGORICAL:
            assert 0 <= out[d] < sp["nch"][d]
        else:
            assert sp["lows"][d] <= out[d] <= sp["highs"][d]


def test_random_fill_ranges_and_uniformity(ctx):
    sp = MIXED
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    n = 30000
    r = ctx.random_fill(seed=5, first=0, n=n)
    assert np.array_equal(ctx.random_fill(seed=5, first=100, n=50), r[100:150])
    for d, (k, lo, hi, st, c) in enumerate(zip(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])):
        x = r[:, d]
        if k == kb.KB_UNIFORM:
            assert x.min() >= lo and x.max() < hi and stats.kstest((x - lo) / (hi - lo), "uniform").statistic < 0.02
        elif k == kb.KB_CATEGORICAL:
            assert set(np.unique(x)) == set(range(c))
        elif k == kb.KB_INT:
            assert x.min() == lo and x.max() == hi and np.all(x == np.round(x))
        else:
            g = (x - lo) / st
            assert np.allclose(g, np.round(g)) and x.min() >= lo and x.max() <= hi


@pytest.mark.parametrize("n_trials,n_cand", [(12, 24), (100, 24), (100, 1), (400, 300), (2000, 100)])
def test_fused_small_suggest_equals_the_staged_pipeline(n_trials, n_cand, monkeypatch):
    """kb_tpe_suggest takes a single fused kernel for small problems (FP64 scoring of every candidate); the staged
    pipeline (FP32 scoring + FP64 re-scoring of the near ties) must pick the same values from the same Philox draws."""
    import katib_b200 as kb
    from test_gpu_tpe import MIXED, make_history
    rng = np.random.Generator(np.random.PCG64(n_trials * 31 + n_cand))
    sp = MIXED
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = (vals[:, 0] - 1.0) ** 2 + 0.05 * (vals[:, 1] - 3) ** 2 + (vals[:, 2] != 1) + rng.normal(0, 0.5, n_trials)
    ctx = kb.Context(0)
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    ctx.tpe_set_history(vals, loss)
    for seed in (1, 2, 3, 99):
        monkeypatch.setenv("KB_SUGGEST_SMALL", "1")
        a = ctx.tpe_suggest(n_cand, seed)
        monkeypatch.setenv("KB_SUGGEST_SMALL", "0")
        b = ctx.tpe_suggest(n_cand, seed)
        assert np.array_equal(a, b), (seed, a, b)
    ctx.close()


def test_step_int_suggestions_stay_on_the_declared_grid():
    """STEP_INT with low not a multiple of the step (low = 1, step = 2): suggestions are {1, 3, 5, 7, 9}, the grid that
    kb_random_fill and the Sobol mapping use -- sampleStepInt shifts the range to [0, high - low] before rounding."""
    rng = np.random.Generator(np.random.PCG64(5))
    ctx = kb.Context(0)
    ctx.space_set([kb.KB_STEP_INT, kb.KB_STEP_INT], [1, -7], [9, 8], [2, 3])
    n = 60
    vals = np.stack([1 + 2 * np.floor(rng.uniform(0, 5, n)), -7 + 3 * np.floor(rng.uniform(0, 6, n))], axis=1)
    loss = (vals[:, 0] - 5) ** 2 + np.abs(vals[:, 1] - 2) + rng.normal(0, 0.1, n)
    ctx.tpe_set_history(vals, loss)
    for n_ei in (24, 3000):  # the fused small-problem kernel and the staged pipeline
        for seed in range(6):
            v = ctx.tpe_suggest(n_ei, seed)
            assert v[0] in (1, 3, 5, 7, 9) and v[1] in (-7, -4, -1, 2, 5, 8), (n_ei, seed, v)
    r = ctx.random_fill(seed=3, first=0, n=500)
    assert set(np.unique(r[:, 0])) <= {1, 3, 5, 7, 9} and set(np.unique(r[:, 1])) <= {-7, -4, -1, 2, 5, 8}
    ctx.close()


def test_suggest_streams_its_candidates_chunk_by_chunk(monkeypatch):
    """kb_tpe_suggest never stores its candidates: each chunk is drawn straight into the staging columns.  With the
    chunk forced down to 2^16 rows a 200 000-candidate call crosses three chunk boundaries and must still return the
    winners of kb_tpe_score_select over the stored candidates of the same seed (near ties regenerate their value from
    the stream)."""
    monkeypatch.setenv("KB_CHUNK_LOG2", "16")
    ctx = kb.Context(0)  # a fresh context: its staging is sized by the first call, i.e. to 2^16 rows
    sp, vals, loss = _setup(ctx, 500, 29)
    n_ei = 200_000
    out = ctx.tpe_suggest(n_ei, seed=4242)
    cand = ctx.tpe_sample_candidates(n_ei, seed=4242)
    idx, val, sc = ctx.tpe_score_select(cand)
    assert np.array_equal(out, val), (out, val)
    ctx.tpe_set_option(2, 1)  # the pruned continuous kernel regenerates through its permutation
    assert np.array_equal(ctx.tpe_suggest(n_ei, seed=4242), val)
    ctx.close()
