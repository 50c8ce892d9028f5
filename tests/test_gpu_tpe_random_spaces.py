"""Randomised spaces and histories: every scoring variant of the library must agree with the FP64 oracle and with each
other -- plain, merged duplicates, pruned, both; the fused small-problem suggest against the staged pipeline."""
import numpy as np
import pytest

import katib_b200 as kb
from test_gpu_tpe import make_history, oracle_scores, rel_err, sample_candidates

import os

pytestmark = pytest.mark.gpu
KB_OPT_MERGE_DUPLICATES, KB_OPT_PRUNE_FAR = 1, 2
N_SEEDS = int(os.environ.get("KB_FUZZ_SEEDS", "40"))  # KB_FUZZ_SEEDS=300 for a longer hunt


def random_space(rng):
    kinds, lows, highs, steps, nch = [], [], [], [], []
    for _ in range(int(rng.integers(2, 7))):
        k = int(rng.choice([kb.KB_UNIFORM, kb.KB_INT, kb.KB_DISCRETE, kb.KB_STEP_INT, kb.KB_CATEGORICAL]))
        if k == kb.KB_UNIFORM:
            lo = float(rng.choice([-1e3, -2.5, 0.0, 1e-4, 7.0]))
            hi = lo + float(rng.choice([1e-5, 0.02, 1.0, 37.5, 4e4]))
            st, c = 0.0, 0
        elif k == kb.KB_INT:
            lo = float(rng.integers(-50, 50))
            hi = lo + float(rng.choice([1, 2, 9, 63, 127, 128, 300]))  # 128 grid points = the look-up limit; 129, 301 beyond it
            st, c = 0.0, 0
        elif k == kb.KB_DISCRETE:
            st = float(rng.choice([0.125, 0.25, 0.5, 3.0]))
            lo = float(rng.choice([-4.0, 0.0, 1.5]))
            hi = lo + st * float(rng.choice([1, 4, 40, 200]))
            c = 0
        elif k == kb.KB_STEP_INT:
            st = float(rng.choice([2, 3, 5]))
            lo = float(rng.integers(-10, 10))
            hi = lo + st * float(rng.choice([1, 3, 20])) + float(rng.choice([0, 1]))  # +1: the last step does not reach `high`
            c = 0
        else:
            lo = hi = st = 0.0
            c = int(rng.choice([2, 3, 7, 16]))
        kinds.append(k); lows.append(lo); highs.append(hi); steps.append(st); nch.append(c)
    return dict(kinds=kinds, lows=lows, highs=highs, steps=steps, nch=nch)


@pytest.mark.parametrize("seed", range(N_SEEDS))
def test_all_scoring_variants_agree_with_the_oracle(seed):
    rng = np.random.Generator(np.random.PCG64(1000 + seed))
    sp = random_space(rng)
    n_trials = int(rng.choice([0, 3, 11, 40, 260, 1500]))
    n_cand = int(rng.choice([1, 24, 700, 5000]))
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    vals = vals.reshape(n_trials, len(sp["kinds"]))
    for d, (k, lo, hi, st) in enumerate(zip(sp["kinds"], sp["lows"], sp["highs"], sp["steps"])):
        if k == kb.KB_STEP_INT:  # make_history may step past `high` when the range is not a multiple of the step
            vals[:, d] = np.minimum(vals[:, d], lo + st * np.floor((hi - lo) / st))
    loss = rng.normal(size=n_trials) + (vals[:, 0] if n_trials else 0.0) * 1e-3
    cand = sample_candidates(rng, n_cand, sp, vals)
    ref, best = oracle_scores(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"], vals, loss, cand)
    ctx = kb.Context(0)
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    ctx.tpe_set_history(vals, loss)
    for merge in (0, 1):
        for prune in (0, 1):
            ctx.tpe_set_option(KB_OPT_MERGE_DUPLICATES, merge)
            ctx.tpe_set_option(KB_OPT_PRUNE_FAR, prune)
            got = ctx.tpe_scores(cand)
            for d in range(len(sp["kinds"])):
                e = rel_err(got[:, d], ref[:, d])
                assert e < 1e-5, (sp, n_trials, n_cand, merge, prune, d, e)
            idx = ctx.tpe_score_select(cand)[0]
            assert list(idx) == best, (sp, n_trials, n_cand, merge, prune)
    ctx.close()


@pytest.mark.parametrize("seed", range(max(8, N_SEEDS // 2)))
def test_fused_and_staged_suggest_agree_on_random_spaces(seed, monkeypatch):
    rng = np.random.Generator(np.random.PCG64(2000 + seed))
    sp = random_space(rng)
    n_trials = int(rng.choice([10, 35, 300]))
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    for d, (k, lo, hi, st) in enumerate(zip(sp["kinds"], sp["lows"], sp["highs"], sp["steps"])):
        if k == kb.KB_STEP_INT:
            vals[:, d] = np.minimum(vals[:, d], lo + st * np.floor((hi - lo) / st))
    loss = rng.normal(size=n_trials)
    ctx = kb.Context(0)
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    ctx.tpe_set_option(KB_OPT_MERGE_DUPLICATES, int(seed % 2))
    ctx.tpe_set_history(vals, loss)
    for s in range(5):
        monkeypatch.setenv("KB_SUGGEST_SMALL", "1")
        a = ctx.tpe_suggest(24, 77 + s)
        monkeypatch.setenv("KB_SUGGEST_SMALL", "0")
        b = ctx.tpe_suggest(24, 77 + s)
        assert np.array_equal(a, b), (sp, n_trials, s, a, b)
        for d, (k, lo, hi) in enumerate(zip(sp["kinds"], sp["lows"], sp["highs"])):
            if k == kb.KB_CATEGORICAL:
                assert 0 <= a[d] < sp["nch"][d]
            else:
                assert lo <= a[d] <= hi, (d, a[d], lo, hi)
    ctx.close()


def test_flat_scores_grow_the_near_tie_table_instead_of_failing():
    """No history: both mixtures are the prior, every candidate scores exactly 0 and every distinct value is a near tie.
    The table (4096 slots for this many candidates) overflows, the pass is repeated with a larger one, index 0 wins."""
    rng = np.random.Generator(np.random.PCG64(9))
    ctx = kb.Context(0)
    ctx.space_set([kb.KB_UNIFORM, kb.KB_INT], [0.0, 0], [1.0, 9])
    ctx.tpe_set_history(np.zeros((0, 2)), np.zeros(0))
    cand = np.stack([rng.uniform(0, 1, 60000), np.floor(rng.uniform(0, 10, 60000))], axis=1)
    idx, val, sc = ctx.tpe_score_select(cand)
    assert list(idx) == [0, 0] and list(val) == [cand[0, 0], cand[0, 1]] and abs(sc[0]) < 1e-12
    ctx.close()
