"""TPE parity on the GPU (through the C ABI): Parzen tables bit-exact, scores within 1e-5 relative of the
FP64 oracle, selected candidates identical.  Tolerance: |gpu - ref| <= 1e-5 * max(1, |ref|) (north_star)."""
import os

import numpy as np
import pytest

import _oracle as O
import katib_b200 as kb

pytestmark = pytest.mark.gpu
TOL = 1e-5


@pytest.fixture(scope="module")
def ctx():
    c = kb.Context(0)
    yield c
    c.close()


def rel_err(got, ref):
    return np.max(np.abs(got - ref) / np.maximum(1.0, np.abs(ref)))


def make_history(rng, n, kinds, lows, highs, steps, nch):
    cols = []
    for k, lo, hi, st, c in zip(kinds, lows, highs, steps, nch):
        if k == kb.KB_UNIFORM:
            cols.append(rng.uniform(lo, hi, n))
        elif k == kb.KB_INT:
            cols.append(np.floor(rng.uniform(lo, hi + 1, n)))
        elif k == kb.KB_STEP_INT:
            cols.append(lo + st * np.floor(rng.uniform(0, (hi - lo) / st + 1, n)))
        elif k == kb.KB_DISCRETE:
            cols.append(lo + st * np.floor(rng.uniform(0, (hi - lo) / st + 1, n)))
        else:
            cols.append(np.floor(rng.uniform(0, c, n)))
    v = np.stack(cols, axis=1)
    return v


def working_domain(k, lo, hi, st):
    """(low, high, q, shift) of sampleNumerical for a distribution kind"""
    if k == kb.KB_UNIFORM:
        return lo, hi, 0.0, 0.0
    if k == kb.KB_INT:
        return lo - 0.5, hi + 0.5, 1.0, 0.0
    return -0.5 * st, (hi - lo) + 0.5 * st, st, lo


def oracle_scores(kinds, lows, highs, steps, nch, vals, loss, cand):
    out = np.zeros_like(cand)
    best = []
    for d, (k, lo, hi, st, c) in enumerate(zip(kinds, lows, highs, steps, nch)):
        if k == kb.KB_CATEGORICAL:
            sc, b = O.tpe_score_categorical(cand[:, d], vals[:, d], loss, c)
        else:
            wl, wh, q, sh = working_domain(k, lo, hi, st)
            sc, b = O.tpe_score_numeric(cand[:, d], vals[:, d] - sh, loss, wl, wh, q)
        out[:, d] = sc
        best.append(b)
    return out, best


MIXED = dict(
    kinds=[kb.KB_UNIFORM, kb.KB_INT, kb.KB_CATEGORICAL, kb.KB_DISCRETE, kb.KB_STEP_INT, kb.KB_UNIFORM, kb.KB_CATEGORICAL],
    lows=[-1.5, -10, 0, 0.0, 0, 1e-3, 0], highs=[5.5, 10, 0, 16.0, 30, 3e-3, 0], steps=[0, 0, 0, 0.25, 3, 0, 0],
    nch=[0, 0, 3, 0, 0, 0, 12])


def sample_candidates(rng, n, sp, vals):
    """working-domain candidates: continuous uniform, quantised on the grid (shifted for DISCRETE), categorical index"""
    cols = []
    for d, (k, lo, hi, st, c) in enumerate(zip(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])):
        if k == kb.KB_UNIFORM:
            cols.append(rng.uniform(lo, hi, n))
        elif k == kb.KB_CATEGORICAL:
            cols.append(np.floor(rng.uniform(0, c, n)))
        else:
            wl, wh, q, sh = working_domain(k, lo, hi, st)
            x = rng.uniform(wl, wh, n)
            cols.append(np.round(x / q) * q)
    return np.stack(cols, axis=1)


@pytest.mark.parametrize("n_trials", [0, 1, 9, 10, 24, 25, 26, 100, 1000, 3000])
def test_parzen_tables_bit_exact(ctx, n_trials):
    rng = np.random.Generator(np.random.PCG64(100 + n_trials))
    sp = MIXED
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = rng.normal(size=n_trials)
    if n_trials >= 26:
        loss[5] = loss[17]  # a tie in the loss: stable order decides
    ctx.tpe_set_history(vals.reshape(n_trials, len(sp["kinds"])), loss)
    for d, (k, lo, hi, st, c) in enumerate(zip(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])):
        if k == kb.KB_CATEGORICAL:
            pl, pg = O.categorical_tables(vals[:, d], loss, c)
            assert np.array_equal(ctx.tpe_get_parzen(d, 0)[0], pl)
            assert np.array_equal(ctx.tpe_get_parzen(d, 1)[0], pg)
            continue
        wl, wh, q, sh = working_domain(k, lo, hi, st)
        below, above = O.tpe_split(vals[:, d] - sh, loss)
        for side, obs in ((0, below), (1, above)):
            w, mu, sg = O.parzen_build(obs, wl, wh)
            gw, gmu, gsg, pa = ctx.tpe_get_parzen(d, side)
            assert np.array_equal(gmu, mu) and np.array_equal(gsg, sg) and np.array_equal(gw, w), (d, side)
            ref_pa = sum(w[j] * (O.lib().kbo_normal_cdf(wh, mu[j], sg[j]) - O.lib().kbo_normal_cdf(wl, mu[j], sg[j]))
                         for j in range(len(w)))
            assert abs(pa - ref_pa) <= 1e-14 * abs(ref_pa)


@pytest.mark.parametrize("n_trials,n_cand", [(10, 24), (100, 24), (100, 1), (300, 5000), (1000, 2049), (2500, 777)])
def test_scores_and_winners_mixed(ctx, n_trials, n_cand):
    rng = np.random.Generator(np.random.PCG64(7 * n_trials + n_cand))
    sp = MIXED
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = (vals[:, 0] - 1.0) ** 2 + 0.05 * (vals[:, 1] - 3) ** 2 + (vals[:, 2] != 1) + rng.normal(0, 0.5, n_trials)
    ctx.tpe_set_history(vals, loss)
    cand = sample_candidates(rng, n_cand, sp, vals)
    ref, best = oracle_scores(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"], vals, loss, cand)
    got = ctx.tpe_scores(cand)
    for d in range(len(sp["kinds"])):
        assert rel_err(got[:, d], ref[:, d]) < TOL, (d, rel_err(got[:, d], ref[:, d]))
    idx, val, sc = ctx.tpe_score_select(cand, index_offset=1000)
    assert list(idx - 1000) == best
    for d, k in enumerate(sp["kinds"]):
        x = cand[best[d], d]
        if k == kb.KB_DISCRETE:
            x = min(max(x + sp["lows"][d], sp["lows"][d]), sp["highs"][d])
        assert val[d] == x
        assert abs(sc[d] - ref[best[d], d]) <= 1e-9 * max(1.0, abs(ref[best[d], d]))


def test_config3_shape_reduced(ctx):
    """BASELINE configs[2] shape (20-dim continuous, 1k trials) at 32768 candidates -- what the oracle finishes fast."""
    rng = np.random.Generator(np.random.PCG64(3))
    D, n, nc = 20, 1000, 32768
    s = np.linspace(0.5, 4.0, D)
    lows, highs = -5 * s, 5 * s
    ctx.space_set([kb.KB_UNIFORM] * D, lows, highs)
    vals = rng.uniform(lows, highs, (n, D))
    loss = np.sum((vals / s) ** 2, axis=1) + rng.normal(0, 1, n)
    ctx.tpe_set_history(vals, loss)
    cand = rng.uniform(lows, highs, (nc, D))
    got = ctx.tpe_scores(cand)
    idx, val, sc = ctx.tpe_score_select(cand)
    worst = 0.0
    for d in range(D):
        ref, b = O.tpe_score_numeric(cand[:, d], vals[:, d], loss, lows[d], highs[d])
        worst = max(worst, rel_err(got[:, d], ref))
        assert idx[d] == b and val[d] == cand[b, d]
    assert worst < TOL, worst
    print("config3-shape max rel err", worst)


def test_duplicates_first_index_wins(ctx):
    rng = np.random.Generator(np.random.PCG64(11))
    ctx.space_set([kb.KB_UNIFORM, kb.KB_INT], [0.0, 0], [1.0, 7])
    vals = np.stack([rng.uniform(0, 1, 200), np.floor(rng.uniform(0, 8, 200))], axis=1)
    loss = (vals[:, 0] - 0.3) ** 2 + 0.01 * vals[:, 1]
    ctx.tpe_set_history(vals, loss)
    base = np.stack([rng.uniform(0, 1, 500), np.floor(rng.uniform(0, 8, 500))], axis=1)
    cand = np.concatenate([base, base[::-1], base])  # every value three times
    r0, b0 = O.tpe_score_numeric(cand[:, 0], vals[:, 0], loss, 0.0, 1.0)
    r1, b1 = O.tpe_score_numeric(cand[:, 1], vals[:, 1], loss, -0.5, 7.5, 1.0)
    idx, val, sc = ctx.tpe_score_select(cand)
    assert list(idx) == [b0, b1]


def test_call_order_and_errors(ctx):
    c2 = kb.Context(0)
    with pytest.raises(kb.KbError):
        c2.tpe_set_history(np.zeros((3, 2)), np.zeros(3))  # no space yet: KB_ERR_STATE
    c2.space_set([kb.KB_UNIFORM], [0.0], [1.0])
    with pytest.raises(kb.KbError):
        c2.tpe_scores(np.zeros((4, 1)))  # no history yet
    with pytest.raises(kb.KbError):
        c2.space_set([kb.KB_DISCRETE], [0.0], [1.0], [0.0])  # step must be > 0
    with pytest.raises(kb.KbError):
        c2.space_set([kb.KB_UNIFORM], [2.0], [1.0])
    c2.close()


def test_config5_shape_reduced(ctx):
    """BASELINE configs[4] shape (50-dim mixed, 10k trials) at 4096 candidates; prints the worst relative error."""
    import bench
    kinds, lows, highs, steps, nch = bench.workload_space()
    vals, loss = bench.workload_history(kinds, lows, highs, steps, nch)
    ctx.space_set(kinds, lows, highs, steps, nch)
    ctx.tpe_set_history(vals, loss)
    rng = np.random.Generator(np.random.PCG64(55))
    sp = dict(kinds=kinds, lows=lows, highs=highs, steps=steps, nch=nch)
    cand = sample_candidates(rng, 4096, sp, vals)
    ref, best = oracle_scores(kinds, lows, highs, steps, nch, vals, loss, cand)
    got = ctx.tpe_scores(cand)
    errs = np.array([rel_err(got[:, d], ref[:, d]) for d in range(len(kinds))])
    print("config5-shape max rel err: cont %.2e quant %.2e cat %.2e" % (errs[:20].max(), errs[20:30].max(), errs[30:].max()))
    assert errs.max() < TOL, errs
    idx, val, sc = ctx.tpe_score_select(cand)
    assert list(idx) == best


@pytest.mark.slow
@pytest.mark.skipif(os.environ.get("KB_SLOW") != "1", reason="about two minutes of CPU-oracle time on 24 cores: KB_SLOW=1")
def test_config5_shape_1Mi_rows():
    """the 1 Mi-row subset of BASELINE configs[4] against the oracle -- every score of all 50
    parameters within 1e-5, every winner identical.  The oracle runs on all host threads (the C port releases the GIL).
    The same rows then go through kb_tpe_score_error_device: the FP64 side computed on the GPU must tell the same story."""
    import threading
    import bench
    kinds, lows, highs, steps, nch = bench.workload_space()
    vals, loss = bench.workload_history(kinds, lows, highs, steps, nch)
    ctx = kb.Context(0)
    ctx.space_set(kinds, lows, highs, steps, nch)
    ctx.tpe_set_history(vals, loss)
    rng = np.random.Generator(np.random.PCG64(55))
    sp = dict(kinds=kinds, lows=lows, highs=highs, steps=steps, nch=nch)
    n = 1 << 20
    cand = sample_candidates(rng, n, sp, vals)
    got = ctx.tpe_scores(cand)
    idx = ctx.tpe_score_select(cand)[0]
    parts = np.array_split(np.arange(n), os.cpu_count() or 8)
    ref = np.empty_like(cand)

    def work(p):
        ref[p] = oracle_scores(kinds, lows, highs, steps, nch, vals, loss, cand[p])[0]
    ths = [threading.Thread(target=work, args=(p,)) for p in parts]
    [t.start() for t in ths]
    [t.join() for t in ths]
    errs = np.array([rel_err(got[:, d], ref[:, d]) for d in range(len(kinds))])
    print("config5-shape, 2^20 rows: max rel err cont %.2e quant %.2e cat %.2e" % (errs[:20].max(), errs[20:30].max(), errs[30:].max()))
    assert errs.max() < TOL, errs
    assert list(idx) == [int(np.argmax(ref[:, d])) for d in range(len(kinds))]
    import torch
    t = torch.from_numpy(cand).to("cuda:0")
    dev_err = ctx.tpe_score_error_device(n, t.data_ptr())
    print("same rows, FP64 side on the device: cont %.2e quant %.2e" % (dev_err[:20].max(), dev_err[20:30].max()))
    assert dev_err.max() < TOL and np.all(np.abs(dev_err[:30] - errs[:30]) < 2e-7)
    ctx.close()


def test_multi_chunk_host_path_and_offsets(monkeypatch):
    """2.5 Mi candidates cross the chunk boundary three times (double-buffered H2D; chunks of 2^20 rows forced on a
    fresh context, the default is 2^22): winners and indices vs the oracle."""
    monkeypatch.setenv("KB_CHUNK_LOG2", "20")
    ctx = kb.Context(0)
    rng = np.random.Generator(np.random.PCG64(8))
    ctx.space_set([kb.KB_UNIFORM, kb.KB_INT, kb.KB_CATEGORICAL], [0.0, 0, 0], [1.0, 9, 0], None, [0, 0, 4])
    n = 60
    vals = np.stack([rng.uniform(0, 1, n), np.floor(rng.uniform(0, 10, n)), np.floor(rng.uniform(0, 4, n))], axis=1)
    loss = (vals[:, 0] - 0.7) ** 2 + 0.01 * vals[:, 1] + 0.1 * (vals[:, 2] != 2) + rng.normal(0, 0.05, n)
    ctx.tpe_set_history(vals, loss)
    nc = (5 << 19) + 12345
    cand = np.stack([rng.uniform(0, 1, nc), np.floor(rng.uniform(0, 10, nc)), np.floor(rng.uniform(0, 4, nc))], axis=1)
    idx, val, sc = ctx.tpe_score_select(cand, index_offset=7_000_000_000)
    _, b0 = O.tpe_score_numeric(cand[:, 0], vals[:, 0], loss, 0.0, 1.0)
    _, b1 = O.tpe_score_numeric(cand[:, 1], vals[:, 1], loss, -0.5, 9.5, 1.0)
    _, b2 = O.tpe_score_categorical(cand[:, 2], vals[:, 2], loss, 4)
    assert list(idx - 7_000_000_000) == [b0, b1, b2]
    assert list(val) == [cand[b0, 0], cand[b1, 1], cand[b2, 2]]
    ctx.tpe_set_option(2, 1)  # the pruned path across the same chunk boundaries
    idx2, val2, _ = ctx.tpe_score_select(cand, index_offset=7_000_000_000)
    assert list(idx2) == list(idx) and list(val2) == list(val)
    ctx.close()


def test_edges_empty_history_empty_candidates_wide_space(ctx):
    # no history at all: both estimators are the prior alone -> every score is 0 -> first candidate wins
    ctx.space_set([kb.KB_UNIFORM, kb.KB_CATEGORICAL], [0.0, 0], [1.0, 0], None, [0, 3])
    ctx.tpe_set_history(np.zeros((0, 2)), np.zeros(0))
    cand = np.array([[0.3, 2.0], [0.9, 0.0], [0.1, 1.0]])
    sc = ctx.tpe_scores(cand)
    assert np.allclose(sc, 0.0, atol=1e-6)
    assert list(ctx.tpe_score_select(cand)[0]) == [0, 0]
    idx, val, _ = ctx.tpe_score_select(np.zeros((0, 2)))
    assert list(idx) == [-1, -1]
    # 300 parameters, 4096-way categorical, long history (tables stream through several shared-memory stages)
    rng = np.random.Generator(np.random.PCG64(77))
    D = 300
    kinds = [kb.KB_UNIFORM] * 298 + [kb.KB_CATEGORICAL, kb.KB_STEP_INT]
    lows = [-1.0] * 298 + [0.0, -50.0]
    highs = [2.0] * 298 + [0.0, 50.0]
    steps = [0.0] * 298 + [0.0, 5.0]
    nch = [0] * 298 + [4096, 0]
    ctx.space_set(kinds, lows, highs, steps, nch)
    n = 5000
    vals = rng.uniform(-1, 2, (n, D))
    vals[:, 298] = np.floor(rng.uniform(0, 4096, n))
    vals[:, 299] = -50 + 5 * np.floor(rng.uniform(0, 21, n))
    loss = np.sum(vals[:, :5] ** 2, axis=1) + rng.normal(0, 0.1, n)
    ctx.tpe_set_history(vals, loss)
    nc = 700
    cand = rng.uniform(-1, 2, (nc, D))
    cand[:, 298] = np.floor(rng.uniform(0, 4096, nc))
    cand[:, 299] = 5 * np.floor(rng.uniform(0, 21, nc))  # working domain of STEP_INT: shifted by -low
    got = ctx.tpe_scores(cand)
    idx = ctx.tpe_score_select(cand)[0]
    for d in (0, 150, 297):
        ref, b = O.tpe_score_numeric(cand[:, d], vals[:, d], loss, -1.0, 2.0)
        assert rel_err(got[:, d], ref) < TOL and idx[d] == b
    ref, b = O.tpe_score_categorical(cand[:, 298], vals[:, 298], loss, 4096)
    assert rel_err(got[:, 298], ref) < 1e-12 and idx[298] == b
    ref, b = O.tpe_score_numeric(cand[:, 299], vals[:, 299] + 50.0, loss, -2.5, 102.5, 5.0)
    assert rel_err(got[:, 299], ref) < TOL and idx[299] == b
