"""KB_OPT_MERGE_DUPLICATES: folding components with identical (mu, sigma) leaves the mixture -- hence every score and
every winner -- unchanged, while the tables of integer / stepped parameters shrink by an order of magnitude."""
import numpy as np
import pytest

import _oracle as O
import katib_b200 as kb
from test_gpu_tpe import MIXED, make_history, oracle_scores, rel_err, sample_candidates

pytestmark = pytest.mark.gpu
KB_OPT_MERGE_DUPLICATES = 1


def test_merged_tables_score_like_unmerged_and_like_the_oracle():
    rng = np.random.Generator(np.random.PCG64(321))
    sp = MIXED
    n = 4000
    vals = make_history(rng, n, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = (vals[:, 0] - 1.0) ** 2 + 0.05 * (vals[:, 1] - 3) ** 2 + (vals[:, 2] != 1) + rng.normal(0, 0.5, n)
    cand = sample_candidates(rng, 3000, sp, vals)
    ref, best = oracle_scores(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"], vals, loss, cand)
    out = {}
    for merge in (0, 1):
        ctx = kb.Context(0)
        ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
        ctx.tpe_set_option(KB_OPT_MERGE_DUPLICATES, merge)
        ctx.tpe_set_history(vals, loss)
        sizes = [len(ctx.tpe_get_parzen(d, 1)[0]) for d in range(len(sp["kinds"]))]
        out[merge] = (ctx.tpe_scores(cand), ctx.tpe_score_select(cand)[0], sizes)
        for d, k in enumerate(sp["kinds"]):
            if k != kb.KB_CATEGORICAL:
                w = ctx.tpe_get_parzen(d, 1)[0]
                assert abs(w.sum() - 1.0) < 1e-12
        ctx.close()
    s0, i0, k0 = out[0]
    s1, i1, k1 = out[1]
    for d, k in enumerate(sp["kinds"]):
        assert rel_err(s1[:, d], ref[:, d]) < 1e-5 and rel_err(s1[:, d], s0[:, d]) < 5e-6, d
    assert list(i0) == best and list(i1) == best
    # INT [-10, 10] (21 levels), DISCRETE 0.25 on [0, 16] (65 levels), STEP_INT 3 on [0, 30] (11 levels): ~3 per level
    assert k0[1] == n - 25 + 1 and k1[1] <= 3 * 21 + 2
    assert k1[3] <= 3 * 65 + 2 and k1[4] <= 3 * 11 + 2
    assert k1[1] >= 21 and k1[3] >= 65 and k1[4] >= 11
    assert k1[0] == k0[0]  # continuous values are all distinct: nothing to merge
    print("above-mixture sizes unmerged", k0, "merged", k1)


def test_service_uses_merged_tables_transparently():
    import _pb
    svc = kb.Service(0)
    space = [_pb.parameter("layers", "INT", max="5", min="2"), _pb.parameter("lr", "DOUBLE", max="0.03", min="0.01")]
    e = _pb.experiment("m", "tpe", [("random_state", "1")], "MINIMIZE", "loss", space)
    trials = []
    for r in range(40):
        code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, trials, 1, r + 1))
        assert code == 0, msg
        a = _pb.parse_reply(reply)[0]
        d = dict(a)
        f = (float(d["lr"]) - 0.02) ** 2 * 1e4 + 0.1 * int(d["layers"])
        trials.append(_pb.trial(f"t{r}", a, "SUCCEEDED", [("loss", repr(f))]))
    assert svc.stat(4) == 39
    svc.close()


@pytest.mark.parametrize("high,offgrid", [(63, False), (63, True), (199, False), (127, True)])
def test_lookup_class_on_and_off_the_grid(high, offgrid):
    """Quantised components and candidates that sit on the grid go through the look-up class of the score kernel;
    candidates off the grid (legal through kb_tpe_scores), and grids of more than 128 points, must fall back to the
    arithmetic classes.  Either way the scores are the oracle's."""
    rng = np.random.Generator(np.random.PCG64(high + offgrid))
    kinds, lows, highs, steps, nch = [kb.KB_INT, kb.KB_DISCRETE], [0, 1.0], [high, 9.0], [0, 0.125], [0, 0]
    n = 1500
    vals = make_history(rng, n, kinds, lows, highs, steps, nch)
    loss = (vals[:, 0] - 0.3 * high) ** 2 / high + (vals[:, 1] - 4) ** 2 + rng.normal(0, 1.0, n)
    sp = dict(kinds=kinds, lows=lows, highs=highs, steps=steps, nch=nch)
    cand = sample_candidates(rng, 4096, sp, vals)
    if offgrid:  # every 7th candidate a little off its grid point, so threads hold both kinds
        cand[::7, 0] += 0.25
        cand[3::7, 1] += 0.03
    ref, best = oracle_scores(kinds, lows, highs, steps, nch, vals, loss, cand)
    ctx = kb.Context(0)
    ctx.space_set(kinds, lows, highs, steps, nch)
    ctx.tpe_set_history(vals, loss)
    got = ctx.tpe_scores(cand)
    for d in range(2):
        assert rel_err(got[:, d], ref[:, d]) < 1e-5, (d, rel_err(got[:, d], ref[:, d]))
    assert list(ctx.tpe_score_select(cand)[0]) == best
    ctx.close()


KB_OPT_PRUNE_FAR = 2


@pytest.mark.parametrize("n_trials,n_cand,clustered", [(3000, 40000, False), (3000, 5000, True), (30, 3000, False),
                                                       (1, 100, False), (12000, 2100, False)])
def test_pruned_continuous_scoring_matches_unpruned(n_trials, n_cand, clustered):
    """KB_OPT_PRUNE_FAR skips (candidate, component) pairs more than 7.75 floor-sigmas apart: scores move by less than
    1e-6 relative, winners (decided in FP64 on the full mixture) not at all."""
    rng = np.random.Generator(np.random.PCG64(n_trials + n_cand))
    sp = MIXED
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    if n_trials >= 3000:  # a dense region and a sparse one: narrow and wide components both
        vals[: n_trials // 2, 0] = rng.uniform(0.0, 0.4, n_trials // 2)
    loss = (vals[:, 0] - 1.0) ** 2 + 0.05 * (vals[:, 1] - 3) ** 2 + rng.normal(0, 0.5, n_trials)
    cand = sample_candidates(rng, n_cand, sp, vals)
    if clustered:
        cand[:, 0] = 0.2 + 1e-4 * rng.uniform(size=n_cand)
        cand[:, 5] = sp["lows"][5]  # all on the lower edge
    ctx = kb.Context(0)
    ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    ctx.tpe_set_history(vals, loss)
    s0, i0 = ctx.tpe_scores(cand), ctx.tpe_score_select(cand, index_offset=7)[0]
    ctx.tpe_set_option(KB_OPT_PRUNE_FAR, 1)
    s1, i1 = ctx.tpe_scores(cand), ctx.tpe_score_select(cand, index_offset=7)[0]
    ctx.tpe_set_option(KB_OPT_PRUNE_FAR, 0)
    s2 = ctx.tpe_scores(cand)
    ctx.close()
    assert np.array_equal(s0, s2)
    for d, k in enumerate(sp["kinds"]):
        # the omitted terms are below 1e-13 of the mixture value; what differs is the FP32 summation order of the two
        # kernels (partial sums over up to 128 components): both stay within 1e-5 of the oracle, 2.5e-6 of each other
        assert rel_err(s1[:, d], s0[:, d]) < 2.5e-6, (d, rel_err(s1[:, d], s0[:, d]))
    assert np.array_equal(i0, i1)
    if n_trials <= 3000 and n_cand <= 5000:
        ref, best = oracle_scores(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"], vals, loss, cand)
        assert list(i1 - 7) == best
