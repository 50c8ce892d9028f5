"""History handling either side of the TPE kernels: incremental upload (kb_tpe_append_history), the trial-matching
index that replaces the reference's O(N^2) cold sync (sample.go:97-127), and the opt-in adoption of trials the service
never suggested."""
import time

import numpy as np
import pytest

import _pb
import katib_b200 as kb
from test_gpu_tpe import MIXED, make_history

pytestmark = pytest.mark.gpu
OK, INTERNAL = 0, 13
KB_SVC_ADOPT_UNKNOWN_TRIALS = 1


def tables(ctx, sp):
    out = []
    for d, k in enumerate(sp["kinds"]):
        for side in (0, 1):
            out.append(ctx.tpe_get_parzen(d, side))
    return out


@pytest.mark.parametrize("cuts", [(0, 7, 8, 30, 31, 400), (50, 51, 3000), (10, 10, 2000)])
def test_append_equals_set_of_the_concatenation(cuts):
    rng = np.random.Generator(np.random.PCG64(sum(cuts)))
    sp = MIXED
    n = cuts[-1]
    vals = make_history(rng, n, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    loss = rng.normal(size=n)
    a, b = kb.Context(0), kb.Context(0)
    for c in (a, b):
        c.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"])
    with pytest.raises(kb.KbError):
        a.tpe_append_history(vals[:1], loss[:1])  # nothing to append to
    a.tpe_set_history(vals[:cuts[0]], loss[:cuts[0]])
    for lo, hi in zip(cuts[:-1], cuts[1:]):
        a.tpe_append_history(vals[lo:hi], loss[lo:hi])  # grows through several reallocations; (10, 10) appends nothing
        b.tpe_set_history(vals[:hi], loss[:hi])
        for ta, tb in zip(tables(a, sp), tables(b, sp)):
            for x, y in zip(ta, tb):
                assert np.array_equal(x, y)
    cand = vals[:64].copy()
    for d, k in enumerate(sp["kinds"]):
        if k == kb.KB_DISCRETE:
            cand[:, d] -= sp["lows"][d]
    assert np.array_equal(a.tpe_scores(cand), b.tpe_scores(cand))
    a.close(); b.close()


def space2():
    return [_pb.parameter("x", "DOUBLE", max="1.0", min="0.0"), _pb.parameter("k", "INT", max="6", min="1"),
            _pb.parameter("c", "CATEGORICAL", list=["a", "b"])]


def run_loop(svc, algo, rounds, per_round=2, straggler=False):
    """Closed loop: suggestions of round r are Running in request r + 1 and Succeeded in request r + 2; with
    `straggler` the first trial of every round finishes one request later than the second (a smaller trial id
    arriving behind a larger one)."""
    e = _pb.experiment("h", algo, [("random_state", "3")], "MINIMIZE", "loss", space2())
    done, running = [], []  # running: [request index at which it finishes, running msg, finished msg]
    for r in range(rounds):
        now = [p for p in running if p[0] <= r]
        running = [p for p in running if p[0] > r]
        done += [p[2] for p in now]
        code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, done + [p[1] for p in running], per_round, 0))
        assert code == OK, msg
        for i, a in enumerate(_pb.parse_reply(reply)):
            d = dict(a)
            assert 0.0 <= float(d["x"]) <= 1.0 and 1 <= int(d["k"]) <= 6 and d["c"] in ("a", "b")
            f = (float(d["x"]) - 0.3) ** 2 + 0.1 * int(d["k"]) + (d["c"] == "b")
            name = f"t{r}-{i}"
            running.append([r + 2 + (1 if straggler and i == 0 else 0), _pb.trial(name, a, "RUNNING"),
                            _pb.trial(name, a, "SUCCEEDED", [("loss", repr(f))])])
    return e, done


def test_only_new_rows_cross_the_boundary():
    svc = kb.Service(0)
    run_loop(svc, "tpe", 40)
    done = svc.stat(4)
    assert done == 2 * 38  # results of round r arrive with the request of round r + 2
    # rows sent: the first upload happens when TPE leaves its start-up phase (10 observations), appends after that
    assert svc.stat(5) == done, (svc.stat(5), done)
    svc.close()


def test_stragglers_force_a_resend_in_trial_order():
    """History rows stay in trial-id order (the weight ramp depends on it), so a late smaller id cannot be appended:
    everything is sent again.  Rows sent then exceed the observations; without stragglers they are equal."""
    svc = kb.Service(0)
    run_loop(svc, "tpe", 30, straggler=True)
    assert svc.stat(4) == 2 * 28 - 1
    assert svc.stat(5) > 4 * svc.stat(4), (svc.stat(5), svc.stat(4))
    svc.close()


def test_unknown_trials_fail_like_the_reference_unless_adopted():
    first = kb.Service(0)
    e, trials = run_loop(first, "tpe", 12)
    first.close()
    fresh = kb.Service(0)  # the pod restarted: the in-memory study is gone (service.go:39-52)
    code, reply, msg = fresh.get_suggestions(_pb.get_suggestions_request(e, trials, 1, 0))
    assert code == INTERNAL and msg.startswith("Same parameter is not found for Trial: t")
    fresh.close()
    fresh = kb.Service(0)
    fresh.set_option(KB_SVC_ADOPT_UNKNOWN_TRIALS, 1)
    code, reply, msg = fresh.get_suggestions(_pb.get_suggestions_request(e, trials, 1, 0))
    assert code == OK, msg
    assert fresh.stat(6) == len(trials) and fresh.stat(4) == len(trials) and fresh.stat(2) == len(trials) + 1
    assert fresh.stat(5) == len(trials)  # >= 10 observations: the adopted history drives TPE at once
    assert len(_pb.parse_reply(reply)) == 1
    with pytest.raises(kb.KbError):
        fresh.set_option(99, 1)
    fresh.close()


def test_cold_sync_of_thousands_of_trials_is_not_quadratic():
    svc = kb.Service(0)
    n = 4000
    e = _pb.experiment("cold", "random", [("random_state", "1")], "MINIMIZE", "loss", space2())
    code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, [], n, n))
    assert code == OK, msg
    trials = [_pb.trial(f"t{i}", a, "SUCCEEDED", [("loss", "1.0")]) for i, a in enumerate(_pb.parse_reply(reply))]
    t0 = time.perf_counter()
    code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, trials, 1, n + 1))
    dt = time.perf_counter() - t0
    assert code == OK, msg
    assert svc.stat(4) == n
    assert dt < 2.0, dt  # parse + convert + match 4000 trials; the quadratic scan took ~16 M map comparisons
    print(f"cold sync of {n} trials: {dt * 1e3:.1f} ms")
    svc.close()


def test_settled_trials_are_recognised_by_their_bytes():
    """Every request resends every trial; a Trial message already synced into a finished state is skipped by digest.
    A finished trial resent with DIFFERENT bytes is parsed again -- and, as in the reference, changes nothing."""
    svc = kb.Service(0)
    e, done = run_loop(svc, "tpe", 20)
    assert svc.stat(7) > 10 * len(done) // 4          # most of what was resent over 20 requests was skipped
    before_obs, before_skipped = svc.stat(4), svc.stat(7)
    code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, done, 0, 0))
    assert code == OK and svc.stat(7) == before_skipped + len(done)
    # the same trial names, now claiming to have failed and carrying another loss: parsed, but finished trials stay as they are
    changed = []
    for r in range(18):
        for i in range(2):
            changed.append(_pb.trial(f"t{r}-{i}", [("x", "0.5"), ("k", "3"), ("c", "a")], "FAILED", [("loss", "123.0")]))
    code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, changed, 1, 0))
    assert code == OK, msg
    assert svc.stat(7) == before_skipped + len(done) and svc.stat(4) == before_obs
    # a malformed value in a message never seen before still fails the whole request, as toGoptunaTrials does
    bad = _pb.trial("t0-0", [("x", "zero point five"), ("k", "3"), ("c", "a")], "SUCCEEDED", [("loss", "1.0")])
    code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, done + [bad], 1, 0))
    assert code == INTERNAL
    svc.close()
