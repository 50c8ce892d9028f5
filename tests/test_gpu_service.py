"""GetSuggestions through the service layer on a B200, in the shape of the reference's own tests:
pkg/suggestion/v1beta1/goptuna/service_test.go:29-187 (zero-value service, 4-parameter mixed space, random_state=10,
CurrentRequestNumber 2, no trials; asserts status code, reply count, assignment count) and the second-call pattern of
test/unit/v1beta1/suggestion/test_optuna_service.py (feed the first reply back as completed trials)."""
import os

import numpy as np
import pytest

import _pb
import katib_b200 as kb

pytestmark = pytest.mark.gpu
OK, INVALID, INTERNAL = 0, 3, 13
CHOICES2, CHOICES3 = ["cat1", "cat2", "cat3"], ["3", "2", "6"]


def params4():
    return [_pb.parameter("param-1", "INT", max="10", min="-10"),
            _pb.parameter("param-2", "CATEGORICAL", list=CHOICES2),
            _pb.parameter("param-3", "DISCRETE", list=CHOICES3),
            _pb.parameter("param-4", "DOUBLE", max="5.5", min="-1.5")]


def exp(algo, settings=(("random_state", "10"),), objective="MINIMIZE", params=None):
    return _pb.experiment("test", algo, list(settings), objective, "metric-1", params or params4())


def check_assignments(a):
    d = dict(a)
    assert set(d) == {"param-1", "param-2", "param-3", "param-4"} and len(a) == 4
    assert -10 <= int(d["param-1"]) <= 10 and str(int(d["param-1"])) == d["param-1"]
    assert d["param-2"] in CHOICES2 and d["param-3"] in CHOICES3
    v = float(d["param-4"])
    assert -1.5 <= v <= 5.5 and "e" not in d["param-4"].lower() and float(repr(v)) == v
    return d


def objective(d):
    return (float(d["param-4"]) - 2.0) ** 2 + 0.02 * (int(d["param-1"]) - 3) ** 2 + (d["param-2"] != "cat2") + 0.3 * (d["param-3"] != "6")


@pytest.mark.parametrize("algo", ["cmaes", "tpe", "random", "sobol"])
def test_reference_unit_test_shape(algo, golden_dir):
    svc = kb.Service(0)
    data = open(os.path.join(golden_dir, f"wire_getsuggestions_{algo}.bin"), "rb").read()  # bytes made by the reference's stubs
    code, reply, msg = svc.get_suggestions(data)
    assert code == OK, msg
    out = _pb.parse_reply(reply)
    assert len(out) == 2
    for a in out:
        check_assignments(a)
    svc.close()


@pytest.mark.parametrize("algo,rounds", [("tpe", 30), ("random", 6), ("sobol", 6), ("cmaes", 24)])
def test_closed_loop_feeds_results_back(algo, rounds):
    """Every call sends ALL trials, as the controller does (suggestionclient.go:111-116, 339-381)."""
    svc = kb.Service(0)
    trials, best, seen = [], [], 0
    for r in range(rounds):
        req = _pb.get_suggestions_request(exp(algo), trials, 2, 2 * (r + 1))
        code, reply, msg = svc.get_suggestions(req)
        assert code == OK, msg
        out = _pb.parse_reply(reply)
        assert len(out) == 2
        for a in out:
            d = check_assignments(a)
            f = objective(d)
            best.append(f)
            trials.append(_pb.trial(f"trial-{seen}", a, "SUCCEEDED", [("other", "1"), ("metric-1", repr(f))],
                                    "2024-01-02T03:04:05Z", "2024-01-02T03:05:05.123456789+09:00"))
            seen += 1
    if algo == "tpe":  # after the 10 start-up trials TPE should beat its own random phase on this easy objective
        assert np.mean(best[-20:]) < np.mean(best[:10])
    svc.close()


def test_states_and_errors():
    svc = kb.Service(0)
    code, reply, _ = svc.get_suggestions(_pb.get_suggestions_request(exp("tpe"), [], 3, 3))
    a = _pb.parse_reply(reply)
    assert code == OK and len(a) == 3
    t_run = _pb.trial("t-run", a[0], "RUNNING")
    t_fail = _pb.trial("t-fail", a[1], "FAILED")
    t_early = _pb.trial("t-early", a[2], "EARLYSTOPPED")
    assert svc.get_suggestions(_pb.get_suggestions_request(exp("tpe"), [t_run, t_fail, t_early], 1, 4))[0] == OK
    # a trial the service never suggested cannot be matched: Internal, as sample.go:126
    alien = _pb.trial("alien", [("param-1", "1"), ("param-2", "cat1"), ("param-3", "2"), ("param-4", "0.123")], "SUCCEEDED",
                      [("metric-1", "1.0")])
    code, _, msg = svc.get_suggestions(_pb.get_suggestions_request(exp("tpe"), [alien], 1, 5))
    assert code == INTERNAL and "Same parameter is not found for Trial" in msg
    # SUCCEEDED without the objective metric, unknown condition, bad time, bad categorical value
    base = _pb.parse_reply(svc.get_suggestions(_pb.get_suggestions_request(exp("tpe"), [], 1, 6))[1])[0]
    for bad, needle in ((_pb.trial("b1", base, "SUCCEEDED", [("other", "1")]), "No objective metric"),
                        (_pb.trial("b2", base, "KILLED"), "Unexpected Trial condition"),
                        (_pb.trial("b3", base, "RUNNING", start_time="yesterday"), "parsing time"),
                        (_pb.trial("b4", [("param-2", "cat9")], "RUNNING"), "Invalid categorical value")):
        code, _, msg = svc.get_suggestions(_pb.get_suggestions_request(exp("tpe"), [bad], 1, 7))
        assert code == INTERNAL and needle in msg, (msg, needle)
    svc.close()


def test_bad_settings_fail_get_suggestions_with_internal():
    svc = kb.Service(0)
    code, _, msg = svc.get_suggestions(_pb.get_suggestions_request(exp("tpe", [("random_state", "x")]), [], 1, 1))
    assert code == INTERNAL and msg.startswith("Failed to create goptuna study and search space: ")
    svc.close()


def test_maximize_direction_is_honoured():
    svc = kb.Service(0)
    space = [_pb.parameter("x", "DOUBLE", max="1", min="0"), _pb.parameter("y", "DOUBLE", max="1", min="0")]
    e = _pb.experiment("m", "tpe", [("random_state", "1"), ("n_startup_trials", "8")], "MAXIMIZE", "acc", space)
    trials, xs = [], []
    for r in range(40):
        code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, trials, 1, r + 1))
        assert code == OK, msg
        a = _pb.parse_reply(reply)[0]
        d = dict(a)
        x = float(d["x"])
        xs.append(x)
        trials.append(_pb.trial(f"t{r}", a, "SUCCEEDED", [("acc", repr(x))]))  # larger x is better
    assert np.mean(xs[-15:]) > np.mean(xs[:8]) + 0.1
    svc.close()


@pytest.mark.parametrize("strategy", ["ipop", "bipop"])
def test_cmaes_restart_strategies(strategy):
    """restart_strategy (converter.go:56-65): on a flat objective the tolfun criterion fires once the history of best /
    worst values spans 10 + ceil(30 n / popsize) generations; IPOP then doubles the population, BIPOP alternates regimes."""
    svc = kb.Service(0)
    space = [_pb.parameter("a", "DOUBLE", max="1", min="0"), _pb.parameter("b", "DOUBLE", max="1", min="0")]
    e = _pb.experiment("r", "cmaes", [("random_state", "3"), ("restart_strategy", strategy)], "MINIMIZE", "loss", space)
    trials = []
    pop0 = None
    for r in range(400):
        code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(e, trials, 1, r + 1))
        assert code == OK, msg
        a = _pb.parse_reply(reply)[0]
        assert all(0.0 <= float(v) <= 1.0 for _, v in a)
        trials.append(_pb.trial(f"t{r}", a, "SUCCEEDED", [("loss", "1.0")]))  # flat objective
        pop0 = pop0 or svc.stat(0)
        if svc.stat(1) >= 2:
            break
    assert pop0 == 6  # 4 + floor(3 ln 2)
    assert svc.stat(1) >= 1, "no restart happened"
    if strategy == "ipop":
        assert svc.stat(0) == pop0 * 2 ** svc.stat(1)
    svc.close()
