"""The service layer without a GPU: ValidateAlgorithmSettings reproduces the reference's status codes
(service.go:173-211, the caller's reading of them: suggestionclient.go:227-245), the wire parser accepts the bytes the
reference's own generated stubs produce (tests/golden/wire_*.bin, tools/gen_wire_golden.py), and GetSuggestions fails
loudly -- codes.Internal, like every GetSuggestions failure in the reference -- when no B200 is present."""
import os

import pytest

import _pb
import katib_b200 as kb

OK, INVALID, INTERNAL = 0, 3, 13


def params4():
    return [_pb.parameter("param-1", "INT", max="10", min="-10"),
            _pb.parameter("param-2", "CATEGORICAL", list=["cat1", "cat2", "cat3"]),
            _pb.parameter("param-3", "DISCRETE", list=["3", "2", "6"]),
            _pb.parameter("param-4", "DOUBLE", max="5.5", min="-1.5")]


def vreq(algo, settings=(), params=None, objective="MINIMIZE"):
    return _pb.validate_request(_pb.experiment("test", algo, list(settings), objective, "metric-1", params or params4()))


@pytest.fixture()
def svc():
    s = kb.Service(0)
    yield s
    s.close()


@pytest.mark.parametrize("algo", ["tpe", "cmaes", "sobol", "random"])
def test_validate_accepts_the_four_algorithms(svc, algo):
    assert svc.validate(vreq(algo, [("random_state", "10")])) == (OK, "")


def test_validate_reference_wire_bytes(svc, golden_dir):
    for algo in ("tpe", "cmaes", "sobol", "random"):
        data = open(os.path.join(golden_dir, f"wire_getsuggestions_{algo}.bin"), "rb").read()
        assert svc.validate(data)[0] == OK  # field 1 = experiment in both request types


def test_validate_rejections(svc):
    assert svc.validate(None) == (INVALID, "request is empty")
    assert svc.validate(vreq("bayesianoptimization")) == (INVALID, "unsupported algorithm")
    assert svc.validate(vreq("")) == (INVALID, "unsupported algorithm")
    one_numeric = [_pb.parameter("a", "DOUBLE", max="1", min="0"), _pb.parameter("b", "CATEGORICAL", list=["x", "y"])]
    code, msg = svc.validate(vreq("cmaes", params=one_numeric))
    assert code == INVALID and msg == "CMA-ES only supports two or more dimensional continuous search space."
    assert svc.validate(vreq("tpe", params=one_numeric))[0] == OK
    dup = params4() + [_pb.parameter("param-1", "DOUBLE", max="1", min="0")]
    assert svc.validate(vreq("tpe", params=dup)) == (INVALID, "Detect duplicated parameter name: param-1")


@pytest.mark.parametrize("algo,settings,needle", [
    ("tpe", [("random_state", "ten")], "strconv.Atoi"),
    ("tpe", [("n_startup_trials", "1.5")], "strconv.Atoi"),
    ("tpe", [("n_ei_candidates", "")], "strconv.Atoi"),
    ("cmaes", [("sigma", "big")], "strconv.ParseFloat"),
    ("cmaes", [("restart_strategy", "sometimes")], "invalid restart_strategy: 'sometimes'"),
    ("random", [("random_state", "x")], "strconv.Atoi"),
])
def test_validate_bad_settings_are_internal(svc, algo, settings, needle):
    code, msg = svc.validate(vreq(algo, settings))
    assert code == INTERNAL and msg.startswith("Failed to create goptuna study and search space: ") and needle in msg


def test_validate_setting_quirks(svc):
    assert svc.validate(vreq("tpe", [("no_such_setting", "1")]))[0] == OK           # unknown names are ignored
    assert svc.validate(vreq("sobol", [("random_state", "x")]))[0] == OK            # sobol reads no setting
    assert svc.validate(vreq("cmaes", [("restart_strategy", "ipop")]))[0] == OK
    assert svc.validate(vreq("cmaes", [("restart_strategy", "bipop"), ("sigma", "0.3")]))[0] == OK
    assert svc.validate(vreq("tpe", [("random_state", "3")], objective="MAXIMIZE"))[0] == OK


def test_validate_bad_feasible_space_is_internal(svc):
    bad = [_pb.parameter("a", "DOUBLE", max="one", min="0"), _pb.parameter("b", "INT", max="3", min="1")]
    assert svc.validate(vreq("tpe", params=bad))[0] == INTERNAL
    bad = [_pb.parameter("a", "DOUBLE", max="1", min="0"), _pb.parameter("b", "INT", max="3.5", min="1")]
    assert svc.validate(vreq("tpe", params=bad))[0] == INTERNAL
    bad = [_pb.parameter("a", "DOUBLE", max="1", min="0", step="q"), _pb.parameter("b", "INT", max="3", min="1")]
    assert svc.validate(vreq("tpe", params=bad))[0] == INTERNAL


def test_get_suggestions_without_gpu_is_internal(svc, golden_dir):
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    data = open(os.path.join(golden_dir, "wire_getsuggestions_tpe.bin"), "rb").read()
    code, reply, msg = svc.get_suggestions(data)
    assert code == INTERNAL and reply == b"" and "no CPU path" in msg


def test_reply_decoder_on_reference_bytes(golden_dir):
    rb = open(os.path.join(golden_dir, "wire_reply.bin"), "rb").read()
    assert _pb.parse_reply(rb) == [[("a", "1"), ("b", "x")]]
