"""kb_service_create_multi: ONE suggestion service over several GPUs (the katib-controller talks to one service per
experiment: composer.go:72-146, main.go:43-57).  A TPE experiment with a large n_ei_candidates setting must get the
same replies from a two-GPU service as from a one-GPU service -- in-process through the C ABI and over the wire
through tools/serve_suggestion.py --devices 0,1 -- and small requests, CMA-ES, Sobol and random keep working on it."""
import os
import sys

import pytest

import _pb
import katib_b200 as kb

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))

SPACE = [_pb.parameter("lr", "DOUBLE", max="0.03", min="0.01"), _pb.parameter("layers", "INT", max="5", min="2"),
         _pb.parameter("opt", "CATEGORICAL", list=["sgd", "adam", "ftrl"]), _pb.parameter("width", "INT", max="512", min="32", step="32"),
         _pb.parameter("drop", "DOUBLE", max="0.5", min="0.0", step="0.05")]


def _need_two_gpus():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")


def _objective(a):
    d = dict(a)
    return ((float(d["lr"]) - 0.02) ** 2 * 1e4 + 0.1 * int(d["layers"]) + (d["opt"] != "adam") + abs(int(d["width"]) - 256) / 256
            + (float(d["drop"]) - 0.2) ** 2)


def _closed_loop(call, algorithm, settings, rounds):
    """the controller's loop: every request carries all trials so far; returns the list of replies"""
    exp = _pb.experiment("multi", algorithm, settings, "MINIMIZE", "loss", SPACE)
    trials, replies = [], []
    for r in range(rounds):
        reply = call(_pb.get_suggestions_request(exp, trials, 1, r + 1))
        a = _pb.parse_reply(reply)[0]
        replies.append(sorted(a))
        trials.append(_pb.trial(f"t{r}", a, "SUCCEEDED", [("loss", repr(_objective(a)))]))
    return replies


def _svc_call(svc):
    def call(req):
        code, reply, msg = svc.get_suggestions(req)
        assert code == 0, msg
        return reply
    return call


@pytest.mark.parametrize("n_ei", ["2000000", "24"])
def test_two_gpu_service_answers_like_one_gpu_service(n_ei):
    _need_two_gpus()
    settings = [("random_state", "10"), ("n_ei_candidates", n_ei), ("n_startup_trials", "5")]
    one, two = kb.Service(0), kb.Service(devices=[0, 1])
    try:
        rounds = 9 if n_ei == "2000000" else 14
        a = _closed_loop(_svc_call(one), "tpe", settings, rounds)
        b = _closed_loop(_svc_call(two), "tpe", settings, rounds)
        assert a == b
    finally:
        one.close(); two.close()


@pytest.mark.parametrize("algorithm", ["cmaes", "sobol", "random"])
def test_other_algorithms_run_on_the_first_device_of_a_multi_service(algorithm):
    _need_two_gpus()
    one, two = kb.Service(0), kb.Service(devices=[0, 1])
    try:
        settings = [("random_state", "3")]
        assert _closed_loop(_svc_call(one), algorithm, settings, 12) == _closed_loop(_svc_call(two), algorithm, settings, 12)
    finally:
        one.close(); two.close()


def test_wire_level_multi_device_listener():
    _need_two_gpus()
    import grpc
    import serve_suggestion
    settings = [("random_state", "10"), ("n_ei_candidates", "2000000"), ("n_startup_trials", "5")]
    out = []
    for devices in ([0], [0, 1]):
        server, port, svc = serve_suggestion.make_server(0, 0, host="127.0.0.1", devices=devices)
        server.start()
        ch = grpc.insecure_channel(f"127.0.0.1:{port}", options=[("grpc.max_receive_message_length", 64 << 20)])
        ident = lambda b: b
        get = ch.unary_unary("/api.v1.beta1.Suggestion/GetSuggestions", request_serializer=ident, response_deserializer=ident)
        out.append(_closed_loop(lambda req: get(req, timeout=120), "tpe", settings, 8))
        ch.close(); server.stop(0); svc.close()
    assert out[0] == out[1]


def test_create_multi_rejects_bad_device_lists():
    import ctypes as C
    h = C.c_void_p()
    arr = (C.c_int * 2)(0, 0)
    assert kb.lib().kb_service_create_multi(2, arr, C.byref(h)) == -1 and "twice" in kb.lib().kb_last_error(None).decode()
    assert kb.lib().kb_service_create_multi(0, arr, C.byref(h)) == -1
