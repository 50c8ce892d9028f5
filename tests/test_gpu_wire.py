"""Wire-level drop-in check: a gRPC listener forwards raw bytes to kb_service_* (tools/serve_suggestion.py), and a
client speaks to it the way the katib-controller does (suggestionclient.go:93-128, 201-246): insecure channel, unary
methods /api.v1.beta1.Suggestion/GetSuggestions and /ValidateAlgorithmSettings, plus the health probe.
Requests are the bytes produced by the reference's generated stubs (tests/golden/wire_*.bin)."""
import os
import sys

import grpc
import pytest

import _pb

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tools"))


@pytest.fixture()
def endpoint():
    import serve_suggestion
    server, port, svc = serve_suggestion.make_server(0, 0, host="127.0.0.1")
    server.start()
    ch = grpc.insecure_channel(f"127.0.0.1:{port}")
    yield ch
    ch.close()
    server.stop(0)
    svc.close()


def test_controller_style_round_trip(endpoint, golden_dir):
    ident = lambda b: b
    get = endpoint.unary_unary("/api.v1.beta1.Suggestion/GetSuggestions", request_serializer=ident, response_deserializer=ident)
    val = endpoint.unary_unary("/api.v1.beta1.Suggestion/ValidateAlgorithmSettings", request_serializer=ident, response_deserializer=ident)
    health = endpoint.unary_unary("/grpc.health.v1.Health/Check", request_serializer=ident, response_deserializer=ident)
    assert health(b"\x0a\x1bmanager.v1beta1.Suggestion", timeout=60) == b"\x08\x01"
    req = open(os.path.join(golden_dir, "wire_getsuggestions_tpe.bin"), "rb").read()
    assert val(req, timeout=60) == b""
    out = _pb.parse_reply(get(req, timeout=60))
    assert len(out) == 2 and all(len(a) == 4 for a in out)  # CurrentRequestNumber = 2, four parameters
    bad = _pb.validate_request(_pb.experiment("x", "gridsearch", [], "MINIMIZE", "m", []))
    with pytest.raises(grpc.RpcError) as e:
        val(bad, timeout=60)
    assert e.value.code() == grpc.StatusCode.INVALID_ARGUMENT and e.value.details() == "unsupported algorithm"
    alien = _pb.trial("alien", [("param-1", "1"), ("param-2", "cat1"), ("param-3", "2"), ("param-4", "0.5")], "SUCCEEDED", [("metric-1", "1")])
    exp = _pb.experiment("test", "tpe", [("random_state", "10")], "MINIMIZE", "metric-1",
                         [_pb.parameter("param-1", "INT", max="10", min="-10"), _pb.parameter("param-2", "CATEGORICAL", list=["cat1", "cat2", "cat3"]),
                          _pb.parameter("param-3", "DISCRETE", list=["3", "2", "6"]), _pb.parameter("param-4", "DOUBLE", max="5.5", min="-1.5")])
    with pytest.raises(grpc.RpcError) as e:
        get(_pb.get_suggestions_request(exp, [alien], 1, 3), timeout=60)
    assert e.value.code() == grpc.StatusCode.INTERNAL
