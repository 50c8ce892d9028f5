#!/usr/bin/env python
"""A stand-in for cmd/suggestion/goptuna/v1beta1/main.go while no Go toolchain is at hand: listens on 0.0.0.0:6789
(main.go:31), registers api.v1.beta1.Suggestion with its two unary methods (api.pb.go:1551, 1569) and
grpc.health.v1.Health/Check -> SERVING (main.go:37-41), and forwards the request BYTES to kb_service_* untouched.
Nothing is parsed or computed here: the protobuf handling and all arithmetic live in libkatib_b200.so.
The production front end is the Go shim of INTEGRATION.md; this file exists for wire-level tests and demos.

    python tools/serve_suggestion.py [--port 6789] [--device 0 | --devices 0,1,2,3]
"""
import argparse
import os
import sys
from concurrent import futures

import grpc

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import katib_b200 as kb  # noqa: E402

CODES = {3: grpc.StatusCode.INVALID_ARGUMENT, 13: grpc.StatusCode.INTERNAL}
SERVING = b"\x08\x01"  # HealthCheckResponse{status: SERVING}
MAX_MSG = 256 << 20    # the reference runs with grpc-go's 4 MiB default; a 10k-trial request needs more


def make_server(port, device=0, host="0.0.0.0", devices=None):
    svc = kb.Service(device, devices=devices)  # several devices: kb_service_create_multi (one service, sharded TPE)

    def get_suggestions(request, context):
        code, reply, msg = svc.get_suggestions(request)
        if code:
            context.abort(CODES.get(code, grpc.StatusCode.UNKNOWN), msg)
        return reply

    def validate(request, context):
        code, msg = svc.validate(request)
        if code:
            context.abort(CODES.get(code, grpc.StatusCode.UNKNOWN), msg)
        return b""

    ident = lambda b: b
    handlers = [
        grpc.method_handlers_generic_handler("api.v1.beta1.Suggestion", {
            "GetSuggestions": grpc.unary_unary_rpc_method_handler(get_suggestions, ident, ident),
            "ValidateAlgorithmSettings": grpc.unary_unary_rpc_method_handler(validate, ident, ident)}),
        grpc.method_handlers_generic_handler("grpc.health.v1.Health", {
            "Check": grpc.unary_unary_rpc_method_handler(lambda r, c: SERVING, ident, ident)}),
    ]
    server = grpc.server(futures.ThreadPoolExecutor(max_workers=4), handlers=handlers,
                         options=[("grpc.max_receive_message_length", MAX_MSG), ("grpc.max_send_message_length", MAX_MSG)])
    bound = server.add_insecure_port(f"{host}:{port}")
    return server, bound, svc


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--port", type=int, default=6789)
    ap.add_argument("--device", type=int, default=0)
    ap.add_argument("--devices", default="", help="comma-separated device list, e.g. 0,1 (overrides --device)")
    a = ap.parse_args()
    devs = [int(x) for x in a.devices.split(",") if x.strip()] or None
    server, port, _ = make_server(a.port, a.device, devices=devs)
    server.start()
    print(f"katib-b200 suggestion service listening on :{port}", flush=True)
    server.wait_for_termination()
