"""The Go shim, its Dockerfile and the katib-config entries cannot be compiled or applied here (no Go toolchain, no
cluster).  What can be checked on any machine: every C symbol the Go file calls is declared in include/katib_b200.h and
exported by the library, the gRPC surface it implements is the reference's (method names of api.proto), the image is a
cgo build on a CUDA base, and the config entries name the four algorithms the service accepts (service.go:181-184)."""
import os
import re
import subprocess

import yaml

import katib_b200 as kb

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GO_SERVICE = os.path.join(ROOT, "go", "pkg", "suggestion", "v1beta1", "goptuna", "service.go")
GO_MAIN = os.path.join(ROOT, "go", "cmd", "suggestion", "goptuna", "v1beta1", "main.go")
DOCKERFILE = os.path.join(ROOT, "go", "cmd", "suggestion", "goptuna", "v1beta1", "Dockerfile")
CONFIG = os.path.join(ROOT, "manifests", "katib-config-b200.yaml")


def test_go_service_calls_only_declared_and_exported_symbols():
    src = open(GO_SERVICE).read()
    called = sorted(set(re.findall(r"\bC\.(kb_[a-z0-9_]+)\(", src)))
    assert {"kb_service_create", "kb_service_create_multi", "kb_service_get_suggestions",
            "kb_service_validate_algorithm_settings", "kb_service_last_message", "kb_service_destroy"} <= set(called)
    declared = set(kb.header_symbols())
    assert not [c for c in called if c not in declared], called
    kb.build()
    out = subprocess.check_output(["nm", "-D", "--defined-only", kb.lib_path()], text=True)
    exported = {line.split()[-1] for line in out.splitlines() if " T " in line}
    assert not [c for c in called if c not in exported]
    # the reference's Go surface is kept: constructor name, both RPC methods, the interface assertion (service.go:39-45, 54, 173, 213-215)
    for needle in ("func NewSuggestionService() *SuggestionService", "func (s *SuggestionService) GetSuggestions(",
                   "func (s *SuggestionService) ValidateAlgorithmSettings(", "var _ api_v1_beta1.SuggestionServer = &SuggestionService{}"):
        assert needle in src, needle
    assert src.count("{") == src.count("}") and src.count("(") == src.count(")")


def test_go_main_keeps_the_reference_listener():
    src = open(GO_MAIN).read()
    assert 'address        = "0.0.0.0:6789"' in src                      # main.go:31
    assert "RegisterSuggestionServer(srv, svc)" in src and "RegisterHealthServer(srv, &healthService{})" in src  # main.go:48-50
    assert "HealthCheckResponse_SERVING" in src                          # main.go:37-41
    assert "grpc.MaxRecvMsgSize(maxRecvMsgSize)" in src                  # the 4 MiB default rejects a 10k-trial request
    assert src.count("{") == src.count("}") and src.count("(") == src.count(")")


def test_dockerfile_is_a_cgo_build_on_a_cuda_base():
    text = open(DOCKERFILE).read()
    runs = [l for l in text.splitlines() if l.startswith("RUN ")]
    assert any("CGO_ENABLED=1" in l and "go build" in l for l in runs)   # the reference builds with CGO_ENABLED=0 (Dockerfile:18)
    assert not any("CGO_ENABLED=0" in l for l in runs)
    assert re.search(r"^FROM nvidia/cuda:[^\s]+-runtime", text, re.M) and "make -C katib_b200/csrc" in text
    assert 'ENTRYPOINT ["./goptuna-suggestion"]' in text                  # same entry point as the reference image


def test_katib_config_routes_the_four_goptuna_algorithms():
    cfg = yaml.safe_load(open(CONFIG))
    entries = cfg["runtime"]["suggestions"]
    assert sorted(e["algorithmName"] for e in entries) == ["cmaes", "random", "sobol", "tpe"]
    for e in entries:
        assert e["image"].endswith("suggestion-goptuna-b200:latest") and e["resources"]["limits"]["nvidia.com/gpu"] >= 1
