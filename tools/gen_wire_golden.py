#!/usr/bin/env python
"""Pin tests/_pb.py (the tests' protobuf encoder) and the service's parser against the reference's own generated
stubs: build the requests of the reference's unit test (pkg/suggestion/v1beta1/goptuna/service_test.go:29-158) with
/root/reference/pkg/apis/manager/v1beta1/python/api_pb2.py, assert tests/_pb.py produces the same bytes, and store
them under tests/golden/wire_*.bin.  Needs /root/reference (this container only) and the pure-python protobuf
runtime:  PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION=python python tools/gen_wire_golden.py"""
import os
import sys

os.environ.setdefault("PROTOCOL_BUFFERS_PYTHON_IMPLEMENTATION", "python")
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, "/root/reference/pkg/apis/manager/v1beta1/python")
sys.path.insert(0, os.path.join(ROOT, "tests"))
import api_pb2 as api  # noqa: E402
import _pb  # noqa: E402

OUT = os.path.join(ROOT, "tests", "golden")


def ref_params():
    P, F = api.ParameterSpec, api.FeasibleSpace
    return [P(name="param-1", parameter_type=api.INT, feasible_space=F(max="10", min="-10")),
            P(name="param-2", parameter_type=api.CATEGORICAL, feasible_space=F(list=["cat1", "cat2", "cat3"])),
            P(name="param-3", parameter_type=api.DISCRETE, feasible_space=F(list=["3", "2", "6"])),
            P(name="param-4", parameter_type=api.DOUBLE, feasible_space=F(max="5.5", min="-1.5"))]


def my_params():
    return [_pb.parameter("param-1", "INT", max="10", min="-10"),
            _pb.parameter("param-2", "CATEGORICAL", list=["cat1", "cat2", "cat3"]),
            _pb.parameter("param-3", "DISCRETE", list=["3", "2", "6"]),
            _pb.parameter("param-4", "DOUBLE", max="5.5", min="-1.5")]


def main():
    for algo, settings in (("cmaes", [("random_state", "10")]), ("tpe", [("random_state", "10")]), ("random", [("random_state", "10")]),
                           ("sobol", [])):
        exp = api.Experiment(name="test", spec=api.ExperimentSpec(
            algorithm=api.AlgorithmSpec(algorithm_name=algo, algorithm_settings=[api.AlgorithmSetting(name=k, value=v) for k, v in settings]),
            objective=api.ObjectiveSpec(type=api.MINIMIZE, goal=0.9, objective_metric_name="metric-1"),
            parameter_specs=api.ExperimentSpec.ParameterSpecs(parameters=ref_params())))
        req = api.GetSuggestionsRequest(experiment=exp, current_request_number=2)
        ref_bytes = req.SerializeToString(deterministic=True)
        open(os.path.join(OUT, f"wire_getsuggestions_{algo}.bin"), "wb").write(ref_bytes)
        # tests/_pb.py omits the (unused) goal field; compare through a re-parse instead of raw bytes
        mine = _pb.get_suggestions_request(_pb.experiment("test", algo, settings, "MINIMIZE", "metric-1", my_params()), [], 2)
        back = api.GetSuggestionsRequest.FromString(mine)
        exp.spec.objective.goal = 0.0
        assert back == api.GetSuggestionsRequest(experiment=exp, current_request_number=2), algo
    # a request with trials (second-call shape) and a reply, both from the reference stubs
    t = api.Trial(name="trial-a", spec=api.TrialSpec(parameter_assignments=api.TrialSpec.ParameterAssignments(
        assignments=[api.ParameterAssignment(name="param-1", value="3"), api.ParameterAssignment(name="param-4", value="0.25")])),
        status=api.TrialStatus(start_time="2024-01-02T03:04:05Z", completion_time="2024-01-02T03:05:05.5Z", condition=api.TrialStatus.SUCCEEDED,
                               observation=api.Observation(metrics=[api.Metric(name="metric-1", value="0.5")])))
    mine_t = _pb.trial("trial-a", [("param-1", "3"), ("param-4", "0.25")], "SUCCEEDED", [("metric-1", "0.5")],
                       "2024-01-02T03:04:05Z", "2024-01-02T03:05:05.5Z")
    assert api.Trial.FromString(mine_t) == t
    open(os.path.join(OUT, "wire_trial.bin"), "wb").write(t.SerializeToString(deterministic=True))
    reply = api.GetSuggestionsReply(parameter_assignments=[api.GetSuggestionsReply.ParameterAssignments(
        assignments=[api.ParameterAssignment(name="a", value="1"), api.ParameterAssignment(name="b", value="x")])])
    rb = reply.SerializeToString(deterministic=True)
    assert _pb.parse_reply(rb) == [[("a", "1"), ("b", "x")]]
    open(os.path.join(OUT, "wire_reply.bin"), "wb").write(rb)
    print("ok:", sorted(f for f in os.listdir(OUT) if f.startswith("wire_")))


if __name__ == "__main__":
    main()
