import sys, os, time, numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import _pb, katib_b200 as kb
svc = kb.Service(0)
space = [_pb.parameter("lr", "DOUBLE", max="0.03", min="0.01"), _pb.parameter("num-layers", "INT", max="5", min="2"),
         _pb.parameter("optimizer", "CATEGORICAL", list=["sgd", "adam", "ftrl"])]
exp = _pb.experiment("hp", "tpe", [("random_state", "10")], "MINIMIZE", "loss", space)
trials, pending = [], None
for i in range(200):
    if pending is not None:
        d = dict(pending)
        f = (float(d["lr"]) - 0.02) ** 2 * 1e4 + 0.1 * int(d["num-layers"]) + (d["optimizer"] != "adam")
        trials.append(_pb.trial(f"t{i}", pending, "SUCCEEDED", [("loss", repr(f))]))
    code, reply, msg = svc.get_suggestions(_pb.get_suggestions_request(exp, trials, 1, i + 1))
    assert code == 0, msg
    pending = _pb.parse_reply(reply)[0]
def timeit(req, n=300):
    ts = []
    for _ in range(n):
        t0 = time.perf_counter(); code, reply, msg = svc.get_suggestions(req); ts.append(time.perf_counter() - t0)
        assert code == 0, msg
    return 1e6 * np.percentile(ts, 50)
print("trials", len(trials), "request bytes", len(_pb.get_suggestions_request(exp, trials, 0, 0)))
print("parse+convert+sync only (0 suggestions): %.1f us" % timeit(_pb.get_suggestions_request(exp, trials, 0, 0)))
print("+1 suggestion, history unchanged:        %.1f us" % timeit(_pb.get_suggestions_request(exp, trials, 1, 0)))
code, m = svc.validate(_pb.validate_request(exp)); print("validate %d" % code)
