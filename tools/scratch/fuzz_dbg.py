import sys, os, numpy as np
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "..", "tests"))
import katib_b200 as kb
from test_gpu_tpe import make_history, oracle_scores, rel_err, sample_candidates
from test_gpu_tpe_random_spaces import random_space
np.set_printoptions(linewidth=200, precision=6)
for seed in (40, 58, 159):
    rng = np.random.Generator(np.random.PCG64(1000 + seed))
    sp = random_space(rng)
    n_trials = int(rng.choice([0, 3, 11, 40, 260, 1500])); n_cand = int(rng.choice([1, 24, 700, 5000]))
    vals = make_history(rng, n_trials, sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"]).reshape(n_trials, len(sp["kinds"]))
    for d, (k, lo, hi, st) in enumerate(zip(sp["kinds"], sp["lows"], sp["highs"], sp["steps"])):
        if k == kb.KB_STEP_INT: vals[:, d] = np.minimum(vals[:, d], lo + st * np.floor((hi - lo) / st))
    loss = rng.normal(size=n_trials) + (vals[:, 0] if n_trials else 0.0) * 1e-3
    cand = sample_candidates(rng, n_cand, sp, vals)
    ref, best = oracle_scores(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"], vals, loss, cand)
    ctx = kb.Context(0); ctx.space_set(sp["kinds"], sp["lows"], sp["highs"], sp["steps"], sp["nch"]); ctx.tpe_set_history(vals, loss)
    print("seed", seed, sp, "n_trials", n_trials, "n_cand", n_cand)
    for merge in (0, 1):
        for prune in (0, 1):
            ctx.tpe_set_option(1, merge); ctx.tpe_set_option(2, prune)
            got = ctx.tpe_scores(cand)
            errs = [rel_err(got[:, d], ref[:, d]) for d in range(len(sp["kinds"]))]
            try:
                idx = list(ctx.tpe_score_select(cand)[0])
            except Exception as e:
                idx = str(e)
            print("  merge", merge, "prune", prune, "errs", ["%.1e" % e for e in errs], "idx", idx, "best", best)
            for d in range(len(sp["kinds"])):
                if errs[d] >= 1e-5:
                    i = int(np.argmax(np.abs(got[:, d] - ref[:, d]) / np.maximum(1, np.abs(ref[:, d]))))
                    print("    dim", d, "kind", sp["kinds"][d], "cand", cand[i, d], "got", got[i, d], "ref", ref[i, d])
                if not isinstance(idx, str) and idx[d] != best[d]:
                    print("    dim", d, "kind", sp["kinds"][d], "picked", idx[d], cand[idx[d], d], ref[idx[d], d], "oracle", best[d], cand[best[d], d], ref[best[d], d])
    ctx.close()
