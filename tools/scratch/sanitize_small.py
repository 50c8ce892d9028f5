"""Small shapes through every kernel touched in round 2, for compute-sanitizer --tool memcheck / racecheck."""
import numpy as np, katib_b200 as kb
rng = np.random.Generator(np.random.PCG64(3))
kinds = [kb.KB_UNIFORM, kb.KB_INT, kb.KB_CATEGORICAL, kb.KB_DISCRETE, kb.KB_STEP_INT]
lows, highs, steps, nch = [-1.5, -10, 0, 0.0, 1], [5.5, 10, 0, 16.0, 31], [0, 0, 0, 0.25, 3], [0, 0, 5, 0, 0]
def hist(n):
    v = np.stack([rng.uniform(-1.5, 5.5, n), np.floor(rng.uniform(-10, 11, n)), np.floor(rng.uniform(0, 5, n)),
                  0.25 * np.floor(rng.uniform(0, 65, n)), 1 + 3 * np.floor(rng.uniform(0, 11, n))], axis=1)
    return v, (v[:, 0] - 1) ** 2 + 0.05 * v[:, 1] ** 2 + rng.normal(0, 0.3, n)
ctx = kb.Context(0)
ctx.space_set(kinds, lows, highs, steps, nch)
for n in (0, 7, 120, 1500):
    v, l = hist(n)
    for merge in (0, 1):
        ctx.tpe_set_option(1, merge)
        ctx.tpe_set_history(v, l)
        a = ctx.tpe_suggest(24, 5)                      # lean kernel (n <= 512) or small kernel
        cand = ctx.tpe_sample_candidates(3000, 9)       # builds the tables (small or large build), sample kernel
        idx, val, sc = ctx.tpe_score_select(cand)       # stage, cont, quant grid + general, compact, rescore, pick
        cand[::7, 1] += 0.3                             # off-grid candidates: tiles fall back to the general quantised kernel
        ctx.tpe_scores(cand)
        b = ctx.tpe_suggest(70000, 5)                   # sample_stage kernel, two chunks when KB_CHUNK_LOG2=16
        ctx.tpe_set_option(2, 1); ctx.tpe_score_select(cand); ctx.tpe_set_option(2, 0)
pts = ctx.sobol_fill(64, 5, 70000, low=np.full(64, -1.0), high=np.full(64, 3.0))
print("ok", a[:2], b[:2], pts[-1, :2])
