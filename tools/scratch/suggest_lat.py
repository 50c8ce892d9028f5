import os, time, numpy as np, katib_b200 as kb
ctx = kb.Context(0)
rng = np.random.Generator(np.random.PCG64(1))
kind = [kb.KB_UNIFORM, kb.KB_INT, kb.KB_CATEGORICAL]
ctx.space_set(kind, [0.01, 2, 0], [0.03, 5, 0], n_choices=[0, 0, 3])
n = 100
vals = np.stack([rng.uniform(0.01, 0.03, n), np.floor(rng.uniform(2, 6, n)), np.floor(rng.uniform(0, 3, n))], axis=1)
loss = (vals[:, 0] - 0.02) ** 2 * 1e4 + 0.1 * vals[:, 1] + (vals[:, 2] != 1) + rng.normal(0, 0.01, n)
ctx.tpe_set_history(vals, loss)
for mode in ("1", "0"):
    os.environ["KB_SUGGEST_SMALL"] = mode
    for i in range(50): ctx.tpe_suggest(24, i)
    ts = []
    for i in range(500):
        t0 = time.perf_counter(); ctx.tpe_suggest(24, i); ts.append(time.perf_counter() - t0)
    ctx.profile(True)
    for i in range(100): ctx.tpe_suggest(24, i)
    ms = {k: ctx.profile_read(k) for k in (0, 1, 2, 3, 6)}
    ctx.profile(False)
    print("small" if mode == "1" else "staged", "wall p50 %.1f us p95 %.1f us" % (1e6 * np.percentile(ts, 50), 1e6 * np.percentile(ts, 95)),
          {k: (round(1e3 * v[0] / max(1, v[1]), 1), v[1]) for k, v in ms.items()})
t0 = time.perf_counter()
for i in range(200): ctx.device_info()
print("trivial C call %.1f us" % (1e6 * (time.perf_counter() - t0) / 200))
