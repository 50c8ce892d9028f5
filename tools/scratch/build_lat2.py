"""GetSuggestions-latency shape, split into its parts: wall time of kb_tpe_set_history and kb_tpe_suggest, and the device
time of the build kernel(s) (profile class 5) and of the fused small suggest kernel (class 3).
Run twice: KB_BUILD_SMALL=1 (single-launch build, default) and KB_BUILD_SMALL=0 (split + parzen + categorical)."""
import os, time, numpy as np, katib_b200 as kb
ctx = kb.Context(0)
rng = np.random.Generator(np.random.PCG64(1))
kind = [kb.KB_UNIFORM, kb.KB_INT, kb.KB_CATEGORICAL]
ctx.space_set(kind, [0.01, 2, 0], [0.03, 5, 0], n_choices=[0, 0, 3])
n = 100
vals = np.stack([rng.uniform(0.01, 0.03, n), np.floor(rng.uniform(2, 6, n)), np.floor(rng.uniform(0, 3, n))], axis=1)
loss = (vals[:, 0] - 0.02) ** 2 * 1e4 + 0.1 * vals[:, 1] + (vals[:, 2] != 1) + rng.normal(0, 0.01, n)
for i in range(50):
    ctx.tpe_set_history(vals, loss); ctx.tpe_suggest(24, i)
a, b = [], []
for i in range(500):
    t0 = time.perf_counter(); ctx.tpe_set_history(vals, loss); t1 = time.perf_counter(); ctx.tpe_suggest(24, i); t2 = time.perf_counter()
    a.append(t1 - t0); b.append(t2 - t1)
ctx.profile(True)
for i in range(200):
    ctx.tpe_set_history(vals, loss); ctx.tpe_suggest(24, i)
ms = {k: ctx.profile_read(k) for k in (3, 5)}
ctx.profile(False)
p = lambda x: 1e6 * np.percentile(x, 50)
print("KB_BUILD_SMALL=%s set_history %.1f us, suggest %.1f us, both %.1f us; device: build %.1f us (%d launches), suggest kernel %.1f us" % (
    os.environ.get("KB_BUILD_SMALL", "1"), p(a), p(b), p(np.array(a) + np.array(b)), 1e3 * ms[5][0] / 200, ms[5][1], 1e3 * ms[3][0] / 200))
