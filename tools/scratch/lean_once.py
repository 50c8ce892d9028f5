"""A few calls of the single-kernel GetSuggestions path at Katib's shape, for an ncu capture (-k regex:lean)."""
import numpy as np, katib_b200 as kb
ctx = kb.Context(0)
rng = np.random.Generator(np.random.PCG64(1))
kind = [kb.KB_UNIFORM, kb.KB_INT, kb.KB_CATEGORICAL]
ctx.space_set(kind, [0.01, 2, 0], [0.03, 5, 0], n_choices=[0, 0, 3])
n = 100
vals = np.stack([rng.uniform(0.01, 0.03, n), np.floor(rng.uniform(2, 6, n)), np.floor(rng.uniform(0, 3, n))], axis=1)
loss = (vals[:, 0] - 0.02) ** 2 * 1e4 + 0.1 * vals[:, 1] + (vals[:, 2] != 1) + rng.normal(0, 0.01, n)
for i in range(6):
    ctx.tpe_set_history(vals, loss); print(ctx.tpe_suggest(24, i))
