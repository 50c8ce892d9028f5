import numpy as np, katib_b200 as kb
ctx = kb.Context(0)
rng = np.random.Generator(np.random.PCG64(1))
n = 512
M = rng.normal(size=(n, n)); A = M @ M.T / n + np.diag(np.logspace(-3, 2, n))
ctx.eigh(A); ctx.eigh(A)
