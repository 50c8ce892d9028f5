"""Jacobi sweeps and eigensolver time per CMA-ES generation at configs[3]'s shape (n = 512, popsize 4096)."""
import time, numpy as np, katib_b200 as kb
ctx = kb.Context(0)
rng = np.random.Generator(np.random.PCG64(4))
n, lam = 512, 4096
Q, _ = np.linalg.qr(rng.normal(size=(n, n)))
scale = np.logspace(0, 6, n)
fobj = lambda X: np.sum((X @ Q) ** 2 * scale, axis=1)
ctx.cmaes_create(np.full(n, 3.0), 2.0, lam, np.full(n, -5.0), np.full(n, 5.0))
out = []
for g in range(80):
    Z = rng.normal(size=(lam, n))
    ctx.profile(True)
    t0 = time.perf_counter(); X = ctx.cmaes_ask(lam, Z); t1 = time.perf_counter()
    eig = ctx.profile_read(9)[0]; ctx.profile(False)
    ctx.cmaes_tell(X, fobj(X))
    out.append((g, ctx.eigh_sweeps(), round(eig, 2), round(1e3 * (t1 - t0), 2)))
print(out[:6]); print(out[6::6])
