"""Eigensolver diagnostics on the GPU box: time, sweeps and accuracy of kb_eigh against LAPACK for a few sizes, then the
sweeps the solver needs inside a CMA-ES run.  KB_EIGH_M=0|8|16 and KB_EIGH_INNER=k select the block step."""
import os
import time

import numpy as np

import katib_b200 as kb

ctx = kb.Context(0)
rng = np.random.Generator(np.random.PCG64(1))
print("KB_EIGH_M", os.environ.get("KB_EIGH_M"), "KB_EIGH_INNER", os.environ.get("KB_EIGH_INNER"))
for n in (7, 64, 130, 256, 512, 1024):
    M = rng.normal(size=(n, n))
    A = M @ M.T / n + np.diag(np.logspace(-3, 2, n))
    ctx.eigh(A)
    t0 = time.perf_counter()
    w, V = ctx.eigh(A)
    dt = time.perf_counter() - t0
    ctx.profile(True)
    ctx.eigh(A)
    ms, _ = ctx.profile_read(9)
    ctx.profile(False)
    sw = ctx.eigh_sweeps()
    wr = np.linalg.eigvalsh(A)
    err_w = np.max(np.abs(w - wr) / np.abs(wr))
    err_o = np.max(np.abs(V.T @ V - np.eye(n)))
    err_r = np.max(np.abs(A @ V - V * w)) / np.max(np.abs(wr))
    print(f"n={n:5d} wall {dt * 1e3:7.2f} ms kernel {ms:7.2f} ms sweeps {sw:2d}  eigenvalue rel err {err_w:.1e} "
          f"orthogonality {err_o:.1e} residual {err_r:.1e}")
n, lam = 512, 4096
ctx.cmaes_create(np.full(n, 3.0), 2.0, lam)
for g in range(4):
    X = ctx.cmaes_ask(lam, rng.normal(size=(lam, n)))
    sw = ctx.eigh_sweeps()
    ctx.profile(True)
    f = np.sum((X * np.logspace(0, 3, n)) ** 2, axis=1)
    ctx.cmaes_tell(X, f)
    ctx.cmaes_ask(lam, rng.normal(size=(lam, n)))
    ms, _ = ctx.profile_read(9)
    ctx.profile(False)
    print("gen", g, "sweeps", sw, f"eig {ms:.2f} ms")
