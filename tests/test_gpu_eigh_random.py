"""kb_eigh on matrices chosen to sit on the solver's path boundaries and weak spots: sizes around the single-CTA /
Gram-kernel switch (32 / 33 columns), odd sizes (padding blocks, scalar loads), clustered and repeated eigenvalues,
rank-deficient and indefinite input (Cholesky fails -> the path with accumulated V), wide dynamic range."""
import numpy as np
import pytest

import katib_b200 as kb

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def ctx():
    c = kb.Context(0)
    yield c
    c.close()


def check(ctx, A, tol=1e-10):
    n = len(A)
    w, V = ctx.eigh(A)
    S = (A + A.T) / 2
    wr = np.linalg.eigvalsh(S)
    scale = max(np.abs(wr).max(), 1e-300)
    assert np.all(np.diff(w) >= 0)
    assert np.max(np.abs(w - wr)) <= tol * scale, np.max(np.abs(w - wr)) / scale
    assert np.max(np.abs(V.T @ V - np.eye(n))) <= 1e-10
    assert np.max(np.abs(S @ V - V * w)) <= tol * scale * np.sqrt(n)


@pytest.mark.parametrize("n", [2, 5, 31, 32, 33, 34, 47, 63, 65, 100, 129, 255, 257, 300])
def test_sizes_around_the_path_switches(ctx, n):
    rng = np.random.Generator(np.random.PCG64(n))
    Q, _ = np.linalg.qr(rng.normal(size=(n, n)))
    check(ctx, (Q * np.logspace(-2, 3, n)) @ Q.T)


@pytest.mark.parametrize("kind", ["clustered", "repeated", "rank_deficient", "indefinite", "wide_range", "nearly_diagonal"])
@pytest.mark.parametrize("n", [40, 150])
def test_hard_spectra(ctx, kind, n):
    rng = np.random.Generator(np.random.PCG64(sum(map(ord, kind)) + n))
    Q, _ = np.linalg.qr(rng.normal(size=(n, n)))
    if kind == "clustered":
        d = 1.0 + 1e-9 * rng.normal(size=n)
    elif kind == "repeated":
        d = np.repeat([1.0, 2.0, 5.0, 5.0, 9.0], (n + 4) // 5)[:n]
    elif kind == "rank_deficient":
        d = np.concatenate([np.zeros(n // 3), np.linspace(1, 2, n - n // 3)])
    elif kind == "indefinite":
        d = np.linspace(-3, 4, n)
    elif kind == "wide_range":
        d = np.logspace(-10, 4, n)
    else:
        A = np.diag(np.linspace(1, 2, n)) + 1e-7 * rng.normal(size=(n, n))
        check(ctx, A)
        return
    check(ctx, (Q * d) @ Q.T, tol=1e-9 if kind in ("rank_deficient", "indefinite") else 1e-10)
