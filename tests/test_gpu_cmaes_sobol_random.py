"""Randomised CMA-ES optimisers and Sobol windows against the oracle."""
import os

import numpy as np
import pytest

import _oracle as O
import katib_b200 as kb

pytestmark = pytest.mark.gpu
N_SEEDS = int(os.environ.get("KB_FUZZ_SEEDS", "10"))


def state_vec(st):
    return np.concatenate([st["mean"], [st["sigma"]], st["C"].ravel(), st["p_sigma"], st["p_c"]])


@pytest.mark.parametrize("seed", range(N_SEEDS))
def test_cmaes_random_shapes_track_the_oracle(seed):
    rng = np.random.Generator(np.random.PCG64(3000 + seed))
    n = int(rng.choice([2, 3, 7, 16, 33, 40, 65]))
    lam = int(rng.choice([0, 4, 2 * n + 1, 6 * n]))
    mean, sigma = rng.normal(size=n) * 3, float(rng.choice([0.05, 1.0, 7.0]))
    ctx = kb.Context(0)
    ctx.cmaes_create(mean, sigma, lam)
    opt = O.Cma(mean, sigma, lam)
    lam = opt.popsize
    scale = np.logspace(0, float(rng.choice([0, 2, 5])), n)
    for gen in range(6):
        B, D = opt.eig()
        ctx.cmaes_set_eig(B, D)  # same factors on both sides: samples must agree to rounding
        z = rng.normal(size=(lam, n))
        X = ctx.cmaes_ask(lam, z)
        Xr = opt.ask_batch(z)
        assert np.max(np.abs(X - Xr)) <= 1e-11 * max(1.0, np.max(np.abs(Xr))), (n, lam, gen)
        f = np.sum((X - 0.5) ** 2 * scale, axis=1)
        if gen == 2:
            f[0] = f[lam - 1]  # a tie: stable order decides
        ctx.cmaes_tell(X, f)
        opt.tell(X, f)
        a, b = state_vec(ctx.cmaes_state()), opt.state_vec()
        assert np.max(np.abs(a - b)) <= 1e-9 * max(1.0, np.max(np.abs(b))), (n, lam, gen)
    ctx.close()


@pytest.mark.parametrize("seed", range(N_SEEDS))
def test_sobol_random_windows_bit_exact(seed, golden_dir):
    rng = np.random.Generator(np.random.PCG64(4000 + seed))
    dim = int(rng.choice([1, 2, 3, 17, 64, 65, 500, 1024]))
    first = int(rng.choice([0, 1, 2 ** 16 - 3, 2 ** 24 + 5, 2 ** 31 - 100, 2 ** 32 - 5000]))
    n = int(rng.choice([1, 31, 1000, 4097]))
    ctx = kb.Context(0)
    got = ctx.sobol_fill(dim, first, n)
    assert np.array_equal(got, O.sobol(golden_dir, dim, first, n)), (dim, first, n)
    ctx.close()
