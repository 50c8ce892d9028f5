"""Oracle (oracle/*.c) against the committed fixtures in tests/golden/ (tools/gen_golden.py).

Sobol is pinned against scipy's independent Joe-Kuo generator (bit-exact).  TPE and CMA-ES fixtures
come from a second, numpy/scipy restatement of the published rules -- agreement shows the C port is
self-consistent, not that it equals goptuna (parity unpinned; see oracle/katib_oracle.h).
"""
import os

import numpy as np
import pytest

import _oracle as O


@pytest.fixture(scope="module")
def g(golden_dir):
    return {k: dict(np.load(os.path.join(golden_dir, f"{k}_golden.npz"))) for k in ("sobol", "tpe", "cmaes")}


# ---------------------------------------------------------------- Sobol: bit-exact
@pytest.mark.parametrize("dim", [1, 2, 3, 64, 1000])
def test_sobol_first16(g, golden_dir, dim):
    got = O.sobol(golden_dir, dim, 0, 16)
    assert np.array_equal(got, g["sobol"][f"first16_d{dim}"])


def test_sobol_offset_and_checksum(g, golden_dir):
    got = O.sobol(golden_dir, 64, (1 << 20) - 3, 8)
    assert np.array_equal(got, g["sobol"]["d64_from_2p20m3"])
    n = 1 << 20
    xi = O.sobol_u32(golden_dir, 64, 0, n).astype(np.uint64)
    idx = (np.arange(n, dtype=np.uint64) + np.uint64(1))[:, None]
    with np.errstate(over="ignore"):
        assert np.array_equal((xi * idx).sum(axis=0, dtype=np.uint64), g["sobol"]["d64_n2p20_wsum"])
    assert np.array_equal(np.bitwise_xor.reduce(xi * np.uint64(2654435761) + idx, axis=0), g["sobol"]["d64_n2p20_xor"])


def test_sobol_chunks_concatenate(golden_dir):
    a = O.sobol(golden_dir, 7, 0, 300)
    b = np.concatenate([O.sobol(golden_dir, 7, 0, 113), O.sobol(golden_dir, 7, 113, 187)])
    assert np.array_equal(a, b)
    assert O.sobol(golden_dir, 5, 12, 0).shape == (0, 5)


def test_sobol_30_bits(golden_dir):
    from_32 = O.sobol_u32(golden_dir, 9, 0, 64)
    got = O.sobol(golden_dir, 9, 0, 64, bits=30)
    assert np.array_equal(got, (from_32 >> 2) / float(1 << 30))


# ---------------------------------------------------------------- TPE
def test_gamma_weights(g):
    t = g["tpe"]
    assert [O.lib().kbo_tpe_gamma(n) for n in range(400)] == list(t["gamma_tab"])
    w = np.zeros(40)
    O.lib().kbo_tpe_weights(40, w)
    np.testing.assert_allclose(w, t["weights_n40"], rtol=0, atol=1e-16)
    for n in (1, 24, 25, 26):
        w = np.zeros(n)
        O.lib().kbo_tpe_weights(n, w)
        assert w[-1] == 1.0 and (n < 26 or w[0] == 1.0 / n)


@pytest.mark.parametrize("tag,low,high,q", [("p12", 0.0, 1.0, 0.0), ("i12", -0.5, 9.5, 1.0)])
def test_parzen_12(g, tag, low, high, q):
    t = g["tpe"]
    w, mu, sg = O.parzen_build(t[f"{tag}_obs"], low, high)
    np.testing.assert_allclose(w, t[f"{tag}_w"], rtol=1e-15)
    assert np.array_equal(mu, t[f"{tag}_mu"])
    np.testing.assert_allclose(sg, t[f"{tag}_sigma"], rtol=1e-15)
    key = f"{tag}_logpdf_q{int(q)}"
    np.testing.assert_allclose(O.gmm_log_pdf(t[f"{tag}_pts"], w, mu, sg, low, high, q), t[key], rtol=1e-12)


def test_parzen_edges():
    w, mu, sg = O.parzen_build([], 2.0, 6.0)  # no observation: prior only
    assert list(w) == [1.0] and list(mu) == [4.0] and list(sg) == [4.0]
    w, mu, sg = O.parzen_build([4.0], 2.0, 6.0)  # observation equal to the prior mean
    assert list(mu) == [4.0, 4.0] and sg[0] == 4.0 and sg[1] == 4.0 / 3.0  # prior first (searchsorted left)
    w, mu, sg = O.parzen_build(np.full(200, 3.0), 2.0, 6.0)  # all equal: sigma clipped to range/100
    assert np.all(sg[:-2] == 4.0 / 100.0) and sg[-2] == 1.0 and sg[-1] == 4.0 and mu[-1] == 4.0


def test_score_numeric_and_split(g):
    t = g["tpe"]
    sc, best = O.tpe_score_numeric(t["s300_cand"], t["s300_vals"], t["s300_loss"], -2.0, 3.0)
    np.testing.assert_allclose(sc, t["s300_scores"], rtol=1e-11, atol=1e-12)
    assert best == int(t["s300_best"])
    sc, best = O.tpe_score_numeric(t["i300_cand"], t["i300_vals"], t["i300_loss"], -0.5, 63.5, 1.0)
    np.testing.assert_allclose(sc, t["i300_scores"], rtol=1e-10, atol=1e-12)
    assert best == int(t["i300_best"])
    b, a = O.tpe_split(t["s300_vals"], t["s300_loss"])
    assert len(b) == 25 and len(a) == 275
    order = np.argsort(t["s300_loss"], kind="stable")
    assert np.array_equal(b, t["s300_vals"][np.sort(order[:25])])


def test_categorical(g):
    t = g["tpe"]
    pl, pg = O.categorical_tables(t["c300_vals"], t["c300_loss"], 5)
    np.testing.assert_allclose(pl, t["c300_pl"], rtol=1e-14)
    np.testing.assert_allclose(pg, t["c300_pg"], rtol=1e-14)
    sc, best = O.tpe_score_categorical(np.arange(5.0), t["c300_vals"], t["c300_loss"], 5)
    np.testing.assert_allclose(sc, np.log(pl) - np.log(pg), rtol=1e-14)
    assert best == int(np.argmax(sc))


def test_gmm_sample_in_range():
    w, mu, sg = O.parzen_build([0.2, 0.4, 0.45, 0.9], 0.0, 1.0)
    out = np.zeros(4000)
    O.lib().kbo_gmm_sample(3, 4000, len(w), w, mu, sg, 0.0, 1.0, 0.0, out)
    assert out.min() >= 0.0 and out.max() < 1.0
    O.lib().kbo_gmm_sample(3, 4000, len(w), w, mu, sg, -0.5, 9.5, 1.0, out)
    assert np.all(out == np.round(out)) and out.min() >= 0 and out.max() <= 10


# ---------------------------------------------------------------- CMA-ES
@pytest.mark.parametrize("n,lam", [(4, 8), (10, 0), (24, 48)])
def test_cma_tell(g, n, lam):
    t = g["cmaes"]
    opt = O.Cma(np.full(n, 0.5), 1.3, lam)
    consts, w = opt.constants()
    np.testing.assert_allclose(consts, t[f"n{n}_consts"], rtol=1e-13)
    np.testing.assert_allclose(w, t[f"n{n}_weights"], rtol=1e-13)
    for gen in range(3):
        opt.tell(t[f"n{n}_xs"][gen], t[f"n{n}_fs"][gen])
        np.testing.assert_allclose(opt.state_vec(), t[f"n{n}_states"][gen], rtol=1e-9, atol=1e-11)


def test_jacobi_vs_lapack():
    rng = np.random.Generator(np.random.PCG64(5))
    for n in (2, 3, 17, 64):
        M = rng.normal(size=(n, n))
        A = M @ M.T + np.eye(n) * 0.1
        w, V = O.jacobi_eigh(A)
        np.testing.assert_allclose(w, np.linalg.eigvalsh(A), rtol=1e-12)
        np.testing.assert_allclose(V @ np.diag(w) @ V.T, A, rtol=0, atol=1e-11 * np.abs(A).max())
        np.testing.assert_allclose(V.T @ V, np.eye(n), atol=1e-13)
        assert np.all(V[np.abs(V).argmax(axis=0), np.arange(n)] > 0)


def test_cma_ask_identity_gen0():
    opt = O.Cma(np.arange(5.0), 2.0, 6)
    B, D = opt.eig()
    assert np.array_equal(B, np.eye(5)) and np.array_equal(D, np.ones(5))
    z = np.arange(30.0).reshape(6, 5) / 7
    np.testing.assert_allclose(opt.ask_batch(z), np.arange(5.0) + 2.0 * z, rtol=1e-15)
