#!/usr/bin/env python
"""Generate the committed fixtures under tests/golden/.

What pins what (see DESIGN.md "Oracle"):
  * joe_kuo_d1024.npz      -- Joe-Kuo `new-joe-kuo-6.21201` direction numbers (first 1024 dims), read from the
                              copy scipy ships.  Input data for oracle/sobol.c (the oracle embeds no table).
  * sobol_golden.npz       -- points produced by scipy.stats.qmc.Sobol(scramble=False, bits=32): an independent
                              implementation of the same published generator.  Pins table + gray-code order.
  * tpe_golden.npz         -- numbers from the numpy/scipy restatement of the Optuna-1.x TPE rules written below
                              (vectorised, scipy.special.erf / logsumexp).  It is a second, independently written
                              restatement, NOT output of goptuna: TPE parity stays "unpinned".
  * cmaes_golden.npz       -- same for one CMA.tell step (numpy.linalg.eigh as the eigensolver).

goptuna itself (Go, un-vendored module) cannot be run in this image, so no fixture here comes from it.
Run:  python tools/gen_golden.py        (needs numpy + scipy; deterministic)
"""
import math
import os

import numpy as np
import scipy
import scipy.special
from scipy.stats import qmc

OUT = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "tests", "golden")
EPS = 1e-12


# ------------------------------------------------------------------ Sobol
def gen_sobol():
    p = os.path.join(os.path.dirname(scipy.stats.__file__), "_sobol_direction_numbers.npz")
    d = np.load(p)
    np.savez_compressed(os.path.join(OUT, "joe_kuo_d1024.npz"), poly=d["poly"][:1024].astype(np.int64),
                        vinit=d["vinit"][:1024].astype(np.int64))
    out = {}
    for dim in (1, 2, 3, 64, 1000):
        s = qmc.Sobol(dim, scramble=False, bits=32)
        out[f"first16_d{dim}"] = s.random(16)
    s = qmc.Sobol(64, scramble=False, bits=32)
    s.fast_forward((1 << 20) - 3)
    out["d64_from_2p20m3"] = s.random(8)
    # position-weighted checksum over 2^20 points x 64 dims: sum_i (i+1) * (x_i * 2^32) mod 2^64, per dim
    s = qmc.Sobol(64, scramble=False, bits=32)
    acc = np.zeros(64, dtype=np.uint64)
    xacc = np.zeros(64, dtype=np.uint64)
    n_total, chunk = 1 << 20, 1 << 18
    for c in range(n_total // chunk):
        x = s.random(chunk)
        xi = (x * 4294967296.0).astype(np.uint64)
        idx = (np.arange(c * chunk, (c + 1) * chunk, dtype=np.uint64) + np.uint64(1))[:, None]
        with np.errstate(over="ignore"):
            acc += (xi * idx).sum(axis=0, dtype=np.uint64)
        xacc ^= np.bitwise_xor.reduce(xi * np.uint64(2654435761) + idx, axis=0)
    out["d64_n2p20_wsum"] = acc
    out["d64_n2p20_xor"] = xacc
    np.savez_compressed(os.path.join(OUT, "sobol_golden.npz"), **out)


# ------------------------------------------------------------------ TPE (numpy restatement of Optuna 1.x)
def default_gamma(n):
    return min(int(math.ceil(0.1 * n)), 25)


def default_weights(n):
    if n == 0:
        return np.asarray([])
    if n < 25:
        return np.ones(n)
    ramp = np.linspace(1.0 / n, 1.0, num=n - 25)
    return np.concatenate([ramp, np.ones(25)])


def parzen(mus, low, high, prior_weight=1.0, magic_clip=True, endpoints=False):
    mus = np.asarray(mus, dtype=float)
    prior_mu, prior_sigma = 0.5 * (low + high), 1.0 * (high - low)
    if mus.size == 0:
        return np.asarray([1.0]), np.asarray([prior_mu]), np.asarray([prior_sigma])
    order = np.argsort(mus, kind="stable")
    ordered = mus[order]
    prior_pos = int(np.searchsorted(ordered, prior_mu))
    lsh = np.zeros(len(mus) + 3)
    sm = lsh[1:-1]
    sm[:prior_pos] = ordered[:prior_pos]
    sm[prior_pos] = prior_mu
    sm[prior_pos + 1:] = ordered[prior_pos:]
    lsh[0], lsh[-1] = low, high
    sigma = np.maximum(lsh[1:-1] - lsh[0:-2], lsh[2:] - lsh[1:-1])
    if not endpoints and lsh.size > 2:
        sigma[0] = lsh[2] - lsh[1]
        sigma[-1] = lsh[-2] - lsh[-3]
    uw = default_weights(mus.size)
    sw = np.zeros_like(sm)
    sw[:prior_pos] = uw[order[:prior_pos]]
    sw[prior_pos] = prior_weight
    sw[prior_pos + 1:] = uw[order[prior_pos:]]
    sw = sw / sw.sum()
    maxsigma = 1.0 * (high - low)
    minsigma = 1.0 * (high - low) / min(100.0, 1.0 + len(sm)) if magic_clip else EPS
    sigma = np.clip(sigma, minsigma, maxsigma)
    sigma[prior_pos] = prior_sigma
    return sw, sm.copy(), sigma


def normal_cdf(x, mu, sigma):
    return 0.5 * (1 + scipy.special.erf((x - mu) / np.maximum(np.sqrt(2) * sigma, EPS)))


def gmm_log_pdf(samples, w, mu, sigma, low, high, q=None):
    samples = np.asarray(samples, dtype=float)
    p_accept = np.sum(w * (normal_cdf(high, mu, sigma) - normal_cdf(low, mu, sigma)))
    if q is None:
        dist = samples[:, None] - mu
        mahal = (dist / np.maximum(sigma, EPS)) ** 2
        coef = w / (np.sqrt(2 * np.pi) * sigma) / p_accept
        return scipy.special.logsumexp(-0.5 * mahal + np.log(coef), axis=1)
    ub = np.minimum(samples + q / 2.0, high)
    lb = np.maximum(samples - q / 2.0, low)
    probs = np.sum(w[:, None] * (normal_cdf(ub[None], mu[:, None], sigma[:, None])
                                 - normal_cdf(lb[None], mu[:, None], sigma[:, None])), axis=0)
    return np.log(probs + EPS) - np.log(p_accept + EPS)


def split(values, loss):
    values, loss = np.asarray(values, float), np.asarray(loss, float)
    nb = default_gamma(len(values))
    asc = np.argsort(loss, kind="stable")
    return values[np.sort(asc[:nb])], values[np.sort(asc[nb:])]


def gen_tpe():
    out = {}
    # (2) of SURVEY 8c: 12 observations on [0,1], estimator + log-pdfs at 5 points, q=0 and quantised
    obs = np.asarray([0.31, 0.05, 0.77, 0.5, 0.52, 0.9, 0.13, 0.66, 0.5, 0.28, 0.97, 0.41])
    w, mu, sg = parzen(obs, 0.0, 1.0)
    pts = np.asarray([0.0, 0.25, 0.5, 0.75, 0.999])
    out["p12_obs"], out["p12_w"], out["p12_mu"], out["p12_sigma"] = obs, w, mu, sg
    out["p12_pts"] = pts
    out["p12_logpdf_q0"] = gmm_log_pdf(pts, w, mu, sg, 0.0, 1.0)
    obs_i = np.asarray([3, 7, 7, 1, 0, 9, 4, 4, 5, 2, 8, 6], dtype=float)
    w, mu, sg = parzen(obs_i, -0.5, 9.5)
    pts_i = np.asarray([0.0, 2.0, 4.0, 7.0, 9.0])
    out["i12_obs"], out["i12_w"], out["i12_mu"], out["i12_sigma"] = obs_i, w, mu, sg
    out["i12_pts"] = pts_i
    out["i12_logpdf_q1"] = gmm_log_pdf(pts_i, w, mu, sg, -0.5, 9.5, q=1.0)
    # weights ramp + gamma
    out["weights_n40"] = default_weights(40)
    out["gamma_tab"] = np.asarray([default_gamma(n) for n in range(0, 400)], dtype=np.int64)
    # full scoring of one parameter: 300 trials, continuous + quantised + categorical
    rng = np.random.Generator(np.random.PCG64(7))
    n = 300
    vals = rng.uniform(-2.0, 3.0, n)
    loss = (vals - 0.5) ** 2 + rng.normal(0, 0.3, n)
    cand = rng.uniform(-2.0, 3.0, 257)
    b, a = split(vals, loss)
    wl, ml, sl = parzen(b, -2.0, 3.0)
    wg, mg, sgg = parzen(a, -2.0, 3.0)
    sc = gmm_log_pdf(cand, wl, ml, sl, -2.0, 3.0) - gmm_log_pdf(cand, wg, mg, sgg, -2.0, 3.0)
    out["s300_vals"], out["s300_loss"], out["s300_cand"], out["s300_scores"] = vals, loss, cand, sc
    out["s300_best"] = np.int64(np.argmax(sc))
    ivals = np.floor(rng.uniform(0, 64, n))
    iloss = (ivals - 20) ** 2 / 100 + rng.normal(0, 0.5, n)
    icand = np.floor(rng.uniform(0, 64, 257))
    b, a = split(ivals, iloss)
    wl, ml, sl = parzen(b, -0.5, 63.5)
    wg, mg, sgg = parzen(a, -0.5, 63.5)
    sc = gmm_log_pdf(icand, wl, ml, sl, -0.5, 63.5, q=1.0) - gmm_log_pdf(icand, wg, mg, sgg, -0.5, 63.5, q=1.0)
    out["i300_vals"], out["i300_loss"], out["i300_cand"], out["i300_scores"] = ivals, iloss, icand, sc
    out["i300_best"] = np.int64(np.argmax(sc))
    # categorical
    C = 5
    cvals = np.floor(rng.uniform(0, C, n))
    closs = (cvals != 2) * 1.0 + rng.normal(0, 0.5, n)
    b, a = split(cvals, closs)
    pl = np.bincount(b.astype(int), minlength=C, weights=default_weights(len(b))) + 1.0
    pl /= pl.sum()
    pg = np.bincount(a.astype(int), minlength=C, weights=default_weights(len(a))) + 1.0
    pg /= pg.sum()
    out["c300_vals"], out["c300_loss"], out["c300_pl"], out["c300_pg"] = cvals, closs, pl, pg
    np.savez_compressed(os.path.join(OUT, "tpe_golden.npz"), **out)


# ------------------------------------------------------------------ CMA-ES (numpy restatement of cmaes.CMA)
class CMA:
    def __init__(self, mean, sigma, popsize=None):
        n = len(mean)
        popsize = popsize or 4 + math.floor(3 * math.log(n))
        mu = popsize // 2
        wp = np.array([math.log((popsize + 1) / 2) - math.log(i + 1) for i in range(popsize)])
        mu_eff = (np.sum(wp[:mu]) ** 2) / np.sum(wp[:mu] ** 2)
        mu_eff_minus = (np.sum(wp[mu:]) ** 2) / np.sum(wp[mu:] ** 2)
        alpha_cov = 2
        c1 = alpha_cov / ((n + 1.3) ** 2 + mu_eff)
        cmu = min(1 - c1 - 1e-8, alpha_cov * (mu_eff - 2 + 1 / mu_eff) / ((n + 2) ** 2 + alpha_cov * mu_eff / 2))
        min_alpha = min(1 + c1 / cmu, 1 + (2 * mu_eff_minus) / (mu_eff + 2), (1 - c1 - cmu) / (n * cmu))
        pos = np.sum(wp[wp > 0])
        neg = np.sum(np.abs(wp[wp < 0]))
        self.weights = np.where(wp >= 0, 1 / pos * wp, min_alpha / neg * wp)
        self.cm = 1
        self.c_sigma = (mu_eff + 2) / (n + mu_eff + 5)
        self.d_sigma = 1 + 2 * max(0, math.sqrt((mu_eff - 1) / (n + 1)) - 1) + self.c_sigma
        self.cc = (4 + mu_eff / n) / (n + 4 + 2 * mu_eff / n)
        self.c1, self.cmu, self.mu_eff, self.mu, self.n, self.popsize = c1, cmu, mu_eff, mu, n, popsize
        self.chi_n = math.sqrt(n) * (1.0 - (1.0 / (4.0 * n)) + 1.0 / (21.0 * (n ** 2)))
        self.p_sigma, self.pc = np.zeros(n), np.zeros(n)
        self.mean, self.C, self.sigma, self.g = np.array(mean, float), np.eye(n), sigma, 0

    def eig(self):
        self.C = (self.C + self.C.T) / 2
        D2, B = np.linalg.eigh(self.C)
        D = np.sqrt(np.where(D2 < 0, 1e-8, D2))
        self.C = np.dot(np.dot(B, np.diag(D ** 2)), B.T)
        return B, D

    def tell(self, xs, fs):
        self.g += 1
        order = np.argsort(fs, kind="stable")
        B, D = self.eig()
        x_k = np.asarray(xs)[order]
        y_k = (x_k - self.mean) / self.sigma
        y_w = np.sum(y_k[: self.mu].T * self.weights[: self.mu], axis=1)
        self.mean = self.mean + self.cm * self.sigma * y_w
        C_2 = B.dot(np.diag(1 / D)).dot(B.T)
        self.p_sigma = (1 - self.c_sigma) * self.p_sigma + math.sqrt(
            self.c_sigma * (2 - self.c_sigma) * self.mu_eff) * C_2.dot(y_w)
        norm_p_sigma = np.linalg.norm(self.p_sigma)
        self.sigma *= np.exp((self.c_sigma / self.d_sigma) * (norm_p_sigma / self.chi_n - 1))
        self.sigma = min(self.sigma, 1e32)
        left = norm_p_sigma / math.sqrt(1 - (1 - self.c_sigma) ** (2 * (self.g + 1)))
        right = (1.4 + 2 / (self.n + 1)) * self.chi_n
        h_sigma = 1.0 if left < right else 0.0
        self.pc = (1 - self.cc) * self.pc + h_sigma * math.sqrt(self.cc * (2 - self.cc) * self.mu_eff) * y_w
        w_io = self.weights * np.where(self.weights >= 0, 1,
                                       self.n / (np.linalg.norm(C_2.dot(y_k.T), axis=0) ** 2 + 1e-8))
        delta = (1 - h_sigma) * self.cc * (2 - self.cc)
        rank_one = np.outer(self.pc, self.pc)
        rank_mu = np.sum(np.array([w * np.outer(y, y) for w, y in zip(w_io, y_k)]), axis=0)
        self.C = ((1 + self.c1 * delta - self.c1 - self.cmu * np.sum(self.weights)) * self.C
                  + self.c1 * rank_one + self.cmu * rank_mu)


def gen_cmaes():
    out = {}
    rng = np.random.Generator(np.random.PCG64(11))
    for n, lam in ((4, 8), (10, 0), (24, 48)):
        opt = CMA(np.full(n, 0.5), 1.3, popsize=lam or None)
        lam = opt.popsize
        tag = f"n{n}"
        out[f"{tag}_weights"] = opt.weights
        out[f"{tag}_consts"] = np.asarray([opt.mu_eff, opt.c1, opt.cmu, opt.c_sigma, opt.d_sigma, opt.cc, opt.chi_n,
                                           opt.cm, np.sum(opt.weights)])
        xs_all, fs_all, states = [], [], []
        for _ in range(3):  # three generations with fixed x (drawn here, not by the optimiser)
            xs = opt.mean + opt.sigma * rng.normal(size=(lam, n)) * np.linspace(0.5, 2.0, n)
            fs = np.sum((xs - 1.0) ** 2 * np.arange(1, n + 1), axis=1)
            opt.tell(xs, fs)
            xs_all.append(xs)
            fs_all.append(fs)
            states.append(np.concatenate([opt.mean, [opt.sigma], opt.C.ravel(), opt.p_sigma, opt.pc]))
        out[f"{tag}_xs"], out[f"{tag}_fs"], out[f"{tag}_states"] = np.asarray(xs_all), np.asarray(fs_all), np.asarray(states)
    np.savez_compressed(os.path.join(OUT, "cmaes_golden.npz"), **out)


if __name__ == "__main__":
    os.makedirs(OUT, exist_ok=True)
    gen_sobol()
    gen_tpe()
    gen_cmaes()
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)))
