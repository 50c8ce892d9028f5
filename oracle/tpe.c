/*
 * oracle/tpe.c -- CPU restatement of goptuna's independent (univariate) TPE sampler.
 * TEST INFRASTRUCTURE ONLY (see katib_oracle.h).  PARITY UNPINNED: goptuna v0.8.0's tpe/ package
 * is not under /root/reference (go.mod:8); this follows the published Optuna-1.x algorithm that
 * package ports.  Reference call sites: converter.go:70-92 (tpe.NewSampler + options),
 * sample.go:48-90 (one Sampler.Sample per parameter), service.go:120-141 (what reaches storage).
 *
 *   kbo_tpe_gamma / kbo_tpe_weights   <- tpe.DefaultGamma, tpe.DefaultWeights
 *   kbo_tpe_split                     <- Sampler.splitObservationPairs
 *   kbo_parzen_build                  <- tpe.NewParzenEstimator / ParzenEstimator.calculate
 *   kbo_normal_cdf                    <- Sampler.normalCDF
 *   kbo_gmm_log_pdf                   <- Sampler.gmmLogPDF (+ logsumRows)
 *   kbo_tpe_score_numeric             <- Sampler.sampleNumerical steps after the draw + compare
 *   kbo_tpe_categorical_*             <- Sampler.sampleCategorical + categoricalLogPDF
 *   kbo_gmm_sample                    <- Sampler.sampleFromGMM (distribution only)
 */
#include "katib_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

kbo_parzen_opts kbo_parzen_defaults(void) {
    kbo_parzen_opts o = {1, 1.0, 1, 0};
    return o;
}

int kbo_tpe_gamma(int n) {
    int a = (int)ceil(0.1 * (double)n);
    return a > 25 ? 25 : a;
}

void kbo_tpe_weights(int n, double *w) {
    if (n <= 0) return;
    if (n < 25) {
        for (int i = 0; i < n; i++) w[i] = 1.0;
        return;
    }
    int r = n - 25; /* linspace(1/n, 1, n-25, endpoint) then 25 ones */
    if (r == 1) {
        w[0] = 1.0 / (double)n;
    } else if (r > 1) {
        double start = 1.0 / (double)n;
        double step = (1.0 - start) / (double)(r - 1);
        for (int i = 0; i < r; i++) w[i] = start + (double)i * step;
    }
    for (int i = r; i < n; i++) w[i] = 1.0;
}

/* stable argsort helpers ----------------------------------------------------------------------- */
typedef struct { double key; int idx; } kv_t;
static int kv_cmp(const void *a, const void *b) {
    const kv_t *x = (const kv_t *)a, *y = (const kv_t *)b;
    if (x->key < y->key) return -1;
    if (x->key > y->key) return 1;
    return (x->idx > y->idx) - (x->idx < y->idx); /* ties: trial order (stable) */
}
static int int_cmp(const void *a, const void *b) {
    int x = *(const int *)a, y = *(const int *)b;
    return (x > y) - (x < y);
}

int kbo_tpe_split(int n, const double *values, const double *loss, double *below, double *above) {
    int nb = kbo_tpe_gamma(n);
    kv_t *kv = (kv_t *)malloc(sizeof(kv_t) * (size_t)(n > 0 ? n : 1));
    int *ix = (int *)malloc(sizeof(int) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) { kv[i].key = loss[i]; kv[i].idx = i; }
    qsort(kv, (size_t)n, sizeof(kv_t), kv_cmp);
    for (int i = 0; i < n; i++) ix[i] = kv[i].idx;
    qsort(ix, (size_t)nb, sizeof(int), int_cmp);            /* below, back in trial order */
    qsort(ix + nb, (size_t)(n - nb), sizeof(int), int_cmp); /* above, back in trial order */
    for (int i = 0; i < nb; i++) below[i] = values[ix[i]];
    for (int i = nb; i < n; i++) above[i - nb] = values[ix[i]];
    free(kv); free(ix);
    return nb;
}

int kbo_parzen_build(const double *obs, int n, double low, double high, kbo_parzen_opts o,
                     double *w, double *mu, double *sigma) {
    if (n == 0) o.consider_prior = 1;
    int K = n + (o.consider_prior ? 1 : 0);
    double prior_mu = 0.5 * (low + high);
    double prior_sigma = 1.0 * (high - low);
    int prior_pos = -1;
    double *uw = (double *)malloc(sizeof(double) * (size_t)(n > 0 ? n : 1));
    kv_t *kv = (kv_t *)malloc(sizeof(kv_t) * (size_t)(n > 0 ? n : 1));
    for (int i = 0; i < n; i++) { kv[i].key = obs[i]; kv[i].idx = i; }
    qsort(kv, (size_t)n, sizeof(kv_t), kv_cmp);
    kbo_tpe_weights(n, uw);

    if (o.consider_prior) {
        /* numpy.searchsorted(..., side='left'): first i with ordered[i] >= prior_mu */
        prior_pos = 0;
        while (prior_pos < n && kv[prior_pos].key < prior_mu) prior_pos++;
        for (int i = 0; i < prior_pos; i++) { mu[i] = kv[i].key; w[i] = uw[kv[i].idx]; }
        mu[prior_pos] = prior_mu; w[prior_pos] = o.prior_weight;
        for (int i = prior_pos; i < n; i++) { mu[i + 1] = kv[i].key; w[i + 1] = uw[kv[i].idx]; }
    } else {
        for (int i = 0; i < n; i++) { mu[i] = kv[i].key; w[i] = uw[kv[i].idx]; }
    }

    if (n == 0) {
        sigma[0] = prior_sigma;
    } else {
        for (int i = 0; i < K; i++) {
            double left = (i == 0) ? low : mu[i - 1];
            double right = (i == K - 1) ? high : mu[i + 1];
            double a = mu[i] - left, b = right - mu[i];
            sigma[i] = a > b ? a : b;
        }
        if (!o.consider_endpoints && K >= 2) { /* low_sorted_mus_high.size > 2 always holds here */
            sigma[0] = mu[1] - mu[0];
            sigma[K - 1] = mu[K - 1] - mu[K - 2];
        }
    }

    double s = 0.0;
    for (int i = 0; i < K; i++) s += w[i]; /* sequential, like a Go for-loop */
    for (int i = 0; i < K; i++) w[i] /= s;

    double maxsigma = 1.0 * (high - low);
    double minsigma;
    if (o.consider_magic_clip) {
        double d = 1.0 + (double)K;
        if (d > 100.0) d = 100.0;
        minsigma = 1.0 * (high - low) / d;
    } else {
        minsigma = KBO_EPS;
    }
    for (int i = 0; i < K; i++) {
        if (sigma[i] < minsigma) sigma[i] = minsigma;
        if (sigma[i] > maxsigma) sigma[i] = maxsigma;
    }
    if (o.consider_prior) sigma[prior_pos] = prior_sigma;
    free(uw); free(kv);
    return K;
}

double kbo_normal_cdf(double x, double mu, double sigma) {
    double num = x - mu;
    double den = sqrt(2.0) * sigma;
    if (den < KBO_EPS) den = KBO_EPS;
    return 0.5 * (1.0 + erf(num / den));
}

void kbo_gmm_log_pdf(const double *x, int64_t nx, int K, const double *w, const double *mu,
                     const double *sigma, double low, double high, double q, double *out) {
    double p_accept = 0.0;
    for (int j = 0; j < K; j++)
        p_accept += w[j] * (kbo_normal_cdf(high, mu[j], sigma[j]) - kbo_normal_cdf(low, mu[j], sigma[j]));

    if (q <= 0.0) {
        double *logc = (double *)malloc(sizeof(double) * (size_t)K);
        double *isig = (double *)malloc(sizeof(double) * (size_t)K);
        for (int j = 0; j < K; j++) {
            double Z = sqrt(2.0 * M_PI) * sigma[j];
            logc[j] = log(w[j] / Z / p_accept);
            isig[j] = sigma[j] > KBO_EPS ? sigma[j] : KBO_EPS;
        }
        double *e = (double *)malloc(sizeof(double) * (size_t)K);
        for (int64_t i = 0; i < nx; i++) {
            double m = -INFINITY;
            for (int j = 0; j < K; j++) {
                double d = (x[i] - mu[j]) / isig[j];
                e[j] = -0.5 * (d * d) + logc[j];
                if (e[j] > m) m = e[j];
            }
            double s = 0.0;
            for (int j = 0; j < K; j++) s += exp(e[j] - m);
            out[i] = log(s) + m;
        }
        free(e); free(logc); free(isig);
    } else {
        for (int64_t i = 0; i < nx; i++) {
            double ub = x[i] + q / 2.0; if (ub > high) ub = high;
            double lb = x[i] - q / 2.0; if (lb < low) lb = low;
            double p = 0.0;
            for (int j = 0; j < K; j++)
                p += w[j] * (kbo_normal_cdf(ub, mu[j], sigma[j]) - kbo_normal_cdf(lb, mu[j], sigma[j]));
            out[i] = log(p + KBO_EPS) - log(p_accept + KBO_EPS);
        }
    }
}

int64_t kbo_tpe_score_numeric(const double *x, int64_t nx, int n, const double *values,
                              const double *loss, double low, double high, double q,
                              double *scores) {
    size_t cap = (size_t)(n + 1);
    double *below = (double *)malloc(sizeof(double) * cap), *above = (double *)malloc(sizeof(double) * cap);
    double *w = (double *)malloc(sizeof(double) * cap), *mu = (double *)malloc(sizeof(double) * cap);
    double *sg = (double *)malloc(sizeof(double) * cap);
    double *ll = (double *)malloc(sizeof(double) * (size_t)(nx > 0 ? nx : 1));
    double *lg = (double *)malloc(sizeof(double) * (size_t)(nx > 0 ? nx : 1));
    kbo_parzen_opts o = kbo_parzen_defaults();
    int nb = kbo_tpe_split(n, values, loss, below, above);
    int K = kbo_parzen_build(below, nb, low, high, o, w, mu, sg);
    kbo_gmm_log_pdf(x, nx, K, w, mu, sg, low, high, q, ll);
    K = kbo_parzen_build(above, n - nb, low, high, o, w, mu, sg);
    kbo_gmm_log_pdf(x, nx, K, w, mu, sg, low, high, q, lg);
    int64_t best = 0; double bs = -INFINITY;
    for (int64_t i = 0; i < nx; i++) {
        double s = ll[i] - lg[i];
        if (scores) scores[i] = s;
        if (i == 0 || s > bs) { bs = s; best = i; } /* first maximum wins */
    }
    free(below); free(above); free(w); free(mu); free(sg); free(ll); free(lg);
    return best;
}

void kbo_tpe_categorical_tables(int n, const double *values, const double *loss, int n_choices,
                                double prior_weight, double *p_below, double *p_above) {
    size_t cap = (size_t)(n + 1);
    double *below = (double *)malloc(sizeof(double) * cap), *above = (double *)malloc(sizeof(double) * cap);
    double *wt = (double *)malloc(sizeof(double) * cap);
    int nb = kbo_tpe_split(n, values, loss, below, above);
    for (int side = 0; side < 2; side++) {
        const double *v = side ? above : below;
        int m = side ? n - nb : nb;
        double *p = side ? p_above : p_below;
        kbo_tpe_weights(m, wt);
        for (int c = 0; c < n_choices; c++) p[c] = 0.0;
        for (int i = 0; i < m; i++) p[(int)v[i]] += wt[i]; /* weighted bincount, trial order */
        double s = 0.0;
        for (int c = 0; c < n_choices; c++) { p[c] += prior_weight; s += p[c]; }
        for (int c = 0; c < n_choices; c++) p[c] /= s;
    }
    free(below); free(above); free(wt);
}

int64_t kbo_tpe_score_categorical(const double *x, int64_t nx, int n, const double *values,
                                  const double *loss, int n_choices, double *scores) {
    double *pl = (double *)malloc(sizeof(double) * (size_t)n_choices);
    double *pg = (double *)malloc(sizeof(double) * (size_t)n_choices);
    kbo_tpe_categorical_tables(n, values, loss, n_choices, 1.0, pl, pg);
    int64_t best = 0; double bs = -INFINITY;
    for (int64_t i = 0; i < nx; i++) {
        int c = (int)x[i];
        double s = log(pl[c]) - log(pg[c]);
        if (scores) scores[i] = s;
        if (i == 0 || s > bs) { bs = s; best = i; }
    }
    free(pl); free(pg);
    return best;
}

/* splitmix64 + Box-Muller: a documented stand-in stream (Go's math/rand is not reproducible). */
static uint64_t sm64(uint64_t *s) {
    uint64_t z = (*s += 0x9E3779B97F4A7C15ull);
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return z ^ (z >> 31);
}
static double u01(uint64_t *s) { return (double)(sm64(s) >> 11) * (1.0 / 9007199254740992.0); }

void kbo_gmm_sample(uint64_t seed, int64_t size, int K, const double *w, const double *mu,
                    const double *sigma, double low, double high, double q, double *out) {
    uint64_t st = seed;
    int64_t got = 0;
    while (got < size) {
        double u = u01(&st), acc = 0.0;
        int a = K - 1;
        for (int j = 0; j < K; j++) { acc += w[j]; if (u < acc) { a = j; break; } }
        double u1 = u01(&st), u2 = u01(&st);
        if (u1 < 1e-300) u1 = 1e-300;
        double z = sqrt(-2.0 * log(u1)) * cos(2.0 * M_PI * u2);
        double d = mu[a] + sigma[a] * z;
        if (low <= d && d < high) out[got++] = d;
    }
    if (q > 0.0)
        for (int64_t i = 0; i < size; i++) out[i] = round(out[i] / q) * q; /* Go math.Round: half away from zero */
}
