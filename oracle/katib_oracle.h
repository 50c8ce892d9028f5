/*
 * katib_oracle.h -- CPU restatement (FP64, scalar C) of the arithmetic behind Katib's
 * goptuna suggestion service.  TEST INFRASTRUCTURE ONLY.
 *
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference leg may
 * link or call anything in oracle/.  The product (katib_b200/, include/) never does.
 *
 * PARITY UNPINNED (TPE, CMA-ES): the arithmetic lives in the third-party Go module
 * github.com/c-bata/goptuna v0.8.0 (reference go.mod:8, go.sum:131-132) with the eigensolver in
 * gonum.org/v1/gonum v0.8.2 (go.mod:141).  Neither module is vendored under /root/reference, no Go
 * toolchain exists in this image, and the reference's own tests for the path
 * (pkg/suggestion/v1beta1/goptuna/service_test.go:160-184, converter_test.go:27-193) assert
 * counts, status codes and type mapping only -- no number.  What follows restates the PUBLISHED
 * algorithms those modules port (Optuna 1.x TPESampler/ParzenEstimator, the `cmaes` library's
 * CMA class, Joe-Kuo Sobol), anchored on the reference's call sites:
 *   converter.go:41-68 (cmaes options)  converter.go:70-92 (tpe options)  converter.go:94 (sobol)
 *   sample.go:39 (relative sampler)     sample.go:48-90 (Suggest* per distribution kind)
 *   service.go:120-141 (only Complete trials carry a value into storage)
 * Rules that could not be checked against module source sit behind the KBO_* switches below.
 * Sobol IS pinned for its direction numbers and gray-code order against scipy's Joe-Kuo
 * generator (tests/golden/sobol_*.npz); the first-point convention stays a switch.
 */
#ifndef KATIB_ORACLE_H
#define KATIB_ORACLE_H
#include <stdint.h>
#ifdef __cplusplus
extern "C" {
#endif

#define KBO_EPS 1e-12 /* goptuna tpe `eps` (Optuna EPS) */

/* ---- Parzen estimator options: goptuna tpe.NewSampler defaults [RECALLED] ------------------ */
typedef struct {
    int consider_prior;     /* 1 */
    double prior_weight;    /* 1.0 */
    int consider_magic_clip;/* 1 */
    int consider_endpoints; /* 0 */
} kbo_parzen_opts;
kbo_parzen_opts kbo_parzen_defaults(void);

/* gamma(n) = min(ceil(0.1 n), 25); weights ramp (tpe.DefaultGamma / tpe.DefaultWeights) */
int kbo_tpe_gamma(int n);
void kbo_tpe_weights(int n, double *w);

/* Split one parameter's observations by loss (already sign-adjusted so lower is better).
 * Stable argsort by loss; below = the n_below best in trial order, above = the rest in trial
 * order.  Returns n_below.  below/above must hold n doubles each. */
int kbo_tpe_split(int n, const double *values, const double *loss, double *below, double *above);

/* Build the estimator over `n` observations on [low, high]; writes K = n+1 (prior inserted)
 * entries, sorted by mu.  Returns K. */
int kbo_parzen_build(const double *obs, int n, double low, double high, kbo_parzen_opts o,
                     double *w, double *mu, double *sigma);

double kbo_normal_cdf(double x, double mu, double sigma);

/* log pdf of the truncated mixture at nx points.  q <= 0: continuous; q > 0: quantised. */
void kbo_gmm_log_pdf(const double *x, int64_t nx, int K, const double *w, const double *mu,
                     const double *sigma, double low, double high, double q, double *out);

/* scores[i] = log l(x_i) - log g(x_i) for one numeric parameter; returns argmax (first max). */
int64_t kbo_tpe_score_numeric(const double *x, int64_t nx, int n, const double *values,
                              const double *loss, double low, double high, double q,
                              double *scores /* nullable */);

/* Categorical: weighted bincount + prior weight, normalised; scores = log pl[c] - log pg[c]. */
void kbo_tpe_categorical_tables(int n, const double *values, const double *loss, int n_choices,
                                double prior_weight, double *p_below, double *p_above);
int64_t kbo_tpe_score_categorical(const double *x, int64_t nx, int n, const double *values,
                                  const double *loss, int n_choices, double *scores);

/* Draw `size` samples from the truncated mixture l(x) (component pick by weights, normal draw,
 * rejected until inside [low, high); then rounded to q if q > 0).  RNG = splitmix64/Box-Muller,
 * NOT Go's math/rand: goptuna's stream cannot be reproduced, so only the distribution is
 * comparable. */
void kbo_gmm_sample(uint64_t seed, int64_t size, int K, const double *w, const double *mu,
                    const double *sigma, double low, double high, double q, double *out);

/* ---- Sobol ---------------------------------------------------------------------------------- */
/* Joe-Kuo direction numbers supplied by the caller (poly[dim], vinit[dim][18], scipy layout).
 * out[p*dim + d] = x_{first+p, d} / 2^bits, gray-code order, point 0 = origin. */
int kbo_sobol_fill(int dim, int bits, const int64_t *poly, const int64_t *vinit, uint64_t first,
                   int64_t n, double *out);
/* raw integer state, same indexing */
int kbo_sobol_fill_u32(int dim, const int64_t *poly, const int64_t *vinit, uint64_t first,
                       int64_t n, uint32_t *out);

/* ---- CMA-ES (the `cmaes` library's CMA class as ported by goptuna/cmaes) [RECALLED] --------- */
typedef struct kbo_cma kbo_cma;
kbo_cma *kbo_cma_create(int n, const double *mean, double sigma, int popsize /*0=default*/,
                        const double *bounds_lo, const double *bounds_hi /* nullable */);
void kbo_cma_destroy(kbo_cma *c);
int kbo_cma_popsize(const kbo_cma *c);
int kbo_cma_mu(const kbo_cma *c);
int kbo_cma_generation(const kbo_cma *c);
/* constants: out[0..8] = mu_eff, c1, cmu, c_sigma, d_sigma, cc, chi_n, cm, sum(weights) */
void kbo_cma_constants(const kbo_cma *c, double *out, double *weights /* popsize, nullable */);
/* eigendecomposition of the current C, canonicalised (ascending eigenvalues, each column's
 * largest-|.| entry positive).  B column-major? no: row-major B[i*n+j] = component i of vector j */
void kbo_cma_eig(kbo_cma *c, double *B, double *D);
/* replace the cached (B, D) (used to feed another solver's factors for sample parity) */
void kbo_cma_set_eig(kbo_cma *c, const double *B, const double *D);
/* x[k*n + i] = mean_i + sigma * (B diag(D) z_k)_i for lam columns z[k*n + i]; no resampling */
void kbo_cma_ask_batch(kbo_cma *c, int lam, const double *z, double *x);
/* x: popsize x n row-major, f: popsize values (lower is better) */
void kbo_cma_tell(kbo_cma *c, const double *x, const double *f);
void kbo_cma_state(const kbo_cma *c, double *mean, double *sigma, double *C, double *p_sigma,
                   double *p_c);
/* cyclic Jacobi on a symmetric n x n matrix (row-major); eigenvalues ascending in w,
 * eigenvectors in columns of V (row-major V[i*n+j]) */
void kbo_jacobi_eigh(int n, const double *A, double *w, double *V);

#ifdef __cplusplus
}
#endif
#endif
