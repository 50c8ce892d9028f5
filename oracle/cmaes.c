/*
 * oracle/cmaes.c -- CPU restatement of goptuna's cmaes.Optimizer (a Go port of the `cmaes`
 * library's CMA class, active CMA-ES with negative weights) and a cyclic Jacobi eigensolver
 * standing in for gonum's mat.EigenSym.  TEST INFRASTRUCTURE ONLY (see katib_oracle.h).
 * PARITY UNPINNED: goptuna v0.8.0 cmaes/ and gonum v0.8.2 are not under /root/reference
 * (go.mod:8, go.mod:141).  Reference call sites: converter.go:41-68 (cmaes.NewSampler + options),
 * sample.go:39 (trial.CallRelativeSampler -> Sampler.SampleRelative -> Optimizer.Ask/Tell),
 * service.go:104-110 (why generations must be tracked).
 *
 *   kbo_cma_create     <- cmaes.NewOptimizer (constants eq.48-56 of Hansen's tutorial)
 *   kbo_cma_eig        <- Optimizer.eigenDecomposition (C=(C+C^T)/2, eigh, D=sqrt(max(D2,eps)))
 *   kbo_cma_ask_batch  <- Optimizer.sampleSolution, batched, z supplied by the caller
 *   kbo_cma_tell       <- Optimizer.Tell
 */
#include "katib_oracle.h"
#include <math.h>
#include <stdlib.h>
#include <string.h>

#define CMA_EPS 1e-8
#define CMA_SIGMA_MAX 1e32

struct kbo_cma {
    int n, popsize, mu, g;
    double mu_eff, c1, cmu, c_sigma, d_sigma, cc, chi_n, cm;
    double sigma;
    double *weights, *mean, *C, *p_sigma, *pc, *B, *D;
    int have_eig;
    double *lo, *hi;
};

static double *dalloc(size_t k) { return (double *)calloc(k ? k : 1, sizeof(double)); }

kbo_cma *kbo_cma_create(int n, const double *mean, double sigma, int popsize,
                        const double *bounds_lo, const double *bounds_hi) {
    kbo_cma *c = (kbo_cma *)calloc(1, sizeof(kbo_cma));
    c->n = n;
    if (popsize <= 0) popsize = 4 + (int)floor(3.0 * log((double)n)); /* eq.48 */
    c->popsize = popsize;
    int mu = popsize / 2;
    c->mu = mu;
    double *wp = dalloc((size_t)popsize);
    double sb = 0, sb2 = 0, sa = 0, sa2 = 0;
    for (int i = 0; i < popsize; i++) { /* eq.49 */
        wp[i] = log(((double)popsize + 1.0) / 2.0) - log((double)(i + 1));
        if (i < mu) { sb += wp[i]; sb2 += wp[i] * wp[i]; } else { sa += wp[i]; sa2 += wp[i] * wp[i]; }
    }
    double mu_eff = sb * sb / sb2, mu_eff_minus = sa * sa / sa2;
    double alpha_cov = 2.0, dn = (double)n;
    double c1 = alpha_cov / ((dn + 1.3) * (dn + 1.3) + mu_eff);
    double cmu_b = alpha_cov * (mu_eff - 2.0 + 1.0 / mu_eff) / ((dn + 2.0) * (dn + 2.0) + alpha_cov * mu_eff / 2.0);
    double cmu = cmu_b < 1.0 - c1 - 1e-8 ? cmu_b : 1.0 - c1 - 1e-8;
    double min_alpha = 1.0 + c1 / cmu;                                       /* eq.50 */
    double a2 = 1.0 + (2.0 * mu_eff_minus) / (mu_eff + 2.0);                 /* eq.51 */
    double a3 = (1.0 - c1 - cmu) / (dn * cmu);                               /* eq.52 */
    if (a2 < min_alpha) min_alpha = a2;
    if (a3 < min_alpha) min_alpha = a3;
    double pos = 0, neg = 0;
    for (int i = 0; i < popsize; i++) { if (wp[i] > 0) pos += wp[i]; else if (wp[i] < 0) neg += fabs(wp[i]); }
    c->weights = dalloc((size_t)popsize);
    for (int i = 0; i < popsize; i++)
        c->weights[i] = wp[i] >= 0 ? 1.0 / pos * wp[i] : min_alpha / neg * wp[i];
    free(wp);
    c->cm = 1.0;
    c->mu_eff = mu_eff; c->c1 = c1; c->cmu = cmu;
    c->c_sigma = (mu_eff + 2.0) / (dn + mu_eff + 5.0);                       /* eq.55 */
    double t = sqrt((mu_eff - 1.0) / (dn + 1.0)) - 1.0;
    c->d_sigma = 1.0 + 2.0 * (t > 0 ? t : 0) + c->c_sigma;
    c->cc = (4.0 + mu_eff / dn) / (dn + 4.0 + 2.0 * mu_eff / dn);            /* eq.56 */
    c->chi_n = sqrt(dn) * (1.0 - (1.0 / (4.0 * dn)) + 1.0 / (21.0 * (dn * dn)));
    c->sigma = sigma;
    c->mean = dalloc((size_t)n); memcpy(c->mean, mean, sizeof(double) * (size_t)n);
    c->C = dalloc((size_t)n * n);
    for (int i = 0; i < n; i++) c->C[(size_t)i * n + i] = 1.0;
    c->p_sigma = dalloc((size_t)n); c->pc = dalloc((size_t)n);
    c->B = dalloc((size_t)n * n); c->D = dalloc((size_t)n);
    if (bounds_lo && bounds_hi) {
        c->lo = dalloc((size_t)n); c->hi = dalloc((size_t)n);
        memcpy(c->lo, bounds_lo, sizeof(double) * (size_t)n);
        memcpy(c->hi, bounds_hi, sizeof(double) * (size_t)n);
    }
    return c;
}

void kbo_cma_destroy(kbo_cma *c) {
    if (!c) return;
    free(c->weights); free(c->mean); free(c->C); free(c->p_sigma); free(c->pc);
    free(c->B); free(c->D); free(c->lo); free(c->hi); free(c);
}
int kbo_cma_popsize(const kbo_cma *c) { return c->popsize; }
int kbo_cma_mu(const kbo_cma *c) { return c->mu; }
int kbo_cma_generation(const kbo_cma *c) { return c->g; }

void kbo_cma_constants(const kbo_cma *c, double *out, double *weights) {
    double sw = 0;
    for (int i = 0; i < c->popsize; i++) sw += c->weights[i];
    out[0] = c->mu_eff; out[1] = c->c1; out[2] = c->cmu; out[3] = c->c_sigma; out[4] = c->d_sigma;
    out[5] = c->cc; out[6] = c->chi_n; out[7] = c->cm; out[8] = sw;
    if (weights) memcpy(weights, c->weights, sizeof(double) * (size_t)c->popsize);
}

/* Cyclic Jacobi (row sweeps) until off-diagonal mass vanishes; ascending, sign-canonical. */
void kbo_jacobi_eigh(int n, const double *Ain, double *w, double *V) {
    size_t nn = (size_t)n * n;
    double *A = (double *)malloc(sizeof(double) * nn);
    memcpy(A, Ain, sizeof(double) * nn);
    for (size_t i = 0; i < nn; i++) V[i] = 0.0;
    for (int i = 0; i < n; i++) V[(size_t)i * n + i] = 1.0;
    for (int sweep = 0; sweep < 60; sweep++) {
        double off = 0, diag = 0;
        for (int i = 0; i < n; i++) {
            diag += A[(size_t)i * n + i] * A[(size_t)i * n + i];
            for (int j = i + 1; j < n; j++) off += A[(size_t)i * n + j] * A[(size_t)i * n + j];
        }
        if (off <= 1e-32 * diag || off == 0.0) break;
        for (int p = 0; p < n - 1; p++)
            for (int q = p + 1; q < n; q++) {
                double apq = A[(size_t)p * n + q];
                if (apq == 0.0) continue;
                double app = A[(size_t)p * n + p], aqq = A[(size_t)q * n + q];
                double theta = (aqq - app) / (2.0 * apq);
                double tt = (theta >= 0 ? 1.0 : -1.0) / (fabs(theta) + sqrt(theta * theta + 1.0));
                double cs = 1.0 / sqrt(tt * tt + 1.0), sn = tt * cs;
                for (int k = 0; k < n; k++) { /* columns p,q */
                    double akp = A[(size_t)k * n + p], akq = A[(size_t)k * n + q];
                    A[(size_t)k * n + p] = cs * akp - sn * akq;
                    A[(size_t)k * n + q] = sn * akp + cs * akq;
                }
                for (int k = 0; k < n; k++) { /* rows p,q */
                    double apk = A[(size_t)p * n + k], aqk = A[(size_t)q * n + k];
                    A[(size_t)p * n + k] = cs * apk - sn * aqk;
                    A[(size_t)q * n + k] = sn * apk + cs * aqk;
                }
                for (int k = 0; k < n; k++) {
                    double vkp = V[(size_t)k * n + p], vkq = V[(size_t)k * n + q];
                    V[(size_t)k * n + p] = cs * vkp - sn * vkq;
                    V[(size_t)k * n + q] = sn * vkp + cs * vkq;
                }
            }
    }
    /* sort ascending (selection), then fix signs */
    int *ord = (int *)malloc(sizeof(int) * (size_t)n);
    for (int i = 0; i < n; i++) ord[i] = i;
    for (int i = 0; i < n; i++) {
        int m = i;
        for (int j = i + 1; j < n; j++)
            if (A[(size_t)ord[j] * n + ord[j]] < A[(size_t)ord[m] * n + ord[m]]) m = j;
        int t = ord[i]; ord[i] = ord[m]; ord[m] = t;
    }
    double *V2 = (double *)malloc(sizeof(double) * nn);
    for (int j = 0; j < n; j++) {
        int src = ord[j];
        w[j] = A[(size_t)src * n + src];
        int big = 0; double bv = 0;
        for (int i = 0; i < n; i++) {
            double v = fabs(V[(size_t)i * n + src]);
            if (v > bv) { bv = v; big = i; }
        }
        double sgn = V[(size_t)big * n + src] < 0 ? -1.0 : 1.0;
        for (int i = 0; i < n; i++) V2[(size_t)i * n + j] = sgn * V[(size_t)i * n + src];
    }
    memcpy(V, V2, sizeof(double) * nn);
    free(V2); free(ord); free(A);
}

static void ensure_eig(kbo_cma *c) {
    if (c->have_eig) return;
    int n = c->n;
    for (int i = 0; i < n; i++)
        for (int j = i + 1; j < n; j++) {
            double s = (c->C[(size_t)i * n + j] + c->C[(size_t)j * n + i]) / 2.0;
            c->C[(size_t)i * n + j] = s; c->C[(size_t)j * n + i] = s;
        }
    kbo_jacobi_eigh(n, c->C, c->D, c->B);
    for (int i = 0; i < n; i++) c->D[i] = sqrt(c->D[i] < 0 ? CMA_EPS : c->D[i]);
    /* C = B diag(D^2) B^T */
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            double s = 0;
            for (int k = 0; k < n; k++) s += c->B[(size_t)i * n + k] * c->D[k] * c->D[k] * c->B[(size_t)j * n + k];
            c->C[(size_t)i * n + j] = s;
        }
    c->have_eig = 1;
}

void kbo_cma_eig(kbo_cma *c, double *B, double *D) {
    ensure_eig(c);
    if (B) memcpy(B, c->B, sizeof(double) * (size_t)c->n * c->n);
    if (D) memcpy(D, c->D, sizeof(double) * (size_t)c->n);
}

void kbo_cma_set_eig(kbo_cma *c, const double *B, const double *D) {
    memcpy(c->B, B, sizeof(double) * (size_t)c->n * c->n);
    memcpy(c->D, D, sizeof(double) * (size_t)c->n);
    c->have_eig = 1;
}

void kbo_cma_ask_batch(kbo_cma *c, int lam, const double *z, double *x) {
    ensure_eig(c);
    int n = c->n;
    for (int k = 0; k < lam; k++)
        for (int i = 0; i < n; i++) {
            double y = 0;
            for (int j = 0; j < n; j++) y += c->B[(size_t)i * n + j] * c->D[j] * z[(size_t)k * n + j];
            x[(size_t)k * n + i] = c->mean[i] + c->sigma * y;
        }
}

typedef struct { double f; int idx; } fi_t;
static int fi_cmp(const void *a, const void *b) {
    const fi_t *x = (const fi_t *)a, *y = (const fi_t *)b;
    if (x->f < y->f) return -1;
    if (x->f > y->f) return 1;
    return (x->idx > y->idx) - (x->idx < y->idx);
}

void kbo_cma_tell(kbo_cma *c, const double *x, const double *f) {
    int n = c->n, lam = c->popsize, mu = c->mu;
    c->g += 1;
    fi_t *o = (fi_t *)malloc(sizeof(fi_t) * (size_t)lam);
    for (int k = 0; k < lam; k++) { o[k].f = f[k]; o[k].idx = k; }
    qsort(o, (size_t)lam, sizeof(fi_t), fi_cmp);
    ensure_eig(c);
    c->have_eig = 0; /* B, D are stale after this update */
    double *B = c->B, *D = c->D;

    double *y = dalloc((size_t)lam * n); /* sorted y_k = (x_k - m)/sigma */
    for (int k = 0; k < lam; k++)
        for (int i = 0; i < n; i++)
            y[(size_t)k * n + i] = (x[(size_t)o[k].idx * n + i] - c->mean[i]) / c->sigma;
    double *yw = dalloc((size_t)n);
    for (int k = 0; k < mu; k++)
        for (int i = 0; i < n; i++) yw[i] += y[(size_t)k * n + i] * c->weights[k]; /* eq.41 */
    for (int i = 0; i < n; i++) c->mean[i] += c->cm * c->sigma * yw[i];

    /* C^(-1/2) = B D^-1 B^T */
    double *C2 = dalloc((size_t)n * n);
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            double s = 0;
            for (int k = 0; k < n; k++) s += B[(size_t)i * n + k] / D[k] * B[(size_t)j * n + k];
            C2[(size_t)i * n + j] = s;
        }
    double coef = sqrt(c->c_sigma * (2.0 - c->c_sigma) * c->mu_eff);
    double norm2 = 0;
    for (int i = 0; i < n; i++) {
        double s = 0;
        for (int j = 0; j < n; j++) s += C2[(size_t)i * n + j] * yw[j];
        c->p_sigma[i] = (1.0 - c->c_sigma) * c->p_sigma[i] + coef * s;
        norm2 += c->p_sigma[i] * c->p_sigma[i];
    }
    double norm_ps = sqrt(norm2);
    c->sigma *= exp((c->c_sigma / c->d_sigma) * (norm_ps / c->chi_n - 1.0));
    if (c->sigma > CMA_SIGMA_MAX) c->sigma = CMA_SIGMA_MAX;

    double hl = norm_ps / sqrt(1.0 - pow(1.0 - c->c_sigma, 2.0 * (double)(c->g + 1)));
    double hr = (1.4 + 2.0 / ((double)n + 1.0)) * c->chi_n;
    double h_sigma = hl < hr ? 1.0 : 0.0;
    double cpc = h_sigma * sqrt(c->cc * (2.0 - c->cc) * c->mu_eff);
    for (int i = 0; i < n; i++) c->pc[i] = (1.0 - c->cc) * c->pc[i] + cpc * yw[i]; /* eq.45 */

    /* eq.46 */
    double *wio = dalloc((size_t)lam);
    double sumw = 0;
    for (int k = 0; k < lam; k++) {
        sumw += c->weights[k];
        if (c->weights[k] >= 0) { wio[k] = c->weights[k]; continue; }
        double nn2 = 0;
        for (int i = 0; i < n; i++) {
            double s = 0;
            for (int j = 0; j < n; j++) s += C2[(size_t)i * n + j] * y[(size_t)k * n + j];
            nn2 += s * s;
        }
        wio[k] = c->weights[k] * ((double)n / (nn2 + CMA_EPS));
    }
    double delta_h = (1.0 - h_sigma) * c->cc * (2.0 - c->cc);
    double keep = 1.0 + c->c1 * delta_h - c->c1 - c->cmu * sumw;
    /* eq.47 */
    for (int i = 0; i < n; i++)
        for (int j = 0; j < n; j++) {
            double rm = 0;
            for (int k = 0; k < lam; k++) rm += wio[k] * y[(size_t)k * n + i] * y[(size_t)k * n + j];
            c->C[(size_t)i * n + j] = keep * c->C[(size_t)i * n + j] + c->c1 * c->pc[i] * c->pc[j] + c->cmu * rm;
        }
    free(o); free(y); free(yw); free(C2); free(wio);
}

void kbo_cma_state(const kbo_cma *c, double *mean, double *sigma, double *C, double *p_sigma,
                   double *p_c) {
    int n = c->n;
    if (mean) memcpy(mean, c->mean, sizeof(double) * (size_t)n);
    if (sigma) *sigma = c->sigma;
    if (C) memcpy(C, c->C, sizeof(double) * (size_t)n * n);
    if (p_sigma) memcpy(p_sigma, c->p_sigma, sizeof(double) * (size_t)n);
    if (p_c) memcpy(p_c, c->pc, sizeof(double) * (size_t)n);
}
