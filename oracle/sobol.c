/*
 * oracle/sobol.c -- CPU restatement of a Joe-Kuo Sobol engine (gray-code order).
 * TEST INFRASTRUCTURE ONLY (see katib_oracle.h).  goptuna v0.8.0's sobol/ package is not under
 * /root/reference (go.mod:8; call site converter.go:94 `sobol.NewSampler()`, driven from
 * sample.go:39 `trial.CallRelativeSampler()`).  Pinned against scipy.stats.qmc.Sobol
 * (scramble=False, bits=32) for the direction-number recurrence and gray-code order
 * (tests/golden/sobol_*.npz).  The caller supplies the Joe-Kuo table, so this file embeds none.
 */
#include "katib_oracle.h"
#include <stdlib.h>

/* v[d*bits + j], already shifted so that bit (bits-1-j) is the leading one */
static int build_v(int dim, int bits, const int64_t *poly, const int64_t *vinit, uint64_t *v) {
    for (int j = 0; j < bits; j++) v[j] = 1;
    for (int d = 1; d < dim; d++) {
        uint64_t p = (uint64_t)poly[d];
        int m = 0;
        for (uint64_t t = p >> 1; t; t >>= 1) m++; /* degree */
        if (m < 1 || m > 18) return -1;
        uint64_t *vd = v + (size_t)d * bits;
        for (int j = 0; j < m && j < bits; j++) vd[j] = (uint64_t)vinit[(size_t)d * 18 + j];
        for (int j = m; j < bits; j++) {
            uint64_t nv = vd[j - m], pow2 = 1;
            for (int k = 0; k < m; k++) {
                pow2 <<= 1;
                if ((p >> (m - 1 - k)) & 1) nv ^= pow2 * vd[j - k - 1];
            }
            vd[j] = nv;
        }
    }
    for (int d = 0; d < dim; d++)
        for (int j = 0; j < bits; j++) v[(size_t)d * bits + j] <<= (bits - 1 - j);
    return 0;
}

static int fill(int dim, int bits, const int64_t *poly, const int64_t *vinit, uint64_t first,
                int64_t n, double *outf, uint32_t *outu) {
    if (dim < 1 || bits < 1 || bits > 32) return -1;
    uint64_t *v = (uint64_t *)malloc(sizeof(uint64_t) * (size_t)dim * bits);
    uint64_t *x = (uint64_t *)calloc((size_t)dim, sizeof(uint64_t));
    if (build_v(dim, bits, poly, vinit, v)) { free(v); free(x); return -1; }
    /* state at index `first`: XOR of v[b] over the set bits of gray(first) */
    uint64_t g = first ^ (first >> 1);
    for (int b = 0; b < bits; b++)
        if ((g >> b) & 1) for (int d = 0; d < dim; d++) x[d] ^= v[(size_t)d * bits + b];
    double scale = 1.0 / (double)(1ull << bits);
    for (int64_t p = 0; p < n; p++) {
        for (int d = 0; d < dim; d++) {
            if (outf) outf[(size_t)p * dim + d] = (double)x[d] * scale;
            if (outu) outu[(size_t)p * dim + d] = (uint32_t)x[d];
        }
        uint64_t idx = first + (uint64_t)p; /* advance: flip direction ctz(~idx) */
        int l = 0;
        while ((idx >> l) & 1) l++;
        if (l < bits) for (int d = 0; d < dim; d++) x[d] ^= v[(size_t)d * bits + l];
    }
    free(v); free(x);
    return 0;
}

int kbo_sobol_fill(int dim, int bits, const int64_t *poly, const int64_t *vinit, uint64_t first,
                   int64_t n, double *out) {
    return fill(dim, bits, poly, vinit, first, n, out, 0);
}
int kbo_sobol_fill_u32(int dim, const int64_t *poly, const int64_t *vinit, uint64_t first,
                       int64_t n, uint32_t *out) {
    return fill(dim, 32, poly, vinit, first, n, 0, out);
}
