// kb_linalg.cuh -- dense FP64 helpers shared by kb_linalg.cu and kb_cmaes.cu (not part of the ABI).
#pragma once
#include "kb_internal.cuh"

// C[M x N] = alpha * (A[M x K] * diag(s)) * Bm[N x K]^T + beta * C; row-major; s may be NULL; DMMA (mma.sync f64)
int kb_dgemm_nt(kb_ctx *ctx, int M, int N, int K, const double *A, int lda, const double *s, const double *Bm, int ldb,
                double *C, int ldc, double alpha, double beta, cudaStream_t st);
int kb_transpose(kb_ctx *ctx, int rows, int cols, const double *in, double *out, cudaStream_t st);
// symmetric eigendecomposition of (A + A^T)/2 + shift*I; w ascending (shift removed), V[i*n+j] = component i of
// eigenvector j, canonical signs; workspace of kb_eigh_workspace_bytes(n) bytes
size_t kb_eigh_workspace_bytes(int n);
// d_V0t (optional): warm start, transposed eigenvectors of a nearby matrix; sweeps_out (optional, host): sweeps used
int kb_eigh_device(kb_ctx *ctx, int n, const double *d_A, double shift, double *d_w, double *d_V, void *workspace,
                   cudaStream_t st, const double *d_V0t, int *sweeps_out);

int kb_eigh_canonical(kb_ctx *ctx, int n, const double *Vt, const double *lam, double *w, double *V, cudaStream_t st);

struct EighShared {   // device scalars of one eigensolve
    int rotated;      // rotations applied in the current sweep
    int sweeps;
};
// In-place Cholesky factor of the symmetric matrix in Gt (column-contiguous), one cooperative kernel (kb_chol.cu);
// *fail (device) is set when a pivot is not positive; *launched = false when a panel does not fit in shared memory
int kb_chol_coop(kb_ctx *ctx, int n, double *Gt, int *fail, cudaStream_t st, bool *launched);
// Gram-form block step of the one-sided Jacobi iteration (kb_eigh_gram.cu); *launched = false when the resident
// columns do not fit into shared memory; hblk: scratch of at least 8 * (n + 16) doubles
int kb_jacobi_gram(kb_ctx *ctx, int n, double *Gt, double *Vt, int *blkstep, EighShared *sh, int *sweep_rot, double *hblk, double tol,
                   int with_v, int max_sweeps, cudaStream_t st, bool *launched);
