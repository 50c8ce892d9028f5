// kb_chol.cu -- Cholesky factorisation A = L L^T in one cooperative kernel (the preconditioner of kb_eigh_device:
// one-sided Jacobi on the columns of L needs no accumulated eigenvector matrix; see kb_linalg.cu).
//
// Right-looking, panels of 32 columns.  Per panel: CTA 0 factorises it out of shared memory -- the 32 x 32 diagonal
// block by one warp with its rows in registers (shuffles, no barriers), the rows below by one thread per row (a
// triangular solve that needs no synchronisation at all) -- then the whole grid applies the rank-32 update to the
// lower triangle of the trailing matrix in 64 x 64 tiles.  Two grid barriers per panel replace the 2 launches per
// panel (a 1024-thread single-CTA kernel + a GEMM) this file's predecessor needed: n = 512 in ~0.2 ms instead of 1.2.
// Storage is the solver's: Gt[c * n + r] = element (r, c), column c contiguous; on exit the strictly upper part is 0.
#include <cooperative_groups.h>
#include <math.h>

#include <algorithm>

#include "kb_internal.cuh"
#include "kb_linalg.cuh"

namespace cg = cooperative_groups;

#define CC_THREADS 512
#define CC_PW 32
#define CC_TILE 64

// The w x w diagonal block, in place in shared memory (element (i, j) at sP[j * m + i]), by the whole CTA: two
// barriers per column, the (i, j) updates of a column step spread over the threads.  Returns false (uniformly) when
// a pivot is not positive.  sInv[c] = 1 / L[c][c].
__device__ __forceinline__ bool cc_diag_block(double *sP, int m, int w, int tid, double *sInv, double *sDiag) {
    for (int c = 0; c < w; c++) {
        const double piv = sP[(size_t)c * m + c];  // not written during this step (the diagonal is stored at the end)
        if (!(piv > 0.0)) return false;
        const double inv = rsqrt(piv);
        if (tid > c && tid < w) sP[(size_t)c * m + tid] *= inv;
        if (tid == c) { sInv[c] = inv; sDiag[c] = piv * inv; }
        __syncthreads();
        for (int e = tid; e < CC_PW * CC_PW; e += CC_THREADS) {
            const int i = e % CC_PW, j = e / CC_PW;
            if (j > c && i >= j && i < w) sP[(size_t)j * m + i] = fma(-sP[(size_t)c * m + i], sP[(size_t)c * m + j], sP[(size_t)j * m + i]);
        }
        __syncthreads();
    }
    if (tid < w) sP[(size_t)tid * m + tid] = sDiag[tid];
    return true;
}

#ifdef KB_CHOL_DEBUG
__device__ long long cc_dbg[8];
#define CC_MARK(i) do { if (tid == 0 && blockIdx.x == 0) { long long now_ = clock64(); cc_dbg[i] += now_ - mark_; mark_ = now_; } } while (0)
extern "C" void kb_dbg_chol_phases(long long *out, int reset) {
    cudaMemcpyFromSymbol(out, cc_dbg, sizeof(long long) * 8);
    if (reset) { long long z[8] = {0}; cudaMemcpyToSymbol(cc_dbg, z, sizeof(z)); }
}
#else
#define CC_MARK(i) do {} while (0)
#endif

__global__ void __launch_bounds__(CC_THREADS, 1)
chol_coop_kernel(int n, double *Gt, int *fail) {
    extern __shared__ __align__(16) double cc_smem[];
    __shared__ double sInv[CC_PW], sDiag[CC_PW];
    __shared__ int s_bad;
    cg::grid_group grid = cg::this_grid();
    const int tid = threadIdx.x;
#ifdef KB_CHOL_DEBUG
    long long mark_ = clock64();
#endif
    for (int j0 = 0; j0 < n; j0 += CC_PW) {
        const int w = min(CC_PW, n - j0), m = n - j0;
        if (blockIdx.x == 0) {
            double *sP = cc_smem;  // [w][m]
            if (tid == 0) s_bad = 0;
            for (int i = tid; i < m; i += CC_THREADS) {  // all of a row's loads in flight before the first store
                double v[CC_PW];
#pragma unroll
                for (int c = 0; c < CC_PW; c++) v[c] = c < w ? __ldcg(Gt + (size_t)(j0 + c) * n + j0 + i) : 0.0;
#pragma unroll
                for (int c = 0; c < CC_PW; c++)
                    if (c < w) sP[(size_t)c * m + i] = v[c];
            }
            __syncthreads();
            CC_MARK(0);
            if (!cc_diag_block(sP, m, w, tid, sInv, sDiag) && tid == 0) s_bad = 1;
            __syncthreads();
            CC_MARK(1);
            if (s_bad) {
                if (tid == 0) *fail = 1;
            } else {
                // rows below the diagonal block: x L11^T = a, one thread per row, everything in registers
                for (int i = w + tid; i < m; i += CC_THREADS) {
                    // right-looking: once x_k is known every later entry is updated at once, so the dependent chain
                    // is 32 (mul, fma) pairs and the 31..1 updates of a step are independent of each other
                    double x[CC_PW];
#pragma unroll
                    for (int c = 0; c < CC_PW; c++) x[c] = c < w ? sP[(size_t)c * m + i] : 0.0;
#pragma unroll
                    for (int k = 0; k < CC_PW; k++) {
                        if (k < w) {
                            x[k] *= sInv[k];
#pragma unroll
                            for (int c = k + 1; c < CC_PW; c++)
                                if (c < w) x[c] = fma(-x[k], sP[(size_t)k * m + c], x[c]);  // L[c][k]
                        }
                    }
#pragma unroll
                    for (int c = 0; c < CC_PW; c++)
                        if (c < w) sP[(size_t)c * m + i] = x[c];
                }
                __syncthreads();
                CC_MARK(2);
                for (int i = tid; i < m; i += CC_THREADS) {
#pragma unroll 8
                    for (int c = 0; c < w; c++)
                        __stcg(Gt + (size_t)(j0 + c) * n + j0 + i, i >= c ? sP[(size_t)c * m + i] : 0.0);
                }
                for (int i = tid; i < j0; i += CC_THREADS) {  // strictly upper part of L
#pragma unroll 8
                    for (int c = 0; c < w; c++) __stcg(Gt + (size_t)(j0 + c) * n + i, 0.0);
                }
            }
            __threadfence();
            CC_MARK(3);
        }
        grid.sync();
        CC_MARK(4);
        if (__ldcg(fail)) return;  // every CTA reads the same value after the barrier
        const int t0 = j0 + w, mt = n - t0;  // trailing matrix: rows / columns t0 .. n-1
        if (mt <= 0) break;
        // ---- rank-w update of the lower triangle: T[r][c] -= sum_k L[r][j0+k] L[c][j0+k]
        double *sA = cc_smem, *sB = cc_smem + CC_PW * CC_TILE;  // [w][64] each: panel rows of the tile's rows / columns
        const int nt = (mt + CC_TILE - 1) / CC_TILE;
        const int n_tiles = nt * (nt + 1) / 2;
        for (int tile = blockIdx.x; tile < n_tiles; tile += gridDim.x) {
            int tr = (int)((sqrt(8.0 * tile + 1.0) - 1.0) * 0.5);  // row-major enumeration of the lower triangle
            while ((tr + 1) * (tr + 2) / 2 <= tile) tr++;
            while (tr * (tr + 1) / 2 > tile) tr--;
            const int tc = tile - tr * (tr + 1) / 2;
            const int r0 = t0 + tr * CC_TILE, c0 = t0 + tc * CC_TILE;
            __syncthreads();
            for (int e = tid; e < w * CC_TILE; e += CC_THREADS) {
                const int k = e / CC_TILE, x = e % CC_TILE;
                sA[k * CC_TILE + x] = r0 + x < n ? __ldcg(Gt + (size_t)(j0 + k) * n + r0 + x) : 0.0;
                sB[k * CC_TILE + x] = c0 + x < n ? __ldcg(Gt + (size_t)(j0 + k) * n + c0 + x) : 0.0;
            }
            __syncthreads();
            const int tx = tid % CC_TILE, ty = tid / CC_TILE;  // row tx, columns ty*8 .. ty*8+7
            const int r = r0 + tx;
            double acc[8], old[8];  // the old values are requested first so that they arrive during the products
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const int c = c0 + ty * 8 + e;
                acc[e] = 0.0;
                old[e] = (r < n && c < n && c <= r) ? __ldcg(Gt + (size_t)c * n + r) : 0.0;
            }
            for (int k = 0; k < w; k++) {
                const double a = sA[k * CC_TILE + tx];
#pragma unroll
                for (int e = 0; e < 8; e++) acc[e] = fma(a, sB[k * CC_TILE + ty * 8 + e], acc[e]);
            }
            if (r < n) {
#pragma unroll
                for (int e = 0; e < 8; e++) {
                    const int c = c0 + ty * 8 + e;
                    if (c < n && c <= r) __stcg(Gt + (size_t)c * n + r, old[e] - acc[e]);
                }
            }
        }
        CC_MARK(5);
        __threadfence();
        grid.sync();
        CC_MARK(6);
    }
}

// *launched = false when the panel does not fit into shared memory (the caller keeps its multi-launch path)
int kb_chol_coop(kb_ctx *ctx, int n, double *Gt, int *fail, cudaStream_t st, bool *launched) {
    *launched = false;
    if (getenv("KB_CHOL_COOP") && atoi(getenv("KB_CHOL_COOP")) == 0) return KB_OK;
    const size_t smem = sizeof(double) * std::max((size_t)CC_PW * n, (size_t)2 * CC_PW * CC_TILE);
    if (smem > (size_t)ctx->prop.sharedMemPerBlockOptin) return KB_OK;
    KB_CUDA(ctx, cudaFuncSetAttribute(chol_coop_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    int per_sm = 0;
    KB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, chol_coop_kernel, CC_THREADS, smem));
    if (per_sm < 1) return KB_OK;
    const int nt = (n + CC_TILE - 1) / CC_TILE;
    int blocks = std::max(1, std::min(nt * (nt + 1) / 2, per_sm * ctx->prop.multiProcessorCount));
    void *args[] = {&n, &Gt, &fail};
    KB_CUDA(ctx, cudaLaunchCooperativeKernel((void *)chol_coop_kernel, dim3((unsigned)blocks), dim3(CC_THREADS), args, smem, st));
    ctx->launches++;
    *launched = true;
    return KB_OK;
}
