// kb_linalg.cu -- the two dense FP64 building blocks of the CMA-ES path.
//
//  kb_dgemm_nt   C[M x N] = alpha * (A[M x K] * diag(s)) * Bm[N x K]^T + beta * C      (row-major, K contiguous in
//                both operands).  FP64 tensor cores: mma.sync.m8n8k4.f64 (DMMA) -- tcgen05 has no FP64 kind, so on
//                sm_100a this is the tensor path for the precision north_star demands (1e-9 parity).  Every dense
//                contraction of ask/tell is this one form (see kb_cmaes.cu).
//  kb_eigh_device  symmetric eigendecomposition by one-sided (Hestenes) parallel Jacobi, standing in for gonum's
//                mat.EigenSym behind goptuna's cmaes.Optimizer.eigenDecomposition (module not in the reference tree,
//                go.mod:141; reached from sample.go:39).  Columns of G = A are orthogonalised by plane rotations in
//                round-robin (chess-tournament) order, n/2 disjoint pairs per step, one warp per pair; V accumulates
//                the rotations.  At convergence G = V Lambda.  One CTA (block barrier) for small n, a cooperative
//                grid (grid.sync) above.
#include <cooperative_groups.h>
#include <math.h>

#include <algorithm>

#include "kb_internal.cuh"
#include "kb_linalg.cuh"

namespace cg = cooperative_groups;

// ---------------------------------------------------------------------------------------------------------------
// DMMA GEMM, NT form.  CTA tile 64 x 64, 4 warps (2 x 2), warp tile 32 x 32 = 4 x 4 m8n8k4 tiles, BK = 16.
// ---------------------------------------------------------------------------------------------------------------
#define GM_BM 64
#define GM_BN 64
#define GM_BK 16
#define GM_LD (GM_BK + 1)  // padded leading dimension (doubles): rows 17 apart -> fragment loads hit distinct banks

__device__ __forceinline__ void dmma_m8n8k4(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}

__global__ void __launch_bounds__(128)
dgemm_nt_kernel(int M, int N, int K, const double *__restrict__ A, int lda, const double *__restrict__ s,
                const double *__restrict__ Bm, int ldb, double *__restrict__ C, int ldc, double alpha, double beta) {
    __shared__ double As[2][GM_BM][GM_LD];
    __shared__ double Bs[2][GM_BN][GM_LD];
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.y * GM_BM, n0 = blockIdx.x * GM_BN;
    const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;
    const int g = lane >> 2, t = lane & 3;
    double acc[4][4][2];
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    // each thread moves 8 doubles of A and 8 of Bm per stage: row = tid / 2, 8 consecutive k
    const int lr = tid >> 1, lk = (tid & 1) * 8;
    auto load_stage = [&](int buf, int k0) {
#pragma unroll
        for (int e = 0; e < 8; e++) {
            const int k = k0 + lk + e;
            double a = 0.0, b = 0.0;
            if (k < K) {
                if (m0 + lr < M) { a = A[(size_t)(m0 + lr) * lda + k]; if (s) a *= s[k]; }
                if (n0 + lr < N) b = Bm[(size_t)(n0 + lr) * ldb + k];
            }
            As[buf][lr][lk + e] = a;
            Bs[buf][lr][lk + e] = b;
        }
    };
    const int nk = (K + GM_BK - 1) / GM_BK;
    load_stage(0, 0);
    __syncthreads();
    for (int kt = 0; kt < nk; kt++) {
        const int buf = kt & 1;
        if (kt + 1 < nk) load_stage(buf ^ 1, (kt + 1) * GM_BK);
#pragma unroll
        for (int kk = 0; kk < GM_BK; kk += 4) {
            double af[4], bf[4];
#pragma unroll
            for (int i = 0; i < 4; i++) af[i] = As[buf][wm + i * 8 + g][kk + t];
#pragma unroll
            for (int j = 0; j < 4; j++) bf[j] = Bs[buf][wn + j * 8 + g][kk + t];
#pragma unroll
            for (int i = 0; i < 4; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) dmma_m8n8k4(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }
        __syncthreads();
    }
    // C fragment: row = g, cols = 2t, 2t+1
#pragma unroll
    for (int i = 0; i < 4; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int r = m0 + wm + i * 8 + g, c = n0 + wn + j * 8 + 2 * t;
            if (r < M) {
#pragma unroll
                for (int e = 0; e < 2; e++)
                    if (c + e < N) {
                        double *p = C + (size_t)r * ldc + c + e;
                        *p = beta == 0.0 ? alpha * acc[i][j][e] : alpha * acc[i][j][e] + beta * *p;
                    }
            }
        }
}

// ---------------------------------------------------------------------------------------------------------------
// the tuned path: CTA tile 128 x 128, 8 warps (2 x 4), warp tile 64 x 32 = 8 x 4 m8n8k4 tiles, BK = 16, three
// cp.async (LDGSTS, 16 B) stages in dynamic shared memory, rows padded to 20 doubles so that both fragment loads are
// bank-conflict free, optional split-K (grid.z) with FP64 atomics when the output has too few tiles to fill 148 SMs.
// Requires even K / lda / ldb and 16-byte aligned operands; anything else takes dgemm_nt_kernel above.
// ---------------------------------------------------------------------------------------------------------------
#define GF_BM 128
#define GF_BN 128
#define GF_BK 16
#define GF_LD 20
#define GF_STAGES 3
#define GF_THREADS 256
#define GF_SMEM (GF_STAGES * (GF_BM + GF_BN) * GF_LD * 8 + GF_STAGES * GF_BK * 8)

__device__ __forceinline__ void cp_async16(void *smem, const void *gmem, int src_bytes) {
    const unsigned int sa = (unsigned int)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(sa), "l"(gmem), "r"(src_bytes));
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

__global__ void __launch_bounds__(GF_THREADS, 1)
dgemm_nt_fast_kernel(int M, int N, int K, const double *__restrict__ A, int lda, const double *__restrict__ s,
                     const double *__restrict__ Bm, int ldb, double *__restrict__ C, int ldc, double alpha, double beta,
                     int k_per_split) {
    extern __shared__ __align__(16) double gsm[];
    double *As = gsm;                                        // [STAGES][BM][LD]
    double *Bs = gsm + GF_STAGES * GF_BM * GF_LD;            // [STAGES][BN][LD]
    double *Ss = Bs + GF_STAGES * GF_BN * GF_LD;             // [STAGES][BK] scale slice
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int m0 = blockIdx.y * GF_BM, n0 = blockIdx.x * GF_BN;
    const int kbeg = blockIdx.z * k_per_split, kend = min(K, kbeg + k_per_split);
    const int wm = (warp >> 2) * 64, wn = (warp & 3) * 32;
    const int g = lane >> 2, t = lane & 3;
    double acc[8][4][2];
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) acc[i][j][0] = acc[i][j][1] = 0.0;

    // 128 rows x 8 chunks (16 B = 2 doubles) per operand and stage: 1024 chunks / 256 threads = 4 each
    auto issue = [&](int stage, int k0) {
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const int chunk = tid + e * GF_THREADS, row = chunk >> 3, kc = (chunk & 7) * 2;
            const int k = k0 + kc;
            const int okA = (m0 + row < M && k < kend) ? 16 : 0, okB = (n0 + row < N && k < kend) ? 16 : 0;
            cp_async16(As + ((size_t)stage * GF_BM + row) * GF_LD + kc, okA ? (const void *)(A + (size_t)(m0 + row) * lda + k) : (const void *)A, okA);
            cp_async16(Bs + ((size_t)stage * GF_BN + row) * GF_LD + kc, okB ? (const void *)(Bm + (size_t)(n0 + row) * ldb + k) : (const void *)Bm, okB);
        }
        if (s && tid < GF_BK / 2) {
            const int k = k0 + tid * 2;
            const int ok = k < kend ? 16 : 0;
            cp_async16(Ss + stage * GF_BK + tid * 2, ok ? (const void *)(s + k) : (const void *)s, ok);
        }
    };
    const int nk = (kend - kbeg + GF_BK - 1) / GF_BK;
    for (int st = 0; st < GF_STAGES - 1; st++) {
        if (st < nk) issue(st, kbeg + st * GF_BK);
        cp_async_commit();
    }
    for (int kt = 0; kt < nk; kt++) {
        cp_async_wait<GF_STAGES - 2>();
        __syncthreads();
        if (kt + GF_STAGES - 1 < nk) issue((kt + GF_STAGES - 1) % GF_STAGES, kbeg + (kt + GF_STAGES - 1) * GF_BK);
        cp_async_commit();
        const int st = kt % GF_STAGES;
        const double *as = As + (size_t)st * GF_BM * GF_LD, *bs = Bs + (size_t)st * GF_BN * GF_LD, *ss = Ss + st * GF_BK;
#pragma unroll
        for (int kk = 0; kk < GF_BK; kk += 4) {
            double af[8], bf[4];
            const double sc = s ? ss[kk + t] : 1.0;
#pragma unroll
            for (int i = 0; i < 8; i++) af[i] = as[(wm + i * 8 + g) * GF_LD + kk + t] * sc;
#pragma unroll
            for (int j = 0; j < 4; j++) bf[j] = bs[(wn + j * 8 + g) * GF_LD + kk + t];
#pragma unroll
            for (int i = 0; i < 8; i++)
#pragma unroll
                for (int j = 0; j < 4; j++) dmma_m8n8k4(acc[i][j][0], acc[i][j][1], af[i], bf[j]);
        }
    }
    const bool split = gridDim.z > 1;
#pragma unroll
    for (int i = 0; i < 8; i++)
#pragma unroll
        for (int j = 0; j < 4; j++) {
            const int r = m0 + wm + i * 8 + g, c = n0 + wn + j * 8 + 2 * t;
            if (r < M) {
#pragma unroll
                for (int e = 0; e < 2; e++)
                    if (c + e < N) {
                        double *p = C + (size_t)r * ldc + c + e;
                        if (split) atomicAdd(p, alpha * acc[i][j][e]);  // C was zeroed (beta = 0) or kept (beta = 1) by the host
                        else *p = beta == 0.0 ? alpha * acc[i][j][e] : alpha * acc[i][j][e] + beta * *p;
                    }
            }
        }
}

__global__ void scale_matrix_kernel(int M, int N, double *C, int ldc, double beta) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= (size_t)M * N) return;
    double *p = C + (idx / N) * ldc + idx % N;
    *p = beta == 0.0 ? 0.0 : *p * beta;
}

int kb_dgemm_nt(kb_ctx *ctx, int M, int N, int K, const double *A, int lda, const double *s, const double *Bm, int ldb,
                double *C, int ldc, double alpha, double beta, cudaStream_t st) {
    if (M <= 0 || N <= 0) return KB_OK;
    const bool aligned = ((K | lda | ldb) & 1) == 0 && (((uintptr_t)A | (uintptr_t)Bm | (uintptr_t)(s ? s : A)) & 15) == 0;
    if (aligned && K >= 32 && (long long)M * N >= 128 * 128) {
        static bool attr_done = false;
        if (!attr_done) {
            KB_CUDA(ctx, cudaFuncSetAttribute(dgemm_nt_fast_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, GF_SMEM));
            attr_done = true;
        }
        const int tiles = ((M + GF_BM - 1) / GF_BM) * ((N + GF_BN - 1) / GF_BN);
        int splits = 1;
        const int sms = ctx->prop.multiProcessorCount;
        if (tiles < sms) {  // too few output tiles: split K so that about one wave of CTAs exists
            splits = std::min((sms + tiles - 1) / tiles, std::max(1, K / 64));
            splits = std::min(splits, 32);
        }
        int kps = (K + splits - 1) / splits;
        kps = (kps + GF_BK - 1) / GF_BK * GF_BK;
        splits = (K + kps - 1) / kps;
        KbProf pf(ctx, 11, st);
        if (splits > 1 && beta != 1.0) {
            scale_matrix_kernel<<<(unsigned)(((size_t)M * N + 255) / 256), 256, 0, st>>>(M, N, C, ldc, beta);
            ctx->launches++;
        }
        dim3 grid((unsigned)((N + GF_BN - 1) / GF_BN), (unsigned)((M + GF_BM - 1) / GF_BM), (unsigned)splits);
        dgemm_nt_fast_kernel<<<grid, GF_THREADS, GF_SMEM, st>>>(M, N, K, A, lda, s, Bm, ldb, C, ldc, alpha, beta, kps);
        KB_LAUNCH_CHECK(ctx);
        return KB_OK;
    }
    dim3 grid((unsigned)((N + GM_BN - 1) / GM_BN), (unsigned)((M + GM_BM - 1) / GM_BM));
    {
        KbProf pf(ctx, 11, st);
        dgemm_nt_kernel<<<grid, 128, 0, st>>>(M, N, K, A, lda, s, Bm, ldb, C, ldc, alpha, beta);
    }
    KB_LAUNCH_CHECK(ctx);
    return KB_OK;
}

__global__ void transpose_kernel(int rows, int cols, const double *__restrict__ in, double *__restrict__ out) {
    __shared__ double tile[32][33];
    const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int r = r0 + j, c = c0 + threadIdx.x;
        if (r < rows && c < cols) tile[j][threadIdx.x] = in[(size_t)r * cols + c];
    }
    __syncthreads();
    for (int j = threadIdx.y; j < 32; j += blockDim.y) {
        const int c = c0 + j, r = r0 + threadIdx.x;
        if (r < rows && c < cols) out[(size_t)c * rows + r] = tile[threadIdx.x][j];
    }
}

int kb_transpose(kb_ctx *ctx, int rows, int cols, const double *in, double *out, cudaStream_t st) {
    dim3 grid((unsigned)((cols + 31) / 32), (unsigned)((rows + 31) / 32));
    transpose_kernel<<<grid, dim3(32, 8), 0, st>>>(rows, cols, in, out);
    KB_LAUNCH_CHECK(ctx);
    return KB_OK;
}

// ---------------------------------------------------------------------------------------------------------------
// one-sided Jacobi.  Storage: Gt[j*n + i] = G[i][j] (column j contiguous), same for Vt.
// ---------------------------------------------------------------------------------------------------------------
#define EG_THREADS 256
#define EG_MAX_SWEEPS 40

__device__ __forceinline__ double warp_sum(double v) {
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// round-robin pairing of N (even) players, round r in [0, N-1): pair index i in [0, N/2)
__device__ __forceinline__ void rr_pair(int N, int r, int i, int &p, int &q) {
    if (i == 0) { p = N - 1; q = r; }
    else { p = (r + i) % (N - 1); q = (r - i + (N - 1)) % (N - 1); }
    if (p > q) { int t = p; p = q; q = t; }
}

template <bool GRID>
__global__ void __launch_bounds__(EG_THREADS)
jacobi_onesided_kernel(int n, double *Gt, double *Vt, EighShared *sh, double tol) {
    const int warps_per_block = EG_THREADS / 32;
    const int gw = blockIdx.x * warps_per_block + (threadIdx.x >> 5);
    const int total_warps = gridDim.x * warps_per_block;
    const int lane = threadIdx.x & 31;
    const int N = n + (n & 1);  // pad to even; player n (if any) is a bye
    cg::grid_group grid = cg::this_grid();
    for (int sweep = 0; sweep < EG_MAX_SWEEPS; sweep++) {
        for (int r = 0; r < N - 1; r++) {
            for (int i = gw; i < N / 2; i += total_warps) {
                int p, q;
                rr_pair(N, r, i, p, q);
                if (q >= n) continue;
                double *gp = Gt + (size_t)p * n, *gq = Gt + (size_t)q * n;
                double a = 0.0, b = 0.0, c = 0.0;
                for (int k = lane; k < n; k += 32) {
                    const double x = gp[k], y = gq[k];
                    a += x * x; b += y * y; c += x * y;
                }
                a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
                if (fabs(c) <= tol * sqrt(a * b) || c == 0.0) continue;
                const double zeta = (b - a) / (2.0 * c);
                const double tt = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                const double cs = 1.0 / sqrt(1.0 + tt * tt), sn = cs * tt;
                double *vp = Vt + (size_t)p * n, *vq = Vt + (size_t)q * n;
                for (int k = lane; k < n; k += 32) {
                    const double x = gp[k], y = gq[k];
                    gp[k] = cs * x - sn * y; gq[k] = sn * x + cs * y;
                    const double u = vp[k], v = vq[k];
                    vp[k] = cs * u - sn * v; vq[k] = sn * u + cs * v;
                }
                if (lane == 0) atomicAdd(&sh->rotated, 1);
            }
            if (GRID) grid.sync(); else __syncthreads();
        }
        const int rot = *(volatile int *)&sh->rotated;
        if (GRID) grid.sync(); else __syncthreads();
        if (blockIdx.x == 0 && threadIdx.x == 0) { sh->rotated = 0; sh->sweeps = sweep + 1; }
        if (GRID) grid.sync(); else __syncthreads();
        if (rot == 0) break;
    }
}

// ---- the multi-CTA solver: one CTA per column pair, data-flow synchronisation instead of a grid barrier ---------
// In round-robin order the pair (p, q) of step t only needs the two pairs of step t-1 that held p and q.  Every
// column carries a counter of completed steps (release store after the CTA's writes, acquire poll before the next
// CTA reads), so a step costs one L2 round trip instead of a grid-wide barrier; the grid synchronises once per sweep
// to agree on convergence.  Cooperative launch guarantees that all CTAs are resident, which the polling needs.
#define EF_THREADS 128
#define EF_NPT 8  // rows per thread held in registers when n <= EF_THREADS * EF_NPT

__device__ __forceinline__ int ld_acquire_gpu(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_release_gpu(int *p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}

__global__ void __launch_bounds__(EF_THREADS)
jacobi_flag_kernel(int n, double *Gt, double *Vt, int *colstep, EighShared *sh, int *sweep_rot, double tol) {
    __shared__ double s_red[3][EF_THREADS / 32];
    const int N = n + (n & 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const bool cached = n <= EF_THREADS * EF_NPT;
    cg::grid_group grid = cg::this_grid();
    int t = 0;  // global step counter across sweeps
    for (int sweep = 0; sweep < EG_MAX_SWEEPS; sweep++) {
        int my_rot = 0;
        for (int r = 0; r < N - 1; r++, t++) {
            for (int i = blockIdx.x; i < N / 2; i += gridDim.x) {
                int p, q;
                rr_pair(N, r, i, p, q);
                if (tid == 0) {
                    while (ld_acquire_gpu(&colstep[p]) < t) {}
                    while (ld_acquire_gpu(&colstep[q]) < t) {}
                }
                __syncthreads();
                if (q < n) {
                    double *gp = Gt + (size_t)p * n, *gq = Gt + (size_t)q * n;
                    double *vp = Vt + (size_t)p * n, *vq = Vt + (size_t)q * n;
                    double x[EF_NPT], y[EF_NPT], u[EF_NPT], v[EF_NPT];
                    double a = 0.0, b = 0.0, c = 0.0;
                    if (cached) {
                        // all four columns are requested up front: the V loads overlap the reduction and the rotation maths
#pragma unroll
                        for (int e = 0; e < EF_NPT; e++) {
                            const int k = tid + e * EF_THREADS;
                            x[e] = k < n ? __ldcg(gp + k) : 0.0;
                            y[e] = k < n ? __ldcg(gq + k) : 0.0;
                        }
#pragma unroll
                        for (int e = 0; e < EF_NPT; e++) {
                            const int k = tid + e * EF_THREADS;
                            u[e] = k < n ? __ldcg(vp + k) : 0.0;
                            v[e] = k < n ? __ldcg(vq + k) : 0.0;
                        }
#pragma unroll
                        for (int e = 0; e < EF_NPT; e++) { a += x[e] * x[e]; b += y[e] * y[e]; c += x[e] * y[e]; }
                    } else {
                        for (int k = tid; k < n; k += EF_THREADS) {
                            const double xv = __ldcg(gp + k), yv = __ldcg(gq + k);
                            a += xv * xv; b += yv * yv; c += xv * yv;
                        }
                    }
                    a = warp_sum(a); b = warp_sum(b); c = warp_sum(c);
                    if (lane == 0) { s_red[0][warp] = a; s_red[1][warp] = b; s_red[2][warp] = c; }
                    __syncthreads();
                    a = b = c = 0.0;
#pragma unroll
                    for (int w2 = 0; w2 < EF_THREADS / 32; w2++) { a += s_red[0][w2]; b += s_red[1][w2]; c += s_red[2][w2]; }
                    if (!(fabs(c) <= tol * sqrt(a * b) || c == 0.0)) {  // block-uniform
                        const double zeta = (b - a) / (2.0 * c);
                        const double tt = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                        const double cs = rsqrt(1.0 + tt * tt), sn = cs * tt;
                        if (cached) {
#pragma unroll
                            for (int e = 0; e < EF_NPT; e++) {
                                const int k = tid + e * EF_THREADS;
                                if (k < n) {
                                    __stcg(gp + k, cs * x[e] - sn * y[e]);
                                    __stcg(gq + k, sn * x[e] + cs * y[e]);
                                    __stcg(vp + k, cs * u[e] - sn * v[e]);
                                    __stcg(vq + k, sn * u[e] + cs * v[e]);
                                }
                            }
                        } else {
                            for (int k = tid; k < n; k += EF_THREADS) {
                                const double xv = __ldcg(gp + k), yv = __ldcg(gq + k);
                                __stcg(gp + k, cs * xv - sn * yv);
                                __stcg(gq + k, sn * xv + cs * yv);
                                const double uu = __ldcg(vp + k), vv = __ldcg(vq + k);
                                __stcg(vp + k, cs * uu - sn * vv);
                                __stcg(vq + k, sn * uu + cs * vv);
                            }
                        }
                        my_rot++;
                    }
                }
                // the release store below, issued after the CTA barrier, publishes every thread's writes (cumulativity)
                __syncthreads();
                if (tid == 0) { st_release_gpu(&colstep[p], t + 1); st_release_gpu(&colstep[q], t + 1); }
            }
        }
        if (tid == 0 && my_rot) atomicAdd(&sweep_rot[sweep], my_rot);
        grid.sync();
        const int rot = __ldcg(&sweep_rot[sweep]);
        if (blockIdx.x == 0 && tid == 0) sh->sweeps = sweep + 1;
        if (rot == 0) break;
    }
}

// ---- block version: a CTA owns 2*JB columns in shared memory for a whole local tournament ------------------------
// Columns are grouped into blocks of JB; the round-robin runs over BLOCK pairs (nblk - 1 outer steps per sweep
// instead of n - 1), and inside an outer step the CTA orthogonalises all pairs among its 2*JB columns (2*JB - 1 local
// steps, JB pairs each, one warp per pair) out of shared memory.  Global traffic and the inter-CTA hand-off (same
// release/acquire counters, per block) are paid once per outer step, so the serial chain is ~JB times shorter.
#define JB 4
#define JB_THREADS (32 * JB)

__global__ void __launch_bounds__(JB_THREADS)
jacobi_block_kernel(int n, double *Gt, double *Vt, int *blkstep, EighShared *sh, int *sweep_rot, double tol, int with_v) {
    extern __shared__ __align__(16) double s_cols[];  // [2*JB][n] of G, then [2*JB][n] of V
    double *sG = s_cols, *sV = s_cols + (size_t)2 * JB * n;
    const int nblk0 = (n + JB - 1) / JB;
    const int NB = nblk0 + (nblk0 & 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    cg::grid_group grid = cg::this_grid();
    int t = 0;
    for (int sweep = 0; sweep < EG_MAX_SWEEPS; sweep++) {
        int my_rot = 0;
        for (int r = 0; r < NB - 1; r++, t++) {
            for (int i = blockIdx.x; i < NB / 2; i += gridDim.x) {
                int P, Q;
                rr_pair(NB, r, i, P, Q);
                if (tid == 0) {
                    while (ld_acquire_gpu(&blkstep[P]) < t) {}
                    while (ld_acquire_gpu(&blkstep[Q]) < t) {}
                }
                __syncthreads();
                // global column of local slot s
                auto col_of = [&](int s) { return (s < JB ? P * JB + s : Q * JB + (s - JB)); };
                if ((n & 1) == 0) {
                    // 16-byte loads, all of a batch in flight before the first shared-memory store (memory-level parallelism)
                    const int n2 = n >> 1;
                    for (int s0 = 0; s0 < 2 * JB; s0 += 2) {
                        for (int k0 = 0; k0 < n2; k0 += JB_THREADS * 4) {
                            double2 g[2][4], v[2][4];
#pragma unroll
                            for (int ss = 0; ss < 2; ss++) {
                                const int c = col_of(s0 + ss);
#pragma unroll
                                for (int e = 0; e < 4; e++) {
                                    const int k = k0 + tid + e * JB_THREADS;
                                    const bool ok = c < n && k < n2;
                                    g[ss][e] = ok ? __ldcg(reinterpret_cast<const double2 *>(Gt + (size_t)c * n) + k) : make_double2(0.0, 0.0);
                                    v[ss][e] = ok && with_v ? __ldcg(reinterpret_cast<const double2 *>(Vt + (size_t)c * n) + k) : make_double2(0.0, 0.0);
                                }
                            }
#pragma unroll
                            for (int ss = 0; ss < 2; ss++)
#pragma unroll
                                for (int e = 0; e < 4; e++) {
                                    const int k = k0 + tid + e * JB_THREADS;
                                    if (k < n2) {
                                        reinterpret_cast<double2 *>(sG + (size_t)(s0 + ss) * n)[k] = g[ss][e];
                                        if (with_v) reinterpret_cast<double2 *>(sV + (size_t)(s0 + ss) * n)[k] = v[ss][e];
                                    }
                                }
                        }
                    }
                } else {
                    for (int s = 0; s < 2 * JB; s++) {
                        const int c = col_of(s);
                        for (int k = tid; k < n; k += JB_THREADS) {
                            sG[(size_t)s * n + k] = c < n ? __ldcg(Gt + (size_t)c * n + k) : 0.0;
                            if (with_v) sV[(size_t)s * n + k] = c < n ? __ldcg(Vt + (size_t)c * n + k) : 0.0;
                        }
                    }
                }
                __syncthreads();
                bool dirty = false;
                for (int lr = 0; lr < 2 * JB - 1; lr++) {
                    int a, b;
                    rr_pair(2 * JB, lr, warp, a, b);
                    if (col_of(a) < n && col_of(b) < n) {
                        double *ga = sG + (size_t)a * n, *gb = sG + (size_t)b * n;
                        double aa = 0.0, bb = 0.0, cc = 0.0, aa2 = 0.0, bb2 = 0.0, cc2 = 0.0;
                        int k = lane;
                        for (; k + 32 < n; k += 64) {  // two independent accumulation chains
                            const double x = ga[k], y = gb[k], x2 = ga[k + 32], y2 = gb[k + 32];
                            aa += x * x; bb += y * y; cc += x * y;
                            aa2 += x2 * x2; bb2 += y2 * y2; cc2 += x2 * y2;
                        }
                        for (; k < n; k += 32) { const double x = ga[k], y = gb[k]; aa += x * x; bb += y * y; cc += x * y; }
                        aa = warp_sum(aa + aa2); bb = warp_sum(bb + bb2); cc = warp_sum(cc + cc2);
                        if (!(fabs(cc) <= tol * sqrt(aa * bb) || cc == 0.0)) {
                            const double zeta = (bb - aa) / (2.0 * cc);
                            const double tt = (zeta >= 0.0 ? 1.0 : -1.0) / (fabs(zeta) + sqrt(1.0 + zeta * zeta));
                            const double cs = rsqrt(1.0 + tt * tt), sn = cs * tt;
                            double *va = sV + (size_t)a * n, *vb = sV + (size_t)b * n;
                            if (with_v) {
                                for (int k = lane; k < n; k += 32) {
                                    const double x = ga[k], y = gb[k];
                                    ga[k] = cs * x - sn * y; gb[k] = sn * x + cs * y;
                                    const double u = va[k], v = vb[k];
                                    va[k] = cs * u - sn * v; vb[k] = sn * u + cs * v;
                                }
                            } else {
                                for (int k = lane; k < n; k += 32) {
                                    const double x = ga[k], y = gb[k];
                                    ga[k] = cs * x - sn * y; gb[k] = sn * x + cs * y;
                                }
                            }
                            dirty = true;
                            if (lane == 0) my_rot++;
                        }
                    }
                    __syncthreads();
                }
                // write back only if something in this CTA rotated (a block-uniform vote keeps the barrier legal)
                const int any = __syncthreads_or(dirty ? 1 : 0);
                if (any) {
                    for (int s = 0; s < 2 * JB; s++) {
                        const int c = col_of(s);
                        if (c >= n) continue;
                        if ((n & 1) == 0) {
                            for (int k = tid; k < (n >> 1); k += JB_THREADS) {
                                __stcg(reinterpret_cast<double2 *>(Gt + (size_t)c * n) + k, reinterpret_cast<const double2 *>(sG + (size_t)s * n)[k]);
                                if (with_v) __stcg(reinterpret_cast<double2 *>(Vt + (size_t)c * n) + k, reinterpret_cast<const double2 *>(sV + (size_t)s * n)[k]);
                            }
                        } else {
                            for (int k = tid; k < n; k += JB_THREADS) {
                                __stcg(Gt + (size_t)c * n + k, sG[(size_t)s * n + k]);
                                if (with_v) __stcg(Vt + (size_t)c * n + k, sV[(size_t)s * n + k]);
                            }
                        }
                    }
                }
                __syncthreads();
                if (tid == 0) { st_release_gpu(&blkstep[P], t + 1); st_release_gpu(&blkstep[Q], t + 1); }
            }
        }
        if (lane == 0 && my_rot) atomicAdd(&sweep_rot[sweep], my_rot);
        grid.sync();
        const int rot = __ldcg(&sweep_rot[sweep]);
        if (blockIdx.x == 0 && tid == 0) sh->sweeps = sweep + 1;
        if (rot == 0) break;
    }
}

// ---- Cholesky preconditioning (Demmel-Veselic / Drmac): for a positive definite A = L L^T, one-sided Jacobi on the
// columns of L converges in far fewer sweeps than on A itself, needs no accumulated V (the eigenvectors are the
// normalised columns of the converged G, the eigenvalues their squared norms) and resolves small eigenvalues to
// high relative accuracy.  Right-looking blocked factorisation: a single-CTA panel kernel (panel resident in shared
// memory) followed by a DMMA rank-pw update of the trailing matrix.  Storage is the column-contiguous Gt.
#define CH_THREADS 1024
__global__ void __launch_bounds__(CH_THREADS)
chol_panel_kernel(int n, int j0, int pw, double *Gt, double *Pt, int *fail) {
    extern __shared__ __align__(16) double sP[];  // [pw][m], m = n - j0
    const int m = n - j0, tid = threadIdx.x;
    if (*fail) return;
    for (int c = 0; c < pw; c++)
        for (int i = tid; i < m; i += CH_THREADS) sP[(size_t)c * m + i] = Gt[(size_t)(j0 + c) * n + j0 + i];
    __syncthreads();
    for (int c = 0; c < pw; c++) {
        const double pivot = sP[(size_t)c * m + c];
        if (!(pivot > 0.0)) {  // not positive definite (or NaN): every thread sees the same value
            if (tid == 0) *fail = 1;
            return;
        }
        __syncthreads();
        const double d = sqrt(pivot), inv = 1.0 / d;
        for (int i = c + tid; i < m; i += CH_THREADS) sP[(size_t)c * m + i] = i == c ? d : sP[(size_t)c * m + i] * inv;
        __syncthreads();
        const int rest = pw - c - 1;
        for (int e = tid; e < rest * m; e += CH_THREADS) {
            const int c2 = c + 1 + e / m, i = e % m;
            if (i >= c2) sP[(size_t)c2 * m + i] -= sP[(size_t)c * m + i] * sP[(size_t)c * m + c2];
        }
        __syncthreads();
    }
    for (int c = 0; c < pw; c++) {
        for (int i = tid; i < m; i += CH_THREADS) {
            const double v = i >= c ? sP[(size_t)c * m + i] : 0.0;
            Gt[(size_t)(j0 + c) * n + j0 + i] = v;
            if (i >= pw) Pt[(size_t)(i - pw) * pw + c] = v;
        }
        for (int i = tid; i < j0; i += CH_THREADS) Gt[(size_t)(j0 + c) * n + i] = 0.0;  // strictly upper part of L
    }
}
// G = V sqrt(Lambda) after Jacobi on L: lambda_j = |g_j|^2 - shift, v_j = g_j / |g_j|
__global__ void eigh_finish_chol_kernel(int n, const double *__restrict__ Gt, double *Vt, double shift, double *lam) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp >= n) return;
    double acc = 0.0;
    for (int k = lane; k < n; k += 32) { const double g = Gt[(size_t)warp * n + k]; acc += g * g; }
    acc = warp_sum(acc);
    const double inv = acc > 0.0 ? rsqrt(acc) : 0.0;
    for (int k = lane; k < n; k += 32) Vt[(size_t)warp * n + k] = Gt[(size_t)warp * n + k] * inv;
    if (lane == 0) lam[warp] = acc - shift;
}

// lambda_j = v_j . g_j (keeps the sign for indefinite input), then rank by value, canonical signs, write outputs:
//   w[rank] = lambda, V[i*n + rank] = +-v_j[i]  (row-major, eigenvector `rank` in column `rank`)
__global__ void eigh_finish_kernel(int n, const double *__restrict__ Gt, const double *__restrict__ Vt, double shift,
                                   double *lam, double *w, double *V) {
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
    if (warp < n) {
        double acc = 0.0;
        for (int k = lane; k < n; k += 32) acc += Vt[(size_t)warp * n + k] * Gt[(size_t)warp * n + k];
        acc = warp_sum(acc);
        if (lane == 0) lam[warp] = acc - shift;
    }
}
__global__ void eigh_order_kernel(int n, const double *__restrict__ Vt, const double *__restrict__ lam, double *w, double *V) {
    const int j = blockIdx.x;  // source column
    __shared__ int s_rank;
    __shared__ double s_sign;
    __shared__ double s_big[4];
    __shared__ int s_at[4];
    // rank = how many eigenvalues sort before this one; sign = that of the first component of largest magnitude
    const double lj = lam[j];
    int cnt = 0;
    double big = 0.0;
    int at = n;
    for (int e = threadIdx.x; e < n; e += blockDim.x) {
        const double le = lam[e];
        cnt += (le < lj || (le == lj && e < j)) ? 1 : 0;
        const double a = fabs(Vt[(size_t)j * n + e]);
        if (a > big) { big = a; at = e; }  // ascending e per thread: the first maximum stays
    }
    if (threadIdx.x == 0) s_rank = 0;
    __syncthreads();
    for (int o = 16; o > 0; o >>= 1) {
        cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        const double ob = __shfl_xor_sync(0xffffffffu, big, o);
        const int oa = __shfl_xor_sync(0xffffffffu, at, o);
        if (ob > big || (ob == big && oa < at)) { big = ob; at = oa; }
    }
    if ((threadIdx.x & 31) == 0) { atomicAdd(&s_rank, cnt); s_big[threadIdx.x >> 5] = big; s_at[threadIdx.x >> 5] = at; }
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int q = 1; q < (int)(blockDim.x >> 5); q++)
            if (s_big[q] > big || (s_big[q] == big && s_at[q] < at)) { big = s_big[q]; at = s_at[q]; }
        s_sign = (at < n && Vt[(size_t)j * n + at] < 0) ? -1.0 : 1.0;
        w[s_rank] = lj;
    }
    __syncthreads();
    const int r = s_rank;
    const double sg = s_sign;
    for (int k = threadIdx.x; k < n; k += blockDim.x) V[(size_t)k * n + r] = sg * Vt[(size_t)j * n + k];
}

__global__ void eigh_init_kernel(int n, const double *__restrict__ A, double shift, double *Gt, double *Vt, EighShared *sh,
                                 int *colstep, int *sweep_rot, int warm) {
    const size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (idx == 0) { sh->rotated = 0; sh->sweeps = 0; }
    if (idx < (size_t)n + 2) colstep[idx] = 0;
    if (idx < EG_MAX_SWEEPS) sweep_rot[idx] = 0;
    if (idx >= (size_t)n * n || warm) return;
    const int j = (int)(idx / n), i = (int)(idx % n);
    // symmetrise as the reference does before factorising: (C + C^T) / 2
    Gt[idx] = 0.5 * (A[(size_t)i * n + j] + A[(size_t)j * n + i]) + (i == j ? shift : 0.0);
    Vt[idx] = i == j ? 1.0 : 0.0;
}

// workspace: Gt, Vt (n*n each), lam (n), EighShared, per-column step counters, per-sweep rotation counts
size_t kb_eigh_workspace_bytes(int n) {
    return sizeof(double) * ((size_t)2 * n * n + n + 2 + (size_t)32 * n) + sizeof(int) * ((size_t)n + 8 + EG_MAX_SWEEPS) + 256;
}

// d_V0t (optional): transposed eigenvectors of a nearby matrix (row j = vector j).  The iteration then starts from
// G = A V0, V = V0, whose columns are already almost orthogonal -- CMA-ES changes C by a few percent per generation,
// so the warm start converges in 2-3 sweeps instead of ~10.
int kb_eigh_device(kb_ctx *ctx, int n, const double *d_A, double shift, double *d_w, double *d_V, void *workspace,
                   cudaStream_t st, const double *d_V0t, int *sweeps_out) {
    double *Gt = (double *)workspace, *Vt = Gt + (size_t)n * n, *lam = Vt + (size_t)n * n;
    EighShared *sh = (EighShared *)(lam + n + (n & 1));
    int *colstep = (int *)(sh + 1), *sweep_rot = colstep + n + 2 + (n & 1), *chol_fail = sweep_rot + EG_MAX_SWEEPS;
    double *Pt = (double *)(chol_fail + 4);  // [n][32] panel copy for the trailing update (8-byte aligned: counts above are even)
    KbProf pf(ctx, 9, st);
    const int warm = d_V0t != nullptr && shift == 0.0;
    eigh_init_kernel<<<(unsigned)(((size_t)n * n + 255) / 256), 256, 0, st>>>(n, d_A, shift, Gt, Vt, sh, colstep, sweep_rot, warm);
    KB_LAUNCH_CHECK(ctx);
    if (warm) {
        KB_CUDA(ctx, cudaMemcpyAsync(Vt, d_V0t, sizeof(double) * (size_t)n * n, cudaMemcpyDeviceToDevice, st));
        int rc = kb_dgemm_nt(ctx, n, n, n, d_V0t, n, nullptr, d_A, n, Gt, n, 1.0, 0.0, st);  // Gt[j,:] = A v_j
        if (rc) return rc;
    }
    const double tol = sqrt((double)n) * 2.220446049250313e-16;  // LAPACK dgesvj-style stopping criterion
    const int pairs = (n + 1) / 2;
    const int warps_per_block = EG_THREADS / 32;
    bool chol_path = false;
    if (pairs <= warps_per_block * 2) {
        jacobi_onesided_kernel<false><<<1, EG_THREADS, 0, st>>>(n, Gt, Vt, sh, tol);
        KB_LAUNCH_CHECK(ctx);
    } else {
        double tolv = tol;
        const size_t smem = sizeof(double) * (size_t)4 * JB * n;  // 2*JB columns of G and of V
        if (smem <= (size_t)ctx->prop.sharedMemPerBlockOptin) {
            // block version: the 2*JB columns of a block pair live in shared memory for a whole local tournament
            const int nblk0 = (n + JB - 1) / JB, NB = nblk0 + (nblk0 & 1);
            int with_v = 1;
            // Cholesky preconditioning when the panel fits in shared memory and no warm start was asked for
            int pw = (int)std::min<size_t>(32, (size_t)ctx->prop.sharedMemPerBlockOptin / (sizeof(double) * (size_t)n));
            if (!warm && pw >= 4) {
                KB_CUDA(ctx, cudaMemsetAsync(chol_fail, 0, sizeof(int), st));
                KB_CUDA(ctx, cudaFuncSetAttribute(chol_panel_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                  (int)(sizeof(double) * (size_t)pw * n)));
                bool coop = false;  // single cooperative kernel (kb_chol.cu) when a 32-column panel fits in shared memory
                int rcc = kb_chol_coop(ctx, n, Gt, chol_fail, st, &coop);
                if (rcc) return rcc;
                for (int j0 = coop ? n : 0; j0 < n; j0 += pw) {
                    const int w = std::min(pw, n - j0), m = n - j0;
                    chol_panel_kernel<<<1, CH_THREADS, sizeof(double) * (size_t)w * m, st>>>(n, j0, w, Gt, Pt, chol_fail);
                    KB_LAUNCH_CHECK(ctx);
                    if (m > w) {
                        double *T22 = Gt + (size_t)(j0 + w) * n + (j0 + w);
                        int rc = kb_dgemm_nt(ctx, m - w, m - w, w, Pt, w, nullptr, Pt, w, T22, n, -1.0, 1.0, st);
                        if (rc) return rc;
                    }
                }
                int failed = 0;
                KB_CUDA(ctx, cudaMemcpyAsync(&failed, chol_fail, sizeof(int), cudaMemcpyDeviceToHost, st));
                KB_CUDA(ctx, cudaStreamSynchronize(st));
                if (failed) {  // not numerically positive definite: start over on A itself with accumulated V
                    eigh_init_kernel<<<(unsigned)(((size_t)n * n + 255) / 256), 256, 0, st>>>(n, d_A, shift, Gt, Vt, sh, colstep, sweep_rot, 0);
                    KB_LAUNCH_CHECK(ctx);
                } else {
                    with_v = 0;
                    chol_path = true;
                }
            }
            bool gram = false;  // Gram-form block step (kb_eigh_gram.cu) when its columns fit; else the column-wise one
            // the Cholesky panel buffer (32 n doubles) is free by now: it carries the per-block Gram diagonal blocks
            int rcg = kb_jacobi_gram(ctx, n, Gt, Vt, colstep, sh, sweep_rot, Pt, tolv, with_v, EG_MAX_SWEEPS, st, &gram);
            if (rcg) return rcg;
            if (!gram) {
                KB_CUDA(ctx, cudaFuncSetAttribute(jacobi_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
                int blocks = NB / 2, per_sm = 0;
                KB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, jacobi_block_kernel, JB_THREADS, smem));
                const int cap = per_sm * ctx->prop.multiProcessorCount;
                if (blocks > cap) blocks = cap;
                void *args[] = {&n, &Gt, &Vt, &colstep, &sh, &sweep_rot, &tolv, &with_v};
                KB_CUDA(ctx, cudaLaunchCooperativeKernel((void *)jacobi_block_kernel, dim3((unsigned)blocks), dim3(JB_THREADS), args, smem, st));
            } else ctx->launches--;  // counted by kb_jacobi_gram
        } else {
            int blocks = pairs, per_sm = 0;
            KB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, jacobi_flag_kernel, EF_THREADS, 0));
            const int cap = per_sm * ctx->prop.multiProcessorCount;
            if (blocks > cap) blocks = cap;
            void *args[] = {&n, &Gt, &Vt, &colstep, &sh, &sweep_rot, &tolv};
            KB_CUDA(ctx, cudaLaunchCooperativeKernel((void *)jacobi_flag_kernel, dim3((unsigned)blocks), dim3(EF_THREADS), args, 0, st));
        }
        ctx->launches++;
    }
    if (chol_path) eigh_finish_chol_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, st>>>(n, Gt, Vt, shift, lam);
    else eigh_finish_kernel<<<(unsigned)((n * 32 + 255) / 256), 256, 0, st>>>(n, Gt, Vt, shift, lam, d_w, d_V);
    KB_LAUNCH_CHECK(ctx);
    eigh_order_kernel<<<(unsigned)n, 128, 0, st>>>(n, Vt, lam, d_w, d_V);
    KB_LAUNCH_CHECK(ctx);
    KB_CUDA(ctx, cudaMemcpyAsync(&ctx->last_eigh_sweeps, &sh->sweeps, sizeof(int), cudaMemcpyDeviceToHost, st));
    if (sweeps_out) KB_CUDA(ctx, cudaMemcpyAsync(sweeps_out, &sh->sweeps, sizeof(int), cudaMemcpyDeviceToHost, st));
    return KB_OK;
}

// canonical order and signs for eigenvectors given as rows of Vt with eigenvalues lam (both device): w ascending,
// V[i*n+j] = component i of eigenvector j, largest-magnitude component positive
int kb_eigh_canonical(kb_ctx *ctx, int n, const double *Vt, const double *lam, double *w, double *V, cudaStream_t st) {
    eigh_order_kernel<<<(unsigned)n, 128, 0, st>>>(n, Vt, lam, w, V);
    KB_LAUNCH_CHECK(ctx);
    return KB_OK;
}

extern "C" int kb_eigh(kb_ctx *ctx, int n, const double *A, double *w, double *V) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (n < 1 || n > 4096) return kb_fail(ctx, KB_ERR_UNSUPPORTED, "kb_eigh: n=%d outside [1,4096]", n);
    if (!A || !w || !V) return kb_fail(ctx, KB_ERR_INVALID, "kb_eigh: NULL buffer");
    // one-sided Jacobi resolves eigenvectors of A^2: shift an indefinite matrix positive (Gershgorin bound) first
    double shift = 0.0, low = INFINITY;
    for (int i = 0; i < n; i++) {
        double r = 0.0;
        for (int j = 0; j < n; j++) if (j != i) r += fabs(0.5 * (A[(size_t)i * n + j] + A[(size_t)j * n + i]));
        low = fmin(low, A[(size_t)i * n + i] - r);
    }
    bool spd_hint = true;  // diagonally "positive enough": keep the relative accuracy of small eigenvalues
    for (int i = 0; i < n; i++) if (!(A[(size_t)i * n + i] > 0.0)) spd_hint = false;
    if (!spd_hint && low < 0.0) shift = -low * 1.0625 + 1e-300;
    cudaStream_t s = ctx->stream;
    double *d = nullptr;
    const size_t nn = (size_t)n * n;
    KB_CUDA(ctx, cudaMalloc(&d, sizeof(double) * (2 * nn + n) + kb_eigh_workspace_bytes(n)));
    double *d_A = d, *d_V = d + nn, *d_w = d_V + nn;
    void *ws = (void *)(d_w + n + (n & 1));
    cudaError_t e = cudaMemcpyAsync(d_A, A, sizeof(double) * nn, cudaMemcpyHostToDevice, s);
    int rc = e == cudaSuccess ? kb_eigh_device(ctx, n, d_A, shift, d_w, d_V, ws, s, nullptr, nullptr) : kb_fail(ctx, KB_ERR_CUDA, "kb_eigh: H2D");
    if (!rc) {
        e = cudaMemcpyAsync(w, d_w, sizeof(double) * n, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaMemcpyAsync(V, d_V, sizeof(double) * nn, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) rc = kb_fail(ctx, KB_ERR_CUDA, "kb_eigh: %s", cudaGetErrorString(e));
    }
    cudaFree(d);
    return rc;
}

extern "C" int kb_eigh_stats(const kb_ctx *ctx, int *sweeps) {
    if (!ctx || !sweeps) return KB_ERR_INVALID;
    *sweeps = ctx->last_eigh_sweeps;
    return KB_OK;
}
