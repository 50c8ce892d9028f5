// kb_eigh_gram.cu -- the block step of the one-sided Jacobi eigensolver, Gram-matrix form.
//
// Stands in (with kb_linalg.cu) for gonum's mat.EigenSym behind goptuna's cmaes.Optimizer (go.mod:141, reached from
// sample.go:39).  A CTA owns the M columns of a block pair in shared memory for one outer step of the round-robin
// tournament, as before; what changed is the work inside the step.  Instead of 3 dot products of length n and an
// O(n) rotation for every one of the M(M-1)/2 column pairs, each behind its own barrier, the step is
//   1. H = G^T G                the M x M Gram matrix of the resident columns, FP64 DMMA (mma.sync.m8n8k4.f64)
//   2. one cyclic sweep of two-sided Jacobi on H (M-1 parallel steps, one thread per entry), rotations accumulated
//      in W -- the same rotations, in the same order, that the column-wise tournament would apply
//   3. G <- G W  (and V <- V W)   one n x M x M DMMA product
// so the O(n) work is two small tensor-core products and the serial chain is M-1 steps on a matrix that lives in
// shared memory.  Block pairs are handed between CTAs by the per-block release/acquire counters of kb_linalg.cu.
#include <cooperative_groups.h>
#include <math.h>

#include <algorithm>

#include "kb_internal.cuh"
#include "kb_linalg.cuh"

namespace cg = cooperative_groups;

#define GJ_THREADS 256
#define GJ_WARPS (GJ_THREADS / 32)

__device__ __forceinline__ void gj_dmma(double &c0, double &c1, double a, double b) {
    asm volatile("mma.sync.aligned.m8n8k4.row.col.f64.f64.f64.f64 {%0,%1}, {%2}, {%3}, {%0,%1};"
                 : "+d"(c0), "+d"(c1) : "d"(a), "d"(b));
}
__device__ __forceinline__ int gj_ld_acquire(const int *p) {
    int v;
    asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void gj_st_release(int *p, int v) {
    asm volatile("st.release.gpu.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ void gj_cp_async16(void *smem, const void *gmem) {
    const unsigned int sa = (unsigned int)__cvta_generic_to_shared(smem);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(sa), "l"(gmem));
}
__device__ __forceinline__ void gj_cp_async_wait_all() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// round-robin pairing of N (even) players, round r in [0, N-1): pair index i in [0, N/2)  (same as kb_linalg.cu)
__device__ __forceinline__ void gj_rr_pair(int N, int r, int i, int &p, int &q) {
    if (i == 0) { p = N - 1; q = r; }
    else { p = (r + i) % (N - 1); q = (r - i + (N - 1)) % (N - 1); }
    if (p > q) { int t = p; p = q; q = t; }
}
// the opponent of player x in round r of the same schedule
__device__ __forceinline__ int gj_partner(int N, int r, int x) {
    if (x == N - 1) return r;
    const int y = (2 * r - x + 2 * (N - 1)) % (N - 1);
    return y == x ? N - 1 : y;
}

// one SFU instruction each, ~20 correct bits, FP64 in and out
__device__ __forceinline__ double gj_rsqrt_approx(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}
__device__ __forceinline__ double gj_rcp_approx(double x) {
    double y;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}

// Rotation of the pair {x, px} from the Gram entries a = H[p][p], b = H[q][q], c = H[p][q] (p < q): columns become
// g_p' = cs g_p - sn g_q, g_q' = sn g_p + cs g_q with the small-angle root, written without divisions:
//   cos 2t = |d| / r, r = sqrt(d^2 + 4 c^2), d = b - a;  cs = sqrt((1 + cos 2t) / 2);  sn = sign(d) c / (r cs).
// Returns the coefficients seen from index x: x' = cx * x + sx * partner.
template <int M>
__device__ __forceinline__ bool gj_rotation(const double *H, int x, int px, double tol2, double &cx, double &sx) {
    const int p = min(x, px), q = max(x, px);
    const double a = H[p * (M + 1) + p], b = H[q * (M + 1) + q], c = H[p * (M + 1) + q];
    cx = 1.0; sx = 0.0;
    if (!(c * c > tol2 * a * b)) return false;
    const double d = b - a;
    double cs, sn;
    const double ad = fabs(d), c2 = 2.0 * c, x2 = fma(d, d, c2 * c2);
    if (x2 < 1e290 && x2 > 1e-290) {
        // The step sits on the solver's serial chain, and an FP64 sqrt or division is ~30 dependent instructions.
        // So: the SFU's 20-bit FP64 approximations (MUFU.RSQ64H / RCP64H) seed everything.  tan(t) = 2c sign(d) /
        // (|d| + r) to 20 bits, one Newton step on c t^2 + d t - c = 0 (40 bits: the angle need not be exact, the
        // rotation must be orthogonal), then 1/sqrt(1 + t^2) from a 20-bit seed and two Newton steps (full FP64).
        const double r = x2 * gj_rsqrt_approx(x2);
        double tt = (d >= 0.0 ? c2 : -c2) * gj_rcp_approx(ad + r);
        const double fp = gj_rcp_approx(fma(c2, tt, d));               // 1 / f'(t), f(t) = c t^2 + d t - c
        tt = fma(-fma(fma(c, tt, d), tt, -c), fp, tt);
        const double s2 = fma(tt, tt, 1.0);
        double y = gj_rsqrt_approx(s2);
        y = fma(0.5 * y, fma(-s2 * y, y, 1.0), y);
        y = fma(0.5 * y, fma(-s2 * y, y, 1.0), y);
        cs = y; sn = tt * y;
    } else {  // close to the ends of the FP64 range: the plain form
        const double inv_r = rsqrt(d * d + 4.0 * c * c);
        const double xh = 0.5 + 0.5 * fabs(d) * inv_r;
        const double rs = rsqrt(xh);
        sn = (d >= 0.0 ? c : -c) * inv_r * rs;
        cs = xh * rs;
    }
    cx = cs;
    sx = x == p ? -sn : sn;
    return true;
}

#ifdef KB_EIGH_DEBUG
__device__ long long gj_dbg[8];
#define GJ_MARK(i) do { if (tid == 0 && blockIdx.x == 0) { long long now_ = clock64(); gj_dbg[i] += now_ - mark_; mark_ = now_; } } while (0)
extern "C" void kb_dbg_eigh_phases(long long *out, int reset) {
    cudaMemcpyFromSymbol(out, gj_dbg, sizeof(long long) * 8);
    if (reset) { long long z[8] = {0}; cudaMemcpyToSymbol(gj_dbg, z, sizeof(z)); }
}
#else
#define GJ_MARK(i) do {} while (0)
#endif

template <int M>
__global__ void __launch_bounds__(GJ_THREADS, 1)
jacobi_gram_kernel(int n, double *Gt, double *Vt, int *blkstep, EighShared *sh, int *sweep_rot, double *hblk, double tol, int with_v,
                   int inner, int max_sweeps) {
    constexpr int JBk = M / 2;         // columns per block
    constexpr int NB8 = M / 8;         // 8-wide tiles across the M columns
    constexpr int HS = M * (M + 1);    // padded M x M
    extern __shared__ __align__(16) double gj_smem[];
    const int n16 = (n + 15) & ~15;
    const int LD = n16 + 4;  // column stride: = 4 mod 16 doubles, so both DMMA fragment patterns are bank-conflict free
    double *sG = gj_smem;
    double *sV = sG + (size_t)M * LD;                        // only when with_v
    double *sH = sV + (with_v ? (size_t)M * LD : 0);         // [2][M][M+1]
    double *sW = sH + 2 * HS;                                // [2][M][M+1]
    double *sPart = sW + 2 * HS;                             // [GJ_WARPS][M*M]
    const int nblk0 = (n + JBk - 1) / JBk;
    const int NB = nblk0 + (nblk0 & 1);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int g = lane >> 2, t4 = lane & 3;
    const double tol2 = tol * tol;
    cg::grid_group grid = cg::this_grid();
    // rows [n, n16) of every resident column stay zero for the whole kernel
    for (int e = tid; e < M * (n16 - n); e += GJ_THREADS) {
        const int s = e / (n16 - n), k = n + e % (n16 - n);
        sG[(size_t)s * LD + k] = 0.0;
        if (with_v) sV[(size_t)s * LD + k] = 0.0;
    }
    int t = 0;
#ifdef KB_EIGH_DEBUG
    long long mark_ = clock64();
#endif
    for (int sweep = 0; sweep < max_sweeps; sweep++) {
        int my_rot = 0;
        for (int r = 0; r < NB - 1; r++, t++) {
            for (int i = blockIdx.x; i < NB / 2; i += gridDim.x) {
                int P, Q;
                gj_rr_pair(NB, r, i, P, Q);
                auto col_of = [&](int s) { return (s < JBk ? P * JBk + s : Q * JBk + (s - JBk)); };
                // ---- wait for and load the two blocks one after the other (L2 -> shared memory directly), so the
                // first block's transfer overlaps the wait for the second
                for (int half = 0; half < 2; half++) {
                    if (tid == 0) while (gj_ld_acquire(&blkstep[half ? Q : P]) < t) {}
                    __syncthreads();
                    if ((n & 1) == 0) {
#pragma unroll 4
                        for (int s = half * JBk; s < (half + 1) * JBk; s++) {
                            const int c = col_of(s);
                            for (int k = tid * 2; k < n; k += GJ_THREADS * 2) {
                                if (c < n) {
                                    gj_cp_async16(sG + (size_t)s * LD + k, Gt + (size_t)c * n + k);
                                    if (with_v) gj_cp_async16(sV + (size_t)s * LD + k, Vt + (size_t)c * n + k);
                                } else {
                                    sG[(size_t)s * LD + k] = 0.0; sG[(size_t)s * LD + k + 1] = 0.0;
                                    if (with_v) { sV[(size_t)s * LD + k] = 0.0; sV[(size_t)s * LD + k + 1] = 0.0; }
                                }
                            }
                        }
                    } else {
                        for (int e = tid; e < JBk * n; e += GJ_THREADS) {
                            const int s = half * JBk + e / n, k = e % n, c = col_of(s);
                            sG[(size_t)s * LD + k] = c < n ? __ldcg(Gt + (size_t)c * n + k) : 0.0;
                            if (with_v) sV[(size_t)s * LD + k] = c < n ? __ldcg(Vt + (size_t)c * n + k) : 0.0;
                        }
                    }
                }
                GJ_MARK(0);
                gj_cp_async_wait_all();
                __syncthreads();
                GJ_MARK(1);
                // ---- 1. Gram matrix: the k range is dealt to the warps four rows at a time, partials meet in smem.
                // Only the cross block G_P^T G_Q is new at every step: H_PP and H_QQ come out of the previous step's
                // two-sided update (H' = W^T H W is the Gram matrix of G W), so they are kept with the block in
                // global memory and recomputed from the columns once per sweep (r == 0) to stop rounding drift.
                const bool full = r == 0 || inner > 1 || NB8 == 1;
                const int hr = tid / M, hc = tid % M;  // this thread's entry of H and W (tid < M*M)
                const int lo = min(hr, hc), hi = max(hr, hc);
                const bool same_blk = (hr < JBk) == (hc < JBk);
                double h_kept = 0.0;
                if (!full && tid < M * M && same_blk)
                    h_kept = __ldcg(hblk + (size_t)(hr < JBk ? P : Q) * (JBk * JBk) + (lo % JBk) * JBk + hi % JBk);
                {
                    // two independent accumulator sets: consecutive DMMAs into one accumulator serialise on its latency
                    double acc[2][NB8][NB8][2];
#pragma unroll
                    for (int u = 0; u < 2; u++)
#pragma unroll
                        for (int a = 0; a < NB8; a++)
#pragma unroll
                            for (int b = 0; b < NB8; b++) acc[u][a][b][0] = acc[u][a][b][1] = 0.0;
                    for (int k0 = warp * 4; k0 < n16; k0 += GJ_WARPS * 8) {
                        double f[2][NB8];
                        const int k1 = k0 + GJ_WARPS * 4;
#pragma unroll
                        for (int b = 0; b < NB8; b++) {
                            f[0][b] = sG[(size_t)(b * 8 + g) * LD + k0 + t4];
                            f[1][b] = k1 < n16 ? sG[(size_t)(b * 8 + g) * LD + k1 + t4] : 0.0;
                        }
#pragma unroll
                        for (int a = 0; a < NB8; a++)
#pragma unroll
                            for (int b = a; b < NB8; b++) {  // upper tiles only: H is symmetric
                                if (!full && a == b) continue;  // diagonal blocks travel with their columns
                                gj_dmma(acc[0][a][b][0], acc[0][a][b][1], f[0][a], f[0][b]);
                                gj_dmma(acc[1][a][b][0], acc[1][a][b][1], f[1][a], f[1][b]);
                            }
                    }
#pragma unroll
                    for (int a = 0; a < NB8; a++)
#pragma unroll
                        for (int b = 0; b < NB8; b++) {
                            sPart[warp * (M * M) + (a * 8 + g) * M + b * 8 + 2 * t4] = acc[0][a][b][0] + acc[1][a][b][0];
                            sPart[warp * (M * M) + (a * 8 + g) * M + b * 8 + 2 * t4 + 1] = acc[0][a][b][1] + acc[1][a][b][1];
                        }
                }
                __syncthreads();
                if (tid < M * M) {
                    // symmetric by construction: entry (r, c) and (c, r) add the same eight numbers in the same order
                    double s = h_kept;
                    if (full || !same_blk) {
                        s = 0.0;
#pragma unroll
                        for (int w = 0; w < GJ_WARPS; w++) s += sPart[w * (M * M) + lo * M + hi];
                    }
                    sH[hr * (M + 1) + hc] = s;
                    sW[hr * (M + 1) + hc] = hr == hc ? 1.0 : 0.0;
                }
                __syncthreads();
                GJ_MARK(2);
                // nothing to rotate in this block pair (the usual case in the last sweeps): skip the serial part
                int todo = 0;
                if (tid < M * M && hr < hc && (full || !same_blk)) {
                    const double c = sH[hr * (M + 1) + hc];
                    todo = c * c > tol2 * sH[hr * (M + 1) + hr] * sH[hc * (M + 1) + hc];
                }
                todo = __syncthreads_or(todo);
                // ---- 2. cyclic two-sided Jacobi on H, ping-pong buffers, one barrier per parallel step
                // Pairs visited: every cross pair (one column of P, one of Q) in JBk parallel steps; the pairs inside a
                // block only in the first round of a sweep, where every block takes part exactly once -- so a sweep
                // still visits each of the n(n-1)/2 column pairs once, and the serial chain per outer step is JBk
                // steps instead of M - 1.  (inner > 1 repeats the full tournament; for experiments.)
                int cur = 0, cnt = 0;
                const int n_within = (r == 0 || inner > 1) ? JBk - 1 : 0;
                const int n_steps = !todo ? 0 : inner > 1 ? inner * (M - 1) : n_within + JBk;
                {
                    for (int lr = 0; lr < n_steps; lr++) {
                        if (tid < M * M) {
                            const double *H = sH + cur * HS, *W = sW + cur * HS;
                            double *Hn = sH + (cur ^ 1) * HS, *Wn = sW + (cur ^ 1) * HS;
                            int pr, pc;
                            if (inner > 1) { pr = gj_partner(M, lr % (M - 1), hr); pc = gj_partner(M, lr % (M - 1), hc); }
                            else if (lr < n_within) {
                                pr = hr < JBk ? gj_partner(JBk, lr, hr) : JBk + gj_partner(JBk, lr, hr - JBk);
                                pc = hc < JBk ? gj_partner(JBk, lr, hc) : JBk + gj_partner(JBk, lr, hc - JBk);
                            } else {
                                const int s = lr - n_within;
                                pr = hr < JBk ? JBk + (hr + s) % JBk : (hr - s + JBk) % JBk;  // (hr - JBk - s) mod JBk
                                pc = hc < JBk ? JBk + (hc + s) % JBk : (hc - s + JBk) % JBk;
                            }
                            // each thread works out the rotation of its COLUMN's pair; the rotation of its row's pair
                            // is the one of the lane (in the same group of M lanes) whose column is this row
                            double cr, sr, cc, sc;
                            const double h00 = H[hr * (M + 1) + hc], h01 = H[hr * (M + 1) + pc];
                            const double h10 = H[pr * (M + 1) + hc], h11 = H[pr * (M + 1) + pc];
                            const double w0 = W[hr * (M + 1) + hc], w1 = W[hr * (M + 1) + pc];
                            const bool rot_c = gj_rotation<M>(H, hc, pc, tol2, cc, sc);
                            const int src = (lane & ~(M - 1)) | hr;
                            cr = __shfl_sync(0xffffffffu, cc, src);
                            sr = __shfl_sync(0xffffffffu, sc, src);
                            double h = cr * fma(cc, h00, sc * h01) + sr * fma(cc, h10, sc * h11);
                            if (pr == hc && rot_c) { h = 0.0; if (hr < hc) cnt++; }
                            Hn[hr * (M + 1) + hc] = h;
                            Wn[hr * (M + 1) + hc] = fma(cc, w0, sc * w1);
                        }
                        __syncthreads();
                        cur ^= 1;
                    }
                }
                const int any = __syncthreads_or(cnt);
                GJ_MARK(3);
                my_rot += cnt;
                // the diagonal blocks of the updated H go with their columns (NB8 == 1 recomputes them every step)
                if (NB8 > 1 && (any || full) && tid < M * M && same_blk)
                    __stcg(hblk + (size_t)(hr < JBk ? P : Q) * (JBk * JBk) + (hr % JBk) * JBk + hc % JBk,
                           sH[cur * HS + hr * (M + 1) + hc]);
                if (any) {
                    // ---- 3. G <- G W (and V <- V W): a warp takes 8 rows at a time, reads all M columns of them
                    // into fragments, multiplies, writes the same 8 rows back
                    const double *W = sW + cur * HS;
                    double wf[M / 4][NB8];
#pragma unroll
                    for (int kb = 0; kb < M / 4; kb++)
#pragma unroll
                        for (int b = 0; b < NB8; b++) wf[kb][b] = W[(kb * 4 + t4) * (M + 1) + b * 8 + g];
                    for (int pass = 0; pass < (with_v ? 2 : 1); pass++) {
                        double *X = pass ? sV : sG;
                        for (int r0 = warp * 8; r0 < n16; r0 += GJ_WARPS * 16) {
                            const int r1 = r0 + GJ_WARPS * 8;  // second, independent row group (latency hiding)
                            const bool two = r1 < n16;
                            double af[2][M / 4], acc[2][NB8][2];
#pragma unroll
                            for (int kb = 0; kb < M / 4; kb++) {
                                af[0][kb] = X[(size_t)(kb * 4 + t4) * LD + r0 + g];
                                af[1][kb] = two ? X[(size_t)(kb * 4 + t4) * LD + r1 + g] : 0.0;
                            }
#pragma unroll
                            for (int b = 0; b < NB8; b++) acc[0][b][0] = acc[0][b][1] = acc[1][b][0] = acc[1][b][1] = 0.0;
#pragma unroll
                            for (int kb = 0; kb < M / 4; kb++)
#pragma unroll
                                for (int b = 0; b < NB8; b++) {
                                    gj_dmma(acc[0][b][0], acc[0][b][1], af[0][kb], wf[kb][b]);
                                    gj_dmma(acc[1][b][0], acc[1][b][1], af[1][kb], wf[kb][b]);
                                }
                            // results go straight from the accumulators to global memory (the next owner of the
                            // block reads them from there; this CTA reloads whatever it needs at its next step)
#pragma unroll
                            for (int b = 0; b < NB8; b++)
#pragma unroll
                                for (int e = 0; e < 2; e++) {
                                    const int c = col_of(b * 8 + 2 * t4 + e);
                                    if (c >= n) continue;
                                    double *dst = (pass ? Vt : Gt) + (size_t)c * n;
                                    if (r0 + g < n) __stcg(dst + r0 + g, acc[0][b][e]);
                                    if (two && r1 + g < n) __stcg(dst + r1 + g, acc[1][b][e]);
                                }
                        }
                    }
                    GJ_MARK(4);
                }
                __syncthreads();
                GJ_MARK(5);
                if (tid == 0) { gj_st_release(&blkstep[P], t + 1); gj_st_release(&blkstep[Q], t + 1); }
            }
        }
        if (my_rot) atomicAdd(&sweep_rot[sweep], my_rot);
        grid.sync();
        GJ_MARK(6);
        const int rot = __ldcg(&sweep_rot[sweep]);
        if (blockIdx.x == 0 && tid == 0) sh->sweeps = sweep + 1;
        if (rot == 0) break;
    }
}

static size_t gj_smem_bytes(int n, int M, int with_v) {
    const int n16 = (n + 15) & ~15, LD = n16 + 4;
    return sizeof(double) * ((size_t)M * LD * (with_v ? 2 : 1) + 4 * (size_t)M * (M + 1) + (size_t)GJ_WARPS * M * M);
}

template <int M>
static int gj_launch(kb_ctx *ctx, int n, double *Gt, double *Vt, int *blkstep, EighShared *sh, int *sweep_rot, double *hblk, double tol,
                     int with_v, int inner, int max_sweeps, cudaStream_t st) {
    const size_t smem = gj_smem_bytes(n, M, with_v);
    KB_CUDA(ctx, cudaFuncSetAttribute(jacobi_gram_kernel<M>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    const int JBk = M / 2, nblk0 = (n + JBk - 1) / JBk, NB = nblk0 + (nblk0 & 1);
    int blocks = NB / 2, per_sm = 0;
    KB_CUDA(ctx, cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, jacobi_gram_kernel<M>, GJ_THREADS, smem));
    const int cap = per_sm * ctx->prop.multiProcessorCount;
    if (cap < 1) return kb_fail(ctx, KB_ERR_CUDA, "jacobi_gram_kernel: no resident CTA with %zu B of shared memory", smem);
    if (blocks > cap) blocks = cap;
    void *args[] = {&n, &Gt, &Vt, &blkstep, &sh, &sweep_rot, &hblk, &tol, &with_v, &inner, &max_sweeps};
    KB_CUDA(ctx, cudaLaunchCooperativeKernel((void *)jacobi_gram_kernel<M>, dim3((unsigned)blocks), dim3(GJ_THREADS), args, smem, st));
    ctx->launches++;
    return KB_OK;
}

// Returns KB_OK and *launched = true when the Gram-form block kernel ran; *launched = false when the columns of a
// block pair do not fit into shared memory (the caller falls back to its streaming kernels).
int kb_jacobi_gram(kb_ctx *ctx, int n, double *Gt, double *Vt, int *blkstep, EighShared *sh, int *sweep_rot, double *hblk, double tol,
                   int with_v, int max_sweeps, cudaStream_t st, bool *launched) {
    *launched = false;
    int M = 16, inner = 1;
    if (const char *e = getenv("KB_EIGH_M")) M = atoi(e);
    if (const char *e = getenv("KB_EIGH_INNER")) inner = std::max(1, atoi(e));
    if (M != 8 && M != 16) return KB_OK;  // KB_EIGH_M=0: the column-wise kernel of kb_linalg.cu
    const size_t lim = (size_t)ctx->prop.sharedMemPerBlockOptin;
    if (M == 16 && gj_smem_bytes(n, 16, with_v) > lim) M = 8;
    if (gj_smem_bytes(n, M, with_v) > lim) return KB_OK;
    const int rc = M == 16 ? gj_launch<16>(ctx, n, Gt, Vt, blkstep, sh, sweep_rot, hblk, tol, with_v, inner, max_sweeps, st)
                           : gj_launch<8>(ctx, n, Gt, Vt, blkstep, sh, sweep_rot, hblk, tol, with_v, inner, max_sweeps, st);
    if (rc == KB_OK) *launched = true;
    return rc;
}
