// kb_tpe_score.cu -- TPE candidate scoring and selection: the hot path.
//
// Replaces, per parameter, the tail of goptuna's tpe.Sampler.sampleNumerical / sampleCategorical
// (module not in the reference tree, go.mod:8; reached from sample.go:48-90):
//   gmmLogPDF(samples, below) - gmmLogPDF(samples, above)   -> tpe_score_cont_kernel (q = 0)
//                                                              tpe_score_quant_kernel (q > 0)
//   categoricalLogPDF                                        -> inside tpe_stage_kernel (table look-up)
//   compare (argmax, first maximum wins)                     -> running FP32 max + near-tie table, then
//                                                              tpe_rescore_kernel (FP64) + tpe_pick_kernel
//
// Data flow per chunk of candidates (row-major f64, as the caller has them):
//   tpe_stage_kernel   transposes through shared memory into per-parameter columns, normalised to
//                      [0,1] and split into an FP32 (hi, lo) pair so that x - mu keeps ~48 bits;
//   tpe_score_*_kernel one CTA = (2048 or 1024 candidates) x (one parameter).  The parameter's folded
//                      Parzen table streams through shared memory in 16 KB stages filled by TMA bulk
//                      copies (cp.async.bulk + mbarrier), every component is read once per CTA as a
//                      broadcast LDS.128 and applied to R candidates held in registers.
//                      Per (candidate, component): 2 FADD + 2 FFMA + MUFU.EX2 + FADD  -> SFU bound.
//                      Partial sums are FP32 over 32 components, then added into FP64.
// Selection keeps, per parameter, every distinct candidate value whose FP32 score lies within
// KB_MARGIN of the running maximum (open-addressing table keyed by the f64 bits, lowest index per
// value); those few are re-scored in FP64 in the oracle's formula order and the best, lowest-index
// one wins -- the "identical selected candidate" rule of the parity contract.
#include <math.h>

#include "kb_internal.cuh"

#define KB_EPS 1e-12
#define KB_MARGIN_REL 3e-5f

// ---------------------------------------------------------------------------------------------------------
// mbarrier / TMA bulk-copy helpers (sm_90+; SASS: UBLKCP + SYNCS)
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() { asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); }
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "WAIT_%=:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra DONE_%=;\n"
        "bra WAIT_%=;\n"
        "DONE_%=:\n"
        "}\n" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tma_bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes, uint64_t *bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

// ---------------------------------------------------------------------------------------------------------
// stage: row-major f64 candidates -> per-parameter normalised FP32 columns; categorical parameters are
// scored here (a table look-up) since every value passes through this kernel anyway.
// ---------------------------------------------------------------------------------------------------------
#define ST_ROWS 128
#define ST_DG 32
#define ST_THREADS 256

__global__ void __launch_bounds__(ST_THREADS)
tpe_stage_kernel(const KbDim *__restrict__ dims, int D, const double *__restrict__ cand, int n, long long chunk_cap,
                 float2 *__restrict__ stage_cont, float4 *__restrict__ stage_quant, const int *__restrict__ slot_of_dim,
                 KbSelect *sel, long long local_base, double *__restrict__ scores_out) {
    __shared__ double tile[ST_DG][ST_ROWS + 1];
    const int row0 = blockIdx.x * ST_ROWS;
    const int rows = min(ST_ROWS, n - row0);
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int d0 = 0; d0 < D; d0 += ST_DG) {
        const int nd = min(ST_DG, D - d0);
        for (int r = warp; r < rows; r += ST_THREADS / 32)
            if (lane < nd) tile[lane][r] = cand[(size_t)(row0 + r) * D + d0 + lane];
        __syncthreads();
        for (int dd = warp; dd < nd; dd += ST_THREADS / 32) {
            const int d = d0 + dd;
            const KbDim &dm = dims[d];
            const int slot = slot_of_dim[d];  // index inside the cont / quant / cat list
            if (dm.kind == KB_CATEGORICAL) {
                unsigned long long best = 0ull;
                for (int r = lane; r < rows; r += 32) {
                    int c = (int)tile[dd][r];
                    c = c < 0 ? 0 : (c >= dm.n_choices ? dm.n_choices - 1 : c);
                    if (scores_out) scores_out[(size_t)(row0 + r) * D + d] = dm.cat_score[c];
                    // (rank+1 : 13 bits) | (~call-local index : 32 bits) | (choice : 12 bits)
                    unsigned long long key = ((unsigned long long)(unsigned)(dm.cat_rank[c] + 1) << 44) |
                                             ((unsigned long long)(~(unsigned)(local_base + row0 + r)) << 12) |
                                             (unsigned long long)c;
                    best = key > best ? key : best;
                }
                if (sel) {
                    for (int o = 16; o > 0; o >>= 1) {
                        unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
                        best = other > best ? other : best;
                    }
                    if (lane == 0 && best) atomicMax(&sel[d].cat_best, best);
                }
            } else {
                const double lo = dm.lo, hi = dm.hi, inv = 1.0 / (hi - lo), q = dm.q;
                for (int r = lane; r < rows; r += 32) {
                    const double x = tile[dd][r];
                    if (q > 0.0) {
                        double ub = x + q / 2.0; ub = ub > hi ? hi : ub;
                        double lb = x - q / 2.0; lb = lb < lo ? lo : lb;
                        double un = (ub - lo) * inv, ln = (lb - lo) * inv;
                        float uh = (float)un, lh = (float)ln;
                        stage_quant[(size_t)slot * chunk_cap + row0 + r] =
                            make_float4(lh, (float)(ln - (double)lh), uh, (float)(un - (double)uh));
                    } else {
                        double xn = (x - lo) * inv;
                        float xh = (float)xn;
                        stage_cont[(size_t)slot * chunk_cap + row0 + r] = make_float2(xh, (float)(xn - (double)xh));
                    }
                }
            }
        }
        __syncthreads();
    }
}

// ---------------------------------------------------------------------------------------------------------
// near-tie table insert
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tie_insert(KbTie *tab, float *tab_score, KbSelect *sl, unsigned long long key,
                                           unsigned long long gidx, float score) {
    unsigned long long h = key * 0x9E3779B97F4A7C15ull;
    unsigned int s = (unsigned int)(h >> 40) & (KB_TIE_SLOTS - 1);
    for (unsigned int probe = 0; probe < KB_TIE_SLOTS / 2; probe++) {
        unsigned long long prev = atomicCAS(&tab[s].key, KB_TIE_EMPTY, key);
        if (prev == KB_TIE_EMPTY || prev == key) {
            atomicMin(&tab[s].minidx, gidx);
            tab_score[s] = score;  // every writer of this key stores the same bits
            return;
        }
        s = (s + 1) & (KB_TIE_SLOTS - 1);
    }
    atomicExch(&sl->overflow, 1u);
}

// block-wide: publish the block maximum, derive the admission threshold, insert the near ties
template <int R, int THREADS>
__device__ __forceinline__ void select_publish(const float (&sc)[R], int tile_base, int n, long long idx_base, int d,
                                               int D, int numeric_slot, KbSelect *sel, KbTie *tie, float *tie_score,
                                               const double *__restrict__ cand_raw) {
    __shared__ float s_wmax[THREADS / 32];
    __shared__ float s_thr;
    float best = -INFINITY;
#pragma unroll
    for (int r = 0; r < R; r++) {
        int i = tile_base + r * THREADS + threadIdx.x;
        if (i < n) best = fmaxf(best, sc[r]);
    }
    for (int o = 16; o > 0; o >>= 1) best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((threadIdx.x & 31) == 0) s_wmax[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        float b = s_wmax[0];
        for (int w = 1; w < THREADS / 32; w++) b = fmaxf(b, s_wmax[w]);
        unsigned int mine = kb_f32_orderable(b);
        unsigned int old = atomicMax(&sel[d].gmax_bits, mine);
        float g = kb_f32_from_orderable(old > mine ? old : mine);
        float thr = g - KB_MARGIN_REL * fmaxf(1.0f, fabsf(g));
        s_thr = b >= thr ? thr : INFINITY;  // block has nothing near the maximum: skip the scan
    }
    __syncthreads();
    const float thr = s_thr;
    if (thr == INFINITY) return;
    KbTie *tab = tie + (size_t)numeric_slot * KB_TIE_SLOTS;
    float *tsc = tie_score + (size_t)numeric_slot * KB_TIE_SLOTS;
#pragma unroll
    for (int r = 0; r < R; r++) {
        int i = tile_base + r * THREADS + threadIdx.x;
        if (i < n && sc[r] >= thr) {
            unsigned long long key = (unsigned long long)__double_as_longlong(cand_raw[(size_t)i * D + d]);
            tie_insert(tab, tsc, &sel[d], key, (unsigned long long)(idx_base + i), sc[r]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// continuous parameters (q = 0): log-sum-exp over the below and above mixtures
// ---------------------------------------------------------------------------------------------------------
#define SC_THREADS 256
#define SC_R 8
#define SC_TILE (SC_THREADS * SC_R)
#define SC_CHUNK 1024  // components per shared-memory stage (16 KB)
#define SC_STAGES 2
#define SC_GROUP 32    // components per FP32 partial sum

__device__ __forceinline__ void table_stream_issue(const KbDim &dm, int chunk, int nchunk0, float4 *dst, uint64_t *bar) {
    const int side = chunk < nchunk0 ? 0 : 1;
    const int c = side == 0 ? chunk : chunk - nchunk0;
    const int cnt = min(SC_CHUNK, dm.Kpad[side] - c * SC_CHUNK);
    const uint32_t bytes = (uint32_t)cnt * (uint32_t)sizeof(float4);
    mbar_expect_tx(bar, bytes);
    tma_bulk_g2s(dst, dm.tab[side] + (size_t)c * SC_CHUNK, bytes, bar);
}

__global__ void __launch_bounds__(SC_THREADS, 3)
tpe_score_cont_kernel(const KbDim *__restrict__ dims, const int *__restrict__ cont_dims,
                      const float2 *__restrict__ stage, long long chunk_cap, int n, long long idx_base, int D,
                      KbSelect *sel, KbTie *tie, float *tie_score, const double *__restrict__ cand_raw,
                      double *__restrict__ scores_out) {
    __shared__ __align__(128) float4 s_tab[SC_STAGES][SC_CHUNK];
    __shared__ __align__(8) uint64_t s_bar[SC_STAGES];
    const int slot = blockIdx.y;
    const int d = cont_dims[slot];
    const KbDim &dm = dims[d];
    const int tile_base = blockIdx.x * SC_TILE;
    const int nchunk0 = (dm.Kpad[0] + SC_CHUNK - 1) / SC_CHUNK, nchunk1 = (dm.Kpad[1] + SC_CHUNK - 1) / SC_CHUNK;
    const int nchunks = nchunk0 + nchunk1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < SC_STAGES; s++) mbar_init(&s_bar[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int c = 0; c < SC_STAGES && c < nchunks; c++) table_stream_issue(dm, c, nchunk0, s_tab[c], &s_bar[c]);

    float xh[SC_R], xl[SC_R];
#pragma unroll
    for (int r = 0; r < SC_R; r++) {
        int i = tile_base + r * SC_THREADS + threadIdx.x;
        float2 v = i < n ? stage[(size_t)slot * chunk_cap + i] : make_float2(0.5f, 0.f);
        xh[r] = v.x; xl[r] = v.y;
    }
    double S[SC_R], logS0[SC_R];
#pragma unroll
    for (int r = 0; r < SC_R; r++) S[r] = 0.0;

    for (int c = 0; c < nchunks; c++) {
        const int st = c % SC_STAGES;
        const int side = c < nchunk0 ? 0 : 1;
        const int cc = side == 0 ? c : c - nchunk0;
        const int cnt = min(SC_CHUNK, dm.Kpad[side] - cc * SC_CHUNK);
        mbar_wait(&s_bar[st], (uint32_t)((c / SC_STAGES) & 1));
        const float4 *tab = s_tab[st];
        for (int j0 = 0; j0 < cnt; j0 += SC_GROUP) {
            float a32[SC_R];
#pragma unroll
            for (int r = 0; r < SC_R; r++) a32[r] = 0.f;
#pragma unroll 4
            for (int j = 0; j < SC_GROUP; j++) {
                const float4 e = tab[j0 + j];  // {mu_hi, c', -mu_lo*c', k'}: one broadcast LDS.128
#pragma unroll
                for (int r = 0; r < SC_R; r++) {
                    float u = xh[r] - e.x;
                    u += xl[r];
                    float t = fmaf(u, e.y, e.z);
                    a32[r] += kb_ex2(fmaf(-t, t, e.w));
                }
            }
#pragma unroll
            for (int r = 0; r < SC_R; r++) S[r] += (double)a32[r];
        }
        __syncthreads();  // every thread is done with stage st
        if (threadIdx.x == 0 && c + SC_STAGES < nchunks)
            table_stream_issue(dm, c + SC_STAGES, nchunk0, s_tab[st], &s_bar[st]);
        if (c == nchunk0 - 1) {
#pragma unroll
            for (int r = 0; r < SC_R; r++) { logS0[r] = log(S[r]) + dm.L[0]; S[r] = 0.0; }
        }
    }
    float sc[SC_R];
#pragma unroll
    for (int r = 0; r < SC_R; r++) {
        double s64 = logS0[r] - (log(S[r]) + dm.L[1]);
        sc[r] = (float)s64;
        int i = tile_base + r * SC_THREADS + threadIdx.x;
        if (scores_out && i < n) scores_out[(size_t)i * D + d] = s64;
    }
    if (sel) select_publish<SC_R, SC_THREADS>(sc, tile_base, n, idx_base, d, D, dm.numeric_slot, sel, tie, tie_score, cand_raw);
}

// ---------------------------------------------------------------------------------------------------------
// quantised parameters (q > 0): sum_j w_j (Phi(ub) - Phi(lb)) per side, written as erfc differences so that
// the tails keep relative accuracy in FP32
// ---------------------------------------------------------------------------------------------------------
#define SQ_R 4
#define SQ_TILE (SC_THREADS * SQ_R)
#define LN2F 0.6931471806f
#define KINVF 0.8325546112f           // 1/sqrt(log2 e)
#define QUANT_WIDE_THR 0.2402245f     // (b - a) <= 0.2 in erf units, scaled by sqrt(log2 e)

__device__ __forceinline__ float kb_rcp(float x) {
    float y;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// erfc(x) for x >= 0, argument pre-scaled: xs = sqrt(log2 e) x.  erfcx(x) = t P6(t), t = 1/(x+2) (near-minimax
// fit, 3.6e-7 relative on [0, 6.6]); erfc = erfcx * 2^{-xs^2}.  1 MUFU.RCP + 1 MUFU.EX2 + 10 FP32 ops.
__device__ __forceinline__ float erfc_pos_scaled(float xs) {
    const float t = kb_rcp(fmaf(xs, KINVF, 2.0f));
    float p = 1.219882546e+01f;
    p = fmaf(p, t, -2.445210690e+01f);
    p = fmaf(p, t, 1.361089618e+01f);
    p = fmaf(p, t, 1.865695796e-01f);
    p = fmaf(p, t, 2.333518575e+00f);
    p = fmaf(p, t, 1.102339077e+00f);
    p = fmaf(p, t, 5.649710102e-01f);
    return (p * t) * kb_ex2(-xs * xs);
}

__global__ void __launch_bounds__(SC_THREADS, 3)
tpe_score_quant_kernel(const KbDim *__restrict__ dims, const int *__restrict__ quant_dims,
                       const float4 *__restrict__ stage, long long chunk_cap, int n, long long idx_base, int D,
                       KbSelect *sel, KbTie *tie, float *tie_score, const double *__restrict__ cand_raw,
                       double *__restrict__ scores_out) {
    __shared__ __align__(128) float4 s_tab[SC_STAGES][SC_CHUNK];
    __shared__ __align__(8) uint64_t s_bar[SC_STAGES];
    const int slot = blockIdx.y;
    const int d = quant_dims[slot];
    const KbDim &dm = dims[d];
    const int tile_base = blockIdx.x * SQ_TILE;
    const int nchunk0 = (dm.Kpad[0] + SC_CHUNK - 1) / SC_CHUNK, nchunk1 = (dm.Kpad[1] + SC_CHUNK - 1) / SC_CHUNK;
    const int nchunks = nchunk0 + nchunk1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < SC_STAGES; s++) mbar_init(&s_bar[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int c = 0; c < SC_STAGES && c < nchunks; c++) table_stream_issue(dm, c, nchunk0, s_tab[c], &s_bar[c]);

    float lh[SQ_R], ll[SQ_R], uh[SQ_R], ul[SQ_R];
#pragma unroll
    for (int r = 0; r < SQ_R; r++) {
        int i = tile_base + r * SC_THREADS + threadIdx.x;
        float4 v = i < n ? stage[(size_t)slot * chunk_cap + i] : make_float4(0.4f, 0.f, 0.6f, 0.f);
        lh[r] = v.x; ll[r] = v.y; uh[r] = v.z; ul[r] = v.w;
    }
    double S[SQ_R], P0[SQ_R];
#pragma unroll
    for (int r = 0; r < SQ_R; r++) S[r] = 0.0;
    const float qn = (float)(dm.q / (dm.hi - dm.lo));

    for (int c = 0; c < nchunks; c++) {
        const int st = c % SC_STAGES;
        const int side = c < nchunk0 ? 0 : 1;
        const int cc = side == 0 ? c : c - nchunk0;
        const int cnt = min(SC_CHUNK, dm.Kpad[side] - cc * SC_CHUNK);
        mbar_wait(&s_bar[st], (uint32_t)((c / SC_STAGES) & 1));
        const float4 *tab = s_tab[st];
        for (int j0 = 0; j0 < cnt; j0 += SC_GROUP) {
            float a32[SQ_R];
#pragma unroll
            for (int r = 0; r < SQ_R; r++) a32[r] = 0.f;
#pragma unroll 2
            for (int j = 0; j < SC_GROUP; j++) {
                const float4 e = tab[j0 + j];  // {mu_hi, r', -mu_lo*r', w},  r' = sqrt(log2 e) / (sqrt2 sigma_n)
                // erfc arguments scaled by sqrt(log2 e): a' = k a, b' = k b, so that e^{-a^2} = 2^{-a'^2}
                if (e.y * qn < QUANT_WIDE_THR) {
                    // bin narrow against sigma (b - a <= 0.2): erfc(a) - erfc(b) would cancel, so integrate the
                    // Gaussian over the bin by its Taylor series about the midpoint (Hermite terms to h^6)
#pragma unroll
                    for (int r = 0; r < SQ_R; r++) {
                        const float a = fmaf((lh[r] - e.x) + ll[r], e.y, e.z);
                        const float b = fmaf((uh[r] - e.x) + ul[r], e.y, e.z);
                        const float m = 0.5f * (a + b), h = 0.5f * (b - a);
                        const float mm = m * m;
                        const float u = mm * LN2F, v = (h * h) * LN2F;
                        const float c1 = fmaf(u, 0.6666666667f, -0.3333333333f);
                        const float c2 = fmaf(fmaf(u, 0.1333333333f, -0.4f), u, 0.1f);
                        const float c3 = fmaf(fmaf(fmaf(u, 0.0126984127f, -0.0952380952f), u, 0.1428571429f), u, -0.0238095238f);
                        const float S = fmaf(fmaf(fmaf(c3, v, c2), v, c1), v, 1.0f);
                        const float d2 = (h * (2.2567583342f * KINVF)) * S * kb_ex2(-mm);  // (4/sqrt(pi)) h S e^{-m^2}
                        a32[r] = fmaf(e.w, d2, a32[r]);
                    }
                } else {
#pragma unroll
                    for (int r = 0; r < SQ_R; r++) {
                        const float a = fmaf((lh[r] - e.x) + ll[r], e.y, e.z);
                        const float b = fmaf((uh[r] - e.x) + ul[r], e.y, e.z);
                        // Phi(b)-Phi(a) = (erfc(a)-erfc(b))/2; mirror so that b >= |a| (erfc small, relative accuracy kept)
                        const bool flip = a + b < 0.f;
                        const float a2 = flip ? -b : a, b2 = flip ? -a : b;
                        const float ea = erfc_pos_scaled(fabsf(a2)), eb = erfc_pos_scaled(b2);
                        const float ea2 = a2 < 0.f ? 2.0f - ea : ea;
                        a32[r] = fmaf(e.w, ea2 - eb, a32[r]);
                    }
                }
            }
#pragma unroll
            for (int r = 0; r < SQ_R; r++) S[r] += (double)a32[r];
        }
        __syncthreads();
        if (threadIdx.x == 0 && c + SC_STAGES < nchunks)
            table_stream_issue(dm, c + SC_STAGES, nchunk0, s_tab[st], &s_bar[st]);
        if (c == nchunk0 - 1) {
#pragma unroll
            for (int r = 0; r < SQ_R; r++) { P0[r] = log(0.5 * S[r] + KB_EPS) - log(dm.p_accept[0] + KB_EPS); S[r] = 0.0; }
        }
    }
    float sc[SQ_R];
#pragma unroll
    for (int r = 0; r < SQ_R; r++) {
        double s64 = P0[r] - (log(0.5 * S[r] + KB_EPS) - log(dm.p_accept[1] + KB_EPS));
        sc[r] = (float)s64;
        int i = tile_base + r * SC_THREADS + threadIdx.x;
        if (scores_out && i < n) scores_out[(size_t)i * D + d] = s64;
    }
    if (sel) select_publish<SQ_R, SC_THREADS>(sc, tile_base, n, idx_base, d, D, dm.numeric_slot, sel, tie, tie_score, cand_raw);
}

// ---------------------------------------------------------------------------------------------------------
// FP64 re-scoring of the near ties (oracle formula order: max, sum of exp, log) and the final pick
// ---------------------------------------------------------------------------------------------------------
#define RS_THREADS 256
#define RS_BLOCKS 8

struct KbPartial { double score; unsigned long long idx; double value; };

__device__ double rescore_value(const KbDim &dm, double x, int lane) {
    double out[2];
    for (int side = 0; side < 2; side++) {
        const int K = dm.K[side];
        const double *mu = dm.mu[side], *sg = dm.sigma[side];
        if (dm.q > 0.0) {
            const double *w = dm.w[side];
            double ub = x + dm.q / 2.0; ub = ub > dm.hi ? dm.hi : ub;
            double lb = x - dm.q / 2.0; lb = lb < dm.lo ? dm.lo : lb;
            double p = 0.0;
            for (int j = lane; j < K; j += 32) {
                double den = sqrt(2.0) * sg[j];
                den = den < KB_EPS ? KB_EPS : den;
                p += w[j] * (0.5 * (1.0 + erf((ub - mu[j]) / den)) - 0.5 * (1.0 + erf((lb - mu[j]) / den)));
            }
            for (int o = 16; o > 0; o >>= 1) p += __shfl_xor_sync(0xffffffffu, p, o);
            out[side] = log(p + KB_EPS) - log(dm.p_accept[side] + KB_EPS);
        } else {
            const double *logc = dm.logc[side];
            double m = -INFINITY;
            for (int j = lane; j < K; j += 32) {
                double s = sg[j] > KB_EPS ? sg[j] : KB_EPS;
                double dd = (x - mu[j]) / s;
                m = fmax(m, -0.5 * (dd * dd) + logc[j]);
            }
            for (int o = 16; o > 0; o >>= 1) m = fmax(m, __shfl_xor_sync(0xffffffffu, m, o));
            double acc = 0.0;
            for (int j = lane; j < K; j += 32) {
                double s = sg[j] > KB_EPS ? sg[j] : KB_EPS;
                double dd = (x - mu[j]) / s;
                acc += exp(-0.5 * (dd * dd) + logc[j] - m);
            }
            for (int o = 16; o > 0; o >>= 1) acc += __shfl_xor_sync(0xffffffffu, acc, o);
            out[side] = log(acc) + m;
        }
    }
    return out[0] - out[1];
}

__global__ void __launch_bounds__(RS_THREADS)
tpe_rescore_kernel(const KbDim *__restrict__ dims, const int *__restrict__ numeric_dims, const KbSelect *__restrict__ sel,
                   const KbTie *__restrict__ tie, const float *__restrict__ tie_score, KbPartial *partial) {
    __shared__ KbPartial s_best[RS_THREADS / 32];
    const int slot = blockIdx.x, d = numeric_dims[slot];
    const KbDim &dm = dims[d];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const float g = kb_f32_from_orderable(sel[d].gmax_bits);
    const float thr = g - KB_MARGIN_REL * fmaxf(1.0f, fabsf(g));
    const KbTie *tab = tie + (size_t)slot * KB_TIE_SLOTS;
    const float *tsc = tie_score + (size_t)slot * KB_TIE_SLOTS;
    KbPartial best = {-INFINITY, ~0ull, 0.0};
    const int wglobal = blockIdx.y * (RS_THREADS / 32) + warp, wtotal = RS_BLOCKS * (RS_THREADS / 32);
    for (unsigned int s = wglobal; s < KB_TIE_SLOTS; s += wtotal) {
        const unsigned long long key = tab[s].key;
        if (key == KB_TIE_EMPTY || !(tsc[s] >= thr)) continue;  // warp-uniform
        const double x = __longlong_as_double((long long)key);
        const double sc = rescore_value(dm, x, lane);
        const unsigned long long idx = tab[s].minidx;
        if (sc > best.score || (sc == best.score && idx < best.idx)) { best.score = sc; best.idx = idx; best.value = x; }
    }
    if (lane == 0) s_best[warp] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < RS_THREADS / 32; w++) {
            KbPartial o = s_best[w];
            if (o.score > best.score || (o.score == best.score && o.idx < best.idx)) best = o;
        }
        partial[(size_t)slot * RS_BLOCKS + blockIdx.y] = best;
    }
}

__global__ void tpe_pick_kernel(const KbDim *__restrict__ dims, int D, const KbSelect *__restrict__ sel,
                                const KbPartial *__restrict__ partial, KbWinner *win, long long idx_base) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    const KbDim &dm = dims[d];
    KbWinner w;
    if (dm.kind == KB_CATEGORICAL) {
        const unsigned long long key = sel[d].cat_best;
        const int c = (int)(key & 0xFFFull);
        const unsigned int local = ~(unsigned int)((key >> 12) & 0xFFFFFFFFull);
        w.score = key ? dm.cat_score[c] : -INFINITY;
        w.index = key ? idx_base + (long long)local : -1;
        w.value = (double)c;
    } else {
        KbPartial best = {-INFINITY, ~0ull, 0.0};
        for (int b = 0; b < RS_BLOCKS; b++) {
            KbPartial o = partial[(size_t)dm.numeric_slot * RS_BLOCKS + b];
            if (o.score > best.score || (o.score == best.score && o.idx < best.idx)) best = o;
        }
        w.score = best.score;
        w.index = best.idx == ~0ull ? -1 : (long long)best.idx;
        double v = best.value;
        if (dm.kind == KB_DISCRETE) {  // sampleDiscreteUniform: shift back and clamp
            v = v + dm.shift;
            v = fmin(fmax(v, dm.ext_lo), dm.ext_hi);
        }
        w.value = v;
    }
    win[d] = w;
}

// pick the best of `world` gathered winner rows (score desc, index asc)
__global__ void tpe_merge_kernel(const KbWinner *__restrict__ all, int world, int D, KbWinner *win) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d >= D) return;
    KbWinner best = all[d];
    for (int r = 1; r < world; r++) {
        KbWinner o = all[(size_t)r * D + d];
        if (o.index < 0) continue;
        if (best.index < 0 || o.score > best.score || (o.score == best.score && o.index < best.index)) best = o;
    }
    win[d] = best;
}

// ---------------------------------------------------------------------------------------------------------
// host orchestration
// ---------------------------------------------------------------------------------------------------------
int kb_tpe_free_history(kb_ctx *ctx);

int kb_tpe_free(kb_ctx *ctx) {
    kb_tpe_free_history(ctx);
    for (int i = 0; i < 2; i++) { cudaFree(ctx->d_cand_chunk[i]); ctx->d_cand_chunk[i] = nullptr; }
    cudaFree(ctx->d_stage_cont); cudaFree(ctx->d_stage_quant); cudaFree(ctx->d_scores_chunk);
    cudaFree(ctx->d_sel); cudaFree(ctx->d_tie); cudaFree(ctx->d_win); cudaFree(ctx->d_win_all);
    ctx->d_stage_cont = nullptr; ctx->d_stage_quant = nullptr; ctx->d_scores_chunk = nullptr;
    ctx->d_sel = nullptr; ctx->d_tie = nullptr; ctx->d_win = nullptr; ctx->d_win_all = nullptr;
    ctx->chunk_cap = 0;
    cudaFree(ctx->d_tie_score); cudaFree(ctx->d_partial); cudaFree(ctx->d_slot_of_dim); cudaFree(ctx->d_numeric_dims);
    if (ctx->h_win) cudaFreeHost(ctx->h_win);
    ctx->d_tie_score = nullptr; ctx->d_partial = nullptr; ctx->d_slot_of_dim = nullptr; ctx->d_numeric_dims = nullptr;
    ctx->h_win = nullptr;
    return KB_OK;
}

static const int64_t KB_CHUNK_ROWS = 1 << 20;

// size the per-chunk scratch for `rows` candidates per chunk
static int tpe_ensure_scratch(kb_ctx *ctx, int64_t rows, bool need_landing, bool need_scores) {
    kb_ctx *sc = ctx;
    const int D = ctx->D;
    if (!ctx->d_sel) {
        KB_CUDA(ctx, cudaMalloc(&ctx->d_sel, sizeof(KbSelect) * (size_t)D));
        KB_CUDA(ctx, cudaMalloc(&ctx->d_win, sizeof(KbWinner) * (size_t)D));
        KB_CUDA(ctx, cudaMalloc(&ctx->d_win_all, sizeof(KbWinner) * (size_t)D * 64));
        const size_t slots = (size_t)(ctx->n_numeric > 0 ? ctx->n_numeric : 1);
        KB_CUDA(ctx, cudaMalloc(&ctx->d_tie, sizeof(KbTie) * slots * KB_TIE_SLOTS));
        KB_CUDA(ctx, cudaMalloc(&sc->d_tie_score, sizeof(float) * slots * KB_TIE_SLOTS));
        KB_CUDA(ctx, cudaMalloc(&sc->d_partial, sizeof(KbPartial) * slots * RS_BLOCKS));
        KB_CUDA(ctx, cudaMallocHost(&sc->h_win, sizeof(KbWinner) * (size_t)D));
        std::vector<int> slot_of((size_t)D, 0), numeric(slots, 0);
        for (int i = 0; i < ctx->n_cont; i++) slot_of[(size_t)ctx->cont_dims[(size_t)i]] = i;
        for (int i = 0; i < ctx->n_quant; i++) slot_of[(size_t)ctx->quant_dims[(size_t)i]] = i;
        for (int i = 0; i < ctx->n_cat; i++) slot_of[(size_t)ctx->cat_dims[(size_t)i]] = i;
        for (int d = 0; d < D; d++) if (ctx->h_dims[(size_t)d].numeric_slot >= 0) numeric[(size_t)ctx->h_dims[(size_t)d].numeric_slot] = d;
        KB_CUDA(ctx, cudaMalloc(&sc->d_slot_of_dim, sizeof(int) * (size_t)D));
        KB_CUDA(ctx, cudaMalloc(&sc->d_numeric_dims, sizeof(int) * slots));
        KB_CUDA(ctx, cudaMemcpy(sc->d_slot_of_dim, slot_of.data(), sizeof(int) * (size_t)D, cudaMemcpyHostToDevice));
        KB_CUDA(ctx, cudaMemcpy(sc->d_numeric_dims, numeric.data(), sizeof(int) * slots, cudaMemcpyHostToDevice));
    }
    if (rows > ctx->chunk_cap) {
        cudaFree(ctx->d_stage_cont); cudaFree(ctx->d_stage_quant);
        for (int i = 0; i < 2; i++) { cudaFree(ctx->d_cand_chunk[i]); ctx->d_cand_chunk[i] = nullptr; }
        cudaFree(ctx->d_scores_chunk);
        ctx->d_stage_cont = nullptr; ctx->d_stage_quant = nullptr; ctx->d_scores_chunk = nullptr;
        ctx->chunk_cap = 0;
        KB_CUDA(ctx, cudaMalloc(&ctx->d_stage_cont, sizeof(float2) * (size_t)rows * (size_t)(ctx->n_cont > 0 ? ctx->n_cont : 1)));
        KB_CUDA(ctx, cudaMalloc(&ctx->d_stage_quant, sizeof(float4) * (size_t)rows * (size_t)(ctx->n_quant > 0 ? ctx->n_quant : 1)));
        ctx->chunk_cap = rows;
    }
    if (need_landing && !ctx->d_cand_chunk[0])
        for (int i = 0; i < 2; i++) KB_CUDA(ctx, cudaMalloc(&ctx->d_cand_chunk[i], sizeof(double) * (size_t)ctx->chunk_cap * D));
    if (need_scores && !ctx->d_scores_chunk)
        KB_CUDA(ctx, cudaMalloc(&ctx->d_scores_chunk, sizeof(double) * (size_t)ctx->chunk_cap * D));
    return KB_OK;
}

// launch stage + scoring kernels for one resident chunk
static int tpe_launch_chunk(kb_ctx *ctx, const double *d_cand, int n, long long idx_base, long long local_base, bool select,
                            double *d_scores, cudaStream_t s) {
    kb_ctx *sc = ctx;
    const int D = ctx->D;
    KbSelect *sel = select ? ctx->d_sel : nullptr;
    { KbProf pf(ctx, 0, s);
    tpe_stage_kernel<<<(n + ST_ROWS - 1) / ST_ROWS, ST_THREADS, 0, s>>>(ctx->d_dims, D, d_cand, n, ctx->chunk_cap,
                                                                        ctx->d_stage_cont, ctx->d_stage_quant,
                                                                        sc->d_slot_of_dim, sel, local_base, d_scores);
    }
    KB_LAUNCH_CHECK(ctx);
    if (ctx->n_cont > 0) {
        dim3 grid((unsigned)((n + SC_TILE - 1) / SC_TILE), (unsigned)ctx->n_cont);
        KbProf pf(ctx, 1, s);
        tpe_score_cont_kernel<<<grid, SC_THREADS, 0, s>>>(ctx->d_dims, ctx->d_cont_dims, ctx->d_stage_cont, ctx->chunk_cap, n,
                                                          idx_base, D, sel, ctx->d_tie, sc->d_tie_score, d_cand, d_scores);
        KB_LAUNCH_CHECK(ctx);
    }
    if (ctx->n_quant > 0) {
        dim3 grid((unsigned)((n + SQ_TILE - 1) / SQ_TILE), (unsigned)ctx->n_quant);
        KbProf pf(ctx, 2, s);
        tpe_score_quant_kernel<<<grid, SC_THREADS, 0, s>>>(ctx->d_dims, ctx->d_quant_dims, ctx->d_stage_quant, ctx->chunk_cap,
                                                           n, idx_base, D, sel, ctx->d_tie, sc->d_tie_score, d_cand, d_scores);
        KB_LAUNCH_CHECK(ctx);
    }
    return KB_OK;
}

static int tpe_reset_selection(kb_ctx *ctx, cudaStream_t s) {
    const size_t slots = (size_t)(ctx->n_numeric > 0 ? ctx->n_numeric : 1);
    KB_CUDA(ctx, cudaMemsetAsync(ctx->d_sel, 0, sizeof(KbSelect) * (size_t)ctx->D, s));
    KB_CUDA(ctx, cudaMemsetAsync(ctx->d_tie, 0xFF, sizeof(KbTie) * slots * KB_TIE_SLOTS, s));
    return KB_OK;
}

static int tpe_finish_selection(kb_ctx *ctx, long long idx_base, int64_t *out_index, double *out_value, double *out_score,
                                cudaStream_t s) {
    kb_ctx *sc = ctx;
    const int D = ctx->D;
    KbProf *pf3 = new KbProf(ctx, 3, s);
    if (ctx->n_numeric > 0) {
        tpe_rescore_kernel<<<dim3((unsigned)ctx->n_numeric, RS_BLOCKS), RS_THREADS, 0, s>>>(
            ctx->d_dims, sc->d_numeric_dims, ctx->d_sel, ctx->d_tie, sc->d_tie_score, (KbPartial *)sc->d_partial);
        KB_LAUNCH_CHECK(ctx);
    }
    tpe_pick_kernel<<<(D + 127) / 128, 128, 0, s>>>(ctx->d_dims, D, ctx->d_sel, (const KbPartial *)sc->d_partial, ctx->d_win, idx_base);
    delete pf3;
    KB_LAUNCH_CHECK(ctx);
    int world = 1, rank = 0;
    kb_comm_world(ctx, &world, &rank);
    if (world > 1) {
        if (world > 64) return kb_fail(ctx, KB_ERR_UNSUPPORTED, "world > 64");
        int rc = kb_comm_allgather(ctx, ctx->d_win, ctx->d_win_all, sizeof(KbWinner) * (size_t)D, s);
        if (rc) return rc;
        tpe_merge_kernel<<<(D + 127) / 128, 128, 0, s>>>(ctx->d_win_all, world, D, ctx->d_win);
        KB_LAUNCH_CHECK(ctx);
    }
    KB_CUDA(ctx, cudaMemcpyAsync(sc->h_win, ctx->d_win, sizeof(KbWinner) * (size_t)D, cudaMemcpyDeviceToHost, s));
    std::vector<KbSelect> hsel((size_t)D);
    KB_CUDA(ctx, cudaMemcpyAsync(hsel.data(), ctx->d_sel, sizeof(KbSelect) * (size_t)D, cudaMemcpyDeviceToHost, s));
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    for (int d = 0; d < D; d++) {
        if (hsel[(size_t)d].overflow)
            return kb_fail(ctx, KB_ERR_OVERFLOW, "kb_tpe_score_select: near-tie table of parameter %d overflowed (%u slots)", d,
                           KB_TIE_SLOTS);
        if (out_index) out_index[d] = sc->h_win[d].index;
        if (out_value) out_value[d] = sc->h_win[d].value;
        if (out_score) out_score[d] = sc->h_win[d].score;
    }
    return KB_OK;
}

static int tpe_precheck(kb_ctx *ctx, int64_t n_cand, const void *cand, const char *who) {
    if (!ctx->history_ready) return kb_fail(ctx, KB_ERR_STATE, "%s: call kb_tpe_set_history first", who);
    if (n_cand < 0 || n_cand >= (1ll << 32)) return kb_fail(ctx, KB_ERR_UNSUPPORTED, "%s: n_cand outside [0, 2^32)", who);
    if (n_cand > 0 && !cand) return kb_fail(ctx, KB_ERR_INVALID, "%s: cand is NULL", who);
    return KB_OK;
}

// candidates on the host: chunked, H2D double-buffered on the copy stream, compute on the ctx stream
static int tpe_run_host(kb_ctx *ctx, int64_t n_cand, const double *cand, int64_t index_offset, bool select, double *scores_host) {
    const int D = ctx->D;
    cudaStream_t s = ctx->stream, cs = ctx->copy_stream;
    const int64_t rows = n_cand < KB_CHUNK_ROWS ? (n_cand > 0 ? n_cand : 1) : KB_CHUNK_ROWS;
    int rc = tpe_ensure_scratch(ctx, rows, true, scores_host != nullptr);
    if (rc) return rc;
    if (select && (rc = tpe_reset_selection(ctx, s))) return rc;
    const int64_t cap = ctx->chunk_cap;
    int buf = 0;
    for (int64_t off = 0; off < n_cand; off += cap, buf ^= 1) {
        const int n = (int)(n_cand - off < cap ? n_cand - off : cap);
        KB_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev_done[buf], 0));  // landing buffer free again
        KB_CUDA(ctx, cudaMemcpyAsync(ctx->d_cand_chunk[buf], cand + (size_t)off * D, sizeof(double) * (size_t)n * D,
                                     cudaMemcpyHostToDevice, cs));
        KB_CUDA(ctx, cudaEventRecord(ctx->ev_h2d[buf], cs));
        KB_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_h2d[buf], 0));
        if ((rc = tpe_launch_chunk(ctx, ctx->d_cand_chunk[buf], n, index_offset + off, off, select,
                                   scores_host ? ctx->d_scores_chunk : nullptr, s)))
            return rc;
        if (scores_host) {
            KB_CUDA(ctx, cudaMemcpyAsync(scores_host + (size_t)off * D, ctx->d_scores_chunk, sizeof(double) * (size_t)n * D,
                                         cudaMemcpyDeviceToHost, s));
        }
        KB_CUDA(ctx, cudaEventRecord(ctx->ev_done[buf], s));
    }
    return KB_OK;
}

extern "C" int kb_tpe_scores(kb_ctx *ctx, int64_t n_cand, const double *cand, double *scores) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    int rc = tpe_precheck(ctx, n_cand, cand, "kb_tpe_scores");
    if (rc) return rc;
    if (n_cand == 0) return KB_OK;
    if (!scores) return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_scores: scores is NULL");
    if ((rc = tpe_run_host(ctx, n_cand, cand, 0, false, scores))) return rc;
    KB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return KB_OK;
}

extern "C" int kb_tpe_score_select(kb_ctx *ctx, int64_t n_cand, const double *cand, int64_t index_offset,
                                   int64_t *out_index, double *out_value, double *out_score) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    int rc = tpe_precheck(ctx, n_cand, cand, "kb_tpe_score_select");
    if (rc) return rc;
    if ((rc = tpe_run_host(ctx, n_cand, cand, index_offset, true, nullptr))) return rc;
    return tpe_finish_selection(ctx, index_offset, out_index, out_value, out_score, ctx->stream);
}

extern "C" int kb_tpe_score_select_device(kb_ctx *ctx, int64_t n_cand, const double *cand_device, int64_t index_offset,
                                          int64_t *out_index, double *out_value, double *out_score, void *stream) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    int rc = tpe_precheck(ctx, n_cand, cand_device, "kb_tpe_score_select_device");
    if (rc) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    const int64_t rows = n_cand < KB_CHUNK_ROWS ? (n_cand > 0 ? n_cand : 1) : KB_CHUNK_ROWS;
    if ((rc = tpe_ensure_scratch(ctx, rows, false, false))) return rc;
    if ((rc = tpe_reset_selection(ctx, s))) return rc;
    const int64_t cap = ctx->chunk_cap;
    for (int64_t off = 0; off < n_cand; off += cap) {
        const int n = (int)(n_cand - off < cap ? n_cand - off : cap);
        if ((rc = tpe_launch_chunk(ctx, cand_device + (size_t)off * ctx->D, n, index_offset + off, off, true, nullptr, s))) return rc;
    }
    return tpe_finish_selection(ctx, index_offset, out_index, out_value, out_score, s);
}
