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

#include <algorithm>

#include "kb_internal.cuh"
#include "kb_tpe_sample.cuh"
#include "kb_tpe_rescore.cuh"

#define KB_EPS 1e-12
// kb_tpe_suggest with a communicator: candidate counts from here on are sharded over the ranks (one collective); below
// it every rank that calls works alone, so that a service may keep answering small requests on one GPU
#define KB_SUGGEST_SHARD_MIN 65536

// Where a kernel finds the f64 value of a candidate it has to remember (near ties): the caller's row-major matrix, or
// -- when the candidates were drawn on the device straight into the staged columns and never stored -- the Philox
// stream they came from (the draw is a pure function of seed, global index and parameter).
struct KbRaw {
    const double *cand;
    unsigned long long seed;
};
__device__ __forceinline__ double kb_raw_value(const KbRaw &raw, const KbDim &dm, long long i_local, long long gidx, int D, int d) {
    if (raw.cand) return raw.cand[(size_t)i_local * D + d];
    const Philox ph(raw.seed);
    return tpe_sample_one(dm, ph, (unsigned long long)gidx, d);
}
// Admission margin of the near-tie table, relative to max(1, |score|): several times the largest FP32 error measured
// against the FP64 decision formula over ALL 2^26 x 30 pairs of the headline step (kb_tpe_score_error_device, bench.py
// score_error_max: continuous 9.6e-7 with 128-component partial sums, quantised 4.4e-7).
#define KB_MARGIN_CONT 5e-6f
#define KB_MARGIN_QUANT 1.5e-5f

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

// one working-domain value -> what the scoring kernels read
__device__ __forceinline__ float2 stage_value_cont(const KbDim &dm, double x) {
    const double xn = (x - dm.lo) * (1.0 / (dm.hi - dm.lo));
    const float xh = (float)xn;
    return make_float2(xh, (float)(xn - (double)xh));
}
__device__ __forceinline__ float4 stage_value_quant(const KbDim &dm, double x) {
    const double lo = dm.lo, hi = dm.hi, inv = 1.0 / (hi - lo), q = dm.q;
    double ub = x + q / 2.0; ub = ub > hi ? hi : ub;
    double lb = x - q / 2.0; lb = lb < lo ? lo : lb;
    // bin midpoint (hi, lo) and half width, normalised; w = 1 when the bin is a full quantum
    const double mn = (0.5 * (ub + lb) - lo) * inv, hn = 0.5 * (ub - lb) * inv;
    const float mh = (float)mn, hwf = (float)hn;
    float regular = fabs(hn - 0.5 * q * inv) <= 1e-9 * (q * inv) ? 1.f : 0.f;
    // a full bin centred ON the quantisation grid: .w = grid index + 2 (look-up class of the score kernels)
    const double gpos = (x - (lo + 0.5 * q)) / q, gidx = rint(gpos);
    if (regular != 0.f && fabs(gpos - gidx) <= 1e-9 && gidx >= 0.0 && gidx < 16384.0) regular = (float)(gidx + 2.0);
    return make_float4(mh, (float)(mn - (double)mh), hwf, regular);
}
// categorical selection key: (rank+1 : 13 bits) | (~call-local index : 32 bits) | (choice : 12 bits)
__device__ __forceinline__ unsigned long long cat_key(const KbDim &dm, int c, long long local_index) {
    return ((unsigned long long)(unsigned)(dm.cat_rank[c] + 1) << 44) | ((unsigned long long)(~(unsigned)local_index) << 12) |
           (unsigned long long)c;
}

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
                    const unsigned long long key = cat_key(dm, c, local_base + row0 + r);
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
                const bool quant = dm.q > 0.0;
                for (int r = lane; r < rows; r += 32) {
                    const double x = tile[dd][r];
                    if (quant) stage_quant[(size_t)slot * chunk_cap + row0 + r] = stage_value_quant(dm, x);
                    else stage_cont[(size_t)slot * chunk_cap + row0 + r] = stage_value_cont(dm, x);
                }
            }
        }
        __syncthreads();
    }
}

// The same staging for candidates that do not exist yet: kb_tpe_suggest draws candidate (first + i) of parameter d from
// the Philox stream (tpe_sample_one, the values kb_tpe_sample_candidates would return for the same seed) and writes it
// straight into the staged column -- the n x D f64 candidate matrix is never materialised.  One CTA = 256 consecutive
// candidates of ONE parameter (blockIdx.y), so the kind branch is uniform and the column stores coalesce.
#define SS_THREADS 256
#define SS_ITEMS 8   // candidates per thread: 64 Mi x 50 one-candidate threads would be 13 M CTAs, and launching them costs
                     // more than the draws (measured: 135 ms per headline step, unchanged by a sampler half as expensive)
__global__ void __launch_bounds__(SS_THREADS)
tpe_sample_stage_kernel(const KbDim *__restrict__ dims, int n, long long chunk_cap, unsigned long long seed, long long first,
                        float2 *__restrict__ stage_cont, float4 *__restrict__ stage_quant, const int *__restrict__ slot_of_dim,
                        KbSelect *sel, long long local_base) {
    const int d = blockIdx.y;
    const KbDim &dm = dims[d];
    const int slot = slot_of_dim[d];
    const Philox ph(seed);
    if (dm.kind == KB_CATEGORICAL) {
        unsigned long long best = 0ull;
        for (int it = 0; it < SS_ITEMS; it++) {
            const int i = (blockIdx.x * SS_ITEMS + it) * SS_THREADS + threadIdx.x;
            if (i >= n) break;
            int c = (int)tpe_sample_one(dm, ph, (unsigned long long)(first + i), d);
            c = c < 0 ? 0 : (c >= dm.n_choices ? dm.n_choices - 1 : c);
            const unsigned long long key = cat_key(dm, c, local_base + i);
            best = key > best ? key : best;
        }
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long other = __shfl_xor_sync(0xffffffffu, best, o);
            best = other > best ? other : best;
        }
        if ((threadIdx.x & 31) == 0 && best) atomicMax(&sel[d].cat_best, best);
        return;
    }
    const bool quant = dm.q > 0.0;
    for (int it = 0; it < SS_ITEMS; it++) {
        const int i = (blockIdx.x * SS_ITEMS + it) * SS_THREADS + threadIdx.x;
        if (i >= n) break;
        const double x = tpe_sample_one(dm, ph, (unsigned long long)(first + i), d);
        if (quant) stage_quant[(size_t)slot * chunk_cap + i] = stage_value_quant(dm, x);
        else stage_cont[(size_t)slot * chunk_cap + i] = stage_value_cont(dm, x);
    }
}

// ---------------------------------------------------------------------------------------------------------
// near-tie table insert
// ---------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void tie_insert(KbTie *tab, float *tab_score, unsigned int slots, KbSelect *sl,
                                           unsigned long long key, unsigned long long gidx, float score) {
    unsigned long long h = key * 0x9E3779B97F4A7C15ull;
    unsigned int s = (unsigned int)(h >> 40) & (slots - 1);
    for (unsigned int probe = 0; probe < 4096u && probe < slots; probe++) {
        unsigned long long prev = atomicCAS(&tab[s].key, KB_TIE_EMPTY, key);
        if (prev == KB_TIE_EMPTY || prev == key) {
            atomicMin(&tab[s].minidx, gidx);
            tab_score[s] = score;  // every writer of this key stores the same bits
            return;
        }
        s = (s + 1) & (slots - 1);
    }
    atomicExch(&sl->overflow, 1u);
}

// block-wide: publish the block maximum, derive the admission threshold, insert the near ties
template <int R, int THREADS>
__device__ __forceinline__ void select_publish(const float (&sc)[R], int tile_base, int n, long long idx_base, int d,
                                               int D, const KbDim &dm, KbSelect *sel, KbTie *tie, float *tie_score,
                                               unsigned int slots, float margin, KbRaw cand_raw) {
    const int numeric_slot = dm.numeric_slot;
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
        float thr = g - margin * fmaxf(1.0f, fabsf(g));
        s_thr = b >= thr ? thr : INFINITY;  // block has nothing near the maximum: skip the scan
    }
    __syncthreads();
    const float thr = s_thr;
    if (thr == INFINITY) return;
    KbTie *tab = tie + (size_t)numeric_slot * slots;
    float *tsc = tie_score + (size_t)numeric_slot * slots;
#pragma unroll
    for (int r = 0; r < R; r++) {
        int i = tile_base + r * THREADS + threadIdx.x;
        if (i < n && sc[r] >= thr) {
            unsigned long long key = (unsigned long long)__double_as_longlong(kb_raw_value(cand_raw, dm, i, idx_base + i, D, d));
            tie_insert(tab, tsc, slots, &sel[d], key, (unsigned long long)(idx_base + i), sc[r]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// continuous parameters (q = 0): log-sum-exp over the below and above mixtures
//
// Table entry = two float4, pre-duplicated so that the packed FP32x2 instructions of sm_100 (FADD2 / FFMA2:
// one issue slot for two candidates) can take them as they come out of one LDS.128 each:
//   A = {-mu_hi, -mu_hi, c', c'}   B = {-mu_lo*c', -mu_lo*c', k', k'}
// Per pair of (candidate, component) evaluations: 2 FADD2 + 2 FFMA2 + 2 MUFU.EX2 + 1 FADD2.
// (This duplicated form feeds the pruned kernel; tpe_score_cont_kernel streams the compact table dm.tabc: one float4
// {-mu_hi, c', -mu_lo*c', k'} per component, each value a both-halves operand of the packed instructions.)
// ---------------------------------------------------------------------------------------------------------
#define SC_THREADS 256
#ifndef SC_R
#define SC_R 8
#endif
#define SC_TILE (SC_THREADS * SC_R)
#ifndef SC_CHUNK
#define SC_CHUNK 512   // components per shared-memory stage (2 float4 each: 16 KB)
#endif
#define SC_STAGES 2
#ifndef SC_GROUP
// components per FP32 partial sum before it is folded into the (hi, lo) pair: the fold is 48 FADDs per thread, and at 32
// components it cost 3 % of the kernel.  Measured (2^24 candidates x 20 parameters, worst score error over 2^21 x 20):
// 32 -> 2240 ms / 5.7e-7, 64 -> 2206 / 6.3e-7, 128 -> 2189 / 8.1e-7, 256 -> 2181 / 1.3e-6, 512 -> 2199 / 2.3e-6.
#define SC_GROUP 128
#endif
#ifndef SC_MINB
#define SC_MINB 4      // CTAs per SM tpe_score_cont_kernel is compiled for (64 registers, 48 KB of shared memory each)
#endif
#define SC_MINB_PRUNED 3   // tpe_score_cont_pruned_kernel keeps its sums in registers (80)

__device__ __forceinline__ void table_stream_issue(const KbDim &dm, int chunk, int nchunk0, float4 *dst, uint64_t *bar) {
    const int side = chunk < nchunk0 ? 0 : 1;
    const int c = side == 0 ? chunk : chunk - nchunk0;
    const int cnt = min(SC_CHUNK, dm.Kpad[side] - c * SC_CHUNK);
    const uint32_t bytes = (uint32_t)cnt * 2u * (uint32_t)sizeof(float4);
    mbar_expect_tx(bar, bytes);
    tma_bulk_g2s(dst, dm.tab[side] + (size_t)c * SC_CHUNK * 2, bytes, bar);
}

#ifndef SCC_CHUNK
#define SCC_CHUNK 512   // components per stage of the compact table (one float4 each: 8 KB)
#endif
__device__ __forceinline__ void table_stream_issue_compact(const KbDim &dm, int chunk, int nchunk0, float4 *dst, uint64_t *bar) {
    const int side = chunk < nchunk0 ? 0 : 1;
    const int c = side == 0 ? chunk : chunk - nchunk0;
    const int cnt = min(SCC_CHUNK, dm.Kpad[side] - c * SCC_CHUNK);
    const uint32_t bytes = (uint32_t)cnt * (uint32_t)sizeof(float4);
    mbar_expect_tx(bar, bytes);
    tma_bulk_g2s(dst, dm.tabc[side] + (size_t)c * SCC_CHUNK, bytes, bar);
}

// (hi, lo) += a with the rounding error of the add recovered (TwoSum): keeps ~48 bits without touching FP64
__device__ __forceinline__ void two_sum_acc(float &hi, float &lo, float a) {
    const float s = hi + a;
    const float bb = s - hi;
    lo += (hi - (s - bb)) + (a - bb);
    hi = s;
}

// 2^e for e <= 0 on the FMA pipe (no MUFU): clamp, split e = n + f with the 1.5 * 2^23 rounding constant (n = round(e),
// |f| <= 1/2), degree-5 polynomial for 2^f (2.0e-7 relative in FP32 -- the accuracy of MUFU.EX2 itself), n added into
// the exponent field.  Packed FP32x2: 3 FADD2 + 5 FFMA2 per candidate pair; the clamp and the exponent insert run on
// the ALU pipe.  e = -inf (padding entries) gives 2^-127 * (1 + 1e-7): a denormal, i.e. nothing.
__device__ __forceinline__ float2 kb_ex2_fma2(float2 e) {
    const float2 em = make_float2(fmaxf(e.x, -127.f), fmaxf(e.y, -127.f));
    const float2 xm = __fadd2_rn(em, make_float2(12582912.f, 12582912.f));
    const float2 nf = __fadd2_rn(xm, make_float2(-12582912.f, -12582912.f));
    const float2 f = __ffma2_rn(nf, make_float2(-1.f, -1.f), em);
    float2 p = __ffma2_rn(make_float2(1.3280353741720319e-3f, 1.3280353741720319e-3f), f,
                          make_float2(9.675574488937855e-3f, 9.675574488937855e-3f));
    p = __ffma2_rn(p, f, make_float2(5.550701171159744e-2f, 5.550701171159744e-2f));
    p = __ffma2_rn(p, f, make_float2(2.4022118747234344e-1f, 2.4022118747234344e-1f));
    p = __ffma2_rn(p, f, make_float2(6.931470036506653e-1f, 6.931470036506653e-1f));
    p = __ffma2_rn(p, f, make_float2(1.0000001192092896f, 1.0000001192092896f));
    return make_float2(__int_as_float(__float_as_int(p.x) + (__float_as_int(xm.x) << 23)),
                       __int_as_float(__float_as_int(p.y) + (__float_as_int(xm.y) << 23)));
}

// One group of at most SC_GROUP components applied to the SC_R candidates of a thread.  The MUFU pipe is the wall of
// this loop (16 ex2 / clk / SM against 128 FP32 lanes): a share of the (component, candidate pair) slots take their 2^e
// from kb_ex2_fma2 instead, which moves that share of the exponentials onto the otherwise half-idle FMA pipe.  Every
// (candidate, component) pair is still evaluated.
// The masks give the slots of a block of 16 components x 4 candidate pairs (bit 4 jj + p) for 0, 1, 2, 3 and 4
// sixteenths.  The slots are spread out -- one pair of one component at a time, the pair rotating -- so that the
// 9-instruction polynomial chain sits between the MUFU work of the other three pairs, and only one chain's temporaries
// are live (the kernel fits 64 registers: 4 CTAs per SM).  Measured on B200, 2^24 candidates x 20 parameters x 10 002
// components, 3 steps (profiles/r02_ncu_tpe_score.md): whole components in lockstep at 3 CTAs/SM 0 -> 2255 ms,
// 1 -> 2172, 2 -> 2123, 3 -> 2157, 4 -> 2266; spread slots at 3 CTAs/SM 2 -> 2094, 2.5 -> 2112, 3 -> 2157; spread slots
// at 4 CTAs/SM (the sums in shared memory) 2 -> 2040, 2.5 -> 2067, 3 -> 2144.  2/16 is the default: 1.07 of the
// MUFU.EX2 peak measured by kb_peak_measure.  KB_CONT_SW = 0..4 picks another at run time (measurement only).
#define SC_SW_MASK_0 0x0000000000000000ull
#define SC_SW_MASK_1 0x0008000400020001ull
#ifndef SC_SW_MASK_2
#define SC_SW_MASK_2 0x0804020108040201ull
#endif
#define SC_SW_MASK_3 0x0421042104210421ull
#define SC_SW_MASK_4 0x8421842184218421ull
#ifndef SC_SW_DEFAULT
#define SC_SW_DEFAULT 2
#endif
static_assert(SC_R == 8, "the slot masks are laid out for 4 candidate pairs per thread");

// component jj (of a block of 16) against the four candidate pairs; the table values arrive as both-halves operands
template <unsigned long long SW_MASK>
__device__ __forceinline__ void cont_component(int jj, float2 nmu, float2 cs, float2 ms, float2 ks,
                                               const float2 (&xh)[SC_R / 2], const float2 (&xl)[SC_R / 2],
                                               float2 (&a32)[SC_R / 2]) {
#pragma unroll
    for (int p = 0; p < SC_R / 2; p++) {
        float2 u = __fadd2_rn(xh[p], nmu);
        u = __fadd2_rn(u, xl[p]);
        const float2 t = __ffma2_rn(u, cs, ms);
        const float2 e = __ffma2_rn(make_float2(-t.x, -t.y), t, ks);
        if ((SW_MASK >> (jj * (SC_R / 2) + p)) & 1ull) a32[p] = __fadd2_rn(a32[p], kb_ex2_fma2(e));
        else a32[p] = __fadd2_rn(a32[p], make_float2(kb_ex2(e.x), kb_ex2(e.y)));
    }
}

// the two-float4 (duplicated) table form: the pruned kernel's windows come out of tabn / tabw
template <unsigned long long SW_MASK>
__device__ __forceinline__ void cont_group_dup(const float4 *__restrict__ tab, int len, const float2 (&xh)[SC_R / 2],
                                               const float2 (&xl)[SC_R / 2], float2 (&a32)[SC_R / 2]) {
#pragma unroll 1
    for (int h = 0; h < len; h += 16) {  // len: a multiple of 16 (tables are padded to KB_TABLE_PAD = 32), at most SC_GROUP
#pragma unroll
        for (int jj = 0; jj < 16; jj++) {
            const float4 A = tab[2 * (h + jj)], B = tab[2 * (h + jj) + 1];  // two broadcast LDS.128
            cont_component<SW_MASK>(jj, make_float2(A.x, A.y), make_float2(A.z, A.w), make_float2(B.x, B.y),
                                    make_float2(B.z, B.w), xh, xl, a32);
        }
    }
}

template <unsigned long long SW_MASK>
__device__ __forceinline__ void cont_group(const float4 *__restrict__ tab, int len, const float2 (&xh)[SC_R / 2],
                                           const float2 (&xl)[SC_R / 2], float2 (&a32)[SC_R / 2]) {
#pragma unroll 1
    for (int h = 0; h < len; h += 16) {
#pragma unroll
        for (int jj = 0; jj < 16; jj++) {
            const float4 T = tab[h + jj];  // one broadcast LDS.128: {-mu_hi, c, -mu_lo c, k}, each a both-halves operand
            cont_component<SW_MASK>(jj, make_float2(T.x, T.x), make_float2(T.y, T.y), make_float2(T.z, T.z),
                                    make_float2(T.w, T.w), xh, xl, a32);
        }
    }
}

// Shared memory (dynamic, SC_CONT_SMEM bytes = 48 KB): two table stages of SCC_CHUNK one-float4 entries (16 KB), then
// the per-candidate (hi, lo) sums of the side in progress and those of the finished below side, [2 SC_R][SC_THREADS]
// floats each (16 KB + 16 KB).  The sums are touched once per SC_GROUP components, so they wait in shared memory and
// the loop fits 64 registers: 4 CTAs = 32 warps per SM (3 CTAs at 80 registers: 4 % slower, see cont_component).
#define SC_CONT_SMEM ((size_t)SC_STAGES * SCC_CHUNK * sizeof(float4) + (size_t)4 * SC_R * SC_THREADS * sizeof(float))
template <unsigned long long SW_MASK>
__global__ void __launch_bounds__(SC_THREADS, SC_MINB)
tpe_score_cont_kernel(const KbDim *__restrict__ dims, const int *__restrict__ cont_dims,
                      const float2 *__restrict__ stage, long long chunk_cap, int n, long long idx_base, int D,
                      KbSelect *sel, KbTie *tie, float *tie_score, unsigned int slots,
                      KbRaw cand_raw, double *__restrict__ scores_out) {
    extern __shared__ __align__(128) float4 s_dyn[];
    float4 (*s_tab)[SCC_CHUNK] = reinterpret_cast<float4 (*)[SCC_CHUNK]>(s_dyn);
    float *s_sum = reinterpret_cast<float *>(s_dyn + SC_STAGES * SCC_CHUNK) + threadIdx.x;  // row 2r: hi, row 2r+1: lo
    float *s_below = s_sum + 2 * SC_R * SC_THREADS;
    __shared__ __align__(8) uint64_t s_bar[SC_STAGES];
    const int slot = blockIdx.y;
    const int d = cont_dims[slot];
    const KbDim &dm = dims[d];
    const int tile_base = blockIdx.x * SC_TILE;
    const int nchunk0 = (dm.Kpad[0] + SCC_CHUNK - 1) / SCC_CHUNK, nchunk1 = (dm.Kpad[1] + SCC_CHUNK - 1) / SCC_CHUNK;
    const int nchunks = nchunk0 + nchunk1;

    if (threadIdx.x == 0) {
        for (int s = 0; s < SC_STAGES; s++) mbar_init(&s_bar[s], 1);
        mbar_fence_init();
    }
    __syncthreads();
    if (threadIdx.x == 0)
        for (int c = 0; c < SC_STAGES && c < nchunks; c++) table_stream_issue_compact(dm, c, nchunk0, s_tab[c], &s_bar[c]);

    float2 xh[SC_R / 2], xl[SC_R / 2];  // candidate pairs (r, r+1)
#pragma unroll
    for (int r = 0; r < SC_R; r++) {
        int i = tile_base + r * SC_THREADS + threadIdx.x;
        float2 v = i < n ? stage[(size_t)slot * chunk_cap + i] : make_float2(0.5f, 0.f);
        if (r & 1) { xh[r / 2].y = v.x; xl[r / 2].y = v.y; } else { xh[r / 2].x = v.x; xl[r / 2].x = v.y; }
    }
#pragma unroll
    for (int k = 0; k < 2 * SC_R; k++) s_sum[k * SC_THREADS] = 0.f;  // each thread reads and writes its own column only

    for (int c = 0; c < nchunks; c++) {
        const int st = c % SC_STAGES;
        const int side = c < nchunk0 ? 0 : 1;
        const int cc = side == 0 ? c : c - nchunk0;
        const int cnt = min(SCC_CHUNK, dm.Kpad[side] - cc * SCC_CHUNK);
        mbar_wait(&s_bar[st], (uint32_t)((c / SC_STAGES) & 1));
        const float4 *tab = s_tab[st];
        for (int j0 = 0; j0 < cnt; j0 += SC_GROUP) {
            float2 a32[SC_R / 2];
#pragma unroll
            for (int p = 0; p < SC_R / 2; p++) a32[p] = make_float2(0.f, 0.f);
            cont_group<SW_MASK>(tab + j0, min(SC_GROUP, cnt - j0), xh, xl, a32);
#pragma unroll
            for (int p = 0; p < SC_R / 2; p++) {  // fold the FP32 partial sums into the (hi, lo) pairs
                float *q = s_sum + 4 * p * SC_THREADS;
                float h0 = q[0], l0 = q[SC_THREADS], h1 = q[2 * SC_THREADS], l1 = q[3 * SC_THREADS];
                two_sum_acc(h0, l0, a32[p].x);
                two_sum_acc(h1, l1, a32[p].y);
                q[0] = h0; q[SC_THREADS] = l0; q[2 * SC_THREADS] = h1; q[3 * SC_THREADS] = l1;
            }
        }
        __syncthreads();  // every thread is done with stage st
        if (threadIdx.x == 0 && c + SC_STAGES < nchunks)
            table_stream_issue_compact(dm, c + SC_STAGES, nchunk0, s_tab[st], &s_bar[st]);
        if (c == nchunk0 - 1) {  // the below side is complete: park its sums, start the above side from zero
#pragma unroll
            for (int k = 0; k < 2 * SC_R; k++) { s_below[k * SC_THREADS] = s_sum[k * SC_THREADS]; s_sum[k * SC_THREADS] = 0.f; }
        }
    }
    float sc[SC_R];
#pragma unroll
    for (int r = 0; r < SC_R; r++) {
        const double below = log((double)s_below[2 * r * SC_THREADS] + (double)s_below[(2 * r + 1) * SC_THREADS]) + dm.L[0];
        const double above = log((double)s_sum[2 * r * SC_THREADS] + (double)s_sum[(2 * r + 1) * SC_THREADS]) + dm.L[1];
        const double s64 = below - above;
        sc[r] = (float)s64;
        int i = tile_base + r * SC_THREADS + threadIdx.x;
        if (scores_out && i < n) scores_out[(size_t)i * D + d] = s64;
    }
    if (sel) select_publish<SC_R, SC_THREADS>(sc, tile_base, n, idx_base, d, D, dm, sel, tie, tie_score, slots,
                                                  KB_MARGIN_CONT, cand_raw);
}

// select_publish for candidates in bucket order: the scores belong to the chunk-local indices orig[] (= perm[] of the
// sorted position, -1 past the end); block maximum and near-tie admission as in select_publish.
__device__ __forceinline__ void select_publish_perm(const float (&sc)[SC_R], const int (&orig)[SC_R], long long idx_base,
                                                    int d, int D, const KbDim &dm, KbSelect *sel, KbTie *tie,
                                                    float *tie_score, unsigned int slots, KbRaw cand_raw) {
    __shared__ float s_wmax[SC_THREADS / 32];
    __shared__ float s_thr;
    float best = -INFINITY;
#pragma unroll
    for (int r = 0; r < SC_R; r++)
        if (orig[r] >= 0) best = fmaxf(best, sc[r]);
    for (int o = 16; o > 0; o >>= 1) best = fmaxf(best, __shfl_xor_sync(0xffffffffu, best, o));
    if ((threadIdx.x & 31) == 0) s_wmax[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x == 0) {
        float b = s_wmax[0];
        for (int w = 1; w < SC_THREADS / 32; w++) b = fmaxf(b, s_wmax[w]);
        const unsigned int mine = kb_f32_orderable(b);
        const unsigned int old = atomicMax(&sel[d].gmax_bits, mine);
        const float g = kb_f32_from_orderable(old > mine ? old : mine);
        const float thr = g - KB_MARGIN_CONT * fmaxf(1.0f, fabsf(g));
        s_thr = b >= thr ? thr : INFINITY;
    }
    __syncthreads();
    const float thr = s_thr;
    if (thr == INFINITY) return;
    KbTie *ttab = tie + (size_t)dm.numeric_slot * slots;
    float *tsc = tie_score + (size_t)dm.numeric_slot * slots;
#pragma unroll
    for (int r = 0; r < SC_R; r++) {
        if (orig[r] >= 0 && sc[r] >= thr) {
            const unsigned long long key = (unsigned long long)__double_as_longlong(kb_raw_value(cand_raw, dm, orig[r], idx_base + orig[r], D, d));
            tie_insert(ttab, tsc, slots, &sel[d], key, (unsigned long long)(idx_base + orig[r]), sc[r]);
        }
    }
}

// ---------------------------------------------------------------------------------------------------------
// KB_OPT_PRUNE_FAR, continuous parameters: candidates are bucketed by value (256 buckets over the normalised range, a
// counting sort per chunk that keeps the chunk-local index), so that the 2048 candidates of a thread block span about
// one bucket; the block then streams only the window of the mu-sorted floor-sigma table within 7.75 sigma of its
// candidates, plus the (few) wide components.  Same arithmetic as tpe_score_cont_kernel on what it does evaluate.
// ---------------------------------------------------------------------------------------------------------
#define KB_PRUNE_BUCKETS 256
#define BK_THREADS 256
#define BK_ITEMS 8

__device__ __forceinline__ int prune_bucket(float x) {
    const int b = (int)(x * (float)KB_PRUNE_BUCKETS);
    return b < 0 ? 0 : (b >= KB_PRUNE_BUCKETS ? KB_PRUNE_BUCKETS - 1 : b);
}

__global__ void __launch_bounds__(BK_THREADS)
tpe_bucket_count_kernel(const float2 *__restrict__ stage, long long chunk_cap, int n, int *__restrict__ hist) {
    __shared__ int s_h[KB_PRUNE_BUCKETS];
    const int slot = blockIdx.y;
    s_h[threadIdx.x] = 0;
    __syncthreads();
#pragma unroll
    for (int it = 0; it < BK_ITEMS; it++) {
        const int i = (blockIdx.x * BK_ITEMS + it) * BK_THREADS + threadIdx.x;
        if (i < n) atomicAdd(&s_h[prune_bucket(stage[(size_t)slot * chunk_cap + i].x)], 1);
    }
    __syncthreads();
    if (s_h[threadIdx.x]) atomicAdd(&hist[slot * KB_PRUNE_BUCKETS + threadIdx.x], s_h[threadIdx.x]);
}

__global__ void __launch_bounds__(KB_PRUNE_BUCKETS)
tpe_bucket_scan_kernel(int *__restrict__ hist) {  // counts -> exclusive start offsets, in place; one block per parameter
    __shared__ int s[KB_PRUNE_BUCKETS];
    const int v = hist[blockIdx.x * KB_PRUNE_BUCKETS + threadIdx.x];
    s[threadIdx.x] = v;
    __syncthreads();
    for (int off = 1; off < KB_PRUNE_BUCKETS; off <<= 1) {
        const int a = threadIdx.x >= off ? s[threadIdx.x - off] : 0;
        __syncthreads();
        s[threadIdx.x] += a;
        __syncthreads();
    }
    hist[blockIdx.x * KB_PRUNE_BUCKETS + threadIdx.x] = s[threadIdx.x] - v;
}

__global__ void __launch_bounds__(BK_THREADS)
tpe_bucket_scatter_kernel(const float2 *__restrict__ stage, long long chunk_cap, int n, int *__restrict__ offs,
                          float2 *__restrict__ sorted, int32_t *__restrict__ perm) {
    __shared__ int s_cnt[KB_PRUNE_BUCKETS], s_base[KB_PRUNE_BUCKETS];
    const int slot = blockIdx.y;
    s_cnt[threadIdx.x] = 0;
    __syncthreads();
    float2 v[BK_ITEMS];
    int b[BK_ITEMS], rk[BK_ITEMS];
#pragma unroll
    for (int it = 0; it < BK_ITEMS; it++) {
        const int i = (blockIdx.x * BK_ITEMS + it) * BK_THREADS + threadIdx.x;
        b[it] = -1;
        if (i < n) {
            v[it] = stage[(size_t)slot * chunk_cap + i];
            b[it] = prune_bucket(v[it].x);
            rk[it] = atomicAdd(&s_cnt[b[it]], 1);
        }
    }
    __syncthreads();
    s_base[threadIdx.x] = s_cnt[threadIdx.x] ? atomicAdd(&offs[slot * KB_PRUNE_BUCKETS + threadIdx.x], s_cnt[threadIdx.x]) : 0;
    __syncthreads();
#pragma unroll
    for (int it = 0; it < BK_ITEMS; it++) {
        if (b[it] < 0) continue;
        const int i = (blockIdx.x * BK_ITEMS + it) * BK_THREADS + threadIdx.x;
        const size_t pos = (size_t)slot * chunk_cap + s_base[b[it]] + rk[it];
        sorted[pos] = v[it];
        perm[pos] = i;
    }
}

struct PruneSeg { const float4 *p; int cnt; int c0; int nch; };  // table segment: entries (multiple of 32), first chunk, chunks

template <unsigned long long SW_MASK>
__global__ void __launch_bounds__(SC_THREADS, SC_MINB_PRUNED)
tpe_score_cont_pruned_kernel(const KbDim *__restrict__ dims, const int *__restrict__ cont_dims,
                             const float2 *__restrict__ sorted, const int32_t *__restrict__ perm, long long chunk_cap, int n,
                             long long idx_base, int D, KbSelect *sel, KbTie *tie, float *tie_score, unsigned int slots,
                             KbRaw cand_raw, double *__restrict__ scores_out) {
    __shared__ __align__(128) float4 s_tab[SC_STAGES][SC_CHUNK * 2];
    __shared__ __align__(8) uint64_t s_bar[SC_STAGES];
    __shared__ PruneSeg s_seg[4];
    __shared__ float s_lo[SC_THREADS / 32], s_hi[SC_THREADS / 32];
    const int slot = blockIdx.y;
    const int d = cont_dims[slot];
    const KbDim &dm = dims[d];
    const int tile_base = blockIdx.x * SC_TILE;
    if (threadIdx.x == 0) {
        for (int s = 0; s < SC_STAGES; s++) mbar_init(&s_bar[s], 1);
        mbar_fence_init();
    }
    float2 xh[SC_R / 2], xl[SC_R / 2];
    float lo = INFINITY, hi = -INFINITY;
#pragma unroll
    for (int r = 0; r < SC_R; r++) {
        const int i = tile_base + r * SC_THREADS + threadIdx.x;
        float2 v = make_float2(0.5f, 0.f);
        if (i < n) { v = sorted[(size_t)slot * chunk_cap + i]; lo = fminf(lo, v.x); hi = fmaxf(hi, v.x); }
        if (r & 1) { xh[r / 2].y = v.x; xl[r / 2].y = v.y; } else { xh[r / 2].x = v.x; xl[r / 2].x = v.y; }
    }
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
    __syncthreads();
    if (threadIdx.x < 2) {  // one thread per side: window of the narrow table by binary search on mu (= -A.x)
        const int side = threadIdx.x;
        float bl = s_lo[0], bh = s_hi[0];
        for (int w = 1; w < SC_THREADS / 32; w++) { bl = fminf(bl, s_lo[w]); bh = fmaxf(bh, s_hi[w]); }
        const float W = dm.prune_w[side] + 2e-6f;  // + the FP32 split of candidate and mu
        const float4 *tn = dm.tabn[side];
        const int Kn = dm.Kn[side];
        int a = 0, b = Kn;  // first j with mu_j >= bl - W
        while (a < b) { const int m = (a + b) >> 1; if (-tn[2 * m].x < bl - W) a = m + 1; else b = m; }
        const int j_lo = a;
        b = Kn;             // first j with mu_j > bh + W
        while (a < b) { const int m = (a + b) >> 1; if (-tn[2 * m].x <= bh + W) a = m + 1; else b = m; }
        const int start = j_lo & ~(KB_TABLE_PAD - 1);
        const int end = (a + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD;  // <= padded Kn
        PruneSeg sn, sw;
        sn.p = tn + (size_t)2 * start; sn.cnt = bl <= bh ? end - start : 0;
        sw.p = dm.tabw[side]; sw.cnt = (dm.Kw[side] + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD;
        sn.nch = (sn.cnt + SC_CHUNK - 1) / SC_CHUNK; sw.nch = (sw.cnt + SC_CHUNK - 1) / SC_CHUNK;
        sn.c0 = 0; sw.c0 = 0;
        s_seg[2 * side] = sn; s_seg[2 * side + 1] = sw;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        int c0 = 0;
        for (int s = 0; s < 4; s++) { s_seg[s].c0 = c0; c0 += s_seg[s].nch; }
    }
    __syncthreads();
    const int nchunks = s_seg[3].c0 + s_seg[3].nch;
    const int side1_first = s_seg[2].c0;
    auto issue = [&](int c, int st) {  // thread 0 only
        int s = 0;
        while (s < 3 && c >= s_seg[s].c0 + s_seg[s].nch) s++;
        const int cc = c - s_seg[s].c0;
        const int cnt = min(SC_CHUNK, s_seg[s].cnt - cc * SC_CHUNK);
        const uint32_t bytes = (uint32_t)cnt * 2u * (uint32_t)sizeof(float4);
        mbar_expect_tx(&s_bar[st], bytes);
        tma_bulk_g2s(s_tab[st], s_seg[s].p + (size_t)cc * SC_CHUNK * 2, bytes, &s_bar[st]);
    };
    if (threadIdx.x == 0)
        for (int c = 0; c < SC_STAGES && c < nchunks; c++) issue(c, c);

    float Shi[SC_R], Slo[SC_R];
    double logS0[SC_R];
#pragma unroll
    for (int r = 0; r < SC_R; r++) { Shi[r] = 0.f; Slo[r] = 0.f; logS0[r] = 0.0; }
    for (int c = 0; c < nchunks; c++) {
        const int st = c % SC_STAGES;
        int s = 0;
        while (s < 3 && c >= s_seg[s].c0 + s_seg[s].nch) s++;
        const int cnt = min(SC_CHUNK, s_seg[s].cnt - (c - s_seg[s].c0) * SC_CHUNK);
        mbar_wait(&s_bar[st], (uint32_t)((c / SC_STAGES) & 1));
        const float4 *tab = s_tab[st];
        for (int j0 = 0; j0 < cnt; j0 += SC_GROUP) {
            float2 a32[SC_R / 2];
#pragma unroll
            for (int p = 0; p < SC_R / 2; p++) a32[p] = make_float2(0.f, 0.f);
            cont_group_dup<SW_MASK>(tab + 2 * j0, min(SC_GROUP, cnt - j0), xh, xl, a32);
#pragma unroll
            for (int p = 0; p < SC_R / 2; p++) {
                two_sum_acc(Shi[2 * p], Slo[2 * p], a32[p].x);
                two_sum_acc(Shi[2 * p + 1], Slo[2 * p + 1], a32[p].y);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0 && c + SC_STAGES < nchunks) issue(c + SC_STAGES, st);
        if (c == side1_first - 1) {
#pragma unroll
            for (int r = 0; r < SC_R; r++) {
                logS0[r] = log((double)Shi[r] + (double)Slo[r]) + dm.L[0];
                Shi[r] = 0.f; Slo[r] = 0.f;
            }
        }
    }
    float sc[SC_R];
    int orig[SC_R];
#pragma unroll
    for (int r = 0; r < SC_R; r++) {
        const double s64 = logS0[r] - (log((double)Shi[r] + (double)Slo[r]) + dm.L[1]);
        sc[r] = (float)s64;
        const int i = tile_base + r * SC_THREADS + threadIdx.x;
        orig[r] = i < n ? perm[(size_t)slot * chunk_cap + i] : -1;
        if (orig[r] >= 0 && scores_out) scores_out[(size_t)orig[r] * D + d] = s64;
    }
    if (sel) select_publish_perm(sc, orig, idx_base, d, D, dm, sel, tie, tie_score, slots, cand_raw);
}

// ---------------------------------------------------------------------------------------------------------
// KB_OPT_LOCAL_ORIGIN, continuous parameters: the same sums as tpe_score_cont_kernel with fewer packed instructions per
// (candidate pair, component): 3 instead of 5 on the FMA pipe for nearly every component.
//
// The (hi, lo) pair of a candidate exists because one FP32 number resolves only 2^-24 of the parameter's range and
// x - mu is needed to about 2^-31 of it.  Here the candidates arrive in bucket order (tpe_bucket_* above), so the 2048
// of a thread block lie within a bucket or two, and the block measures everything from its own origin B (the middle of
// its candidates) in units of the narrow components' sigma: xs = (x - B) c0, |xs| < 0.3, one FP32 number good to 2^-26.
// The table cannot be prepared for every B, so a stage arrives as {mu (FP64), c', k'} (dm.tabs, TMA) and the block
// rewrites it in place -- a few FP64 operations per component per block, against 2048 evaluations of it:
//   narrow component (sigma on the magic-clip floor, c' = c0: all but a handful):  {-(mu-B) c0, k'}
//        t = xs + (-M)                e = -t t + k'       FADD2, FFMA2 (+ the FADD2 that accumulates 2^e)
//   wide component (the prior, sparse regions):                                    {-(mu-B) c0, c'/c0, k'}
//        t = (xs + (-M)) (c'/c0)      e = -t t + k'       FADD2, FMUL2, FFMA2
// M is rounded to FP32 where it is large, i.e. for components far from every candidate of the block, whose terms are
// negligible; near ones (|t| < 6) have |M| < 7 and an absolute error below 2^-25 * 8 in t.  With the FMA pipe
// relieved, more of the exponentials fit on it (SW_N / SW_W slots of 64).
// A block whose candidates span more than KB_LOCAL_SPAN (thin tails) keeps B = 0 and the (hi, lo) form of
// tpe_score_cont_kernel: still every pair, just not faster.
// ---------------------------------------------------------------------------------------------------------
#define KB_LOCAL_SPAN 0.0078125f   // 2 buckets of 1/256: what a block in a well-filled region spans at most
#define KB_LOCAL_MIN_ROWS 32768    // KB_OPT_LOCAL_ORIGIN = 1: chunks below this keep their order (sorting would not pay)
// ... and so do histories below this: bucketing and the per-block table rewrite cost about what 140 components cost a
// candidate, the shorter loop saves 12 % (measured: 1 002 components -2.5 %, 10 002 components +11 %)
#define KB_LOCAL_MIN_TRIALS 2048
#ifndef SC_SW_LOCAL_DEFAULT
#define SC_SW_LOCAL_DEFAULT 16     // of 64 slots, narrow components (measured: 12 -> 1882 ms, 16 -> 1812, 18 -> 1828, 20 -> 1884, 24 -> 2043)
#endif
#define SC_SW_LOCAL_WIDE 12        // of 64 slots, wide components

template <unsigned long long SW_MASK>
__device__ __forceinline__ void cont_group_own_narrow(const float4 *__restrict__ tab, int len, const float2 (&xs)[SC_R / 2],
                                                      float2 (&a32)[SC_R / 2]) {
#pragma unroll 1
    for (int h = 0; h < len; h += 16) {
#pragma unroll
        for (int jj = 0; jj < 16; jj++) {
            const float2 T = *reinterpret_cast<const float2 *>(tab + h + jj);  // {-M, k'}
            const float2 nm = make_float2(T.x, T.x), ks = make_float2(T.y, T.y);
#pragma unroll
            for (int p = 0; p < SC_R / 2; p++) {
                const float2 tt = __fadd2_rn(xs[p], nm);
                const float2 e = __ffma2_rn(make_float2(-tt.x, -tt.y), tt, ks);
                if ((SW_MASK >> (jj * (SC_R / 2) + p)) & 1ull) a32[p] = __fadd2_rn(a32[p], kb_ex2_fma2(e));
                else a32[p] = __fadd2_rn(a32[p], make_float2(kb_ex2(e.x), kb_ex2(e.y)));
            }
        }
    }
}

template <unsigned long long SW_MASK>
__device__ __forceinline__ void cont_group_own_wide(const float4 *__restrict__ tab, int len, const float2 (&xs)[SC_R / 2],
                                                    float2 (&a32)[SC_R / 2]) {
#pragma unroll 1
    for (int h = 0; h < len; h += 16) {
#pragma unroll
        for (int jj = 0; jj < 16; jj++) {
            const float4 T = tab[h + jj];  // {-M, c'/c0, k', -}
            const float2 nm = make_float2(T.x, T.x), rho = make_float2(T.y, T.y), ks = make_float2(T.z, T.z);
#pragma unroll
            for (int p = 0; p < SC_R / 2; p++) {
                const float2 tt = __fmul2_rn(__fadd2_rn(xs[p], nm), rho);
                const float2 e = __ffma2_rn(make_float2(-tt.x, -tt.y), tt, ks);
                if ((SW_MASK >> (jj * (SC_R / 2) + p)) & 1ull) a32[p] = __fadd2_rn(a32[p], kb_ex2_fma2(e));
                else a32[p] = __fadd2_rn(a32[p], make_float2(kb_ex2(e.x), kb_ex2(e.y)));
            }
        }
    }
}

// n of the 64 (component, pair) slots of a 16-component block on the FMA pipe: component floor(16 k / n), pair k mod 4
__host__ __device__ constexpr unsigned long long kb_sw_slots(int n) {
    unsigned long long m = 0;
    for (int k = 0; k < n; k++) m |= 1ull << (((k * 16) / n) * (SC_R / 2) + (k % (SC_R / 2)));
    return m;
}
static_assert(kb_sw_slots(8) == SC_SW_MASK_2 && kb_sw_slots(16) == SC_SW_MASK_4 && kb_sw_slots(4) == SC_SW_MASK_1, "slot layout");

template <unsigned long long SW_N>
__global__ void __launch_bounds__(SC_THREADS, SC_MINB)
tpe_score_cont_local_kernel(const KbDim *__restrict__ dims, const int *__restrict__ cont_dims,
                            const float2 *__restrict__ sorted, const int32_t *__restrict__ perm, long long chunk_cap, int n,
                            long long idx_base, int D, KbSelect *sel, KbTie *tie, float *tie_score, unsigned int slots,
                            KbRaw cand_raw, double *__restrict__ scores_out, unsigned long long *__restrict__ stats) {
    extern __shared__ __align__(128) float4 s_dyn[];
    float4 (*s_tab)[SCC_CHUNK] = reinterpret_cast<float4 (*)[SCC_CHUNK]>(s_dyn);
    float *s_sum = reinterpret_cast<float *>(s_dyn + SC_STAGES * SCC_CHUNK) + threadIdx.x;
    float *s_below = s_sum + 2 * SC_R * SC_THREADS;
    __shared__ __align__(8) uint64_t s_bar[SC_STAGES];
    __shared__ float s_lo[SC_THREADS / 32], s_hi[SC_THREADS / 32];
    __shared__ PruneSeg s_seg[4];  // below narrow, below wide, above narrow, above wide
    const int slot = blockIdx.y;
    const int d = cont_dims[slot];
    const KbDim &dm = dims[d];
    const int tile_base = blockIdx.x * SC_TILE;
    const float2 *mine = sorted + (size_t)slot * chunk_cap + tile_base + threadIdx.x;
    if (threadIdx.x == 0) {
        for (int s = 0; s < SC_STAGES; s++) mbar_init(&s_bar[s], 1);
        mbar_fence_init();
        int c0 = 0;
        for (int side = 0; side < 2; side++) {
            const int Knp = (dm.Kn[side] + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD;
            const int Kwp = (dm.Kw[side] + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD;
            PruneSeg a, b;
            a.p = dm.tabs[side]; a.cnt = Knp; a.nch = (Knp + SCC_CHUNK - 1) / SCC_CHUNK; a.c0 = c0; c0 += a.nch;
            b.p = dm.tabs[side] + Knp; b.cnt = Kwp; b.nch = (Kwp + SCC_CHUNK - 1) / SCC_CHUNK; b.c0 = c0; c0 += b.nch;
            s_seg[2 * side] = a; s_seg[2 * side + 1] = b;
        }
    }
    // the range of the block's candidates
    float lo = INFINITY, hi = -INFINITY;
#pragma unroll
    for (int r = 0; r < SC_R; r++)
        if (tile_base + r * SC_THREADS + threadIdx.x < n) { const float v = mine[r * SC_THREADS].x; lo = fminf(lo, v); hi = fmaxf(hi, v); }
    for (int o = 16; o > 0; o >>= 1) {
        lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, o));
        hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, o));
    }
    if ((threadIdx.x & 31) == 0) { s_lo[threadIdx.x >> 5] = lo; s_hi[threadIdx.x >> 5] = hi; }
    __syncthreads();
    const int nchunks = s_seg[3].c0 + s_seg[3].nch;
    const int side1_first = s_seg[2].c0;
    auto seg_of = [&](int c) { int s = 0; while (s < 3 && c >= s_seg[s].c0 + s_seg[s].nch) s++; return s; };
    auto issue = [&](int c, int st) {  // thread 0 only
        const int s = seg_of(c);
        const int cc = c - s_seg[s].c0;
        const uint32_t bytes = (uint32_t)min(SCC_CHUNK, s_seg[s].cnt - cc * SCC_CHUNK) * (uint32_t)sizeof(float4);
        mbar_expect_tx(&s_bar[st], bytes);
        tma_bulk_g2s(s_tab[st], s_seg[s].p + (size_t)cc * SCC_CHUNK, bytes, &s_bar[st]);
    };
    if (threadIdx.x == 0)
        for (int c = 0; c < SC_STAGES && c < nchunks; c++) issue(c, c);
    float bl = s_lo[0], bh = s_hi[0];
#pragma unroll
    for (int w = 1; w < SC_THREADS / 32; w++) { bl = fminf(bl, s_lo[w]); bh = fmaxf(bh, s_hi[w]); }
    const bool own = bh - bl <= KB_LOCAL_SPAN;   // block-uniform
    const double B = own ? (double)(0.5f * (bl + bh)) : 0.0;
    if (stats && threadIdx.x == 0) atomicAdd(&stats[own ? 0 : 1], 1ull);

    // own origin: xh = xs = (x - B) c0 of the side in progress (one rounding, from FP64), xl unused;
    // otherwise (hi, lo) of x as staged.  Rows past the end sit on the origin.
    float2 xh[SC_R / 2], xl[SC_R / 2];
    auto load_candidates = [&](int side) {
        const double c0 = (double)dm.c0[side];
#pragma unroll
        for (int r = 0; r < SC_R; r++) {
            float2 v = make_float2(0.5f, 0.f);
            const bool valid = tile_base + r * SC_THREADS + threadIdx.x < n;
            if (valid) v = mine[r * SC_THREADS];
            if (own) v = make_float2(valid ? (float)((((double)v.x - B) + (double)v.y) * c0) : 0.f, 0.f);
            if (r & 1) { xh[r / 2].y = v.x; xl[r / 2].y = v.y; } else { xh[r / 2].x = v.x; xl[r / 2].x = v.y; }
        }
    };
    load_candidates(0);
#pragma unroll
    for (int k = 0; k < 2 * SC_R; k++) s_sum[k * SC_THREADS] = 0.f;

    for (int c = 0; c < nchunks; c++) {
        const int st = c % SC_STAGES;
        const int s = seg_of(c);
        const int cnt = min(SCC_CHUNK, s_seg[s].cnt - (c - s_seg[s].c0) * SCC_CHUNK);
        const bool seg_narrow = (s & 1) == 0;
        mbar_wait(&s_bar[st], (uint32_t)((c / SC_STAGES) & 1));
        {   // the stage, rewritten about B
            const double c0 = (double)dm.c0[s >> 1];
            for (int e = threadIdx.x; e < cnt; e += SC_THREADS) {
                const float4 S = s_tab[st][e];
                const double m = __hiloint2double(__float_as_int(S.y), __float_as_int(S.x)) - B;
                float4 T;
                if (!own) { const float mh = (float)m; T = make_float4(-mh, S.z, (float)(-(m - (double)mh) * (double)S.z), S.w); }
                else if (seg_narrow) T = make_float4((float)(-m * (double)S.z), S.w, 0.f, 0.f);
                else T = make_float4((float)(-m * c0), (float)((double)S.z / c0), S.w, 0.f);
                s_tab[st][e] = T;
            }
        }
        __syncthreads();
        const float4 *tab = s_tab[st];
        for (int j0 = 0; j0 < cnt; j0 += SC_GROUP) {
            float2 a32[SC_R / 2];
#pragma unroll
            for (int p = 0; p < SC_R / 2; p++) a32[p] = make_float2(0.f, 0.f);
            const int len = min(SC_GROUP, cnt - j0);
            if (!own) cont_group<SC_SW_MASK_2>(tab + j0, len, xh, xl, a32);
            else if (seg_narrow) cont_group_own_narrow<SW_N>(tab + j0, len, xh, a32);
            else cont_group_own_wide<kb_sw_slots(SC_SW_LOCAL_WIDE)>(tab + j0, len, xh, a32);
#pragma unroll
            for (int p = 0; p < SC_R / 2; p++) {
                float *q = s_sum + 4 * p * SC_THREADS;
                float h0 = q[0], l0 = q[SC_THREADS], h1 = q[2 * SC_THREADS], l1 = q[3 * SC_THREADS];
                two_sum_acc(h0, l0, a32[p].x);
                two_sum_acc(h1, l1, a32[p].y);
                q[0] = h0; q[SC_THREADS] = l0; q[2 * SC_THREADS] = h1; q[3 * SC_THREADS] = l1;
            }
        }
        __syncthreads();  // every thread is done with stage st
        if (threadIdx.x == 0 && c + SC_STAGES < nchunks) {
            // the stage was written through the generic proxy; the bulk copy writes it through the async proxy
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            issue(c + SC_STAGES, st);
        }
        if (c == side1_first - 1) {  // the below side is complete: park its sums, rescale the candidates for the above side
#pragma unroll
            for (int k = 0; k < 2 * SC_R; k++) { s_below[k * SC_THREADS] = s_sum[k * SC_THREADS]; s_sum[k * SC_THREADS] = 0.f; }
            if (own) load_candidates(1);
        }
    }
    float sc[SC_R];
    int orig[SC_R];
#pragma unroll
    for (int r = 0; r < SC_R; r++) {
        const double below = log((double)s_below[2 * r * SC_THREADS] + (double)s_below[(2 * r + 1) * SC_THREADS]) + dm.L[0];
        const double above = log((double)s_sum[2 * r * SC_THREADS] + (double)s_sum[(2 * r + 1) * SC_THREADS]) + dm.L[1];
        const double s64 = below - above;
        sc[r] = (float)s64;
        const int i = tile_base + r * SC_THREADS + threadIdx.x;
        orig[r] = i < n ? perm[(size_t)slot * chunk_cap + i] : -1;
        if (orig[r] >= 0 && scores_out) scores_out[(size_t)orig[r] * D + d] = s64;
    }
    if (sel) select_publish_perm(sc, orig, idx_base, d, D, dm, sel, tie, tie_score, slots, cand_raw);
}

// ---------------------------------------------------------------------------------------------------------
// quantised parameters (q > 0): p(x) = sum_j w_j (Phi(ub) - Phi(lb)) per side.
//
// Table entry (same duplicated layout): A = {-mu_hi, -mu_hi, r', r'}  B = {-mu_lo*r', -mu_lo*r', +-w, +-w},
// r' = sqrt(log2 e)/(sqrt2 sigma_n).  Three warp-uniform component classes:
//   w < 0  sigma sits on the magic-clip floor (nearly every component of a dense history): the bin integral is one
//          function of y = |x-mu| r', fitted at build time: 2^(-(y-h0)^2 + R12(tau)); packed FP32x2, 1 MUFU each
//   narrow bin (b-a <= 0.2): Taylor/Hermite series about the bin midpoint (erfc differences would cancel)
//   else   erfc(a) - erfc(b) with a hand-fitted erfcx (1 MUFU.RCP + 1 MUFU.EX2 each), mirrored onto y >= 0
// ---------------------------------------------------------------------------------------------------------
#ifndef SQ_R
#define SQ_R 8
#endif
#ifndef SQ_MINB
#define SQ_MINB 2
#endif
#define SQ_TILE (SC_THREADS * SQ_R)
#define SQ_GROUP 32   // components per FP32 partial sum of tpe_score_quant_kernel; its run cutting works on one warp-wide group
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

__global__ void __launch_bounds__(SC_THREADS, SQ_MINB)
tpe_score_quant_kernel(const KbDim *__restrict__ dims, const int *__restrict__ quant_dims,
                       const float4 *__restrict__ stage, long long chunk_cap, int n, long long idx_base, int D,
                       KbSelect *sel, KbTie *tie, float *tie_score, unsigned int slots,
                       KbRaw cand_raw, double *__restrict__ scores_out,
                       const int *__restrict__ tile_flags, int tiles_stride) {
    __shared__ __align__(128) float4 s_tab[SC_STAGES][SC_CHUNK * 2];
    __shared__ __align__(8) uint64_t s_bar[SC_STAGES];
    extern __shared__ __align__(128) float s_lut[];  // [2G - 1][32]: the current side's look-up table, one copy per bank
    // this tile was scored by tpe_score_quant_grid_kernel (all of its candidates on the grid): nothing left to do
    if (tile_flags && tile_flags[blockIdx.y * tiles_stride + blockIdx.x] == 0) return;
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

    float2 xm[SQ_R / 2], xl[SQ_R / 2], hw[SQ_R / 2];
    bool all_regular = true, all_grid = true;
    // look-up class: shared-memory address of this lane's copy of row (grid index + G - 1), +64 (see KbDim::lut);
    // the table is replicated over the 32 banks (row pitch 128 B), so any mix of rows in a warp is conflict-free
    uint32_t xo[SQ_R];
    const int lutG = max(dm.lut_G[0], dm.lut_G[1]);
    const uint32_t lut_lane = ((threadIdx.x & 31u) << 2) + 64u;  // byte offset into s_lut
#pragma unroll
    for (int r = 0; r < SQ_R; r++) {
        int i = tile_base + r * SC_THREADS + threadIdx.x;
        float4 v = i < n ? stage[(size_t)slot * chunk_cap + i] : make_float4(0.5f, 0.f, dm.qf_hw[0], 2.f);
        if (r & 1) { xm[r / 2].y = v.x; xl[r / 2].y = v.y; hw[r / 2].y = v.z; }
        else { xm[r / 2].x = v.x; xl[r / 2].x = v.y; hw[r / 2].x = v.z; }
        all_regular = all_regular && v.w != 0.f;
        const int gi = (int)v.w - 2;
        all_grid = all_grid && gi >= 0 && gi < lutG;
        xo[r] = lut_lane + ((uint32_t)(gi + lutG - 1) << 7);
    }
    float Shi[SQ_R], Slo[SQ_R];
    double P0[SQ_R];
#pragma unroll
    for (int r = 0; r < SQ_R; r++) { Shi[r] = 0.f; Slo[r] = 0.f; }
    const float qn = (float)(dm.q / (dm.hi - dm.lo));
    float2 qc[KB_QF_DEG + 1];
    float2 qnh0 = make_float2(0.f, 0.f), qts = make_float2(0.f, 0.f);
    bool fast_ok = false, grid_ok = false;
    float2 tvp[SQ_R / 2];      // look-up class: table values of the candidate pairs for grid code tv_code
    uint32_t tv_code = 0u;
#pragma unroll
    for (int p = 0; p < SQ_R / 2; p++) tvp[p] = make_float2(0.f, 0.f);

    for (int c = 0; c < nchunks; c++) {
        const int st = c % SC_STAGES;
        const int side = c < nchunk0 ? 0 : 1;
        const int cc = side == 0 ? c : c - nchunk0;
        const int cnt = min(SC_CHUNK, dm.Kpad[side] - cc * SC_CHUNK);
        mbar_wait(&s_bar[st], (uint32_t)((c / SC_STAGES) & 1));
        const float4 *tab = s_tab[st];
        if (cc == 0) {  // per-side constants of the fitted class, and the side's look-up table
#pragma unroll
            for (int i = 0; i <= KB_QF_DEG; i++) { const float v = dm.qf_coef[side][i]; qc[i] = make_float2(v, v); }
            qnh0 = make_float2(-dm.qf_h0[side], -dm.qf_h0[side]);
            qts = make_float2(dm.qf_tau_s[side], dm.qf_tau_s[side]);
            fast_ok = all_regular && dm.qf_h0[side] > 0.f;
            // (every thread passed the barrier that closed the previous chunk, so the other side's table is dead)
            const int rows = dm.lut_G[side] > 0 ? 2 * dm.lut_G[side] - 1 : 0;
            for (int e = threadIdx.x; e < rows * 32; e += SC_THREADS) s_lut[e] = dm.lut[side][e >> 5];
            grid_ok = all_grid && rows > 0;
            tv_code = 0u;  // the cached table values belong to the other side's table
            __syncthreads();
        }
        for (int j0 = 0; j0 < cnt; j0 += SQ_GROUP) {
            float2 a32[SQ_R / 2];
#pragma unroll
            for (int p = 0; p < SQ_R / 2; p++) a32[p] = make_float2(0.f, 0.f);
            // Runs of components on ONE grid point (the rule in a dense history of an integer parameter: the table is
            // sorted by mu) take a branch-free inner loop: one broadcast LDS of the weight and four packed FFMAs per
            // component.  The group is cut into runs of equal grid code with one ballot per run.
            const uint32_t cj = (uint32_t)__float_as_int(tab[2 * (j0 + (int)(threadIdx.x & 31u)) + 1].w);
            for (int jb = 0, je; jb < SQ_GROUP; jb = je) {
                const uint32_t rcode = __shfl_sync(0xffffffffu, cj, jb);
                const uint32_t eq = __ballot_sync(0xffffffffu, cj == rcode) >> jb;  // bit 0 = component jb itself
                je = jb + (~eq == 0u ? 32 : __ffs((int)~eq) - 1);
                if (rcode != 0u && grid_ok) {
                    if (rcode != tv_code) {
#pragma unroll
                        for (int r = 0; r < SQ_R; r++) {
                            const float tv = *reinterpret_cast<const float *>(reinterpret_cast<const char *>(s_lut) + (xo[r] - rcode));
                            if (r & 1) tvp[r / 2].y = tv; else tvp[r / 2].x = tv;
                        }
                        tv_code = rcode;
                    }
#pragma unroll 4
                    for (int j = jb; j < je; j++) {
                        const float wj = fabsf(tab[2 * (j0 + j) + 1].z);
                        const float2 w2 = make_float2(wj, wj);
#pragma unroll
                        for (int p = 0; p < SQ_R / 2; p++) a32[p] = __ffma2_rn(w2, tvp[p], a32[p]);
                    }
                } else
#pragma unroll 2
            for (int j = jb; j < je; j++) {
                // the class is decided by the weight's sign and the grid code: one broadcast LDS.64; the other six
                // floats of the entry are fetched only by the classes that do arithmetic on mu and sigma
                const float2 Bzw = *reinterpret_cast<const float2 *>(&tab[2 * (j0 + j) + 1].z);
                const float wj = fabsf(Bzw.x);
                if (wj == 0.f) continue;  // padding entry
                const uint32_t code = (uint32_t)__float_as_int(Bzw.y);
                if (code != 0u && grid_ok) {
                    // component and candidate both on the grid: the table value is fetched (one conflict-free LDS per
                    // candidate) only when the component's grid point differs from the previous component's -- the
                    // table is sorted by mu, so a dense history repeats each grid point many times in a row -- and
                    // every (candidate, component) pair costs half a packed FFMA
                    if (code != tv_code) {
#pragma unroll
                        for (int r = 0; r < SQ_R; r++) {
                            const float tv = *reinterpret_cast<const float *>(reinterpret_cast<const char *>(s_lut) + (xo[r] - code));
                            if (r & 1) tvp[r / 2].y = tv; else tvp[r / 2].x = tv;
                        }
                        tv_code = code;
                    }
                    const float2 w2 = make_float2(wj, wj);
#pragma unroll
                    for (int p = 0; p < SQ_R / 2; p++) a32[p] = __ffma2_rn(w2, tvp[p], a32[p]);
                    continue;
                }
                const float4 A = tab[2 * (j0 + j)];
                const float2 Bxy = *reinterpret_cast<const float2 *>(&tab[2 * (j0 + j) + 1].x);
                const float4 B = make_float4(Bxy.x, Bxy.y, Bzw.x, Bzw.y);
                const float2 nmu = make_float2(A.x, A.y), rs = make_float2(A.z, A.w), ms = make_float2(B.x, B.y);
                if (B.z < 0.f && fast_ok) {
                    const float2 w2 = make_float2(wj, wj);
                    // Horner steps run across the four candidate pairs in lockstep: consecutive FFMA2s then share the
                    // coefficient operand (operand-reuse cache) and read two fresh register pairs each
                    float2 tau[SQ_R / 2], z[SQ_R / 2], pp[SQ_R / 2];
#pragma unroll
                    for (int p = 0; p < SQ_R / 2; p++) {
                        const float2 m = __ffma2_rn(__fadd2_rn(__fadd2_rn(xm[p], nmu), xl[p]), rs, ms);
                        const float2 y = make_float2(fabsf(m.x), fabsf(m.y));
                        const float2 t0 = __ffma2_rn(y, qts, make_float2(-1.f, -1.f));
                        tau[p] = make_float2(fminf(t0.x, 1.f), fminf(t0.y, 1.f));
                        z[p] = __fadd2_rn(y, qnh0);
                        pp[p] = __ffma2_rn(qc[KB_QF_DEG], tau[p], qc[KB_QF_DEG - 1]);
                    }
#pragma unroll
                    for (int i = KB_QF_DEG - 2; i >= 0; i--) {
#pragma unroll
                        for (int p = 0; p < SQ_R / 2; p++) pp[p] = __ffma2_rn(pp[p], tau[p], qc[i]);
                    }
#pragma unroll
                    for (int p = 0; p < SQ_R / 2; p++) {
                        const float2 e = __ffma2_rn(make_float2(-z[p].x, -z[p].y), z[p], pp[p]);
                        a32[p] = __ffma2_rn(w2, make_float2(kb_ex2(e.x), kb_ex2(e.y)), a32[p]);
                    }
                } else if (A.z * qn < QUANT_WIDE_THR) {
#pragma unroll
                    for (int r = 0; r < SQ_R; r++) {
                        const float xmr = (r & 1) ? xm[r / 2].y : xm[r / 2].x, xlr = (r & 1) ? xl[r / 2].y : xl[r / 2].x;
                        const float hwr = (r & 1) ? hw[r / 2].y : hw[r / 2].x;
                        const float m = fmaf((xmr + A.x) + xlr, A.z, B.x), h = hwr * A.z;
                        const float mm = m * m;
                        const float u = mm * LN2F, v = (h * h) * LN2F;
                        const float c1 = fmaf(u, 0.6666666667f, -0.3333333333f);
                        const float c2 = fmaf(fmaf(u, 0.1333333333f, -0.4f), u, 0.1f);
                        const float c3 = fmaf(fmaf(fmaf(u, 0.0126984127f, -0.0952380952f), u, 0.1428571429f), u, -0.0238095238f);
                        const float S = fmaf(fmaf(fmaf(c3, v, c2), v, c1), v, 1.0f);
                        const float d2 = (h * (2.2567583342f * KINVF)) * S * kb_ex2(-mm);  // (4/sqrt(pi)) h S e^{-m^2}
                        if (r & 1) a32[r / 2].y = fmaf(wj, d2, a32[r / 2].y); else a32[r / 2].x = fmaf(wj, d2, a32[r / 2].x);
                    }
                } else {
#pragma unroll
                    for (int r = 0; r < SQ_R; r++) {
                        const float xmr = (r & 1) ? xm[r / 2].y : xm[r / 2].x, xlr = (r & 1) ? xl[r / 2].y : xl[r / 2].x;
                        const float hwr = (r & 1) ? hw[r / 2].y : hw[r / 2].x;
                        const float ma = fabsf(fmaf((xmr + A.x) + xlr, A.z, B.x)), h = hwr * A.z;
                        // Phi(b)-Phi(a) = (erfc(a)-erfc(b))/2 with a = m-h, b = m+h, mirrored so that b >= |a|
                        const float a2 = ma - h, b2 = ma + h;
                        const float ea = erfc_pos_scaled(fabsf(a2)), eb = erfc_pos_scaled(b2);
                        const float ea2 = a2 < 0.f ? 2.0f - ea : ea;
                        if (r & 1) a32[r / 2].y = fmaf(wj, ea2 - eb, a32[r / 2].y); else a32[r / 2].x = fmaf(wj, ea2 - eb, a32[r / 2].x);
                    }
                }
            }
            }  // runs of the group
#pragma unroll
            for (int p = 0; p < SQ_R / 2; p++) {
                two_sum_acc(Shi[2 * p], Slo[2 * p], a32[p].x);
                two_sum_acc(Shi[2 * p + 1], Slo[2 * p + 1], a32[p].y);
            }
        }
        __syncthreads();
        if (threadIdx.x == 0 && c + SC_STAGES < nchunks)
            table_stream_issue(dm, c + SC_STAGES, nchunk0, s_tab[st], &s_bar[st]);
        if (c == nchunk0 - 1) {
#pragma unroll
            for (int r = 0; r < SQ_R; r++) {
                P0[r] = log(0.5 * ((double)Shi[r] + (double)Slo[r]) + KB_EPS) - log(dm.p_accept[0] + KB_EPS);
                Shi[r] = 0.f; Slo[r] = 0.f;
            }
        }
    }
    float sc[SQ_R];
#pragma unroll
    for (int r = 0; r < SQ_R; r++) {
        double s64 = P0[r] - (log(0.5 * ((double)Shi[r] + (double)Slo[r]) + KB_EPS) - log(dm.p_accept[1] + KB_EPS));
        sc[r] = (float)s64;
        int i = tile_base + r * SC_THREADS + threadIdx.x;
        if (scores_out && i < n) scores_out[(size_t)i * D + d] = s64;
    }
    if (sel) select_publish<SQ_R, SC_THREADS>(sc, tile_base, n, idx_base, d, D, dm, sel, tie, tie_score, slots,
                                                  KB_MARGIN_QUANT, cand_raw);
}

// ---------------------------------------------------------------------------------------------------------
// quantised parameters, the common case: every candidate of the tile sits ON the quantisation grid (what the sampler
// produces) and the parameter has a look-up table.  Nearly all components are look-up-class then (observations of an
// integer / stepped parameter are on the grid, and a dense history clips their sigma to the floor); the build kernel
// has laid them out as a dense weight array cut into runs of equal grid point (KbDim::qw / qruns).  Per run the table
// value of each candidate is fetched once (one conflict-free LDS from the bank-replicated table), then every
// component of the run is one broadcast weight (LDS.128 per four components) and one packed FFMA per candidate pair:
// the FP32 FMA pipe is the bound, no shuffles, ballots or class branches in the loop.  Every (candidate, component)
// pair is visited.  The few other components (prior, wide sigmas) come from KbDim::tabw and take the arithmetic
// classes.  A tile with an off-grid candidate, or a parameter without a table, is flagged instead and left to
// tpe_score_quant_kernel, which runs right after this kernel and skips the tiles done here.
// ---------------------------------------------------------------------------------------------------------
#define SQG_R 8
#define SQG_WCHUNK 4096     // weights per shared-memory stage (16 KB), two stages
#ifndef SQG_FLUSH
#define SQG_FLUSH 256       // components per FP32 partial sum before it is folded into the (hi, lo) pair (multiple of 4);
                            // measured 128 -> 226 ms, 256 -> 219 ms, 512 -> 216 ms (2^24 candidates x 10 parameters, 3 steps)
#endif
#define SQG_OTHER 64        // other-class entries per shared-memory pass
#ifndef SQG_UNROLL
#define SQG_UNROLL 2        // LDS.128s of weights per trip of the run loop
#endif
constexpr int kSqgUnroll = SQG_UNROLL;
static_assert(SQG_R == SQ_R, "the grid kernel and tpe_score_quant_kernel must agree on the tile size");

__global__ void __launch_bounds__(SC_THREADS, 3)
tpe_score_quant_grid_kernel(const KbDim *__restrict__ dims, const int *__restrict__ quant_dims,
                            const float4 *__restrict__ stage, long long chunk_cap, int n, long long idx_base, int D,
                            KbSelect *sel, KbTie *tie, float *tie_score, unsigned int slots,
                            KbRaw cand_raw, double *__restrict__ scores_out, int *tile_flags,
                            int tiles_stride) {
    __shared__ __align__(128) float s_w[2][SQG_WCHUNK];
    __shared__ __align__(16) float4 s_other[SQG_OTHER * 2];
    __shared__ int4 s_runs[KB_LUT_MAX / 2];
    __shared__ __align__(8) uint64_t s_bar[2];
    extern __shared__ __align__(128) float s_lut[];  // [2G - 1][32]: the current side's look-up table, one copy per bank
    const int slot = blockIdx.y;
    const int d = quant_dims[slot];
    const KbDim &dm = dims[d];
    const int tile_base = blockIdx.x * SQ_TILE;
    const int G = dm.lut_G[0];
    const float4 *my_stage = stage + (size_t)slot * chunk_cap;

    uint32_t xo[SQG_R];  // byte offset of this lane's copy of table row (grid index + G - 1), + 64 (see KbDim::lut)
    bool all_grid = G > 0 && dm.lut_G[1] == G;
    const uint32_t lut_lane = ((threadIdx.x & 31u) << 2) + 64u;
#pragma unroll
    for (int r = 0; r < SQG_R; r++) {
        const int i = tile_base + r * SC_THREADS + threadIdx.x;
        const float code = i < n ? my_stage[i].w : 2.f;
        const int gi = (int)code - 2;
        all_grid = all_grid && gi >= 0 && gi < G;
        xo[r] = lut_lane + ((uint32_t)(gi + G - 1) << 7);
    }
    const int ok = __syncthreads_and(all_grid ? 1 : 0);
    if (threadIdx.x == 0) {
        tile_flags[slot * tiles_stride + blockIdx.x] = ok ? 0 : 1;
        if (ok) { mbar_init(&s_bar[0], 1); mbar_init(&s_bar[1], 1); mbar_fence_init(); }
    }
    if (!ok) return;
    __syncthreads();

    float Shi[SQG_R], Slo[SQG_R];
    float2 a32[SQG_R / 2], tvp[SQG_R / 2];
    double P0[SQG_R];
    int it = 0;  // weight stages consumed so far (both sides): stage = it & 1, mbarrier parity = (it >> 1) & 1
    for (int side = 0; side < 2; side++) {
        const int Kg = dm.Kn[side], Ko = dm.Kw[side], nr = dm.n_runs[side];
        const float *qw = dm.qw[side];
        const int nst = (Kg + SQG_WCHUNK - 1) / SQG_WCHUNK;
        // the side's look-up table (one copy per bank) and run table; every thread is past the previous side's loops
        const int rows = 2 * G - 1;
        for (int e = threadIdx.x; e < rows * 32; e += SC_THREADS) s_lut[e] = dm.lut[side][e >> 5];
        for (int e = threadIdx.x; e < nr; e += SC_THREADS) s_runs[e] = dm.qruns[side][e];
        if (threadIdx.x == 0)
            for (int c = 0; c < 2 && c < nst; c++) {
                const uint32_t bytes = (uint32_t)((min(SQG_WCHUNK, Kg - c * SQG_WCHUNK) + 3) & ~3) * 4u;  // qw is zero-padded
                mbar_expect_tx(&s_bar[(it + c) & 1], bytes);
                tma_bulk_g2s(s_w[(it + c) & 1], qw + (size_t)c * SQG_WCHUNK, bytes, &s_bar[(it + c) & 1]);
            }
#pragma unroll
        for (int r = 0; r < SQG_R; r++) { Shi[r] = 0.f; Slo[r] = 0.f; }
#pragma unroll
        for (int p = 0; p < SQG_R / 2; p++) a32[p] = make_float2(0.f, 0.f);
        auto flush = [&]() {
#pragma unroll
            for (int p = 0; p < SQG_R / 2; p++) {
                two_sum_acc(Shi[2 * p], Slo[2 * p], a32[p].x);
                two_sum_acc(Shi[2 * p + 1], Slo[2 * p + 1], a32[p].y);
                a32[p] = make_float2(0.f, 0.f);
            }
        };
        // ---- the other components (arithmetic classes), while the first weight stages are in flight
        {
            float xm[SQG_R], xl[SQG_R], hw[SQG_R];
            int gi[SQG_R];  // grid index of the candidate
#pragma unroll
            for (int r = 0; r < SQG_R; r++) {
                const int i = tile_base + r * SC_THREADS + threadIdx.x;
                const float4 v = i < n ? my_stage[i] : make_float4(0.5f, 0.f, dm.qf_hw[0], 2.f);
                xm[r] = v.x; xl[r] = v.y; hw[r] = v.z;
                gi[r] = (int)v.w - 2;
            }
            const float qn = (float)(dm.q / (dm.hi - dm.lo));
            const int Kop = (Ko + KB_TABLE_PAD - 1) / KB_TABLE_PAD * KB_TABLE_PAD;
            for (int o0 = 0; o0 < Kop; o0 += SQG_OTHER) {
                const int cnt = min(SQG_OTHER, Kop - o0);
                __syncthreads();
                for (int e = threadIdx.x; e < cnt * 2; e += SC_THREADS) s_other[e] = dm.tabw[side][(size_t)o0 * 2 + e];
                __syncthreads();
                for (int j = 0; j < cnt; j++) {
                    const float4 A = s_other[2 * j], B = s_other[2 * j + 1];
                    const float wj = fabsf(B.z);
                    if (wj == 0.f) continue;  // padding
                    const int code2 = __float_as_int(B.w);
                    if (code2 & 0x40000000) {  // mu on the grid: the component's own table row (KbDim::lut2), one load each
                        const float *row = dm.lut2[side] + ((code2 >> 16) & 0xFF) * (2 * G - 1) + (G - 1) - (code2 & 0xFFFF);
#pragma unroll
                        for (int r = 0; r < SQG_R; r++) {
                            const float tv = __ldg(row + gi[r]);
                            if (r & 1) a32[r / 2].y = fmaf(wj, tv, a32[r / 2].y); else a32[r / 2].x = fmaf(wj, tv, a32[r / 2].x);
                        }
                        continue;
                    }
                    if (A.z * qn < QUANT_WIDE_THR) {  // narrow bin: series about the midpoint
#pragma unroll
                        for (int r = 0; r < SQG_R; r++) {
                            const float m = fmaf((xm[r] + A.x) + xl[r], A.z, B.x), h = hw[r] * A.z;
                            const float mm = m * m;
                            const float u = mm * LN2F, v = (h * h) * LN2F;
                            const float c1 = fmaf(u, 0.6666666667f, -0.3333333333f);
                            const float c2 = fmaf(fmaf(u, 0.1333333333f, -0.4f), u, 0.1f);
                            const float c3 = fmaf(fmaf(fmaf(u, 0.0126984127f, -0.0952380952f), u, 0.1428571429f), u, -0.0238095238f);
                            const float S = fmaf(fmaf(fmaf(c3, v, c2), v, c1), v, 1.0f);
                            const float d2 = (h * (2.2567583342f * KINVF)) * S * kb_ex2(-mm);
                            if (r & 1) a32[r / 2].y = fmaf(wj, d2, a32[r / 2].y); else a32[r / 2].x = fmaf(wj, d2, a32[r / 2].x);
                        }
                    } else {  // erfc(a) - erfc(b), mirrored so that b >= |a|
#pragma unroll
                        for (int r = 0; r < SQG_R; r++) {
                            const float ma = fabsf(fmaf((xm[r] + A.x) + xl[r], A.z, B.x)), h = hw[r] * A.z;
                            const float a2 = ma - h, b2 = ma + h;
                            const float ea = erfc_pos_scaled(fabsf(a2)), eb = erfc_pos_scaled(b2);
                            const float ea2 = a2 < 0.f ? 2.0f - ea : ea;
                            if (r & 1) a32[r / 2].y = fmaf(wj, ea2 - eb, a32[r / 2].y); else a32[r / 2].x = fmaf(wj, ea2 - eb, a32[r / 2].x);
                        }
                    }
                }
                flush();
            }
        }
        __syncthreads();  // s_lut and s_runs are complete (the loop above may have run zero times)
        // ---- the look-up-class components, run by run
        int ri = 0, since = 0;
        for (int c = 0; c < nst; c++, it++) {
            const int st = it & 1;
            mbar_wait(&s_bar[st], (uint32_t)((it >> 1) & 1));
            const float *w = s_w[st];
            const int base = c * SQG_WCHUNK, cnt = min(SQG_WCHUNK, Kg - base);
            while (ri < nr) {
                const int4 run = s_runs[ri];
                if (run.y >= base + cnt) break;
                const int a = max(run.y, base) - base, b = min(run.y + run.z, base + cnt) - base;
#pragma unroll
                for (int r = 0; r < SQG_R; r++) {
                    const float tv = *reinterpret_cast<const float *>(reinterpret_cast<const char *>(s_lut) + (xo[r] - (uint32_t)run.x));
                    if (r & 1) tvp[r / 2].y = tv; else tvp[r / 2].x = tv;
                }
                // runs start on multiples of four and are zero-padded to multiples of four: whole LDS.128s only
                int j = a;
                while (j < b) {
                    const int m = min(b - j, SQG_FLUSH - since);
                    const float4 *wp = reinterpret_cast<const float4 *>(w + j);
#pragma unroll kSqgUnroll
                    for (int k = 0; k < m / 4; k++) {
                        const float4 w4 = wp[k];  // one broadcast LDS.128: four weights
                        const float2 wa = make_float2(w4.x, w4.x), wb = make_float2(w4.y, w4.y);
                        const float2 wc = make_float2(w4.z, w4.z), wd = make_float2(w4.w, w4.w);
#pragma unroll
                        for (int p = 0; p < SQG_R / 2; p++) a32[p] = __ffma2_rn(wa, tvp[p], a32[p]);
#pragma unroll
                        for (int p = 0; p < SQG_R / 2; p++) a32[p] = __ffma2_rn(wb, tvp[p], a32[p]);
#pragma unroll
                        for (int p = 0; p < SQG_R / 2; p++) a32[p] = __ffma2_rn(wc, tvp[p], a32[p]);
#pragma unroll
                        for (int p = 0; p < SQG_R / 2; p++) a32[p] = __ffma2_rn(wd, tvp[p], a32[p]);
                    }
                    j += m; since += m;
                    if (since >= SQG_FLUSH) { flush(); since = 0; }
                }
                if (run.y + run.z <= base + cnt) ri++; else break;
            }
            __syncthreads();  // every thread is done with stage st
            if (threadIdx.x == 0 && c + 2 < nst) {
                const uint32_t bytes = (uint32_t)((min(SQG_WCHUNK, Kg - (c + 2) * SQG_WCHUNK) + 3) & ~3) * 4u;
                mbar_expect_tx(&s_bar[st], bytes);
                tma_bulk_g2s(s_w[st], qw + (size_t)(c + 2) * SQG_WCHUNK, bytes, &s_bar[st]);
            }
        }
        flush();
        if (side == 0) {
#pragma unroll
            for (int r = 0; r < SQG_R; r++)
                P0[r] = log(0.5 * ((double)Shi[r] + (double)Slo[r]) + KB_EPS) - log(dm.p_accept[0] + KB_EPS);
        }
        __syncthreads();  // the next side overwrites s_lut / s_runs
    }
    float sc[SQG_R];
#pragma unroll
    for (int r = 0; r < SQG_R; r++) {
        const double s64 = P0[r] - (log(0.5 * ((double)Shi[r] + (double)Slo[r]) + KB_EPS) - log(dm.p_accept[1] + KB_EPS));
        sc[r] = (float)s64;
        const int i = tile_base + r * SC_THREADS + threadIdx.x;
        if (scores_out && i < n) scores_out[(size_t)i * D + d] = s64;
    }
    if (sel) select_publish<SQG_R, SC_THREADS>(sc, tile_base, n, idx_base, d, D, dm, sel, tie, tie_score, slots,
                                                   KB_MARGIN_QUANT, cand_raw);
}

// ---------------------------------------------------------------------------------------------------------
// FP64 re-scoring of the near ties (oracle formula order: max, sum of exp, log) and the final pick
// ---------------------------------------------------------------------------------------------------------
#define RS_THREADS 256
#define RS_GRID_X 64

struct KbCand { double x; unsigned long long idx; double score; };

// table -> dense list of the entries still within the margin of the FINAL maximum
__global__ void __launch_bounds__(RS_THREADS)
tpe_compact_kernel(const KbDim *__restrict__ dims, const int *__restrict__ numeric_dims, const KbSelect *__restrict__ sel,
                   const KbTie *__restrict__ tie, const float *__restrict__ tie_score, unsigned int slots, KbCand *list,
                   unsigned int *count) {
    const int slot = blockIdx.y, d = numeric_dims[slot];
    const float g = kb_f32_from_orderable(sel[d].gmax_bits);
    const float margin = dims[d].q > 0.0 ? KB_MARGIN_QUANT : KB_MARGIN_CONT;
    const float thr = g - margin * fmaxf(1.0f, fabsf(g));
    const KbTie *tab = tie + (size_t)slot * slots;
    const float *tsc = tie_score + (size_t)slot * slots;
    for (unsigned int s = blockIdx.x * blockDim.x + threadIdx.x; s < slots; s += gridDim.x * blockDim.x) {
        const unsigned long long key = tab[s].key;
        if (key == KB_TIE_EMPTY || !(tsc[s] >= thr)) continue;
        const unsigned int pos = atomicAdd(&count[slot], 1u);
        KbCand c = {__longlong_as_double((long long)key), tab[s].minidx, 0.0};
        list[(size_t)slot * slots + pos] = c;
    }
}

// one warp per listed value: FP64 score in the oracle's formula order
__global__ void __launch_bounds__(RS_THREADS)
tpe_rescore_kernel(const KbDim *__restrict__ dims, const int *__restrict__ numeric_dims, unsigned int slots, KbCand *list,
                   const unsigned int *__restrict__ count) {
    const int slot = blockIdx.y, d = numeric_dims[slot];
    const KbDim &dm = dims[d];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const unsigned int n = count[slot];
    KbCand *l = list + (size_t)slot * slots;
    for (unsigned int e = blockIdx.x * (RS_THREADS / 32) + warp; e < n; e += gridDim.x * (RS_THREADS / 32)) {
        const double sc = rescore_value(dm, l[e].x, lane);
        if (lane == 0) l[e].score = sc;
    }
}

// one CTA per parameter: best listed value (score desc, index asc); categorical winners decode their packed key
// win[D] is the status row of the call: .index = 1 + the highest parameter whose near-tie table overflowed (0 = none).
// It travels with the winners through the all-gather, so that every rank learns of an overflow on any rank.
__global__ void __launch_bounds__(RS_THREADS)
tpe_pick_kernel(const KbDim *__restrict__ dims, const KbSelect *__restrict__ sel, unsigned int slots,
                const KbCand *__restrict__ list, const unsigned int *__restrict__ count, KbWinner *win, long long idx_base,
                int D) {
    __shared__ KbCand s_best[RS_THREADS];
    const int d = blockIdx.x;
    const KbDim &dm = dims[d];
    if (threadIdx.x == 0 && sel[d].overflow)
        atomicMax(reinterpret_cast<unsigned long long *>(&win[D].index), (unsigned long long)(d + 1));
    KbWinner w;
    if (dm.kind == KB_CATEGORICAL) {
        if (threadIdx.x != 0) return;
        const unsigned long long key = sel[d].cat_best;
        const int c = (int)(key & 0xFFFull);
        const unsigned int local = ~(unsigned int)((key >> 12) & 0xFFFFFFFFull);
        w.score = key ? dm.cat_score[c] : -INFINITY;
        w.index = key ? idx_base + (long long)local : -1;
        w.value = (double)c;
        win[d] = w;
        return;
    }
    const KbCand *l = list + (size_t)dm.numeric_slot * slots;
    const unsigned int n = count[dm.numeric_slot];
    KbCand best = {0.0, ~0ull, -INFINITY};
    for (unsigned int e = threadIdx.x; e < n; e += RS_THREADS) {
        KbCand o = l[e];
        if (o.score > best.score || (o.score == best.score && o.idx < best.idx)) best = o;
    }
    s_best[threadIdx.x] = best;
    __syncthreads();
    for (int off = RS_THREADS / 2; off > 0; off >>= 1) {
        if (threadIdx.x < off) {
            KbCand o = s_best[threadIdx.x + off], b = s_best[threadIdx.x];
            if (o.score > b.score || (o.score == b.score && o.idx < b.idx)) s_best[threadIdx.x] = o;
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        best = s_best[0];
        w.score = best.score;
        w.index = best.idx == ~0ull ? -1 : (long long)best.idx;
        double v = best.x;
        if (dm.kind == KB_DISCRETE || dm.kind == KB_STEP_INT) {  // sampleDiscreteUniform: shift back and clamp
            v = v + dm.shift;
            v = fmin(fmax(v, dm.ext_lo), dm.ext_hi);
        }
        w.value = v;
        win[d] = w;
    }
}

// ---------------------------------------------------------------------------------------------------------
// The latency path of kb_tpe_suggest (Katib's real shape: 24 EI candidates, tens to hundreds of trials): ONE kernel,
// one CTA per parameter, that draws the candidates (the same Philox stream as tpe_sample_kernel), scores each in FP64
// in the oracle's formula order (rescore_value -- what the big path falls back to for its near ties) and picks the
// best, lowest index first.  Same winners as the staged pipeline, one launch and one 24 B x D read-back instead of a
// dozen stream operations.
// ---------------------------------------------------------------------------------------------------------
#define SM_THREADS 256
#define SM_MAX_CAND 1024
__global__ void __launch_bounds__(SM_THREADS)
tpe_suggest_small_kernel(const KbDim *__restrict__ dims, int n, uint64_t seed, KbWinner *__restrict__ win) {
    __shared__ double s_x[SM_MAX_CAND], s_sc[SM_MAX_CAND];
    __shared__ double s_bs[SM_THREADS];
    __shared__ int s_bi[SM_THREADS];
    const int d = blockIdx.x;
    const KbDim &dm = dims[d];
    const Philox ph(seed);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < n; i += SM_THREADS) {
        const double x = tpe_sample_one(dm, ph, (unsigned long long)i, d);
        s_x[i] = x;
        if (dm.kind == KB_CATEGORICAL) {
            int c = (int)x;
            c = c < 0 ? 0 : (c >= dm.n_choices ? dm.n_choices - 1 : c);
            s_sc[i] = dm.cat_score[c];
        }
    }
    __syncthreads();
    if (dm.kind != KB_CATEGORICAL) {
        for (int i = warp; i < n; i += SM_THREADS / 32) {
            const double v = rescore_value(dm, s_x[i], lane);
            if (lane == 0) s_sc[i] = v;
        }
        __syncthreads();
    }
    double bs = -INFINITY;
    int bi = n;
    for (int i = threadIdx.x; i < n; i += SM_THREADS) {  // ascending i per thread: ">" keeps the first maximum
        const double v = s_sc[i];
        if (bi == n || v > bs) { bs = v; bi = i; }
    }
    s_bs[threadIdx.x] = bs; s_bi[threadIdx.x] = bi;
    __syncthreads();
    for (int off = SM_THREADS / 2; off > 0; off >>= 1) {
        if (threadIdx.x < off) {
            const double os = s_bs[threadIdx.x + off];
            const int oi = s_bi[threadIdx.x + off];
            if (oi < n && (s_bi[threadIdx.x] == n || os > s_bs[threadIdx.x] || (os == s_bs[threadIdx.x] && oi < s_bi[threadIdx.x]))) {
                s_bs[threadIdx.x] = os; s_bi[threadIdx.x] = oi;
            }
        }
        __syncthreads();
    }
    if (threadIdx.x == 0) {
        KbWinner w;
        const int b = s_bi[0];
        w.score = s_bs[0];
        w.index = b < n ? (long long)b : -1;
        double v = b < n ? s_x[b] : 0.0;
        if (dm.kind == KB_DISCRETE || dm.kind == KB_STEP_INT) {  // sampleDiscreteUniform: shift back and clamp
            v = v + dm.shift;
            v = fmin(fmax(v, dm.ext_lo), dm.ext_hi);
        }
        if (dm.kind == KB_CATEGORICAL) {
            int c = (int)v;
            v = (double)(c < 0 ? 0 : (c >= dm.n_choices ? dm.n_choices - 1 : c));
        }
        w.value = v;
        win[d] = w;
    }
}

// pick the best of `world` gathered winner rows (score desc, index asc)
__global__ void tpe_merge_kernel(const KbWinner *__restrict__ all, int world, int D, KbWinner *win) {
    const int d = blockIdx.x * blockDim.x + threadIdx.x;
    if (d > D) return;
    if (d == D) {  // status row: an overflow on any rank is an overflow of the call, on every rank
        KbWinner st = all[D];
        for (int r = 1; r < world; r++) {
            const KbWinner o = all[(size_t)r * (D + 1) + D];
            if (o.index > st.index) st = o;
        }
        win[D] = st;
        return;
    }
    KbWinner best = all[d];
    for (int r = 1; r < world; r++) {
        KbWinner o = all[(size_t)r * (D + 1) + d];
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
    cudaFree(ctx->d_stage_sorted); cudaFree(ctx->d_perm); cudaFree(ctx->d_bucket);
    ctx->d_stage_sorted = nullptr; ctx->d_perm = nullptr; ctx->d_bucket = nullptr;
    cudaFree(ctx->d_qflags); ctx->d_qflags = nullptr;
    cudaFree(ctx->d_tie_score); cudaFree(ctx->d_tie_list); cudaFree(ctx->d_tie_count); cudaFree(ctx->d_slot_of_dim); cudaFree(ctx->d_numeric_dims);
    if (ctx->h_win) cudaFreeHost(ctx->h_win);
    ctx->d_tie_score = nullptr; ctx->d_tie_list = nullptr; ctx->d_tie_count = nullptr; ctx->tie_slots = 0; ctx->d_slot_of_dim = nullptr; ctx->d_numeric_dims = nullptr;
    ctx->h_win = nullptr;
    return KB_OK;
}

static const int64_t KB_CHUNK_ROWS_DEFAULT = 1 << 22;  // 2^20 -> 2^22: four times fewer kernel tails, +1.2 % on configs[4]
// candidates per staged chunk; KB_CHUNK_LOG2 (16..24) overrides it for experiments
static int64_t kb_chunk_rows() {
    if (const char *e = getenv("KB_CHUNK_LOG2")) { const int l = atoi(e); if (l >= 16 && l <= 24) return (int64_t)1 << l; }
    return KB_CHUNK_ROWS_DEFAULT;
}
#define KB_CHUNK_ROWS kb_chunk_rows()

// size the per-chunk scratch for `rows` candidates per chunk
static int tpe_ensure_scratch(kb_ctx *ctx, int64_t rows, int64_t n_cand, bool need_landing, bool need_scores) {
    kb_ctx *sc = ctx;
    const int D = ctx->D;
    if (!ctx->d_sel) {
        KB_CUDA(ctx, cudaMalloc(&ctx->d_sel, sizeof(KbSelect) * (size_t)D));
        if (!ctx->d_win) KB_CUDA(ctx, cudaMalloc(&ctx->d_win, sizeof(KbWinner) * (size_t)(D + 1)));  // the small path may have made these two
        KB_CUDA(ctx, cudaMalloc(&ctx->d_win_all, sizeof(KbWinner) * (size_t)(D + 1) * 64));
        const size_t slots = (size_t)(ctx->n_numeric > 0 ? ctx->n_numeric : 1);
        KB_CUDA(ctx, cudaMalloc(&ctx->d_tie_count, sizeof(unsigned int) * slots));
        if (!sc->h_win) KB_CUDA(ctx, cudaMallocHost(&sc->h_win, sizeof(KbWinner) * (size_t)(D + 1)));
        std::vector<int> slot_of((size_t)D, 0), numeric(slots, 0);
        for (int i = 0; i < ctx->n_cont; i++) slot_of[(size_t)ctx->cont_dims[(size_t)i]] = i;
        for (int i = 0; i < ctx->n_quant; i++) slot_of[(size_t)ctx->quant_dims[(size_t)i]] = i;
        for (int i = 0; i < ctx->n_cat; i++) slot_of[(size_t)ctx->cat_dims[(size_t)i]] = i;
        for (int d = 0; d < D; d++) if (ctx->h_dims[(size_t)d].numeric_slot >= 0) numeric[(size_t)ctx->h_dims[(size_t)d].numeric_slot] = d;
        KB_CUDA(ctx, cudaMalloc(&sc->d_slot_of_dim, sizeof(int) * (size_t)D));
        if (!sc->d_numeric_dims) KB_CUDA(ctx, cudaMalloc(&sc->d_numeric_dims, sizeof(int) * slots));
        KB_CUDA(ctx, cudaMemcpy(sc->d_slot_of_dim, slot_of.data(), sizeof(int) * (size_t)D, cudaMemcpyHostToDevice));
        KB_CUDA(ctx, cudaMemcpy(sc->d_numeric_dims, numeric.data(), sizeof(int) * slots, cudaMemcpyHostToDevice));
    }
    {   // near-tie table: one slot per 256 candidates, within [2^12, 2^20] per numeric parameter (it holds the distinct
        // values seen within the margin of the running maximum -- hundreds at the headline size; a flat landscape
        // overflows it and the pass is repeated with a table 16 times larger, tpe_retry_overflow)
        unsigned int want = KB_TIE_SLOTS_MIN;
        while (want < KB_TIE_SLOTS_MAX && (int64_t)want * 256 < n_cand) want <<= 1;
        if (want < ctx->tie_slots_floor) want = ctx->tie_slots_floor;  // raised after an overflow (tpe_retry_overflow)
        if (want > ctx->tie_slots) {
            const size_t nn = (size_t)(ctx->n_numeric > 0 ? ctx->n_numeric : 1);
            cudaFree(ctx->d_tie); cudaFree(ctx->d_tie_score); cudaFree(ctx->d_tie_list);
            ctx->d_tie = nullptr; ctx->d_tie_score = nullptr; ctx->d_tie_list = nullptr; ctx->tie_slots = 0;
            KB_CUDA(ctx, cudaMalloc(&ctx->d_tie, sizeof(KbTie) * nn * want));
            KB_CUDA(ctx, cudaMalloc(&ctx->d_tie_score, sizeof(float) * nn * want));
            KB_CUDA(ctx, cudaMalloc(&ctx->d_tie_list, sizeof(KbCand) * nn * want));
            ctx->tie_slots = want;
        }
    }
    if (rows > ctx->chunk_cap) {
        cudaFree(ctx->d_stage_sorted); cudaFree(ctx->d_perm);  // sized by chunk_cap; re-made on demand by the pruned path
        ctx->d_stage_sorted = nullptr; ctx->d_perm = nullptr;
        cudaFree(ctx->d_stage_cont); cudaFree(ctx->d_stage_quant);
        for (int i = 0; i < 2; i++) { cudaFree(ctx->d_cand_chunk[i]); ctx->d_cand_chunk[i] = nullptr; }
        cudaFree(ctx->d_scores_chunk);
        ctx->d_stage_cont = nullptr; ctx->d_stage_quant = nullptr; ctx->d_scores_chunk = nullptr;
        ctx->chunk_cap = 0;
        KB_CUDA(ctx, cudaMalloc(&ctx->d_stage_cont, sizeof(float2) * (size_t)rows * (size_t)(ctx->n_cont > 0 ? ctx->n_cont : 1)));
        KB_CUDA(ctx, cudaMalloc(&ctx->d_stage_quant, sizeof(float4) * (size_t)rows * (size_t)(ctx->n_quant > 0 ? ctx->n_quant : 1)));
        cudaFree(ctx->d_qflags); ctx->d_qflags = nullptr;
        KB_CUDA(ctx, cudaMalloc(&ctx->d_qflags, sizeof(int) * (size_t)((rows + SQ_TILE - 1) / SQ_TILE) * (size_t)(ctx->n_quant > 0 ? ctx->n_quant : 1)));
        ctx->chunk_cap = rows;
    }
    if (need_landing && !ctx->d_cand_chunk[0])
        for (int i = 0; i < 2; i++) KB_CUDA(ctx, cudaMalloc(&ctx->d_cand_chunk[i], sizeof(double) * (size_t)ctx->chunk_cap * D));
    if (need_scores && !ctx->d_scores_chunk)
        KB_CUDA(ctx, cudaMalloc(&ctx->d_scores_chunk, sizeof(double) * (size_t)ctx->chunk_cap * D));
    return KB_OK;
}

// launch stage + scoring kernels for one resident chunk
// d_cand_in == nullptr: the chunk's candidates are global indices [idx_base, idx_base + n) of the Philox stream `seed`,
// drawn straight into the staged columns (kb_tpe_suggest)
static int tpe_launch_chunk(kb_ctx *ctx, const double *d_cand_in, int n, long long idx_base, long long local_base, bool select,
                            double *d_scores, cudaStream_t s, unsigned long long seed = 0, bool as_selection = false) {
    kb_ctx *sc = ctx;
    const int D = ctx->D;
    KbSelect *sel = select ? ctx->d_sel : nullptr;
    const KbRaw d_cand = {d_cand_in, seed};
    if (d_cand_in) {
        KbProf pf(ctx, 0, s);
        tpe_stage_kernel<<<(n + ST_ROWS - 1) / ST_ROWS, ST_THREADS, 0, s>>>(ctx->d_dims, D, d_cand_in, n, ctx->chunk_cap,
                                                                            ctx->d_stage_cont, ctx->d_stage_quant,
                                                                            sc->d_slot_of_dim, sel, local_base, d_scores);
    } else {
        KbProf pf(ctx, 6, s);
        tpe_sample_stage_kernel<<<dim3((unsigned)((n + SS_THREADS * SS_ITEMS - 1) / (SS_THREADS * SS_ITEMS)), (unsigned)D), SS_THREADS, 0, s>>>(
            ctx->d_dims, n, ctx->chunk_cap, seed, idx_base, ctx->d_stage_cont, ctx->d_stage_quant, sc->d_slot_of_dim, sel, local_base);
    }
    KB_LAUNCH_CHECK(ctx);
    // KB_CONT_LOCAL=0 / 2 overrides KB_OPT_LOCAL_ORIGIN for every context of the process (measurement only)
    static const int local_env = [] { const char *e = getenv("KB_CONT_LOCAL"); return e ? atoi(e) : -1; }();
    const int local_opt = local_env >= 0 ? local_env : ctx->tpe_local;
    // bucketing uses atomics, so the composition of a block -- and with it the last bits of an FP32 score -- changes from
    // run to run (the winners, decided in FP64, do not): kb_tpe_scores keeps the ordered kernel unless the option says 2
    const bool ordered_scores = d_scores != nullptr && !as_selection;
    const bool local = !ctx->tpe_prune && (local_opt == 2 || (local_opt == 1 && n >= KB_LOCAL_MIN_ROWS &&
                                                              ctx->n_trials >= KB_LOCAL_MIN_TRIALS && !ordered_scores));
    if (ctx->n_cont > 0 && (ctx->tpe_prune || local)) {
        const size_t per = (size_t)ctx->chunk_cap * (size_t)ctx->n_cont;
        if (!ctx->d_stage_sorted) KB_CUDA(ctx, cudaMalloc(&ctx->d_stage_sorted, sizeof(float2) * per));
        if (!ctx->d_perm) KB_CUDA(ctx, cudaMalloc(&ctx->d_perm, sizeof(int32_t) * per));
        if (!ctx->d_bucket) {  // bucket counters of the counting sort, then the two block counters of kb_tpe_local_stats
            const size_t b = sizeof(int32_t) * KB_PRUNE_BUCKETS * (size_t)ctx->n_cont;
            KB_CUDA(ctx, cudaMalloc(&ctx->d_bucket, b + 2 * sizeof(unsigned long long)));
            KB_CUDA(ctx, cudaMemsetAsync((char *)ctx->d_bucket + b, 0, 2 * sizeof(unsigned long long), s));
        }
        KbProf pf(ctx, 1, s);
        KB_CUDA(ctx, cudaMemsetAsync(ctx->d_bucket, 0, sizeof(int32_t) * KB_PRUNE_BUCKETS * (size_t)ctx->n_cont, s));
        dim3 bgrid((unsigned)((n + BK_THREADS * BK_ITEMS - 1) / (BK_THREADS * BK_ITEMS)), (unsigned)ctx->n_cont);
        tpe_bucket_count_kernel<<<bgrid, BK_THREADS, 0, s>>>(ctx->d_stage_cont, ctx->chunk_cap, n, ctx->d_bucket);
        KB_LAUNCH_CHECK(ctx);
        tpe_bucket_scan_kernel<<<(unsigned)ctx->n_cont, KB_PRUNE_BUCKETS, 0, s>>>(ctx->d_bucket);
        KB_LAUNCH_CHECK(ctx);
        tpe_bucket_scatter_kernel<<<bgrid, BK_THREADS, 0, s>>>(ctx->d_stage_cont, ctx->chunk_cap, n, ctx->d_bucket,
                                                              ctx->d_stage_sorted, ctx->d_perm);
        KB_LAUNCH_CHECK(ctx);
        dim3 grid((unsigned)((n + SC_TILE - 1) / SC_TILE), (unsigned)ctx->n_cont);
        if (ctx->tpe_prune) {
            tpe_score_cont_pruned_kernel<SC_SW_MASK_2><<<grid, SC_THREADS, 0, s>>>(ctx->d_dims, ctx->d_cont_dims, ctx->d_stage_sorted, ctx->d_perm,
                                                                     ctx->chunk_cap, n, idx_base, D, sel, ctx->d_tie, sc->d_tie_score,
                                                                     ctx->tie_slots, d_cand, d_scores);
        } else {
            // slots of 64 on the FMA pipe in blocks that use their own origin (KB_CONT_SW_LOCAL: measurement only)
            static const int swl = [] { const char *e = getenv("KB_CONT_SW_LOCAL"); return e ? atoi(e) : SC_SW_LOCAL_DEFAULT; }();
            unsigned long long *stats = (unsigned long long *)((char *)ctx->d_bucket + sizeof(int32_t) * KB_PRUNE_BUCKETS * (size_t)ctx->n_cont);
            if (!ctx->local_attr_set) {
#define KB_LOCAL_ATTR(N) KB_CUDA(ctx, cudaFuncSetAttribute(tpe_score_cont_local_kernel<kb_sw_slots(N)>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SC_CONT_SMEM))
                KB_LOCAL_ATTR(12); KB_LOCAL_ATTR(16); KB_LOCAL_ATTR(18); KB_LOCAL_ATTR(20); KB_LOCAL_ATTR(22); KB_LOCAL_ATTR(24);
#undef KB_LOCAL_ATTR
                ctx->local_attr_set = true;
            }
#define KB_LOCAL_LAUNCH(N)                                                                                               \
            tpe_score_cont_local_kernel<kb_sw_slots(N)><<<grid, SC_THREADS, SC_CONT_SMEM, s>>>(                          \
                ctx->d_dims, ctx->d_cont_dims, ctx->d_stage_sorted, ctx->d_perm, ctx->chunk_cap, n, idx_base, D, sel,    \
                ctx->d_tie, sc->d_tie_score, ctx->tie_slots, d_cand, d_scores, stats)
            if (swl <= 12) KB_LOCAL_LAUNCH(12);
            else if (swl <= 16) KB_LOCAL_LAUNCH(16);
            else if (swl <= 18) KB_LOCAL_LAUNCH(18);
            else if (swl <= 20) KB_LOCAL_LAUNCH(20);
            else if (swl <= 22) KB_LOCAL_LAUNCH(22);
            else KB_LOCAL_LAUNCH(24);
#undef KB_LOCAL_LAUNCH
        }
        KB_LAUNCH_CHECK(ctx);
    } else if (ctx->n_cont > 0) {
        dim3 grid((unsigned)((n + SC_TILE - 1) / SC_TILE), (unsigned)ctx->n_cont);
        KbProf pf(ctx, 1, s);
        static const int sw = [] { const char *e = getenv("KB_CONT_SW"); return e ? atoi(e) : SC_SW_DEFAULT; }();
        if (!ctx->cont_attr_set) {  // once per context: 48 KB of dynamic shared memory + the barriers is over the default
#define KB_CONT_ATTR(MASK) KB_CUDA(ctx, cudaFuncSetAttribute(tpe_score_cont_kernel<MASK>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)SC_CONT_SMEM))
            KB_CONT_ATTR(SC_SW_MASK_0); KB_CONT_ATTR(SC_SW_MASK_1); KB_CONT_ATTR(SC_SW_MASK_2); KB_CONT_ATTR(SC_SW_MASK_3); KB_CONT_ATTR(SC_SW_MASK_4);
#undef KB_CONT_ATTR
            ctx->cont_attr_set = true;
        }
#define KB_CONT_LAUNCH(MASK)                                                                                              \
        tpe_score_cont_kernel<MASK><<<grid, SC_THREADS, SC_CONT_SMEM, s>>>(ctx->d_dims, ctx->d_cont_dims, ctx->d_stage_cont,           \
                                                                ctx->chunk_cap, n, idx_base, D, sel, ctx->d_tie,           \
                                                                sc->d_tie_score, ctx->tie_slots, d_cand, d_scores)
        if (sw == 0) KB_CONT_LAUNCH(SC_SW_MASK_0);
        else if (sw == 4) KB_CONT_LAUNCH(SC_SW_MASK_4);
        else if (sw == 3) KB_CONT_LAUNCH(SC_SW_MASK_3);
        else if (sw == 1) KB_CONT_LAUNCH(SC_SW_MASK_1);
        else KB_CONT_LAUNCH(SC_SW_MASK_2);
#undef KB_CONT_LAUNCH
        KB_LAUNCH_CHECK(ctx);
    }
    if (ctx->n_quant > 0) {
        dim3 grid((unsigned)((n + SQ_TILE - 1) / SQ_TILE), (unsigned)ctx->n_quant);
        KbProf pf(ctx, 2, s);
        const size_t lut_bytes = (size_t)KB_LUT_MAX * 32 * sizeof(float);  // bank-replicated look-up table
        if (!ctx->quant_attr_set) {  // once per context (static 32 KB + dynamic 32 KB > the 48 KB default)
            KB_CUDA(ctx, cudaFuncSetAttribute(tpe_score_quant_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lut_bytes));
            KB_CUDA(ctx, cudaFuncSetAttribute(tpe_score_quant_grid_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)lut_bytes));
            ctx->quant_attr_set = true;
        }
        // on-grid tiles (the rule) in the grid kernel; it flags the others for the general kernel.  KB_QUANT_GRID=0
        // sends everything through the general kernel (measurement only)
        static const bool use_grid = [] { const char *e = getenv("KB_QUANT_GRID"); return !(e && e[0] == '0'); }();
        const int tiles_stride = (int)((ctx->chunk_cap + SQ_TILE - 1) / SQ_TILE);
        if (use_grid) {
            tpe_score_quant_grid_kernel<<<grid, SC_THREADS, lut_bytes, s>>>(ctx->d_dims, ctx->d_quant_dims, ctx->d_stage_quant,
                                                                            ctx->chunk_cap, n, idx_base, D, sel, ctx->d_tie,
                                                                            sc->d_tie_score, ctx->tie_slots, d_cand, d_scores,
                                                                            ctx->d_qflags, tiles_stride);
            KB_LAUNCH_CHECK(ctx);
        }
        tpe_score_quant_kernel<<<grid, SC_THREADS, lut_bytes, s>>>(ctx->d_dims, ctx->d_quant_dims, ctx->d_stage_quant, ctx->chunk_cap,
                                                           n, idx_base, D, sel, ctx->d_tie, sc->d_tie_score, ctx->tie_slots, d_cand, d_scores,
                                                           use_grid ? ctx->d_qflags : nullptr, tiles_stride);
        KB_LAUNCH_CHECK(ctx);
    }
    return KB_OK;
}

static int tpe_reset_selection(kb_ctx *ctx, cudaStream_t s) {
    const size_t slots = (size_t)(ctx->n_numeric > 0 ? ctx->n_numeric : 1);
    KB_CUDA(ctx, cudaMemsetAsync(ctx->d_sel, 0, sizeof(KbSelect) * (size_t)ctx->D, s));
    KB_CUDA(ctx, cudaMemsetAsync(ctx->d_tie, 0xFF, sizeof(KbTie) * slots * ctx->tie_slots, s));
    KB_CUDA(ctx, cudaMemsetAsync(ctx->d_tie_count, 0, sizeof(unsigned int) * slots, s));
    KB_CUDA(ctx, cudaMemsetAsync(ctx->d_win + ctx->D, 0, sizeof(KbWinner), s));  // status row (tpe_pick_kernel)
    return KB_OK;
}

static int tpe_finish_selection(kb_ctx *ctx, long long idx_base, int64_t *out_index, double *out_value, double *out_score,
                                cudaStream_t s, bool collective = true) {
    kb_ctx *sc = ctx;
    const int D = ctx->D;
    {
        KbProf pf(ctx, 3, s);
        if (ctx->n_numeric > 0) {
            const unsigned int slots = ctx->tie_slots;
            KbCand *list = (KbCand *)ctx->d_tie_list;
            tpe_compact_kernel<<<dim3((slots + RS_THREADS * 8 - 1) / (RS_THREADS * 8), (unsigned)ctx->n_numeric), RS_THREADS, 0, s>>>(
                ctx->d_dims, sc->d_numeric_dims, ctx->d_sel, ctx->d_tie, sc->d_tie_score, slots, list, ctx->d_tie_count);
            KB_LAUNCH_CHECK(ctx);
            tpe_rescore_kernel<<<dim3(RS_GRID_X, (unsigned)ctx->n_numeric), RS_THREADS, 0, s>>>(ctx->d_dims, sc->d_numeric_dims, slots,
                                                                                             list, ctx->d_tie_count);
            KB_LAUNCH_CHECK(ctx);
        }
        tpe_pick_kernel<<<D, RS_THREADS, 0, s>>>(ctx->d_dims, ctx->d_sel, ctx->tie_slots, (const KbCand *)ctx->d_tie_list,
                                                  ctx->d_tie_count, ctx->d_win, idx_base, D);
        KB_LAUNCH_CHECK(ctx);
    }
    int world = 1, rank = 0;
    kb_comm_world(ctx, &world, &rank);
    if (!collective) world = 1;
    if (world > 1) {
        // winners and the status row go through ONE collective: a rank whose near-tie table overflowed is seen by all
        // ranks before any of them trusts the merged winners, and all of them repeat the pass (or fail) together
        if (world > 64) return kb_fail(ctx, KB_ERR_UNSUPPORTED, "world > 64");
        KbProf pf(ctx, 10, s);
        int rc = kb_comm_allgather(ctx, ctx->d_win, ctx->d_win_all, sizeof(KbWinner) * (size_t)(D + 1), s);
        if (rc) return rc;
        tpe_merge_kernel<<<(D + 1 + 127) / 128, 128, 0, s>>>(ctx->d_win_all, world, D, ctx->d_win);
        KB_LAUNCH_CHECK(ctx);
    }
    KB_CUDA(ctx, cudaMemcpyAsync(sc->h_win, ctx->d_win, sizeof(KbWinner) * (size_t)(D + 1), cudaMemcpyDeviceToHost, s));
    KB_CUDA(ctx, cudaStreamSynchronize(s));
    if (sc->h_win[D].index > 0)
        return kb_fail(ctx, KB_ERR_OVERFLOW, "kb_tpe_score_select: near-tie table of parameter %lld overflowed (%u slots%s)",
                       sc->h_win[D].index - 1, ctx->tie_slots, world > 1 ? "; on at least one rank" : "");
    for (int d = 0; d < D; d++) {
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
    return kb_tpe_ensure_tables(ctx);
}

// candidates on the host: chunked, H2D double-buffered on the copy stream, compute on the ctx stream
static int tpe_run_host(kb_ctx *ctx, int64_t n_cand, const double *cand, int64_t index_offset, bool select, double *scores_host) {
    const int D = ctx->D;
    cudaStream_t s = ctx->stream, cs = ctx->copy_stream;
    // chunk size of the host path: at least eight chunks when the set allows it, so that the copy of chunk k + 1
    // hides behind the kernels of chunk k (two chunks would leave half of the H2D time exposed), in 64 Ki-row steps
    int64_t rows = (((n_cand + 7) / 8 + 65535) / 65536) * 65536;
    rows = std::min<int64_t>(std::max<int64_t>(rows, 1 << 18), KB_CHUNK_ROWS);
    if (rows > n_cand) rows = n_cand > 0 ? n_cand : 1;
    int rc = tpe_ensure_scratch(ctx, rows, n_cand, true, scores_host != nullptr);
    if (rc) return rc;
    if (select && (rc = tpe_reset_selection(ctx, s))) return rc;
    const int64_t cap = std::min<int64_t>(ctx->chunk_cap, rows);
    // The first copy has nothing to hide behind: start with a small chunk (64 Ki rows) and double up to the full
    // chunk size, so that the exposed H2D time is that of the small one
    int64_t cur = std::min<int64_t>(cap, 65536);
    int buf = 0;
    for (int64_t off = 0, n64 = 0; off < n_cand; off += n64, buf ^= 1, cur = std::min<int64_t>(cap, cur * 2)) {
        n64 = std::min<int64_t>(cur, n_cand - off);
        const int n = (int)n64;
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

// The near-tie table is sized for the usual case (a handful of distinct values near the maximum).  A flat score
// landscape -- e.g. a history so small that both mixtures are the prior alone -- puts every distinct candidate value
// within the margin; then the pass is repeated with a table 16 times larger (twice at most, up to 2^20 slots per
// parameter) before KB_ERR_OVERFLOW is reported.  With a communicator the overflow status is part of the gathered
// winner rows (tpe_pick_kernel / tpe_merge_kernel), so every rank takes the same decision here and the collectives
// stay balanced; the decision must not depend on this rank's own table size (shards may differ by a row), only on
// the attempt count.
static bool tpe_retry_overflow(kb_ctx *ctx, int rc, int attempt) {
    if (rc != KB_ERR_OVERFLOW || attempt >= 2) return false;
    const unsigned long long next = (unsigned long long)std::max(ctx->tie_slots, (unsigned int)KB_TIE_SLOTS_MIN) * 16ull;
    ctx->tie_slots_floor = next > KB_TIE_SLOTS_MAX ? KB_TIE_SLOTS_MAX : (unsigned int)next;
    return true;
}

extern "C" int kb_tpe_score_select(kb_ctx *ctx, int64_t n_cand, const double *cand, int64_t index_offset,
                                   int64_t *out_index, double *out_value, double *out_score) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    int rc = tpe_precheck(ctx, n_cand, cand, "kb_tpe_score_select");
    if (rc) return rc;
    for (int attempt = 0;; attempt++) {
        rc = tpe_run_host(ctx, n_cand, cand, index_offset, true, nullptr);
        if (!rc) rc = tpe_finish_selection(ctx, index_offset, out_index, out_value, out_score, ctx->stream);
        if (!tpe_retry_overflow(ctx, rc, attempt)) return rc;
    }
}

extern "C" int kb_tpe_score_select_device(kb_ctx *ctx, int64_t n_cand, const double *cand_device, int64_t index_offset,
                                          int64_t *out_index, double *out_value, double *out_score, void *stream) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    int rc = tpe_precheck(ctx, n_cand, cand_device, "kb_tpe_score_select_device");
    if (rc) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    KB_CUDA(ctx, kb_order_after_build(ctx, s));
    const int64_t rows = n_cand < KB_CHUNK_ROWS ? (n_cand > 0 ? n_cand : 1) : KB_CHUNK_ROWS;
    for (int attempt = 0;; attempt++) {
        if ((rc = tpe_ensure_scratch(ctx, rows, n_cand, false, false))) return rc;
        if ((rc = tpe_reset_selection(ctx, s))) return rc;
        const int64_t cap = ctx->chunk_cap;
        for (int64_t off = 0; off < n_cand; off += cap) {
            const int n = (int)(n_cand - off < cap ? n_cand - off : cap);
            if ((rc = tpe_launch_chunk(ctx, cand_device + (size_t)off * ctx->D, n, index_offset + off, off, true, nullptr, s))) return rc;
        }
        rc = tpe_finish_selection(ctx, index_offset, out_index, out_value, out_score, s);
        if (!tpe_retry_overflow(ctx, rc, attempt)) return rc;
    }
}

// Full tpe.Sampler.Sample for every parameter: draw on the device (kb_tpe_sample.cu), score, select.

static int tpe_suggest_range(kb_ctx *ctx, uint64_t seed, int64_t first, int64_t count, bool collective, int64_t *out_index,
                             double *out_value, double *out_score);

extern "C" int kb_tpe_suggest(kb_ctx *ctx, int64_t n_ei_candidates, uint64_t seed, double *out_value) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (!ctx->history_ready) return kb_fail(ctx, KB_ERR_STATE, "kb_tpe_suggest: call kb_tpe_set_history first");
    if (n_ei_candidates < 1 || n_ei_candidates >= (1ll << 32))
        return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_suggest: n_ei_candidates outside [1, 2^32)");
    if (!out_value) return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_suggest: out_value is NULL");
    cudaStream_t s = ctx->stream;
    const int64_t n = n_ei_candidates;
    int world = 1, rank = 0;
    kb_comm_world(ctx, &world, &rank);
    const bool sharded = world > 1 && n >= KB_SUGGEST_SHARD_MIN;
    // a fresh small history and a small request: one kernel builds both estimators in shared memory, draws, scores and
    // picks (kb_tpe_small.cu); the tables in HBM are not built for it
    if (ctx->tables_dirty && kb_tpe_lean_fits(ctx, n)) return kb_tpe_suggest_lean(ctx, n, seed, out_value);
    { int rc = kb_tpe_ensure_tables(ctx); if (rc) return rc; }
    {   // small problems (Katib's own: 24 candidates): the fused single-kernel path
        int64_t kmax = 1;
        for (int d = 0; d < ctx->D; d++) {
            const KbDim &k = ctx->h_dims[(size_t)d];
            if (k.kind != KB_CATEGORICAL) kmax = std::max<int64_t>(kmax, (int64_t)k.K[0] + k.K[1]);
        }
        const char *off = getenv("KB_SUGGEST_SMALL");
        if (n <= SM_MAX_CAND && n * kmax <= (1 << 18) && !(off && off[0] == '0')) {
            if (!ctx->d_win) KB_CUDA(ctx, cudaMalloc(&ctx->d_win, sizeof(KbWinner) * (size_t)(ctx->D + 1)));
            if (!ctx->h_win) KB_CUDA(ctx, cudaMallocHost(&ctx->h_win, sizeof(KbWinner) * (size_t)(ctx->D + 1)));
            {
                KbProf pf(ctx, 3, s);
                tpe_suggest_small_kernel<<<(unsigned)ctx->D, SM_THREADS, 0, s>>>(ctx->d_dims, (int)n, seed, ctx->d_win);
                KB_LAUNCH_CHECK(ctx);
            }
            KB_CUDA(ctx, cudaMemcpyAsync(ctx->h_win, ctx->d_win, sizeof(KbWinner) * (size_t)ctx->D, cudaMemcpyDeviceToHost, s));
            KB_CUDA(ctx, cudaStreamSynchronize(s));
            for (int d = 0; d < ctx->D; d++) out_value[d] = ctx->h_win[d].value;
            return KB_OK;
        }
    }
    // The staged pipeline, candidates drawn chunk by chunk straight into the staged columns (tpe_sample_stage_kernel): no
    // n x D candidate matrix exists, the scratch is the staging of one chunk whatever n is.  With a communicator and
    // n >= KB_SUGGEST_SHARD_MIN this rank draws and scores only its index range of the one global Philox stream; the
    // winners are merged by the collective in tpe_finish_selection, so every rank returns the same values -- those of
    // the unsharded call.
    int64_t first = 0, count = n;
    if (sharded) {
        const int64_t base = n / world, extra = n % world;  // katib_b200/sharding.py: shard_range
        first = rank * base + std::min<int64_t>(rank, extra);
        count = base + (rank < extra ? 1 : 0);
    }
    return tpe_suggest_range(ctx, seed, first, count, sharded, nullptr, out_value, nullptr);
}

// Candidates [first, first + count) of the Philox stream `seed`, drawn into the staged columns, scored, selected.
static int tpe_suggest_range(kb_ctx *ctx, uint64_t seed, int64_t first, int64_t count, bool collective, int64_t *out_index,
                             double *out_value, double *out_score) {
    cudaStream_t s = ctx->stream;
    const int64_t rows = std::max<int64_t>(1, std::min<int64_t>(count, KB_CHUNK_ROWS));
    int rc = KB_OK;
    for (int attempt = 0; !rc; attempt++) {
        rc = tpe_ensure_scratch(ctx, rows, count, false, false);
        if (!rc) rc = tpe_reset_selection(ctx, s);
        const int64_t cap = ctx->chunk_cap;
        for (int64_t off = 0; !rc && off < count; off += cap) {
            const int m = (int)(count - off < cap ? count - off : cap);
            rc = tpe_launch_chunk(ctx, nullptr, m, first + off, off, true, nullptr, s, seed);
        }
        if (!rc) rc = tpe_finish_selection(ctx, first, out_index, out_value, out_score, s, collective);
        if (!tpe_retry_overflow(ctx, rc, attempt)) break;
        rc = KB_OK;
    }
    return rc;
}

// In-process sharding (kb_service_create_multi): one rank's share of a suggestion WITHOUT a collective -- the local
// winners (FP64 score, global candidate index, value) come back to the host, which holds every rank's answer in one
// address space and merges them there (tpe_merge_kernel's rule: best score, lowest index).
int kb_tpe_suggest_range(kb_ctx *ctx, uint64_t seed, int64_t first, int64_t count, int64_t *out_index, double *out_value,
                         double *out_score) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (!ctx->history_ready) return kb_fail(ctx, KB_ERR_STATE, "kb_tpe_suggest_range: no history set");
    if (first < 0 || count < 0 || first + count >= (1ll << 32)) return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_suggest_range: range outside [0, 2^32)");
    { int rc = kb_tpe_ensure_tables(ctx); if (rc) return rc; }
    return tpe_suggest_range(ctx, seed, first, count, false, out_index, out_value, out_score);
}

// ---------------------------------------------------------------------------------------------------------
// measurement helper: worst FP32-path score error of a candidate set, FP64 side computed on the device
// ---------------------------------------------------------------------------------------------------------
// one warp per (candidate, numeric parameter): rescore_value (the FP64 decision formula, oracle-tested) against the
// score the FP32 pipeline produced for the same pair, as selection sees it (rounded to float)
__global__ void __launch_bounds__(RS_THREADS)
tpe_score_error_kernel(const KbDim *__restrict__ dims, const int *__restrict__ numeric_dims, int n_numeric, int D, int n,
                       const double *__restrict__ cand, const double *__restrict__ scores, unsigned long long *err_bits) {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long total = (long long)n * n_numeric;
    for (long long e = (long long)blockIdx.x * (RS_THREADS / 32) + warp; e < total; e += (long long)gridDim.x * (RS_THREADS / 32)) {
        const int slot = (int)(e % n_numeric);
        const long long i = e / n_numeric;
        const int d = numeric_dims[slot];
        const double ref = rescore_value(dims[d], cand[(size_t)i * D + d], lane);
        if (lane == 0) {
            const double got = (double)(float)scores[(size_t)i * D + d];
            double err = fabs(got - ref) / fmax(1.0, fabs(ref));
            if (!(err == err)) err = INFINITY;
            atomicMax(&err_bits[d], (unsigned long long)__double_as_longlong(err));  // err >= 0: bit order = value order
        }
    }
}

extern "C" int kb_tpe_local_stats(kb_ctx *ctx, uint64_t *out_blocks) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (!out_blocks) return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_local_stats: out_blocks is NULL");
    out_blocks[0] = out_blocks[1] = 0;
    if (!ctx->d_bucket) return KB_OK;  // the bucketed path has not run on this context
    const char *stats = (const char *)ctx->d_bucket + sizeof(int32_t) * KB_PRUNE_BUCKETS * (size_t)ctx->n_cont;
    KB_CUDA(ctx, cudaMemcpyAsync(out_blocks, stats, 2 * sizeof(uint64_t), cudaMemcpyDeviceToHost, ctx->stream));
    KB_CUDA(ctx, cudaStreamSynchronize(ctx->stream));
    return KB_OK;
}

extern "C" int kb_tpe_score_error_device(kb_ctx *ctx, int64_t n_cand, const double *cand_device, double *out_max_err) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    int rc = tpe_precheck(ctx, n_cand, cand_device, "kb_tpe_score_error_device");
    if (rc) return rc;
    if (!out_max_err) return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_score_error_device: out_max_err is NULL");
    const int D = ctx->D;
    for (int d = 0; d < D; d++) out_max_err[d] = 0.0;
    if (n_cand == 0 || ctx->n_numeric == 0) return KB_OK;
    cudaStream_t s = ctx->stream;
    const int64_t rows = std::min<int64_t>(n_cand, 1 << 20);
    if ((rc = tpe_ensure_scratch(ctx, rows, n_cand, false, true))) return rc;
    unsigned long long *d_err = nullptr;
    KB_CUDA(ctx, cudaMalloc(&d_err, sizeof(unsigned long long) * (size_t)D));
    cudaMemsetAsync(d_err, 0, sizeof(unsigned long long) * (size_t)D, s);
    for (int64_t off = 0; off < n_cand && !rc; off += rows) {
        const int n = (int)std::min<int64_t>(rows, n_cand - off);
        rc = tpe_launch_chunk(ctx, cand_device + (size_t)off * D, n, off, off, false, ctx->d_scores_chunk, s, 0, true);  // the kernels selection uses
        if (rc) break;
        tpe_score_error_kernel<<<148 * 8, RS_THREADS, 0, s>>>(ctx->d_dims, ctx->d_numeric_dims, ctx->n_numeric, D, n,
                                                              cand_device + (size_t)off * D, ctx->d_scores_chunk, d_err);
        ctx->launches++;
        if (cudaGetLastError() != cudaSuccess) rc = kb_fail(ctx, KB_ERR_CUDA, "kb_tpe_score_error_device: launch failed");
    }
    std::vector<unsigned long long> h((size_t)D, 0ull);
    if (!rc) {
        cudaError_t e = cudaMemcpyAsync(h.data(), d_err, sizeof(unsigned long long) * (size_t)D, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) rc = kb_fail(ctx, KB_ERR_CUDA, "kb_tpe_score_error_device: %s", cudaGetErrorString(e));
    }
    cudaFree(d_err);
    if (rc) return rc;
    for (int d = 0; d < D; d++) memcpy(&out_max_err[d], &h[(size_t)d], sizeof(double));
    return KB_OK;
}
