// kb_sobol.cu -- Sobol points on the GPU: gray-code XOR of Joe-Kuo direction numbers.
// Replaces sobol.Engine.Draw() behind sobol.NewSampler() (reference converter.go:94, reached from
// sample.go:39).  Integer work + an 8-byte store per value: HBM-write bound.
//
// Layout: out[p*dim + d].  A CTA covers 32 dimensions x (8 point groups of SOBOL_P consecutive points);
// lanes run along the dimension so that every store instruction writes 256 contiguous bytes.  The
// state at a group's first index comes from the closed form x_i = XOR_{b in gray(i)} v[b]; the next
// SOBOL_P-1 points flip one direction number each (v[ctz(~i)]), the low four of which sit in registers.
#include <algorithm>
#include <thread>

#include "kb_internal.cuh"
#include "sobol_joe_kuo.inc"

#define SOBOL_P 16        // consecutive points per thread (group start is 16-aligned)
#define SOBOL_GROUPS 8    // point groups per CTA
#define SOBOL_DT 32       // dimensions per CTA

__global__ void __launch_bounds__(SOBOL_DT *SOBOL_GROUPS)
sobol_fill_kernel(const uint32_t *__restrict__ V, int dim, unsigned long long first, long long n_points,
                  const double *__restrict__ low, const double *__restrict__ high, double *__restrict__ out) {
    __shared__ uint32_t sv[32][SOBOL_DT];
    const int dl = threadIdx.x % SOBOL_DT, pg = threadIdx.x / SOBOL_DT;
    const int d = blockIdx.y * SOBOL_DT + dl;
    for (int b = pg; b < 32; b += SOBOL_GROUPS) sv[b][dl] = d < dim ? V[b * KB_SOBOL_MAX_DIM + d] : 0u;
    __syncthreads();
    if (d >= dim) return;
    const uint32_t v0 = sv[0][dl], v1 = sv[1][dl], v2 = sv[2][dl], v3 = sv[3][dl];
    const bool scaled = low != nullptr;
    const double lo = scaled ? low[d] : 0.0, width = scaled ? high[d] - low[d] : 1.0;

    // groups are aligned to SOBOL_P in absolute sequence index so the low-bit flip pattern is static
    const unsigned long long a0 = first & ~(unsigned long long)(SOBOL_P - 1);
    const unsigned long long last = first + (unsigned long long)n_points;  // exclusive
    const long long n_groups = (long long)((last - a0 + SOBOL_P - 1) / SOBOL_P);
    for (long long gi = (long long)blockIdx.x * SOBOL_GROUPS + pg; gi < n_groups;
         gi += (long long)gridDim.x * SOBOL_GROUPS) {
        const unsigned long long i0 = a0 + (unsigned long long)gi * SOBOL_P;
        uint32_t g = (uint32_t)(i0 ^ (i0 >> 1)), x = 0;
        while (g) {
            int b = __ffs(g) - 1;
            x ^= sv[b][dl];
            g &= g - 1;
        }
#pragma unroll
        for (int p = 0; p < SOBOL_P; p++) {
            const unsigned long long i = i0 + p;
            if (i >= first && i < last) {
                // u = x / 2^32 exactly: build 1 + x*2^-32 from bits, subtract 1
                double u = __hiloint2double((int)(0x3FF00000u | (x >> 12)), (int)(x << 20)) - 1.0;
                if (scaled) u = __dadd_rn(lo, __dmul_rn(u, width));
                __stcs(out + (size_t)(i - first) * dim + d, u);
            }
            // flip for i -> i+1: direction ctz(~i); static for the low four bits of a 16-aligned group
            const int c = (p & 1) == 0 ? 0 : (p & 2) == 0 ? 1 : (p & 4) == 0 ? 2 : (p & 8) == 0 ? 3 : -1;
            if (c == 0) x ^= v0; else if (c == 1) x ^= v1; else if (c == 2) x ^= v2; else if (c == 3) x ^= v3;
        }
    }
}

static int sobol_tables(kb_ctx *ctx) {
    if (ctx->d_sobol_v) return KB_OK;
    const int bits = 32;
    std::vector<uint32_t> V((size_t)32 * KB_SOBOL_MAX_DIM);
    for (int d = 0; d < KB_SOBOL_MAX_DIM; d++) {
        uint64_t v[32];
        if (d == 0) {
            for (int j = 0; j < bits; j++) v[j] = 1;
        } else {
            const uint32_t p = kJoeKuoPoly[d];
            int m = 0;
            for (uint32_t t = p >> 1; t; t >>= 1) m++;
            for (int j = 0; j < m; j++) v[j] = kJoeKuoInit[d][j];
            for (int j = m; j < bits; j++) {  // m_j = m_{j-s} ^ 2^s m_{j-s} ^ XOR_k 2^k a_k m_{j-k}
                uint64_t nv = v[j - m], pow2 = 1;
                for (int k = 0; k < m; k++) {
                    pow2 <<= 1;
                    if ((p >> (m - 1 - k)) & 1) nv ^= pow2 * v[j - k - 1];
                }
                v[j] = nv;
            }
        }
        for (int j = 0; j < bits; j++) V[(size_t)j * KB_SOBOL_MAX_DIM + d] = (uint32_t)(v[j] << (bits - 1 - j));
    }
    KB_CUDA(ctx, cudaMalloc(&ctx->d_sobol_v, V.size() * sizeof(uint32_t)));
    KB_CUDA(ctx, cudaMemcpy(ctx->d_sobol_v, V.data(), V.size() * sizeof(uint32_t), cudaMemcpyHostToDevice));
    return KB_OK;
}

int kb_sobol_free(kb_ctx *ctx) {
    cudaFree(ctx->d_sobol_v);
    ctx->d_sobol_v = nullptr;
    for (int i = 0; i < 2; i++) {
        cudaFree(ctx->d_sobol_buf[i]); ctx->d_sobol_buf[i] = nullptr;
        if (ctx->h_sobol_pin[i]) cudaFreeHost(ctx->h_sobol_pin[i]);
        ctx->h_sobol_pin[i] = nullptr;
    }
    ctx->sobol_buf_bytes = ctx->sobol_pin_bytes = 0;
    return KB_OK;
}

static int sobol_launch(kb_ctx *ctx, int dim, uint64_t first, int64_t n, const double *d_low, const double *d_high,
                        double *d_out, cudaStream_t s) {
    if (n == 0) return KB_OK;
    const long long n_groups = (long long)(((first + (uint64_t)n) - (first & ~(uint64_t)(SOBOL_P - 1)) + SOBOL_P - 1) / SOBOL_P);
    long long bx = (n_groups + SOBOL_GROUPS - 1) / SOBOL_GROUPS;
    const int by = (dim + SOBOL_DT - 1) / SOBOL_DT;
    // grid-stride over point groups: a whole number of waves (8 CTAs of 256 threads fit one SM)
    const long long wave = (long long)ctx->prop.multiProcessorCount * 8;
    const long long cap = wave * 8 / by > 0 ? wave * 8 / by : 1;
    if (bx > cap) bx = cap;
    dim3 grid((unsigned)bx, (unsigned)by);
    KbProf pf(ctx, 4, s);
    sobol_fill_kernel<<<grid, SOBOL_DT * SOBOL_GROUPS, 0, s>>>(ctx->d_sobol_v, dim, first, n, d_low, d_high, d_out);
    KB_LAUNCH_CHECK(ctx);
    return KB_OK;
}

static int sobol_check(kb_ctx *ctx, int dim, uint64_t first, int64_t n, const double *low, const double *high,
                       const void *out) {
    if (dim < 1 || dim > KB_SOBOL_MAX_DIM)
        return kb_fail(ctx, KB_ERR_UNSUPPORTED, "kb_sobol_fill: dim=%d outside [1,%d]", dim, KB_SOBOL_MAX_DIM);
    if (n < 0) return kb_fail(ctx, KB_ERR_INVALID, "kb_sobol_fill: n_points < 0");
    if (first + (uint64_t)n > (1ull << 32))
        return kb_fail(ctx, KB_ERR_UNSUPPORTED, "kb_sobol_fill: sequence index beyond 2^32 (32-bit generator)");
    if ((low == nullptr) != (high == nullptr)) return kb_fail(ctx, KB_ERR_INVALID, "kb_sobol_fill: low/high must both be given");
    if (n > 0 && !out) return kb_fail(ctx, KB_ERR_INVALID, "kb_sobol_fill: out is NULL");
    return KB_OK;
}

extern "C" int kb_sobol_fill_device(kb_ctx *ctx, int dim, uint64_t first_index, int64_t n_points, const double *low,
                                    const double *high, double *out_device, void *stream) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    int rc = sobol_check(ctx, dim, first_index, n_points, low, high, out_device);
    if (rc) return rc;
    if ((rc = sobol_tables(ctx))) return rc;
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    double *d_lh = nullptr;
    if (low) {  // low/high are host arrays of `dim` doubles
        KB_CUDA(ctx, cudaMallocAsync(&d_lh, sizeof(double) * 2 * (size_t)dim, s));
        KB_CUDA(ctx, cudaMemcpyAsync(d_lh, low, sizeof(double) * (size_t)dim, cudaMemcpyHostToDevice, s));
        KB_CUDA(ctx, cudaMemcpyAsync(d_lh + dim, high, sizeof(double) * (size_t)dim, cudaMemcpyHostToDevice, s));
    }
    rc = sobol_launch(ctx, dim, first_index, n_points, d_lh, d_lh ? d_lh + dim : nullptr, out_device, s);
    if (d_lh) cudaFreeAsync(d_lh, s);
    return rc;
}

// Host-buffer fill.  The block is produced in chunks of about 32 MB into two alternating device buffers and leaves
// through the copy stream while the next chunk is being generated; no allocation per call (the buffers belong to the
// context and only grow), which matters for the service: it draws ONE point per suggestion (sample.go:39).  A
// destination in pinned / registered host memory is written by the copies directly; a pageable one goes through two
// pinned staging buffers, and a few host threads move each finished chunk on while the GPU works on the next ones.
#define SOBOL_CHUNK_BYTES ((size_t)32 << 20)
#define SOBOL_COPY_THREADS 4

static void sobol_host_move(double *dst, const double *src, size_t bytes) {
    if (bytes < ((size_t)4 << 20)) { memcpy(dst, src, bytes); return; }
    std::thread th[SOBOL_COPY_THREADS - 1];
    const size_t part = (bytes / SOBOL_COPY_THREADS) & ~(size_t)4095;
    for (int t = 0; t < SOBOL_COPY_THREADS - 1; t++)
        th[t] = std::thread([=] { memcpy((char *)dst + (size_t)t * part, (const char *)src + (size_t)t * part, part); });
    const size_t done = (size_t)(SOBOL_COPY_THREADS - 1) * part;
    memcpy((char *)dst + done, (const char *)src + done, bytes - done);
    for (auto &t : th) t.join();
}

extern "C" int kb_sobol_fill(kb_ctx *ctx, int dim, uint64_t first_index, int64_t n_points, const double *low,
                             const double *high, double *out) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    int rc = sobol_check(ctx, dim, first_index, n_points, low, high, out);
    if (rc) return rc;
    if (n_points == 0) return KB_OK;
    if ((rc = sobol_tables(ctx))) return rc;
    cudaStream_t s = ctx->stream, cs = ctx->copy_stream;
    const size_t row = sizeof(double) * (size_t)dim;
    const int64_t chunk_pts = std::max<int64_t>(1, std::min<int64_t>(n_points, (int64_t)(SOBOL_CHUNK_BYTES / row)));
    const size_t chunk_bytes = (size_t)chunk_pts * row;
    cudaPointerAttributes attr;
    const bool pinned_dst = cudaPointerGetAttributes(&attr, out) == cudaSuccess && attr.type == cudaMemoryTypeHost;
    cudaGetLastError();  // an unregistered pointer is not an error here
    const size_t lh_bytes = low ? 2 * row : 0;
    if (chunk_bytes + lh_bytes > ctx->sobol_buf_bytes) {
        for (int i = 0; i < 2; i++) { cudaFree(ctx->d_sobol_buf[i]); ctx->d_sobol_buf[i] = nullptr; }
        ctx->sobol_buf_bytes = 0;
        const size_t want = std::max<size_t>(chunk_bytes + lh_bytes, (size_t)64 << 10);
        for (int i = 0; i < 2; i++) KB_CUDA(ctx, cudaMalloc(&ctx->d_sobol_buf[i], want));
        ctx->sobol_buf_bytes = want;
    }
    if (!pinned_dst && chunk_bytes > ctx->sobol_pin_bytes) {
        for (int i = 0; i < 2; i++) { if (ctx->h_sobol_pin[i]) cudaFreeHost(ctx->h_sobol_pin[i]); ctx->h_sobol_pin[i] = nullptr; }
        ctx->sobol_pin_bytes = 0;
        const size_t want = std::max<size_t>(chunk_bytes, (size_t)64 << 10);
        for (int i = 0; i < 2; i++) KB_CUDA(ctx, cudaMallocHost(&ctx->h_sobol_pin[i], want));
        ctx->sobol_pin_bytes = want;
    }
    const double *d_low = nullptr, *d_high = nullptr;
    if (low) {  // the bounds ride behind the chunk area of buffer 0 (never overwritten by a chunk)
        double *d_lh = (double *)((char *)ctx->d_sobol_buf[0] + (ctx->sobol_buf_bytes - lh_bytes));
        KB_CUDA(ctx, cudaMemcpyAsync(d_lh, low, row, cudaMemcpyHostToDevice, s));
        KB_CUDA(ctx, cudaMemcpyAsync(d_lh + dim, high, row, cudaMemcpyHostToDevice, s));
        d_low = d_lh; d_high = d_lh + dim;
    }
    const int64_t n_chunks = (n_points + chunk_pts - 1) / chunk_pts;
    auto drain = [&](int64_t k) -> cudaError_t {  // chunk k has left the device: move it on if it sits in staging
        const int buf = (int)(k & 1);
        cudaError_t e = cudaEventSynchronize(ctx->ev_h2d[buf]);
        if (e == cudaSuccess && !pinned_dst) {
            const int64_t off = k * chunk_pts, m = std::min<int64_t>(chunk_pts, n_points - off);
            sobol_host_move(out + (size_t)off * dim, ctx->h_sobol_pin[buf], (size_t)m * row);
        }
        return e;
    };
    for (int64_t k = 0; k < n_chunks; k++) {
        const int buf = (int)(k & 1);
        const int64_t off = k * chunk_pts, m = std::min<int64_t>(chunk_pts, n_points - off);
        if (k >= 2) KB_CUDA(ctx, drain(k - 2));  // device buffer and staging buffer `buf` are free again
        if ((rc = sobol_launch(ctx, dim, first_index + (uint64_t)off, m, d_low, d_high, ctx->d_sobol_buf[buf], s))) return rc;
        KB_CUDA(ctx, cudaEventRecord(ctx->ev_done[buf], s));
        KB_CUDA(ctx, cudaStreamWaitEvent(cs, ctx->ev_done[buf], 0));
        double *dst = pinned_dst ? out + (size_t)off * dim : ctx->h_sobol_pin[buf];
        KB_CUDA(ctx, cudaMemcpyAsync(dst, ctx->d_sobol_buf[buf], (size_t)m * row, cudaMemcpyDeviceToHost, cs));
        KB_CUDA(ctx, cudaEventRecord(ctx->ev_h2d[buf], cs));
        // the fill of chunk k + 2 reuses this device buffer: it must wait for this copy
        KB_CUDA(ctx, cudaStreamWaitEvent(s, ctx->ev_h2d[buf], 0));
    }
    for (int64_t k = std::max<int64_t>(0, n_chunks - 2); k < n_chunks; k++) KB_CUDA(ctx, drain(k));
    return KB_OK;
}
