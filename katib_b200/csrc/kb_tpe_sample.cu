// kb_tpe_sample.cu -- candidate generation on the device: the first half of tpe.Sampler.Sample.
//
// Replaces (goptuna v0.8.0, not in the reference tree, go.mod:8; reached from sample.go:48-90):
//   Sampler.sampleFromGMM              -> tpe_sample_kernel, numeric parameters: pick a component of the BELOW
//                                         mixture by its weight, draw N(mu, sigma), reject until low <= x < high,
//                                         then round to the quantum (Go math.Round: half away from zero)
//   Sampler.sampleFromCategoricalDist  -> same kernel, categorical parameters: inverse CDF over p_below
//   goptuna.RandomSampler.Sample       -> random_fill_kernel (converter.go:96-106, and TPE's start-up phase)
// goptuna draws from Go's math/rand (not reproducible outside Go), so only the distribution can be matched; the
// stream here is Philox4x32-10 keyed by the caller's seed with counter (candidate index, parameter, attempt).
#include <math.h>

#include "kb_internal.cuh"
#include "kb_tpe_sample.cuh"

// one thread per (candidate, parameter); lanes run along the parameter so row-major stores coalesce
__global__ void tpe_sample_kernel(const KbDim *__restrict__ dims, int D, long long n, uint64_t seed, long long first,
                                  double *__restrict__ out) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * D) return;
    const long long i = t / D;
    const int d = (int)(t % D);
    const KbDim &dm = dims[d];
    const Philox ph(seed);
    out[i * D + d] = tpe_sample_one(dm, ph, (unsigned long long)(first + i), d);
}

// goptuna.RandomSampler: uniform over each distribution, internal representation
__global__ void random_fill_kernel(const KbDim *__restrict__ dims, int D, long long n, uint64_t seed, unsigned long long first,
                                   double *__restrict__ out) {
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= n * D) return;
    const long long i = t / D;
    const int d = (int)(t % D);
    const KbDim &dm = dims[d];
    const Philox ph(seed ^ 0x5bd1e995a5a5a5a5ull);
    const unsigned long long gi = first + (unsigned long long)i;
    const uint4 r = ph((uint32_t)gi, (uint32_t)(gi >> 32), (uint32_t)d, 0u);
    const double u = u01(r.x, r.y);
    double v;
    switch (dm.kind) {
        case KB_UNIFORM: v = u * (dm.ext_hi - dm.ext_lo) + dm.ext_lo; break;
        case KB_INT: v = floor(u * (dm.ext_hi - dm.ext_lo + 1.0)) + dm.ext_lo; v = fmin(v, dm.ext_hi); break;
        case KB_STEP_INT: {
            const double steps = floor((dm.ext_hi - dm.ext_lo) / dm.q);
            v = fmin(floor(u * (steps + 1.0)), steps) * dm.q + dm.ext_lo;
            break;
        }
        case KB_DISCRETE: {  // sample on the shifted grid [ -q/2, r + q/2 ), round, shift back, clamp
            const double s = u * (dm.hi - dm.lo) + dm.lo;
            v = round(s / dm.q) * dm.q + dm.shift;
            v = fmin(fmax(v, dm.ext_lo), dm.ext_hi);
            break;
        }
        default: v = fmin(floor(u * (double)dm.n_choices), (double)(dm.n_choices - 1)); break;
    }
    out[i * D + d] = v;
}

static int sample_launch(kb_ctx *ctx, int64_t n, uint64_t seed, double *d_out, cudaStream_t s) {
    const long long total = (long long)n * ctx->D;
    if (total == 0) return KB_OK;
    KbProf pf(ctx, 6, s);
    tpe_sample_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(ctx->d_dims, ctx->D, n, seed, 0, d_out);
    KB_LAUNCH_CHECK(ctx);
    return KB_OK;
}

extern "C" int kb_tpe_sample_candidates(kb_ctx *ctx, int64_t n, uint64_t seed, double *cand_host, double *cand_device,
                                        void *stream) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (!ctx->history_ready) return kb_fail(ctx, KB_ERR_STATE, "kb_tpe_sample_candidates: call kb_tpe_set_history first");
    if (n < 0 || n >= (1ll << 32)) return kb_fail(ctx, KB_ERR_UNSUPPORTED, "kb_tpe_sample_candidates: n outside [0, 2^32)");
    if (n > 0 && !cand_host && !cand_device) return kb_fail(ctx, KB_ERR_INVALID, "kb_tpe_sample_candidates: no output buffer");
    if (n == 0) return KB_OK;
    { int rc = kb_tpe_ensure_tables(ctx); if (rc) return rc; }
    cudaStream_t s = stream ? (cudaStream_t)stream : ctx->stream;
    KB_CUDA(ctx, kb_order_after_build(ctx, s));
    double *d_out = cand_device;
    if (!d_out) KB_CUDA(ctx, cudaMalloc(&d_out, sizeof(double) * (size_t)n * ctx->D));
    int rc = sample_launch(ctx, n, seed, d_out, s);
    if (rc == KB_OK && cand_host) {
        cudaError_t e = cudaMemcpyAsync(cand_host, d_out, sizeof(double) * (size_t)n * ctx->D, cudaMemcpyDeviceToHost, s);
        if (e == cudaSuccess) e = cudaStreamSynchronize(s);
        if (e != cudaSuccess) rc = kb_fail(ctx, KB_ERR_CUDA, "kb_tpe_sample_candidates: D2H: %s", cudaGetErrorString(e));
    }
    if (!cand_device) cudaFree(d_out);
    return rc;
}

extern "C" int kb_random_fill(kb_ctx *ctx, uint64_t seed, uint64_t first_index, int64_t n, double *out) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (ctx->D == 0) return kb_fail(ctx, KB_ERR_STATE, "kb_random_fill: call kb_space_set first");
    if (n < 0) return kb_fail(ctx, KB_ERR_INVALID, "kb_random_fill: n < 0");
    if (n == 0) return KB_OK;
    if (!out) return kb_fail(ctx, KB_ERR_INVALID, "kb_random_fill: out is NULL");
    cudaStream_t s = ctx->stream;
    double *d_out = nullptr;
    KB_CUDA(ctx, cudaMalloc(&d_out, sizeof(double) * (size_t)n * ctx->D));
    const long long total = (long long)n * ctx->D;
    random_fill_kernel<<<(unsigned)((total + 255) / 256), 256, 0, s>>>(ctx->d_dims, ctx->D, n, seed, first_index, d_out);
    ctx->launches++;
    cudaError_t e = cudaGetLastError();
    if (e == cudaSuccess) e = cudaMemcpyAsync(out, d_out, sizeof(double) * (size_t)total, cudaMemcpyDeviceToHost, s);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s);
    cudaFree(d_out);
    if (e != cudaSuccess) return kb_fail(ctx, KB_ERR_CUDA, "kb_random_fill: %s", cudaGetErrorString(e));
    return KB_OK;
}
