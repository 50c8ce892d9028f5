// kb_ctx.cu -- context lifetime, error text and the search-space table.
// Reference counterparts: goptuna.CreateStudy (converter.go:369) for the ctx,
// toGoptunaSearchSpace (converter.go:110-184) for kb_space_set.
#include <stdarg.h>

#include "kb_internal.cuh"

static thread_local std::string g_create_err;
static void prof_drain(kb_ctx *ctx);

int kb_fail(const kb_ctx *ctx, int code, const char *fmt, ...) {
    char buf[1024];
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(buf, sizeof buf, fmt, ap);
    va_end(ap);
    if (ctx) ctx->err = buf; else g_create_err = buf;
    return code;
}

extern "C" int kb_version(void) { return 100; }

extern "C" const char *kb_last_error(const kb_ctx *ctx) {
    return ctx ? ctx->err.c_str() : g_create_err.c_str();
}

extern "C" int kb_create(int device, kb_ctx **out) {
    if (!out) return kb_fail(nullptr, KB_ERR_INVALID, "kb_create: out is NULL");
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0)
        return kb_fail(nullptr, KB_ERR_CUDA, "kb_create: no CUDA device (%s); this library has no CPU path",
                       cudaGetErrorString(e));
    if (device < 0 || device >= count)
        return kb_fail(nullptr, KB_ERR_INVALID, "kb_create: device %d out of range [0,%d)", device, count);
    kb_ctx *ctx = new kb_ctx();
    ctx->device = device;
    KbDeviceGuard g(device);
    if ((e = cudaGetDeviceProperties(&ctx->prop, device)) != cudaSuccess) {
        delete ctx;
        return kb_fail(nullptr, KB_ERR_CUDA, "kb_create: cudaGetDeviceProperties: %s", cudaGetErrorString(e));
    }
    if (ctx->prop.major != 10) {
        int mj = ctx->prop.major, mn = ctx->prop.minor;
        delete ctx;
        return kb_fail(nullptr, KB_ERR_UNSUPPORTED,
                       "kb_create: device is sm_%d%d; this library is built for sm_100a (B200) only", mj, mn);
    }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess ||
        cudaStreamCreateWithFlags(&ctx->copy_stream, cudaStreamNonBlocking) != cudaSuccess) {
        delete ctx;
        return kb_fail(nullptr, KB_ERR_CUDA, "kb_create: cudaStreamCreate failed");
    }
    for (int i = 0; i < 2; i++) {
        cudaEventCreateWithFlags(&ctx->ev_h2d[i], cudaEventDisableTiming);
        cudaEventCreateWithFlags(&ctx->ev_done[i], cudaEventDisableTiming);
    }
    cudaEventCreateWithFlags(&ctx->ev_build, cudaEventDisableTiming);
    *out = ctx;
    return KB_OK;
}

extern "C" int kb_destroy(kb_ctx *ctx) {
    if (!ctx) return KB_OK;
    {
        KbDeviceGuard g(ctx->device);
        cudaDeviceSynchronize();
        kb_tpe_free(ctx);
        kb_sobol_free(ctx);
        kb_cmaes_free(ctx);
        kb_comm_free(ctx);
        cudaFree(ctx->d_dims);
        cudaFree(ctx->d_cont_dims);
        cudaFree(ctx->d_quant_dims);
        cudaFree(ctx->d_cat_dims);
        for (int i = 0; i < 2; i++) {
            if (ctx->ev_h2d[i]) cudaEventDestroy(ctx->ev_h2d[i]);
            if (ctx->ev_done[i]) cudaEventDestroy(ctx->ev_done[i]);
        }
        if (ctx->ev_build) cudaEventDestroy(ctx->ev_build);
        if (ctx->ev_up) cudaEventDestroy(ctx->ev_up);
        if (ctx->h_up) cudaFreeHost(ctx->h_up);
        prof_drain(ctx);
        for (auto e : ctx->prof_pool) cudaEventDestroy(e);
        if (ctx->stream) cudaStreamDestroy(ctx->stream);
        if (ctx->copy_stream) cudaStreamDestroy(ctx->copy_stream);
    }
    delete ctx;
    return KB_OK;
}

extern "C" int kb_device_info(const kb_ctx *ctx, int64_t out[5]) {
    if (!ctx || !out) return KB_ERR_INVALID;
    int khz = 0;
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, ctx->device);
    out[0] = ctx->prop.multiProcessorCount;
    out[1] = ctx->prop.major;
    out[2] = ctx->prop.minor;
    out[3] = khz;
    out[4] = ctx->prop.l2CacheSize;
    return KB_OK;
}

extern "C" void *kb_stream(const kb_ctx *ctx) { return ctx ? (void *)ctx->stream : nullptr; }
extern "C" int64_t kb_launch_count(const kb_ctx *ctx) { return ctx ? ctx->launches : 0; }

static void prof_drain(kb_ctx *ctx) {
    for (int k = 0; k < KB_PROF_N; k++) {
        for (auto &pr : ctx->prof_pending[k]) {
            float ms = 0.f;
            cudaEventSynchronize(pr.second);
            if (cudaEventElapsedTime(&ms, pr.first, pr.second) == cudaSuccess) { ctx->prof_ms[k] += ms; ctx->prof_launches[k]++; }
            ctx->prof_pool.push_back(pr.first);
            ctx->prof_pool.push_back(pr.second);
        }
        ctx->prof_pending[k].clear();
    }
}

extern "C" int kb_profile(kb_ctx *ctx, int enable) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    prof_drain(ctx);
    if (enable) for (int k = 0; k < KB_PROF_N; k++) { ctx->prof_ms[k] = 0; ctx->prof_launches[k] = 0; }
    ctx->profile = enable != 0;
    return KB_OK;
}

extern "C" int kb_profile_read(kb_ctx *ctx, int kernel_class, double *ms_total, int64_t *launches) {
    if (!ctx || kernel_class < 0 || kernel_class >= KB_PROF_N) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    prof_drain(ctx);
    if (ms_total) *ms_total = ctx->prof_ms[kernel_class];
    if (launches) *launches = ctx->prof_launches[kernel_class];
    return KB_OK;
}

extern "C" int kb_space_set(kb_ctx *ctx, int n_params, const int32_t *kind, const double *low,
                            const double *high, const double *step, const int32_t *n_choices) {
    if (!ctx) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    if (n_params < 1 || n_params > KB_MAX_PARAMS)
        return kb_fail(ctx, KB_ERR_UNSUPPORTED, "kb_space_set: n_params=%d outside [1,%d]", n_params, KB_MAX_PARAMS);
    if (!kind) return kb_fail(ctx, KB_ERR_INVALID, "kb_space_set: kind is NULL");
    std::vector<KbDim> dims((size_t)n_params);
    std::vector<int> cont, quant, cat;
    int numeric = 0;
    for (int d = 0; d < n_params; d++) {
        KbDim k;
        memset(&k, 0, sizeof k);
        k.kind = kind[d];
        k.numeric_slot = -1;
        if (k.kind == KB_CATEGORICAL) {
            if (!n_choices || n_choices[d] < 1 || n_choices[d] > KB_MAX_CHOICES)
                return kb_fail(ctx, KB_ERR_INVALID, "kb_space_set: param %d: n_choices outside [1,%d]", d, KB_MAX_CHOICES);
            k.n_choices = n_choices[d];
            k.lo = 0; k.hi = (double)k.n_choices; k.ext_lo = 0; k.ext_hi = (double)(k.n_choices - 1);
            cat.push_back(d);
        } else {
            if (!low || !high) return kb_fail(ctx, KB_ERR_INVALID, "kb_space_set: low/high NULL");
            double lo = low[d], hi = high[d];
            k.ext_lo = lo; k.ext_hi = hi;
            if (!(lo <= hi)) return kb_fail(ctx, KB_ERR_INVALID, "kb_space_set: param %d: low > high", d);
            // working domain of tpe.Sampler.sample{Uniform,Int,StepInt,DiscreteUniform}
            switch (k.kind) {
                case KB_UNIFORM: k.lo = lo; k.hi = hi; k.q = 0; k.shift = 0; break;
                case KB_INT: k.q = 1.0; k.lo = lo - 0.5; k.hi = hi + 0.5; k.shift = 0; break;
                case KB_STEP_INT:
                case KB_DISCRETE: {
                    if (!step || !(step[d] > 0))
                        return kb_fail(ctx, KB_ERR_INVALID, "kb_space_set: param %d: step must be > 0", d);
                    k.q = step[d];
                    // sampleStepInt goes through sampleDiscreteUniform like DISCRETE: the range is shifted to
                    // [0, high - low] so that rounding to multiples of the step lands on the grid {low + k step}
                    // whatever low is (random_fill_kernel and the Sobol mapping use the same grid)
                    k.lo = 0 - 0.5 * k.q; k.hi = (hi - lo) + 0.5 * k.q; k.shift = lo;
                    break;
                }
                default: return kb_fail(ctx, KB_ERR_INVALID, "kb_space_set: param %d: unknown kind %d", d, k.kind);
            }
            if (!(k.lo < k.hi)) return kb_fail(ctx, KB_ERR_INVALID, "kb_space_set: param %d: empty range", d);
            k.numeric_slot = numeric++;
            (k.q > 0 ? quant : cont).push_back(d);
        }
        dims[(size_t)d] = k;
    }
    kb_tpe_free(ctx);  // tables depend on the space
    cudaFree(ctx->d_dims); cudaFree(ctx->d_cont_dims); cudaFree(ctx->d_quant_dims); cudaFree(ctx->d_cat_dims);
    ctx->d_dims = nullptr; ctx->d_cont_dims = ctx->d_quant_dims = ctx->d_cat_dims = nullptr;
    ctx->D = n_params;
    ctx->h_dims = dims;
    ctx->h_dims_sent.clear();  // d_dims is a new allocation
    ctx->cont_dims = cont; ctx->quant_dims = quant; ctx->cat_dims = cat;
    ctx->n_cont = (int)cont.size(); ctx->n_quant = (int)quant.size(); ctx->n_cat = (int)cat.size();
    ctx->n_numeric = numeric;
    KB_CUDA(ctx, cudaMalloc(&ctx->d_dims, sizeof(KbDim) * (size_t)n_params));
    KB_CUDA(ctx, cudaMemcpy(ctx->d_dims, dims.data(), sizeof(KbDim) * (size_t)n_params, cudaMemcpyHostToDevice));
    auto up = [&](std::vector<int> &v, int **dst) -> cudaError_t {
        size_t b = sizeof(int) * (v.empty() ? 1 : v.size());
        cudaError_t e = cudaMalloc(dst, b);
        if (e == cudaSuccess && !v.empty()) e = cudaMemcpy(*dst, v.data(), sizeof(int) * v.size(), cudaMemcpyHostToDevice);
        return e;
    };
    KB_CUDA(ctx, up(cont, &ctx->d_cont_dims));
    KB_CUDA(ctx, up(quant, &ctx->d_quant_dims));
    KB_CUDA(ctx, up(cat, &ctx->d_cat_dims));
    return KB_OK;
}
