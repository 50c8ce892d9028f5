// kb_peaks.cu -- saturating micro-kernels that measure this GPU's own ceilings for the pipes the hot
// kernels are bound by (MUFU.EX2, FP32 FMA, FP64 FMA, HBM write / copy).  MEASURED_PEAKS.json holds
// only HBM copy and bf16 GEMM; the TPE scoring kernels are SFU- / FP32-bound, so their roofline
// denominators are measured here, on the same device and clocks as the run that quotes them.
#include "kb_internal.cuh"

#define PK_THREADS 256
#define PK_ITERS 4096

__global__ void __launch_bounds__(PK_THREADS) peak_mufu_kernel(float *out, float seed) {
    float a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed + (float)(threadIdx.x + i) * 1e-3f;
    for (int it = 0; it < PK_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = kb_ex2(a[i]) - 1.0f;  // MUFU.EX2 + FADD, 8 independent chains
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i];
    if (s == 123.456f) out[0] = s;
}

__global__ void __launch_bounds__(PK_THREADS) peak_ffma_kernel(float *out, float seed) {
    float a[16];
#pragma unroll
    for (int i = 0; i < 16; i++) a[i] = seed + (float)(threadIdx.x + i);
    const float m = 1.000001f, c = seed;
    for (int it = 0; it < PK_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 16; i++) a[i] = fmaf(a[i], m, c);
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i++) s += a[i];
    if (s == 123.456f) out[0] = s;
}

__global__ void __launch_bounds__(PK_THREADS) peak_ffma2_kernel(float *out, float seed) {
    float2 a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = make_float2(seed + (float)(threadIdx.x + i), seed - (float)i);
    const float2 m = make_float2(1.000001f, 0.999999f), c = make_float2(seed, seed);
    for (int it = 0; it < PK_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = __ffma2_rn(a[i], m, c);  // FFMA2: two FMAs per lane per instruction
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i].x + a[i].y;
    if (s == 123.456f) out[0] = s;
}

__global__ void __launch_bounds__(PK_THREADS) peak_dfma_kernel(double *out, double seed) {
    double a[8];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = seed + (double)(threadIdx.x + i);
    const double m = 1.000001, c = seed;
    for (int it = 0; it < PK_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = fma(a[i], m, c);
    }
    double s = 0.0;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i];
    if (s == 123.456) out[0] = s;
}

__global__ void __launch_bounds__(PK_THREADS) peak_write_kernel(double2 *out, size_t n2) {
    const double2 v = make_double2(1.0, 2.0);
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (size_t)gridDim.x * blockDim.x) __stcs(out + i, v);
}
__global__ void __launch_bounds__(PK_THREADS) peak_copy_kernel(const double2 *__restrict__ in, double2 *__restrict__ out, size_t n2) {
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n2; i += (size_t)gridDim.x * blockDim.x) __stcs(out + i, __ldcs(in + i));
}
// FFMA2 and scalar FFMA interleaved (8 packed + 8 scalar chains): do the two forms share one pipe or run side by side?
__global__ void __launch_bounds__(PK_THREADS) peak_ffma_mix_kernel(float *out, float seed) {
    float2 a[8];
    float b[8];
#pragma unroll
    for (int i = 0; i < 8; i++) { a[i] = make_float2(seed + (float)(threadIdx.x + i), seed - (float)i); b[i] = seed * (float)(i + 1); }
    const float2 m = make_float2(1.000001f, 0.999999f), c = make_float2(seed, seed);
    for (int it = 0; it < PK_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) { a[i] = __ffma2_rn(a[i], m, c); b[i] = fmaf(b[i], m.x, c.x); }
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i].x + a[i].y + b[i];
    if (s == 123.456f) out[0] = s;
}
// FFMA2 with two varying 64-bit register operands (accumulator and multiplicand), the grid kernel's operand pattern
__global__ void __launch_bounds__(PK_THREADS) peak_ffma2_2op_kernel(float *out, float seed) {
    float2 a[8], t[4];
#pragma unroll
    for (int i = 0; i < 8; i++) a[i] = make_float2(seed + (float)(threadIdx.x + i), seed - (float)i);
#pragma unroll
    for (int i = 0; i < 4; i++) t[i] = make_float2(1.0f + 1e-7f * (float)(threadIdx.x + i), 1.0f - 1e-7f * (float)i);
    float w = seed * 1e-9f;
    for (int it = 0; it < PK_ITERS; it++) {
#pragma unroll
        for (int i = 0; i < 8; i++) a[i] = __ffma2_rn(make_float2(w, w), t[i & 3], a[i]);
        w += 1e-12f;
    }
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; i++) s += a[i].x + a[i].y;
    if (s == 123.456f) out[0] = s;
}

extern "C" int kb_peak_measure(kb_ctx *ctx, int which, double *out) {
    if (!ctx || !out) return KB_ERR_INVALID;
    std::lock_guard<std::mutex> lk(ctx->mu);
    KbDeviceGuard g(ctx->device);
    cudaStream_t s = ctx->stream;
    const int sms = ctx->prop.multiProcessorCount;
    cudaEvent_t e0, e1;
    KB_CUDA(ctx, cudaEventCreate(&e0));
    KB_CUDA(ctx, cudaEventCreate(&e1));
    void *buf = nullptr;
    const size_t bytes = (size_t)4 << 30;  // 4 GiB: far beyond the 126 MB L2
    if (which == 3 || which == 4) KB_CUDA(ctx, cudaMalloc(&buf, which == 4 ? 2 * bytes : bytes));
    else KB_CUDA(ctx, cudaMalloc(&buf, 256));
    const int grid = sms * 8;
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {
        KB_CUDA(ctx, cudaEventRecord(e0, s));
        double work = 0.0;
        switch (which) {
            case 0: peak_mufu_kernel<<<grid, PK_THREADS, 0, s>>>((float *)buf, 0.5f); work = (double)grid * PK_THREADS * PK_ITERS * 8; break;
            case 1: peak_ffma_kernel<<<grid, PK_THREADS, 0, s>>>((float *)buf, 0.5f); work = (double)grid * PK_THREADS * PK_ITERS * 16; break;
            case 2: peak_dfma_kernel<<<grid, PK_THREADS, 0, s>>>((double *)buf, 0.5); work = (double)grid * PK_THREADS * PK_ITERS * 8; break;
            case 3: peak_write_kernel<<<sms * 16, PK_THREADS, 0, s>>>((double2 *)buf, bytes / 16); work = (double)bytes; break;
            case 5: peak_ffma2_kernel<<<grid, PK_THREADS, 0, s>>>((float *)buf, 0.5f); work = (double)grid * PK_THREADS * PK_ITERS * 16; break;
            case 6: peak_ffma_mix_kernel<<<grid, PK_THREADS, 0, s>>>((float *)buf, 0.5f); work = (double)grid * PK_THREADS * PK_ITERS * 24; break;
            case 7: peak_ffma2_2op_kernel<<<grid, PK_THREADS, 0, s>>>((float *)buf, 0.5f); work = (double)grid * PK_THREADS * PK_ITERS * 16; break;
            case 4: peak_copy_kernel<<<sms * 16, PK_THREADS, 0, s>>>((const double2 *)buf, (double2 *)((char *)buf + bytes), bytes / 16); work = 2.0 * (double)bytes; break;
            default: cudaFree(buf); return kb_fail(ctx, KB_ERR_INVALID, "kb_peak_measure: which=%d", which);
        }
        ctx->launches++;
        KB_CUDA(ctx, cudaEventRecord(e1, s));
        KB_CUDA(ctx, cudaEventSynchronize(e1));
        float ms = 0.f;
        KB_CUDA(ctx, cudaEventElapsedTime(&ms, e0, e1));
        double rate = work / ((double)ms * 1e-3) / 1e9;  // Gop/s or GB/s
        if (rep > 0 && rate > best) best = rate;
    }
    cudaFree(buf);
    cudaEventDestroy(e0); cudaEventDestroy(e1);
    KB_CUDA(ctx, cudaGetLastError());
    *out = best;
    return KB_OK;
}
