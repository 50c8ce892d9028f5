#include <cstdio>
#include <cuda_runtime.h>
__global__ void k(double *out, long long *cyc, double x0, int n) {
    double x = x0, y = 1.000000001;
    long long t0 = clock64();
    for (int i = 0; i < n; i++) x = fma(x, y, 1e-9);
    long long t1 = clock64();
    cyc[0] = t1 - t0;
    double z = x0 + 2.0;
    t0 = clock64();
    for (int i = 0; i < n; i++) z = rsqrt(z) + 1.5;
    t1 = clock64();
    cyc[1] = t1 - t0;
    double w = x0 + 2.0;
    t0 = clock64();
    for (int i = 0; i < n; i++) w = (double)((float)w * 1.0001f);
    t1 = clock64();
    cyc[2] = t1 - t0;
    float f = (float)x0 + 2.f;
    t0 = clock64();
    for (int i = 0; i < n; i++) f = fmaf(f, 1.0001f, 1e-9f);
    t1 = clock64();
    cyc[3] = t1 - t0;
    double u = x0 + 2.0;
    t0 = clock64();
    for (int i = 0; i < n; i++) u = 1.0 / u + 1.5;
    t1 = clock64();
    cyc[4] = t1 - t0;
    double v = x0 + 2.0;
    t0 = clock64();
    for (int i = 0; i < n; i++) v = sqrt(v) + 1.5;
    t1 = clock64();
    cyc[5] = t1 - t0;
    float q = (float)x0 + 2.f;
    t0 = clock64();
    for (int i = 0; i < n; i++) q = rsqrtf(q) + 1.5f;
    t1 = clock64();
    cyc[6] = t1 - t0;
    t0 = clock64();
    for (int i = 0; i < n; i++) __syncthreads();
    t1 = clock64();
    cyc[7] = t1 - t0;
    out[threadIdx.x] = x + z + w + f + u + v + q;
}
int main() {
    double *o; long long *c; cudaMalloc(&o, 8 * 1024); cudaMalloc(&c, 64);
    for (int threads : {32, 256}) {
        k<<<1, threads>>>(o, c, 1.0, 1000); cudaDeviceSynchronize();
        k<<<1, threads>>>(o, c, 1.0, 1000); cudaDeviceSynchronize();
        long long h[8]; cudaMemcpy(h, c, 64, cudaMemcpyDeviceToHost);
        const char *nm[8] = {"dfma", "rsqrt64+add", "cvt f64->f32 mul ->f64", "ffma", "1/x f64 + add", "sqrt64+add", "rsqrtf+add", "syncthreads"};
        printf("threads %d\n", threads);
        for (int i = 0; i < 8; i++) printf("  %-26s %.1f cycles/iter\n", nm[i], h[i] / 1000.0);
    }
    return 0;
}
