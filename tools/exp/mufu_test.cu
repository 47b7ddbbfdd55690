// Error statistics of the reciprocal-square-root variants against double precision (experiment, not product code).
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>
__device__ __forceinline__ float mufu_rsqrt(float x) { float r; asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float newton(float x) { const float y = mufu_rsqrt(x); const float e = fmaf(-x * y, y, 1.f); return fmaf(0.5f * y, e, y); }
__device__ __forceinline__ float newton2(float x) { const float y = mufu_rsqrt(x); const float h = 0.5f * y; const float s = x * y; const float e = fmaf(-s, h, 0.5f); return fmaf(y, e, y); }
__device__ __forceinline__ float ieee(float x) { return __fdiv_rn(1.f, __fsqrt_rn(x)); }
__global__ void k(int n, double* out) {
    // out: per variant {sum signed rel err, sum abs rel err, max abs rel err} ; also |b|^2 - 1 bias of normalised vectors
    double s[4][3] = {};
    double nb[4] = {};
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < n; i += gridDim.x * blockDim.x) {
        unsigned h = i * 2654435761u; h ^= h >> 15; h *= 2246822519u; h ^= h >> 13;
        const float a = 0.2f + 1.6f * (h & 0xffffff) / 16777216.f;
        unsigned h2 = h * 3266489917u; h2 ^= h2 >> 16;
        const float b = -0.8f + 1.6f * (h2 & 0xffffff) / 16777216.f, c = 0.3f * b + 0.1f;
        const float x = fmaf(c, c, fmaf(b, b, a * a));
        const double t = 1.0 / sqrt((double)x);
        const float v[4] = {mufu_rsqrt(x), newton(x), newton2(x), ieee(x)};
        for (int q = 0; q < 4; ++q) {
            const double r = ((double)v[q] - t) / t;
            s[q][0] += r; s[q][1] += fabs(r); s[q][2] = fmax(s[q][2], fabs(r));
            const float b0 = a * v[q], b1 = b * v[q], b2 = c * v[q];
            nb[q] += (double)b0 * b0 + (double)b1 * b1 + (double)b2 * b2 - 1.0;
        }
    }
    for (int q = 0; q < 4; ++q) {
        atomicAdd(out + q * 4, s[q][0]); atomicAdd(out + q * 4 + 1, s[q][1]); atomicAdd(out + q * 4 + 3, nb[q]);
        // max via atomicMax on bits (non-negative doubles order like unsigned long long)
        atomicMax((unsigned long long*)(out + q * 4 + 2), (unsigned long long)__double_as_longlong(s[q][2]));
    }
}
int main() {
    const int n = 1 << 24;
    double* d; cudaMalloc(&d, 16 * sizeof(double)); cudaMemset(d, 0, 16 * sizeof(double));
    k<<<296, 256>>>(n, d);
    double h[16]; cudaMemcpy(h, d, sizeof(h), cudaMemcpyDeviceToHost);
    const char* names[4] = {"rsqrt.approx", "approx + Newton (e = 1 - x y y)", "approx + Newton (s = x y form)", "1 / sqrt.rn (div.rn)"};
    for (int q = 0; q < 4; ++q)
        printf("%-34s mean signed rel err % .3e  mean |rel err| %.3e  max %.3e  mean(|v|^2 - 1) % .3e\n", names[q], h[q * 4] / n,
               h[q * 4 + 1] / n, h[q * 4 + 2], h[q * 4 + 3] / n);
    return 0;
}
