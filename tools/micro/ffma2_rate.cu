#include <cstdio>
#include <cuda_runtime.h>
__device__ __forceinline__ float2 ffma2(float2 a, float2 b, float2 c) {
    unsigned long long ra, rb, rc, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rc) : "f"(c.x), "f"(c.y));
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(rd) : "l"(ra), "l"(rb), "l"(rc));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}
template <int MODE>
__global__ void k(float *out, float s, int n) {
    float2 a0 = make_float2(threadIdx.x * 1e-3f, 1.0f), a1 = make_float2(2.0f, threadIdx.x * 2e-3f), a2 = make_float2(0.5f, 0.25f), a3 = make_float2(0.125f, 3.0f);
    float2 a4 = a0, a5 = a1, a6 = a2, a7 = a3;
    const float2 m = make_float2(s, s * 0.999f), c = make_float2(1e-3f, 2e-3f);
    for (int i = 0; i < n; i++) {
        if (MODE == 0) {
#define F1(v) v.x = fmaf(v.x, m.x, c.x); v.y = fmaf(v.y, m.y, c.y);
            F1(a0) F1(a1) F1(a2) F1(a3) F1(a4) F1(a5) F1(a6) F1(a7)
        } else {
            a0 = ffma2(a0, m, c); a1 = ffma2(a1, m, c); a2 = ffma2(a2, m, c); a3 = ffma2(a3, m, c);
            a4 = ffma2(a4, m, c); a5 = ffma2(a5, m, c); a6 = ffma2(a6, m, c); a7 = ffma2(a7, m, c);
        }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = a0.x + a0.y + a1.x + a1.y + a2.x + a2.y + a3.x + a3.y + a4.x + a4.y + a5.x + a5.y + a6.x + a6.y + a7.x + a7.y;
}
int main() {
    float *d; cudaMalloc(&d, 148 * 8 * 256 * 4);
    const int n = 20000;
    for (int mode = 0; mode < 2; mode++) {
        cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
        for (int rep = 0; rep < 3; rep++) {
            cudaEventRecord(e0);
            if (mode == 0) k<0><<<148 * 8, 256>>>(d, 0.9999f, n); else k<1><<<148 * 8, 256>>>(d, 0.9999f, n);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1);
            double fl = 2.0 * 16 * n * 148.0 * 8 * 256;
            if (rep == 2) printf("mode %d (%s): %.3f ms, %.1f TFLOP/s\n", mode, mode ? "fma.rn.f32x2" : "fmaf", ms, fl / ms / 1e9);
        }
    }
    return 0;
}
