typedef unsigned long long f2_t;
__global__ void k(const f2_t *in, f2_t *out) {
    f2_t a = in[threadIdx.x], b = in[threadIdx.x + 32], c = in[threadIdx.x + 64], m, s;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(m) : "l"(a), "l"(b));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(s) : "l"(m), "l"(c));
    out[threadIdx.x] = s;
}
