// Micro-benchmark of the sm_100a issue model for the instruction mix of the render kernels.
// Loop body = NF scalar FFMA + NP packed FFMA2 + NL LOP3 + NM integer min + NX MUFU.RCP, every op an
// `asm volatile` on one of 8..16 independent chains (the compiler cannot merge or drop them),
// 16 warps per SMSP.  Prints cycles per loop body per warp scheduler.
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o issue_mix issue_mix.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int NF, int NP, int NL, int NM, int NX>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed) {
    float f[16]; unsigned long long p[8]; unsigned u[8]; float x[4];
#pragma unroll
    for (int j = 0; j < 16; ++j) f[j] = seed + threadIdx.x + j;
#pragma unroll
    for (int j = 0; j < 8; ++j) { p[j] = (unsigned long long)__float_as_uint(seed + j) << 32 | __float_as_uint(seed - threadIdx.x); u[j] = threadIdx.x * (2 * j + 1); }
#pragma unroll
    for (int j = 0; j < 4; ++j) x[j] = seed + j + threadIdx.x;
    const float m = 0.999f, c = 0.001f;
    const unsigned long long m2 = (unsigned long long)__float_as_uint(0.999f) << 32 | __float_as_uint(0.998f);
    const unsigned long long c2 = (unsigned long long)__float_as_uint(0.001f) << 32 | __float_as_uint(0.002f);
    for (int i = 0; i < iters; ++i) {
        constexpr int N = NF + NP + NL + NM + NX;
        int nf = 0, np = 0, nl = 0, nm = 0, nx = 0;
#pragma unroll
        for (int s = 0; s < N; ++s) {   // interleave the op kinds proportionally
            if (nf * N <= s * NF && nf < NF) { asm volatile("fma.rn.ftz.f32 %0, %0, %1, %2;" : "+f"(f[nf & 15]) : "f"(m), "f"(c)); ++nf; }
            else if (np * N <= s * NP && np < NP) { asm volatile("fma.rn.ftz.f32x2 %0, %0, %1, %2;" : "+l"(p[np & 7]) : "l"(m2), "l"(c2)); ++np; }
            else if (nl * N <= s * NL && nl < NL) { asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(u[nl & 7]) : "r"(0xffffffc0u), "r"(u[(nl + 1) & 7])); ++nl; }
            else if (nm * N <= s * NM && nm < NM) { asm volatile("min.u32 %0, %0, %1;" : "+r"(u[nm & 7]) : "r"(u[(nm + 3) & 7])); ++nm; }
            else if (nx < NX) { asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(x[nx & 3])); ++nx; }
            else if (nf < NF) { asm volatile("fma.rn.ftz.f32 %0, %0, %1, %2;" : "+f"(f[nf & 15]) : "f"(m), "f"(c)); ++nf; }
            else if (np < NP) { asm volatile("fma.rn.ftz.f32x2 %0, %0, %1, %2;" : "+l"(p[np & 7]) : "l"(m2), "l"(c2)); ++np; }
            else if (nl < NL) { asm volatile("lop3.b32 %0, %0, %1, %2, 0xEA;" : "+r"(u[nl & 7]) : "r"(0xffffffc0u), "r"(u[(nl + 1) & 7])); ++nl; }
            else if (nm < NM) { asm volatile("min.u32 %0, %0, %1;" : "+r"(u[nm & 7]) : "r"(u[(nm + 3) & 7])); ++nm; }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 16; ++j) s += f[j];
#pragma unroll
    for (int j = 0; j < 8; ++j) s += __uint_as_float((unsigned)p[j]) + __uint_as_float((unsigned)(p[j] >> 32)) + __uint_as_float(u[j]);
#pragma unroll
    for (int j = 0; j < 4; ++j) s += x[j];
    if (s == -12345.678f) out[0] = s;
}

template <int NF, int NP, int NL, int NM, int NX>
void run() {
    float* d; cudaMalloc(&d, 4);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    int clk; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
    const int blocks = sms * 8, iters = 4096;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 4; ++rep) {
        cudaEventRecord(e0); k<NF, NP, NL, NM, NX><<<blocks, 256>>>(d, iters, 1.0f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep) best = best < ms ? best : ms;
    }
    double cyc = best * 1e-3 * clk * 1e3 / (16.0 * iters);   // 16 warps per scheduler
    printf("FFMA %2d FFMA2 %2d LOP3 %2d MIN %2d MUFU %2d : %7.2f cycles/body  slots %3d  slots/clk %.2f  fma-flops/clk/SMSP %.1f\n", NF, NP, NL, NM, NX, cyc,
           NF + NP + NL + NM + NX, (NF + NP + NL + NM + NX) / cyc, (NF + 2 * NP) * 64.0 / cyc);
    cudaFree(d);
}

int main() {
    run<32, 0, 0, 0, 0>();
    run<0, 16, 0, 0, 0>();
    run<0, 0, 32, 0, 0>();
    run<0, 0, 0, 32, 0>();
    run<0, 0, 16, 16, 0>();
    run<0, 0, 0, 0, 8>();
    run<32, 0, 8, 8, 0>();
    run<32, 0, 16, 16, 0>();
    run<32, 0, 4, 4, 0>();
    run<0, 16, 8, 8, 0>();
    run<0, 16, 16, 16, 0>();
    run<0, 16, 4, 4, 0>();
    run<32, 0, 8, 8, 4>();
    run<0, 16, 8, 8, 4>();
    run<16, 8, 8, 8, 4>();
    return 0;
}
