// Micro-benchmark: FP32 non-tensor throughput of FFMA (scalar) vs FFMA2 (packed f32x2) on sm_100a,
// alone and mixed with ALU-pipe instructions (does FFMA2 free issue slots for other pipes?).
// build: nvcc -O3 -gencode arch=compute_100a,code=sm_100a -o ffma2_peak ffma2_peak.cu
#include <cstdio>
#include <cuda_runtime.h>

template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed) {
    float2 a[8];
    unsigned u[4] = {threadIdx.x, threadIdx.x * 3u, threadIdx.x * 5u, threadIdx.x * 7u};
#pragma unroll
    for (int j = 0; j < 8; ++j) a[j] = make_float2(seed + threadIdx.x + j, seed - j);
    const float2 m = make_float2(0.999f, 0.998f), c = make_float2(0.001f, 0.002f);
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int r = 0; r < 8; ++r) {
#pragma unroll
            for (int j = 0; j < 8; ++j) {
                if (MODE == 0 || MODE == 3) { a[j].x = fmaf(a[j].x, m.x, c.x); a[j].y = fmaf(a[j].y, m.y, c.y); }      // 2 FFMA
                else a[j] = __ffma2_rn(a[j], m, c);                                                      // 1 FFMA2
            }
            if (MODE == 2 || MODE == 3) {  // + ALU-pipe work: 8 LOP3/IADD per 8 FFMA2 (MODE 2) or per 16 FFMA (MODE 3 = scalar)
#pragma unroll
                for (int q = 0; q < 4; ++q) { u[q] = (u[q] ^ 0x9e3779b9u) + (u[(q + 1) & 3] & 0xfffffff0u); u[q] = min(u[q], u[(q + 2) & 3] | 5u); u[q] = __funnelshift_l(u[q], u[(q + 3) & 3], 7); u[q] ^= u[(q + 1) & 3] >> 3; }
            }
        }
    }
    float s = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) s += a[j].x + a[j].y;
    if (s == -12345.678f) out[0] = s;
    if (iters < 0) ((unsigned*)out)[threadIdx.x] = u[0] ^ u[1] ^ u[2] ^ u[3];
}

template <int MODE>
double run(const char* name, double fma_per_thread_iter) {
    float* d; cudaMalloc(&d, 4);
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int blocks = sms * 8, iters = 2048;
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0); k<MODE><<<blocks, 256>>>(d, iters, 1.0f); cudaEventRecord(e1); cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1); if (rep) best = best < ms ? best : ms;
    }
    double flops = 2.0 * fma_per_thread_iter * iters * 256.0 * blocks;
    double tf = flops / (best * 1e-3) / 1e12;
    printf("%-48s %8.3f ms  %7.2f TFLOP/s\n", name, best, tf);
    cudaFree(d); return tf;
}

int main() {
    run<0>("FFMA  scalar, 16 independent chains", 8 * 8 * 2);
    run<1>("FFMA2 packed, 8 independent chains x2", 8 * 8 * 2);
    run<2>("FFMA2 + ALU mix (8 FFMA2 : ~N ALU)", 8 * 8 * 2);
    run<3>("FFMA scalar + ALU mix (16 FFMA : ~N ALU)", 8 * 8 * 2);
    return 0;
}
