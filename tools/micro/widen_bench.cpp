// Host-side f32 -> f64 widening of one 4K frame (24.9 M floats): plain loop vs AVX2 with streaming stores,
// T threads.  g++ -O3 -pthread -o widen_bench widen_bench.cpp ; ./widen_bench [threads]
#include <immintrin.h>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>
static void widen_plain(double* d, const float* s, size_t n) { for (size_t k = 0; k < n; ++k) d[k] = (double)s[k]; }
__attribute__((target("avx2"))) static void widen_nt(double* d, const float* s, size_t n) {
    size_t k = 0;
    while (k < n && ((uintptr_t)(d + k) & 31)) { d[k] = (double)s[k]; ++k; }
    for (; k + 8 <= n; k += 8) {
        __m256 v = _mm256_loadu_ps(s + k);
        _mm256_stream_pd(d + k, _mm256_cvtps_pd(_mm256_castps256_ps128(v)));
        _mm256_stream_pd(d + k + 4, _mm256_cvtps_pd(_mm256_extractf128_ps(v, 1)));
    }
    for (; k < n; ++k) d[k] = (double)s[k];
    _mm_sfence();
}
__attribute__((target("avx512f,avx512dq"))) static void widen_nt512(double* d, const float* s, size_t n) {
    size_t k = 0;
    while (k < n && ((uintptr_t)(d + k) & 63)) { d[k] = (double)s[k]; ++k; }
    for (; k + 16 <= n; k += 16) {   // two full 64-byte lines per trip
        __m512 v = _mm512_loadu_ps(s + k);
        _mm512_stream_pd(d + k, _mm512_cvtps_pd(_mm512_castps512_ps256(v)));
        _mm512_stream_pd(d + k + 8, _mm512_cvtps_pd(_mm512_extractf32x8_ps(v, 1)));
    }
    for (; k < n; ++k) d[k] = (double)s[k];
    _mm_sfence();
}
int main(int argc, char** argv) {
    const size_t n = 3840ull * 2160 * 3;
    float* s = (float*)aligned_alloc(64, n * 4);
    double* d = (double*)aligned_alloc(64, n * 8);
    memset(s, 1, n * 4); memset(d, 0, n * 8);
    for (int T : {1, 4, 8, 12, 15, 16}) {
        for (int mode = 0; mode < 4; ++mode) {
            if (mode == 3 && !__builtin_cpu_supports("avx512dq")) continue;
            double best = 1e9;
            for (int rep = 0; rep < 5; ++rep) {
                auto t0 = std::chrono::steady_clock::now();
                std::vector<std::thread> th;
                for (int t = 0; t < T; ++t) th.emplace_back([=] {
                    size_t b = n * t / T, e = n * (t + 1) / T;
                    if (mode == 0) widen_plain(d + b, s + b, e - b);
                    else if (mode == 1) widen_nt(d + b, s + b, e - b);
                    else if (mode == 3) widen_nt512(d + b, s + b, e - b);
                    else memcpy((char*)d + 4 * b, (char*)s + 4 * b, 4 * (e - b));
                });
                for (auto& x : th) x.join();
                best = std::min(best, std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count());
            }
            printf("threads %2d %-22s %7.2f ms\n", T, mode == 0 ? "widen plain" : mode == 1 ? "widen avx2 + nt stores" : mode == 3 ? "widen avx512 + nt stores" : "memcpy 100 MB", best);
        }
    }
    return 0;
}
