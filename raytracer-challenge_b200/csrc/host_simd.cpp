// Host side of rtc_render's delivery step: f32 -> f64 widening of staged frame rows into the caller's
// Canvas.pixels buffer (src/canvas.rs:17-22 is Vec<Color>, three f64 per pixel).  The conversion is exact.
// With AVX2 the destination is written with streaming stores: the caller's buffer is write-only here, so
// reading its lines first (write-allocate) would add a third of the memory traffic — measured on the
// bench box, 24.9 M elements with 12 threads: 3.1 ms plain, 1.8 ms streamed (tools/micro/widen_bench.cpp).
#include <immintrin.h>
#include <cstddef>
#include <cstdint>

namespace {
void widen_plain(double* d, const float* s, size_t n) {
    for (size_t k = 0; k < n; ++k) d[k] = (double)s[k];
}
__attribute__((target("avx2"))) void widen_avx2_stream(double* d, const float* s, size_t n) {
    size_t k = 0;
    while (k < n && (reinterpret_cast<uintptr_t>(d + k) & 31u)) { d[k] = (double)s[k]; ++k; }
    for (; k + 8 <= n; k += 8) {
        const __m256 v = _mm256_loadu_ps(s + k);
        _mm256_stream_pd(d + k, _mm256_cvtps_pd(_mm256_castps256_ps128(v)));
        _mm256_stream_pd(d + k + 4, _mm256_cvtps_pd(_mm256_extractf128_ps(v, 1)));
    }
    for (; k < n; ++k) d[k] = (double)s[k];
    _mm_sfence();
}
}  // namespace

extern "C" void rtc_host_widen_f32_f64(double* dst, const float* src, size_t n) {
    static const bool have_avx2 = __builtin_cpu_supports("avx2");
    if (have_avx2) widen_avx2_stream(dst, src, n);
    else widen_plain(dst, src, n);
}
