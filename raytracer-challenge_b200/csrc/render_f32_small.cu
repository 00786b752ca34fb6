// F32 FAST kernel, small-scene variant (<= 16 planes, <= 48 similarity-transformed spheres, no
// genuinely affine sphere, light outside every sphere): the path configs C1/C2/C3/C5 of BASELINE.json take.  Same algorithm as render_f32.cu —
// ray_for_pixel, World::intersect over every shape, hit selection, prepare_computations, Phong
// lighting with a shadow ray, reflection + refraction with Schlick, as path weights
// (src/camera.rs:64-114, src/shape.rs:677-781, src/material.rs:319-361) — but built so that the
// brute-force loops issue almost nothing except FP32 math:
//   * the intersect streams travel as a __grid_constant__ kernel parameter, i.e. they live in
//     constant bank 0, and the loops are fully unrolled in groups of two with a warp-uniform exit
//     per group: coefficients arrive through LDC/LDCU at compile-time offsets (uniform registers feed
//     the FFMA/FADD directly) — no address arithmetic, no loop counter, stream positions are immediates;
//   * hit selection is ONE unsigned min per candidate root: for t >= 0 the IEEE bit pattern orders
//     like the value, negative and NaN roots compare above +inf, so
//         best = min(best, (bits(t) & ~mask) | stream_position)        (one LOP3 + half a VIMNMX3)
//     keeps the nearest non-negative root together with its shape in a single register (the low
//     6 mantissa bits carry the position: t is truncated by <= 2^-17 relative); there is no branch
//     on the discriminant (sqrt of a negative discriminant is NaN, which loses the min);
//   * shadow rays are any-hit tests reduced to "min over shapes of g <= 0" (SURVEY.md A4);
//   * persistent warps, one pixel per lane, lanes refill from the warp's 8x4 tile by ballot rank;
//     pending refraction rays sit on a small per-lane stack (north-star item 2).
// EPSILON: kEps32 = 1e-4 instead of the reference's 1e-8 (below f32 resolution).
#include "rtc_common.h"

namespace rtc {
namespace {

constexpr float kEps32 = 1e-4f;
constexpr int kThreads = 128;
constexpr int kTileX = 8, kTileY = 4;
constexpr int kBatch = kTileX * kTileY;
constexpr int kMaxStack = RTC_MAX_DEPTH;
constexpr unsigned kInfBits = 0x7f800000u;
constexpr int kSubOutside = 1 << 20;   // sub-sample cursor of a pixel outside the rect: nothing to trace

struct F3 { float x, y, z; };
__device__ __forceinline__ F3 mk(float x, float y, float z) { F3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ F3 operator-(F3 a, F3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ F3 operator+(F3 a, F3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ F3 operator*(F3 a, float m) { return mk(a.x * m, a.y * m, a.z * m); }
__device__ __forceinline__ F3 operator*(F3 a, F3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float dot(F3 a, F3 b) { return fmaf(a.z, b.z, fmaf(a.y, b.y, a.x * b.x)); }
__device__ __forceinline__ F3 fma3(F3 a, float m, F3 c) { return mk(fmaf(a.x, m, c.x), fmaf(a.y, m, c.y), fmaf(a.z, m, c.z)); }
__device__ __forceinline__ F3 xyz(float4 v) { return mk(v.x, v.y, v.z); }
__device__ __forceinline__ float row4(float4 r, F3 p) { return fmaf(r.z, p.z, fmaf(r.y, p.y, fmaf(r.x, p.x, r.w))); }
__device__ __forceinline__ float row3(float4 r, F3 p) { return fmaf(r.z, p.z, fmaf(r.y, p.y, r.x * p.x)); }
__device__ __forceinline__ float rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }
__device__ __forceinline__ float sqrt_nan(float x) { float r; asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }  // NaN for x < 0
__device__ __forceinline__ bool is_even(float v) { float h = v * 0.5f; return h == floorf(h); }

// Stream positions are fixed at compile time: planes occupy [0, 16), spheres [16, 64) — so the
// position a candidate carries in the low kIdxBits bits of its key is an immediate.
constexpr int kIdxBits = 6;
constexpr unsigned kMask = (1u << kIdxBits) - 1u;

// key = (bits(t) & ~kMask) | pos as ONE LOP3: the mask comes in a register so that the position can be
// the instruction's immediate (with both as immediates ptxas needs two LOP3s; ALU-pipe instructions
// dispatch at half rate on sm_100a, so each one costs as much as two FFMAs: tools/micro/issue_mix.cu).
template <unsigned kPos>
__device__ __forceinline__ unsigned key_of(float t, unsigned mask_reg) {
    unsigned r;
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(r) : "r"(__float_as_uint(t)), "r"(mask_reg), "n"(kPos));
    return r;
}

// Group dispatch.  Shapes are tested in groups of two; the group count is warp-uniform.  Default: a
// chain with one uniform compare + branch per group.  RTC_F32S_DUFF: jump once into a fall-through
// switch (groups run from the last to the first; min / top-2 do not depend on the order).
template <int V> struct IntC { static constexpr int value = V; };
#ifdef RTC_F32S_DUFF
#define RTC_DUFF_CASE(k) case (k) + 1: if constexpr ((k) < N) f(IntC<((k) < N ? (k) : 0)>{});
template <int N, class F>
__device__ __forceinline__ void for_groups(int n, F&& f) {
    switch (n) {
        RTC_DUFF_CASE(23) RTC_DUFF_CASE(22) RTC_DUFF_CASE(21) RTC_DUFF_CASE(20) RTC_DUFF_CASE(19) RTC_DUFF_CASE(18)
        RTC_DUFF_CASE(17) RTC_DUFF_CASE(16) RTC_DUFF_CASE(15) RTC_DUFF_CASE(14) RTC_DUFF_CASE(13) RTC_DUFF_CASE(12)
        RTC_DUFF_CASE(11) RTC_DUFF_CASE(10) RTC_DUFF_CASE(9) RTC_DUFF_CASE(8) RTC_DUFF_CASE(7) RTC_DUFF_CASE(6)
        RTC_DUFF_CASE(5) RTC_DUFF_CASE(4) RTC_DUFF_CASE(3) RTC_DUFF_CASE(2) RTC_DUFF_CASE(1) RTC_DUFF_CASE(0)
        default: break;
    }
}
#else
template <int N, int G = 0, class F>
__device__ __forceinline__ void for_groups(int n, F&& f) {
    if constexpr (G < N) {
        if (G >= n) return;
        f(IntC<G>{});
        for_groups<N, G + 1>(n, f);
    }
}
#endif

// One sphere against a unit-length radiance ray (Sphere::intersect_local, src/shape.rs:361-375, with
// a = d.d = 1 and b halved).  Only one root can be the hit: t1 when the origin is outside (c > 0),
// t2 when it is inside (c < 0; then t1 < 0 < t2)  ->  t = -b - copysign(s, c).
template <unsigned kPos>
__device__ __forceinline__ unsigned sphere_key(const float4 sp, F3 o, F3 d, unsigned mask_reg) {
    F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
    float b = dot(d, L);
    float c = fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w)));
    float s = sqrt_nan(fmaf(b, b, -c));                   // NaN when the ray misses: loses the min
    unsigned ss;                                          // (c & 0x80000000) | s
    asm("lop3.b32 %0, %1, %2, %3, 0xEA;" : "=r"(ss) : "r"(__float_as_uint(c)), "n"(0x80000000), "r"(__float_as_uint(s)));
    return key_of<kPos>(-b - __uint_as_float(ss), mask_reg);
}
template <unsigned kPos>
__device__ __forceinline__ unsigned plane_key(const float4 pl, F3 o, F3 d, unsigned mask_reg) {
    return key_of<kPos>(-row4(pl, o) * rcp(row3(pl, d)), mask_reg);   // Plane::intersect_local, src/shape.rs:462-471; parallel: +-inf / NaN
}

// World::intersect + Intersections::get_hit for a unit-length ray: nearest root in [0, +inf).
// Stream positions are compile-time constants (plane i -> i, sphere i -> 16 + i).  Returns the packed
// key; >= kInfBits = no hit.
__device__ __forceinline__ unsigned intersect_small(const SmallParams& P, F3 o, F3 d, unsigned mask_reg) {
    unsigned best = kInfBits;
    for_groups<kSmallMaxPlanes / 2>(P.npg, [&](auto gc) {
        constexpr int g = decltype(gc)::value;
        best = min(best, min(plane_key<2 * g>(P.planes[2 * g], o, d, mask_reg), plane_key<2 * g + 1>(P.planes[2 * g + 1], o, d, mask_reg)));
    });
    for_groups<kSmallMaxUsph / 2>(P.nug, [&](auto gc) {
        constexpr int g = decltype(gc)::value;
        best = min(best, min(sphere_key<kSmallMaxPlanes + 2 * g>(P.usph[2 * g], o, d, mask_reg),
                             sphere_key<kSmallMaxPlanes + 2 * g + 1>(P.usph[2 * g + 1], o, d, mask_reg)));
    });
    // an odd count's last shape: the last slot of its array (a padded pair would cost a whole second test per ray)
    if (P.np_odd) best = min(best, plane_key<kSmallMaxPlanes - 1>(P.planes[kSmallMaxPlanes - 1], o, d, mask_reg));
    if (P.nu_odd) best = min(best, sphere_key<kSmallMaxPlanes + kSmallMaxUsph - 1>(P.usph[kSmallMaxUsph - 1], o, d, mask_reg));
    return best;
}

// World::is_shadowed_by_light (src/shape.rs:716-727) as an any-hit test over the segment
// o -> light (d = light - o): some shape has a root in [0, 1)  <=>  min over shapes of g <= 0.
//   plane : f(0) f(1) <= 0; f(1) depends on the light only and is folded into the coefficients
//   sphere: f(t) = a t^2 + 2 b t + f0 is convex, its minimum over [0,1] is at clamp(-b/a, 0, 1)
__device__ __forceinline__ bool shadow_small(const SmallParams& P, F3 o, F3 d) {
    float g = 1.0f;
    for_groups<kSmallMaxPlanes / 2>(P.npg, [&](auto gc) {
        constexpr int k = decltype(gc)::value;
        g = fminf(g, fminf(row4(P.splanes[2 * k], o), row4(P.splanes[2 * k + 1], o)));
    });
    const float a = dot(d, d);
    const float nhia = -0.5f * rcp(a);
    const F3 d2 = mk(d.x + d.x, d.y + d.y, d.z + d.z);
    for_groups<kSmallMaxUsph / 2>(P.nug, [&](auto gc) {
        constexpr int k = decltype(gc)::value;
        float fm[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const float4 sp = P.usph_s[2 * k + j];
            F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
            float f0 = fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w)));
            float b2 = dot(d2, L);                        // f(t) = a t^2 + b2 t + f0
            float tc = __saturatef(b2 * nhia);
            fm[j] = fmaf(tc, fmaf(a, tc, b2), f0);
        }
        g = fminf(g, fminf(fm[0], fm[1]));
    });
    if (P.np_odd) g = fminf(g, row4(P.splanes[kSmallMaxPlanes - 1], o));
    if (P.nu_odd) {
        const float4 sp = P.usph_s[kSmallMaxUsph - 1];
        F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
        float f0 = fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w)));
        float b2 = dot(d2, L);
        float tc = __saturatef(b2 * nhia);
        g = fminf(g, fmaf(tc, fmaf(a, tc, b2), f0));
    }
    return g <= 0.f;
}

// Intersection::compute_refractive (src/shape.rs:115-141) in its list-free form (SURVEY.md A6;
// world_ids are unique on this path): a shape is a container iff it has exactly one negative-t
// entry (plane: t < 0; sphere: origin inside), containers are ordered by (t, stream position), and
//   n1 = ri(last container)            n2 = hit is a container ? ri(last container other than it) : ri(hit)
// "Latest negative t" is the NEAREST hit of the reversed ray (o, -d), so the same unsigned-min keys as
// in intersect_small apply with t' = -t: a plane is a container when t' > 0, a sphere when the origin
// is inside it (c < 0; then t' = b + sqrt(b^2 - c) > 0) — for c >= 0 the sign bit of ~c is OR-ed into
// the key, which then loses the min like a negative root.  Ties go to the later stream position
// (the low bits carry kMask - pos).  Two smallest keys = last container and the one before it.
struct Top2 { unsigned a, b; };   // smallest, second smallest
__device__ __forceinline__ Top2 offer2(Top2 t, unsigned k0, unsigned k1) {
    const unsigned lo = min(k0, k1), hi = max(k0, k1);
    Top2 r;
    r.b = min(t.b, min(hi, max(t.a, lo)));
    r.a = min(t.a, lo);
    return r;
}
template <unsigned kPos>
__device__ __forceinline__ unsigned plane_back_key(const float4 pl, F3 o, F3 d, unsigned mask_reg) {
    return key_of<kMask - kPos>(row4(pl, o) * rcp(row3(pl, d)), mask_reg);
}
template <unsigned kPos>
__device__ __forceinline__ unsigned sphere_back_key(const float4 sp, F3 o, F3 d, unsigned mask_reg) {
    F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
    float b = dot(d, L);
    float c = fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w)));
    float tb = b + sqrt_nan(fmaf(b, b, -c));
    unsigned u;                                           // (~bits(c) & 0x80000000) | bits(tb)
    asm("lop3.b32 %0, %1, %2, %3, 0xAE;" : "=r"(u) : "r"(__float_as_uint(c)), "n"(0x80000000), "r"(__float_as_uint(tb)));
    return key_of<kMask - kPos>(__uint_as_float(u), mask_reg);
}
__device__ __forceinline__ Top2 containers_small(const SmallParams& P, F3 o, F3 d, unsigned mask_reg) {
    Top2 t; t.a = t.b = 0xffffffffu;
    for_groups<kSmallMaxPlanes / 2>(P.npg, [&](auto gc) {
        constexpr int g = decltype(gc)::value;
        t = offer2(t, plane_back_key<2 * g>(P.planes_d[2 * g], o, d, mask_reg), plane_back_key<2 * g + 1>(P.planes_d[2 * g + 1], o, d, mask_reg));
    });
    for_groups<kSmallMaxUsph / 2>(P.nug, [&](auto gc) {
        constexpr int g = decltype(gc)::value;
        t = offer2(t, sphere_back_key<kSmallMaxPlanes + 2 * g>(P.usph[2 * g], o, d, mask_reg),
                   sphere_back_key<kSmallMaxPlanes + 2 * g + 1>(P.usph[2 * g + 1], o, d, mask_reg));
    });
    if (P.np_odd) t = offer2(t, plane_back_key<kSmallMaxPlanes - 1>(P.planes_d[kSmallMaxPlanes - 1], o, d, mask_reg), 0xffffffffu);
    if (P.nu_odd) t = offer2(t, sphere_back_key<kSmallMaxPlanes + kSmallMaxUsph - 1>(P.usph[kSmallMaxUsph - 1], o, d, mask_reg), 0xffffffffu);
    return t;
}

// Pattern::pattern_at for the rarely used kinds (src/material.rs:67-69,133-135,163-170,233-241)
__device__ __noinline__ F3 pattern_rare(int kind, F3 a, F3 b, F3 p) {
    switch (kind) {
        case RTC_PATTERN_TEST: return p;
        case RTC_PATTERN_GRADIENT: return fma3(b - a, p.x - floorf(p.x), a);
        case RTC_PATTERN_RING: return is_even(floorf(sqrtf(fmaf(p.x, p.x, p.y * p.y)))) ? a : b;
        default: return (fabsf(p.x - floorf(p.x)) < 0.01f || fabsf(p.z - floorf(p.z)) < 0.01f) ? b : a;
    }
}

// Shading records are read through explicit 32-bit shared-window addresses: indexing the generic
// `extern __shared__` pointer makes ptxas rebuild the window base (S2R CgaCtaId + MOV + LEA) every
// loop iteration; a base computed once plus `hpos * sizeof(SmallRec32)` is one IMAD.
__device__ __forceinline__ float4 lds4(unsigned addr) {
    float4 v;
    asm("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr));
    return v;
}
__device__ __forceinline__ float lds1(unsigned addr) {
    float v;
    asm("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(addr));
    return v;
}
constexpr unsigned kRecBytes = (unsigned)sizeof(SmallRec32);
constexpr unsigned kOffNrm = 0, kOffColAmb = 16, kOffDifSpec = 32, kOffTrRi = 48, kOffP0 = 64, kOffP1 = 80, kOffP2 = 96, kOffPa = 112, kOffPb = 128;
static_assert(sizeof(SmallRec32) == 144, "SmallRec32 layout");

#ifdef RTC_F32S_TRACE   // diagnostic build only (tools/trace_tail.py): per-warp start / tiles-exhausted / end times
__device__ unsigned long long g_trace[16384 * 6];   // per warp: t0, t_exhausted, t_end, iterations | per warp: last tile, its hand-out rank
__device__ __forceinline__ unsigned long long gtime() { unsigned long long t; asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t)); return t; }
#endif

// Blocks per SM (128 threads each).  Measured after the odd-slot change, C3 4K / 4K with samples = 4: 5 blocks 1.261 ms, 6
// blocks (74 registers) 1.268 / 6.03, 7 blocks (71 registers, no spills) 1.214 / 5.91, 8 blocks (64 registers, 36 B of
// spills) 1.217 / 5.78: 7 for the plain frame, 8 for the resample-pass kernels.
#ifndef RTC_F32S_MIN_BLOCKS
#define RTC_F32S_MIN_BLOCKS 7
#endif
#ifndef RTC_F32S_MIN_BLOCKS_AA
#define RTC_F32S_MIN_BLOCKS_AA 8
#endif

// kAux: the launch wants the work counters and/or the hit mask; without them (the plain frame) the
// per-iteration bookkeeping for both is compiled out.
// kAA: antialiasing_samples >= 2 (Camera::render_pixel + resample, src/camera.rs:84-114): the lane keeps the four fixed
// samples apart (local memory, touched once per sample), compares them with their mean and, where they disagree, goes on
// with `samples` more at counter-based offsets (include/rtc.h); rays carry weight 1 and the sum is divided at the end.
// Frames with samples 0 / 1 run the kAA = false code, unchanged.
template <int kFmt, bool kAux, bool kAA>
__global__ void __launch_bounds__(kThreads, kAA ? RTC_F32S_MIN_BLOCKS_AA : RTC_F32S_MIN_BLOCKS)
render_f32_small_kernel(const __grid_constant__ SmallParams P, void* __restrict__ out_rgb, uint8_t* __restrict__ out_hit,
                        Counters* __restrict__ counters, unsigned* __restrict__ work) {
    extern __shared__ float4 smem4[];
    const unsigned rbase = (unsigned)__cvta_generic_to_shared(smem4);
    {   // stage the shading records (indexed by stream position): coalesced 16-byte loads
        constexpr int kRec4 = (int)(sizeof(SmallRec32) / 16);
        const float4* g_rec = reinterpret_cast<const float4*>(P.rec);
        for (int i = threadIdx.x; i < kRec4 * 2 * P.npg; i += kThreads) smem4[i] = g_rec[i];
        for (int i = threadIdx.x; i < kRec4 * 2 * P.nug; i += kThreads) smem4[kRec4 * kSmallMaxPlanes + i] = g_rec[kRec4 * kSmallMaxPlanes + i];
        if (P.np_odd) for (int i = threadIdx.x; i < kRec4; i += kThreads) smem4[kRec4 * (kSmallMaxPlanes - 1) + i] = g_rec[kRec4 * (kSmallMaxPlanes - 1) + i];
        if (P.nu_odd) for (int i = threadIdx.x; i < kRec4; i += kThreads) smem4[kRec4 * (kSmallMaxShapes - 1) + i] = g_rec[kRec4 * (kSmallMaxShapes - 1) + i];
        if (P.wait_flag && threadIdx.x == 0) {   // back-pressure: the destination frame is free once the consumer says so
            unsigned long long t0, t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
            for (;;) {
                unsigned v;
                asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(P.wait_flag) : "memory");
                if ((int)(v - P.wait_value) >= 0) break;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > P.wait_timeout_ns) { atomicMin(P.wait_status, 1u); break; }   // recorded, not trapped (rtc_wait_result)
                __nanosleep(200);
            }
        }
        __syncthreads();
    }
    const F3 light_pos = mk(P.light_pos[0], P.light_pos[1], P.light_pos[2]);
    const F3 light_int = mk(P.light_int[0], P.light_int[1], P.light_int[2]);
    unsigned mask_reg;   // ~kMask in a register (see key_of)
    asm volatile("mov.u32 %0, %1;" : "=r"(mask_reg) : "n"(~kMask));
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int n_sub = P.n_sub;
    const float w0 = (n_sub == 1 || kAA) ? 1.0f : 0.25f;
    const uint64_t aa_seed_mixed = kAA ? splitmix64((uint64_t)P.aa_seed) : 0ull;

    // warp-uniform tile cursor
    int wnext = kBatch;        // next pixel of the current tile
    int wtx = 0, wty = 0;      // tile origin (pixels / compact rows)
    bool exhausted = false;
    // per-lane path state
    bool has_ray = false, have = false, primary = false;
    F3 ro = mk(0.f, 0.f, 0.f), rd = mk(0.f, 0.f, 1.f);
    float rw = 0.f;
    int remaining = 0;
    int px = 0, pr = 0, py = 0, sub = 0;
    F3 acc = mk(0.f, 0.f, 0.f);
    float aa_s[12];                   // kAA: the four fixed samples of the pixel
    F3 aa_tot = mk(0.f, 0.f, 0.f);    // kAA: Color::average_over's running sums
    int aa_ntot = n_sub;              // samples this pixel takes: 4, or 4 + n_extra once the fixed four disagree
    unsigned hitbits = 0;
    float4 stk_a[kMaxStack], stk_b[kMaxStack];  // pending refraction rays: (o, weight), (d, remaining)
    int sp = 0;
    unsigned n_rad = 0, n_hits = 0, n_phits = 0;
#ifdef RTC_F32S_TRACE
    const unsigned long long tr_t0 = gtime();
    unsigned long long tr_t1 = 0;
    unsigned tr_iters = 0, tr_tail_iters = 0, tr_last_tile = 0, tr_last_rank = 0;
#endif

    for (;;) {
        // ------------------------------------------------------------ lanes without a ray
        if (!has_ray && sp == 0 && have) {
            if (kAA && sub < aa_ntot) {                       // a sample of the pixel is complete
                aa_tot = aa_tot + acc;
                if (sub < 4) { aa_s[3 * sub] = acc.x; aa_s[3 * sub + 1] = acc.y; aa_s[3 * sub + 2] = acc.z; }
                if (sub == 3) {                               // any(|sample - mean| > 0.01) -> resample (src/camera.rs:107-110)
                    const F3 mean = aa_tot * 0.25f;
                    bool any = false;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const F3 dl = mk(aa_s[3 * k], aa_s[3 * k + 1], aa_s[3 * k + 2]) - mean;
                        any = any || dot(dl, dl) > 1.0e-4f;
                    }
                    if (any) aa_ntot = 4 + P.n_extra;
                }
                acc = mk(0.f, 0.f, 0.f);
            }
            if (sub + 1 < (kAA ? aa_ntot : n_sub)) ++sub;
            else {                                            // Canvas::write_pixel (src/canvas.rs:43-46)
                if (kAA) acc = aa_tot * (1.0f / (float)aa_ntot);
                size_t p = (size_t)(P.dst_full ? py : (P.shard_count > 1 ? P.row0 + pr : pr)) * (size_t)P.hsize + (size_t)px;
                RTC_CHECK(px >= 0 && px < P.hsize && pr >= 0 && pr < P.shard_rows && py >= 0);
                if (kFmt == RTC_OUT_F64_RGB) {
                    double* o = (double*)out_rgb + 3 * p;
                    o[0] = (double)acc.x; o[1] = (double)acc.y; o[2] = (double)acc.z;
                } else if (kFmt == RTC_OUT_F32_RGB) {
                    float* o = (float*)out_rgb + 3 * p;
                    o[0] = acc.x; o[1] = acc.y; o[2] = acc.z;
                } else {                                      // Color::scale, src/color.rs:100-114
                    uint8_t* o = (uint8_t*)out_rgb + 3 * p;
                    float c3[3] = {acc.x, acc.y, acc.z};
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        float s = c3[k] * 255.0f;
                        int q = s != s ? 0 : (s >= 2147483520.0f ? 2147483647 : (s <= -2147483648.0f ? (-2147483647 - 1) : (int)s));
                        o[k] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
                    }
                }
                if (kAux && out_hit) out_hit[p] = (uint8_t)hitbits;
                have = false;
            }
        }
        // lanes without a pixel claim the next ones of the warp's tile (ballot + popc rank)
        for (;;) {
            const bool need = !has_ray && sp == 0 && !have;
            const unsigned m = __ballot_sync(full, need);
            if (m == 0u || exhausted) break;
            if (wnext >= kBatch) {
                unsigned t = 0;
                if (lane == 0) t = atomicAdd(work, 1u);
                t = __shfl_sync(full, t, 0);
                if (t >= (unsigned)P.total_tiles) { exhausted = true; break; }
#ifdef RTC_F32S_TRACE
                tr_last_rank = t;
#endif
                if (P.order) t = P.order[t];                  // heavy tiles first (classify_tiles_small_kernel)
#ifdef RTC_F32S_TRACE
                tr_last_tile = t;
#endif
                RTC_CHECK(t < (unsigned)P.total_tiles);
                const int ty = (int)(t / (unsigned)P.tiles_x);
                wtx = ((int)t - ty * P.tiles_x) * kTileX;
                wty = ty * kTileY;
                wnext = 0;
            }
            const int rank = __popc(m & lt_mask);
            const int take = min(__popc(m), kBatch - wnext);
            if (need && rank < take) {
                const int k = wnext + rank;
                px = wtx + (k & (kTileX - 1));
                pr = wty + (k >> 3);
                if (px < P.hsize && pr < P.shard_rows) {
                    py = P.shard_count > 1 ? shard_row_to_frame_row(P.row0 + pr, P.tile_rows, P.shard_index, P.shard_count) : P.row0 + pr;
                    have = true; sub = 0; acc = mk(0.f, 0.f, 0.f); hitbits = 0;
                    if (kAA) { aa_tot = mk(0.f, 0.f, 0.f); aa_ntot = n_sub; }
                    if (!(px < P.xlim && py < P.ylim)) sub = kSubOutside;   // outside the rect: stays black
                }
            }
            wnext += take;
        }
        if (!has_ray) {
            if (sp > 0) {                                     // next pending refraction ray of this pixel
                --sp;
                RTC_CHECK(sp >= 0 && sp < kMaxStack);
                const float4 a = stk_a[sp], b = stk_b[sp];
                ro = xyz(a); rw = a.w; rd = xyz(b); remaining = __float_as_int(b.w);
                primary = false; has_ray = true;
            } else if (have && sub < (kAA ? aa_ntot : n_sub)) {   // Camera::ray_for_pixel_offset (src/camera.rs:64-76)
                float xo = n_sub == 1 ? 0.5f : ((sub & 1) ? 0.75f : 0.25f);
                float yo = n_sub == 1 ? 0.5f : ((sub & 2) ? 0.75f : 0.25f);
                if (kAA && sub >= 4) {                        // Camera::resample's offsets (src/camera.rs:84-92; include/rtc.h)
                    const uint64_t pixel_index = (uint64_t)py * (uint64_t)P.hsize + (uint64_t)px;
                    xo = (float)(unsigned)(resample_bits(aa_seed_mixed, pixel_index, (unsigned)(sub - 4), 0u) >> 40) * 5.9604644775390625e-8f;
                    yo = (float)(unsigned)(resample_bits(aa_seed_mixed, pixel_index, (unsigned)(sub - 4), 1u) >> 40) * 5.9604644775390625e-8f;
                }
                const float fx = (float)px + xo;
                const float fy = (float)py + yo;
                F3 dir = mk(fmaf(P.cam_dx[0], fx, fmaf(P.cam_dy[0], fy, P.cam_d0[0])),
                            fmaf(P.cam_dx[1], fx, fmaf(P.cam_dy[1], fy, P.cam_d0[1])),
                            fmaf(P.cam_dx[2], fx, fmaf(P.cam_dy[2], fy, P.cam_d0[2])));
                rd = dir * rsqrtf(dot(dir, dir));
                ro = mk(P.cam_o[0], P.cam_o[1], P.cam_o[2]);
                rw = w0; remaining = P.max_depth; primary = true; has_ray = true;
            }
        }
        if (__all_sync(full, !has_ray && !have && exhausted)) break;
#ifdef RTC_F32S_TRACE
        if (exhausted && tr_t1 == 0) tr_t1 = gtime();
        ++tr_iters; if (exhausted) ++tr_tail_iters;
#endif

        // ------------------------------------------------------------ phase A: World::intersect (all lanes, idle ones trace a dummy)
        const unsigned key = intersect_small(P, ro, rd, mask_reg);
        const bool hit = has_ray && key < kInfBits;
        if (kAux) {
            n_rad += has_ray ? 1u : 0u;
            n_hits += hit ? 1u : 0u;
            if (hit && primary) { if (!kAA || sub < 4) hitbits |= 1u << sub; ++n_phits; }
        }
        has_ray = hit;                                        // miss: World::color_at -> black (src/shape.rs:708)

        // ------------------------------------------------------------ phase B: prepare_computations (src/shape.rs:75-96,144-152)
        const int hpos = hit ? (int)(key & kMask) : 0;
        RTC_CHECK(!hit || hpos < 2 * P.npg || (P.np_odd && hpos == kSmallMaxPlanes - 1) || (hpos >= kSmallMaxPlanes && hpos < kSmallMaxPlanes + 2 * P.nug) ||
                  (P.nu_odd && hpos == kSmallMaxShapes - 1));
        const float t = __uint_as_float((key & ~kMask) | ((kMask + 1u) >> 1));   // centre of the truncation interval
        const unsigned rec = rbase + (unsigned)hpos * kRecBytes;
        const float4 tr = lds4(rec + kOffTrRi);
        F3 p = fma3(rd, hit ? t : 0.f, ro);                   // Ray::position
        F3 hn;
        {
            const float4 nr = lds4(rec + kOffNrm);            // plane: constant normal; sphere: (p - c) / r
            hn = mk(fmaf(p.x, nr.w, nr.x), fmaf(p.y, nr.w, nr.y), fmaf(p.z, nr.w, nr.z));
            // p carries t's truncation error (<= 2^-18 relative), so |hn| is only ~1e-4 from 1; the
            // reflected / refracted directions must stay unit length (the sphere test assumes a = 1)
            hn = hn * rsqrtf(dot(hn, hn));
        }
        float hn1 = 1.f, hn2 = tr.y;
        if (hit && tr.x != 0.f) {                             // Intersection::compute_refractive
            bool inside_glass = P.n_glass < 0;                // fast test unavailable -> general form
#pragma unroll
            for (int k = 0; k < kSmallMaxGlass; ++k) {
                if (k >= P.n_glass) break;
                const float4 sp4 = P.glass[k];
                F3 L = mk(ro.x - sp4.x, ro.y - sp4.y, ro.z - sp4.z);
                inside_glass |= fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp4.w))) < 0.f;
            }
            // origin inside no shape with ri != 1: every container has ri 1.0 -> n1 = 1, n2 = ri(hit)
            if (inside_glass) {
                const Top2 top = containers_small(P, ro, rd, mask_reg);
                const int p1 = top.a < kInfBits ? (int)(kMask - (top.a & kMask)) : -1;
                const int p2 = top.b < kInfBits ? (int)(kMask - (top.b & kMask)) : -1;
                // the hit shape is a container iff it is a sphere with the origin inside: |n(ro)| < 1
                const float4 nr = lds4(rec + kOffNrm);
                const F3 no = mk(fmaf(ro.x, nr.w, nr.x), fmaf(ro.y, nr.w, nr.y), fmaf(ro.z, nr.w, nr.z));
                const bool hit_is_container = hpos >= kSmallMaxPlanes && dot(no, no) < 1.f;
                hn1 = p1 >= 0 ? lds1(rbase + (unsigned)p1 * kRecBytes + kOffTrRi + 4) : 1.0f;
                if (p1 == hpos) hn2 = p2 >= 0 ? lds1(rbase + (unsigned)p2 * kRecBytes + kOffTrRi + 4) : 1.0f;
                else if (hit_is_container) hn2 = hn1;
            }
        }
        const float dn = dot(rd, hn);
        if (dn > 0.f) hn = mk(-hn.x, -hn.y, -hn.z);           // inside: normal . eyev < 0
        const float cos_i = fabsf(dn);                        // eyev . normal after the flip
        const F3 hp = fma3(hn, kEps32, p);                    // over_point

        // ------------------------------------------------------------ phase C: is_shadowed (all lanes)
        const F3 lv0 = light_pos - hp;
        const bool shadowed = shadow_small(P, hp, lv0);

        // ------------------------------------------------------------ phase D: shade_hit
        if (hit) {
            const float4 ca = lds4(rec + kOffColAmb), ds = lds4(rec + kOffDifSpec);
            const int pkind = __float_as_int(tr.z);
            F3 base = xyz(ca);
            if (pkind != RTC_PATTERN_NONE) {                  // Pattern::pattern_at_shape at over_point
                const F3 pp = mk(row4(lds4(rec + kOffP0), hp), row4(lds4(rec + kOffP1), hp), row4(lds4(rec + kOffP2), hp));
                const float4 pa = lds4(rec + kOffPa), pb = lds4(rec + kOffPb);
                if (pkind == RTC_PATTERN_STRIPE || pkind == RTC_PATTERN_CHECKER) {
                    // stripe: floor(x) % 2 ; checker: (floor x + floor y + floor z) % 2   (pb.w = 0 / 1)
                    const float s = fmaf(pb.w, floorf(pp.y) + floorf(pp.z), floorf(pp.x));
                    base = is_even(s) ? xyz(pa) : xyz(pb);
                } else base = pattern_rare(pkind, xyz(pa), xyz(pb), pp);
            }
            const F3 eff = base * light_int;                  // Material::lighting, src/material.rs:319-361
            F3 surface = eff * ca.w;
            if (!shadowed) {
                const F3 lv = lv0 * rsqrtf(dot(lv0, lv0));
                const float ldn = dot(lv, hn);
                if (ldn >= 0.f) {
                    surface = fma3(eff, ds.x * ldn, surface);
                    // reflect(-lightv, n) . eye  with eye = -rd:  (2 ldn n - lv) . (-rd)
                    const float rde = fmaf(2.f * ldn, cos_i, dot(lv, rd));
                    if (rde > 0.f) surface = fma3(light_int, ds.y * __powf(rde, ds.z), surface);
                }
            }
            acc = fma3(surface, rw, acc);
            // children (src/shape.rs:689-699, 729-781) as weighted rays
            const float k_refl = ds.w, k_tr = tr.x;
            float refl_w = rw * k_refl, refr_w = rw * k_tr;
            const float ratio = hn1 * rcp(hn2);
            const float sin2_t = ratio * ratio * fmaf(-cos_i, cos_i, 1.f);
            if (k_refl > 0.f && k_tr > 0.f) {                 // World::reflectance (Schlick), src/shape.rs:768-781
                float refl = 1.f;                             // total internal reflection
                float cs = cos_i;
                bool tir = false;
                if (hn1 > hn2) { if (sin2_t > 1.f) tir = true; else cs = sqrtf(1.f - sin2_t); }
                if (!tir) {
                    const float q = (hn1 - hn2) * rcp(hn1 + hn2), r0 = q * q, m = 1.f - cs, m2 = m * m;
                    refl = fmaf(1.f - r0, m2 * m2 * m, r0);
                }
                refl_w *= refl; refr_w *= 1.f - refl;
            }
            const bool want_refl = remaining != 0 && k_refl > 0.f;
            const bool want_refr = remaining != 0 && k_tr != 0.f && !(sin2_t > 1.f);   // refracted_color, :751-766
            F3 fo = mk(0.f, 0.f, 0.f), fd = mk(0.f, 0.f, 1.f);
            if (want_refr) {
                const float cos_t = sqrtf(1.f - sin2_t);
                fd = fma3(hn, ratio * cos_i - cos_t, rd * ratio);   // n (r cos_i - cos_t) - eyev r
                fo = fma3(hn, -2.f * kEps32, hp);                   // under_point
            }
            primary = false;
            --remaining;
            if (want_refl) {
                if (want_refr && sp < kMaxStack) {
                    stk_a[sp] = make_float4(fo.x, fo.y, fo.z, refr_w);
                    stk_b[sp] = make_float4(fd.x, fd.y, fd.z, __int_as_float(remaining));
                    ++sp;
                }
                rd = fma3(hn, 2.f * cos_i, rd);                     // reflectv = d.reflect(n); d . n = -cos_i
                ro = hp; rw = refl_w;
            } else if (want_refr) {
                ro = fo; rd = fd; rw = refr_w;
            } else {
                has_ray = false;
            }
        }
    }

#ifdef RTC_F32S_TRACE
    if (lane == 0) {
        const unsigned w = blockIdx.x * (kThreads / 32) + (threadIdx.x >> 5);
        if (w < 16384) { g_trace[4 * w] = tr_t0; g_trace[4 * w + 1] = tr_t1; g_trace[4 * w + 2] = gtime(); g_trace[4 * w + 3] = ((unsigned long long)tr_iters << 32) | tr_tail_iters;
                         g_trace[4 * 16384 + 2 * w] = tr_last_tile; g_trace[4 * 16384 + 2 * w + 1] = tr_last_rank; }
    }
#endif
    // The tile counter resets itself: the last block to finish zeroes work[0] (tiles handed out) and
    // work[1] (blocks done), so a launch needs no memset in front of it.
    // With a completion flag the block's pixels (possibly peer stores) are fenced at system scope before the block counts
    // itself done; the last block then fences again and release-stores the flag: whoever acquires the flag sees the frame.
    __syncthreads();
    if (threadIdx.x == 0) {
        if (P.done_flag) __threadfence_system(); else __threadfence();
        if (atomicAdd(work + 1, 1u) == gridDim.x - 1) {
            work[0] = 0u; work[1] = 0u; work[2] = 0u; work[3] = 0u; __threadfence();
            if (P.done_flag) {
                __threadfence_system();
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(P.done_flag), "r"(P.done_value) : "memory");
            }
        }
    }
    if (kAux && counters) {
        // every shaded hit casts exactly one shadow ray: rays = radiance rays + shaded hits
        const unsigned r = __reduce_add_sync(full, n_rad), h = __reduce_add_sync(full, n_hits), ph = __reduce_add_sync(full, n_phits);
        if (lane == 0) {
            atomicAdd(&counters->rays, (unsigned long long)r + h);
            atomicAdd(&counters->shadow_rays, (unsigned long long)h);
            atomicAdd(&counters->shaded_hits, (unsigned long long)h);
            atomicAdd(&counters->primary_hits, (unsigned long long)ph);
        }
    }
}

// Tile scheduling pre-pass.  The frame time of a persistent kernel ends with the longest ray chain that
// started last: a tile of glass pixels (up to 2^6 - 1 radiance rays per pixel) claimed at the very end
// runs ~40 loop iterations on an almost empty GPU.  One primary ray through the centre of every 8x4
// tile tells whether it sees a transparent surface; those tiles are handed out first
// (order[0 .. heavy)), the others after them (filled from the back).  It changes the order in which
// pixels are rendered, not what is rendered.  work[2] / work[3] count heavy / light tiles.
__global__ void __launch_bounds__(kThreads)
classify_tiles_small_kernel(const __grid_constant__ SmallParams P, unsigned* __restrict__ work) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    unsigned mask_reg;
    asm volatile("mov.u32 %0, %1;" : "=r"(mask_reg) : "n"(~kMask));
    const int t = blockIdx.x * kThreads + threadIdx.x;
    const bool valid = t < P.total_tiles;
    bool heavy = false;
    if (valid) {
        const int ty = t / P.tiles_x, tx = t - ty * P.tiles_x;
        const int px = min(tx * kTileX + kTileX / 2, P.hsize - 1);
        const int pr = min(ty * kTileY + kTileY / 2, P.shard_rows - 1);
        const int py = P.shard_count > 1 ? shard_row_to_frame_row(P.row0 + pr, P.tile_rows, P.shard_index, P.shard_count) : P.row0 + pr;
        const float fx = (float)px, fy = (float)py;          // the corner shared by the four centre pixels
        F3 dir = mk(fmaf(P.cam_dx[0], fx, fmaf(P.cam_dy[0], fy, P.cam_d0[0])),
                    fmaf(P.cam_dx[1], fx, fmaf(P.cam_dy[1], fy, P.cam_d0[1])),
                    fmaf(P.cam_dx[2], fx, fmaf(P.cam_dy[2], fy, P.cam_d0[2])));
        dir = dir * rsqrtf(dot(dir, dir));
        const unsigned key = intersect_small(P, mk(P.cam_o[0], P.cam_o[1], P.cam_o[2]), dir, mask_reg);
        if (key < kInfBits) heavy = P.rec[key & kMask].tr_ri.x != 0.f;
    }
    const unsigned mh = __ballot_sync(full, valid && heavy), ml = __ballot_sync(full, valid && !heavy);
    unsigned bh = 0, bl = 0;
    if (lane == 0) { if (mh) bh = atomicAdd(work + 2, (unsigned)__popc(mh)); if (ml) bl = atomicAdd(work + 3, (unsigned)__popc(ml)); }
    bh = __shfl_sync(full, bh, 0); bl = __shfl_sync(full, bl, 0);
    const unsigned below = (1u << lane) - 1u;
    if (valid && heavy) P.order[bh + __popc(mh & below)] = (unsigned)t;
    if (valid && !heavy) P.order[P.total_tiles - 1 - (int)(bl + __popc(ml & below))] = (unsigned)t;
}

}  // namespace

#ifdef RTC_F32S_TRACE
extern "C" int rtc_debug_read_trace(unsigned long long* dst, int n_words) {
    return (int)cudaMemcpyFromSymbol(dst, g_trace, sizeof(unsigned long long) * (size_t)n_words);
}
#endif

cudaError_t preload_render_f32_small() {
    cudaFuncAttributes at;
    cudaError_t e = cudaFuncGetAttributes(&at, classify_tiles_small_kernel);
    auto touch = [&](auto kernel) { if (e == cudaSuccess) e = cudaFuncGetAttributes(&at, kernel); };
#define RTC_TOUCHS(F) touch(render_f32_small_kernel<F, true, true>); touch(render_f32_small_kernel<F, false, true>); \
                      touch(render_f32_small_kernel<F, true, false>); touch(render_f32_small_kernel<F, false, false>);
    RTC_TOUCHS(RTC_OUT_F64_RGB) RTC_TOUCHS(RTC_OUT_F32_RGB) RTC_TOUCHS(RTC_OUT_U8_RGB)
#undef RTC_TOUCHS
    return e;
}

// ---- rtc_render_device_noalloc: scratch in a static device array --------------------------------------------------------
// Words 0..7: the persistent kernel's work counters (tiles handed out, blocks done, heavy / light tile counts); words
// 64.. : the tile hand-out order written by classify_tiles_small_kernel (one word per tile, all total_tiles of them
// rewritten by every call that uses them).  Every GPU that loads the library holds the array until the process exits; every
// call on a GPU shares it, so calls that overlap on that GPU race (include/rtc.h).  In host code the NAME of the array is
// a host-side stand-in, not a device address: the address a kernel may be given comes from cudaGetSymbolAddress, taken once
// per device with that device current.
static __device__ __align__(256) unsigned g_noalloc_scratch[64 + kNoAllocMaxTiles];
static unsigned* g_noalloc_ptr[64] = {};   // by device ordinal

cudaError_t noalloc_setup_current_device() {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (g_noalloc_ptr[dev]) return cudaSuccess;
    unsigned* ptr = nullptr;
    e = cudaGetSymbolAddress((void**)&ptr, g_noalloc_scratch);
    if (e != cudaSuccess) return e;
    if (!ptr) return cudaErrorInvalidSymbol;
    g_noalloc_ptr[dev] = ptr;
    return cudaSuccess;
}

// Same kernels, same launch shape as launch_render_f32_small; nothing here allocates.  The counter words are cleared on
// the stream first: the array keeps the last call's contents, and a call that was cut short would have left them non-zero.
cudaError_t launch_render_f32_small_noalloc(SmallParams P, bool tile_order, int out_format, void* d_rgb, uint8_t* d_hit, cudaStream_t stream) {
    if (P.total_tiles > kNoAllocMaxTiles) return cudaErrorInvalidValue;
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    if (!g_noalloc_ptr[dev] && (e = noalloc_setup_current_device()) != cudaSuccess) return e;
    unsigned* scratch = g_noalloc_ptr[dev];
    e = cudaMemsetAsync(scratch, 0, 64 * sizeof(unsigned), stream);
    if (e != cudaSuccess) return e;
    P.order = tile_order ? scratch + 64 : nullptr;
    return launch_render_f32_small(P, out_format, d_rgb, d_hit, nullptr, scratch, stream);
}

cudaError_t launch_render_f32_small(const SmallParams& P, int out_format, void* d_rgb, uint8_t* d_hit, Counters* d_counters,
                                    unsigned* d_work, cudaStream_t stream) {
    if (P.shard_rows <= 0) return cudaSuccess;
    // d_work[0..3] are zero between launches (the kernel resets them)
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const size_t smem = (size_t)(P.nu_odd ? kSmallMaxShapes : kSmallMaxPlanes + P.nu) * sizeof(SmallRec32) + 16;   // records by stream position
    const int warps = kThreads / 32;
    if (P.order) classify_tiles_small_kernel<<<(P.total_tiles + kThreads - 1) / kThreads, kThreads, 0, stream>>>(P, d_work);
    auto launch = [&](auto kernel) -> cudaError_t {
        int blocks_per_sm = 0;
        cudaError_t er = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kernel, kThreads, smem);
        if (er != cudaSuccess) return er;
        const int grid = max(1, min(sms * max(blocks_per_sm, 1), (P.total_tiles + warps - 1) / warps));
        kernel<<<grid, kThreads, smem, stream>>>(P, d_rgb, d_hit, d_counters, d_work);
        return cudaGetLastError();
    };
    const bool aux = d_counters != nullptr || d_hit != nullptr;
    auto pick = [&](auto fmt) -> cudaError_t {
        constexpr int kF = decltype(fmt)::value;
        if (P.n_extra > 0) return aux ? launch(render_f32_small_kernel<kF, true, true>) : launch(render_f32_small_kernel<kF, false, true>);
        return aux ? launch(render_f32_small_kernel<kF, true, false>) : launch(render_f32_small_kernel<kF, false, false>);
    };
    switch (out_format) {
        case RTC_OUT_F64_RGB: return pick(IntC<RTC_OUT_F64_RGB>{});
        case RTC_OUT_F32_RGB: return pick(IntC<RTC_OUT_F32_RGB>{});
        default:              return pick(IntC<RTC_OUT_U8_RGB>{});
    }
}

}  // namespace rtc
