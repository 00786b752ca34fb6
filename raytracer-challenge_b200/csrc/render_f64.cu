// F64 PARITY kernel — compiled with -fmad=false (see __graft_entry__.build) so that no a*b+c is
// contracted: Rust never fuses, and the hit/miss mask has to be bit-exact.
// f64 '/' and sqrt() are IEEE-rounded in CUDA by construction.
//
// One thread per pixel sample.  The reference's recursion
//   World::color_at -> shade_hit -> reflected_color / refracted_color -> color_at ...
// (src/shape.rs:685-766) becomes an explicit per-thread stack of at most max_depth frames,
// evaluated in the reference's post-order so every colour is combined with the same operations
// in the same order.
#include <type_traits>
#include "rtc_common.h"
#include "parity64.cuh"   // the reference's geometry statements in its operation order (shared with render_f32.cu)

namespace rtc {
namespace {
using namespace p64;

#ifndef RTC_F64_UNROLL
#define RTC_F64_UNROLL 1
#endif
constexpr int kUnroll = RTC_F64_UNROLL;   // shapes tested per loop trip in World::intersect (independent FP64 chains in flight)
constexpr int kBlockX = 16, kBlockY = 8, kThreads = kBlockX * kBlockY;

struct Streams {  // intersect streams, in shared memory when they fit, else global
    const double2* sph; const int32_t* sph_index; int ns;   // spheres [0, ns), then cubes [ns, ns + nc): same 3x4 inverse rows
    int nc;
    const double2* pl;  const int32_t* pl_index;  int np;
    const float4* bound; int nb4; float cmax, rmax;         // f32 bounding spheres of [0, ns + nc) (DevScene64::bound); nb4 = 0: none
    const float4* gbound; float gcmax, grmax;               // one per eight entries of `bound` (DevScene64::gbound), padded to a multiple of 32
    const float4* sbound; float scmax, srmax;               // one per 32 groups (DevScene64::sbound), padded to a multiple of 32
};
#ifndef RTC_F64_ONE_COPY
#define RTC_F64_ONE_COPY(ext) true   // one copy of the search for both ray kinds in every kernel (see intersect_world_impl)
#endif
// The block's copy, written once by thread 0: the searches below are out-of-line functions, and a Streams built in the
// kernel's registers would reach them through a reference, i.e. through the thread's local memory (every field a local load
// on every call; ncu counted 130 local loads per loop iteration before this).
__shared__ Streams g_streams;
// Result of a search, returned in registers (reference parameters of an out-of-line function live in local memory too).
struct Found { double t; int idx; int n_neg; };

// kCollect (radiance rays): the list entries with t < 0 that the loops come across are kept in `neg`
// (t, 2 * shape index + root): they are all Intersection::compute_refractive looks at, so a transparent
// hit needs no further pass over the shapes.  n_neg counts all of them; beyond kNegCap they are dropped
// and compute_refractive falls back to its two passes.
struct Neg { double t; int ord; int pad; };
constexpr int kNegCap = 48;
// kGroups: two-level filter (streams in global memory: big scenes, Morton order); otherwise entries four at a time
// kMode 0 / 1: kCollect fixed at compile time (two copies of the search in a kernel, one per ray kind); kMode 2: kCollect is
// the run-time argument — ONE copy serves radiance and shadow rays.  The kernel is bound by instruction fetch as much as by
// anything (5.7 k instructions, `no_instruction` 1.3 stall cycles per issue at the time): sharing the copy took 6 KB off the
// hot loop and the room frames from 7.39 to 7.12 ms (C2 2.21 -> 2.10, C4 71.9 -> 69.9).  The kExt kernels (resample pass,
// several lights) first lost 2-3 % to the extra live flag; once the streams' descriptor had moved to shared memory they
// gain too (samples = 4: 33.2 -> 32.2 ms, two lights 10.0 -> 9.7), so every kernel uses mode 2 (RTC_F64_ONE_COPY).
template <int kMode, bool kCubes, bool kGroups>
__device__ __noinline__ Found intersect_world_impl(V3 o, V3 d, Neg* neg, const bool collect) {
    const Streams& S = g_streams;
    const bool kCollect = kMode == 2 ? collect : kMode == 1;
    double best_t = __longlong_as_double(0x7ff0000000000000LL);
    int best_idx = 0x7fffffff;
    int n_neg = 0;
    auto keep = [&](double t, int ord) { if (n_neg < kNegCap) { neg[n_neg].t = t; neg[n_neg].ord = ord; } ++n_neg; };
#pragma unroll kUnroll
    for (int i = 0; i < S.np; ++i) {
        double t;
        if (plane_root(S.pl + 2 * i, o, d, t)) {
            take(t, S.pl_index[i], best_t, best_idx);
            if (kCollect && t < 0.0) keep(t, 2 * S.pl_index[i]);
        }
    }
    auto sphere = [&](int i) {
        double t1, t2;
        if (sphere_roots(S.sph + 6 * i, o, d, t1, t2)) {
            int idx = S.sph_index[i];
            if (t1 >= 0.0) take(t1, idx, best_t, best_idx);
            else {
                take(t2, idx, best_t, best_idx);
                if (kCollect) { if (t1 < 0.0) keep(t1, 2 * idx); if (t2 < 0.0) keep(t2, 2 * idx + 1); }
            }
        }
    };
    auto cube = [&](int i) {   // cubes: behind the spheres in the same stream
        double t1, t2;
        if (cube_roots(S.sph + 6 * i, o, d, t1, t2)) {
            int idx = S.sph_index[i];
            if (t1 >= 0.0) take(t1, idx, best_t, best_idx);
            else {
                take(t2, idx, best_t, best_idx);
                if (kCollect) { if (t1 < 0.0) keep(t1, 2 * idx); if (t2 < 0.0) keep(t2, 2 * idx + 1); }
            }
        }
    };
    // f32 line-distance filter in front of the exact tests (DevScene64::bound).  |(c - o) x d^| is the distance of
    // the centre from the ray's line; with m = o x d^ it is |c x d^ - m|.  E bounds the rounding of this test (o, d
    // rounded to f32, the cross products, an approximate 1/|d|) with a factor ~10 to spare; a shape is skipped only
    // when its gap is provably positive, i.e. when the exact test below would find disc < 0 (no list entry, no
    // negative-t entry either).  A NaN gap (degenerate ray) counts as a candidate.  Without bounds (nb4 = 0:
    // RTC_OPT_NO_BOUND_FILTER or a shape with no finite bound) every shape is a candidate: one loop, one copy of the
    // exact tests either way.
    const bool filter = S.nb4 > 0;
    const float fox = (float)o.x, foy = (float)o.y, foz = (float)o.z;
    float fdx = (float)d.x, fdy = (float)d.y, fdz = (float)d.z;
    const float inv_len = rsqrtf(fmaf(fdz, fdz, fmaf(fdy, fdy, fdx * fdx)));
    fdx *= inv_len; fdy *= inv_len; fdz *= inv_len;
    const float mx = foy * fdz - foz * fdy, my = foz * fdx - fox * fdz, mz = fox * fdy - foy * fdx;
    const float E = 3.8e-6f * (fmaxf(fmaxf(S.cmax, S.gcmax), S.scmax) + fabsf(fox) + fabsf(foy) + fabsf(foz));
    const float margin = fmaf(E, 2.f * S.rmax + E, 2e-6f * S.rmax * S.rmax);
    const float gmargin = fmaf(E, 2.f * S.grmax + E, 2e-6f * S.grmax * S.grmax);
    const float smargin = fmaf(E, 2.f * S.srmax + E, 2e-6f * S.srmax * S.srmax);
    auto gap = [&](float4 b) {
        const float vx = fmaf(-b.z, fdy, fmaf(b.y, fdz, -mx));
        const float vy = fmaf(-b.x, fdz, fmaf(b.z, fdx, -my));
        const float vz = fmaf(-b.y, fdx, fmaf(b.x, fdy, -mz));
        return fmaf(vz, vz, fmaf(vy, vy, fmaf(vx, vx, b.w)));
    };
    // Two levels: eight consecutive entries (big scenes are stored in Morton order: neighbours in space) share a
    // bounding sphere, four such groups per trip; inside a candidate group, four entries per trip.
    const int n_solid = S.ns + (kCubes ? S.nc : 0);
    auto members = [&](int g) {   // entries g .. g+3: the exact tests of the candidates
        unsigned cand;
        if (filter) {   // the padding entries of the last group have gap = +inf: never candidates
            const float4 b0 = S.bound[g], b1 = S.bound[g + 1], b2 = S.bound[g + 2], b3 = S.bound[g + 3];
            const float s0 = gap(b0), s1 = gap(b1), s2 = gap(b2), s3 = gap(b3);
            cand = (!(s0 > margin) ? 1u : 0u) | (!(s1 > margin) ? 2u : 0u) | (!(s2 > margin) ? 4u : 0u) | (!(s3 > margin) ? 8u : 0u);
        } else {
            cand = n_solid - g >= 4 ? 15u : (1u << (n_solid - g)) - 1u;
        }
        while (cand) {
            const int i = g + __ffs(cand) - 1;
            cand &= cand - 1u;
            if (!kCubes || i < S.ns) sphere(i); else cube(i);
        }
    };
    if (!kGroups) {
        for (int g = 0; g < n_solid; g += 4) members(g);   // (unroll 2 measured: room 7.35 -> 8.8 ms)
    } else {
        // Three levels, each a line-distance test against a bounding sphere: blocks of 32 groups (256 entries), groups
        // of eight entries, entries.  32 tests of a level run branch-free into one bit mask per lane and the lanes then
        // drain their masks side by side, every lane walking its own candidate blocks and groups (per-lane addresses
        // below the block level): a warp pays for the largest number of candidates any of its lanes has.  The bound
        // arrays are padded to multiples of 32 with never-hit entries.  Without the filter everything is a candidate.
        const int n_groups = (n_solid + 7) >> 3;
        const int n_blocks = (n_groups + 31) >> 5;
        for (int sb = 0; sb < n_blocks; sb += 32) {
            unsigned sm = n_blocks - sb >= 32 ? 0xffffffffu : (1u << (n_blocks - sb)) - 1u;
            if (filter) {
                sm = 0u;
#pragma unroll 2
                for (int j = 0; j < 32; j += 4) {
                    const float4 q0 = S.sbound[sb + j], q1 = S.sbound[sb + j + 1], q2 = S.sbound[sb + j + 2], q3 = S.sbound[sb + j + 3];
                    const float t0 = gap(q0), t1 = gap(q1), t2 = gap(q2), t3 = gap(q3);
                    sm |= ((!(t0 > smargin) ? 1u : 0u) | (!(t1 > smargin) ? 2u : 0u) | (!(t2 > smargin) ? 4u : 0u) | (!(t3 > smargin) ? 8u : 0u)) << j;
                }
            }
            while (sm) {
                const int gb = 32 * (sb + __ffs(sm) - 1);
                sm &= sm - 1u;
                unsigned gm = n_groups - gb >= 32 ? 0xffffffffu : (1u << (n_groups - gb)) - 1u;
                if (filter) {
                    gm = 0u;
#pragma unroll 2
                    for (int j = 0; j < 32; j += 4) {
                        const float4 q0 = S.gbound[gb + j], q1 = S.gbound[gb + j + 1], q2 = S.gbound[gb + j + 2], q3 = S.gbound[gb + j + 3];
                        const float t0 = gap(q0), t1 = gap(q1), t2 = gap(q2), t3 = gap(q3);
                        gm |= ((!(t0 > gmargin) ? 1u : 0u) | (!(t1 > gmargin) ? 2u : 0u) | (!(t2 > gmargin) ? 4u : 0u) | (!(t3 > gmargin) ? 8u : 0u)) << j;
                    }
                }
                while (gm) {
                    const int g = 8 * (gb + __ffs(gm) - 1);
                    gm &= gm - 1u;
                    unsigned cm = n_solid - g >= 8 ? 255u : (1u << (n_solid - g)) - 1u;
                    if (filter) {   // the padding entries of the last group have gap = +inf: never candidates
                        cm = 0u;
#pragma unroll
                        for (int j = 0; j < 8; j += 4) {
                            const float4 b0 = S.bound[g + j], b1 = S.bound[g + j + 1], b2 = S.bound[g + j + 2], b3 = S.bound[g + j + 3];
                            const float s0 = gap(b0), s1 = gap(b1), s2 = gap(b2), s3 = gap(b3);
                            cm |= ((!(s0 > margin) ? 1u : 0u) | (!(s1 > margin) ? 2u : 0u) | (!(s2 > margin) ? 4u : 0u) | (!(s3 > margin) ? 8u : 0u)) << j;
                        }
                    }
                    while (cm) {
                        const int i = g + __ffs(cm) - 1;
                        cm &= cm - 1u;
                        if (!kCubes || i < S.ns) sphere(i); else cube(i);
                    }
                }
            }
        }
    }
    if (best_idx == 0x7fffffff) best_idx = -1;
    Found f; f.t = best_t; f.idx = best_idx; f.n_neg = n_neg;
    return f;
}
template <bool kCollect, bool kCubes, bool kGroups, bool kOneCopy>
__device__ __forceinline__ void intersect_world(V3 o, V3 d, double& best_t, int& best_idx, Neg* neg, int& n_neg) {
    const Found f = intersect_world_impl<kOneCopy ? 2 : (kCollect ? 1 : 0), kCubes, kGroups>(o, d, neg, kCollect);
    best_t = f.t; best_idx = f.idx;
    if (kCollect) n_neg = f.n_neg;
}

// Intersection::compute_refractive (src/shape.rs:115-141) without materialising the list: see parity64.cuh (Cont, offer).
template <bool kCubes>
__device__ __noinline__ void compute_refractive(const DevScene64& sc, V3 o, V3 d, int hit_idx,
                                                double& n1, double& n2) {
    const Streams& S = g_streams;
    unsigned parity[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) parity[k] = 0u;
    // pass A: parity of the number of negative-t entries per world_id
    for (int i = 0; i < S.np; ++i) {
        double t;
        if (plane_root(S.pl + 2 * i, o, d, t) && t < 0.0) {
            int id = sc.rec[S.pl_index[i]].world_id & 255;
            parity[id >> 5] ^= 1u << (id & 31);
        }
    }
    for (int i = 0; i < S.ns; ++i) {
        double t1, t2;
        if (sphere_roots(S.sph + 6 * i, o, d, t1, t2)) {
            int id = sc.rec[S.sph_index[i]].world_id & 255;
            if (t1 < 0.0) parity[id >> 5] ^= 1u << (id & 31);
            if (t2 < 0.0) parity[id >> 5] ^= 1u << (id & 31);
        }
    }
    if (kCubes) for (int i = S.ns; i < S.ns + S.nc; ++i) {   // cubes: behind the spheres in the same stream
        double t1, t2;
        if (cube_roots(S.sph + 6 * i, o, d, t1, t2)) {
            int id = sc.rec[S.sph_index[i]].world_id & 255;
            if (t1 < 0.0) parity[id >> 5] ^= 1u << (id & 31);
            if (t2 < 0.0) parity[id >> 5] ^= 1u << (id & 31);
        }
    }
    // pass B: the last two containers
    Cont top1, top2;
    top1.idx = -1; top2.idx = -1; top1.t = top2.t = 0.0; top1.ord = top2.ord = 0; top1.id = top2.id = -1;
    for (int i = 0; i < S.np; ++i) {
        double t;
        if (plane_root(S.pl + 2 * i, o, d, t) && t < 0.0) {
            int idx = S.pl_index[i];
            int id = sc.rec[idx].world_id & 255;
            if ((parity[id >> 5] >> (id & 31)) & 1u) { Cont e; e.t = t; e.ord = 2 * idx; e.idx = idx; e.id = id; offer(e, top1, top2); }
        }
    }
    for (int i = 0; i < S.ns; ++i) {
        double t1, t2;
        if (sphere_roots(S.sph + 6 * i, o, d, t1, t2)) {
            int idx = S.sph_index[i];
            int id = sc.rec[idx].world_id & 255;
            if ((parity[id >> 5] >> (id & 31)) & 1u) {
                if (t1 < 0.0) { Cont e; e.t = t1; e.ord = 2 * idx; e.idx = idx; e.id = id; offer(e, top1, top2); }
                if (t2 < 0.0) { Cont e; e.t = t2; e.ord = 2 * idx + 1; e.idx = idx; e.id = id; offer(e, top1, top2); }
            }
        }
    }
    if (kCubes) for (int i = S.ns; i < S.ns + S.nc; ++i) {   // cubes: behind the spheres in the same stream
        double t1, t2;
        if (cube_roots(S.sph + 6 * i, o, d, t1, t2)) {
            int idx = S.sph_index[i];
            int id = sc.rec[idx].world_id & 255;
            if ((parity[id >> 5] >> (id & 31)) & 1u) {
                if (t1 < 0.0) { Cont e; e.t = t1; e.ord = 2 * idx; e.idx = idx; e.id = id; offer(e, top1, top2); }
                if (t2 < 0.0) { Cont e; e.t = t2; e.ord = 2 * idx + 1; e.idx = idx; e.id = id; offer(e, top1, top2); }
            }
        }
    }
    n1 = top1.idx >= 0 ? sc.rec[top1.idx].refractive_index : 1.0;
    int hid = sc.rec[hit_idx].world_id & 255;
    bool hit_is_container = (parity[hid >> 5] >> (hid & 31)) & 1u;
    if (hit_is_container) {
        if (top1.idx >= 0 && top1.id == hid) n2 = top2.idx >= 0 ? sc.rec[top2.idx].refractive_index : 1.0;
        else n2 = top1.idx >= 0 ? sc.rec[top1.idx].refractive_index : 1.0;
    } else {
        n2 = sc.rec[hit_idx].refractive_index;
    }
}

// The same from the entries intersect_world<true> kept for this ray.
struct N12 { double n1, n2; };
__device__ __noinline__ N12 compute_refractive_kept(const DevScene64& sc, int hit_idx, const Neg* neg, int n_neg) {
    double n1, n2;
    if (sc.ids_unique) {
        // Unique world_ids (every scene of at most 256 shapes): an id is toggled once per negative entry of ITS shape, so a
        // shape is a container iff exactly one of its entries is negative.  A sphere or cube with both roots negative was
        // kept as two consecutive entries: skipped as a pair — no parity words (a dynamically indexed local array).
        Cont top1, top2;
        top1.idx = -1; top2.idx = -1; top1.t = top2.t = 0.0; top1.ord = top2.ord = 0; top1.id = top2.id = -1;
        bool hit_is_container = false;
        for (int k = 0; k < n_neg; ++k) {
            const int idx = neg[k].ord >> 1;
            if (k + 1 < n_neg && (neg[k + 1].ord >> 1) == idx) { ++k; continue; }
            if (idx == hit_idx) hit_is_container = true;
            Cont e; e.t = neg[k].t; e.ord = neg[k].ord; e.idx = idx; e.id = idx;
            if (top1.idx < 0 || later(e, top1)) { top2 = top1; top1 = e; }
            else if (top2.idx < 0 || later(e, top2)) top2 = e;
        }
        n1 = top1.idx >= 0 ? sc.rec[top1.idx].refractive_index : 1.0;
        if (hit_is_container) {
            if (top1.idx == hit_idx) n2 = top2.idx >= 0 ? sc.rec[top2.idx].refractive_index : 1.0;
            else n2 = n1;
        } else {
            n2 = sc.rec[hit_idx].refractive_index;
        }
        N12 r; r.n1 = n1; r.n2 = n2;
        return r;
    }
    unsigned parity[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) parity[k] = 0u;
    for (int k = 0; k < n_neg; ++k) {
        int id = sc.rec[neg[k].ord >> 1].world_id & 255;
        parity[id >> 5] ^= 1u << (id & 31);
    }
    Cont top1, top2;
    top1.idx = -1; top2.idx = -1; top1.t = top2.t = 0.0; top1.ord = top2.ord = 0; top1.id = top2.id = -1;
    for (int k = 0; k < n_neg; ++k) {
        const int idx = neg[k].ord >> 1;
        const int id = sc.rec[idx].world_id & 255;
        if ((parity[id >> 5] >> (id & 31)) & 1u) { Cont e; e.t = neg[k].t; e.ord = neg[k].ord; e.idx = idx; e.id = id; offer(e, top1, top2); }
    }
    n1 = top1.idx >= 0 ? sc.rec[top1.idx].refractive_index : 1.0;
    int hid = sc.rec[hit_idx].world_id & 255;
    bool hit_is_container = (parity[hid >> 5] >> (hid & 31)) & 1u;
    if (hit_is_container) {
        if (top1.idx >= 0 && top1.id == hid) n2 = top2.idx >= 0 ? sc.rec[top2.idx].refractive_index : 1.0;
        else n2 = top1.idx >= 0 ? sc.rec[top1.idx].refractive_index : 1.0;
    } else {
        n2 = sc.rec[hit_idx].refractive_index;
    }
    N12 r; r.n1 = n1; r.n2 = n2;
    return r;
}

// `n % 2.0 == 0.0` (Rust's % on f64 is fmod) for an INTEGER-VALUED n — a floor(), or a sum of floor()s
// (src/material.rs:97-103,163-170,198-205) — without the ~170-instruction fmod routine: the remainder is 0 iff n is even,
// n * 0.5 is exact, and it is an integer iff n is even (from 2^53 on every double is an even integer and its half an
// integer).  fmod(+-inf) and fmod(NaN) are NaN, which never equals 0: the magnitude test and the NaN compare keep that.
__device__ __forceinline__ bool even_int(double n) {
    const double h = n * 0.5;
    return h == floor(h) && fabs(n) <= 1.7976931348623157e308;
}

// Pattern::pattern_at for the six patterns (src/material.rs:67-69,97-103,133-135,163-170,198-205,233-241)
__device__ __forceinline__ V3 pattern_at(const ShadeRec64& r, V3 p) {
    V3 a = mk(r.pat_a[0], r.pat_a[1], r.pat_a[2]), b = mk(r.pat_b[0], r.pat_b[1], r.pat_b[2]);
    switch (r.pattern_kind) {
        case RTC_PATTERN_TEST: return p;
        case RTC_PATTERN_STRIPE: return even_int(floor(p.x)) ? a : b;
        case RTC_PATTERN_GRADIENT: return add(a, mul(sub(b, a), p.x - floor(p.x)));
        case RTC_PATTERN_RING: return even_int(floor(sqrt(p.x * p.x + p.y * p.y))) ? a : b;
        case RTC_PATTERN_CHECKER: return even_int(floor(p.x) + floor(p.y) + floor(p.z)) ? a : b;
        default: /* GRID: a = base, b = grid lines */
            return (fabs(p.x - floor(p.x)) < 0.01 || fabs(p.z - floor(p.z)) < 0.01) ? b : a;
    }
}

// Material::lighting (src/material.rs:319-361); `point` is over_point (src/shape.rs:688)
__device__ __forceinline__ V3 lighting(const ShadeRec64& r, V3 light_pos, V3 light_int, V3 point, V3 eye, V3 normal,
                                       bool in_shadow) {
    V3 base;
    if (r.pattern_kind != RTC_PATTERN_NONE) {
        V3 object_point = xpoint(r.inv, point);            // src/material.rs:41-45
        V3 pattern_point = xpoint(r.pat_inv, object_point);
        base = pattern_at(r, pattern_point);
    } else {
        base = mk(r.color[0], r.color[1], r.color[2]);
    }
    V3 effective = mulc(base, light_int);
    V3 lightv = normalize(sub(light_pos, point));
    V3 ambient = mul(effective, r.ambient);
    if (in_shadow) return ambient;
    double ldn = dot(lightv, normal);
    V3 diffuse = mk(0., 0., 0.), specular = mk(0., 0., 0.);
    if (!(ldn < 0.)) {
        diffuse = mul(effective, r.diffuse * ldn);
        V3 rv = reflect(neg(lightv), normal);
        double rde = dot(rv, eye);
        if (!(rde <= 0.)) {
            double factor = pow(rde, r.shininess);
            specular = mul(light_int, r.specular * factor);
        }
    }
    return add(add(ambient, diffuse), specular);
}

struct Frame {  // one suspended shade_hit (src/shape.rs:685-700)
    V3 surface;
    V3 refl;            // reflected_color once its subtree is done
    V3 ro, rd;          // pending refraction ray
    double k_refl, k_transp, reflectance;
    unsigned char both, has_refr, waiting_refr;
};

// thread -> sample mapping: a warp covers an 8x4 patch of the sample grid, a block 16x8.
__device__ __forceinline__ void thread_sample(int& sx, int& sy) {
    int w = threadIdx.x >> 5, l = threadIdx.x & 31;
    sx = blockIdx.x * kBlockX + (w & 1) * 8 + (l & 7);
    sy = blockIdx.y * kBlockY + (w >> 1) * 4 + (l >> 3);
}

// kCubes: the scene has cubes (src/shape.rs:500-630); scenes without them run exactly the sphere / plane code
// kExt: antialiasing_samples >= 2 (the resample passes of Camera::render_pixel, src/camera.rs:84-114) and / or more than one
// light (rtc_scene_upload_lights); frames with neither run the kExt = false code, which compiles none of it
template <bool kInSmem, bool kCubes, bool kExt>
#ifndef RTC_F64_MIN_BLOCKS
#define RTC_F64_MIN_BLOCKS 4
#endif
#ifndef RTC_F64_MIN_BLOCKS_GLOBAL
#define RTC_F64_MIN_BLOCKS_GLOBAL 8
#endif
// Blocks per SM.  Streams in shared memory (the room scenes): the kernel waits for its own FP64 chains and for instruction
// fetch, and 4 blocks at 128 registers beat 5 / 6 blocks at 96 / 80 registers with spills (C3 4K 6.69 / 6.71 / 6.83 ms).
// Streams in global memory (big scenes): it waits for the hierarchy's loads, and more resident warps pay for the
// spills (C4 4K 69.7 / 65.3 / 61.1 / 58.0 / 56.2 / 62.9 / 74.4 ms at 4 / 5 / 6 / 7 / 8 / 10 / 12 blocks: 8 blocks, 64 registers).
__global__ void __launch_bounds__(kThreads, kInSmem ? RTC_F64_MIN_BLOCKS : RTC_F64_MIN_BLOCKS_GLOBAL)
render_f64_kernel(const __grid_constant__ DevScene64 sc, const __grid_constant__ CamParams cam, double* __restrict__ out_f64, float* __restrict__ out_f32,
                  uint8_t* __restrict__ out_u8, uint8_t* __restrict__ out_hit, Counters* __restrict__ counters) {
    extern __shared__ double2 smem[];
    Streams S;
    S.ns = sc.n_spheres; S.nc = sc.n_cubes; S.np = sc.n_planes;
    const int n_solid = sc.n_spheres + sc.n_cubes;
    if (kInSmem) {
        // stage the intersect streams: coalesced 16-byte loads, one pass per block
        double2* s_sph = smem;
        double2* s_pl = s_sph + 6 * n_solid;
        float4* s_bound = reinterpret_cast<float4*>(s_pl + 2 * sc.n_planes);
        float4* s_gbound = s_bound + sc.n_bound4;
        const int n_gb = sc.n_bound4 > 0 ? ((sc.n_bound4 >> 3) + 3) / 4 * 4 : 0;
        int32_t* s_si = reinterpret_cast<int32_t*>(s_gbound + n_gb);
        int32_t* s_pi = s_si + n_solid;
        for (int i = threadIdx.x; i < 6 * n_solid; i += kThreads) s_sph[i] = sc.sph_rows[i];
        for (int i = threadIdx.x; i < 2 * sc.n_planes; i += kThreads) s_pl[i] = sc.pl_rows[i];
        for (int i = threadIdx.x; i < sc.n_bound4; i += kThreads) s_bound[i] = sc.bound[i];
        for (int i = threadIdx.x; i < n_gb; i += kThreads) s_gbound[i] = sc.gbound[i];
        for (int i = threadIdx.x; i < n_solid; i += kThreads) s_si[i] = sc.sph_index[i];
        for (int i = threadIdx.x; i < sc.n_planes; i += kThreads) s_pi[i] = sc.pl_index[i];
        __syncthreads();
        S.sph = s_sph; S.pl = s_pl; S.sph_index = s_si; S.pl_index = s_pi; S.bound = s_bound; S.gbound = s_gbound;
    } else {
        S.sph = sc.sph_rows; S.pl = sc.pl_rows; S.sph_index = sc.sph_index; S.pl_index = sc.pl_index; S.bound = sc.bound; S.gbound = sc.gbound;
    }
    S.nb4 = sc.n_bound4; S.cmax = sc.bound_cmax; S.rmax = sc.bound_rmax; S.gcmax = sc.gbound_cmax; S.grmax = sc.gbound_rmax;
    S.sbound = sc.sbound; S.scmax = sc.sbound_cmax; S.srmax = sc.sbound_rmax;
    if (threadIdx.x == 0) g_streams = S;   // (the staged data is already visible: the barrier above; this one publishes the struct)
    __syncthreads();

    const int ss = cam.samples == 1 ? 1 : 2;  // samples per pixel edge
    int sx, sy;
    thread_sample(sx, sy);
    const int px = sx / ss, pr = sy / ss;     // pixel x, compact shard row
    const bool in_image = px < (int)cam.hsize && pr < cam.shard_rows;
    int py = 0;
    if (in_image) py = cam.shard_count > 1 ? shard_row_to_frame_row(cam.row0 + pr, cam.tile_rows, cam.shard_index, cam.shard_count) : cam.row0 + pr;
    const bool active = in_image && px < (int)cam.xlim && py < (int)cam.ylim;

    V3 result = mk(0., 0., 0.);
    bool primary_hit = false;
    unsigned n_rays = 0, n_shadow = 0, n_shaded = 0;

    // Camera::render_pixel (src/camera.rs:94-114).  Pass 0 traces this thread's fixed sample; with
    // antialiasing_samples >= 2 the four threads of a pixel then share the resample pass (:84-92): extra sample
    // k = 4 (pass - 1) + q goes to quad member q, and the colours are summed in k order like Color::average_over.
    const int n_extra = (kExt && cam.samples >= 2) ? (int)cam.samples : 0;
    const int n_lights = kExt ? sc.n_lights : 1;
    const unsigned full_mask = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const int quad_base = lane & 0x16;                   // lane of the pixel's (.25, .25) sample in the 8x4 warp patch
    const int quad_member = (lane & 1) | ((lane >> 2) & 2);
    double xo = ss == 1 ? 0.5 : ((sx & 1) ? 0.75 : 0.25);
    double yo = ss == 1 ? 0.5 : ((sy & 1) ? 0.75 : 0.25);
    bool trace_now = active;
    bool need_resample = false;
    double reds = 0., greens = 0., blues = 0.;           // Color::average_over's running sums (src/color.rs:128-139)
    unsigned hit_bits = 0;

    for (int pass = 0;; ++pass) {
    V3 sample_color = mk(0., 0., 0.);
    if (trace_now) {
        // Camera::ray_for_pixel_offset (src/camera.rs:64-76)
        double xoffset = ((double)px + xo) * cam.pixel_size;
        double yoffset = ((double)py + yo) * cam.pixel_size;
        double world_x = cam.half_width - xoffset;
        double world_y = cam.half_height - yoffset;
        V3 pixel = xpoint(cam.vinv, mk(world_x, world_y, -1.));
        V3 origin = xpoint(cam.vinv, mk(0., 0., 0.));
        V3 ro = origin, rd = normalize(sub(pixel, origin));

        const V3 light_pos = mk(sc.light_pos[0], sc.light_pos[1], sc.light_pos[2]);
        const V3 light_int = mk(sc.light_int[0], sc.light_int[1], sc.light_int[2]);
        Frame stack[RTC_MAX_DEPTH];
        Neg kept[kNegCap];   // negative-t entries of the current radiance ray
        int n_neg = 0, n_shadow_neg = 0;
        int level = 0;
        int remaining = cam.max_depth;
        bool first = pass == 0;
        for (;;) {
            // ---- World::color_at(ray, remaining)  (src/shape.rs:702-710) ----
            V3 color;
            bool descended = false;
            double t; int idx;
            ++n_rays;
            intersect_world<true, kCubes, !kInSmem, RTC_F64_ONE_COPY(kExt)>(ro, rd, t, idx, kept, n_neg);
            if (first) { primary_hit = idx >= 0; first = false; }
            if (idx < 0) {
                color = mk(0., 0., 0.);  // background_color, src/shape.rs:652
            } else {
                RTC_CHECK(idx >= 0 && idx < sc.n_shapes);
                const ShadeRec64& rec = sc.rec[idx];
                // Intersection::compute_vectors + CachedVectors::new (src/shape.rs:144-152, 75-96)
                V3 point = add(ro, mul(rd, t));
                V3 eyev = neg(rd);
                V3 lp = xpoint(rec.inv, point);                                    // Shape::normal_at :34-40
                V3 ln = rec.kind == RTC_SHAPE_PLANE ? mk(0., 1., 0.) : lp;
                if (kCubes && rec.kind == RTC_SHAPE_CUBE) {                                  // Cube::normal_at_local :601-610; rec.invt = transpose(inverse), :623-629
                    const double ax = fabs(lp.x), ay = fabs(lp.y), az = fabs(lp.z);
                    const double maxc = fmax(ax, fmax(ay, az));
                    if (maxc == ax) ln = mk(lp.x, 0., 0.);
                    else if (maxc == ay) ln = mk(0., lp.y, 0.);
                    else ln = mk(0., 0., lp.z);
                }
                V3 normal = normalize(xvec3(rec.invt, ln));
                double n1 = 1.0, n2 = 1.0;
                if (rec.transparency != 0.0) {
                    if (n_neg <= kNegCap) { const N12 r = compute_refractive_kept(sc, idx, kept, n_neg); n1 = r.n1; n2 = r.n2; }
                    else compute_refractive<kCubes>(sc, ro, rd, idx, n1, n2);   // more entries than were kept: the two passes
                }
                bool inside = dot(normal, eyev) < 0.0;
                if (inside) normal = neg(normal);
                V3 over_point = add(point, mul(normal, kEps));
                V3 under_point = sub(point, mul(normal, kEps));
                V3 reflectv = reflect(rd, normal);
                // surface = lighting(light, .., is_shadowed(over_point)) (src/shape.rs:686-688); with rtc_scene_upload_lights
                // one such term per light, added in list order (the "FIXME -- multiple lights" of :686; include/rtc.h)
                ++n_shaded;
                V3 surface = mk(0., 0., 0.);
                for (int li = 0; li < n_lights; ++li) {
                    V3 lpos = light_pos, lint = light_int;
                    if (li > 0) {
                        const double* L = sc.lights + 6 * li;
                        lpos = mk(L[0], L[1], L[2]); lint = mk(L[3], L[4], L[5]);
                    }
                    // World::is_shadowed_by_light (src/shape.rs:716-727)
                    V3 v = sub(lpos, over_point);
                    double distance = dsqrt(v.x * v.x + v.y * v.y + v.z * v.z);
                    V3 sdir = normalize(v);
                    double st; int sidx;
                    ++n_rays; ++n_shadow;
                    intersect_world<false, kCubes, !kInSmem, RTC_F64_ONE_COPY(kExt)>(over_point, sdir, st, sidx, kept, n_shadow_neg);
                    bool shadowed = sidx >= 0 && st < distance;
                    V3 term = lighting(rec, lpos, lint, over_point, eyev, normal, shadowed);
                    surface = li == 0 ? term : add(surface, term);
                }
                // shade_hit's children (src/shape.rs:689-699, 729-766)
                bool want_refl = remaining != 0 && !(rec.reflective <= 0.);
                bool want_refr = false;
                V3 fo = mk(0., 0., 0.), fd = mk(0., 0., 0.);
                if (remaining != 0 && rec.transparency != 0.0) {
                    double n_ratio = ddiv(n1, n2);
                    double cos_i = dot(eyev, normal);
                    double sin2_t = (n_ratio * n_ratio) * (1.0 - (cos_i * cos_i));
                    if (!(sin2_t > 1.0)) {
                        double cos_t = dsqrt(1.0 - sin2_t);
                        fd = sub(mul(normal, n_ratio * cos_i - cos_t), mul(eyev, n_ratio));
                        fo = under_point;
                        want_refr = true;
                    }
                }
                bool both = rec.reflective > 0.0 && rec.transparency > 0.0;
                double reflectance = 0.0;
                if (both) {  // World::reflectance, src/shape.rs:768-781
                    double cs = dot(eyev, normal);
                    bool tir = false;
                    if (n1 > n2) {
                        double n = ddiv(n1, n2);
                        double sin2_t = (n * n) * (1.0 - (cs * cs));
                        if (sin2_t > 1.0) tir = true;
                        else cs = dsqrt(1.0 - sin2_t);
                    }
                    if (tir) reflectance = 1.0;
                    else {
                        double q = ddiv(n1 - n2, n1 + n2);
                        double r0 = q * q;
                        double m = 1.0 - cs;
                        double m2 = m * m;
                        double m4 = m2 * m2;
                        reflectance = r0 + (1.0 - r0) * (m4 * m);
                    }
                }
                if (!want_refl && !want_refr) {
                    // both children are BLACK; surface + 0 is surface
                    if (both) {
                        V3 z = mk(0., 0., 0.);
                        color = add(surface, add(mul(z, reflectance), mul(z, 1.0 - reflectance)));
                    } else {
                        color = surface;
                    }
                } else {
                    RTC_CHECK(level >= 0 && level < RTC_MAX_DEPTH);
                    Frame& f = stack[level];
                    f.surface = surface;
                    f.refl = mk(0., 0., 0.);
                    f.k_refl = rec.reflective; f.k_transp = rec.transparency; f.reflectance = reflectance;
                    f.both = both; f.has_refr = want_refr; f.ro = fo; f.rd = fd;
                    if (want_refl) { f.waiting_refr = 0; ro = over_point; rd = reflectv; }
                    else { f.waiting_refr = 1; ro = fo; rd = fd; }
                    ++level; --remaining;
                    descended = true;
                }
            }
            if (descended) continue;
            // ---- return `color` to the suspended callers ----
            bool resumed = false;
            while (level > 0) {
                Frame& f = stack[level - 1];
                if (!f.waiting_refr) {
                    f.refl = mul(color, f.k_refl);                   // reflected_color, :736
                    if (f.has_refr) {
                        f.waiting_refr = 1;
                        ro = f.ro; rd = f.rd;
                        resumed = true;
                        break;
                    }
                    V3 z = mk(0., 0., 0.);
                    color = f.both ? add(f.surface, add(mul(f.refl, f.reflectance), mul(z, 1.0 - f.reflectance)))
                                   : add(add(f.surface, f.refl), z);
                } else {
                    V3 refr = mul(color, f.k_transp);               // refracted_color, :765
                    color = f.both ? add(f.surface, add(mul(f.refl, f.reflectance), mul(refr, 1.0 - f.reflectance)))
                                   : add(add(f.surface, f.refl), refr);
                }
                --level; ++remaining;
            }
            if (resumed) continue;
            sample_color = color;
            break;
        }
    }

    if (pass == 0) {
        result = sample_color;
        hit_bits = primary_hit ? 1u : 0u;
        if (ss == 1) break;
        // ---- combine the fixed sub-samples (src/camera.rs:96-113; Color::average_over) ----
        const int src[4] = {quad_base, quad_base + 1, quad_base + 8, quad_base + 9};
        bool any = false;
        double sr[4], sg[4], sb[4];
        unsigned bits = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            sr[k] = __shfl_sync(full_mask, sample_color.x, src[k]);
            sg[k] = __shfl_sync(full_mask, sample_color.y, src[k]);
            sb[k] = __shfl_sync(full_mask, sample_color.z, src[k]);
            bits |= __shfl_sync(full_mask, hit_bits, src[k]) << k;
            reds += sr[k]; greens += sg[k]; blues += sb[k];
        }
        hit_bits = bits;
        result = mk(reds / 4.0, greens / 4.0, blues / 4.0);
        if (n_extra == 0) break;
#pragma unroll
        for (int k = 0; k < 4; ++k) {                    // Color::distance_from, src/color.rs:122-126
            const double dr = sr[k] - result.x, dg = sg[k] - result.y, db = sb[k] - result.z;
            if (sqrt(dr * dr + dg * dg + db * db) > 0.01) any = true;
        }
        need_resample = active && any;
        if (!__any_sync(full_mask, need_resample)) break;
    } else {
        // the four extra samples of this pass, in list order
        const int src[4] = {quad_base, quad_base + 1, quad_base + 8, quad_base + 9};
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            const double r = __shfl_sync(full_mask, sample_color.x, src[q]);
            const double g = __shfl_sync(full_mask, sample_color.y, src[q]);
            const double b = __shfl_sync(full_mask, sample_color.z, src[q]);
            if (4 * (pass - 1) + q < n_extra) { reds += r; greens += g; blues += b; }
        }
        if (4 * pass >= n_extra) {
            if (need_resample) {
                const double l = (double)(4 + n_extra);
                result = mk(reds / l, greens / l, blues / l);
            }
            break;
        }
    }
    {   // next pass: extra sample k of this pixel (Camera::resample, src/camera.rs:84-92; offsets: include/rtc.h)
        const int k = 4 * pass + quad_member;
        trace_now = need_resample && k < n_extra;
        const uint64_t seed_mixed = splitmix64((uint64_t)cam.aa_seed);
        const uint64_t pixel_index = (uint64_t)py * cam.hsize + (uint64_t)px;
        xo = (double)(resample_bits(seed_mixed, pixel_index, (unsigned)k, 0u) >> 11) * (1.0 / 9007199254740992.0);
        yo = (double)(resample_bits(seed_mixed, pixel_index, (unsigned)k, 1u) >> 11) * (1.0 / 9007199254740992.0);
    }
    }

    const bool writer = in_image && (ss == 1 || (((sx | sy) & 1) == 0));
    if (writer) {
        size_t p = (size_t)(cam.dst_full ? py : (cam.shard_count > 1 ? cam.row0 + pr : pr)) * cam.hsize + px;
        RTC_CHECK(px >= 0 && px < (int)cam.hsize && (cam.dst_full ? (py >= 0 && py < (int)cam.vsize) : (pr >= 0 && pr < cam.shard_rows)));
        if (cam.out_format == RTC_OUT_F64_RGB) {
            out_f64[3 * p + 0] = result.x; out_f64[3 * p + 1] = result.y; out_f64[3 * p + 2] = result.z;
        } else if (cam.out_format == RTC_OUT_F32_RGB) {
            out_f32[3 * p + 0] = (float)result.x; out_f32[3 * p + 1] = (float)result.y; out_f32[3 * p + 2] = (float)result.z;
        } else {
            // Color::scale: clamp((c*255.0) as i32, 0, 255)  (src/color.rs:100-114)
            double c3[3] = {result.x, result.y, result.z};
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                double s = c3[k] * 255.0;
                int q = s != s ? 0 : (s >= 2147483647.0 ? 2147483647 : (s <= -2147483648.0 ? (-2147483647 - 1) : (int)s));
                out_u8[3 * p + k] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
            }
        }
        if (out_hit) out_hit[p] = (uint8_t)hit_bits;
    }
    if (counters) {
        const unsigned full = 0xffffffffu;
        unsigned r = __reduce_add_sync(full, n_rays), s = __reduce_add_sync(full, n_shadow);
        unsigned h = __reduce_add_sync(full, n_shaded), ph = __reduce_add_sync(full, primary_hit ? 1u : 0u);
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&counters->rays, (unsigned long long)r);
            atomicAdd(&counters->shadow_rays, (unsigned long long)s);
            atomicAdd(&counters->shaded_hits, (unsigned long long)h);
            atomicAdd(&counters->primary_hits, (unsigned long long)ph);
        }
    }
}

}  // namespace

cudaError_t preload_render_f64() {
    cudaFuncAttributes at;
    cudaError_t e = cudaSuccess;
    auto touch = [&](auto kernel) { if (e == cudaSuccess) e = cudaFuncGetAttributes(&at, kernel); };
    touch(render_f64_kernel<true, true, true>); touch(render_f64_kernel<true, true, false>); touch(render_f64_kernel<true, false, true>); touch(render_f64_kernel<true, false, false>);
    touch(render_f64_kernel<false, true, true>); touch(render_f64_kernel<false, true, false>); touch(render_f64_kernel<false, false, true>); touch(render_f64_kernel<false, false, false>);
    return e;
}

cudaError_t launch_render_f64(const DevScene64& scene, const CamParams& cam, void* d_rgb, uint8_t* d_hit,
                              Counters* d_counters, cudaStream_t stream) {
    const int ss = cam.samples == 1 ? 1 : 2;
    dim3 grid((cam.hsize * ss + kBlockX - 1) / kBlockX, (cam.shard_rows * ss + kBlockY - 1) / kBlockY);
    if (grid.x == 0 || grid.y == 0) return cudaSuccess;
    double* o64 = cam.out_format == RTC_OUT_F64_RGB ? (double*)d_rgb : nullptr;
    float* o32 = cam.out_format == RTC_OUT_F32_RGB ? (float*)d_rgb : nullptr;
    uint8_t* o8 = cam.out_format == RTC_OUT_U8_RGB ? (uint8_t*)d_rgb : nullptr;
    size_t smem = (size_t)(scene.n_spheres + scene.n_cubes) * (6 * 16 + 4) + (size_t)scene.n_planes * (2 * 16 + 4) + (size_t)scene.n_bound4 * 16 + (size_t)(scene.n_bound4 > 0 ? ((scene.n_bound4 >> 3) + 3) / 4 * 4 : 0) * 16 + 16;
    auto launch = [&](auto kernel, size_t bytes) -> cudaError_t {
        if (bytes) {
            cudaError_t e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
            if (e != cudaSuccess) return e;
        }
        kernel<<<grid, kThreads, bytes, stream>>>(scene, cam, o64, o32, o8, d_hit, d_counters);
        return cudaSuccess;
    };
    const bool cubes = scene.n_cubes > 0, resample = cam.samples >= 2 || scene.n_lights > 1;
    cudaError_t e;
    auto pick = [&](auto in_smem, size_t bytes) -> cudaError_t {
        constexpr bool kS = decltype(in_smem)::value;
        if (resample) return cubes ? launch(render_f64_kernel<kS, true, true>, bytes) : launch(render_f64_kernel<kS, false, true>, bytes);
        return cubes ? launch(render_f64_kernel<kS, true, false>, bytes) : launch(render_f64_kernel<kS, false, false>, bytes);
    };
    // Streams in shared memory + the flat search (four bounding spheres per trip) for small scenes; from ~64 spheres / cubes on
    // the hierarchical search over streams in global memory wins, and by a growing factor (1920x1080, n random spheres, flat /
    // hierarchical: n = 48 1.9 / 2.3 ms, 64 2.2 / 2.2, 100 2.8 / 2.2, 400 8.5 / 3.3, 1500 99.6 / 6.2: the flat search is
    // linear in n and large streams leave room for one block per SM; profiles/r02_mid_scenes.txt).  Same frames either way.
    const bool in_smem = scene.n_spheres + scene.n_cubes <= 64 && smem <= 48 * 1024;
    if (in_smem) e = pick(std::true_type{}, smem);
    else e = pick(std::false_type{}, 0);
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}

}  // namespace rtc
