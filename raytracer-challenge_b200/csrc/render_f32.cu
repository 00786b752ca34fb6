// F32 FAST kernel (compiled with --use_fast_math; FMA contraction on).
//
// Same path as render_f64.cu — ray_for_pixel, World::intersect over every shape, hit selection,
// prepare_computations, Phong lighting with a shadow ray, reflection + refraction with Schlick
// (src/camera.rs:64-114, src/shape.rs:677-781, src/material.rs:319-361) — re-designed for
// throughput instead of bit parity:
//   * shape tests are done in WORLD space (planes as n.x+w, similarity-transformed spheres as
//     centre/radius; only genuinely affine spheres go through their inverse), so the brute-force
//     loop is almost pure FFMA;
//   * the recursion is unrolled into path weights: every factor the reference multiplies a child
//     colour by (reflectiveness, transparency, Schlick reflectance; src/shape.rs:692-699,736,765)
//     is a scalar, so colour = sum over tree nodes of (path weight x surface colour) and a node
//     needs no frame — only pending refraction rays sit on a small per-lane stack;
//   * persistent warps: each lane is a little state machine doing exactly ONE World::intersect per
//     loop iteration (a radiance ray or the shadow ray of its pending hit), and lanes that run out
//     of rays pull the next pixel from the warp's batch with a ballot/popc rank, so reflected,
//     refracted and shadow rays of different pixels share the same brute-force loop instead of
//     idling each other (north-star item 2);
//   * EPSILON: the reference offsets over_point/under_point by 1e-8, below f32 resolution
//     (SURVEY.md F4).  Fast mode uses kEps32 = 1e-4 (the book's value) for the same purpose.
#include "rtc_common.h"
#include "parity64.cuh"

namespace rtc {
namespace {

constexpr float kEps32 = 1e-4f;
constexpr int kThreads = 128;
constexpr int kTileX = 8, kTileY = 4;
constexpr int kBatch = kTileX * kTileY;  // pixels a warp claims per atomic: one 8x4 tile
constexpr int kMaxStack = RTC_MAX_DEPTH;
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
__device__ __forceinline__ float rcp(float x) { float r; asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x)); return r; }  // MUFU.RCP
__device__ __forceinline__ float row3(float4 r, F3 p) { return fmaf(r.z, p.z, fmaf(r.y, p.y, r.x * p.x)); }

struct Streams32 {
    const float4* planes; const int32_t* pl_index; int np;
    const float4* usph;   const int32_t* us_index; int nu;
    const float4* gsph;   const int32_t* gs_index; int ng;   // general spheres [0, ng), then cubes [ng, ng + nc): 3 inverse rows each
    int nc;
    const float4* bound; int nb4; float cmax, rmax;   // bounding-sphere filter (DevScene32::bound), kFilter kernels only
    const float4* gbound; float gcmax, grmax;         // one bound per eight entries of `bound` (DevScene32::gbound)
    const float4* sbound; float scmax, srmax;         // one bound per 32 groups (DevScene32::sbound)
    const float4* gbox; const float4* sbox;           // the same two levels as axis-aligned boxes (lo, hi interleaved)
};

// Cube::intersect_local (src/shape.rs:540-591) in object space: slab test per axis.  A zero direction
// component gives +-inf (or NaN for a ray in a face plane, which fminf / fmaxf skip, like f64::max / min).
__device__ __forceinline__ bool cube_slab32(const float4* r, F3 o, F3 d, float& tmin, float& tmax) {
    const float4 r0 = r[0], r1 = r[1], r2 = r[2];
    const F3 lo = mk(row4(r0, o), row4(r1, o), row4(r2, o));
    const F3 ld = mk(row3(r0, d), row3(r1, d), row3(r2, d));
    const float ix = rcp(ld.x), iy = rcp(ld.y), iz = rcp(ld.z);
    const float x0 = (-1.f - lo.x) * ix, x1 = (1.f - lo.x) * ix;
    const float y0 = (-1.f - lo.y) * iy, y1 = (1.f - lo.y) * iy;
    const float z0 = (-1.f - lo.z) * iz, z1 = (1.f - lo.z) * iz;
    tmin = fmaxf(fminf(x0, x1), fmaxf(fminf(y0, y1), fminf(z0, z1)));
    tmax = fminf(fmaxf(x0, x1), fminf(fmaxf(y0, y1), fmaxf(z0, z1)));
    return tmin < tmax;
}

// World::intersect + get_hit for one ray: nearest t in [0, best_t).  best_t comes in as the
// upper bound (+inf for radiance rays, 1 for shadow rays whose direction is light - point, which
// is the reference's `h.t < distance` in units of the unnormalised direction, src/shape.rs:716-727).
// `hit` is a position in the concatenated stream order planes | uniform spheres | general spheres.
// kCollect (scenes with aliased world_ids): every list entry with t < 0 that the loops come across is
// appended to `neg` as (t, 2 * stream position + root) — Intersection::compute_refractive needs exactly
// those entries (src/shape.rs:115-141), and the line of a ray meets only a few of 10^4 spheres, so keeping
// them costs a few instructions inside the rare "line meets sphere" branch instead of two more passes
// over every shape.  n_neg counts all of them; entries beyond kNegCap are dropped and the caller falls
// back to the passes.
constexpr int kNegCap = 48;

// Bounding-sphere filter (kFilter kernels: 10^3+ shapes, streams in global memory) — a RAY query.  |(c - o) x d^| is
// the distance of a centre from the ray's line; with m = o x d^ it is |c x d^ - m|, six FFMA per shape and three more
// for the squared length minus R^2 (the stream stores -R^2).  s = (c - o) . d^ is the centre's position along the ray
// (three more FFMA): a shape can only matter to a query interested in positions [lo, hi] along the ray when the
// interval [s - R, s + R] reaches into it, i.e. when q = max(lo - s, s - hi, 0) has q^2 <= R^2.
//   radiance rays, World::intersect + get_hit (src/shape.rs:677-683, 195-232): [0, current nearest hit] — hi shrinks
//     while the sweep runs (the exact test of a candidate updates it), so everything behind a hit already found and
//     everything behind the origin is skipped;
//   shadow rays, is_shadowed_by_light (:716-727): [0, distance to the light];
//   the negative-t entries Intersection::compute_refractive needs (:115-141): (-inf, 0], run only for hits on a
//     transparent material.
// `exact(i)` runs for every entry i of the bound stream that passes.  E bounds the error of these tests plus that of
// the exact tests (each a few dozen ulps of |c| + |o|), so no shape an exact test would accept for the interval is
// skipped and the image is the one the unfiltered loops produce (RTC_OPT_NO_BOUND_FILTER renders that one for
// comparison).  A NaN never culls.  Four shapes per trip, one branch per trip: the exact tests sit behind it.
struct RayQ {
    F3 dh, m;        // unit direction, o x dh
    float od;        // o . dh
    float len;       // |d|: position along the ray = t * len
    float E;         // absolute error bound of the positions computed here
    float margin;    // of the entry-level sphere test (squared-distance units)
    F3 inv, c0, c1;  // box levels: position of the ray in the plane x_a = b is fmaf(b, inv_a, c_a); c0 / c1 carry the box's
                     // lower / upper faces moved outwards by E
};
__device__ __forceinline__ RayQ make_rayq(const Streams32& S, F3 o, F3 d) {
    RayQ q;
    const float dd = dot(d, d), il = rsqrtf(dd);
    q.dh = d * il;
    q.len = dd * il;
    q.m = mk(o.y * q.dh.z - o.z * q.dh.y, o.z * q.dh.x - o.x * q.dh.z, o.x * q.dh.y - o.y * q.dh.x);
    q.od = dot(o, q.dh);
    q.E = 3.8e-6f * (S.cmax + S.rmax + fabsf(o.x) + fabsf(o.y) + fabsf(o.z));
    q.margin = fmaf(q.E, 2.f * S.rmax + q.E, 2e-6f * S.rmax * S.rmax);
    // a direction component of (almost) zero keeps a finite reciprocal: the slab then holds the whole ray or none of it
    auto inv_of = [](float v) { return rcp(copysignf(fmaxf(fabsf(v), 1e-30f), v)); };
    q.inv = mk(inv_of(q.dh.x), inv_of(q.dh.y), inv_of(q.dh.z));
    q.c0 = mk(-(o.x + q.E) * q.inv.x, -(o.y + q.E) * q.inv.y, -(o.z + q.E) * q.inv.z);
    q.c1 = mk(-(o.x - q.E) * q.inv.x, -(o.y - q.E) * q.inv.y, -(o.z - q.E) * q.inv.z);
    return q;
}
// ---- warp-cooperative queries (kFilter kernels) -------------------------------------------------------------------
// Thirty-two rays walking the hierarchy side by side diverge: every lane has other candidate blocks, groups and
// members, and a warp pays for the union (round 1: 12 of 32 lanes busy in the member and exact tests).  Here the warp
// takes its rays ONE AT A TIME and spends its 32 lanes on the hierarchy instead: lane j tests block j, then group j of
// the chosen block, then member (j & 7) of four candidate groups at once; the exact tests of a step run on the lanes
// that found the candidates.  Every bound is loaded once, as coalesced rows, and the loop nest is uniform.
// The two upper levels are axis-aligned boxes (slab test against the query's interval of positions [lo, hi] along the
// ray), the entries bounding spheres (distance from the ray's line + position, as before).  Blocks are visited NEAREST
// FIRST: each lane turns the position where the ray enters its block into a key (float bits, low 6 bits = block slot),
// a warp min picks the next block, and the walk stops when that block starts beyond `hi` — for radiance rays the
// nearest hit so far, which the exact tests keep shrinking.
// A ray's query constants travel through shared memory: its lane publishes them, all lanes read them back (broadcast).
// The slots are addressed through explicit 32-bit shared-window addresses: indexing generic pointers made ptxas rebuild
// the window base inside every divergent region (7 % of the executed instructions in the first version).
struct WarpSlots { unsigned base; unsigned* parity; };   // byte address of the warp's slab; the containers query's parity words (generic)
constexpr unsigned kOffQa = 0, kOffQb = 512, kOffQc = 1024, kOffQd = 1536, kOffQe = 2048, kOffQf = 2560;   // RayQ: (dh, od) (m, len) (E, margin, -, -) inv c0 c1
constexpr unsigned kOffD0 = 3072, kOffD1 = 3584, kOffD2 = 4096;                                          // the ray in f64: (ox oy) (oz dx) (dy dz)
constexpr unsigned kOffSo = 4608, kOffSd = 5120;                                                          // shadow ray in f32: (o, 0) (d, 0)
constexpr unsigned kOffQueue = 5632, kOffParity = 5760;                                                   // candidate groups of the block being visited; parity words
constexpr unsigned kOffKept = 6144;      // the negative-t entries the containers query keeps: kKeptCap x (t : f64, (id, ord) : 2 x s32)
constexpr int kKeptCap = 64;
constexpr int kWarpSlotBytes = 6144 + kKeptCap * 16;
__device__ __forceinline__ float4 slds4(unsigned addr) {
    float4 v;
    asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void ssts4(unsigned addr, float4 v) {
    asm volatile("st.shared.v4.f32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ double2 sldsd2(unsigned addr) {
    double2 v;
    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sstsd2(unsigned addr, double2 v) {
    asm volatile("st.shared.v2.f64 [%0], {%1, %2};" ::"r"(addr), "d"(v.x), "d"(v.y) : "memory");
}
__device__ __forceinline__ int sldsi(unsigned addr) {
    int v;
    asm volatile("ld.shared.s32 %0, [%1];" : "=r"(v) : "r"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void sstsi(unsigned addr, int v) { asm volatile("st.shared.s32 [%0], %1;" ::"r"(addr), "r"(v) : "memory"); }
__device__ __forceinline__ WarpSlots warp_slots(void* smem_base) {
    WarpSlots W;
    char* base = reinterpret_cast<char*>(smem_base) + (threadIdx.x >> 5) * kWarpSlotBytes;
    W.base = (unsigned)__cvta_generic_to_shared(base);
    W.parity = reinterpret_cast<unsigned*>(base + kOffParity);
    return W;
}
__device__ __forceinline__ void publish_rayq(const WarpSlots& W, int lane, const RayQ& q) {
    const unsigned a = W.base + 16u * (unsigned)lane;
    ssts4(a + kOffQa, make_float4(q.dh.x, q.dh.y, q.dh.z, q.od));
    ssts4(a + kOffQb, make_float4(q.m.x, q.m.y, q.m.z, q.len));
    ssts4(a + kOffQc, make_float4(q.E, q.margin, 0.f, 0.f));
    ssts4(a + kOffQd, make_float4(q.inv.x, q.inv.y, q.inv.z, 0.f));
    ssts4(a + kOffQe, make_float4(q.c0.x, q.c0.y, q.c0.z, 0.f));
    ssts4(a + kOffQf, make_float4(q.c1.x, q.c1.y, q.c1.z, 0.f));
}
__device__ __forceinline__ RayQ load_rayq(const WarpSlots& W, int w) {
    const unsigned x = W.base + 16u * (unsigned)w;
    const float4 a = slds4(x + kOffQa), b = slds4(x + kOffQb), c = slds4(x + kOffQc), d = slds4(x + kOffQd), e = slds4(x + kOffQe), f = slds4(x + kOffQf);
    RayQ q;
    q.dh = mk(a.x, a.y, a.z); q.od = a.w; q.m = mk(b.x, b.y, b.z); q.len = b.w;
    q.E = c.x; q.margin = c.y;
    q.inv = mk(d.x, d.y, d.z); q.c0 = mk(e.x, e.y, e.z); q.c1 = mk(f.x, f.y, f.z);
    return q;
}
__device__ __forceinline__ void publish_ray64(const WarpSlots& W, int lane, p64::V3 o, p64::V3 d) {
    const unsigned a = W.base + 16u * (unsigned)lane;
    sstsd2(a + kOffD0, make_double2(o.x, o.y)); sstsd2(a + kOffD1, make_double2(o.z, d.x)); sstsd2(a + kOffD2, make_double2(d.y, d.z));
}
__device__ __forceinline__ void load_ray64(const WarpSlots& W, int w, p64::V3& o, p64::V3& d) {
    const unsigned x = W.base + 16u * (unsigned)w;
    const double2 a = sldsd2(x + kOffD0), b = sldsd2(x + kOffD1), c = sldsd2(x + kOffD2);
    o = p64::mk(a.x, a.y, b.x); d = p64::mk(b.y, c.x, c.y);
}

// One ray (Q, warp-uniform), all 32 lanes.  [lo, hi]: the positions along the ray the query cares about (rounding margin E
// included by the caller); `hi` is read again before every test, so `step` may shrink it.  step(c, i) is called by ALL
// lanes once per member step — c says this lane holds candidate entry i of the bound stream — and returns true
// (warp-uniform) to end the query.  `queue`: shared-window address of 32 ints of the warp's slab.
template <typename Step>
__device__ __forceinline__ void warp_query(const Streams32& S, const RayQ& Q, float lo, float& hi, unsigned queue, Step&& step) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const F3 dh = Q.dh, m = Q.m;
    // entry level: bounding sphere (cx, cy, cz, -R^2) — the ray's line passes within R, and [s - R, s + R] reaches into [lo, hi]
    auto sphere_test = [&](float4 b) -> bool {
        const float vx = fmaf(-b.z, dh.y, fmaf(b.y, dh.z, -m.x));
        const float vy = fmaf(-b.x, dh.z, fmaf(b.z, dh.x, -m.y));
        const float vz = fmaf(-b.y, dh.x, fmaf(b.x, dh.y, -m.z));
        const float gap = fmaf(vz, vz, fmaf(vy, vy, fmaf(vx, vx, b.w)));
        const float s = fmaf(b.z, dh.z, fmaf(b.y, dh.y, fmaf(b.x, dh.x, -Q.od)));
        const float q = fmaxf(fmaxf(lo - s, s - hi), 0.f);
        return !(gap > Q.margin) && !(fmaf(q, q, b.w) > 0.f);   // a NaN never culls; padding entries (w = +inf) never pass
    };
    // upper levels: axis-aligned box, faces moved outwards by E — the positions at which the ray is inside all three slabs,
    // cut to [lo, hi]; returns whether any are left and where they start
    auto box_test = [&](float4 bl, float4 bh, float& enter) -> bool {
        const float x0 = fmaf(bl.x, Q.inv.x, Q.c0.x), x1 = fmaf(bh.x, Q.inv.x, Q.c1.x);
        const float y0 = fmaf(bl.y, Q.inv.y, Q.c0.y), y1 = fmaf(bh.y, Q.inv.y, Q.c1.y);
        const float z0 = fmaf(bl.z, Q.inv.z, Q.c0.z), z1 = fmaf(bh.z, Q.inv.z, Q.c1.z);
        enter = fmaxf(fmaxf(fminf(x0, x1), fminf(y0, y1)), fmaxf(fminf(z0, z1), lo));
        const float leave = fminf(fminf(fmaxf(x0, x1), fmaxf(y0, y1)), fminf(fmaxf(z0, z1), hi));
        return !(enter > leave);
    };
    const int n_groups = S.nb4 >> 3;              // `bound` is padded to whole groups of eight
    const int n_blocks = (n_groups + 31) >> 5;
    auto block_key = [&](int slot, int blk) -> unsigned {   // where the ray enters the block, rounded down, | its slot
        if (blk >= n_blocks) return 0xffffffffu;
        float enter;
        if (!box_test(S.sbox[2 * blk], S.sbox[2 * blk + 1], enter)) return 0xffffffffu;
        return (__float_as_uint(fmaxf(enter, 0.f)) & ~63u) | (unsigned)slot;
    };
    for (int sb0 = 0; sb0 < n_blocks; sb0 += 64) {
        unsigned k0 = block_key(lane, sb0 + lane);
        unsigned k1 = block_key(32 + lane, sb0 + 32 + lane);
        for (;;) {
            const unsigned mine = min(k0, k1);
            const unsigned k = __reduce_min_sync(full, mine);
            if (k == 0xffffffffu) break;
            if (__uint_as_float(k & ~63u) > hi) break;       // every block left starts beyond the interval
            if (mine == k) { if (k0 == k) k0 = 0xffffffffu; else k1 = 0xffffffffu; }
            const int gidx = 32 * (sb0 + (int)(k & 63u)) + lane;      // lane j <-> group j of the block
            float ge;
            RTC_CHECK(gidx >= 0 && gidx < ((n_groups + 31) & ~31));
            const bool gp = gidx < n_groups && box_test(S.gbox[2 * gidx], S.gbox[2 * gidx + 1], ge);
            const unsigned gm = __ballot_sync(full, gp);
            if (gm == 0u) continue;
            // the candidate groups, compacted into the warp's queue; then four groups per step: lane l tests member (l & 7)
            const int n_c = __popc(gm);
            if (gp) sstsi(queue + 4u * (unsigned)__popc(gm & ((1u << lane) - 1u)), gidx);
            __syncwarp();
            for (int s0 = 0; s0 < n_c; s0 += 4) {
                const int qi = s0 + (lane >> 3);
                bool c = false;
                int i = 0;
                if (qi < n_c) { i = 8 * sldsi(queue + 4u * (unsigned)qi) + (lane & 7); RTC_CHECK(i >= 0 && i < S.nb4); c = sphere_test(S.bound[i]); }
                if (step(c, i)) return;
            }
            __syncwarp();   // the queue is free again
        }
    }
}

template <bool kCollect, bool kCubes>
__device__ __forceinline__ void intersect32(const Streams32& S, F3 o, F3 d, float& best_t, int& hit, float2* neg, int& n_neg) {
    hit = -1;
    if (kCollect) n_neg = 0;
    auto keep = [&](float t, int code) { if (n_neg < kNegCap) neg[n_neg] = make_float2(t, __int_as_float(code)); ++n_neg; };
    const float a = dot(d, d);
    const float inv_a = rcp(a);
#pragma unroll 2
    for (int i = 0; i < S.np; ++i) {                     // Plane::intersect_local, src/shape.rs:462-471
        float4 pl = S.planes[i];
        float dy = row3(pl, d);
        float oy = row4(pl, o);
        float t = -oy * rcp(dy);
        if (t >= 0.f && t < best_t) { best_t = t; hit = i; }  // dy ~ 0 -> t = +-inf or NaN: never taken
    }
    auto usph = [&](int i) {                             // Sphere::intersect_local, src/shape.rs:361-375
        float4 sp = S.usph[i];
        F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
        // b^2 - a c cancels catastrophically in f32 for a small sphere far away (b^2 ~ c ~ 10^3, r^2 ~ 10^-2):
        // use the distance of the centre from the ray instead, disc / a = r^2 - |L - (b/a) d|^2
        float bia = dot(d, L) * inv_a;                   // half of the reference's b, over a
        F3 q = mk(fmaf(-bia, d.x, L.x), fmaf(-bia, d.y, L.y), fmaf(-bia, d.z, L.z));
        float dn = fmaf(-q.z, q.z, fmaf(-q.y, q.y, fmaf(-q.x, q.x, sp.w)));
        if (dn >= 0.f) {                                 // most rays miss most spheres: skip the roots
            float s = sqrtf(dn * inv_a);
            float t1 = -bia - s, t2 = s - bia;
            float t = t1 >= 0.f ? t1 : t2;
            if (t >= 0.f && t < best_t) { best_t = t; hit = S.np + i; }
            if (kCollect && t1 < 0.f) { keep(t1, 2 * (S.np + i)); if (t2 < 0.f) keep(t2, 2 * (S.np + i) + 1); }
        }
    };
    auto gsph = [&](int i) {                             // Shape::intersect: ray.transform(inverse) first
        float4 r0 = S.gsph[3 * i], r1 = S.gsph[3 * i + 1], r2 = S.gsph[3 * i + 2];
        F3 lo = mk(row4(r0, o), row4(r1, o), row4(r2, o));
        F3 ld = mk(row3(r0, d), row3(r1, d), row3(r2, d));
        float inv = rcp(dot(ld, ld)), bia = dot(ld, lo) * inv;
        F3 q = mk(fmaf(-bia, ld.x, lo.x), fmaf(-bia, ld.y, lo.y), fmaf(-bia, ld.z, lo.z));   // closest approach to the centre
        float dn = fmaf(-q.z, q.z, fmaf(-q.y, q.y, fmaf(-q.x, q.x, 1.f)));                    // disc / la, without cancellation
        float s = sqrtf(fmaxf(dn, 0.f) * inv);
        float t1 = -bia - s, t2 = s - bia;
        float t = t1 >= 0.f ? t1 : t2;
        if (dn >= 0.f && t >= 0.f && t < best_t) { best_t = t; hit = S.np + S.nu + i; }
        if (kCollect && dn >= 0.f && t1 < 0.f) { keep(t1, 2 * (S.np + S.nu + i)); if (t2 < 0.f) keep(t2, 2 * (S.np + S.nu + i) + 1); }
    };
    auto cube = [&](int i) {                             // Cube
        float t1, t2;
        if (cube_slab32(S.gsph + 3 * i, o, d, t1, t2)) {
            float t = t1 >= 0.f ? t1 : t2;
            if (t >= 0.f && t < best_t) { best_t = t; hit = S.np + S.nu + i; }
            if (kCollect && t1 < 0.f) { keep(t1, 2 * (S.np + S.nu + i)); if (t2 < 0.f) keep(t2, 2 * (S.np + S.nu + i) + 1); }
        }
    };
#pragma unroll 2
    for (int i = 0; i < S.nu; ++i) usph(i);
    for (int i = 0; i < S.ng; ++i) gsph(i);
    if (kCubes) for (int i = S.ng; i < S.ng + S.nc; ++i) cube(i);
}

__device__ __forceinline__ int stream_to_shape(const Streams32& S, int pos) {
    if (pos < S.np) return S.pl_index[pos];
    pos -= S.np;
    if (pos < S.nu) return S.us_index[pos];
    return S.gs_index[pos - S.nu];
}

// Intersection::compute_refractive (src/shape.rs:115-141), list-free, as in render_f64.cu:
// containers = per world_id with an odd number of negative-t entries, that id's last entry.
struct Cont32 { float t; int ord; int idx; int id; };
__device__ __forceinline__ bool later32(const Cont32& a, const Cont32& b) { return a.t > b.t || (a.t == b.t && a.ord > b.ord); }
__device__ __forceinline__ void offer32(Cont32 e, Cont32& top1, Cont32& top2) {
    if (top1.idx >= 0 && e.id == top1.id) { if (later32(e, top1)) top1 = e; return; }
    if (top2.idx >= 0 && e.id == top2.id) {
        if (later32(e, top2)) top2 = e;
        if (top1.idx < 0 || later32(top2, top1)) { Cont32 s = top1; top1 = top2; top2 = s; }
        return;
    }
    if (top1.idx < 0 || later32(e, top1)) { top2 = top1; top1 = e; }
    else if (top2.idx < 0 || later32(e, top2)) top2 = e;
}

// Calls f(t, shape_index, root, single) for every list entry of the ray with t < 0; `single` says the
// shape has exactly one negative entry (it contains the ray origin).
template <typename F>
__device__ __forceinline__ void for_each_negative(const Streams32& S, F3 o, F3 d, F f) {
    const float a = dot(d, d);
    const float inv_a = rcp(a);
    for (int i = 0; i < S.np; ++i) {
        float4 pl = S.planes[i];
        float dy = row3(pl, d), oy = row4(pl, o);
        float t = -oy * rcp(dy);
        if (t < 0.f) f(t, S.pl_index[i], 0, true);
    }
    for (int i = 0; i < S.nu; ++i) {
        float4 sp = S.usph[i];
        F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
        float b = dot(d, L);
        float c = fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w)));
        float disc = fmaf(b, b, -a * c);
        if (disc >= 0.f) {
            float s = sqrtf(disc);
            float t1 = (-b - s) * inv_a, t2 = (s - b) * inv_a;
            if (t1 < 0.f) f(t1, S.us_index[i], 0, !(t2 < 0.f));
            if (t2 < 0.f) f(t2, S.us_index[i], 1, false);
        }
    }
    for (int i = 0; i < S.ng; ++i) {
        float4 r0 = S.gsph[3 * i], r1 = S.gsph[3 * i + 1], r2 = S.gsph[3 * i + 2];
        F3 lo = mk(row4(r0, o), row4(r1, o), row4(r2, o));
        F3 ld = mk(row3(r0, d), row3(r1, d), row3(r2, d));
        float la = dot(ld, ld), b = dot(ld, lo), c = dot(lo, lo) - 1.f;
        float disc = fmaf(b, b, -la * c);
        if (disc >= 0.f) {
            float s = sqrtf(disc), inv = rcp(la);
            float t1 = (-b - s) * inv, t2 = (s - b) * inv;
            if (t1 < 0.f) f(t1, S.gs_index[i], 0, !(t2 < 0.f));
            if (t2 < 0.f) f(t2, S.gs_index[i], 1, false);
        }
    }
    for (int i = S.ng; i < S.ng + S.nc; ++i) {
        float t1, t2;
        if (cube_slab32(S.gsph + 3 * i, o, d, t1, t2)) {
            if (t1 < 0.f) f(t1, S.gs_index[i], 0, !(t2 < 0.f));
            if (t2 < 0.f) f(t2, S.gs_index[i], 1, false);
        }
    }
}

__device__ __noinline__ void compute_refractive32(const Streams32& S, const ShadeRec32* rec, F3 o, F3 d, int hit_idx,
                                                  float& n1, float& n2) {
    unsigned parity[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) parity[k] = 0u;
    for_each_negative(S, o, d, [&](float, int idx, int, bool) {
        int id = __float_as_int(rec[idx].pa.w) & 255;
        parity[id >> 5] ^= 1u << (id & 31);
    });
    Cont32 top1, top2;
    top1.idx = -1; top2.idx = -1; top1.t = top2.t = 0.f; top1.ord = top2.ord = 0; top1.id = top2.id = -1;
    for_each_negative(S, o, d, [&](float t, int idx, int root, bool) {
        int id = __float_as_int(rec[idx].pa.w) & 255;
        if ((parity[id >> 5] >> (id & 31)) & 1u) {
            Cont32 e; e.t = t; e.ord = 2 * idx + root; e.idx = idx; e.id = id;
            offer32(e, top1, top2);
        }
    });
    n1 = top1.idx >= 0 ? rec[top1.idx].tr_ri.y : 1.0f;
    int hid = __float_as_int(rec[hit_idx].pa.w) & 255;
    bool hit_is_container = (parity[hid >> 5] >> (hid & 31)) & 1u;
    if (hit_is_container) {
        if (top1.idx >= 0 && top1.id == hid) n2 = top2.idx >= 0 ? rec[top2.idx].tr_ri.y : 1.0f;
        else n2 = top1.idx >= 0 ? rec[top1.idx].tr_ri.y : 1.0f;
    } else {
        n2 = rec[hit_idx].tr_ri.y;
    }
}

// The same, from the negative entries phase A kept for this ray (intersect32<true>) plus the planes.
__device__ __noinline__ void compute_refractive32_kept(const Streams32& S, const ShadeRec32* rec, F3 o, F3 d, int hit_idx,
                                                       const float2* neg, int n_neg, float& n1, float& n2) {
    unsigned parity[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) parity[k] = 0u;
    auto for_each = [&](auto f) {
        for (int i = 0; i < S.np; ++i) {
            float4 pl = S.planes[i];
            float t = -row4(pl, o) * rcp(row3(pl, d));
            if (t < 0.f) f(t, S.pl_index[i], 0);
        }
        for (int k = 0; k < n_neg; ++k) {
            const float2 e = neg[k];
            const int code = __float_as_int(e.y);
            f(e.x, stream_to_shape(S, code >> 1), code & 1);
        }
    };
    for_each([&](float, int idx, int) {
        int id = __float_as_int(rec[idx].pa.w) & 255;
        parity[id >> 5] ^= 1u << (id & 31);
    });
    Cont32 top1, top2;
    top1.idx = -1; top2.idx = -1; top1.t = top2.t = 0.f; top1.ord = top2.ord = 0; top1.id = top2.id = -1;
    for_each([&](float t, int idx, int root) {
        int id = __float_as_int(rec[idx].pa.w) & 255;
        if ((parity[id >> 5] >> (id & 31)) & 1u) {
            Cont32 e; e.t = t; e.ord = 2 * idx + root; e.idx = idx; e.id = id;
            offer32(e, top1, top2);
        }
    });
    n1 = top1.idx >= 0 ? rec[top1.idx].tr_ri.y : 1.0f;
    int hid = __float_as_int(rec[hit_idx].pa.w) & 255;
    bool hit_is_container = (parity[hid >> 5] >> (hid & 31)) & 1u;
    if (hit_is_container) {
        if (top1.idx >= 0 && top1.id == hid) n2 = top2.idx >= 0 ? rec[top2.idx].tr_ri.y : 1.0f;
        else n2 = top1.idx >= 0 ? rec[top1.idx].tr_ri.y : 1.0f;
    } else {
        n2 = rec[hit_idx].tr_ri.y;
    }
}

// floor(x) % 2.0 == 0.0 (Rust % = fmod): is the integer-valued float v even?
__device__ __forceinline__ bool is_even(float v) { float h = v * 0.5f; return h == floorf(h); }

// Pattern::pattern_at (src/material.rs:67-69,97-103,133-135,163-170,198-205,233-241)
__device__ __forceinline__ F3 pattern_at32(int kind, F3 a, F3 b, F3 p) {
    switch (kind) {
        case RTC_PATTERN_TEST: return p;
        case RTC_PATTERN_STRIPE: return is_even(floorf(p.x)) ? a : b;
        case RTC_PATTERN_GRADIENT: return fma3(b - a, p.x - floorf(p.x), a);
        case RTC_PATTERN_RING: return is_even(floorf(sqrtf(fmaf(p.x, p.x, p.y * p.y)))) ? a : b;
        case RTC_PATTERN_CHECKER: return is_even(floorf(p.x) + floorf(p.y) + floorf(p.z)) ? a : b;
        default: return (fabsf(p.x - floorf(p.x)) < 0.01f || fabsf(p.z - floorf(p.z)) < 0.01f) ? b : a;
    }
}


// World::is_shadowed_by_light (src/shape.rs:716-727) as an ANY-hit test.  The reference takes the
// nearest hit of the shadow ray and asks `h.t < distance`; that is true iff some shape has a root
// in [0, distance).  With d = light - p the question is "a root in [0,1)", which for
// a plane is a sign change of f(t) = n.(p + t d) + w between t=0 and t=1, and for a sphere needs
// no square root: f(0) f(1) < 0 (one root inside) or both roots inside.  f(1) depends only on the
// light and is precomputed per shape (plL / usL, staged beside the streams).
// One entry i of the bound stream (uniform spheres | general spheres | cubes) against the segment o -> o + d: does it have a
// root in [0, 1)?  The per-shape terms of shadow32 below, as a predicate (kFilter kernels: the candidates of a query).
template <bool kCubes>
__device__ __forceinline__ bool shadow_exact(const Streams32& S, const float* __restrict__ usL, bool light_inside, int i, F3 o, F3 d, float inv_a) {
    if (i < S.nu) {
        const float4 sp = S.usph[i];
        const F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
        const float tc = __saturatef(-dot(d, L) * inv_a);
        const F3 q = fma3(d, tc, L);
        const float fmin = fmaf(q.z, q.z, fmaf(q.y, q.y, fmaf(q.x, q.x, -sp.w)));
        if (!light_inside) return fmin <= 0.f;
        const float f0 = fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w)));
        return fmaxf(fmin, -fmaxf(f0, usL[i])) <= 0.f;
    }
    const int k = i - S.nu;
    float t1, t2;
    if (!kCubes || k < S.ng) {
        const float4 r0 = S.gsph[3 * k], r1 = S.gsph[3 * k + 1], r2 = S.gsph[3 * k + 2];
        const F3 lo = mk(row4(r0, o), row4(r1, o), row4(r2, o));
        const F3 ld = mk(row3(r0, d), row3(r1, d), row3(r2, d));
        const float inv = rcp(dot(ld, ld)), bia = dot(ld, lo) * inv;
        const F3 q = mk(fmaf(-bia, ld.x, lo.x), fmaf(-bia, ld.y, lo.y), fmaf(-bia, ld.z, lo.z));
        const float dn = fmaf(-q.z, q.z, fmaf(-q.y, q.y, fmaf(-q.x, q.x, 1.f)));
        if (!(dn >= 0.f)) return false;
        const float sq = sqrtf(dn * inv);
        t1 = -bia - sq; t2 = sq - bia;
    } else if (!cube_slab32(S.gsph + 3 * k, o, d, t1, t2)) return false;
    const float t = t1 >= 0.f ? t1 : t2;
    return t >= 0.f && t < 1.f;
}

template <bool kCubes>
__device__ __forceinline__ bool shadow32(const Streams32& S, const float* __restrict__ plL, const float* __restrict__ usL,
                                         bool light_inside, F3 o, F3 d) {
    // every test is reduced to "g <= 0" and the g's are min-accumulated: FMNMX instead of predicates
    float g = 1.0f;
    const float a = dot(d, d);
    const float inv_a = rcp(a);
#pragma unroll 2
    for (int i = 0; i < S.np; ++i) {
        float4 pl = S.planes[i];
        g = fminf(g, row4(pl, o) * plL[i]);            // f(0) f(1) <= 0: the two ends on opposite sides
    }
    bool hit = false;
    auto usph = [&](int i) {
        float4 sp = S.usph[i];
        F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
        // f(t) = |L + t d|^2 - r^2 is convex: its minimum over [0,1] is at tc = clamp(-b/a, 0, 1); evaluated
        // as |L + tc d|^2 - r^2 (no cancellation between large terms)
        float tc = __saturatef(-dot(d, L) * inv_a);
        F3 q = fma3(d, tc, L);
        g = fminf(g, fmaf(q.z, q.z, fmaf(q.y, q.y, fmaf(q.x, q.x, -sp.w))));
    };
    auto usph_in = [&](int i) {  // the light sits inside some sphere: a segment entirely inside it (f0 < 0 and f1 < 0) has no root
        float4 sp = S.usph[i];
        F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
        float f0 = fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w)));
        float tc = __saturatef(-dot(d, L) * inv_a);
        F3 q = fma3(d, tc, L);
        float fmin = fmaf(q.z, q.z, fmaf(q.y, q.y, fmaf(q.x, q.x, -sp.w)));
        g = fminf(g, fmaxf(fmin, -fmaxf(f0, usL[i])));
    };
    if (!light_inside) {
#pragma unroll 2
        for (int i = 0; i < S.nu; ++i) usph(i);
    } else {
        for (int i = 0; i < S.nu; ++i) usph_in(i);
    }
    for (int i = S.nu; i < S.nu + S.ng + (kCubes ? S.nc : 0); ++i) hit |= shadow_exact<kCubes>(S, usL, light_inside, i, o, d, inv_a);
    return hit || g <= 0.f;
}

// kFilter kernels: the plane terms of the shadow test stay with the ray's lane ...
__device__ __forceinline__ bool shadow_planes32(const Streams32& S, const float* __restrict__ plL, F3 o) {
    float g = 1.0f;
#pragma unroll 2
    for (int i = 0; i < S.np; ++i) g = fminf(g, row4(S.planes[i], o) * plL[i]);
    return g <= 0.f;
}
// ... and the solids are an any-hit segment query by the whole warp: [0, |light - p|], over at the first occluder.
template <bool kCubes>
__device__ __forceinline__ bool coop_shadow(const Streams32& S, const float* __restrict__ usL, bool light_inside, const RayQ& Q, F3 o, F3 d, unsigned queue) {
    const float inv_a = rcp(dot(d, d));
    float hi = fmaf(Q.len, 1.000002f, Q.E);
    bool shadowed = false;
    warp_query(S, Q, -Q.E, hi, queue, [&](bool c, int i) -> bool {
        bool occ = false;
        if (c) occ = shadow_exact<kCubes>(S, usL, light_inside, i, o, d, inv_a);
        if (__any_sync(0xffffffffu, occ)) { shadowed = true; return true; }
        return false;
    });
    return shadowed;
}

// ---- kRefine kernels (scenes whose streams stay in global memory: 10^3+ shapes, e.g. config C4) ----------------------
// Thousands of small spheres are strongly diverging mirrors: a hit-point error of one f32 ulp is amplified ~10^3 per
// sphere-to-sphere bounce, so an all-f32 path leaves the reference's ray tree after two or three bounces.  These kernels
// therefore keep the GEOMETRY of the radiance rays in f64, in the reference's operation order (parity64.cuh, the
// statements the parity kernel runs): the ray (o, d) is carried in f64 from bounce to bounce; the f32 bounding-sphere
// hierarchy only selects candidates (a conservative filter), each candidate's exact test is Sphere / Cube / Plane::
// intersect_local in f64 on the stored inverse of the shape's f64 record, hit selection follows the insertion sort's tie
// rule, and hit point, normal, over / under point, reflected and refracted directions and n1 / n2 are the f64 ones.  The
// ray tree is then the parity kernel's, node for node; what stays in f32 is what does not feed back into geometry:
// the filter, the any-hit shadow test, patterns, Phong lighting, Schlick weights and the colour sums.
__device__ __forceinline__ void planes_mixed(const Streams32& S, const ShadeRec64* __restrict__ R64, p64::V3 od, p64::V3 dd, double& best_t, int& best_idx) {
    best_t = __longlong_as_double(0x7ff0000000000000LL);
    best_idx = 0x7fffffff;
    for (int i = 0; i < S.np; ++i) {                     // Plane::intersect_local, src/shape.rs:462-471
        const int idx = S.pl_index[i];
        double t;
        if (p64::plane_root(reinterpret_cast<const double2*>(R64[idx].inv + 4), od, dd, t)) p64::take(t, idx, best_t, best_idx);
    }
}
// entry i of the bound stream -> its shape, its two list entries (t1 <= t2)
template <bool kCubes>
__device__ __forceinline__ bool roots_mixed(const Streams32& S, const ShadeRec64* __restrict__ R64, int i, p64::V3 od, p64::V3 dd, int& idx, double& t1, double& t2) {
    RTC_CHECK(i >= 0 && i < S.nu + S.ng + S.nc);
    idx = i < S.nu ? S.us_index[i] : S.gs_index[i - S.nu];
    RTC_CHECK(idx >= 0 && idx < S.np + S.nu + S.ng + S.nc);
    const double2* rows = reinterpret_cast<const double2*>(R64[idx].inv);
    return (kCubes && i - S.nu >= S.ng) ? p64::cube_roots(rows, od, dd, t1, t2) : p64::sphere_roots(rows, od, dd, t1, t2);
}

// kGlobal == 1 (RTC_OPT_NO_BOUND_FILTER): every shape's exact test, per lane
template <bool kCubes>
__device__ __forceinline__ void intersect_mixed(const Streams32& S, const ShadeRec64* __restrict__ R64, p64::V3 od, p64::V3 dd,
                                                double& best_t, int& best_idx) {
    planes_mixed(S, R64, od, dd, best_t, best_idx);
    for (int i = 0; i < S.nu + S.ng + (kCubes ? S.nc : 0); ++i) {
        int idx; double t1, t2;
        if (roots_mixed<kCubes>(S, R64, i, od, dd, idx, t1, t2)) p64::take(t1 >= 0.0 ? t1 : t2, idx, best_t, best_idx);
    }
    if (best_idx == 0x7fffffff) best_idx = -1;
}

// kFilter: the solids of one ray by the whole warp — a forward query over [0, nearest hit so far].  best_t / best_idx come
// in as the planes' result and are warp-uniform throughout: every candidate's root is broadcast and taken by all lanes.
template <bool kCubes>
__device__ __forceinline__ void coop_hit(const Streams32& S, const ShadeRec64* __restrict__ R64, const RayQ& Q, p64::V3 od, p64::V3 dd,
                                         unsigned queue, double& best_t, int& best_idx) {
    const unsigned full = 0xffffffffu;
    // positions beyond the nearest hit found so far cannot hold the hit (an equal t loses or wins by shape index: kept by the margin)
    auto bound_of = [&](double t) { return fmaf(__double2float_ru(t) * Q.len, 1.000002f, Q.E); };
    float hi = bound_of(best_t);
    warp_query(S, Q, -Q.E, hi, queue, [&](bool c, int i) -> bool {
        double t = 0.0;
        int idx = -1;
        if (c) {
            int sidx; double t1, t2;
            if (roots_mixed<kCubes>(S, R64, i, od, dd, sidx, t1, t2)) {
                const double tt = t1 >= 0.0 ? t1 : t2;
                if (tt >= 0.0) { t = tt; idx = sidx; }
            }
        }
        unsigned m = __ballot_sync(full, idx >= 0);
        if (m) {
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1u;
                p64::take(__shfl_sync(full, t, src), __shfl_sync(full, idx, src), best_t, best_idx);
            }
            hi = bound_of(best_t);
        }
        return false;
    });
}

// Intersection::compute_refractive (src/shape.rs:115-141) for a hit on a transparent material, list-free as in the
// parity kernel (parity64.cuh): only entries with t < 0 matter, i.e. shapes that reach behind the ray origin — a second
// query over (-inf, 0], run twice: pass 0 takes the parity of the negative entries per world_id (u8 ids alias beyond 255
// shapes), pass 1 keeps the last two containers.
template <bool kCubes>
__device__ __noinline__ void containers_mixed(const Streams32& S, const ShadeRec64* __restrict__ R64, p64::V3 od, p64::V3 dd,
                                              int hit_idx, double& n1, double& n2) {
    unsigned parity[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) parity[k] = 0u;
    p64::Cont top1, top2;
    top1.idx = -1; top2.idx = -1; top1.t = top2.t = 0.0; top1.ord = top2.ord = 0; top1.id = top2.id = -1;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        auto entry = [&](double t, int idx, int root) {
            const int id = R64[idx].world_id & 255;
            if (pass == 0) parity[id >> 5] ^= 1u << (id & 31);
            else if ((parity[id >> 5] >> (id & 31)) & 1u) { p64::Cont e; e.t = t; e.ord = 2 * idx + root; e.idx = idx; e.id = id; p64::offer(e, top1, top2); }
        };
        for (int i = 0; i < S.np; ++i) {
            const int idx = S.pl_index[i];
            double t;
            if (p64::plane_root(reinterpret_cast<const double2*>(R64[idx].inv + 4), od, dd, t) && t < 0.0) entry(t, idx, 0);
        }
        for (int i = 0; i < S.nu + S.ng + (kCubes ? S.nc : 0); ++i) {
            int idx; double t1, t2;
            if (roots_mixed<kCubes>(S, R64, i, od, dd, idx, t1, t2)) { if (t1 < 0.0) entry(t1, idx, 0); if (t2 < 0.0) entry(t2, idx, 1); }
        }
    }
    n1 = top1.idx >= 0 ? R64[top1.idx].refractive_index : 1.0;
    const int hid = R64[hit_idx].world_id & 255;
    const bool hit_is_container = (parity[hid >> 5] >> (hid & 31)) & 1u;
    if (hit_is_container) {
        if (top1.idx >= 0 && top1.id == hid) n2 = top2.idx >= 0 ? R64[top2.idx].refractive_index : 1.0;
        else n2 = top1.idx >= 0 ? R64[top1.idx].refractive_index : 1.0;
    } else {
        n2 = R64[hit_idx].refractive_index;
    }
}

// The same by the whole warp for one ray (kFilter): planes one per lane, solids as a backward query; every negative entry
// is broadcast, the parity words live in the warp's shared memory (lane 0 toggles), the last two containers in
// warp-uniform registers.
template <bool kCubes>
__device__ __forceinline__ void coop_containers(const Streams32& S, const ShadeRec64* __restrict__ R64, const RayQ& Q, p64::V3 od, p64::V3 dd,
                                                int hit_idx, unsigned* parity, unsigned queue, unsigned kept, double& n1, double& n2) {
    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    int n_kept = 0;   // pass 0 also keeps the entries (warp-uniform count): with at most kKeptCap of them pass 1 reads them back instead of querying again
    __syncwarp();
    if (lane < 8) parity[lane] = 0u;
    __syncwarp();
    p64::Cont top1, top2;
    top1.idx = -1; top2.idx = -1; top1.t = top2.t = 0.0; top1.ord = top2.ord = 0; top1.id = top2.id = -1;
#pragma unroll 1
    for (int pass = 0; pass < 2; ++pass) {
        auto entry = [&](double t, int idx, int root) {     // called with warp-uniform arguments
            const int id = R64[idx].world_id & 255;
            if (pass == 0) {
                if (lane == 0) {
                    parity[id >> 5] ^= 1u << (id & 31);
                    if (n_kept < kKeptCap) sstsd2(kept + 16u * (unsigned)n_kept, make_double2(t, __hiloint2double(id, 2 * idx + root)));
                }
                ++n_kept;
            }
            else if ((parity[id >> 5] >> (id & 31)) & 1u) { p64::Cont e; e.t = t; e.ord = 2 * idx + root; e.idx = idx; e.id = id; p64::offer(e, top1, top2); }
        };
        if (pass == 1 && n_kept <= kKeptCap) {   // the kept entries, in the order pass 0 met them (the container rule is order-free)
            for (int k = 0; k < n_kept; ++k) {
                const double2 e = sldsd2(kept + 16u * (unsigned)k);
                const int ord = __double2loint(e.y), id = __double2hiint(e.y);
                if ((parity[id >> 5] >> (id & 31)) & 1u) { p64::Cont c; c.t = e.x; c.ord = ord; c.idx = ord >> 1; c.id = id; p64::offer(c, top1, top2); }
            }
            break;
        }
        for (int p0 = 0; p0 < S.np; p0 += 32) {
            double t = 0.0;
            int idx = 0;
            bool neg = false;
            if (p0 + lane < S.np) {
                idx = S.pl_index[p0 + lane];
                neg = p64::plane_root(reinterpret_cast<const double2*>(R64[idx].inv + 4), od, dd, t) && t < 0.0;
            }
            unsigned m = __ballot_sync(full, neg);
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1u;
                entry(__shfl_sync(full, t, src), __shfl_sync(full, idx, src), 0);
            }
        }
        float hi = Q.E;
        warp_query(S, Q, -__int_as_float(0x7f800000), hi, queue, [&](bool c, int i) -> bool {
            double t1 = 0.0, t2 = 0.0;
            int idx = 0;
            bool neg = false;
            if (c) neg = roots_mixed<kCubes>(S, R64, i, od, dd, idx, t1, t2) && t1 < 0.0;   // t1 <= t2
            unsigned m = __ballot_sync(full, neg);
            while (m) {
                const int src = __ffs(m) - 1;
                m &= m - 1u;
                const double a = __shfl_sync(full, t1, src), b = __shfl_sync(full, t2, src);
                const int is = __shfl_sync(full, idx, src);
                entry(a, is, 0);
                if (b < 0.0) entry(b, is, 1);
            }
            return false;
        });
        __syncwarp();   // pass 0's toggles are visible to every lane before pass 1 reads them
    }
    n1 = top1.idx >= 0 ? R64[top1.idx].refractive_index : 1.0;
    const int hid = R64[hit_idx].world_id & 255;
    const bool hit_is_container = (parity[hid >> 5] >> (hid & 31)) & 1u;
    if (hit_is_container) {
        if (top1.idx >= 0 && top1.id == hid) n2 = top2.idx >= 0 ? R64[top2.idx].refractive_index : 1.0;
        else n2 = top1.idx >= 0 ? R64[top1.idx].refractive_index : 1.0;
    } else {
        n2 = R64[hit_idx].refractive_index;
    }
}

// compute_refractive when every world_id is unique (<= 255 shapes): a shape is a container iff it
// has exactly one negative-t entry (plane: t < 0; sphere: t1 < 0 <= t2), so one pass that keeps the
// last two containers by (t, shape index) is enough.
// Fast path of compute_refractive: only shapes whose refractive index is not 1.0 ("gset", stream
// positions) can make n1/n2 differ from (1.0, ri(hit)): if the ray origin is inside none of them
// (no negative single entry), every container has ri 1.0, so n1 = 1 and n2 = ri(hit) either way.
__device__ __forceinline__ bool origin_in_gset(const Streams32& S, const int32_t* gset, int n_gset, F3 o, F3 d) {
    bool any = false;
    for (int k = 0; k < n_gset; ++k) {
        int pos = gset[k];
        if (pos < S.np) { float4 pl = S.planes[pos]; any |= row4(pl, o) * row3(pl, d) > 0.f; }      // t = -oy/dy < 0
        else if (pos < S.np + S.nu) {
            float4 sp = S.usph[pos - S.np];
            F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
            any |= fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w))) < 0.f;                      // origin inside
        } else if (pos < S.np + S.nu + S.ng) {
            const float4* r = S.gsph + 3 * (pos - S.np - S.nu);
            F3 lo = mk(row4(r[0], o), row4(r[1], o), row4(r[2], o));
            any |= dot(lo, lo) < 1.f;
        } else {                                                                                       // cube: one negative entry
            float t1, t2;
            any |= cube_slab32(S.gsph + 3 * (pos - S.np - S.nu), o, d, t1, t2) && t1 < 0.f && !(t2 < 0.f);
        }
    }
    return any;
}

__device__ __noinline__ void compute_refractive32_unique(const Streams32& S, const ShadeRec32* rec, F3 o, F3 d, int hit_idx,
                                                         float& n1, float& n2) {
    const float ninf = -__int_as_float(0x7f800000);
    float t1st = ninf, t2nd = ninf;
    int i1st = -1, i2nd = -1;
    bool hit_is_container = false;
    auto offer = [&](float t, int idx) {
        hit_is_container |= idx == hit_idx;
        if (t > t1st || (t == t1st && idx > i1st)) { t2nd = t1st; i2nd = i1st; t1st = t; i1st = idx; }
        else if (t > t2nd || (t == t2nd && idx > i2nd)) { t2nd = t; i2nd = idx; }
    };
    const float a = dot(d, d);
#pragma unroll 1
    for (int i = 0; i < S.np; ++i) {           // a plane behind the origin (t < 0) counts as a container
        float4 pl = S.planes[i];
        float t = -row4(pl, o) * rcp(row3(pl, d));
        if (t < 0.f) offer(t, S.pl_index[i]);
    }
#pragma unroll 1
    for (int i = 0; i < S.nu; ++i) {           // t1 < 0 <= t2  <=>  the origin is inside the sphere (c < 0)
        float4 sp = S.usph[i];
        F3 L = mk(o.x - sp.x, o.y - sp.y, o.z - sp.z);
        float c = fmaf(L.z, L.z, fmaf(L.y, L.y, fmaf(L.x, L.x, -sp.w)));
        if (c < 0.f) {
            float b = dot(d, L);
            offer((-b - sqrtf(fmaf(b, b, -a * c))) * rcp(a), S.us_index[i]);
        }
    }
#pragma unroll 1
    for (int i = 0; i < S.ng; ++i) {
        float4 r0 = S.gsph[3 * i], r1 = S.gsph[3 * i + 1], r2 = S.gsph[3 * i + 2];
        F3 lo = mk(row4(r0, o), row4(r1, o), row4(r2, o));
        float c = dot(lo, lo) - 1.f;
        if (c < 0.f) {
            F3 ld = mk(row3(r0, d), row3(r1, d), row3(r2, d));
            float la = dot(ld, ld), b = dot(ld, lo);
            offer((-b - sqrtf(fmaf(b, b, -la * c))) * rcp(la), S.gs_index[i]);
        }
    }
#pragma unroll 1
    for (int i = S.ng; i < S.ng + S.nc; ++i) {  // a cube is a container iff exactly one of its two entries is negative
        float t1, t2;
        if (cube_slab32(S.gsph + 3 * i, o, d, t1, t2) && t1 < 0.f && !(t2 < 0.f)) offer(t1, S.gs_index[i]);
    }
    n1 = i1st >= 0 ? rec[i1st].tr_ri.y : 1.0f;
    if (i1st == hit_idx) n2 = i2nd >= 0 ? rec[i2nd].tr_ri.y : 1.0f;
    else if (hit_is_container) n2 = i1st >= 0 ? rec[i1st].tr_ri.y : 1.0f;
    else n2 = rec[hit_idx].tr_ri.y;
}

#ifndef RTC_F32_MIN_BLOCKS
#define RTC_F32_MIN_BLOCKS 6
#endif
#ifndef RTC_F32_REFINE_MIN_BLOCKS
#define RTC_F32_REFINE_MIN_BLOCKS 4   // kRefine kernels carry the ray and the hit geometry in f64: 128 registers, like the parity kernel
#endif
// kSmem: shading records staged to shared memory; kUnique: world_ids are unique (<= 255 shapes);
// kGlobal: the intersect streams stay in global memory (read through L1/L2 as warp-uniform loads) —
// scenes of any size, e.g. the 10k-shape config C4; 2 = with the bounding-sphere hierarchy in front (warp_query).
// kCubes: the scene has cubes (object-space slab tests behind the general spheres); scenes without them compile none of it
// kGlobal != 0 is also kRefine: the geometry of the radiance rays in f64 (see intersect_mixed)
template <bool kSmem, bool kUnique, int kGlobal, bool kCubes>
__global__ void __launch_bounds__(kThreads, kGlobal ? RTC_F32_REFINE_MIN_BLOCKS : RTC_F32_MIN_BLOCKS)
render_f32_kernel(DevScene32 sc, CamParams cam, void* __restrict__ out_rgb, uint8_t* __restrict__ out_hit,
                  Counters* __restrict__ counters, unsigned* __restrict__ work) {
    extern __shared__ float4 smem4[];
    Streams32 S;
    S.np = sc.n_planes; S.nu = sc.n_usph; S.ng = sc.n_gsph; S.nc = sc.n_cube;
    const int n_obj = sc.n_gsph + sc.n_cube;   // object-space shapes: general spheres, then cubes
    const ShadeRec32* R = sc.rec;
    const float* plL;  // f(1) of the shadow test per plane / per uniform sphere (depends on the light only)
    const float* usL;
    const int32_t* gset;
    const int n_gset = sc.n_gset;
    const F3 light_pos = mk(sc.light_pos[0], sc.light_pos[1], sc.light_pos[2]);
    const F3 light_int = mk(sc.light_int[0], sc.light_int[1], sc.light_int[2]);
    constexpr bool kFilter = kGlobal == 2;
    constexpr bool kRefine = kGlobal != 0;
    const ShadeRec64* __restrict__ R64 = sc.rec64;
    const WarpSlots W = warp_slots(smem4);   // kFilter kernels: one published ray per lane (kWarpSlotBytes per warp of dynamic shared memory)
    S.bound = sc.bound; S.nb4 = sc.n_bound4; S.cmax = sc.bound_cmax; S.rmax = sc.bound_rmax;
    S.gbound = sc.gbound; S.gcmax = sc.gbound_cmax; S.grmax = sc.gbound_rmax;
    S.sbound = sc.sbound; S.scmax = sc.sbound_cmax; S.srmax = sc.sbound_rmax;
    S.gbox = sc.gbox; S.sbox = sc.sbox;
    if (kGlobal) {
        S.planes = sc.planes; S.usph = sc.usph; S.gsph = sc.gsph; S.pl_index = sc.pl_index; S.us_index = sc.us_index; S.gs_index = sc.gs_index;
        plL = sc.plL; usL = sc.usL; gset = sc.gset;
    } else {
        // stage the intersect streams (and, when small, the shading records): coalesced 16-byte loads
        float4* s_pl = smem4;
        float4* s_us = s_pl + S.np;
        float4* s_gs = s_us + S.nu;
        float4* s_rec = s_gs + 3 * n_obj;
        const int rec4 = kSmem ? (int)(sizeof(ShadeRec32) / 16) * sc.n_shapes : 0;
        int32_t* s_pi = reinterpret_cast<int32_t*>(s_rec + rec4);
        int32_t* s_ui = s_pi + S.np;
        int32_t* s_gi = s_ui + S.nu;
        float* s_plL = reinterpret_cast<float*>(s_gi + n_obj);
        float* s_usL = s_plL + S.np;
        int32_t* s_gset = reinterpret_cast<int32_t*>(s_usL + S.nu);
        for (int i = threadIdx.x; i < sc.n_gset && i < 16; i += kThreads) s_gset[i] = sc.gset[i];
        for (int i = threadIdx.x; i < S.np; i += kThreads) { s_pl[i] = sc.planes[i]; s_pi[i] = sc.pl_index[i]; s_plL[i] = sc.plL[i]; }
        for (int i = threadIdx.x; i < S.nu; i += kThreads) { s_us[i] = sc.usph[i]; s_ui[i] = sc.us_index[i]; s_usL[i] = sc.usL[i]; }
        for (int i = threadIdx.x; i < 3 * n_obj; i += kThreads) s_gs[i] = sc.gsph[i];
        for (int i = threadIdx.x; i < n_obj; i += kThreads) s_gi[i] = sc.gs_index[i];
        if (kSmem) {
            const float4* g_rec = reinterpret_cast<const float4*>(sc.rec);
            for (int i = threadIdx.x; i < rec4; i += kThreads) s_rec[i] = g_rec[i];
            R = reinterpret_cast<const ShadeRec32*>(s_rec);
        }
        __syncthreads();
        S.planes = s_pl; S.usph = s_us; S.gsph = s_gs; S.pl_index = s_pi; S.us_index = s_ui; S.gs_index = s_gi;
        plL = s_plL; usL = s_usL; gset = s_gset;
    }

    const unsigned full = 0xffffffffu;
    const int lane = threadIdx.x & 31;
    const unsigned lt_mask = (1u << lane) - 1u;
    const int tiles_x = ((int)cam.hsize + kTileX - 1) / kTileX;
    const int tiles_y = (cam.shard_rows + kTileY - 1) / kTileY;
    const unsigned total_tiles = (unsigned)(tiles_x * tiles_y);
    const int n_sub = cam.samples == 1 ? 1 : 4;
    // antialiasing_samples >= 2 (Camera::render_pixel + resample, src/camera.rs:84-114): the lane keeps the four fixed
    // samples apart (local memory: touched once per sample), compares them with their mean and, where they disagree,
    // goes on with `samples` more at counter-based offsets (include/rtc.h); rays then carry weight 1 and the sum is
    // divided at the end.  samples 0 / 1 keep the weighted accumulation (x 0.25 is exact, so 0 gives the same bits either way).
    const bool aa = cam.samples >= 2;
    const int n_extra = aa ? (int)cam.samples : 0;
    const uint64_t aa_seed_mixed = splitmix64((uint64_t)cam.aa_seed);
    const float w0 = (cam.samples == 1 || aa) ? 1.0f : 0.25f;
    const F3 cam_origin = mk((float)cam.vinv[3], (float)cam.vinv[7], (float)cam.vinv[11]);  // Vinv * (0,0,0)
    // Camera::ray_for_pixel_offset (src/camera.rs:64-76): pixel - origin = Vinv3 * (wx, wy, -1) with
    // wx = half_width - (x + xo) ps, wy = half_height - (y + yo) ps  ==  cam_dx (x+xo) + cam_dy (y+yo) + cam_d0
    const F3 cam_dx = mk((float)(-cam.pixel_size * cam.vinv[0]), (float)(-cam.pixel_size * cam.vinv[4]), (float)(-cam.pixel_size * cam.vinv[8]));
    const F3 cam_dy = mk((float)(-cam.pixel_size * cam.vinv[1]), (float)(-cam.pixel_size * cam.vinv[5]), (float)(-cam.pixel_size * cam.vinv[9]));
    const F3 cam_d0 = mk((float)(cam.half_width * cam.vinv[0] + cam.half_height * cam.vinv[1] - cam.vinv[2]),
                         (float)(cam.half_width * cam.vinv[4] + cam.half_height * cam.vinv[5] - cam.vinv[6]),
                         (float)(cam.half_width * cam.vinv[8] + cam.half_height * cam.vinv[9] - cam.vinv[10]));

    unsigned wtile = 0;       // warp-uniform: the tile being handed out, next pixel in it
    int wnext = kBatch;
    bool exhausted = false;
    bool has_ray = false, dead = false;
    F3 ro = mk(0, 0, 0), rd = mk(0, 0, 1);
    float rw = 0.f;
    int remaining = 0;
    bool is_primary = false;
    bool have = false;
    int px = 0, pr = 0, py = 0, sub = 0;
    F3 acc = mk(0, 0, 0);
    float aa_s[12];                 // the four fixed samples of the pixel (aa only)
    F3 aa_tot = mk(0, 0, 0);        // Color::average_over's running sums
    int aa_ntot = n_sub;            // samples this pixel takes: 4, or 4 + n_extra once the fixed four disagree
    unsigned hitbits = 0;
    float4 stk_a[kRefine ? 1 : kMaxStack], stk_b[kRefine ? 1 : kMaxStack];  // pending refraction rays: (o, w), (d, remaining)
    // kRefine: the ray in f64 (ro / rd are its f32 roundings, refreshed every iteration) and the pending refraction rays in f64
    p64::V3 rod = p64::mk(0., 0., 0.), rdd = p64::mk(0., 0., 1.);
    double2 stk_d0[kRefine ? kMaxStack : 1], stk_d1[kRefine ? kMaxStack : 1], stk_d2[kRefine ? kMaxStack : 1];   // (ox oy) (oz dx) (dy dz)
    float2 stk_w[kRefine ? kMaxStack : 1];                                                                       // (weight, remaining)
    int sp = 0;
    unsigned n_rays = 0, n_shaded = 0, n_phits = 0;
    float2 neg[(kUnique || kRefine) ? 1 : kNegCap];   // negative-t entries of the current ray (aliased world_ids only)
    int n_neg = 0;

    for (;;) {
        // ---------------------------------------------------------------- refill lanes without a ray
        if (!has_ray && sp > 0) {
            --sp;
            RTC_CHECK(sp >= 0 && sp < kMaxStack);
            if (kRefine) {
                const double2 a = stk_d0[sp], b = stk_d1[sp], c = stk_d2[sp];
                const float2 w = stk_w[sp];
                rod = p64::mk(a.x, a.y, b.x); rdd = p64::mk(b.y, c.x, c.y); rw = w.x; remaining = __float_as_int(w.y);
            } else {
                float4 a = stk_a[sp], b = stk_b[sp];
                ro = xyz(a); rw = a.w; rd = xyz(b); remaining = __float_as_int(b.w);
            }
            is_primary = false; has_ray = true;
        }
        if (!has_ray && have) {
            if (aa && sub < aa_ntot) {          // a sample of the pixel is complete
                aa_tot = aa_tot + acc;
                if (sub < 4) { aa_s[3 * sub] = acc.x; aa_s[3 * sub + 1] = acc.y; aa_s[3 * sub + 2] = acc.z; }
                if (sub == 3) {                 // any(|sample - mean| > 0.01) -> resample (src/camera.rs:107-110)
                    const F3 mean = aa_tot * 0.25f;
                    bool any = false;
#pragma unroll
                    for (int k = 0; k < 4; ++k) {
                        const F3 dl = mk(aa_s[3 * k], aa_s[3 * k + 1], aa_s[3 * k + 2]) - mean;
                        any = any || dot(dl, dl) > 1.0e-4f;
                    }
                    if (any) aa_ntot = 4 + n_extra;
                }
                acc = mk(0, 0, 0);
            }
            if (sub + 1 < aa_ntot) ++sub;   // next sub-sample of the same pixel
            else {
                if (aa) acc = aa_tot * (1.0f / (float)aa_ntot);
                // pixel finished: Canvas::write_pixel (src/canvas.rs:43-46)
                size_t p = (size_t)(cam.dst_full ? py : (cam.shard_count > 1 ? cam.row0 + pr : pr)) * cam.hsize + px;
                RTC_CHECK(px >= 0 && px < (int)cam.hsize && (cam.dst_full ? (py >= 0 && py < (int)cam.vsize) : (pr >= 0 && pr < cam.shard_rows)));
                if (cam.out_format == RTC_OUT_F64_RGB) {
                    double* o = (double*)out_rgb + 3 * p;
                    o[0] = (double)acc.x; o[1] = (double)acc.y; o[2] = (double)acc.z;
                } else if (cam.out_format == RTC_OUT_F32_RGB) {
                    float* o = (float*)out_rgb + 3 * p;
                    o[0] = acc.x; o[1] = acc.y; o[2] = acc.z;
                } else {
                    uint8_t* o = (uint8_t*)out_rgb + 3 * p;  // Color::scale, src/color.rs:100-114
                    float c3[3] = {acc.x, acc.y, acc.z};
#pragma unroll
                    for (int k = 0; k < 3; ++k) {
                        float s = c3[k] * 255.0f;
                        int q = s != s ? 0 : (s >= 2147483520.0f ? 2147483647 : (s <= -2147483648.0f ? (-2147483647 - 1) : (int)s));
                        o[k] = (uint8_t)(q < 0 ? 0 : (q > 255 ? 255 : q));
                    }
                }
                if (out_hit) out_hit[p] = (uint8_t)hitbits;
                have = false;
            }
        }
        // lanes without a pixel claim the next ones of the warp's tile (ballot + popc rank)
        for (;;) {
            bool need = !has_ray && !have && !dead;
            unsigned m = __ballot_sync(full, need);
            if (m == 0u || exhausted) break;
            if (wnext >= kBatch) {
                unsigned t = 0;
                if (lane == 0) t = atomicAdd(work, 1u);
                wtile = __shfl_sync(full, t, 0);
                wnext = 0;
                if (wtile >= total_tiles) { exhausted = true; break; }
            }
            int rank = __popc(m & lt_mask);
            int take = min(__popc(m), kBatch - wnext);
            if (need && rank < take) {
                int k = wnext + rank;
                int tx = (int)(wtile % (unsigned)tiles_x), ty = (int)(wtile / (unsigned)tiles_x);
                px = tx * kTileX + (k & (kTileX - 1));
                pr = ty * kTileY + k / kTileX;
                if (px < (int)cam.hsize && pr < cam.shard_rows) {
                    py = cam.shard_count > 1 ? shard_row_to_frame_row(cam.row0 + pr, cam.tile_rows, cam.shard_index, cam.shard_count) : cam.row0 + pr;
                    have = true; sub = 0; acc = mk(0, 0, 0); hitbits = 0;
                    aa_tot = mk(0, 0, 0); aa_ntot = n_sub;
                    if (!(px < (int)cam.xlim && py < (int)cam.ylim)) sub = kSubOutside;  // outside the rect: stays black
                }
            }
            wnext += take;
        }
        if (!has_ray && have && sub < aa_ntot) {
            float xo = n_sub == 1 ? 0.5f : ((sub & 1) ? 0.75f : 0.25f);
            float yo = n_sub == 1 ? 0.5f : ((sub & 2) ? 0.75f : 0.25f);
            if (aa && sub >= 4) {               // Camera::resample's offsets (src/camera.rs:84-92; include/rtc.h)
                const uint64_t pixel_index = (uint64_t)py * cam.hsize + (uint64_t)px;
                xo = (float)(unsigned)(resample_bits(aa_seed_mixed, pixel_index, (unsigned)(sub - 4), 0u) >> 40) * 5.9604644775390625e-8f;
                yo = (float)(unsigned)(resample_bits(aa_seed_mixed, pixel_index, (unsigned)(sub - 4), 1u) >> 40) * 5.9604644775390625e-8f;
            }
            if (kRefine) {                      // Camera::ray_for_pixel_offset in f64, the reference's statements (src/camera.rs:64-76)
                double xod = n_sub == 1 ? 0.5 : ((sub & 1) ? 0.75 : 0.25);
                double yod = n_sub == 1 ? 0.5 : ((sub & 2) ? 0.75 : 0.25);
                if (aa && sub >= 4) {
                    const uint64_t pixel_index = (uint64_t)py * cam.hsize + (uint64_t)px;
                    xod = (double)(resample_bits(aa_seed_mixed, pixel_index, (unsigned)(sub - 4), 0u) >> 11) * (1.0 / 9007199254740992.0);
                    yod = (double)(resample_bits(aa_seed_mixed, pixel_index, (unsigned)(sub - 4), 1u) >> 11) * (1.0 / 9007199254740992.0);
                }
                const double xoffset = ((double)px + xod) * cam.pixel_size;
                const double yoffset = ((double)py + yod) * cam.pixel_size;
                const double world_x = cam.half_width - xoffset;
                const double world_y = cam.half_height - yoffset;
                const p64::V3 pixel = p64::xpoint(cam.vinv, p64::mk(world_x, world_y, -1.));
                const p64::V3 origin = p64::xpoint(cam.vinv, p64::mk(0., 0., 0.));
                rod = origin; rdd = p64::normalize(p64::sub(pixel, origin));
            } else {
                float fx = (float)px + xo;
                float fy = (float)py + yo;
                F3 dir = fma3(cam_dx, fx, fma3(cam_dy, fy, cam_d0));
                rd = dir * rsqrtf(dot(dir, dir));
                ro = cam_origin;
            }
            rw = w0; remaining = cam.max_depth; is_primary = true; has_ray = true;
        }
        if (!has_ray && !have && exhausted) dead = true;
        if (__all_sync(full, dead)) break;

        // ---------------------------------------------------------------- phase A: World::intersect, radiance rays
        float t = __int_as_float(0x7f800000);
        int hit = -1;
        double t64 = 0.0;
        if (kRefine) {
            ro = mk((float)rod.x, (float)rod.y, (float)rod.z);   // f32 roundings of the ray: the filter, the shading
            rd = mk((float)rdd.x, (float)rdd.y, (float)rdd.z);
            if (kFilter) {
                // planes with the ray's own lane; the solids of every ray by the whole warp, one ray at a time (warp_query)
                double pt = __longlong_as_double(0x7ff0000000000000LL);
                int pidx = 0x7fffffff;
                __syncwarp();
                if (has_ray) {
                    ++n_rays;
                    planes_mixed(S, R64, rod, rdd, pt, pidx);
                    publish_rayq(W, lane, make_rayq(S, ro, rd));
                    publish_ray64(W, lane, rod, rdd);
                }
                __syncwarp();
                unsigned todo = __ballot_sync(full, has_ray);
                while (todo) {
                    const int w = __ffs(todo) - 1;
                    todo &= todo - 1u;
                    const RayQ Q = load_rayq(W, w);
                    p64::V3 o64, d64;
                    load_ray64(W, w, o64, d64);
                    double bt = __shfl_sync(full, pt, w);
                    int bi = __shfl_sync(full, pidx, w);
                    coop_hit<kCubes>(S, R64, Q, o64, d64, W.base + kOffQueue, bt, bi);
                    if (lane == w) { pt = bt; pidx = bi; }
                }
                t64 = pt;
                hit = pidx == 0x7fffffff ? -1 : pidx;                                                          // hit = shape index
            } else if (has_ray) { intersect_mixed<kCubes>(S, R64, rod, rdd, t64, hit); ++n_rays; }
        } else {
            if (has_ray) { intersect32<!kUnique, kCubes>(S, ro, rd, t, hit, neg, n_neg); ++n_rays; }            // hit = stream position
        }
        if (has_ray && is_primary && hit >= 0) { if (sub < 4) hitbits |= 1u << sub; ++n_phits; }
        if (hit < 0) has_ray = false;   // World::color_at -> background_color (black), src/shape.rs:708
        double n1c = 1.0, n2c = 1.0;    // kFilter: compute_refractive of the transparent hits, by the whole warp (the rays' slots are still published)
        if (kRefine && kFilter) {
            const bool need = hit >= 0 && R[hit].tr_ri.x != 0.f;
            unsigned todo = __ballot_sync(full, need);
            while (todo) {
                const int w = __ffs(todo) - 1;
                todo &= todo - 1u;
                const RayQ Q = load_rayq(W, w);
                p64::V3 o64, d64;
                load_ray64(W, w, o64, d64);
                double n1, n2;
                coop_containers<kCubes>(S, R64, Q, o64, d64, __shfl_sync(full, hit, w), W.parity, W.base + kOffQueue, W.base + kOffKept, n1, n2);
                if (lane == w) { n1c = n1; n2c = n2; }
            }
        }

        // ---------------------------------------------------------------- phase B: prepare_computations
        F3 hp = mk(0, 0, 0), hn = mk(0, 1, 0);
        float hn1 = 1.f, hn2 = 1.f;
        int hidx = 0;
        float rw_hit = rw;          // weight of the ray that produced this hit (kRefine sets up the children before shading)
        if (kRefine && hit >= 0) {
            // prepare_computations in f64, the parity kernel's statements (src/shape.rs:75-96, 144-152, 34-40), then
            // shade_hit's children (:689-699, 729-781): the next ray and the pending refraction ray leave this block in f64;
            // the shadow test and the shading below get f32 roundings (hp with the f32 epsilon, hn, rd).
            hidx = hit;
            RTC_CHECK(hidx >= 0 && hidx < sc.n_shapes);
            const ShadeRec64& r64 = R64[hidx];
            const ShadeRec32& rec = R[hidx];
            const float4 ds = rec.dif_spec, tr = rec.tr_ri;
            const p64::V3 point = p64::add(rod, p64::mul(rdd, t64));
            const p64::V3 eyev = p64::neg(rdd);
            const p64::V3 lp = p64::xpoint(r64.inv, point);
            p64::V3 ln = r64.kind == RTC_SHAPE_PLANE ? p64::mk(0., 1., 0.) : lp;
            if (kCubes && r64.kind == RTC_SHAPE_CUBE) {            // Cube::normal_at_local :601-610; r64.invt = transpose(inverse), :623-629
                const double ax = fabs(lp.x), ay = fabs(lp.y), az = fabs(lp.z);
                const double maxc = fmax(ax, fmax(ay, az));
                if (maxc == ax) ln = p64::mk(lp.x, 0., 0.);
                else if (maxc == ay) ln = p64::mk(0., lp.y, 0.);
                else ln = p64::mk(0., 0., lp.z);
            }
            p64::V3 normal = p64::normalize(p64::xvec3(r64.invt, ln));
            double n1 = 1.0, n2 = 1.0;
            if (tr.x != 0.f) {
                if (kFilter) { n1 = n1c; n2 = n2c; }
                else containers_mixed<kCubes>(S, R64, rod, rdd, hidx, n1, n2);
            }
            if (p64::dot(normal, eyev) < 0.0) normal = p64::neg(normal);
            const p64::V3 over_point = p64::add(point, p64::mul(normal, p64::kEps));
            const double cos_i = p64::dot(eyev, normal);
            hn = mk((float)normal.x, (float)normal.y, (float)normal.z);
            hp = fma3(hn, kEps32, mk((float)point.x, (float)point.y, (float)point.z));
            hn1 = (float)n1; hn2 = (float)n2;
            const float k_refl = ds.w, k_tr = tr.x;
            float refl_w = rw * k_refl, refr_w = rw * k_tr;
            if (k_refl > 0.f && k_tr > 0.f) {                          // World::reflectance (Schlick): a weight, f32
                float cs = (float)cos_i, refl = 0.f;
                bool tir = false;
                if (hn1 > hn2) {
                    float nn = hn1 / hn2;
                    float sin2_t = nn * nn * (1.f - cs * cs);
                    if (sin2_t > 1.f) tir = true; else cs = sqrtf(1.f - sin2_t);
                }
                if (tir) refl = 1.f;
                else {
                    float q = (hn1 - hn2) / (hn1 + hn2), r0 = q * q, m1 = 1.f - cs, m2 = m1 * m1;
                    refl = fmaf(1.f - r0, m2 * m2 * m1, r0);
                }
                refl_w *= refl; refr_w *= 1.f - refl;
            }
            const bool want_refl = remaining != 0 && k_refl > 0.f;
            bool want_refr = false;
            p64::V3 fo = p64::mk(0., 0., 0.), fd = p64::mk(0., 0., 1.);
            if (remaining != 0 && k_tr != 0.f) {                       // refracted_color, src/shape.rs:751-766
                const double n_ratio = p64::ddiv(n1, n2);
                const double sin2_t = (n_ratio * n_ratio) * (1.0 - (cos_i * cos_i));
                if (!(sin2_t > 1.0)) {
                    const double cos_t = p64::dsqrt(1.0 - sin2_t);
                    fd = p64::sub(p64::mul(normal, n_ratio * cos_i - cos_t), p64::mul(eyev, n_ratio));
                    fo = p64::sub(point, p64::mul(normal, p64::kEps));   // under_point
                    want_refr = true;
                }
            }
            if (want_refl) {
                if (want_refr && sp < kMaxStack) {
                    stk_d0[sp] = make_double2(fo.x, fo.y); stk_d1[sp] = make_double2(fo.z, fd.x); stk_d2[sp] = make_double2(fd.y, fd.z);
                    stk_w[sp] = make_float2(refr_w, __int_as_float(remaining - 1));
                    ++sp;
                }
                rdd = p64::reflect(rdd, normal);                       // reflectv, src/shape.rs:86
                rod = over_point; rw = refl_w; --remaining;
            } else if (want_refr) {
                rod = fo; rdd = fd; rw = refr_w; --remaining;
            } else {
                has_ray = false;
            }
            is_primary = false;
        }
        if (!kRefine && hit >= 0) {
            RTC_CHECK(hit < S.np + S.nu + S.ng + S.nc);
            hidx = stream_to_shape(S, hit);
            RTC_CHECK(hidx >= 0 && hidx < sc.n_shapes);
            const ShadeRec32& rec = R[hidx];
            const float4 tr = rec.tr_ri;
            const int geom = __float_as_int(tr.w);
            F3 p = fma3(rd, t, ro);                              // Ray::position
            if (geom == GEOM_PLANE) hn = xyz(rec.g0);
            else if (geom == GEOM_USPHERE) { float4 g = rec.g0; hn = (p - xyz(g)) * g.w; }
            else {                                                // Shape::normal_at, src/shape.rs:34-40
                F3 lp = mk(row4(rec.g0, p), row4(rec.g1, p), row4(rec.g2, p));
                if (kCubes && geom == GEOM_CUBE) {                          // Cube::normal_at_local, src/shape.rs:601-610 (t rows = transpose(inverse), :623-629)
                    const float ax = fabsf(lp.x), ay = fabsf(lp.y), az = fabsf(lp.z);
                    const float maxc = fmaxf(ax, fmaxf(ay, az));
                    if (maxc == ax) lp = mk(lp.x, 0.f, 0.f);
                    else if (maxc == ay) lp = mk(0.f, lp.y, 0.f);
                    else lp = mk(0.f, 0.f, lp.z);
                }
                F3 wn = mk(row3(rec.t0, lp), row3(rec.t1, lp), row3(rec.t2, lp));
                hn = wn * rsqrtf(dot(wn, wn));
            }
            if (tr.x != 0.f) {                                    // Intersection::compute_refractive
                if (kUnique && n_gset <= 16 && !origin_in_gset(S, gset, n_gset, ro, rd)) { hn1 = 1.f; hn2 = tr.y; }
                else if (kUnique) compute_refractive32_unique(S, R, ro, rd, hidx, hn1, hn2);
                else if (n_neg <= kNegCap) compute_refractive32_kept(S, R, ro, rd, hidx, neg, n_neg, hn1, hn2);
                else compute_refractive32(S, R, ro, rd, hidx, hn1, hn2);   // more entries than were kept: the two passes
            }
            if (dot(hn, rd) > 0.f) hn = mk(-hn.x, -hn.y, -hn.z);  // inside: normal . eyev < 0 (src/shape.rs:77-78)
            hp = fma3(hn, kEps32, p);                             // over_point
        }

        // ---------------------------------------------------------------- phases C + D, part 1: is_shadowed (any-hit) and
        // Material::lighting (src/material.rs:319-361) at over_point — once per light (rtc_scene_upload_lights: the terms
        // add up in list order, include/rtc.h; the reference's World has one light, src/shape.rs:686).  The loop bound is
        // warp-uniform, so all lanes still run each shadow test in lock step.
        F3 surface = mk(0, 0, 0);
        F3 base = mk(0, 0, 0);
        if (hit >= 0) {
            ++n_shaded;
            const ShadeRec32& rec = R[hidx];
            const int pkind = __float_as_int(rec.tr_ri.z);
            base = xyz(rec.col_amb);
            if (pkind != RTC_PATTERN_NONE) {
                F3 pp = mk(row4(rec.p0, hp), row4(rec.p1, hp), row4(rec.p2, hp));
                base = pattern_at32(pkind, xyz(rec.pa), xyz(rec.pb), pp);
            }
        }
        for (int li = 0; li < sc.n_lights; ++li) {
            F3 lpos = light_pos, lint = light_int;
            const float* pL = plL; const float* uL = usL;
            if (li > 0) {
                lpos = xyz(sc.lights[2 * li]); lint = xyz(sc.lights[2 * li + 1]);
                pL = sc.plL + li * S.np; uL = sc.usL + li * S.nu;
            }
            const bool inside = ((sc.light_inside >> li) & 1) != 0;
            bool shadowed = false;
            if (kFilter) {
                // is_shadowed: plane terms with the hit's lane, the solids as an any-hit segment query by the whole warp
                const F3 sd = lpos - hp;
                bool todo_me = false;
                __syncwarp();
                if (hit >= 0) {
                    ++n_rays;
                    shadowed = shadow_planes32(S, pL, hp);
                    if (!shadowed) {
                        todo_me = true;
                        publish_rayq(W, lane, make_rayq(S, hp, sd));
                        ssts4(W.base + kOffSo + 16u * (unsigned)lane, make_float4(hp.x, hp.y, hp.z, 0.f));
                        ssts4(W.base + kOffSd + 16u * (unsigned)lane, make_float4(sd.x, sd.y, sd.z, 0.f));
                    }
                }
                __syncwarp();
                unsigned todo = __ballot_sync(full, todo_me);
                while (todo) {
                    const int w = __ffs(todo) - 1;
                    todo &= todo - 1u;
                    const RayQ Q = load_rayq(W, w);
                    const bool sh = coop_shadow<kCubes>(S, uL, inside, Q, xyz(slds4(W.base + kOffSo + 16u * (unsigned)w)),
                                                        xyz(slds4(W.base + kOffSd + 16u * (unsigned)w)), W.base + kOffQueue);
                    if (lane == w) shadowed = sh;
                }
            } else if (hit >= 0) { shadowed = shadow32<kCubes>(S, pL, uL, inside, hp, lpos - hp); ++n_rays; }
            if (hit >= 0) {
                const ShadeRec32& rec = R[hidx];
                const float4 ca = rec.col_amb, ds = rec.dif_spec;
                const F3 eff = base * lint;
                surface = fma3(eff, ca.w, surface);                        // ambient
                if (!shadowed) {
                    F3 lv = lpos - hp;
                    lv = lv * rsqrtf(dot(lv, lv));
                    float ldn = dot(lv, hn);
                    if (ldn >= 0.f) {
                        surface = fma3(eff, ds.x * ldn, surface);          // diffuse
                        F3 rv = fma3(hn, 2.f * ldn, mk(-lv.x, -lv.y, -lv.z));  // (-lightv).reflect(normal)
                        float rde = -dot(rv, rd);                          // rv . eye
                        if (rde > 0.f) surface = fma3(lint, ds.y * __powf(rde, ds.z), surface);
                    }
                }
            }
        }

        // ---------------------------------------------------------------- phase D, part 2: shade_hit's children
        if (hit >= 0) {
            const ShadeRec32& rec = R[hidx];
            const float4 ds = rec.dif_spec, tr = rec.tr_ri;
            const F3 eye = mk(-rd.x, -rd.y, -rd.z);
            acc = fma3(surface, rw_hit, acc);
            if (!kRefine) {   // (kRefine: the children were set up in phase B)
            // children (src/shape.rs:689-699, 729-781) as weighted rays
            const float k_refl = ds.w, k_tr = tr.x;
            const bool both = k_refl > 0.f && k_tr > 0.f;
            const float cos_i = dot(eye, hn);
            float refl_w = rw * k_refl, refr_w = rw * k_tr;
            if (both) {                                                // World::reflectance (Schlick)
                float cs = cos_i, refl = 0.f;
                bool tir = false;
                if (hn1 > hn2) {
                    float nn = hn1 / hn2;
                    float sin2_t = nn * nn * (1.f - cs * cs);
                    if (sin2_t > 1.f) tir = true; else cs = sqrtf(1.f - sin2_t);
                }
                if (tir) refl = 1.f;
                else {
                    float q = (hn1 - hn2) / (hn1 + hn2), r0 = q * q, m = 1.f - cs, m2 = m * m;
                    refl = fmaf(1.f - r0, m2 * m2 * m, r0);
                }
                refl_w *= refl; refr_w *= 1.f - refl;
            }
            const bool want_refl = remaining != 0 && k_refl > 0.f;
            bool want_refr = false;
            F3 fo = mk(0, 0, 0), fd = mk(0, 0, 1);
            if (remaining != 0 && k_tr != 0.f) {                       // refracted_color
                float ratio = hn1 / hn2;
                float sin2_t = ratio * ratio * (1.f - cos_i * cos_i);
                if (!(sin2_t > 1.f)) {
                    float cos_t = sqrtf(1.f - sin2_t);
                    fd = fma3(hn, ratio * cos_i - cos_t, eye * (-ratio));
                    fo = fma3(hn, -2.f * kEps32, hp);                  // under_point
                    want_refr = true;
                }
            }
            is_primary = false;
            if (want_refl) {
                if (want_refr && sp < kMaxStack) {
                    stk_a[sp] = make_float4(fo.x, fo.y, fo.z, refr_w);
                    stk_b[sp] = make_float4(fd.x, fd.y, fd.z, __int_as_float(remaining - 1));
                    ++sp;
                }
                float dn = dot(rd, hn);
                rd = fma3(hn, -2.f * dn, rd);                          // reflectv = d.reflect(n)
                ro = hp; rw = refl_w; --remaining;
            } else if (want_refr) {
                ro = fo; rd = fd; rw = refr_w; --remaining;
            } else {
                has_ray = false;
            }
            }
        }
    }

    // The tile counter resets itself: the last block to finish zeroes work[0] (tiles handed out) and
    // work[1] (blocks done), so a launch needs no memset in front of it.
    __syncthreads();
    if (threadIdx.x == 0) {
        __threadfence();
        if (atomicAdd(work + 1, 1u) == gridDim.x - 1) { work[0] = 0u; work[1] = 0u; __threadfence(); }
    }
    if (counters) {
        unsigned r = __reduce_add_sync(full, n_rays), h = __reduce_add_sync(full, n_shaded), ph = __reduce_add_sync(full, n_phits);
        if (lane == 0) {
            atomicAdd(&counters->rays, (unsigned long long)r);
            atomicAdd(&counters->shadow_rays, (unsigned long long)h);
            atomicAdd(&counters->shaded_hits, (unsigned long long)h);
            atomicAdd(&counters->primary_hits, (unsigned long long)ph);
        }
    }
}

}  // namespace

cudaError_t preload_render_f32() {
    cudaFuncAttributes at;
    cudaError_t e = cudaSuccess;
    auto touch = [&](auto kernel) { if (e == cudaSuccess) e = cudaFuncGetAttributes(&at, kernel); };
#define RTC_TOUCH32(SM, GL) touch(render_f32_kernel<SM, true, GL, true>); touch(render_f32_kernel<SM, true, GL, false>); \
                            touch(render_f32_kernel<SM, false, GL, true>); touch(render_f32_kernel<SM, false, GL, false>);
    RTC_TOUCH32(false, 2) RTC_TOUCH32(false, 1) RTC_TOUCH32(true, 0) RTC_TOUCH32(false, 0)
#undef RTC_TOUCH32
    return e;
}

// Beyond this many shapes the general kernel runs in its kRefine form (hierarchy queries in f32, the radiance rays' geometry
// in f64) even when its streams would fit shared memory: an all-f32 path drifts on dense sphere fields — every bounce
// amplifies the hit point's rounding — and falls under the 99.9 % bar (1500 random spheres: 99.79 % of the pixels within
// 1/255; tests/test_gpu_parity.py::test_fast_mode_dispatch_by_scene_size).  Fast mode's default for such scenes is the
// parity kernel (rtc_api.cu, launch_on); this is what RTC_OPT_F32_GENERAL_KERNEL gets.
constexpr int kRefineFromShapes = 128;
static size_t f32_stream_bytes(const DevScene32& scene) {
    const int n_obj = scene.n_gsph + scene.n_cube;
    return (size_t)(scene.n_planes + scene.n_usph + 3 * n_obj) * 16 + (size_t)(scene.n_planes + scene.n_usph + n_obj) * 4 +
           (size_t)(scene.n_planes + scene.n_usph) * 4 + 16 * 4 + 16;
}
bool f32_streams_are_global(const DevScene32& scene) {
    return f32_stream_bytes(scene) + (size_t)scene.n_shapes * sizeof(ShadeRec32) > 40 * 1024 && (f32_stream_bytes(scene) > 96 * 1024 || scene.n_shapes > kRefineFromShapes);
}

cudaError_t launch_render_f32(const DevScene32& scene, const CamParams& cam, void* d_rgb, uint8_t* d_hit,
                              Counters* d_counters, unsigned* d_work, cudaStream_t stream) {
    if (cam.shard_rows <= 0) return cudaSuccess;
    cudaError_t e = cudaSuccess;   // d_work[0..1] are zero between launches (the kernel resets them)
    int dev = 0, sms = 0;
    cudaGetDevice(&dev);
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    const int n_obj = scene.n_gsph + scene.n_cube;
    const size_t streams = (size_t)(scene.n_planes + scene.n_usph + 3 * n_obj) * 16 +
                           (size_t)(scene.n_planes + scene.n_usph + n_obj) * 4 + (size_t)(scene.n_planes + scene.n_usph) * 4 + 16 * 4 + 16;
    const size_t recs = (size_t)scene.n_shapes * sizeof(ShadeRec32);
    const int tiles = ((cam.hsize + kTileX - 1) / kTileX) * ((cam.shard_rows + kTileY - 1) / kTileY);
    const int warps = kThreads / 32;
    const bool small = streams + recs <= 40 * 1024;
    const bool global_streams = !small && (streams > 96 * 1024 || scene.n_shapes > kRefineFromShapes);   // big scenes: streams stay in global memory behind L1/L2
    const size_t smem = global_streams ? (size_t)(kThreads / 32) * kWarpSlotBytes : (small ? streams + recs : streams);
    auto launch = [&](auto kernel) -> cudaError_t {
        cudaError_t er = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (er != cudaSuccess) return er;
        int blocks_per_sm = 0;
        cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kernel, kThreads, smem);
        int grid = max(1, min(sms * max(blocks_per_sm, 1), (tiles + warps - 1) / warps));
        kernel<<<grid, kThreads, smem, stream>>>(scene, cam, d_rgb, d_hit, d_counters, d_work);
        return cudaSuccess;
    };
    const bool uniq = scene.ids_unique != 0, cubes = scene.n_cube > 0;
#define RTC_LAUNCH32(SM, GL) (uniq ? (cubes ? launch(render_f32_kernel<SM, true, GL, true>) : launch(render_f32_kernel<SM, true, GL, false>)) \
                                   : (cubes ? launch(render_f32_kernel<SM, false, GL, true>) : launch(render_f32_kernel<SM, false, GL, false>)))
    if (global_streams && scene.n_bound4 > 0) e = RTC_LAUNCH32(false, 2);
    else if (global_streams) e = RTC_LAUNCH32(false, 1);
    else if (small) e = RTC_LAUNCH32(true, 0);
    else e = RTC_LAUNCH32(false, 0);
#undef RTC_LAUNCH32
    if (e != cudaSuccess) return e;
    return cudaGetLastError();
}

}  // namespace rtc
