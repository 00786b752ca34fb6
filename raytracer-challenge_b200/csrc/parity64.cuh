// f64 statements of the reference's geometry in the reference's operation order — shared by the parity kernel
// (render_f64.cu) and by the geometry half of the mixed-precision big-scene kernel (render_f32.cu, kRefine).
// Every translation unit that includes this is compiled with -fmad=false: Rust never contracts a*b+c, and a value
// computed here can decide a branch (hit / miss, t >= 0, tie rules), so the roundings have to be the reference's.
// f64 '/' and sqrt() are IEEE-rounded in CUDA by construction (--use_fast_math only touches f32).
#pragma once
#include "rtc_common.h"

namespace rtc {
namespace p64 {

constexpr double kEps = 0.00000001;  // Vector::EPSILON, src/vec.rs:16

struct V3 { double x, y, z; };
__device__ __forceinline__ V3 mk(double x, double y, double z) { V3 v; v.x = x; v.y = y; v.z = z; return v; }
__device__ __forceinline__ V3 add(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 sub(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 mul(V3 a, double m) { return mk(a.x * m, a.y * m, a.z * m); }
__device__ __forceinline__ V3 mulc(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ V3 neg(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ double dot(V3 a, V3 b) { return a.x * b.x + a.y * b.y + a.z * b.z; }   // src/vec.rs:78
// IEEE division and square root of the path.  By default they are inlined where they stand (30 + 11 copies of the
// ~15-instruction sequences in the C3 variant of the parity kernel); -DRTC_F64_SHARED_DIV makes each a single
// non-inlined routine (same operations, same bits) to shrink the instruction footprint — an experiment knob,
// measured slower.
#ifdef RTC_F64_SHARED_DIV
__device__ __noinline__ double ddiv(double a, double b) { return a / b; }
__device__ __noinline__ double dsqrt(double a) { return sqrt(a); }
#else
__device__ __forceinline__ double ddiv(double a, double b) { return a / b; }
__device__ __forceinline__ double dsqrt(double a) { return sqrt(a); }
#endif
__device__ __forceinline__ V3 normalize(V3 a) {                                                    // src/vec.rs:65-76
    double m = dsqrt(a.x * a.x + a.y * a.y + a.z * a.z);
    return mk(ddiv(a.x, m), ddiv(a.y, m), ddiv(a.z, m));
}
__device__ __forceinline__ V3 reflect(V3 v, V3 n) { return sub(v, mul(n, 2. * dot(v, n))); }       // src/vec.rs:106
// Matrix::transform_point / transform_vector, src/transform.rs:107-128 (sums left to right)
__device__ __forceinline__ V3 xpoint(const double* m, V3 p) {
    return mk(m[0] * p.x + m[1] * p.y + m[2] * p.z + m[3], m[4] * p.x + m[5] * p.y + m[6] * p.z + m[7],
              m[8] * p.x + m[9] * p.y + m[10] * p.z + m[11]);
}
__device__ __forceinline__ V3 xvec3(const double* m9, V3 p) {
    return mk(m9[0] * p.x + m9[1] * p.y + m9[2] * p.z, m9[3] * p.x + m9[4] * p.y + m9[5] * p.z,
              m9[6] * p.x + m9[7] * p.y + m9[8] * p.z);
}

// Shape::intersect + Sphere::intersect_local (src/shape.rs:23-26, 361-375); r = rows 0..2 of the stored inverse as
// six double2 (r0.xy r0.zw r1.xy r1.zw r2.xy r2.zw)
__device__ __forceinline__ bool sphere_roots(const double2* r, V3 o, V3 d, double& t1, double& t2) {
    double2 a0 = r[0], a1 = r[1], b0 = r[2], b1 = r[3], c0 = r[4], c1 = r[5];
    double ox = a0.x * o.x + a0.y * o.y + a1.x * o.z + a1.y;
    double oy = b0.x * o.x + b0.y * o.y + b1.x * o.z + b1.y;
    double oz = c0.x * o.x + c0.y * o.y + c1.x * o.z + c1.y;
    double dx = a0.x * d.x + a0.y * d.y + a1.x * d.z;
    double dy = b0.x * d.x + b0.y * d.y + b1.x * d.z;
    double dz = c0.x * d.x + c0.y * d.y + c1.x * d.z;
    double a = dx * dx + dy * dy + dz * dz;
    double b = 2. * (dx * ox + dy * oy + dz * oz);
    double c = (ox * ox + oy * oy + oz * oz) - 1.;
    double disc = (b * b) - 4. * a * c;
    if (disc < 0.) return false;
    double s = dsqrt(disc);
    t1 = ddiv(-b - s, 2. * a);
    t2 = ddiv(-b + s, 2. * a);
    return true;
}
// Cube::check_axis (src/shape.rs:540-564)
__device__ __forceinline__ void check_axis(double origin, double direction, double& tmin, double& tmax) {
    const double tmin_numerator = -1.0 - origin;
    const double tmax_numerator = 1. - origin;
    if (fabs(direction) >= kEps) {
        tmin = ddiv(tmin_numerator, direction);
        tmax = ddiv(tmax_numerator, direction);
    } else {
        const double inf = __longlong_as_double(0x7ff0000000000000LL);
        tmin = tmin_numerator >= 0.0 ? inf : -inf;
        tmax = tmax_numerator >= 0.0 ? inf : -inf;
    }
    if (tmin > tmax) { const double t = tmin; tmin = tmax; tmax = t; }
}
// Shape::intersect + Cube::intersect_local (src/shape.rs:618-621, 577-591): same ray transform as the
// sphere, slab test per axis; f64::max / f64::min return the other operand for a NaN, like fmax / fmin.
__device__ __forceinline__ bool cube_roots(const double2* r, V3 o, V3 d, double& t1, double& t2) {
    double2 a0 = r[0], a1 = r[1], b0 = r[2], b1 = r[3], c0 = r[4], c1 = r[5];
    double ox = a0.x * o.x + a0.y * o.y + a1.x * o.z + a1.y;
    double oy = b0.x * o.x + b0.y * o.y + b1.x * o.z + b1.y;
    double oz = c0.x * o.x + c0.y * o.y + c1.x * o.z + c1.y;
    double dx = a0.x * d.x + a0.y * d.y + a1.x * d.z;
    double dy = b0.x * d.x + b0.y * d.y + b1.x * d.z;
    double dz = c0.x * d.x + c0.y * d.y + c1.x * d.z;
    double xmin, xmax, ymin, ymax, zmin, zmax;
    check_axis(ox, dx, xmin, xmax);
    check_axis(oy, dy, ymin, ymax);
    check_axis(oz, dz, zmin, zmax);
    const double tmin = fmax(xmin, fmax(ymin, zmin));
    const double tmax = fmin(xmax, fmin(ymax, zmax));
    if (!(tmin < tmax)) return false;
    t1 = tmin; t2 = tmax;
    return true;
}
// Plane::intersect_local (src/shape.rs:462-471); r = row 1 of the stored inverse as two double2
__device__ __forceinline__ bool plane_root(const double2* r, V3 o, V3 d, double& t) {
    double2 b0 = r[0], b1 = r[1];
    double dy = b0.x * d.x + b0.y * d.y + b1.x * d.z;
    if (fabs(dy) < kEps) return false;
    double oy = b0.x * o.x + b0.y * o.y + b1.x * o.z + b1.y;
    t = ddiv(-oy, dy);
    return true;
}

// World::intersect + Intersections::get_hit (src/shape.rs:677-683, 195-232): the hit is the
// smallest t >= 0; among equal t the earlier shape (then t1 before t2) wins — the order the
// reference's stable insertion sort leaves them in.
__device__ __forceinline__ void take(double t, int idx, double& best_t, int& best_idx) {
    if (t >= 0.0 && (t < best_t || (t == best_t && idx < best_idx))) { best_t = t; best_idx = idx; }
}

// Intersection::compute_refractive (src/shape.rs:115-141) without materialising the list.
// Every list entry before the hit has t < 0.  Containers are toggled by world_id, so after the
// negative entries the container list holds, per id with an ODD number of negative entries, that
// id's LAST entry, ordered by list position (t, shape index, root).  n1 = refractive index of the
// last container; then the hit toggles its own id and n2 = the last container after that.
struct Cont { double t; int ord; int idx; int id; };  // ord = 2*shape_index + root
__device__ __forceinline__ bool later(const Cont& a, const Cont& b) { return a.t > b.t || (a.t == b.t && a.ord > b.ord); }
__device__ __forceinline__ void offer(Cont e, Cont& top1, Cont& top2) {
    if (top1.idx >= 0 && e.id == top1.id) { if (later(e, top1)) top1 = e; return; }
    if (top2.idx >= 0 && e.id == top2.id) {
        if (later(e, top2)) top2 = e;
        if (top1.idx < 0 || later(top2, top1)) { Cont s = top1; top1 = top2; top2 = s; }
        return;
    }
    if (top1.idx < 0 || later(e, top1)) { top2 = top1; top1 = e; }
    else if (top2.idx < 0 || later(e, top2)) top2 = e;
}

}  // namespace p64
}  // namespace rtc
