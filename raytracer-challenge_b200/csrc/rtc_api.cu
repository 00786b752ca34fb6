// C ABI of include/rtc.h: context management, the scene-flattening pass (AoS rtc_shape[] ->
// device SoA intersect streams + shading records), render dispatch over the context's GPUs,
// row-tile de-interleave, and the FMA peak micro-benchmark.
// No CPU path exists here: every compute entry point needs a CUDA device.
#include <algorithm>
#include <array>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <condition_variable>
#include <functional>
#include <limits>
#include <mutex>
#include <thread>
#include <vector>
#include "rtc_common.h"

using namespace rtc;

extern "C" void rtc_host_widen_f32_f64(double* dst, const float* src, size_t n);   // host_simd.cpp

namespace {

thread_local std::string g_err;

int fail(int code, const std::string& msg) {
    g_err = msg;
    return code;
}
#define CUDA_TRY(expr)                                                                              \
    do {                                                                                            \
        cudaError_t _e = (expr);                                                                    \
        if (_e != cudaSuccess)                                                                      \
            return fail(RTC_ERR_CUDA, std::string(#expr) + ": " + cudaGetErrorString(_e));          \
    } while (0)

template <typename T>
struct DevBuf {
    T* p = nullptr;
    size_t cap = 0;  // elements
    cudaError_t ensure(size_t n) {
        if (n <= cap) return cudaSuccess;
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
        cudaError_t e = cudaMalloc(&p, std::max<size_t>(n, 1) * sizeof(T));
        if (e == cudaSuccess) cap = n;
        return e;
    }
    void release() {
        if (p) cudaFree(p);
        p = nullptr;
        cap = 0;
    }
};

constexpr int kMaxBands = 16;   // row bands of one rtc_render frame (launches whose copies overlap the next launch)

struct DeviceState {
    int dev = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // D2H of finished bands while later bands render
    cudaEvent_t ev_k0 = nullptr, ev_k1 = nullptr, ev_c0 = nullptr, ev_c1 = nullptr;   // timing events: kernels, copies of a frame
    cudaEvent_t ev_band[kMaxBands] = {}, ev_copy[kMaxBands] = {};
    // scene
    DevBuf<double2> sph64, pl64;
    DevBuf<float4> pl32, us32, gs32, bound;
    DevBuf<int32_t> sph_index, pl_index, us_index, gs_index, gset;
    DevBuf<float> plL, usL;
    DevBuf<float4> bound64, gbound, gbound64, sbound, sbound64, gbox, sbox;
    DevBuf<double> lights64;   // [n_lights][6]
    DevBuf<float4> lights32;   // [n_lights][2]
    DevBuf<ShadeRec64> rec64;
    DevBuf<ShadeRec32> rec32;
    DevBuf<SmallRec32> srec;   // small-scene shading records, by stream position
    SmallParams small{};       // scene part of the small-scene kernel parameter (camera part filled per launch)
    bool small_ok = false;
    DevScene64 s64{};
    DevScene32 s32{};
    // frame
    DevBuf<unsigned char> rgb;
    DevBuf<unsigned char> rgb_wide;   // f64 rows of the bands a mixed delivery sends by DMA straight into the caller's frame
    DevBuf<unsigned char> hit;
    DevBuf<Counters> counters;
    DevBuf<unsigned> work;  // tile counter of the persistent f32 kernel (+ blocks done, heavy / light tile counts)
    DevBuf<unsigned> order; // tile hand-out order of the small-scene kernel
    DevBuf<unsigned char> upload_arena;   // device copy of rtc_context::pinned_upload, scattered to the scene buffers by one kernel
    cudaEvent_t upload_done = nullptr;    // recorded on `stream` behind that kernel: the blob may be packed again, and launches on
                                          // OTHER streams (rtc_render_device with a caller's stream) wait for it before they read the scene
    DevBuf<unsigned> wait_status;   // rtc_wait_device: 0, or 1 + index of the first flag that did not arrive in time
    bool peer_to_first = false;     // this GPU can store into the context's first GPU (rtc_render_device on several GPUs)
};

// Allocation base of a device pointer (cuMemGetAddressRange, fetched through the runtime so that the
// library does not link libcuda directly).
typedef unsigned long long CUdeviceptr_v;
int cuMemGetAddressRange_shim(void** base, size_t* size, const void* ptr) {
    typedef int (*fn_t)(CUdeviceptr_v*, size_t*, CUdeviceptr_v);
    static fn_t fn = nullptr;
    if (!fn) {
        void* f = nullptr;
        cudaDriverEntryPointQueryResult qr;
        if (cudaGetDriverEntryPoint("cuMemGetAddressRange", &f, cudaEnableDefault, &qr) != cudaSuccess || !f) return -1;
        fn = (fn_t)f;
    }
    CUdeviceptr_v b = 0;
    size_t sz = 0;
    int rc = fn(&b, &sz, (CUdeviceptr_v)(uintptr_t)ptr);
    *base = (void*)(uintptr_t)b;
    *size = sz;
    return rc;
}

size_t elem_size(int out_format) {
    return out_format == RTC_OUT_F64_RGB ? 8 : (out_format == RTC_OUT_F32_RGB ? 4 : 1);
}

}  // namespace

// Host worker pool of a context: splits the pinned-staging -> caller-buffer step (memcpy, or the
// lossless f32 -> f64 widening of fast-mode frames) over the host cores while later bands are still
// in flight over PCIe.
class HostPool {
public:
    explicit HostPool(int n) {
        for (int i = 0; i < n; ++i) workers_.emplace_back([this] { loop(); });
    }
    ~HostPool() {
        { std::lock_guard<std::mutex> l(m_); stop_ = true; }
        stop_flag_.store(true);
        cv_.notify_all();
        for (auto& t : workers_) t.join();
    }
    int size() const { return (int)workers_.size(); }
    // Between begin_burst() and end_burst() (one rtc_render call: a few milliseconds) idle workers poll for work instead
    // of sleeping on the condition variable: a band or a chunk of tiles is ~0.1-0.3 ms of work per thread, the same order
    // as a sleeping thread's wake-up latency (measured at 8 ranks x 4 threads: 12 MB chunks were in effect delivered by
    // the calling thread alone, 1.4-2.0 ms instead of ~0.3 ms).
    void begin_burst() {
        { std::lock_guard<std::mutex> l(m_); burst_.store(true); }
        cv_.notify_all();
    }
    void end_burst() { burst_.store(false); }
    // run fn(i) for i in [0, n) on the pool and the calling thread; returns when all are done
    void parallel_for(int n, const std::function<void(int)>& fn) {
        if (n <= 0) return;
        {
            std::lock_guard<std::mutex> l(m_);
            fn_ = &fn; next_ = 0; total_ = n;
            pending_.store(n);
            avail_.store(1, std::memory_order_release);
        }
        if (!burst_.load()) cv_.notify_all();
        for (;;) {
            int i;
            { std::lock_guard<std::mutex> l(m_); if (next_ >= total_) break; i = next_++; if (next_ >= total_) avail_.store(0); }
            fn(i);
            pending_.fetch_sub(1, std::memory_order_acq_rel);
        }
        while (pending_.load(std::memory_order_acquire) != 0) cpu_relax();   // the last tasks are running on the workers
        std::lock_guard<std::mutex> l(m_);
        fn_ = nullptr; total_ = 0; next_ = 0;
    }
private:
    static void cpu_relax() {
#if defined(__x86_64__) || defined(__i386__)
        __builtin_ia32_pause();
#endif
    }
    void loop() {
        for (;;) {
            while (burst_.load(std::memory_order_relaxed) && !stop_flag_.load(std::memory_order_relaxed) &&
                   avail_.load(std::memory_order_acquire) == 0)
                cpu_relax();
            int i;
            const std::function<void(int)>* fn;
            {
                std::unique_lock<std::mutex> l(m_);
                if (!burst_.load()) cv_.wait(l, [this] { return stop_ || burst_.load() || (fn_ && next_ < total_); });
                if (stop_) return;
                if (!(fn_ && next_ < total_)) continue;
                i = next_++; fn = fn_;
                if (next_ >= total_) avail_.store(0);
            }
            (*fn)(i);
            pending_.fetch_sub(1, std::memory_order_acq_rel);
        }
    }
    std::vector<std::thread> workers_;
    std::mutex m_;
    std::condition_variable cv_;
    const std::function<void(int)>* fn_ = nullptr;
    int next_ = 0, total_ = 0;
    std::atomic<int> pending_{0}, avail_{0};
    std::atomic<bool> burst_{false}, stop_flag_{false};
    bool stop_ = false;
};

struct rtc_context {
    std::vector<DeviceState> devs;
    HostPool* pool = nullptr;
    bool has_scene = false;
    // pinned staging for D2H (full frame), grown on demand
    unsigned char* pinned_rgb = nullptr;
    size_t pinned_rgb_cap = 0;
    unsigned char* pinned_hit = nullptr;
    size_t pinned_hit_cap = 0;
    Counters* pinned_counters = nullptr;  // one per device
    unsigned* pinned_status = nullptr;    // rtc_wait_result reads the wait status word through this
    unsigned char* pinned_upload = nullptr;   // rtc_scene_upload on a one-GPU context: table + all scene arrays, one H2D copy
    size_t pinned_upload_cap = 0;
    std::vector<std::pair<unsigned char*, size_t>> registered;   // rtc_register_host_buffer ranges
    std::vector<bool> registered_owned;                          // false: the range was already page-locked by another owner (not ours to unregister)
    double wait_timeout_s = 60.0;
    uint64_t last_d2h_bytes = 0;   // rtc_delivery_info: what the last rtc_render moved over PCIe, and how
    int last_bands = 0, last_bands_direct = 0;
    double mix_lag_ms = 0.0;    // running mean of (host threads done) - (last byte landed) of the mixed delivery
    int mix_direct_bands = 1;   // mixed delivery (rtc_render, one GPU): how many of the last row bands go by f64 DMA; follows the host's speed
};

namespace {

int ensure_pinned(unsigned char** p, size_t* cap, size_t n) {
    if (n <= *cap) return RTC_OK;
    if (*p) cudaFreeHost(*p);
    *p = nullptr;
    *cap = 0;
    CUDA_TRY(cudaMallocHost((void**)p, n));
    *cap = n;
    return RTC_OK;
}

int validate_opts(const rtc_camera* cam, const rtc_render_opts* o) {
    if (!cam || !o) return fail(RTC_ERR_INVALID_ARG, "null camera or opts");
    if (cam->hsize == 0 || cam->vsize == 0) return fail(RTC_ERR_INVALID_ARG, "empty image");
    if ((uint64_t)cam->hsize * cam->vsize > (1ull << 31)) return fail(RTC_ERR_INVALID_ARG, "image too large");
    if (cam->samples > 255) return fail(RTC_ERR_UNSUPPORTED, "antialiasing_samples is a u8 in the reference (src/camera.rs:26)");
    if (o->precision != RTC_PRECISION_F64_PARITY && o->precision != RTC_PRECISION_F32_FAST)
        return fail(RTC_ERR_INVALID_ARG, "bad precision");
    if (o->rect != RTC_RECT_REFERENCE_RENDER && o->rect != RTC_RECT_FULL) return fail(RTC_ERR_INVALID_ARG, "bad rect");
    if (o->max_depth < 0 || o->max_depth > RTC_MAX_DEPTH) return fail(RTC_ERR_INVALID_ARG, "max_depth out of range");
    if (o->out_format < RTC_OUT_F64_RGB || o->out_format > RTC_OUT_U8_RGB) return fail(RTC_ERR_INVALID_ARG, "bad out_format");
    if (o->shard_count > 1 && (o->shard_index < 0 || o->shard_index >= o->shard_count || o->tile_rows <= 0))
        return fail(RTC_ERR_INVALID_ARG, "bad shard spec");
    return RTC_OK;
}

CamParams make_cam(const rtc_camera* cam, const rtc_render_opts* o, int shard_index, int shard_count, int tile_rows) {
    CamParams c{};
    c.hsize = cam->hsize;
    c.vsize = cam->vsize;
    c.half_width = cam->half_width;
    c.half_height = cam->half_height;
    c.pixel_size = cam->pixel_size;
    for (int i = 0; i < 12; ++i) c.vinv[i] = cam->view_inv[i];
    c.samples = cam->samples;
    c.aa_seed = cam->aa_seed;
    // Camera::render loops 0..vsize-1 x 0..hsize-1 (src/camera.rs:120-121)
    c.xlim = o->rect == RTC_RECT_REFERENCE_RENDER ? cam->hsize - 1 : cam->hsize;
    c.ylim = o->rect == RTC_RECT_REFERENCE_RENDER ? cam->vsize - 1 : cam->vsize;
    c.max_depth = o->max_depth;
    c.shard_index = shard_count > 1 ? shard_index : 0;
    c.shard_count = shard_count > 1 ? shard_count : 1;
    c.tile_rows = tile_rows > 0 ? tile_rows : 16;
    c.shard_rows = shard_num_rows((int)cam->vsize, c.tile_rows, c.shard_index, c.shard_count);
    c.out_format = o->out_format;
    c.dst_full = 0;
    c.row0 = 0;
    return c;
}

// ---- scene flattening ---------------------------------------------------------------------------
struct HostScene {
    std::vector<double2> sph64, pl64;
    std::vector<float4> pl32, us32, gs32;
    std::vector<int32_t> sph_index, pl_index, us_index, gs_index, gset;
    std::vector<double2> cub64; std::vector<float4> cb32; std::vector<int32_t> cub_index;   // cubes: appended to sph64 / gs32 after the spheres
    int n_spheres = 0, n_gsph = 0, n_cubes = 0;
    std::vector<float> plL, usL;          // [n_lights][np], [n_lights][nu]
    std::vector<double> lights64; std::vector<float4> lights32;
    int light_inside = 0;                 // bit k: light k inside some uniform sphere
    std::vector<float4> bound;            // bounding spheres of us32 | gs32 objects (DevScene32::bound)
    float bound_cmax = 0.f, bound_rmax = 0.f;
    std::vector<float4> bound64;          // bounding spheres of the sph64 entries (spheres | cubes; DevScene64::bound)
    float bound64_cmax = 0.f, bound64_rmax = 0.f;
    std::vector<float4> gbound, gbound64; // one bounding sphere per kGroup consecutive entries of bound / bound64
    float gbound_cmax = 0.f, gbound_rmax = 0.f, gbound64_cmax = 0.f, gbound64_rmax = 0.f;
    std::vector<float4> sbound, sbound64; // one bounding sphere per 32 consecutive groups (a block of 256 entries)
    std::vector<float4> gbox, sbox;       // the two upper levels of `bound` as axis-aligned boxes (lo, hi interleaved; DevScene32::gbox)
    float sbound_cmax = 0.f, sbound_rmax = 0.f, sbound64_cmax = 0.f, sbound64_rmax = 0.f;
    std::vector<ShadeRec64> rec64;
    std::vector<ShadeRec32> rec32;
    int ids_unique = 1;
    // small-scene form (render_f32_small.cu)
    bool small_ok = false;
    SmallParams small{};
    std::vector<SmallRec32> srec;
};

inline float __int_as_float_host(int v) {
    float f;
    std::memcpy(&f, &v, 4);
    return f;
}

void mul34(const double* a, const double* b, double* out) {  // (a as 3x4 affine) * (b as 3x4 affine)
    for (int r = 0; r < 3; ++r)
        for (int c = 0; c < 4; ++c) {
            double s = 0.0;
            for (int k = 0; k < 3; ++k) s += a[4 * r + k] * b[4 * k + c];
            if (c == 3) s += a[4 * r + 3];
            out[4 * r + c] = s;
        }
}

// Smallest and largest eigenvalue of a symmetric 3x3 matrix (cyclic Jacobi rotations).
void sym3_eig_minmax(double g[3][3], double& lo, double& hi) {
    for (int sweep = 0; sweep < 32; ++sweep) {
        double off = std::fabs(g[0][1]) + std::fabs(g[0][2]) + std::fabs(g[1][2]);
        if (off <= 1e-300 || off <= 1e-18 * (std::fabs(g[0][0]) + std::fabs(g[1][1]) + std::fabs(g[2][2]))) break;
        for (int p = 0; p < 2; ++p)
            for (int q = p + 1; q < 3; ++q) {
                if (g[p][q] == 0.0) continue;
                double theta = (g[q][q] - g[p][p]) / (2.0 * g[p][q]);
                double t = (theta >= 0 ? 1.0 : -1.0) / (std::fabs(theta) + std::sqrt(theta * theta + 1.0));
                double c = 1.0 / std::sqrt(t * t + 1.0), sn = t * c;
                for (int k = 0; k < 3; ++k) { double a = g[k][p], b = g[k][q]; g[k][p] = c * a - sn * b; g[k][q] = sn * a + c * b; }
                for (int k = 0; k < 3; ++k) { double a = g[p][k], b = g[q][k]; g[p][k] = c * a - sn * b; g[q][k] = sn * a + c * b; }
            }
    }
    lo = std::min(g[0][0], std::min(g[1][1], g[2][2]));
    hi = std::max(g[0][0], std::max(g[1][1], g[2][2]));
}

// World-space bounding sphere of {p : |A p + t|_2 <= unit_radius} (A, t = the shape's inverse, rows 0..2):
// centre = -A^-1 t, radius = unit_radius / sigma_min(A), padded for the f32 rounding of the coefficients
// the exact tests read.  Returns false when A is too close to singular for the bound to mean anything.
bool bounding_sphere(const double* inv, double unit_radius, double c[3], double& radius) {
    const double a[3][3] = {{inv[0], inv[1], inv[2]}, {inv[4], inv[5], inv[6]}, {inv[8], inv[9], inv[10]}};
    const double t[3] = {inv[3], inv[7], inv[11]};
    const double det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1]) - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0]) +
                       a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    if (!(std::fabs(det) > 0.0) || !std::isfinite(det)) return false;
    double m[3][3];   // A^-1 by cofactors
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) {
            const int i1 = (i + 1) % 3, i2 = (i + 2) % 3, j1 = (j + 1) % 3, j2 = (j + 2) % 3;
            m[j][i] = (a[i1][j1] * a[i2][j2] - a[i1][j2] * a[i2][j1]) / det;
        }
    for (int i = 0; i < 3; ++i) c[i] = -(m[i][0] * t[0] + m[i][1] * t[1] + m[i][2] * t[2]);
    double g[3][3];
    for (int i = 0; i < 3; ++i)
        for (int j = 0; j < 3; ++j) g[i][j] = a[0][i] * a[0][j] + a[1][i] * a[1][j] + a[2][i] * a[2][j];
    double lo, hi;
    sym3_eig_minmax(g, lo, hi);
    if (!(lo > 0.0) || !std::isfinite(hi)) return false;
    const double cond = std::sqrt(hi / lo);
    const double cn = std::fabs(c[0]) + std::fabs(c[1]) + std::fabs(c[2]);
    radius = unit_radius / std::sqrt(lo) * (1.0 + 1e-4) + 4.0 * 5.97e-8 * cn * cond;
    return std::isfinite(radius) && std::isfinite(cn);
}

// ---- spatial order + group bounds for big scenes -----------------------------------------------------
// The intersect streams of a scene are read front to back by every ray (brute force, World::intersect
// src/shape.rs:677-683), so their order is free as long as the index arrays follow: hit selection takes the smallest t
// (ties: the parity kernel compares shape indices, the fast kernels take the first in stream order — equal f32 roots of
// different shapes are not a case the tolerance cares about).  Streams with >= kSortMin entries are put in a spatial order
// of the shapes' centres (spatial_order), so that kGroup consecutive entries are neighbours in space and one bounding sphere around
// them (DevScene32::gbound / DevScene64::gbound) lets a ray skip all of them with one nine-FFMA line-distance test.
constexpr size_t kSortMin = 64;
constexpr int kGroup = 8;

std::vector<int> spatial_order(const std::vector<std::array<double, 3>>& c) {
    // Recursive median split along the longest extent of the centres (a k-d order).  Splits fall on multiples of the
    // block size (256 entries) while a node is larger than a block, and on multiples of the group size (8) below, so
    // that every block and every group of the stream is one compact subtree: its bounding box is near-cubic, where runs
    // of a Morton curve are elongated and jump (simulated on C4: 5.4 instead of 11.5 candidate groups per ray;
    // measured: C4 4K fast mode 142 -> 109 ms, parity mode 177 -> 72 ms).
    const size_t n = c.size();
    std::vector<int> perm(n);
    for (size_t i = 0; i < n; ++i) perm[i] = (int)i;
    if (n < kSortMin) return perm;
    for (const auto& p : c)
        for (int k = 0; k < 3; ++k) if (!std::isfinite(p[k])) return perm;
    struct Node { size_t b, e; };
    std::vector<Node> stack{{0, n}};
    while (!stack.empty()) {
        const Node nd = stack.back();
        stack.pop_back();
        const size_t len = nd.e - nd.b;
        if (len <= (size_t)kGroup) continue;
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        for (size_t i = nd.b; i < nd.e; ++i)
            for (int k = 0; k < 3; ++k) { lo[k] = std::min(lo[k], c[perm[i]][k]); hi[k] = std::max(hi[k], c[perm[i]][k]); }
        int ax = 0;
        for (int k = 1; k < 3; ++k) if (hi[k] - lo[k] > hi[ax] - lo[ax]) ax = k;
        auto before = [&](int x, int y) { return c[x][ax] < c[y][ax] || (c[x][ax] == c[y][ax] && x < y); };   // a strict total order
        const size_t unit = len > 256 ? 256 : (size_t)kGroup;
        size_t half = ((len / 2 + unit - 1) / unit) * unit;     // a multiple of the unit at or just above the median
        if (half >= len) half = (len - 1) / unit * unit;        // (len > unit: both sides stay non-empty)
        // Only the SETS on either side of the split matter here: a child that is split again orders its own range, and a
        // child that is a leaf (one group) is sorted below — the permutation is the one a full sort of every node gives
        // (which this was: 20 of the 32 ms rtc_scene_upload spent on 10^4 shapes), at O(n log n).
        std::nth_element(perm.begin() + (long)nd.b, perm.begin() + (long)(nd.b + half), perm.begin() + (long)nd.e, before);
        if (half <= (size_t)kGroup) std::sort(perm.begin() + (long)nd.b, perm.begin() + (long)(nd.b + half), before);
        if (len - half <= (size_t)kGroup) std::sort(perm.begin() + (long)(nd.b + half), perm.begin() + (long)nd.e, before);
        stack.push_back({nd.b, nd.b + half});
        stack.push_back({nd.b + half, nd.e});
    }
    return perm;
}
template <typename T>
void apply_perm(std::vector<T>& v, const std::vector<int>& perm, size_t stride) {
    std::vector<T> out(v.size());
    for (size_t i = 0; i < perm.size(); ++i)
        for (size_t k = 0; k < stride; ++k) out[i * stride + k] = v[(size_t)perm[i] * stride + k];
    v.swap(out);
}
// Bounding spheres (cx, cy, cz, -R^2) of each `group` consecutive entries of `bound` (padding entries have w = +inf and
// are ignored; `bound` is padded to a multiple of `group` first, the result to a multiple of `pad_out`, with never-hit
// entries).  Returns false when there is nothing to group.
bool group_bounds(std::vector<float4>& bound, int group, int pad_out, std::vector<float4>& gbound, float& gcmax, float& grmax) {
    const float4 never = make_float4(0.f, 0.f, 0.f, std::numeric_limits<float>::infinity());
    gbound.clear(); gcmax = 0.f; grmax = 0.f;
    if (bound.empty()) return false;
    while (bound.size() % (size_t)group) bound.push_back(never);
    double cm = 0.0, rm = 0.0;
    for (size_t g = 0; g < bound.size(); g += (size_t)group) {
        double c[3] = {0, 0, 0}; int m = 0;
        for (int k = 0; k < group; ++k) { const float4& b = bound[g + k]; if (std::isinf(b.w)) continue; c[0] += b.x; c[1] += b.y; c[2] += b.z; ++m; }
        if (m == 0) { gbound.push_back(never); continue; }
        for (int k = 0; k < 3; ++k) c[k] /= m;
        double r = 0.0;
        for (int k = 0; k < group; ++k) {
            const float4& b = bound[g + k];
            if (std::isinf(b.w)) continue;
            const double dx = b.x - c[0], dy = b.y - c[1], dz = b.z - c[2];
            r = std::max(r, std::sqrt(dx * dx + dy * dy + dz * dz) + std::sqrt(std::max(0.0, -(double)b.w)));
        }
        const double cn = std::fabs(c[0]) + std::fabs(c[1]) + std::fabs(c[2]);
        r = r * (1.0 + 1e-6) + 2e-7 * cn;   // the centre is rounded to f32
        gbound.push_back(make_float4((float)c[0], (float)c[1], (float)c[2], -(float)(r * r * (1.0 + 1e-6))));
        cm = std::max(cm, cn); rm = std::max(rm, r);
    }
    if (!std::isfinite(cm) || !std::isfinite(rm)) { gbound.clear(); return false; }
    while (gbound.size() % (size_t)pad_out) gbound.push_back(never);
    gcmax = (float)(cm * (1.0 + 1e-6)); grmax = (float)(rm * (1.0 + 1e-6));
    return true;
}

// Axis-aligned boxes around each `group` consecutive bounding spheres (cx, cy, cz, -R^2; w = +inf: padding, ignored), rounded
// outwards to f32; an empty group gets an inverted box.  Output: lo, hi interleaved, padded to a multiple of pad_out entries.
void sphere_boxes(const std::vector<float4>& bound, int group, int pad_out, std::vector<float4>& box) {
    const float inf = std::numeric_limits<float>::infinity();
    box.clear();
    for (size_t g = 0; g < bound.size(); g += (size_t)group) {
        double lo[3] = {1e300, 1e300, 1e300}, hi[3] = {-1e300, -1e300, -1e300};
        bool any = false;
        for (size_t k = g; k < std::min(bound.size(), g + (size_t)group); ++k) {
            const float4& b = bound[k];
            if (std::isinf(b.w)) continue;
            const double r = std::sqrt(std::max(0.0, -(double)b.w)), c[3] = {b.x, b.y, b.z};
            for (int a = 0; a < 3; ++a) { lo[a] = std::min(lo[a], c[a] - r); hi[a] = std::max(hi[a], c[a] + r); }
            any = true;
        }
        if (!any) { box.push_back(make_float4(inf, inf, inf, 0.f)); box.push_back(make_float4(-inf, -inf, -inf, 0.f)); continue; }
        auto down = [](double v) { float f = (float)v; return (double)f > v ? std::nextafterf(f, -std::numeric_limits<float>::infinity()) : f; };
        auto up = [](double v) { float f = (float)v; return (double)f < v ? std::nextafterf(f, std::numeric_limits<float>::infinity()) : f; };
        box.push_back(make_float4(down(lo[0]), down(lo[1]), down(lo[2]), 0.f));
        box.push_back(make_float4(up(hi[0]), up(hi[1]), up(hi[2]), 0.f));
    }
    while ((box.size() / 2) % (size_t)pad_out) { box.push_back(make_float4(inf, inf, inf, 0.f)); box.push_back(make_float4(-inf, -inf, -inf, 0.f)); }
}
// boxes around each `group` consecutive boxes
void box_boxes(const std::vector<float4>& in, int group, int pad_out, std::vector<float4>& box) {
    const float inf = std::numeric_limits<float>::infinity();
    box.clear();
    for (size_t g = 0; g < in.size() / 2; g += (size_t)group) {
        float lo[3] = {inf, inf, inf}, hi[3] = {-inf, -inf, -inf};
        for (size_t k = g; k < std::min(in.size() / 2, g + (size_t)group); ++k) {
            const float4 &l = in[2 * k], &h = in[2 * k + 1];
            if (l.x > h.x) continue;   // inverted: empty
            lo[0] = std::min(lo[0], l.x); lo[1] = std::min(lo[1], l.y); lo[2] = std::min(lo[2], l.z);
            hi[0] = std::max(hi[0], h.x); hi[1] = std::max(hi[1], h.y); hi[2] = std::max(hi[2], h.z);
        }
        box.push_back(make_float4(lo[0], lo[1], lo[2], 0.f));
        box.push_back(make_float4(hi[0], hi[1], hi[2], 0.f));
    }
    while ((box.size() / 2) % (size_t)pad_out) { box.push_back(make_float4(inf, inf, inf, 0.f)); box.push_back(make_float4(-inf, -inf, -inf, 0.f)); }
}

int flatten(const rtc_light* lights, size_t n_lights, const rtc_shape* shapes, size_t n, HostScene& hs) {
    const rtc_light* light = lights;   // light 0 = World.light
    bool seen[256] = {false};
    hs.rec64.resize(n);
    hs.rec32.resize(n);
    for (size_t i = 0; i < n; ++i) {
        const rtc_shape& s = shapes[i];
        if (s.kind != RTC_SHAPE_SPHERE && s.kind != RTC_SHAPE_PLANE && s.kind != RTC_SHAPE_CUBE)
            return fail(RTC_ERR_INVALID_ARG, "shape " + std::to_string(i) + ": bad kind");
        if (s.pattern_kind < RTC_PATTERN_NONE || s.pattern_kind > RTC_PATTERN_GRID)
            return fail(RTC_ERR_INVALID_ARG, "shape " + std::to_string(i) + ": bad pattern kind");
        if (s.pattern_kind == RTC_PATTERN_NONE && !s.has_color)
            return fail(RTC_ERR_UNSUPPORTED, "shape " + std::to_string(i) +
                                                 ": material has neither colour nor pattern (the reference panics, src/material.rs:331)");
        int id = s.world_id & 255;
        if (seen[id]) hs.ids_unique = 0;
        seen[id] = true;

        ShadeRec64& r = hs.rec64[i];
        std::memset(&r, 0, sizeof r);
        for (int k = 0; k < 12; ++k) r.inv[k] = s.inv[k];
        for (int rr = 0; rr < 3; ++rr)
            for (int cc = 0; cc < 3; ++cc)   // Cube::normal_at derives the transpose from the inverse per hit (src/shape.rs:623-629): the stored one is not read
                r.invt[3 * rr + cc] = s.kind == RTC_SHAPE_CUBE ? s.inv[4 * cc + rr] : s.inv_t[4 * rr + cc];
        for (int k = 0; k < 3; ++k) { r.color[k] = s.color[k]; r.pat_a[k] = s.pat_a[k]; r.pat_b[k] = s.pat_b[k]; }
        r.ambient = s.ambient; r.diffuse = s.diffuse; r.specular = s.specular; r.shininess = s.shininess;
        r.reflective = s.reflective; r.transparency = s.transparency; r.refractive_index = s.refractive_index;
        for (int k = 0; k < 12; ++k) r.pat_inv[k] = s.pat_inv[k];
        r.kind = s.kind; r.world_id = id; r.pattern_kind = s.pattern_kind;

        ShadeRec32& q = hs.rec32[i];
        std::memset(&q, 0, sizeof q);
        q.col_amb = make_float4((float)s.color[0], (float)s.color[1], (float)s.color[2], (float)s.ambient);
        q.dif_spec = make_float4((float)s.diffuse, (float)s.specular, (float)s.shininess, (float)s.reflective);
        double comb[12];
        mul34(s.pat_inv, s.inv, comb);  // world point -> pattern space in one apply (fast mode only)
        q.p0 = make_float4((float)comb[0], (float)comb[1], (float)comb[2], (float)comb[3]);
        q.p1 = make_float4((float)comb[4], (float)comb[5], (float)comb[6], (float)comb[7]);
        q.p2 = make_float4((float)comb[8], (float)comb[9], (float)comb[10], (float)comb[11]);
        q.pa = make_float4((float)s.pat_a[0], (float)s.pat_a[1], (float)s.pat_a[2], __int_as_float_host(id));
        q.pb = make_float4((float)s.pat_b[0], (float)s.pat_b[1], (float)s.pat_b[2], 0.f);
        int geom;
        if (s.kind == RTC_SHAPE_SPHERE) {
            for (int rr = 0; rr < 3; ++rr) {
                hs.sph64.push_back(make_double2(s.inv[4 * rr + 0], s.inv[4 * rr + 1]));
                hs.sph64.push_back(make_double2(s.inv[4 * rr + 2], s.inv[4 * rr + 3]));
            }
            hs.sph_index.push_back((int32_t)i);
            // similarity test: M^T M == k I  (M = upper 3x3 of the inverse)  ->  world-space sphere
            const double* m = s.inv;
            double g[3][3];
            for (int a2 = 0; a2 < 3; ++a2)
                for (int b2 = 0; b2 < 3; ++b2) g[a2][b2] = m[a2] * m[b2] + m[4 + a2] * m[4 + b2] + m[8 + a2] * m[8 + b2];
            double k2 = (g[0][0] + g[1][1] + g[2][2]) / 3.0;
            double dev = 0.0;
            for (int a2 = 0; a2 < 3; ++a2)
                for (int b2 = 0; b2 < 3; ++b2) dev = std::max(dev, std::fabs(g[a2][b2] - (a2 == b2 ? k2 : 0.0)));
            // the stored inverse-transpose must also be the true one (Sphere keeps it in sync, src/shape.rs:319-322)
            double devt = 0.0;
            for (int a2 = 0; a2 < 3; ++a2)
                for (int b2 = 0; b2 < 3; ++b2) devt = std::max(devt, std::fabs(s.inv_t[4 * a2 + b2] - s.inv[4 * b2 + a2]));
            if (k2 > 0.0 && dev <= 1e-12 * k2 && devt <= 1e-12 * std::sqrt(k2)) {
                geom = GEOM_USPHERE;
                double r2 = 1.0 / k2, c3[3];
                for (int a2 = 0; a2 < 3; ++a2) c3[a2] = -r2 * (m[a2] * m[3] + m[4 + a2] * m[7] + m[8 + a2] * m[11]);
                hs.us32.push_back(make_float4((float)c3[0], (float)c3[1], (float)c3[2], (float)r2));
                hs.us_index.push_back((int32_t)i);
                q.g0 = make_float4((float)c3[0], (float)c3[1], (float)c3[2], (float)std::sqrt(k2));
            } else {
                geom = GEOM_GSPHERE;
                for (int rr = 0; rr < 3; ++rr) {
                    float4 row = make_float4((float)m[4 * rr], (float)m[4 * rr + 1], (float)m[4 * rr + 2], (float)m[4 * rr + 3]);
                    hs.gs32.push_back(row);
                    (rr == 0 ? q.g0 : rr == 1 ? q.g1 : q.g2) = row;
                    float4 trow = make_float4((float)s.inv_t[4 * rr], (float)s.inv_t[4 * rr + 1], (float)s.inv_t[4 * rr + 2], 0.f);
                    (rr == 0 ? q.t0 : rr == 1 ? q.t1 : q.t2) = trow;
                }
                hs.gs_index.push_back((int32_t)i);
            }
        } else if (s.kind == RTC_SHAPE_CUBE) {
            geom = GEOM_CUBE;
            const double* m = s.inv;
            for (int rr = 0; rr < 3; ++rr) {
                hs.cub64.push_back(make_double2(m[4 * rr + 0], m[4 * rr + 1]));
                hs.cub64.push_back(make_double2(m[4 * rr + 2], m[4 * rr + 3]));
                float4 row = make_float4((float)m[4 * rr], (float)m[4 * rr + 1], (float)m[4 * rr + 2], (float)m[4 * rr + 3]);
                hs.cb32.push_back(row);
                (rr == 0 ? q.g0 : rr == 1 ? q.g1 : q.g2) = row;
                float4 trow = make_float4((float)m[rr], (float)m[4 + rr], (float)m[8 + rr], 0.f);   // row rr of transpose(inverse 3x3)
                (rr == 0 ? q.t0 : rr == 1 ? q.t1 : q.t2) = trow;
            }
            hs.cub_index.push_back((int32_t)i);
        } else {
            geom = GEOM_PLANE;
            hs.pl64.push_back(make_double2(s.inv[4], s.inv[5]));
            hs.pl64.push_back(make_double2(s.inv[6], s.inv[7]));
            hs.pl32.push_back(make_float4((float)s.inv[4], (float)s.inv[5], (float)s.inv[6], (float)s.inv[7]));
            hs.pl_index.push_back((int32_t)i);
            // world normal = normalize(stored inverse-transpose * (0,1,0))  (src/shape.rs:34-40, 481)
            double nx = s.inv_t[1], ny = s.inv_t[5], nz = s.inv_t[9];
            double nm = std::sqrt(nx * nx + ny * ny + nz * nz);
            q.g0 = make_float4((float)(nx / nm), (float)(ny / nm), (float)(nz / nm), 0.f);
        }
        q.tr_ri = make_float4((float)s.transparency, (float)s.refractive_index, __int_as_float_host(s.pattern_kind),
                              __int_as_float_host(geom));
    }
    // big scenes: spatial (k-d) order per stream (see spatial_order)
    {
        auto centres_of = [&](const std::vector<int32_t>& index, double unit_radius) {
            std::vector<std::array<double, 3>> c(index.size());
            for (size_t k = 0; k < index.size(); ++k) {
                double r = 0.0;
                if (!bounding_sphere(shapes[index[k]].inv, unit_radius, c[k].data(), r)) c[k] = {NAN, NAN, NAN};
            }
            return c;
        };
        if (hs.us_index.size() >= kSortMin) {
            std::vector<std::array<double, 3>> c(hs.us32.size());
            for (size_t k = 0; k < c.size(); ++k) c[k] = {hs.us32[k].x, hs.us32[k].y, hs.us32[k].z};
            const std::vector<int> perm = spatial_order(c);
            apply_perm(hs.us32, perm, 1); apply_perm(hs.us_index, perm, 1);
        }
        if (hs.gs_index.size() >= kSortMin) {
            const std::vector<int> perm = spatial_order(centres_of(hs.gs_index, 1.0));
            apply_perm(hs.gs32, perm, 3); apply_perm(hs.gs_index, perm, 1);
        }
        if (hs.cub_index.size() >= kSortMin) {
            const std::vector<int> perm = spatial_order(centres_of(hs.cub_index, std::sqrt(3.0)));
            apply_perm(hs.cb32, perm, 3); apply_perm(hs.cub64, perm, 6); apply_perm(hs.cub_index, perm, 1);
        }
        if (hs.sph_index.size() >= kSortMin) {
            const std::vector<int> perm = spatial_order(centres_of(hs.sph_index, 1.0));
            apply_perm(hs.sph64, perm, 6); apply_perm(hs.sph_index, perm, 1);
        }
    }
    // cubes travel behind the spheres in the same streams (same 3x4 inverse rows; the kernels switch test at the count)
    hs.n_spheres = (int)hs.sph_index.size(); hs.n_gsph = (int)hs.gs_index.size(); hs.n_cubes = (int)hs.cub_index.size();
    hs.sph64.insert(hs.sph64.end(), hs.cub64.begin(), hs.cub64.end());
    hs.sph_index.insert(hs.sph_index.end(), hs.cub_index.begin(), hs.cub_index.end());
    hs.gs32.insert(hs.gs32.end(), hs.cb32.begin(), hs.cb32.end());
    hs.gs_index.insert(hs.gs_index.end(), hs.cub_index.begin(), hs.cub_index.end());
    // stream positions (planes | uniform spheres | general spheres | cubes) of the shapes whose refractive
    // index is not 1.0: the only ones that can change n1/n2 (fast path of compute_refractive)
    const int np = (int)hs.pl_index.size(), nu = (int)hs.us_index.size();
    for (int k = 0; k < np; ++k) if (shapes[hs.pl_index[k]].refractive_index != 1.0) hs.gset.push_back(k);
    for (int k = 0; k < nu; ++k) if (shapes[hs.us_index[k]].refractive_index != 1.0) hs.gset.push_back(np + k);
    for (int k = 0; k < (int)hs.gs_index.size(); ++k) if (shapes[hs.gs_index[k]].refractive_index != 1.0) hs.gset.push_back(np + nu + k);

    // bounding spheres in stream order (uniform spheres are their own bound); a shape without a finite bound
    // switches the filter off for the scene
    {
        bool ok = true;
        double cmax = 0.0, rmax = 0.0;
        for (int k = 0; k < nu; ++k) {
            const float4& u = hs.us32[k];
            hs.bound.push_back(make_float4(u.x, u.y, u.z, -u.w));
            cmax = std::max(cmax, (double)std::fabs(u.x) + std::fabs(u.y) + std::fabs(u.z));
            rmax = std::max(rmax, std::sqrt((double)u.w));
            ok = ok && std::isfinite(u.w) && std::isfinite(cmax);
        }
        for (int k = 0; ok && k < (int)hs.gs_index.size(); ++k) {
            double c[3], r = 0.0;
            ok = bounding_sphere(shapes[hs.gs_index[k]].inv, k < hs.n_gsph ? 1.0 : std::sqrt(3.0), c, r);
            if (!ok) break;
            hs.bound.push_back(make_float4((float)c[0], (float)c[1], (float)c[2], -(float)(r * r)));
            cmax = std::max(cmax, std::fabs(c[0]) + std::fabs(c[1]) + std::fabs(c[2]));
            rmax = std::max(rmax, r);
        }
        ok = ok && std::isfinite(cmax) && std::isfinite(rmax) && cmax < 1e15 && rmax < 1e15;
        if (!ok) hs.bound.clear();
        else {
            hs.bound_cmax = (float)(cmax * (1.0 + 1e-6)); hs.bound_rmax = (float)(rmax * (1.0 + 1e-6));
            // eight entries per group, 32 groups per block (256 entries), padded so that a kernel can read whole blocks
            if (!group_bounds(hs.bound, kGroup, 32, hs.gbound, hs.gbound_cmax, hs.gbound_rmax) ||
                !group_bounds(hs.gbound, 32, 32, hs.sbound, hs.sbound_cmax, hs.sbound_rmax)) hs.bound.clear();
            else { sphere_boxes(hs.bound, kGroup, 32, hs.gbox); box_boxes(hs.gbox, 32, 32, hs.sbox); }
        }
    }

    // the same for the parity kernel's stream (spheres | cubes in sph64 order)
    {
        bool ok = true;
        double cmax = 0.0, rmax = 0.0;
        for (int k = 0; ok && k < (int)hs.sph_index.size(); ++k) {
            double c[3], r = 0.0;
            ok = bounding_sphere(shapes[hs.sph_index[k]].inv, k < hs.n_spheres ? 1.0 : std::sqrt(3.0), c, r);
            if (!ok) break;
            // the centre is rounded to f32 (<= 6e-8 |c| per component): widen the radius by that much
            r += 2e-7 * (std::fabs(c[0]) + std::fabs(c[1]) + std::fabs(c[2]));
            hs.bound64.push_back(make_float4((float)c[0], (float)c[1], (float)c[2], -(float)(r * r * (1.0 + 1e-6))));
            cmax = std::max(cmax, std::fabs(c[0]) + std::fabs(c[1]) + std::fabs(c[2]));
            rmax = std::max(rmax, r);
        }
        ok = ok && std::isfinite(cmax) && std::isfinite(rmax) && cmax < 1e15 && rmax < 1e15;
        if (!ok) hs.bound64.clear();
        else {
            hs.bound64_cmax = (float)(cmax * (1.0 + 1e-6)); hs.bound64_rmax = (float)(rmax * (1.0 + 1e-6));
            if (!group_bounds(hs.bound64, kGroup, 32, hs.gbound64, hs.gbound64_cmax, hs.gbound64_rmax) ||
                !group_bounds(hs.gbound64, 32, 32, hs.sbound64, hs.sbound64_cmax, hs.sbound64_rmax)) hs.bound64.clear();
        }
    }

    // light-only terms of the shadow test (f32, from the f32 coefficients the kernels use), one row per light
    for (size_t li = 0; li < n_lights; ++li) {
        const rtc_light& L = lights[li];
        const float lx = (float)L.position[0], ly = (float)L.position[1], lz = (float)L.position[2];
        for (int k = 0; k < np; ++k) {
            // f(light) == 0: the light lies in the plane.  The reference's shadow ray then meets the plane at t == distance
            // (up to rounding), which fails `h.t < distance` (src/shape.rs:724): such a plane never shadows.  A NaN factor
            // drops out of the kernels' fminf accumulation.
            const rtc_shape& ps = shapes[hs.pl_index[k]];
            const double f1 = ps.inv[4] * L.position[0] + ps.inv[5] * L.position[1] + ps.inv[6] * L.position[2] + ps.inv[7];
            const float4& q = hs.pl32[k];
            hs.plL.push_back(f1 == 0.0 ? std::numeric_limits<float>::quiet_NaN() : fmaf(q.z, lz, fmaf(q.y, ly, fmaf(q.x, lx, q.w))));
        }
        for (int k = 0; k < nu; ++k) {
            const float4& q = hs.us32[k];
            const float dx = lx - q.x, dy = ly - q.y, dz = lz - q.z;
            const float v = fmaf(dz, dz, fmaf(dy, dy, dx * dx)) - q.w;
            hs.usL.push_back(v);
            if (v < 0.f) hs.light_inside |= 1 << li;
        }
        for (int k = 0; k < 3; ++k) hs.lights64.push_back(L.position[k]);
        for (int k = 0; k < 3; ++k) hs.lights64.push_back(L.intensity[k]);
        hs.lights32.push_back(make_float4(lx, ly, lz, 0.f));
        hs.lights32.push_back(make_float4((float)L.intensity[0], (float)L.intensity[1], (float)L.intensity[2], 0.f));
    }

    // ---- small-scene form: constant-bank streams + records by stream position ----
    SmallParams& sp = hs.small;
    std::memset(&sp, 0, sizeof sp);
    bool ok = np <= kSmallMaxPlanes && nu <= kSmallMaxUsph && hs.gs_index.empty() && hs.ids_unique && np + nu > 0 &&
              n_lights == 1;   // the shadow coefficients of this form are folded with the one light
    const double lp[3] = {light->position[0], light->position[1], light->position[2]};
    for (int k = 0; ok && k < nu; ++k) {   // light inside a sphere: the convex-minimum shadow test does not apply
        const float4& u = hs.us32[k];
        double dx = lp[0] - u.x, dy = lp[1] - u.y, dz = lp[2] - u.z;
        if (dx * dx + dy * dy + dz * dz - u.w <= 0.0) ok = false;
    }
    if (ok) {
        for (int k = 0; k < kSmallMaxPlanes; ++k) sp.planes[k] = sp.planes_d[k] = sp.splanes[k] = make_float4(0.f, 0.f, 0.f, 1.f);
        for (int k = 0; k < kSmallMaxUsph; ++k) sp.usph[k] = sp.usph_s[k] = make_float4(0.f, 0.f, 0.f, -1.f);
        hs.srec.assign(kSmallMaxShapes, SmallRec32{});
        sp.np = np; sp.nu = nu; sp.npg = np / 2; sp.nug = nu / 2; sp.np_odd = np & 1; sp.nu_odd = nu & 1;
        // stream position of plane k / sphere k: an odd count's last shape goes to the last slot (SmallParams::np_odd)
        auto plane_slot = [&](int k) { return (k == np - 1 && (np & 1)) ? kSmallMaxPlanes - 1 : k; };
        auto sphere_slot = [&](int k) { return (k == nu - 1 && (nu & 1)) ? kSmallMaxUsph - 1 : k; };
        int n_glass = 0;
        auto fill_common = [&](SmallRec32& r, const rtc_shape& s, const ShadeRec32& q) {
            r.col_amb = q.col_amb; r.dif_spec = q.dif_spec;
            r.tr_ri = make_float4((float)s.transparency, (float)s.refractive_index, __int_as_float_host(s.pattern_kind), 0.f);
            r.p0 = q.p0; r.p1 = q.p1; r.p2 = q.p2;
            r.pa = make_float4(q.pa.x, q.pa.y, q.pa.z, 0.f);
            r.pb = make_float4(q.pb.x, q.pb.y, q.pb.z, s.pattern_kind == RTC_PATTERN_CHECKER ? 1.f : 0.f);
        };
        for (int k = 0; k < np; ++k) {
            const int i = hs.pl_index[k];
            const float4 pl = hs.pl32[k];
            const int slot = plane_slot(k);
            sp.planes[slot] = sp.planes_d[slot] = pl;
            // f(1) = n.light + w: fold it into the coefficients so that the shadow test is splane.(o,1) <= 0
            const double f1 = shapes[i].inv[4] * lp[0] + shapes[i].inv[5] * lp[1] + shapes[i].inv[6] * lp[2] + shapes[i].inv[7];
            const double sc = f1 / (std::fabs(f1) > 0.0 ? std::fabs(f1) : 1.0);   // sign only: keeps the magnitudes of the plane
            sp.splanes[slot] = make_float4((float)(shapes[i].inv[4] * sc), (float)(shapes[i].inv[5] * sc), (float)(shapes[i].inv[6] * sc),
                                           (float)(shapes[i].inv[7] * sc));
            if (f1 == 0.0) sp.splanes[slot] = make_float4(0.f, 0.f, 0.f, 1.f);   // the light lies in the plane: it never shadows (see plL above)
            SmallRec32& r = hs.srec[slot];
            const float4 g0 = hs.rec32[i].g0;
            r.nrm = make_float4(g0.x, g0.y, g0.z, 0.f);
            fill_common(r, shapes[i], hs.rec32[i]);
            if (shapes[i].refractive_index != 1.0) n_glass = -1000;   // a glass plane: no fast n1/n2 test
        }
        for (int k = 0; k < nu; ++k) {
            const int i = hs.us_index[k];
            const int slot = sphere_slot(k);
            sp.usph[slot] = sp.usph_s[slot] = hs.us32[k];
            SmallRec32& r = hs.srec[kSmallMaxPlanes + slot];
            const float4 g0 = hs.rec32[i].g0;   // (c, 1/r)
            r.nrm = make_float4(-g0.x * g0.w, -g0.y * g0.w, -g0.z * g0.w, g0.w);
            fill_common(r, shapes[i], hs.rec32[i]);
            if (shapes[i].refractive_index != 1.0) {
                if (n_glass >= 0 && n_glass < kSmallMaxGlass) sp.glass[n_glass] = hs.us32[k];
                ++n_glass;
            }
        }
        sp.n_glass = (n_glass < 0 || n_glass > kSmallMaxGlass) ? -1 : n_glass;
        for (int k = 0; k < 3; ++k) { sp.light_pos[k] = (float)light->position[k]; sp.light_int[k] = (float)light->intensity[k]; }
    }
    hs.small_ok = ok;
    return RTC_OK;
}

template <typename T>
cudaError_t upload(DevBuf<T>& b, const std::vector<T>& v, cudaStream_t st) {
    cudaError_t e = b.ensure(v.size());
    if (e != cudaSuccess || v.empty()) return e;
    return cudaMemcpyAsync(b.p, v.data(), v.size() * sizeof(T), cudaMemcpyHostToDevice, st);
}

// rtc_scene_upload on a one-GPU context: the ~20 scene arrays of a small scene are ~10 KB in all, and 20 separate copies
// (each finished before the call returns: the host vectors die) were 60 us on the critical path of every end-to-end step.
// They travel as ONE pinned blob — a table of (destination, offset, bytes) followed by the arrays — and one kernel scatters
// the blob to the scene buffers; the call returns without waiting (the blob is the context's, packed again only after
// upload_done).  All arrays are 4-byte granular and the offsets 256-byte aligned.
struct ScatterEntry { unsigned long long dst, off, bytes; };
__global__ void scatter_upload_kernel(const unsigned char* __restrict__ blob, int n) {
    const ScatterEntry* tab = reinterpret_cast<const ScatterEntry*>(blob);
    for (int e = blockIdx.y; e < n; e += gridDim.y) {
        const ScatterEntry t = tab[e];
        const uint4* s = reinterpret_cast<const uint4*>(blob + t.off);
        uint4* d = reinterpret_cast<uint4*>(t.dst);
        const size_t n16 = t.bytes / 16;
        for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) d[i] = s[i];
        const size_t tail4 = (t.bytes - n16 * 16) / 4;
        if (blockIdx.x == 0 && threadIdx.x < tail4)
            reinterpret_cast<unsigned*>(t.dst)[n16 * 4 + threadIdx.x] = reinterpret_cast<const unsigned*>(blob + t.off)[n16 * 4 + threadIdx.x];
    }
}
struct UploadBatch {
    struct Item { void* dst; const void* src; size_t bytes; };
    std::vector<Item> items;
    cudaError_t err = cudaSuccess;
    template <typename T>
    void add(DevBuf<T>& b, const std::vector<T>& v) {
        static_assert(sizeof(T) % 4 == 0, "4-byte granular arrays");
        if (err != cudaSuccess) return;
        err = b.ensure(v.size());
        if (err == cudaSuccess && !v.empty()) items.push_back({b.p, v.data(), v.size() * sizeof(T)});
    }
};

}  // namespace

// ---- de-interleave kernel (rtc_assemble_device) --------------------------------------------------
namespace {
template <typename U>
__global__ void assemble_rows_kernel(const U* __restrict__ gathered, U* __restrict__ frame, int vsize, size_t row_units,
                                     int tile_rows, int shard_count, size_t shard_stride_units) {
    int y = blockIdx.y;
    if (y >= vsize) return;
    int k = y / tile_rows;              // frame tile
    int shard = k % shard_count;
    int local_tile = k / shard_count;
    int compact_row = local_tile * tile_rows + (y - k * tile_rows);
    const U* src = gathered + (size_t)shard * shard_stride_units + (size_t)compact_row * row_units;
    U* dst = frame + (size_t)y * row_units;
    for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < row_units; i += (size_t)gridDim.x * blockDim.x) dst[i] = src[i];
}

// ---- RGBA8 pack (rtc_pack_rgba8_device): Canvas::to_imgbuf, src/canvas.rs:61-79 ----------------------
// channel = Color::scale(component.powf(1/gamma), 255) (src/color.rs:55-65,100-114): `as i32` truncates toward
// zero, saturates and maps NaN to 0, then the value is clamped to [0, 255]; alpha = 255.  HBM-bound:
// 24 (f64) or 12 (f32) bytes read and 4 written per pixel; a thread packs four pixels — 16-byte loads, one
// 16-byte store.
__device__ __forceinline__ unsigned pack_channel(double c, double inv_gamma, bool use_gamma) {
    if (use_gamma) c = pow(c, inv_gamma);
    const double v = c * 255.0;
    int q = v != v ? 0 : (v >= 2147483647.0 ? 2147483647 : (v <= -2147483648.0 ? (-2147483647 - 1) : (int)v));
    return (unsigned)(q < 0 ? 0 : (q > 255 ? 255 : q));
}
template <typename T>
__global__ void __launch_bounds__(256) pack_rgba8_kernel(const T* __restrict__ rgb, uint32_t* __restrict__ out, size_t n_pixels,
                                                         double inv_gamma, int use_gamma) {
    const size_t n_quads = n_pixels / 4;
    for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < n_quads; q += (size_t)gridDim.x * blockDim.x) {
        double c[12];
        if (sizeof(T) == 8) {
            const double2* src = reinterpret_cast<const double2*>(rgb) + 6 * q;
#pragma unroll
            for (int k = 0; k < 6; ++k) { const double2 v = src[k]; c[2 * k] = v.x; c[2 * k + 1] = v.y; }
        } else {
            const float4* src = reinterpret_cast<const float4*>(rgb) + 3 * q;
#pragma unroll
            for (int k = 0; k < 3; ++k) { const float4 v = src[k]; c[4 * k] = v.x; c[4 * k + 1] = v.y; c[4 * k + 2] = v.z; c[4 * k + 3] = v.w; }
        }
        uint4 px;
        unsigned* w = &px.x;
#pragma unroll
        for (int k = 0; k < 4; ++k)
            w[k] = pack_channel(c[3 * k], inv_gamma, use_gamma) | (pack_channel(c[3 * k + 1], inv_gamma, use_gamma) << 8) |
                   (pack_channel(c[3 * k + 2], inv_gamma, use_gamma) << 16) | 0xff000000u;
        reinterpret_cast<uint4*>(out)[q] = px;
    }
    // the last n_pixels % 4 pixels
    const size_t tail = n_quads * 4 + (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (tail < n_pixels)
        out[tail] = pack_channel((double)rgb[3 * tail], inv_gamma, use_gamma) | (pack_channel((double)rgb[3 * tail + 1], inv_gamma, use_gamma) << 8) |
                    (pack_channel((double)rgb[3 * tail + 2], inv_gamma, use_gamma) << 16) | 0xff000000u;
}

// ---- completion flags (rtc_signal_device / rtc_wait_device) ---------------------------------------
__global__ void signal_kernel(unsigned* flag, unsigned value) {
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}
// A wait that runs out of time records which flag is missing (1 + its index, the smallest one wins) and returns: the
// stream goes on, the CUDA context stays usable (a trap would poison it: every buffer and the scene of the waiting
// process lost), and the host learns about it from rtc_wait_result.
// then_flag (rtc_wait_signal_device, one block): once every flag has arrived, then_value is release-stored to it by the same
// kernel — "frame consumed" for a rank 0 that has no consumer of its own between frames, without a second launch.
__global__ void wait_kernel(const unsigned* flags, int n, unsigned value, unsigned long long timeout_ns, unsigned* status,
                            unsigned* then_flag = nullptr, unsigned then_value = 0u) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    bool arrived = true;
    if (i < n) {
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (;;) {
            unsigned v;
            asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(flags + i) : "memory");
            if ((int)(v - value) >= 0) break;
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > timeout_ns) { atomicMin(status, (unsigned)i + 1u); arrived = false; break; }
            __nanosleep(200);
        }
    }
    if (then_flag) {   // (single block)
        const int all = __syncthreads_and(arrived ? 1 : 0);
        if (threadIdx.x == 0 && all) {
            __threadfence_system();
            asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(then_flag), "r"(then_value) : "memory");
        }
    }
}

// ---- FMA peak micro-benchmark --------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(256) fma_peak_kernel(T* out, int iters, T seed) {
    T a0 = seed + (T)threadIdx.x, a1 = a0 + (T)1, a2 = a0 + (T)2, a3 = a0 + (T)3, a4 = a0 + (T)4, a5 = a0 + (T)5, a6 = a0 + (T)6, a7 = a0 + (T)7;
    const T m = (T)0.999, c = (T)0.001;
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
            a0 = fma(a0, m, c); a1 = fma(a1, m, c); a2 = fma(a2, m, c); a3 = fma(a3, m, c);
            a4 = fma(a4, m, c); a5 = fma(a5, m, c); a6 = fma(a6, m, c); a7 = fma(a7, m, c);
        }
    }
    T s = ((a0 + a1) + (a2 + a3)) + ((a4 + a5) + (a6 + a7));
    if (s == (T)-12345.678) out[0] = s;  // never true: keeps the chains alive
}

template <typename T>
int run_peak(DeviceState& d, double* out_tflops) {
    T* dummy = nullptr;
    CUDA_TRY(cudaMalloc(&dummy, sizeof(T)));
    int sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, d.dev));
    const int blocks = sms * 8, iters = 4096;
    cudaEvent_t e0, e1;
    CUDA_TRY(cudaEventCreate(&e0));
    CUDA_TRY(cudaEventCreate(&e1));
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        CUDA_TRY(cudaEventRecord(e0, d.stream));
        fma_peak_kernel<T><<<blocks, 256, 0, d.stream>>>(dummy, iters, (T)1.0);
        CUDA_TRY(cudaEventRecord(e1, d.stream));
        CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, e0, e1));
        if (rep > 0) best = std::min(best, ms);
    }
    CUDA_TRY(cudaGetLastError());
    double flops = 2.0 * 64.0 * (double)iters * 256.0 * (double)blocks;
    *out_tflops = flops / ((double)best * 1e-3) / 1e12;
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(dummy);
    return RTC_OK;
}

// Staging -> caller buffer for `n_elems` elements: plain copy, or f32 -> f64 widening (exact) when the
// device produced f32 for an F64_RGB request (fast mode: halves the PCIe bytes).
// The context's host workers: all hardware threads (at most 32) for a process that owns the machine, a share of them
// when `share` rank processes render shards of one frame side by side.  RTC_B200_HOST_THREADS overrides.
void ensure_pool(rtc_context* ctx, int share) {
    if (ctx->pool) return;
    unsigned hw = std::thread::hardware_concurrency();
    if (const char* env = std::getenv("RTC_B200_HOST_THREADS")) hw = (unsigned)std::max(1, std::atoi(env));
    else hw = std::max(1u, std::min(hw, 32u) / (unsigned)std::max(1, share));
    ctx->pool = new HostPool((int)hw - 1);
}
struct Burst {   // idle workers poll instead of sleeping while one render call is in flight (HostPool::begin_burst)
    HostPool* p;
    explicit Burst(HostPool* pool) : p(pool) { if (p) p->begin_burst(); }
    ~Burst() { if (p) p->end_burst(); }
    Burst(const Burst&) = delete;
    Burst& operator=(const Burst&) = delete;
};

void host_deliver(rtc_context* ctx, unsigned char* dst, const unsigned char* src, size_t n_elems, size_t src_elem, size_t dst_elem) {
    ensure_pool(ctx, 1);
    const size_t kChunk = 256u << 10;   // elements per task
    const int n_tasks = (int)((n_elems + kChunk - 1) / kChunk);
    const bool widen = src_elem == 4 && dst_elem == 8;
    auto task = [=](int i) {
        const size_t b = (size_t)i * kChunk, e = std::min(n_elems, b + kChunk);
        if (widen) {
            rtc_host_widen_f32_f64(reinterpret_cast<double*>(dst) + b, reinterpret_cast<const float*>(src) + b, e - b);   // host_simd.cpp
        } else {
            std::memcpy(dst + b * dst_elem, src + b * src_elem, (e - b) * dst_elem);
        }
    };
    if (n_tasks <= 1) { if (n_tasks == 1) task(0); return; }
    ctx->pool->parallel_for(n_tasks, task);
}

// Camera part of the small-scene kernel parameter: Camera::ray_for_pixel_offset (src/camera.rs:64-76) reduced to
//   dir = normalize(cam_dx (x + xo) + cam_dy (y + yo) + cam_d0),   origin = Vinv * (0,0,0)
void fill_small_camera(SmallParams& P, const CamParams& cam) {
    const double* v = cam.vinv;
    for (int k = 0; k < 3; ++k) {
        P.cam_o[k] = (float)v[4 * k + 3];
        P.cam_dx[k] = (float)(-cam.pixel_size * v[4 * k + 0]);
        P.cam_dy[k] = (float)(-cam.pixel_size * v[4 * k + 1]);
        P.cam_d0[k] = (float)(cam.half_width * v[4 * k + 0] + cam.half_height * v[4 * k + 1] - v[4 * k + 2]);
    }
    P.hsize = (int)cam.hsize; P.shard_rows = cam.shard_rows; P.xlim = (int)cam.xlim; P.ylim = (int)cam.ylim;
    P.shard_index = cam.shard_index; P.shard_count = cam.shard_count; P.tile_rows = cam.tile_rows; P.dst_full = cam.dst_full; P.row0 = cam.row0;
    P.n_sub = cam.samples == 1 ? 1 : 4;
    P.n_extra = cam.samples >= 2 ? (int)cam.samples : 0;
    P.aa_seed = cam.aa_seed;
    P.max_depth = cam.max_depth;
    P.tiles_x = ((int)cam.hsize + 7) / 8;
    P.total_tiles = P.tiles_x * ((cam.shard_rows + 3) / 4);
}

// rtc_render_device_synced: start after *wait_flag >= wait_value, release-store done_value to *done_flag when the pixels are out
struct SyncSpec {
    const unsigned* wait_flag; unsigned wait_value;
    unsigned* done_flag; unsigned done_value;
    unsigned long long timeout_ns;
};

int launch_on(DeviceState& d, const CamParams& cp, int precision, int flags, void* d_rgb, uint8_t* d_hit, Counters* d_counters,
              cudaStream_t st, const SyncSpec* sync = nullptr) {
    cudaError_t e;
    if (st != d.stream && d.upload_done && (e = cudaStreamWaitEvent(st, d.upload_done, 0)) != cudaSuccess)   // rtc_scene_upload no longer waits
        return fail(RTC_ERR_CUDA, std::string("stream order behind the scene upload: ") + cudaGetErrorString(e));
    const bool fused = sync && precision == RTC_PRECISION_F32_FAST && d.small_ok && !(flags & RTC_OPT_F32_GENERAL_KERNEL);
    if (sync && !fused && sync->wait_flag) {   // kernels without the fused form: the flag kernels around the launch
        wait_kernel<<<1, 32, 0, st>>>(sync->wait_flag, 1, sync->wait_value, sync->timeout_ns, d.wait_status.p);
    }
    // Fast mode promises the tolerance, not a data type.  Beyond ~128 shapes the fastest kernel that meets it is the parity
    // kernel in its hierarchical form (per-lane three-level walk over the k-d ordered streams, 8 blocks per SM), and it is
    // exact: the general f32 kernel tests every bounding sphere of a stream held in shared memory, linear in the scene
    // (1920x1080, n random spheres, general f32 kernel / parity kernel: n = 100 1.9 / 2.2 ms, 150 2.9 / 2.2, 400 7.4 / 3.3,
    // 1500 29 / 6.2, 3000 96 / 8.7; profiles/r02_mid_scenes.txt), and on streams too large for shared memory its
    // mixed-precision kRefine form needs 109 ms where the parity kernel needs 56 (C4 4K; profiles/r02_experiments.txt).
    // Both stay reachable through RTC_OPT_F32_GENERAL_KERNEL.  The parity kernel converts to the requested output format.
    const bool big_to_parity = precision == RTC_PRECISION_F32_FAST && !(flags & RTC_OPT_F32_GENERAL_KERNEL) && !d.small_ok &&
                               d.s64.n_bound4 > 0 && (f32_streams_are_global(d.s32) || d.s64.n_shapes > 128);
    if (precision == RTC_PRECISION_F64_PARITY || big_to_parity) {
        DevScene64 s64 = d.s64;
        if (flags & RTC_OPT_NO_BOUND_FILTER) s64.n_bound4 = 0;
        e = launch_render_f64(s64, cp, d_rgb, d_hit, d_counters, st);
    }
    else if (d.small_ok && !(flags & RTC_OPT_F32_GENERAL_KERNEL)) {
        SmallParams P = d.small;
        fill_small_camera(P, cp);
        P.wait_flag = nullptr; P.done_flag = nullptr; P.wait_status = d.wait_status.p; P.wait_value = P.done_value = 0u; P.wait_timeout_ns = 0ull;
        if (fused) {
            P.wait_flag = sync->wait_flag; P.wait_value = sync->wait_value; P.done_flag = sync->done_flag; P.done_value = sync->done_value;
            P.wait_timeout_ns = sync->timeout_ns;
        }
        P.order = nullptr;
        if (P.max_depth > 0 && P.n_glass != 0 && P.total_tiles >= 4096 && !(flags & RTC_OPT_LINEAR_TILE_ORDER)) {
            if (d.order.ensure((size_t)P.total_tiles) != cudaSuccess) return fail(RTC_ERR_CUDA, "tile order buffer");
            P.order = d.order.p;
        }
        e = launch_render_f32_small(P, cp.out_format, d_rgb, d_hit, d_counters, d.work.p, st);
    } else {
        DevScene32 s32 = d.s32;
        if (flags & RTC_OPT_NO_BOUND_FILTER) s32.n_bound4 = 0;
        e = launch_render_f32(s32, cp, d_rgb, d_hit, d_counters, d.work.p, st);
    }
    if (e != cudaSuccess) return fail(RTC_ERR_CUDA, std::string("kernel launch: ") + cudaGetErrorString(e));
    if (sync && !fused && sync->done_flag) {
        signal_kernel<<<1, 1, 0, st>>>(sync->done_flag, sync->done_value);
        if ((e = cudaGetLastError()) != cudaSuccess) return fail(RTC_ERR_CUDA, std::string("signal launch: ") + cudaGetErrorString(e));
    }
    return RTC_OK;
}

}  // namespace

extern "C" {

int rtc_abi_version(void) { return RTC_ABI_VERSION; }

const char* rtc_last_error(void) { return g_err.c_str(); }

int rtc_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

int rtc_context_create(const int* device_ids, int n_gpus, rtc_context** out_ctx) {
    if (!out_ctx || n_gpus < 1) return fail(RTC_ERR_INVALID_ARG, "rtc_context_create: bad arguments");
    *out_ctx = nullptr;
    int avail = rtc_device_count();
    if (avail == 0) return fail(RTC_ERR_NO_DEVICE, "no CUDA device available (there is no CPU fallback)");
    rtc_context* ctx = new rtc_context();
    ctx->devs.resize(n_gpus);
    for (int g = 0; g < n_gpus; ++g) {
        DeviceState& d = ctx->devs[g];
        d.dev = device_ids ? device_ids[g] : g;
        if (d.dev < 0 || d.dev >= avail) {
            delete ctx;
            return fail(RTC_ERR_INVALID_ARG, "device id out of range");
        }
        cudaError_t e = cudaSetDevice(d.dev);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d.stream, cudaStreamNonBlocking);
        if (e == cudaSuccess) e = cudaEventCreate(&d.ev_k0);
        if (e == cudaSuccess) e = cudaEventCreate(&d.ev_k1);
        if (e == cudaSuccess) e = cudaEventCreate(&d.ev_c0);
        if (e == cudaSuccess) e = cudaEventCreate(&d.ev_c1);
        if (e == cudaSuccess) e = cudaStreamCreateWithFlags(&d.copy_stream, cudaStreamNonBlocking);
        for (int b = 0; b < kMaxBands && e == cudaSuccess; ++b) {
            e = cudaEventCreateWithFlags(&d.ev_band[b], cudaEventDisableTiming);
            if (e == cudaSuccess) e = cudaEventCreateWithFlags(&d.ev_copy[b], cudaEventDisableTiming);
        }
        if (e == cudaSuccess) e = d.counters.ensure(1);
        if (e == cudaSuccess) e = d.work.ensure(8);
        if (e == cudaSuccess) e = cudaMemset(d.work.p, 0, 8 * sizeof(unsigned));
        if (e == cudaSuccess) e = preload_render_f64();      // no launch of the library ever loads code (rtc_common.h)
        if (e == cudaSuccess) e = preload_render_f32();
        if (e == cudaSuccess) e = preload_render_f32_small();
        if (e == cudaSuccess) { cudaFuncAttributes at; e = cudaFuncGetAttributes(&at, scatter_upload_kernel); }
        if (e == cudaSuccess) {
            cudaFuncAttributes at;
            const void* own[] = {(const void*)signal_kernel, (const void*)wait_kernel, (const void*)assemble_rows_kernel<uint4>,
                                 (const void*)assemble_rows_kernel<uint32_t>, (const void*)assemble_rows_kernel<unsigned char>,
                                 (const void*)pack_rgba8_kernel<double>, (const void*)pack_rgba8_kernel<float>,
                                 (const void*)fma_peak_kernel<float>, (const void*)fma_peak_kernel<double>};
            for (const void* k : own) if (e == cudaSuccess) e = cudaFuncGetAttributes(&at, k);
        }
        if (e == cudaSuccess) e = noalloc_setup_current_device();   // rtc_render_device_noalloc's static scratch: its address on this device
        if (e == cudaSuccess) e = d.wait_status.ensure(1);
        if (e == cudaSuccess) e = cudaMemset(d.wait_status.p, 0xff, sizeof(unsigned));
        if (e == cudaSuccess && g > 0) {   // peer stores into the first GPU's frame (rtc_render_device on several GPUs)
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, d.dev, ctx->devs[0].dev) == cudaSuccess && can) {
                cudaError_t pe = cudaDeviceEnablePeerAccess(ctx->devs[0].dev, 0);
                if (pe == cudaSuccess || pe == cudaErrorPeerAccessAlreadyEnabled) d.peer_to_first = true;
            }
            cudaGetLastError();
        }
        if (e != cudaSuccess) {
            std::string m = cudaGetErrorString(e);
            rtc_context_destroy(ctx);
            return fail(RTC_ERR_CUDA, "rtc_context_create: " + m);
        }
    }
    cudaError_t e = cudaMallocHost((void**)&ctx->pinned_counters, sizeof(Counters) * n_gpus);
    if (e == cudaSuccess) e = cudaMallocHost((void**)&ctx->pinned_status, sizeof(unsigned));
    if (e != cudaSuccess) {
        rtc_context_destroy(ctx);
        return fail(RTC_ERR_CUDA, std::string("rtc_context_create: ") + cudaGetErrorString(e));
    }
    *out_ctx = ctx;
    return RTC_OK;
}

void rtc_context_destroy(rtc_context* ctx) {
    if (!ctx) return;
    for (DeviceState& d : ctx->devs) {
        cudaSetDevice(d.dev);
        if (d.stream) cudaStreamSynchronize(d.stream);
        d.sph64.release(); d.pl64.release(); d.pl32.release(); d.us32.release(); d.gs32.release(); d.us_index.release(); d.gs_index.release(); d.gset.release(); d.plL.release(); d.usL.release(); d.lights64.release(); d.lights32.release(); d.bound64.release(); d.gbound.release(); d.gbound64.release(); d.sbound.release(); d.sbound64.release(); d.bound.release(); d.gbox.release(); d.sbox.release();
        d.sph_index.release(); d.pl_index.release(); d.rec64.release(); d.rec32.release(); d.srec.release();
        d.rgb.release(); d.rgb_wide.release(); d.hit.release(); d.counters.release(); d.work.release(); d.order.release(); d.wait_status.release(); d.upload_arena.release();
        if (d.ev_k0) cudaEventDestroy(d.ev_k0);
        if (d.ev_k1) cudaEventDestroy(d.ev_k1);
        if (d.ev_c0) cudaEventDestroy(d.ev_c0);
        if (d.upload_done) cudaEventDestroy(d.upload_done);
        if (d.ev_c1) cudaEventDestroy(d.ev_c1);
        for (int b = 0; b < kMaxBands; ++b) {
            if (d.ev_band[b]) cudaEventDestroy(d.ev_band[b]);
            if (d.ev_copy[b]) cudaEventDestroy(d.ev_copy[b]);
        }
        if (d.copy_stream) { cudaStreamSynchronize(d.copy_stream); cudaStreamDestroy(d.copy_stream); }
        if (d.stream) cudaStreamDestroy(d.stream);
    }
    if (ctx->pinned_rgb) cudaFreeHost(ctx->pinned_rgb);
    if (ctx->pinned_upload) cudaFreeHost(ctx->pinned_upload);
    if (ctx->pinned_hit) cudaFreeHost(ctx->pinned_hit);
    if (ctx->pinned_counters) cudaFreeHost(ctx->pinned_counters);
    if (ctx->pinned_status) cudaFreeHost(ctx->pinned_status);
    for (size_t i = 0; i < ctx->registered.size(); ++i) if (ctx->registered_owned[i]) cudaHostUnregister(ctx->registered[i].first);
    delete ctx->pool;
    delete ctx;
}

int rtc_scene_upload(rtc_context* ctx, const rtc_light* light, const rtc_shape* shapes, size_t n_shapes) {
    return rtc_scene_upload_lights(ctx, light, 1, shapes, n_shapes);
}

int rtc_scene_upload_lights(rtc_context* ctx, const rtc_light* light, size_t n_lights, const rtc_shape* shapes, size_t n_shapes) {
    if (!ctx || !light || (!shapes && n_shapes)) return fail(RTC_ERR_INVALID_ARG, "rtc_scene_upload: null argument");
    if (n_lights < 1 || n_lights > RTC_MAX_LIGHTS) return fail(RTC_ERR_INVALID_ARG, "rtc_scene_upload_lights: 1 <= n_lights <= RTC_MAX_LIGHTS");
    HostScene hs;
    int rc = flatten(light, n_lights, shapes, n_shapes, hs);
    if (rc != RTC_OK) return rc;
    for (DeviceState& d : ctx->devs) {
        CUDA_TRY(cudaSetDevice(d.dev));
        const bool batched = ctx->devs.size() == 1;   // several GPUs: the plain copies (every device needs its own table)
        UploadBatch batch;
        if (batched) {
            batch.add(d.sph64, hs.sph64);
            batch.add(d.pl64, hs.pl64);
            batch.add(d.pl32, hs.pl32);
            batch.add(d.us32, hs.us32);
            batch.add(d.gs32, hs.gs32);
            batch.add(d.us_index, hs.us_index);
            batch.add(d.gs_index, hs.gs_index);
            batch.add(d.gset, hs.gset);
            batch.add(d.plL, hs.plL);
            batch.add(d.usL, hs.usL);
            batch.add(d.lights64, hs.lights64);
            batch.add(d.lights32, hs.lights32);
            batch.add(d.bound, hs.bound);
            batch.add(d.bound64, hs.bound64);
            batch.add(d.gbound, hs.gbound);
            batch.add(d.gbound64, hs.gbound64);
            batch.add(d.sbound, hs.sbound);
            batch.add(d.sbound64, hs.sbound64);
            batch.add(d.gbox, hs.gbox);
            batch.add(d.sbox, hs.sbox);
            batch.add(d.sph_index, hs.sph_index);
            batch.add(d.pl_index, hs.pl_index);
            batch.add(d.rec64, hs.rec64);
            batch.add(d.rec32, hs.rec32);
            if (hs.small_ok) batch.add(d.srec, hs.srec);
            CUDA_TRY(batch.err);
        } else {
            CUDA_TRY(upload(d.sph64, hs.sph64, d.stream));
            CUDA_TRY(upload(d.pl64, hs.pl64, d.stream));
            CUDA_TRY(upload(d.pl32, hs.pl32, d.stream));
            CUDA_TRY(upload(d.us32, hs.us32, d.stream));
            CUDA_TRY(upload(d.gs32, hs.gs32, d.stream));
            CUDA_TRY(upload(d.us_index, hs.us_index, d.stream));
            CUDA_TRY(upload(d.gs_index, hs.gs_index, d.stream));
            CUDA_TRY(upload(d.gset, hs.gset, d.stream));
            CUDA_TRY(upload(d.plL, hs.plL, d.stream));
            CUDA_TRY(upload(d.usL, hs.usL, d.stream));
            CUDA_TRY(upload(d.lights64, hs.lights64, d.stream));
            CUDA_TRY(upload(d.lights32, hs.lights32, d.stream));
            CUDA_TRY(upload(d.bound, hs.bound, d.stream));
            CUDA_TRY(upload(d.bound64, hs.bound64, d.stream));
            CUDA_TRY(upload(d.gbound, hs.gbound, d.stream));
            CUDA_TRY(upload(d.gbound64, hs.gbound64, d.stream));
            CUDA_TRY(upload(d.sbound, hs.sbound, d.stream));
            CUDA_TRY(upload(d.sbound64, hs.sbound64, d.stream));
            CUDA_TRY(upload(d.gbox, hs.gbox, d.stream));
            CUDA_TRY(upload(d.sbox, hs.sbox, d.stream));
            CUDA_TRY(upload(d.sph_index, hs.sph_index, d.stream));
            CUDA_TRY(upload(d.pl_index, hs.pl_index, d.stream));
            CUDA_TRY(upload(d.rec64, hs.rec64, d.stream));
            CUDA_TRY(upload(d.rec32, hs.rec32, d.stream));
            if (hs.small_ok) CUDA_TRY(upload(d.srec, hs.srec, d.stream));
        }
        d.small_ok = hs.small_ok;
        if (hs.small_ok) {
            d.small = hs.small;
            d.small.rec = d.srec.p;
        }
        d.s64.n_shapes = d.s32.n_shapes = (int)n_shapes;
        d.s64.n_spheres = hs.n_spheres; d.s64.n_cubes = hs.n_cubes;
        d.s32.n_usph = (int)hs.us_index.size();
        d.s32.n_gsph = hs.n_gsph; d.s32.n_cube = hs.n_cubes;
        d.s64.n_planes = d.s32.n_planes = (int)hs.pl_index.size();
        d.s64.ids_unique = d.s32.ids_unique = hs.ids_unique;
        d.s64.sph_rows = d.sph64.p; d.s64.pl_rows = d.pl64.p; d.s64.sph_index = d.sph_index.p; d.s64.pl_index = d.pl_index.p;
        d.s64.rec = d.rec64.p;
        d.s32.planes = d.pl32.p; d.s32.pl_index = d.pl_index.p; d.s32.usph = d.us32.p; d.s32.us_index = d.us_index.p;
        d.s32.gsph = d.gs32.p; d.s32.gs_index = d.gs_index.p;
        d.s32.gset = d.gset.p; d.s32.n_gset = (int)hs.gset.size();
        d.s32.plL = d.plL.p; d.s32.usL = d.usL.p; d.s32.light_inside = hs.light_inside;
        d.s64.n_lights = d.s32.n_lights = (int)n_lights;
        d.s64.lights = d.lights64.p; d.s32.lights = d.lights32.p;
        d.s32.rec = d.rec32.p;
        d.s32.rec64 = d.rec64.p;
        d.s64.bound = d.bound64.p; d.s64.n_bound4 = (int)hs.bound64.size(); d.s64.bound_cmax = hs.bound64_cmax; d.s64.bound_rmax = hs.bound64_rmax;
        d.s64.gbound = d.gbound64.p; d.s64.gbound_cmax = hs.gbound64_cmax; d.s64.gbound_rmax = hs.gbound64_rmax;
        d.s32.gbound = d.gbound.p; d.s32.gbound_cmax = hs.gbound_cmax; d.s32.gbound_rmax = hs.gbound_rmax;
        d.s64.sbound = d.sbound64.p; d.s64.sbound_cmax = hs.sbound64_cmax; d.s64.sbound_rmax = hs.sbound64_rmax;
        d.s32.sbound = d.sbound.p; d.s32.sbound_cmax = hs.sbound_cmax; d.s32.sbound_rmax = hs.sbound_rmax;
        d.s32.gbox = d.gbox.p; d.s32.sbox = d.sbox.p;
        d.s32.bound = d.bound.p; d.s32.n_bound4 = (int)hs.bound.size(); d.s32.bound_cmax = hs.bound_cmax; d.s32.bound_rmax = hs.bound_rmax;
        for (int k = 0; k < 3; ++k) {
            d.s64.light_pos[k] = light->position[k]; d.s64.light_int[k] = light->intensity[k];
            d.s32.light_pos[k] = (float)light->position[k]; d.s32.light_int[k] = (float)light->intensity[k];
        }
        if (batched) {
            size_t total = (batch.items.size() * sizeof(ScatterEntry) + 255) / 256 * 256, max_bytes = 0;
            std::vector<ScatterEntry> table;
            for (const auto& it : batch.items) { table.push_back({(unsigned long long)(uintptr_t)it.dst, (unsigned long long)total, (unsigned long long)it.bytes}); total += (it.bytes + 255) / 256 * 256; max_bytes = std::max(max_bytes, it.bytes); }
            if (!d.upload_done) CUDA_TRY(cudaEventCreateWithFlags(&d.upload_done, cudaEventDisableTiming));
            else CUDA_TRY(cudaEventSynchronize(d.upload_done));          // the previous upload's copy has left the blob
            if (total > ctx->pinned_upload_cap) {
                if (ctx->pinned_upload) cudaFreeHost(ctx->pinned_upload);
                ctx->pinned_upload = nullptr; ctx->pinned_upload_cap = 0;
                CUDA_TRY(cudaMallocHost((void**)&ctx->pinned_upload, total));
                ctx->pinned_upload_cap = total;
            }
            CUDA_TRY(d.upload_arena.ensure(total));
            if (!table.empty()) {
                std::memcpy(ctx->pinned_upload, table.data(), table.size() * sizeof(ScatterEntry));
                for (size_t k = 0; k < batch.items.size(); ++k) std::memcpy(ctx->pinned_upload + table[k].off, batch.items[k].src, batch.items[k].bytes);
                CUDA_TRY(cudaMemcpyAsync(d.upload_arena.p, ctx->pinned_upload, total, cudaMemcpyHostToDevice, d.stream));
                const unsigned gx = (unsigned)std::max<size_t>(1, std::min<size_t>(64, (max_bytes / 16 + 255) / 256));
                scatter_upload_kernel<<<dim3(gx, (unsigned)std::min<size_t>(table.size(), 64)), 256, 0, d.stream>>>(d.upload_arena.p, (int)table.size());
                CUDA_TRY(cudaGetLastError());
            }
            CUDA_TRY(cudaEventRecord(d.upload_done, d.stream));
            // no wait: the blob is the context's, and everything that reads the scene is ordered behind d.stream
        } else {
            // the host vectors die at return: finish the copies now
            CUDA_TRY(cudaStreamSynchronize(d.stream));
        }
    }
    ctx->has_scene = true;
    return RTC_OK;
}

int rtc_shard_rows(const rtc_camera* cam, const rtc_render_opts* opts) {
    if (!cam || !opts) return -1;
    if (opts->shard_count <= 1) return (int)cam->vsize;
    return shard_num_rows((int)cam->vsize, opts->tile_rows > 0 ? opts->tile_rows : 16, opts->shard_index, opts->shard_count);
}

int rtc_render_device(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts, void* d_out_rgb,
                      uint8_t* d_out_hit_mask, void* stream, rtc_stats* stats) {
    if (!ctx || !d_out_rgb) return fail(RTC_ERR_INVALID_ARG, "rtc_render_device: null argument");
    int rc = validate_opts(cam, opts);
    if (rc != RTC_OK) return rc;
    if (!ctx->has_scene) return fail(RTC_ERR_NO_SCENE, "rtc_render_device before rtc_scene_upload");
    DeviceState& d = ctx->devs[0];
    const int G = (int)ctx->devs.size();
    if (G > 1) {
        // Several GPUs in one process: GPU g renders tiles g, g + G, ... and its kernel stores the pixels straight into the
        // frame on the first GPU through peer access (NVLink) while it is still tracing; events order the launches after
        // what `stream` held before and let `stream` continue when every GPU is done.  No staging buffer, no collective.
        for (int g = 1; g < G; ++g)
            if (!ctx->devs[g].peer_to_first)
                return fail(RTC_ERR_UNSUPPORTED, "rtc_render_device: GPU " + std::to_string(ctx->devs[g].dev) + " has no peer access to GPU " + std::to_string(d.dev));
        const int tile_rows = opts->tile_rows > 0 ? opts->tile_rows : 16;
        CUDA_TRY(cudaSetDevice(d.dev));
        cudaStream_t s0 = stream ? (cudaStream_t)stream : d.stream;
        auto t0 = std::chrono::steady_clock::now();
        const bool counters = !stream && stats;
        CUDA_TRY(cudaEventRecord(d.ev_band[0], s0));                    // "everything enqueued before this call"
        for (int g = G - 1; g >= 0; --g) {                              // the first GPU last: its stream then waits for the others
            DeviceState& dg = ctx->devs[g];
            CUDA_TRY(cudaSetDevice(dg.dev));
            cudaStream_t sg = g == 0 ? s0 : dg.stream;
            if (g > 0) CUDA_TRY(cudaStreamWaitEvent(sg, d.ev_band[0], 0));
            if (counters) CUDA_TRY(cudaMemsetAsync(dg.counters.p, 0, sizeof(Counters), sg));
            CamParams cp = make_cam(cam, opts, g, G, tile_rows);
            cp.dst_full = 1;
            if (g == 0) CUDA_TRY(cudaEventRecord(d.ev_k0, s0));
            rc = launch_on(dg, cp, opts->precision, opts->flags | (g > 0 && G > 4 ? RTC_OPT_LINEAR_TILE_ORDER : 0)   /* more than four GPUs: an even pixel stream into the first GPU's NVLink ingress */, d_out_rgb, d_out_hit_mask,
                           counters ? dg.counters.p : nullptr, sg);
            if (rc != RTC_OK) return rc;
            if (counters) CUDA_TRY(cudaMemcpyAsync(&ctx->pinned_counters[g], dg.counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, sg));
            if (g > 0) CUDA_TRY(cudaEventRecord(dg.ev_k1, sg));
        }
        for (int g = 1; g < G; ++g) CUDA_TRY(cudaStreamWaitEvent(s0, ctx->devs[g].ev_k1, 0));
        CUDA_TRY(cudaEventRecord(d.ev_k1, s0));
        if (stream) return RTC_OK;
        CUDA_TRY(cudaStreamSynchronize(s0));
        if (stats) {
            std::memset(stats, 0, sizeof *stats);
            float ms = 0.f;
            CUDA_TRY(cudaEventElapsedTime(&ms, d.ev_k0, d.ev_k1));
            for (int g = 0; g < G; ++g) {
                const Counters& c = ctx->pinned_counters[g];
                stats->rays += c.rays; stats->primary_hits += c.primary_hits; stats->shaded_hits += c.shaded_hits; stats->shadow_rays += c.shadow_rays;
            }
            stats->kernel_launches = (uint64_t)G;
            stats->kernel_ms = ms;
            stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        }
        return RTC_OK;
    }
    CUDA_TRY(cudaSetDevice(d.dev));
    CamParams cp = make_cam(cam, opts, opts->shard_index, opts->shard_count, opts->tile_rows);
    cp.dst_full = (opts->flags & RTC_OPT_DST_FULL_FRAME) ? 1 : 0;
    if (stream) {
        return launch_on(d, cp, opts->precision, opts->flags, d_out_rgb, d_out_hit_mask, nullptr, (cudaStream_t)stream);
    }
    auto t0 = std::chrono::steady_clock::now();
    CUDA_TRY(cudaMemsetAsync(d.counters.p, 0, sizeof(Counters), d.stream));
    CUDA_TRY(cudaEventRecord(d.ev_k0, d.stream));
    rc = launch_on(d, cp, opts->precision, opts->flags, d_out_rgb, d_out_hit_mask, d.counters.p, d.stream);
    if (rc != RTC_OK) return rc;
    CUDA_TRY(cudaEventRecord(d.ev_k1, d.stream));
    CUDA_TRY(cudaMemcpyAsync(&ctx->pinned_counters[0], d.counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, d.stream));
    CUDA_TRY(cudaStreamSynchronize(d.stream));
    if (stats) {
        std::memset(stats, 0, sizeof *stats);
        float ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&ms, d.ev_k0, d.ev_k1));
        const Counters& c = ctx->pinned_counters[0];
        stats->rays = c.rays; stats->primary_hits = c.primary_hits; stats->shaded_hits = c.shaded_hits; stats->shadow_rays = c.shadow_rays;
        stats->kernel_launches = 1;
        stats->kernel_ms = ms;
        stats->total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    }
    return RTC_OK;
}

int rtc_assemble_device(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts, const void* d_gathered,
                        int shard_stride_rows, int elem_kind, void* d_frame, void* stream) {
    if (!ctx || !cam || !opts || !d_gathered || !d_frame) return fail(RTC_ERR_INVALID_ARG, "rtc_assemble_device: null argument");
    if (opts->shard_count < 1 || opts->tile_rows <= 0 || shard_stride_rows <= 0) return fail(RTC_ERR_INVALID_ARG, "rtc_assemble_device: bad shard spec");
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    cudaStream_t st = stream ? (cudaStream_t)stream : d.stream;
    size_t row_bytes = elem_kind < 0 ? (size_t)cam->hsize : (size_t)cam->hsize * 3 * elem_size(elem_kind);
    size_t stride_bytes = row_bytes * (size_t)shard_stride_rows;
    dim3 grid(8, cam->vsize);
    bool a16 = row_bytes % 16 == 0 && ((uintptr_t)d_gathered % 16 == 0) && ((uintptr_t)d_frame % 16 == 0);
    bool a4 = row_bytes % 4 == 0 && ((uintptr_t)d_gathered % 4 == 0) && ((uintptr_t)d_frame % 4 == 0);
    if (a16)
        assemble_rows_kernel<uint4><<<grid, 256, 0, st>>>((const uint4*)d_gathered, (uint4*)d_frame, (int)cam->vsize, row_bytes / 16,
                                                          opts->tile_rows, opts->shard_count, stride_bytes / 16);
    else if (a4)
        assemble_rows_kernel<uint32_t><<<grid, 256, 0, st>>>((const uint32_t*)d_gathered, (uint32_t*)d_frame, (int)cam->vsize,
                                                             row_bytes / 4, opts->tile_rows, opts->shard_count, stride_bytes / 4);
    else
        assemble_rows_kernel<unsigned char><<<grid, 256, 0, st>>>((const unsigned char*)d_gathered, (unsigned char*)d_frame,
                                                                  (int)cam->vsize, row_bytes, opts->tile_rows, opts->shard_count, stride_bytes);
    CUDA_TRY(cudaGetLastError());
    if (!stream) CUDA_TRY(cudaStreamSynchronize(st));
    return RTC_OK;
}

int rtc_ipc_export(rtc_context* ctx, const void* d_ptr, void* handle, uint64_t* offset) {
    if (!ctx || !d_ptr || !handle || !offset) return fail(RTC_ERR_INVALID_ARG, "rtc_ipc_export: null argument");
    static_assert(sizeof(cudaIpcMemHandle_t) == RTC_IPC_HANDLE_BYTES, "handle size");
    CUDA_TRY(cudaSetDevice(ctx->devs[0].dev));
    CUdeviceptr_v base = 0;  // allocation base: the handle always refers to the start of the allocation
    {
        cudaPointerAttributes at;
        CUDA_TRY(cudaPointerGetAttributes(&at, d_ptr));
        if (at.type != cudaMemoryTypeDevice) return fail(RTC_ERR_INVALID_ARG, "rtc_ipc_export: not a device pointer");
        void* b = nullptr;
        size_t sz = 0;
        if (cuMemGetAddressRange_shim(&b, &sz, d_ptr) != 0) return fail(RTC_ERR_CUDA, "rtc_ipc_export: cannot find the allocation base");
        base = (CUdeviceptr_v)(uintptr_t)b;
    }
    cudaIpcMemHandle_t h;
    CUDA_TRY(cudaIpcGetMemHandle(&h, (void*)(uintptr_t)base));
    std::memcpy(handle, &h, sizeof h);
    *offset = (uint64_t)((uintptr_t)d_ptr - (uintptr_t)base);
    return RTC_OK;
}

int rtc_ipc_open(rtc_context* ctx, const void* handle, uint64_t offset, void** out_d_ptr) {
    if (!ctx || !handle || !out_d_ptr) return fail(RTC_ERR_INVALID_ARG, "rtc_ipc_open: null argument");
    CUDA_TRY(cudaSetDevice(ctx->devs[0].dev));
    cudaIpcMemHandle_t h;
    std::memcpy(&h, handle, sizeof h);
    void* base = nullptr;
    CUDA_TRY(cudaIpcOpenMemHandle(&base, h, cudaIpcMemLazyEnablePeerAccess));
    *out_d_ptr = (char*)base + offset;
    return RTC_OK;
}

int rtc_ipc_close(rtc_context* ctx, void* d_ptr, uint64_t offset) {
    if (!ctx || !d_ptr) return fail(RTC_ERR_INVALID_ARG, "rtc_ipc_close: null argument");
    CUDA_TRY(cudaSetDevice(ctx->devs[0].dev));
    CUDA_TRY(cudaIpcCloseMemHandle((char*)d_ptr - offset));
    return RTC_OK;
}

int rtc_signal_device(rtc_context* ctx, void* d_flag, uint32_t value, void* stream) {
    if (!ctx || !d_flag) return fail(RTC_ERR_INVALID_ARG, "rtc_signal_device: null argument");
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    cudaStream_t st = stream ? (cudaStream_t)stream : d.stream;
    signal_kernel<<<1, 1, 0, st>>>((unsigned*)d_flag, value);
    CUDA_TRY(cudaGetLastError());
    return RTC_OK;
}

int rtc_wait_device(rtc_context* ctx, const void* d_flags, int n, uint32_t value, void* stream) {
    if (!ctx || !d_flags || n < 0) return fail(RTC_ERR_INVALID_ARG, "rtc_wait_device: bad argument");
    if (n == 0) return RTC_OK;
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    cudaStream_t st = stream ? (cudaStream_t)stream : d.stream;
    const unsigned long long timeout_ns = (unsigned long long)(ctx->wait_timeout_s * 1e9);
    wait_kernel<<<(n + 31) / 32, 32, 0, st>>>((const unsigned*)d_flags, n, value, timeout_ns, d.wait_status.p);
    CUDA_TRY(cudaGetLastError());
    return RTC_OK;
}

int rtc_wait_signal_device(rtc_context* ctx, const void* d_flags, int n, uint32_t value, void* d_then_flag, uint32_t then_value, void* stream) {
    if (!ctx || !d_flags || !d_then_flag || n < 1 || n > 1024) return fail(RTC_ERR_INVALID_ARG, "rtc_wait_signal_device: bad argument (1 <= n <= 1024)");
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    cudaStream_t st = stream ? (cudaStream_t)stream : d.stream;
    const unsigned long long timeout_ns = (unsigned long long)(ctx->wait_timeout_s * 1e9);
    wait_kernel<<<1, (n + 31) / 32 * 32, 0, st>>>((const unsigned*)d_flags, n, value, timeout_ns, d.wait_status.p, (unsigned*)d_then_flag, then_value);
    CUDA_TRY(cudaGetLastError());
    return RTC_OK;
}

int rtc_set_wait_timeout(rtc_context* ctx, double seconds) {
    if (!ctx || !(seconds > 0.0) || seconds > 86400.0) return fail(RTC_ERR_INVALID_ARG, "rtc_set_wait_timeout: bad argument");
    ctx->wait_timeout_s = seconds;
    return RTC_OK;
}

int rtc_wait_result(rtc_context* ctx, void* stream, int* first_missing) {
    if (!ctx) return fail(RTC_ERR_INVALID_ARG, "rtc_wait_result: null context");
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    cudaStream_t st = stream ? (cudaStream_t)stream : d.stream;
    CUDA_TRY(cudaMemcpyAsync(ctx->pinned_status, d.wait_status.p, sizeof(unsigned), cudaMemcpyDeviceToHost, st));
    CUDA_TRY(cudaStreamSynchronize(st));
    const unsigned v = *ctx->pinned_status;
    if (first_missing) *first_missing = v == 0xffffffffu ? -1 : (int)v - 1;
    if (v == 0xffffffffu) return RTC_OK;
    CUDA_TRY(cudaMemsetAsync(d.wait_status.p, 0xff, sizeof(unsigned), st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return fail(RTC_ERR_TIMEOUT, "rtc_wait_device: flag " + std::to_string(v - 1) + " did not arrive within " + std::to_string(ctx->wait_timeout_s) + " s");
}

int rtc_render_device_synced(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts, void* d_out_rgb, uint8_t* d_out_hit_mask,
                             void* stream, const void* d_wait_flag, uint32_t wait_value, void* d_done_flag, uint32_t done_value) {
    if (!ctx || !d_out_rgb || !stream) return fail(RTC_ERR_INVALID_ARG, "rtc_render_device_synced: null argument (a stream is required)");
    int rc = validate_opts(cam, opts);
    if (rc != RTC_OK) return rc;
    if (!ctx->has_scene) return fail(RTC_ERR_NO_SCENE, "rtc_render_device_synced before rtc_scene_upload");
    if (ctx->devs.size() != 1) return fail(RTC_ERR_UNSUPPORTED, "rtc_render_device_synced: one-GPU contexts (one process per GPU)");
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    CamParams cp = make_cam(cam, opts, opts->shard_index, opts->shard_count, opts->tile_rows);
    cp.dst_full = (opts->flags & RTC_OPT_DST_FULL_FRAME) ? 1 : 0;
    SyncSpec sy{(const unsigned*)d_wait_flag, wait_value, (unsigned*)d_done_flag, done_value, (unsigned long long)(ctx->wait_timeout_s * 1e9)};
    return launch_on(d, cp, opts->precision, opts->flags, d_out_rgb, d_out_hit_mask, nullptr, (cudaStream_t)stream, &sy);
}

int rtc_render_device_noalloc(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts, void* d_out_rgb, uint8_t* d_out_hit_mask,
                              void* stream) {
    if (!ctx || !d_out_rgb || !stream) return fail(RTC_ERR_INVALID_ARG, "rtc_render_device_noalloc: null argument (a stream is required)");
    int rc = validate_opts(cam, opts);
    if (rc != RTC_OK) return rc;
    if (!ctx->has_scene) return fail(RTC_ERR_NO_SCENE, "rtc_render_device_noalloc before rtc_scene_upload");
    if (ctx->devs.size() != 1) return fail(RTC_ERR_UNSUPPORTED, "rtc_render_device_noalloc: one-GPU contexts only");
    DeviceState& d = ctx->devs[0];
    if (opts->precision != RTC_PRECISION_F32_FAST || !d.small_ok || (opts->flags & RTC_OPT_F32_GENERAL_KERNEL))
        return fail(RTC_ERR_UNSUPPORTED, "rtc_render_device_noalloc: only the small-scene fast kernel has a variant without allocations "
                                         "(RTC_PRECISION_F32_FAST, a scene that fits its constant-bank form: <= 16 planes, <= 48 similarity-transformed spheres, one light)");
    CUDA_TRY(cudaSetDevice(d.dev));
    CamParams cp = make_cam(cam, opts, opts->shard_index, opts->shard_count, opts->tile_rows);
    cp.dst_full = (opts->flags & RTC_OPT_DST_FULL_FRAME) ? 1 : 0;
    if (cp.shard_rows <= 0) return RTC_OK;
    SmallParams P = d.small;
    fill_small_camera(P, cp);
    P.wait_flag = nullptr; P.done_flag = nullptr; P.wait_status = d.wait_status.p; P.wait_value = P.done_value = 0u; P.wait_timeout_ns = 0ull;
    if (P.total_tiles > kNoAllocMaxTiles)
        return fail(RTC_ERR_UNSUPPORTED, "rtc_render_device_noalloc: the frame has " + std::to_string(P.total_tiles) + " tiles of 8x4 pixels, the static scratch holds " +
                                             std::to_string(kNoAllocMaxTiles) + " (RTC_NOALLOC_MAX_TILES; 7680x4320 fits)");
    const bool tile_order = P.max_depth > 0 && P.n_glass != 0 && P.total_tiles >= 4096 && !(opts->flags & RTC_OPT_LINEAR_TILE_ORDER);
    if ((cudaStream_t)stream != d.stream && d.upload_done) CUDA_TRY(cudaStreamWaitEvent((cudaStream_t)stream, d.upload_done, 0));   // the scene upload does not wait
    cudaError_t e = launch_render_f32_small_noalloc(P, tile_order, cp.out_format, d_out_rgb, d_out_hit_mask, (cudaStream_t)stream);
    if (e != cudaSuccess) return fail(RTC_ERR_CUDA, std::string("rtc_render_device_noalloc: ") + cudaGetErrorString(e));
    return RTC_OK;
}

int rtc_register_host_buffer(rtc_context* ctx, void* ptr, size_t bytes) {
    if (!ctx || !ptr || bytes == 0) return fail(RTC_ERR_INVALID_ARG, "rtc_register_host_buffer: bad argument");
    for (auto& r : ctx->registered)
        if (r.first == (unsigned char*)ptr) return r.second >= bytes ? RTC_OK : fail(RTC_ERR_INVALID_ARG, "rtc_register_host_buffer: already registered with a smaller size");
    CUDA_TRY(cudaSetDevice(ctx->devs[0].dev));
    cudaError_t e = cudaHostRegister(ptr, bytes, cudaHostRegisterPortable);
    bool owned = true;
    if (e == cudaErrorHostMemoryAlreadyRegistered) {
        // Page-locked already — by another context of this process or by the caller (cudaHostRegister / cudaMallocHost).  That
        // is all the direct delivery needs when the whole range is: check both ends, and leave the registration to its owner.
        cudaGetLastError();
        cudaPointerAttributes a0, a1;
        if (cudaPointerGetAttributes(&a0, ptr) != cudaSuccess || cudaPointerGetAttributes(&a1, (unsigned char*)ptr + bytes - 1) != cudaSuccess ||
            a0.type != cudaMemoryTypeHost || a1.type != cudaMemoryTypeHost) {
            cudaGetLastError();
            return fail(RTC_ERR_INVALID_ARG, "rtc_register_host_buffer: the range overlaps memory another owner registered only in part");
        }
        owned = false;
    } else if (e != cudaSuccess) { cudaGetLastError(); return fail(RTC_ERR_CUDA, std::string("cudaHostRegister: ") + cudaGetErrorString(e)); }
    ctx->registered.emplace_back((unsigned char*)ptr, bytes);
    ctx->registered_owned.push_back(owned);
    return RTC_OK;
}

int rtc_unregister_host_buffer(rtc_context* ctx, void* ptr) {
    if (!ctx || !ptr) return fail(RTC_ERR_INVALID_ARG, "rtc_unregister_host_buffer: bad argument");
    for (size_t i = 0; i < ctx->registered.size(); ++i)
        if (ctx->registered[i].first == (unsigned char*)ptr) {
            for (DeviceState& d : ctx->devs) { cudaSetDevice(d.dev); cudaStreamSynchronize(d.stream); cudaStreamSynchronize(d.copy_stream); }
            if (ctx->registered_owned[i]) CUDA_TRY(cudaHostUnregister(ptr));
            ctx->registered.erase(ctx->registered.begin() + (long)i);
            ctx->registered_owned.erase(ctx->registered_owned.begin() + (long)i);
            return RTC_OK;
        }
    return fail(RTC_ERR_INVALID_ARG, "rtc_unregister_host_buffer: not a registered buffer");
}

int rtc_scene_selfcheck(const rtc_light* lights, size_t n_lights, const rtc_shape* shapes, size_t n_shapes, rtc_selfcheck* out) {
    if (!lights || !out || (!shapes && n_shapes)) return fail(RTC_ERR_INVALID_ARG, "rtc_scene_selfcheck: null argument");
    if (n_lights < 1 || n_lights > RTC_MAX_LIGHTS) return fail(RTC_ERR_INVALID_ARG, "rtc_scene_selfcheck: 1 <= n_lights <= RTC_MAX_LIGHTS");
    HostScene hs;
    int rc = flatten(lights, n_lights, shapes, n_shapes, hs);
    if (rc != RTC_OK) return rc;
    std::memset(out, 0, sizeof *out);
    out->n_planes = (int32_t)hs.pl_index.size();
    out->n_uniform_spheres = (int32_t)hs.us_index.size();
    out->n_general_spheres = hs.n_gsph;
    out->n_cubes = hs.n_cubes;
    out->small_scene_form = hs.small_ok ? 1 : 0;
    auto is_perm = [&](std::initializer_list<const std::vector<int32_t>*> parts) {
        std::vector<int> seen(n_shapes, 0);
        size_t total = 0;
        for (const auto* v : parts)
            for (int32_t i : *v) { if (i < 0 || (size_t)i >= n_shapes) return false; ++seen[i]; ++total; }
        if (total != n_shapes) return false;
        for (int c : seen) if (c != 1) return false;
        return true;
    };
    out->streams_are_permutations = (is_perm({&hs.pl_index, &hs.us_index, &hs.gs_index}) && is_perm({&hs.pl_index, &hs.sph_index})) ? 1 : 0;
    auto ascending = [](const std::vector<int32_t>& v, size_t b, size_t e) { for (size_t k = b + 1; k < e; ++k) if (v[k] < v[k - 1]) return false; return true; };
    out->morton_sorted = (!ascending(hs.us_index, 0, hs.us_index.size()) || !ascending(hs.sph_index, 0, (size_t)hs.n_spheres) ||
                          !ascending(hs.gs_index, 0, (size_t)hs.n_gsph)) ? 1 : 0;
    auto worst = [](const std::vector<float4>& child, const std::vector<float4>& parent, size_t per) {
        double w = -1.0;
        for (size_t i = 0; i < child.size(); ++i) {
            const float4& c = child[i];
            if (std::isinf(c.w)) continue;
            const float4& p = parent[i / per];
            if (std::isinf(p.w)) return 1e300;   // a real entry under a never-hit parent
            const double dx = (double)c.x - p.x, dy = (double)c.y - p.y, dz = (double)c.z - p.z;
            const double rc = std::sqrt(std::max(0.0, -(double)c.w)), rp = std::sqrt(std::max(0.0, -(double)p.w));
            w = std::max(w, (std::sqrt(dx * dx + dy * dy + dz * dz) + rc - rp) / std::max(rp, 1e-300));
        }
        return w;
    };
    if (!hs.bound.empty()) {
        out->n_groups_f32 = (int32_t)(hs.bound.size() / kGroup); out->n_blocks_f32 = (int32_t)((hs.bound.size() / kGroup + 31) / 32);
        out->worst_enclosure_f32 = std::max(worst(hs.bound, hs.gbound, kGroup), worst(hs.gbound, hs.sbound, 32));
        // the box form of the two upper levels: every entry's bounding sphere inside its group's box, every group's box inside its block's
        double wb = -1.0;
        for (size_t i = 0; i < hs.bound.size(); ++i) {
            const float4& c = hs.bound[i];
            if (std::isinf(c.w)) continue;
            const float4 &l = hs.gbox[2 * (i / kGroup)], &h = hs.gbox[2 * (i / kGroup) + 1];
            const double r = std::sqrt(std::max(0.0, -(double)c.w)), p[3] = {c.x, c.y, c.z}, lo[3] = {l.x, l.y, l.z}, hi[3] = {h.x, h.y, h.z};
            for (int a = 0; a < 3; ++a) wb = std::max(wb, std::max(lo[a] - (p[a] - r), (p[a] + r) - hi[a]) / std::max(r, 1e-300));
        }
        for (size_t g = 0; g < hs.gbox.size() / 2; ++g) {
            const float4 &l = hs.gbox[2 * g], &h = hs.gbox[2 * g + 1];
            if (l.x > h.x) continue;
            const float4 &L = hs.sbox[2 * (g / 32)], &H = hs.sbox[2 * (g / 32) + 1];
            const double over = std::max({(double)L.x - l.x, (double)L.y - l.y, (double)L.z - l.z, (double)h.x - H.x, (double)h.y - H.y, (double)h.z - H.z});
            wb = std::max(wb, over > 0.0 ? 1.0 : -1.0);
        }
        out->worst_enclosure_f32 = std::max(out->worst_enclosure_f32, wb);
    }
    if (!hs.bound64.empty()) {
        out->n_groups_f64 = (int32_t)(hs.bound64.size() / kGroup); out->n_blocks_f64 = (int32_t)((hs.bound64.size() / kGroup + 31) / 32);
        out->worst_enclosure_f64 = std::max(worst(hs.bound64, hs.gbound64, kGroup), worst(hs.gbound64, hs.sbound64, 32));
    }
    return RTC_OK;
}

int rtc_measure_peak_tflops(rtc_context* ctx, int precision, double* out_tflops) {
    if (!ctx || !out_tflops) return fail(RTC_ERR_INVALID_ARG, "rtc_measure_peak_tflops: null argument");
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    return precision == RTC_PRECISION_F64_PARITY ? run_peak<double>(d, out_tflops) : run_peak<float>(d, out_tflops);
}

}  // extern "C"

// ---- rtc_render: host buffers in, host buffers out ------------------------------------------------
extern "C" int rtc_pack_rgba8_device(rtc_context* ctx, const void* d_frame, int frame_format, size_t n_pixels, float gamma,
                                     void* d_out_rgba, void* stream) {
    if (!ctx || !d_frame || !d_out_rgba) return fail(RTC_ERR_INVALID_ARG, "rtc_pack_rgba8_device: null argument");
    if (frame_format != RTC_OUT_F64_RGB && frame_format != RTC_OUT_F32_RGB) return fail(RTC_ERR_INVALID_ARG, "rtc_pack_rgba8_device: frame must be F64_RGB or F32_RGB");
    if (((uintptr_t)d_frame | (uintptr_t)d_out_rgba) % 16 != 0) return fail(RTC_ERR_INVALID_ARG, "rtc_pack_rgba8_device: buffers must be 16-byte aligned");
    if (n_pixels == 0) return RTC_OK;
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    cudaStream_t st = stream ? (cudaStream_t)stream : d.stream;
    int sms = 0;
    CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, d.dev));
    const size_t want = (n_pixels / 4 + 255) / 256 + 1;
    const int grid = (int)std::max<size_t>(1, std::min<size_t>(want, (size_t)sms * 8));   // a multiple of the SM count, grid-stride beyond
    // gamma.recip() is computed in f32 and widened (src/color.rs:56: `gamma.recip().into()`)
    const double inv_gamma = (double)(1.0f / gamma);
    const int use_gamma = gamma != 1.0f;   // powf(x, 1.0) == x
    if (frame_format == RTC_OUT_F64_RGB) pack_rgba8_kernel<double><<<grid, 256, 0, st>>>((const double*)d_frame, (uint32_t*)d_out_rgba, n_pixels, inv_gamma, use_gamma);
    else pack_rgba8_kernel<float><<<grid, 256, 0, st>>>((const float*)d_frame, (uint32_t*)d_out_rgba, n_pixels, inv_gamma, use_gamma);
    CUDA_TRY(cudaGetLastError());
    if (!stream) CUDA_TRY(cudaStreamSynchronize(st));
    return RTC_OK;
}

extern "C" int rtc_download_frame(rtc_context* ctx, const void* d_frame, int dev_format, uint32_t hsize, uint32_t vsize,
                                  int out_format, void* out_rgb, void* stream) {
    if (!ctx || !d_frame || !out_rgb || hsize == 0 || vsize == 0) return fail(RTC_ERR_INVALID_ARG, "rtc_download_frame: null argument");
    if (dev_format < RTC_OUT_F64_RGB || dev_format > RTC_OUT_U8_RGB || out_format < RTC_OUT_F64_RGB || out_format > RTC_OUT_U8_RGB)
        return fail(RTC_ERR_INVALID_ARG, "rtc_download_frame: bad format");
    if (dev_format != out_format && !(dev_format == RTC_OUT_F32_RGB && out_format == RTC_OUT_F64_RGB))
        return fail(RTC_ERR_UNSUPPORTED, "rtc_download_frame: only equal formats or F32 -> F64");
    DeviceState& d = ctx->devs[0];
    CUDA_TRY(cudaSetDevice(d.dev));
    const size_t dev_elem = elem_size(dev_format), out_elem = elem_size(out_format);
    const size_t row_elems = (size_t)hsize * 3, row_bytes = row_elems * dev_elem, frame_bytes = row_bytes * vsize;
    int rc = ensure_pinned(&ctx->pinned_rgb, &ctx->pinned_rgb_cap, frame_bytes);
    if (rc != RTC_OK) return rc;
    cudaStream_t src_stream = stream ? (cudaStream_t)stream : d.stream;
    ensure_pool(ctx, 1);
    Burst burst(frame_bytes >= (4u << 20) ? ctx->pool : nullptr);
    const int B = frame_bytes >= (32u << 20) ? 8 : (frame_bytes >= (4u << 20) ? 4 : 1);
    const int band_rows = ((int)vsize + B - 1) / B;
    CUDA_TRY(cudaEventRecord(d.ev_band[0], src_stream));
    CUDA_TRY(cudaStreamWaitEvent(d.copy_stream, d.ev_band[0], 0));
    int n_bands = 0;
    for (int b = 0; b < B; ++b) {
        const int r0 = b * band_rows, r1 = std::min((int)vsize, r0 + band_rows);
        if (r0 >= r1) break;
        ++n_bands;
        CUDA_TRY(cudaMemcpyAsync(ctx->pinned_rgb + (size_t)r0 * row_bytes, (const unsigned char*)d_frame + (size_t)r0 * row_bytes,
                                 (size_t)(r1 - r0) * row_bytes, cudaMemcpyDeviceToHost, d.copy_stream));
        CUDA_TRY(cudaEventRecord(d.ev_copy[b], d.copy_stream));
    }
    for (int b = 0; b < n_bands; ++b) {
        const int r0 = b * band_rows, r1 = std::min((int)vsize, r0 + band_rows);
        CUDA_TRY(cudaEventSynchronize(d.ev_copy[b]));
        host_deliver(ctx, (unsigned char*)out_rgb + (size_t)r0 * row_elems * out_elem, ctx->pinned_rgb + (size_t)r0 * row_bytes,
                     (size_t)(r1 - r0) * row_elems, dev_elem, out_elem);
    }
    return RTC_OK;
}

namespace {

bool is_registered(const rtc_context* ctx, const void* p, size_t bytes) {
    const unsigned char* q = (const unsigned char*)p;
    for (const auto& r : ctx->registered)
        if (q >= r.first && q + bytes <= r.first + r.second) return true;
    return false;
}

// One shard of the frame (interleaved row tiles: tile k -> shard k % shard_count) on GPU `d`, delivered into the caller's
// FULL host frame `user` at its frame rows — the unit of work of one rank process (one process per GPU, frame in shared
// host memory) and of one GPU of a several-GPU context.  The shard is rendered into a compact device buffer by one launch.
//   direct : the kernel writes out_format; pitched 2-D DMA copies place the tiles at their frame rows in the (registered)
//            caller buffer — no staging, no host pass; the last, shorter tile of the frame goes by a 1-D copy
//   staged : D2H of the compact buffer into pinned staging (at row `stage_row0`) in up to four chunks of whole tiles; the
//            tiles of chunk c are placed (and, f32 -> f64, widened) by the host while chunk c+1 crosses PCIe — by `pool`,
//            or by the calling thread alone when pool is null (the GPUs of a several-GPU context deliver side by side)
int render_shard(rtc_context* ctx, DeviceState& d, const rtc_camera* cam, const rtc_render_opts* opts, int dev_fmt, bool direct,
                 int shard_index, int shard_count, int tile_rows, unsigned char* user, uint8_t* out_hit_mask, size_t stage_row0,
                 HostPool* pool, bool want_counters, Counters* pinned_counter, rtc_stats& st) {
    auto t0 = std::chrono::steady_clock::now();
    CUDA_TRY(cudaSetDevice(d.dev));
    rtc_render_opts dopts = *opts;
    dopts.out_format = dev_fmt;
    const size_t dev_elem = elem_size(dev_fmt), out_elem = elem_size(opts->out_format);
    const size_t row_elems = (size_t)cam->hsize * 3, row_bytes = row_elems * dev_elem, hsize = cam->hsize;
    CamParams cp = make_cam(cam, &dopts, shard_index, shard_count, tile_rows);
    const int shard_rows = cp.shard_rows;
    if (shard_rows <= 0) return RTC_OK;
    const size_t shard_bytes = (size_t)shard_rows * row_bytes;
    CUDA_TRY(d.rgb.ensure(shard_bytes));
    if (out_hit_mask) CUDA_TRY(d.hit.ensure((size_t)shard_rows * hsize));
    if (want_counters) CUDA_TRY(cudaMemsetAsync(d.counters.p, 0, sizeof(Counters), d.stream));
    CUDA_TRY(cudaEventRecord(d.ev_k0, d.stream));
    const int n_tiles = (shard_rows + tile_rows - 1) / tile_rows;
    double deliver_ms = 0.0;
    unsigned char* stage_hit = ctx->pinned_hit ? ctx->pinned_hit + stage_row0 * hsize : nullptr;
    int rc = RTC_OK;
    int n_launches = 1;
    if (direct) {
        // Row bands of whole tiles, one launch each: band b's tiles cross PCIe (pitched 2-D DMA: compact rows -> their frame
        // rows in the registered caller buffer) while band b+1 renders.  Small-scene frames only: every launch of a big
        // scene ends in a long tail of deep ray trees.
        const int n_full = shard_rows / tile_rows;                       // only the frame's last tile can be short
        const size_t tile_bytes = (size_t)tile_rows * row_bytes;
        const bool small_kernel = opts->precision == RTC_PRECISION_F32_FAST && d.small_ok && !(opts->flags & RTC_OPT_F32_GENERAL_KERNEL);
        const int want_bands = d.s64.n_shapes > 256 ? 1 : (shard_bytes >= (32u << 20) ? 4 : (shard_bytes >= (8u << 20) ? 2 : 1));
        const int n_bands = std::max(1, std::min(want_bands, small_kernel || opts->precision == RTC_PRECISION_F64_PARITY ? n_tiles : 1));
        n_launches = 0;
        for (int b = 0; b < n_bands; ++b) {
            const int t0b = (int)((long long)n_tiles * b / n_bands), t1b = (int)((long long)n_tiles * (b + 1) / n_bands);
            if (t1b <= t0b) continue;
            const int r0 = t0b * tile_rows, r1 = std::min(shard_rows, t1b * tile_rows);
            CamParams cb = cp;
            cb.row0 = r0; cb.shard_rows = r1 - r0;
            // buffers are passed at their base: the kernels address compact row row0 + r (shard_count > 1) themselves; a
            // one-shard "shard" (shard_count == 1 never comes here) is not banded this way
            rc = launch_on(d, cb, opts->precision, opts->flags, d.rgb.p, out_hit_mask ? d.hit.p : nullptr, want_counters ? d.counters.p : nullptr, d.stream);
            if (rc != RTC_OK) return rc;
            ++n_launches;
            CUDA_TRY(cudaEventRecord(d.ev_band[b], d.stream));
            CUDA_TRY(cudaStreamWaitEvent(d.copy_stream, d.ev_band[b], 0));
            const int full_hi = std::min(t1b, n_full);
            if (full_hi > t0b)
                CUDA_TRY(cudaMemcpy2DAsync(user + ((size_t)t0b * shard_count + shard_index) * tile_bytes, tile_bytes * (size_t)shard_count,
                                           d.rgb.p + (size_t)t0b * tile_bytes, tile_bytes, tile_bytes, (size_t)(full_hi - t0b), cudaMemcpyDeviceToHost, d.copy_stream));
            if (t1b > n_full) {   // the frame's last, shorter tile
                const int rs = n_full * tile_rows;
                const size_t fr = (size_t)shard_row_to_frame_row(rs, tile_rows, shard_index, shard_count);
                CUDA_TRY(cudaMemcpyAsync(user + fr * row_bytes, d.rgb.p + (size_t)rs * row_bytes, (size_t)(shard_rows - rs) * row_bytes, cudaMemcpyDeviceToHost, d.copy_stream));
            }
        }
        CUDA_TRY(cudaEventRecord(d.ev_k1, d.stream));
        if (want_counters) CUDA_TRY(cudaMemcpyAsync(pinned_counter, d.counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, d.stream));
        CUDA_TRY(cudaStreamWaitEvent(d.copy_stream, d.ev_k1, 0));
        if (out_hit_mask) CUDA_TRY(cudaMemcpyAsync(stage_hit, d.hit.p, (size_t)shard_rows * hsize, cudaMemcpyDeviceToHost, d.copy_stream));
        CUDA_TRY(cudaStreamSynchronize(d.copy_stream));
        if (out_hit_mask)
            for (int t = 0; t < n_tiles; ++t) {
                const int r0 = t * tile_rows, r1 = std::min(shard_rows, r0 + tile_rows);
                std::memcpy(out_hit_mask + (size_t)shard_row_to_frame_row(r0, tile_rows, shard_index, shard_count) * hsize, stage_hit + (size_t)r0 * hsize, (size_t)(r1 - r0) * hsize);
            }
    } else {
        rc = launch_on(d, cp, opts->precision, opts->flags, d.rgb.p, out_hit_mask ? d.hit.p : nullptr, want_counters ? d.counters.p : nullptr, d.stream);
        if (rc != RTC_OK) return rc;
        CUDA_TRY(cudaEventRecord(d.ev_k1, d.stream));
        if (want_counters) CUDA_TRY(cudaMemcpyAsync(pinned_counter, d.counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, d.stream));
        CUDA_TRY(cudaStreamWaitEvent(d.copy_stream, d.ev_k1, 0));
        unsigned char* staged = ctx->pinned_rgb + stage_row0 * row_bytes;
        const int n_chunks = std::max(1, std::min(n_tiles, shard_bytes >= (4u << 20) ? 4 : 1));
        int chunk_t0[5];
        for (int c = 0; c <= n_chunks; ++c) chunk_t0[c] = (int)((long long)n_tiles * c / n_chunks);
        for (int c = 0; c < n_chunks; ++c) {
            const int r0 = chunk_t0[c] * tile_rows, r1 = std::min(shard_rows, chunk_t0[c + 1] * tile_rows);
            CUDA_TRY(cudaMemcpyAsync(staged + (size_t)r0 * row_bytes, d.rgb.p + (size_t)r0 * row_bytes, (size_t)(r1 - r0) * row_bytes, cudaMemcpyDeviceToHost, d.copy_stream));
            if (out_hit_mask)
                CUDA_TRY(cudaMemcpyAsync(stage_hit + (size_t)r0 * hsize, d.hit.p + (size_t)r0 * hsize, (size_t)(r1 - r0) * hsize, cudaMemcpyDeviceToHost, d.copy_stream));
            CUDA_TRY(cudaEventRecord(d.ev_copy[c], d.copy_stream));
        }
        const bool widen = dev_elem == 4 && out_elem == 8;
        for (int c = 0; c < n_chunks; ++c) {
            CUDA_TRY(cudaEventSynchronize(d.ev_copy[c]));
            auto h0 = std::chrono::steady_clock::now();
            const int t0c = chunk_t0[c];
            auto task = [=](int i) {   // one tile: compact rows [r0, r1) -> their frame rows (contiguous in both)
                const int r0 = (t0c + i) * tile_rows, r1 = std::min(shard_rows, r0 + tile_rows);
                const size_t fr = (size_t)shard_row_to_frame_row(r0, tile_rows, shard_index, shard_count);
                const size_t n = (size_t)(r1 - r0) * row_elems;
                if (widen) rtc_host_widen_f32_f64(reinterpret_cast<double*>(user) + fr * row_elems, reinterpret_cast<const float*>(staged) + (size_t)r0 * row_elems, n);
                else std::memcpy(user + fr * row_elems * out_elem, staged + (size_t)r0 * row_bytes, n * out_elem);
                if (out_hit_mask) std::memcpy(out_hit_mask + fr * hsize, stage_hit + (size_t)r0 * hsize, (size_t)(r1 - r0) * hsize);
            };
            const int n_c = chunk_t0[c + 1] - t0c;
            if (pool && n_c > 1) pool->parallel_for(n_c, task);
            else for (int i = 0; i < n_c; ++i) task(i);
            deliver_ms += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - h0).count();
        }
    }
    CUDA_TRY(cudaStreamSynchronize(d.stream));
    float k_ms = 0.f;
    CUDA_TRY(cudaEventElapsedTime(&k_ms, d.ev_k0, d.ev_k1));
    if (want_counters) { st.rays = pinned_counter->rays; st.primary_hits = pinned_counter->primary_hits; st.shaded_hits = pinned_counter->shaded_hits; st.shadow_rays = pinned_counter->shadow_rays; }
    st.kernel_launches = (uint64_t)n_launches;
    st.kernel_ms = k_ms;
    st.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    st.d2h_ms = st.total_ms - k_ms;
    st.gather_ms = deliver_ms;
    return RTC_OK;
}

}  // namespace

extern "C" int rtc_render(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts, void* out_rgb,
                          uint8_t* out_hit_mask, rtc_stats* stats) {
    if (!ctx || !out_rgb) return fail(RTC_ERR_INVALID_ARG, "rtc_render: null argument");
    int rc = validate_opts(cam, opts);
    if (rc != RTC_OK) return rc;
    if (!ctx->has_scene) return fail(RTC_ERR_NO_SCENE, "rtc_render before rtc_scene_upload");
    auto t0 = std::chrono::steady_clock::now();
    const int G = (int)ctx->devs.size();
    const size_t out_elem = elem_size(opts->out_format);
    const size_t row_elems = (size_t)cam->hsize * 3;
    const size_t mask_bytes = (size_t)cam->hsize * cam->vsize;
    const bool sharded = G > 1 || opts->shard_count > 1;
    // Delivery.  A registered caller buffer (rtc_register_host_buffer) takes finished rows by DMA in out_format: no staging,
    // no host pass.  The one case where staging wins: a whole fast-mode frame on ONE GPU into an F64_RGB Canvas crosses
    // PCIe as f32 (99.5 MB instead of 199 MB at 4K over the one link) and is widened by the host threads (exact).
    const bool registered = is_registered(ctx, out_rgb, row_elems * cam->vsize * out_elem);
    if ((opts->flags & RTC_OPT_DELIVER_DIRECT) && !registered)
        return fail(RTC_ERR_INVALID_ARG, "rtc_render: RTC_OPT_DELIVER_DIRECT needs a buffer registered with rtc_register_host_buffer");
    const bool f32_wins = !sharded && opts->precision == RTC_PRECISION_F32_FAST && opts->out_format == RTC_OUT_F64_RGB;
    const bool direct = registered && !(opts->flags & RTC_OPT_DELIVER_STAGED) && ((opts->flags & RTC_OPT_DELIVER_DIRECT) || !f32_wins);
    // Staged fast mode computes in f32: an F64_RGB (Canvas.pixels) request is produced as f32 on the device,
    // crosses PCIe as f32 and is widened on the host (exact), halving the bytes of the dominant copy.
    const int dev_fmt = (!direct && opts->precision == RTC_PRECISION_F32_FAST && opts->out_format == RTC_OUT_F64_RGB) ? RTC_OUT_F32_RGB : opts->out_format;
    const size_t dev_elem = elem_size(dev_fmt);
    const size_t row_bytes = row_elems * dev_elem;            // device / staging row
    const size_t frame_bytes = row_bytes * cam->vsize;
    if (!direct) {
        rc = ensure_pinned(&ctx->pinned_rgb, &ctx->pinned_rgb_cap, frame_bytes);
        if (rc != RTC_OK) return rc;
    }
    if (out_hit_mask) {
        rc = ensure_pinned(&ctx->pinned_hit, &ctx->pinned_hit_cap, mask_bytes);
        if (rc != RTC_OK) return rc;
    }
    rtc_render_opts dopts = *opts;
    dopts.out_format = dev_fmt;
    rtc_stats st;
    std::memset(&st, 0, sizeof st);
    unsigned char* user = (unsigned char*)out_rgb;
    ensure_pool(ctx, G == 1 && opts->shard_count > 1 ? opts->shard_count : 1);
    Burst burst(frame_bytes >= (4u << 20) ? ctx->pool : nullptr);   // small frames are delivered by the calling thread

    if (G == 1 && opts->shard_count > 1) {
        // One shard of the frame, one process per GPU (torchrun style): out_rgb is the FULL frame — typically memory the
        // rank processes share — and this call fills in the rows of its shard only.  Every GPU delivers its rows over its
        // own PCIe link (and, staged, by its own host threads): no gather step, no collective.
        const int tile_rows = opts->tile_rows > 0 ? opts->tile_rows : 16;
        rc = render_shard(ctx, ctx->devs[0], cam, opts, dev_fmt, direct, opts->shard_index, opts->shard_count, tile_rows, user, out_hit_mask, 0,
                          ctx->pool, stats != nullptr, &ctx->pinned_counters[0], st);
        if (rc != RTC_OK) return rc;
        st.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        ctx->last_d2h_bytes = (uint64_t)shard_num_rows((int)cam->vsize, tile_rows, opts->shard_index, opts->shard_count) * row_bytes;
        ctx->last_bands = (int)st.kernel_launches; ctx->last_bands_direct = direct ? ctx->last_bands : 0;
        if (stats) *stats = st;
        return RTC_OK;
    }

    if (G == 1) {
        // One GPU: the frame is rendered as up to 8 contiguous row bands (one launch each, rows
        // [row0, row0 + rows) written at their frame positions).  Band b's D2H runs on the copy stream
        // while band b+1 renders, and (staged) the host delivers band b while band b+1 crosses PCIe.
        //  * fast mode is copy- and delivery-bound (99.5 MB at ~56 GB/s and a ~1.8 ms host widening against a
        //    1.2 ms kernel): eight equal bands keep all three stages busy (measured 2.5 ms; small-large-small
        //    bands 3.2 ms, because a large band is delivered only after its whole copy); into a registered buffer:
        //    nine bands, the first and the last half as high, and the mixed delivery;
        //  * parity mode is kernel-bound: few large bands first (every launch has its own tail), and what is
        //    left after the last kernel is the copy and delivery of a sixteenth of the frame.
        DeviceState& d = ctx->devs[0];
        CUDA_TRY(cudaSetDevice(d.dev));
        static const int kPlanEq8[] = {1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1, 1}, kPlanF64[] = {6, 5, 3, 1, 1}, kPlanMid[] = {1, 2, 1}, kPlanOne[] = {1},
                         kPlanMix9[] = {1, 2, 2, 2, 2, 2, 2, 2, 1};
        const int* plan = kPlanOne; int n_plan = 1, plan_sum = 1;
        if (frame_bytes >= (32u << 20)) {
            // kernel-bound frames: parity mode, and fast mode outside the small-scene kernel (every launch has its own
            // tail of deep ray trees: eight bands of the 10k-shape frame took 173 ms against 117 ms for one launch)
            const bool kernel_bound = opts->precision == RTC_PRECISION_F64_PARITY || !d.small_ok || (opts->flags & RTC_OPT_F32_GENERAL_KERNEL);
            if (d.s64.n_shapes > 256) { plan = kPlanOne; n_plan = 1; plan_sum = 1; }   // tens of ms of kernel, ms-long tails: one launch, then the copy
            else if (kernel_bound) { plan = kPlanF64; n_plan = 5; plan_sum = 16; }
            else if (registered && f32_wins && !(opts->flags & (RTC_OPT_DELIVER_DIRECT | RTC_OPT_DELIVER_STAGED))) {
                // mixed delivery (below): a short first band starts the copies after a sixteenth of the kernel work instead of an
                // eighth, a short last band is what goes by DMA first; 2.44 -> 2.34 ms against eight equal bands
                // (profiles/r02_e2e_mix_probe.txt: seven plans x four splits)
                plan = kPlanMix9; n_plan = 9; plan_sum = 16;
            }
            else { plan = kPlanEq8; n_plan = 8; plan_sum = 8; }
#ifdef RTC_PROBE_KNOBS   // `make probe`: measurement build only (tools/e2e_mix_probe.py); the product library reads no environment
            if (plan == kPlanEq8 && getenv("RTC_MIX_BANDS")) { n_plan = plan_sum = std::max(1, std::min(kMaxBands, atoi(getenv("RTC_MIX_BANDS")))); }
            static int probe_plan[kMaxBands];
            if (plan == kPlanEq8 && getenv("RTC_MIX_PLAN")) {   // e.g. "1,2,2,2,2,2,2,2,1": band heights in sixteenths or whatever they sum to
                n_plan = 0; plan_sum = 0;
                for (const char* c = getenv("RTC_MIX_PLAN"); *c && n_plan < kMaxBands;) {
                    probe_plan[n_plan] = std::max(1, atoi(c)); plan_sum += probe_plan[n_plan++];
                    while (*c && *c != ',') ++c;
                    if (*c == ',') ++c;
                }
                plan = probe_plan;
            }
#endif
        } else if (frame_bytes >= (4u << 20)) { plan = kPlanMid; n_plan = 3; plan_sum = 4; }
        const bool mix_plan = plan != kPlanOne && plan != kPlanMid && plan != kPlanF64;   // the copy- and delivery-bound plan of a big fast-mode frame
        int band_r0[kMaxBands + 1];
        int n_bands = 0;
        {
            int acc = 0, prev = 0;
            for (int b = 0; b < n_plan; ++b) {
                acc += plan[b];
                int r = b + 1 == n_plan ? (int)cam->vsize : (int)((long long)cam->vsize * acc / plan_sum) / 4 * 4;   // whole 8x4 tiles
                if (r > prev) { band_r0[n_bands++] = prev; prev = r; }
            }
            band_r0[n_bands] = (int)cam->vsize;
        }
        // Mixed delivery (a fast-mode F64_RGB frame into a REGISTERED buffer): the first bands cross PCIe as f32 and are
        // widened by the host threads, the last k_wide bands are written as f64 by the kernel and go by DMA straight into
        // the caller's frame — the link and the host threads work side by side instead of the slower of the two setting
        // the frame time.  f32 -> f64 is exact on either side, so the frame does not depend on k_wide.  Chosen per frame
        // from how the previous frame ended (host threads still widening after the last byte had landed: one band more
        // goes by DMA; DMA still running after the host had finished: one band fewer), or half the bands when both
        // delivery flags are set (tests).
        const bool mix_forced = registered && f32_wins && (opts->flags & RTC_OPT_DELIVER_DIRECT) && (opts->flags & RTC_OPT_DELIVER_STAGED);
        const bool mix_auto = registered && f32_wins && !direct && !(opts->flags & (RTC_OPT_DELIVER_DIRECT | RTC_OPT_DELIVER_STAGED)) && mix_plan;
        int k_wide = mix_forced ? n_bands / 2 : (mix_auto ? std::min(ctx->mix_direct_bands, n_bands / 2) : 0);
#ifdef RTC_PROBE_KNOBS
        if (mix_auto && getenv("RTC_MIX_K")) k_wide = std::max(0, std::min(n_bands - 1, atoi(getenv("RTC_MIX_K"))));
#endif
        const size_t wide_row_bytes = row_elems * 8;
        CUDA_TRY(d.rgb.ensure(frame_bytes));
        if (k_wide) CUDA_TRY(d.rgb_wide.ensure(wide_row_bytes * cam->vsize));
        if (out_hit_mask) CUDA_TRY(d.hit.ensure(mask_bytes));
        CUDA_TRY(cudaMemsetAsync(d.counters.p, 0, sizeof(Counters), d.stream));
        CUDA_TRY(cudaEventRecord(d.ev_k0, d.stream));
        for (int b = 0; b < n_bands; ++b) {
            const int r0 = band_r0[b], r1 = band_r0[b + 1];
            const bool wide = b >= n_bands - k_wide;
            rtc_render_opts bopts = dopts;
            if (wide) bopts.out_format = RTC_OUT_F64_RGB;
            CamParams cp = make_cam(cam, &bopts, 0, 1, 16);
            cp.dst_full = 1; cp.row0 = r0; cp.shard_rows = r1 - r0;
            // work counters only when the caller asked for stats: the plain kernel instantiation carries no book-keeping
            rc = launch_on(d, cp, opts->precision, opts->flags, wide ? d.rgb_wide.p : d.rgb.p, out_hit_mask ? d.hit.p : nullptr, stats ? d.counters.p : nullptr, d.stream);
            if (rc != RTC_OK) return rc;
            CUDA_TRY(cudaEventRecord(d.ev_band[b], d.stream));
            CUDA_TRY(cudaStreamWaitEvent(d.copy_stream, d.ev_band[b], 0));
            // direct (and the wide bands of a mixed delivery): the band goes straight to its rows of the caller's (registered) frame
            if (wide)
                CUDA_TRY(cudaMemcpyAsync(user + (size_t)r0 * wide_row_bytes, d.rgb_wide.p + (size_t)r0 * wide_row_bytes, (size_t)(r1 - r0) * wide_row_bytes,
                                         cudaMemcpyDeviceToHost, d.copy_stream));
            else
            CUDA_TRY(cudaMemcpyAsync((direct ? user : ctx->pinned_rgb) + (size_t)r0 * row_bytes, d.rgb.p + (size_t)r0 * row_bytes, (size_t)(r1 - r0) * row_bytes,
                                     cudaMemcpyDeviceToHost, d.copy_stream));
            if (out_hit_mask)
                CUDA_TRY(cudaMemcpyAsync(ctx->pinned_hit + (size_t)r0 * cam->hsize, d.hit.p + (size_t)r0 * cam->hsize, (size_t)(r1 - r0) * cam->hsize,
                                         cudaMemcpyDeviceToHost, d.copy_stream));
            CUDA_TRY(cudaEventRecord(d.ev_copy[b], d.copy_stream));
            if (mix_auto && b == 0) CUDA_TRY(cudaEventRecord(d.ev_c0, d.copy_stream));
        }
        if (mix_auto) CUDA_TRY(cudaEventRecord(d.ev_c1, d.copy_stream));
        CUDA_TRY(cudaEventRecord(d.ev_k1, d.stream));
        CUDA_TRY(cudaMemcpyAsync(&ctx->pinned_counters[0], d.counters.p, sizeof(Counters), cudaMemcpyDeviceToHost, d.stream));
        double deliver_ms = 0.0;
        auto first_landed = std::chrono::steady_clock::now(), host_done = first_landed;
        for (int b = 0; b < n_bands; ++b) {
            const int r0 = band_r0[b], r1 = band_r0[b + 1];
            const bool wide = b >= n_bands - k_wide;
            CUDA_TRY(cudaEventSynchronize(d.ev_copy[b]));
            auto h0 = std::chrono::steady_clock::now();
            if (b == 0) first_landed = h0;
            if (!direct && !wide) host_deliver(ctx, user + (size_t)r0 * row_elems * out_elem, ctx->pinned_rgb + (size_t)r0 * row_bytes, (size_t)(r1 - r0) * row_elems, dev_elem, out_elem);
            if (out_hit_mask) std::memcpy(out_hit_mask + (size_t)r0 * cam->hsize, ctx->pinned_hit + (size_t)r0 * cam->hsize, (size_t)(r1 - r0) * cam->hsize);
            auto h1 = std::chrono::steady_clock::now();
            if (!wide) host_done = h1;
            deliver_ms += std::chrono::duration<double, std::milli>(h1 - h0).count();
        }
        if (mix_auto) {
            // how long after the frame's last byte had landed the host threads finished their last band (negative: before)
            float dma_ms = 0.f;
            CUDA_TRY(cudaEventSynchronize(d.ev_c1));
            CUDA_TRY(cudaEventElapsedTime(&dma_ms, d.ev_c0, d.ev_c1));
            const double lag_ms = std::chrono::duration<double, std::milli>(host_done - first_landed).count() - (double)dma_ms;
            const double band_ms = (double)dma_ms / 8.0;   // about an eighth of the frame over the link
#ifdef RTC_PROBE_KNOBS
            if (getenv("RTC_MIX_TRACE"))
                fprintf(stderr, "mix: bands %d wide %d  first copy -> last copy %.3f ms  first landed -> host done %.3f ms  lag %.3f  deliver %.3f\n", n_bands, k_wide,
                        dma_ms, std::chrono::duration<double, std::milli>(host_done - first_landed).count(), lag_ms, deliver_ms);
#endif
            // single frames are noisy (0.2 ... 0.8 ms at the same split): follow a running mean, and move one band at a time
            ctx->mix_lag_ms = 0.7 * ctx->mix_lag_ms + 0.3 * lag_ms;
            if (ctx->mix_lag_ms > 2.5 * band_ms && ctx->mix_direct_bands < n_bands / 2) { ++ctx->mix_direct_bands; ctx->mix_lag_ms = 0.0; }
            else if (ctx->mix_lag_ms < -1.5 * band_ms && ctx->mix_direct_bands > 0) { --ctx->mix_direct_bands; ctx->mix_lag_ms = 0.0; }
        }
        CUDA_TRY(cudaStreamSynchronize(d.stream));
        float k_ms = 0.f;
        CUDA_TRY(cudaEventElapsedTime(&k_ms, d.ev_k0, d.ev_k1));
        const Counters& c = ctx->pinned_counters[0];
        st.rays = c.rays; st.primary_hits = c.primary_hits; st.shaded_hits = c.shaded_hits; st.shadow_rays = c.shadow_rays;
        st.kernel_launches = n_bands;
        st.kernel_ms = k_ms;   // all bands (later bands share the GPU with the D2H of earlier ones)
        st.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
        st.d2h_ms = st.total_ms - k_ms;   // whatever was not hidden behind the kernel: PCIe + host delivery
        st.gather_ms = deliver_ms;
        ctx->last_bands = n_bands; ctx->last_bands_direct = direct ? n_bands : k_wide;
        ctx->last_d2h_bytes = (uint64_t)band_r0[n_bands - k_wide] * row_bytes + (uint64_t)((int)cam->vsize - band_r0[n_bands - k_wide]) * (k_wide ? wide_row_bytes : row_bytes);
        if (stats) *stats = st;
        return RTC_OK;
    }

    // Several GPUs in ONE process (what a Rust camera.render(&world) on a multi-GPU box gets): interleaved row tiles, tile
    // k -> GPU k % G.  One host thread per GPU runs that GPU's shard — launch, D2H over its own PCIe link, delivery into
    // the caller's frame — side by side with the others: launches are concurrent, nothing waits for the slowest GPU
    // before its own rows are delivered, and there is no gather (the shards' rows are disjoint).
    const int tile_rows = opts->tile_rows > 0 ? opts->tile_rows : 16;
    std::vector<size_t> stage_row0(G, 0);
    for (int g = 1; g < G; ++g) stage_row0[g] = stage_row0[g - 1] + (size_t)shard_num_rows((int)cam->vsize, tile_rows, g - 1, G);
    std::vector<rtc_stats> gst(G);
    std::vector<int> grc(G, RTC_OK);
    std::vector<std::string> gerr(G);
    for (auto& x : gst) std::memset(&x, 0, sizeof x);
    auto job = [&](int g) {
        grc[g] = render_shard(ctx, ctx->devs[g], cam, opts, dev_fmt, direct, g, G, tile_rows, user, out_hit_mask, stage_row0[g], nullptr,
                              stats != nullptr, &ctx->pinned_counters[g], gst[g]);
        if (grc[g] != RTC_OK) gerr[g] = g_err;   // g_err is thread_local: carry the worker's message to the caller
    };
    ctx->pool->parallel_for(G, job);
    for (int g = 0; g < G; ++g)
        if (grc[g] != RTC_OK) return fail(grc[g], "GPU " + std::to_string(ctx->devs[g].dev) + ": " + gerr[g]);
    for (int g = 0; g < G; ++g) {
        st.kernel_ms = std::max(st.kernel_ms, gst[g].kernel_ms);
        st.d2h_ms = std::max(st.d2h_ms, gst[g].d2h_ms);
        st.gather_ms = std::max(st.gather_ms, gst[g].gather_ms);
        st.rays += gst[g].rays; st.primary_hits += gst[g].primary_hits; st.shaded_hits += gst[g].shaded_hits; st.shadow_rays += gst[g].shadow_rays;
        st.kernel_launches += gst[g].kernel_launches;
    }
    st.total_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count();
    ctx->last_d2h_bytes = (uint64_t)cam->vsize * row_bytes;
    ctx->last_bands = (int)st.kernel_launches; ctx->last_bands_direct = direct ? ctx->last_bands : 0;
    if (stats) *stats = st;
    return RTC_OK;
}

extern "C" int rtc_delivery_info(rtc_context* ctx, uint64_t* d2h_bytes, int* bands, int* bands_direct) {
    if (!ctx) return fail(RTC_ERR_INVALID_ARG, "rtc_delivery_info: null context");
    if (d2h_bytes) *d2h_bytes = ctx->last_d2h_bytes;
    if (bands) *bands = ctx->last_bands;
    if (bands_direct) *bands_direct = ctx->last_bands_direct;
    return RTC_OK;
}
