// Shared between the C-ABI translation unit (rtc_api.cu) and the kernel translation units
// (render_f64.cu: parity mode, built with -fmad=false; render_f32.cu: fast mode).
//
// Device scene layout ("scene flattening pass", north-star item 1).  World.shapes arrives as an
// AoS of rtc_shape (include/rtc.h); rtc_scene_upload splits it into
//   * INTERSECT STREAMS (structure of arrays, one entry per sphere / per plane, in add_shape
//     order): exactly what the brute-force World::intersect loop (src/shape.rs:677-683) reads for
//     every ray.  All lanes of a warp walk these arrays in lock step, so they are staged into
//     shared memory per block with 16-byte vector loads and read back as broadcasts.
//       spheres: 3 rows x (m0 m1 m2 m3) of the inverse transform  -> 6 x double2 / 3 x float4
//       planes : row 1 of the inverse only (the plane test uses just o'.y and d'.y,
//                src/shape.rs:462-471)                             -> 2 x double2 / 1 x float4
//   * SHADING RECORDS (one 16-byte-aligned record per shape, original order): inverse,
//     inverse-transpose 3x3, material scalars, pattern.  Gathered by individual lanes at hit time
//     (one record per hit), so they stay in global memory behind L1/L2.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include "../../include/rtc.h"

// `make checked` builds librtc_b200_checked.so with -DRTC_CHECKED: every index the kernels form into a stream, a record
// array, a per-lane stack or a frame is asserted in range (a failed device assert prints file:line and stops the kernel).
// tools/sanitize_run.py runs the whole entry-point matrix on it between guard bands — the stand-in for compute-sanitizer,
// which this GPU pool refuses to run (profiles/r02_sanitizer.txt).  The product build compiles none of it.
#ifdef RTC_CHECKED
#include <assert.h>
#define RTC_CHECK(c) assert(c)
#else
#define RTC_CHECK(c) ((void)0)
#endif

namespace rtc {

// ---- f64 (parity) layout ----------------------------------------------------------------------
struct __align__(16) ShadeRec64 {
    double inv[12];     // rows 0..2 of Shape::get_transform_inverse()            (src/shape.rs:30)
    double invt[9];     // upper-left 3x3 of get_transform_inverse_transpose()    (src/shape.rs:32)
    double color[3];    // Material.color                                         (src/material.rs:246)
    double ambient, diffuse, specular, shininess;
    double reflective, transparency, refractive_index;
    double pat_inv[12]; // Pattern::get_transform_inverse()                       (src/material.rs:37)
    double pat_a[3], pat_b[3];
    int32_t kind;         // RTC_SHAPE_*
    int32_t world_id;     // u8 in the reference (src/shape.rs:287)
    int32_t pattern_kind; // RTC_PATTERN_*
    int32_t pad;
};

struct DevScene64 {
    int n_shapes, n_spheres, n_planes, ids_unique;
    int n_cubes, pad0;         // cubes follow the spheres in sph_rows / sph_index (same 3x4 inverse rows, src/shape.rs:618-621)
    const double2* sph_rows;   // [n_spheres + n_cubes][6]: r0.xy r0.zw r1.xy r1.zw r2.xy r2.zw
    const int32_t* sph_index;  // [n_spheres + n_cubes] original shape index
    const double2* pl_rows;    // [n_planes][2]: r1.xy r1.zw
    const int32_t* pl_index;   // [n_planes]
    const ShadeRec64* rec;     // [n_shapes]
    double light_pos[3];       // light 0 (World.light, src/shape.rs:635)
    double light_int[3];
    int n_lights, pad1;        // rtc_scene_upload_lights: lights 1..n_lights-1 live in `lights`
    const double* lights;      // [n_lights][6]: position xyz, intensity rgb
    // f32 bounding spheres (cx, cy, cz, -R^2) of the spheres | cubes in sph_rows order, padded to a multiple of four with
    // (0, 0, 0, +inf): the parity kernel runs a shape's exact f64 test only when the ray's LINE passes within R plus a
    // rounding margin of the centre (nine FFMA on the otherwise idle FP32 pipe against ~54 unfused FP64 operations).  A
    // shape the filter skips is one whose exact test finds no root, so the list of intersections — and with it every
    // bit of the frame and every counter — is unchanged.  n_bound4 = 0: no filter (RTC_OPT_NO_BOUND_FILTER, or a shape
    // without a finite bound).
    const float4* bound;
    int n_bound4;              // entries of `bound` (a multiple of eight)
    float bound_cmax, bound_rmax;
    // one bounding sphere per eight consecutive entries of `bound` (big scenes are stored in Morton order, so eight
    // neighbours in the stream are neighbours in space): a ray whose line misses it skips the eight member tests
    const float4* gbound;
    float gbound_cmax, gbound_rmax;
    const float4* sbound;      // and one per 32 consecutive groups (a block of 256 entries), padded to a multiple of 32
    float sbound_cmax, sbound_rmax;
};

// ---- f32 (fast) layout ------------------------------------------------------------------------
// Fast mode re-derives each shape test in WORLD space so that almost every instruction is an FFMA:
//   planes          (nx, ny, nz, w) = row 1 of the inverse: o'.y = n.o + w, d'.y = n.d
//   uniform spheres (cx, cy, cz, r^2): spheres whose transform is a similarity (uniform scale +
//                   rotation + translation) are ordinary world-space spheres
//   general spheres 3 rows of the inverse (any affine transform), tested in object space
//   cubes           3 rows of the inverse, slab test in object space (Cube::intersect_local, src/shape.rs:577-591)
// Geometry kinds of a shading record:
enum { GEOM_PLANE = 0, GEOM_USPHERE = 1, GEOM_GSPHERE = 2, GEOM_CUBE = 3 };

struct __align__(16) ShadeRec32 {
    float4 g0, g1, g2;  // plane: g0.xyz = unit world normal | usphere: g0 = (c, 1/r) | gsphere, cube: inverse rows
    float4 t0, t1, t2;  // gsphere: rows of the stored inverse-transpose 3x3 | cube: rows of transpose(inverse 3x3) (src/shape.rs:623-629)
    float4 col_amb;     // colour rgb, ambient
    float4 dif_spec;    // diffuse, specular, shininess, reflective
    float4 tr_ri;       // transparency, refractive_index, (int) pattern_kind, (int) geom kind
    float4 p0, p1, p2;  // world -> pattern space (pattern_inv * shape_inv), rows
    float4 pa, pb;      // pattern colours a, b ; pa.w = (int) world_id
};

struct DevScene32 {
    int n_shapes, n_planes, n_usph, n_gsph, ids_unique;
    int n_cube;                // cubes follow the general spheres in gsph / gs_index (object-space slab test)
    const float4* planes;      // [n_planes]
    const int32_t* pl_index;   // original shape index
    const float4* usph;        // [n_usph]
    const int32_t* us_index;
    const float4* gsph;        // [n_gsph][3]
    const int32_t* gs_index;
    const ShadeRec32* rec;     // [n_shapes]
    const ShadeRec64* rec64;   // [n_shapes] the parity kernel's records: big scenes (kRefine kernels) take the geometry of a hit from them in f64
    const int32_t* gset;       // stream positions of the shapes with refractive_index != 1.0
    int n_gset;
    const float* plL;          // [n_planes] f(light) = n.light + w   (shadow test: f(0) f(1) <= 0)
    const float* usL;          // [n_usph]   |light - c|^2 - r^2
    int light_inside;          // bit k: light k sits inside some uniform sphere
    // Bounding spheres (cx, cy, cz, -R^2) of the uniform spheres | general spheres | cubes, in stream order,
    // padded to a multiple of four with (0, 0, 0, +inf).  Scenes whose streams stay in global memory (10^3+
    // shapes) test a ray against these first — 9 FFMA per shape — and run the exact test of a shape only
    // when the ray's line passes within R + rounding margin of its centre.  n_bound4 = 0: no filter.
    const float4* bound;
    int n_bound4;                   // entries of `bound` (a multiple of eight)
    float bound_cmax, bound_rmax;   // max |c|_1 and max R over the entries (rounding margin of the filter)
    const float4* gbound;           // one bounding sphere per eight consecutive entries of `bound` (Morton-ordered streams)
    float gbound_cmax, gbound_rmax;
    const float4* sbound;           // one per 32 consecutive groups (a block of 256 entries), padded to a multiple of 32
    float sbound_cmax, sbound_rmax;
    // The same two upper levels as axis-aligned boxes (lo, hi interleaved: two float4 per entry; padded to multiples of 32
    // entries, padding never read as a candidate).  Eight scattered small spheres fill well under 1 % of the sphere around
    // them and Morton runs are elongated, so boxes turn away several times more rays than the bounding spheres do; the
    // warp-cooperative queries of the fast kernel test boxes at the block and group level, spheres at the entries.
    const float4* gbox;             // [2 * groups]
    const float4* sbox;             // [2 * blocks]
    float light_pos[3];        // light 0
    float light_int[3];
    int n_lights;              // rtc_scene_upload_lights: plL / usL hold n_lights rows; lights 1.. read them from global memory
    const float4* lights;      // [n_lights][2]: position, intensity
};

// ---- f32 small-scene layout (render_f32_small.cu) ------------------------------------------------
// Scenes with <= 16 planes and <= 48 spheres, all of them similarity-transformed, travel to the kernel
// as a __grid_constant__ parameter: the intersect streams sit in constant bank 0 and every
// coefficient is a direct c[0x0][imm] operand of the FP32 instruction that uses it.
constexpr int kSmallMaxPlanes = 16, kSmallMaxUsph = 48, kSmallMaxShapes = 64, kSmallMaxGlass = 8;

struct __align__(16) SmallRec32 {   // shading record, indexed by stream position
    float4 nrm;       // world normal n = p * nrm.w + nrm.xyz   (plane: w = 0, xyz = n; sphere: w = 1/r, xyz = -c/r)
    float4 col_amb;   // colour rgb, ambient
    float4 dif_spec;  // diffuse, specular, shininess, reflective
    float4 tr_ri;     // transparency, refractive_index, (int) pattern_kind, unused
    float4 p0, p1, p2;  // world -> pattern space rows (pattern_inv * shape_inv)
    float4 pa, pb;    // pattern colours; pb.w = 1 for CHECKER, 0 for STRIPE
};

struct SmallParams {
    // Stream positions: plane i -> i, sphere i -> 16 + i.  Arrays that two different loops read are
    // stored twice so that each copy has a single use and stays a direct constant operand.
    float4 planes[kSmallMaxPlanes];    // (nx, ny, nz, w) = row 1 of the inverse; padding (0,0,0,1): t = -inf
    float4 planes_d[kSmallMaxPlanes];  // same rows again (the n.d use)
    float4 splanes[kSmallMaxPlanes];   // planes * f(light): shadow test g = splane . (o,1); padding (0,0,0,1)
    float4 usph[kSmallMaxUsph];        // (cx, cy, cz, r^2); padding (0,0,0,-1): never hit
    float4 usph_s[kSmallMaxUsph];      // same again (the shadow loop's copy)
    float4 glass[kSmallMaxGlass];      // the spheres with refractive_index != 1
    int npg, nug;                      // FULL plane / sphere groups of two: the streams' positions [0, 2 npg) and [0, 2 nug)
    int np_odd, nu_odd;                // an odd count's last shape sits in the LAST slot of its array (position kSmallMaxPlanes - 1 /
                                       // kSmallMaxUsph - 1: a compile-time position like every other) and is tested on its own
    int np, nu;
    int n_glass;                       // -1: fast n1/n2 test unavailable (a glass plane, or more than 8 glass spheres)
    int n_extra;                       // antialiasing_samples when >= 2 (resample passes, src/camera.rs:84-92), else 0
    const SmallRec32* rec;             // [64] by stream position (global, staged to shared memory per block)
    unsigned* order;                   // [total_tiles] tile hand-out order (heavy tiles first), or null = linear
    float light_pos[3], light_int[3];
    float cam_o[3], cam_dx[3], cam_dy[3], cam_d0[3];   // ray dir = normalize(cam_dx (x+xo) + cam_dy (y+yo) + cam_d0)
    int hsize, shard_rows, xlim, ylim;
    int shard_index, shard_count, tile_rows, dst_full;
    int n_sub, max_depth, tiles_x, total_tiles;
    int row0;                          // first frame row of the launch (row bands of rtc_render)
    unsigned aa_seed;                  // seed of the resample offsets (include/rtc.h)
    // rtc_render_device_synced, fused: every block starts only once *wait_flag >= wait_value (the consumer is done with the
    // previous frame: the destination may be overwritten), and the last block to finish release-stores done_value to
    // *done_flag at system scope after every block's pixels are visible there (peer stores over NVLink included)
    const unsigned* wait_flag; unsigned* done_flag; unsigned* wait_status;
    unsigned wait_value, done_value;
    unsigned long long wait_timeout_ns;
};

// ---- per-launch parameters --------------------------------------------------------------------
struct CamParams {
    uint32_t hsize, vsize;
    double half_width, half_height, pixel_size;
    double vinv[12];  // rows 0..2 of view_transform_inv (src/camera.rs:25)
    uint32_t samples; // 1 = centre sample; else 2x2 fixed sub-samples + `samples` resamples where they disagree (src/camera.rs:84-114)
    uint32_t aa_seed; // seed of the resample offsets (include/rtc.h)
    uint32_t xlim, ylim;  // rect: pixels with x >= xlim or y >= ylim are written black
    int32_t max_depth;
    // row-tile sharding: tile k (rows [k*tile_rows, ...)) belongs to shard k % shard_count
    int32_t shard_index, shard_count, tile_rows;
    int32_t shard_rows;   // rows in this shard's compact buffer
    int32_t out_format;   // RTC_OUT_*
    int32_t dst_full;     // 1: the destination is a full frame (rows at their frame positions), not the compact shard buffer
    int32_t row0;         // row bands (one launch each): shard_count == 1: first FRAME row, the launch renders rows [row0, row0 + shard_rows);
                          // shard_count > 1: first COMPACT row of the shard, the launch renders compact rows [row0, row0 + shard_rows)
};

struct Counters {
    unsigned long long rays, primary_hits, shaded_hits, shadow_rays;
};

// Resample offsets (include/rtc.h, "Resample offsets"): 64 random bits for (pixel, extra sample k, axis).
__host__ __device__ inline uint64_t splitmix64(uint64_t x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
__host__ __device__ inline uint64_t resample_bits(uint64_t seed_mixed, uint64_t pixel_index, unsigned k, unsigned axis) {
    return splitmix64(seed_mixed ^ ((pixel_index << 9) | ((uint64_t)k << 1) | (uint64_t)axis));
}

// Map a compact shard row to the frame row (tiles stacked in frame order).
__host__ __device__ inline int shard_row_to_frame_row(int r, int tile_rows, int shard_index, int shard_count) {
    int t = r / tile_rows;                 // shard-local tile
    int in = r - t * tile_rows;
    return (t * shard_count + shard_index) * tile_rows + in;
}

// Rows owned by `shard_index` in a frame of vsize rows.
__host__ __device__ inline int shard_num_rows(int vsize, int tile_rows, int shard_index, int shard_count) {
    if (shard_count <= 1) return vsize;
    int n_tiles = (vsize + tile_rows - 1) / tile_rows;
    int rows = 0;
    for (int t = shard_index; t < n_tiles; t += shard_count) {
        int r0 = t * tile_rows;
        int r1 = r0 + tile_rows < vsize ? r0 + tile_rows : vsize;
        rows += r1 - r0;
    }
    return rows;
}

// Loads every kernel of a translation unit on the current device (cudaFuncGetAttributes).  With CUDA's lazy module loading
// the FIRST launch of a kernel may need a device-wide synchronisation; behind a kernel that spins on a completion flag
// (rtc_wait_device, the fused wait of rtc_render_device_synced) that launch would block until the wait times out.
// rtc_context_create calls these once per GPU, so no launch of the library ever loads code.
cudaError_t preload_render_f64();
cudaError_t preload_render_f32();
cudaError_t preload_render_f32_small();

// rtc_render_device_noalloc (include/rtc.h): the small-scene fast kernel with its scratch — the tile counter words and the
// tile hand-out order — in a static device array of render_f32_small.cu instead of cudaMalloc'd buffers.
constexpr int kNoAllocMaxTiles = 1 << 20;            // 8x4-pixel tiles: covers 7680x4320 (1,036,800 tiles); 4 MiB + 256 B per GPU
cudaError_t noalloc_setup_current_device();          // cudaGetSymbolAddress of the array for the CURRENT device (once per device)
cudaError_t launch_render_f32_small_noalloc(SmallParams P, bool tile_order, int out_format, void* d_rgb, uint8_t* d_hit, cudaStream_t stream);

// kernel launchers implemented in the kernel TUs
cudaError_t launch_render_f64(const DevScene64& scene, const CamParams& cam, void* d_rgb, uint8_t* d_hit,
                              Counters* d_counters, cudaStream_t stream);
cudaError_t launch_render_f32(const DevScene32& scene, const CamParams& cam, void* d_rgb, uint8_t* d_hit,
                              Counters* d_counters, unsigned* d_work, cudaStream_t stream);

// true: the scene's fast-mode streams exceed shared memory (the general f32 kernel would run its kRefine form: streams in
// global memory, geometry of the radiance rays in f64).  rtc_api.cu's dispatch sends such scenes to the parity kernel instead.
bool f32_streams_are_global(const DevScene32& scene);

cudaError_t launch_render_f32_small(const SmallParams& P, int out_format, void* d_rgb, uint8_t* d_hit, Counters* d_counters,
                                    unsigned* d_work, cudaStream_t stream);

}  // namespace rtc
