/*
 * rtc.h — C ABI of the B200-native `Camera::render` hot path.
 *
 * One shared library (librtc_b200.so), C linkage, plain pointers and sizes, no
 * exceptions across the boundary, `int` status + rtc_last_error().  The caller owns
 * every buffer; descriptors are copied during the call.  The reference
 * (joedane/raytracer-challenge, crate `ch1`) has no FFI/plugin interface: the
 * boundary replaced is the Rust method
 *     pub fn render(&self, w:&World) -> Canvas          src/camera.rs:116
 * and its siblings render_async (src/camera.rs:144) / render_async1 (:128).
 * Each entry point below cites the reference code whose work it takes over.
 * INTEGRATION.md shows the Rust `extern "C"` block + `flatten()` shim that binds them.
 *
 * All matrices are 4x4 row-major f64 (src/transform.rs:23-27); only the affine 3x4
 * part is used by the path (src/transform.rs:107-128).
 *
 * There is NO CPU fallback behind this ABI: without a CUDA device every compute
 * entry point returns RTC_ERR_NO_DEVICE.
 */
#ifndef RTC_B200_H
#define RTC_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define RTC_ABI_VERSION 1

/* status codes */
enum {
    RTC_OK = 0,
    RTC_ERR_INVALID_ARG = 1,  /* null pointer, bad enum, bad size */
    RTC_ERR_UNSUPPORTED = 2,  /* e.g. material with neither colour nor pattern (the reference
                                 panics: src/material.rs:331), samples > 255 (a u8 in the
                                 reference, src/camera.rs:26) */
    RTC_ERR_NO_DEVICE = 3,    /* no CUDA device / driver: the path never falls back to the CPU */
    RTC_ERR_CUDA = 4,         /* a CUDA runtime call failed; see rtc_last_error() */
    RTC_ERR_NO_SCENE = 5,     /* rtc_render* before rtc_scene_upload */
    RTC_ERR_NCCL = 6,
    RTC_ERR_TIMEOUT = 7       /* rtc_wait_result: a completion flag did not arrive within the context's wait timeout */
};

/* src/shape.rs:281 (Sphere), :405 (Plane), :500 (Cube) */
enum { RTC_SHAPE_SPHERE = 0, RTC_SHAPE_PLANE = 1, RTC_SHAPE_CUBE = 2 };

/* src/material.rs:48 Test, :72 Stripe, :106 Gradient, :138 Ring, :173 Checker, :208 Grid */
enum {
    RTC_PATTERN_NONE = 0,
    RTC_PATTERN_TEST = 1,
    RTC_PATTERN_STRIPE = 2,
    RTC_PATTERN_GRADIENT = 3,
    RTC_PATTERN_RING = 4,
    RTC_PATTERN_CHECKER = 5,
    RTC_PATTERN_GRID = 6
};

/* One `Box<dyn Shape>` of World.shapes (src/shape.rs:633-637), flattened:
 *   inv     = Shape::get_transform_inverse()            src/shape.rs:30
 *   inv_t   = Shape::get_transform_inverse_transpose()  src/shape.rs:32   (read verbatim; the
 *             stale value left by Plane/Cube::set_transform, :446-449, is kept as is)
 *   world_id= Shape::get_world_id()                     src/shape.rs:21   (u8, wraps at 256)
 *   material fields                                     src/material.rs:244-254
 *   pattern: kind + Pattern::get_transform_inverse() + its two colours  src/material.rs:33-242
 *            (GRID: pat_a = base colour, pat_b = grid colour)
 */
typedef struct rtc_shape {
    int32_t kind;
    int32_t world_id;
    double inv[16];
    double inv_t[16];
    int32_t has_color;
    int32_t pattern_kind;
    double color[3];
    double ambient;
    double diffuse;
    double specular;
    double shininess;
    double reflective;
    double transparency;
    double refractive_index;
    double pat_inv[16];
    double pat_a[3];
    double pat_b[3];
} rtc_shape;

/* World.light (src/shape.rs:635, src/material.rs:10-14): a point light.  The reference's World holds one. */
typedef struct rtc_light {
    double intensity[3];
    double position[3];
} rtc_light;

/* rtc_scene_upload_lights accepts up to this many lights. */
#define RTC_MAX_LIGHTS 8

/* Camera (src/camera.rs:16-27): the derived fields computed by Camera::new (:39-58) and
 * the private view_transform_inv (:25).  samples = antialiasing_samples (a u8, 0..255), Camera::render_pixel
 * (:94-114): 1 = one centre sample; otherwise four fixed sub-samples at (.25/.75)^2, their mean, and — when
 * any of the four is further than 0.01 from the mean (Color::distance_from, src/color.rs:122-126) — `samples`
 * extra samples at random sub-pixel offsets, the mean taken over all 4 + samples (Camera::resample, :84-92).
 * samples = 0 is therefore the deterministic 2x2 supersample.
 *
 * Resample offsets.  The reference draws them from rand::thread_rng() (:85-89; x offset first, then y), so two
 * of its own runs differ.  Here they come from a counter-based stream, reproducible and independent of the
 * order pixels are rendered in:
 *     u(pixel, k, axis) = splitmix64( splitmix64(aa_seed) ^ ((pixel_index << 9) | (k << 1) | axis) )
 *     splitmix64(x): x += 0x9E3779B97F4A7C15; x = (x ^ x>>30) * 0xBF58476D1CE4E5B9;
 *                    x = (x ^ x>>27) * 0x94D049BB133111EB; return x ^ x>>31          (64-bit wrap-around)
 *     offset = (u >> 11) * 2^-53 in [0, 1)      (f32 fast mode: (u >> 40) * 2^-24)
 * with pixel_index = y*hsize + x, k = 0..samples-1 the extra sample, axis 0 = x, 1 = y. */
typedef struct rtc_camera {
    uint32_t hsize;
    uint32_t vsize;
    double half_width;
    double half_height;
    double pixel_size;
    double view_inv[16];
    uint32_t samples;
    uint32_t aa_seed;   /* seed of the resample offsets (samples >= 2); any value */
} rtc_camera;

enum { RTC_PRECISION_F64_PARITY = 0, RTC_PRECISION_F32_FAST = 1 };
/* RTC_RECT_REFERENCE_RENDER reproduces Camera::render's loop bounds, which skip the last row
 * and the last column (src/camera.rs:120-121); RTC_RECT_FULL is render_async's (:149). */
enum { RTC_RECT_REFERENCE_RENDER = 0, RTC_RECT_FULL = 1 };
/* F64_RGB is Canvas.pixels' layout: row-major y*width+x, 3 x f64 (src/canvas.rs:17-22,43-51).
 * U8_RGB applies Color::scale's truncate-then-clamp rule (src/color.rs:100-114). */
enum { RTC_OUT_F64_RGB = 0, RTC_OUT_F32_RGB = 1, RTC_OUT_U8_RGB = 2 };

typedef struct rtc_render_opts {
    int32_t precision;   /* RTC_PRECISION_* */
    int32_t rect;        /* RTC_RECT_* */
    int32_t max_depth;   /* Camera::MAX_REFLECTIONS = 5 (src/camera.rs:31); 0..RTC_MAX_DEPTH */
    int32_t out_format;  /* RTC_OUT_* */
    /* Row-tile sharding.  Tile k = rows [k*tile_rows, (k+1)*tile_rows) belongs to shard k % shard_count.
     * shard_count <= 1 renders the whole frame.  rtc_render_device: the shard's compact buffer (or its rows of a
     * full frame with RTC_OPT_DST_FULL_FRAME).  rtc_render on a one-GPU context: out_rgb / out_hit_mask are the FULL
     * frame and only the shard's rows are written (one process per GPU filling a frame in shared host memory, every
     * GPU over its own PCIe link); a context with several GPUs shards over them itself and ignores these fields. */
    int32_t shard_index;
    int32_t shard_count;
    int32_t tile_rows;
    int32_t flags;       /* RTC_OPT_* bits */
} rtc_render_opts;

/* rtc_render_device only: the destination buffers are FULL frames (hsize*vsize pixels) and the
 * shard's rows are written at their frame positions instead of into a compact shard buffer.
 * With a peer-mapped frame of another GPU (rtc_ipc_open) this is the fused render + gather:
 * finished pixels travel over NVLink as the kernel produces them, no collective afterwards. */
#define RTC_OPT_DST_FULL_FRAME 1
/* f32 fast mode: use the general f32 kernel even when the library would choose another one — the constant-bank small-scene
 * kernel for scenes that fit it, the parity kernel for scenes of more than 128 shapes (it is the fastest kernel that meets
 * the fast-mode tolerance there, and exact; the general kernel then runs its mixed-precision form: f32 hierarchy and
 * shading, geometry of the radiance rays in f64).  For testing and measuring one against the other. */
#define RTC_OPT_F32_GENERAL_KERNEL 2
/* f32 small-scene kernel: hand tiles out in linear order instead of "tiles that see a transparent
 * surface first" (the scheduling pre-pass).  Same image either way; for measuring the difference. */
#define RTC_OPT_LINEAR_TILE_ORDER 4
/* Both precisions, scenes with a bounding-sphere hierarchy (the parity kernel: any scene with spheres or cubes; fast
 * mode: scenes whose streams stay in global memory): run every shape's exact test instead of filtering by bounding
 * sphere first.  Same image, hit mask and counters either way (the filter is conservative); for checking that. */
#define RTC_OPT_NO_BOUND_FILTER 8
/* rtc_render: deliver by DMA straight into the caller's buffer in out_format (no pinned staging, no host pass).  Needs
 * the buffer registered (rtc_register_host_buffer), else RTC_ERR_INVALID_ARG.  Without a delivery flag the library chooses:
 * registered buffers get the direct delivery except for a whole fast-mode frame on one GPU into an F64_RGB Canvas, where
 * the one link is the limit: most of that frame crosses PCIe as f32 (half the bytes) and is widened by host threads, and
 * the last few row bands are written as f64 by the kernel and go by DMA straight into the buffer, so that link and host
 * threads finish together (how many bands follows the previous frame's timing; f32 -> f64 is exact on either side, the
 * frame does not depend on the split). */
#define RTC_OPT_DELIVER_DIRECT 16
/* rtc_render: never deliver directly (measure the staged path with a registered buffer).  Both delivery flags together:
 * the mixed delivery above with a fixed split (half the bands each way) — for tests. */
#define RTC_OPT_DELIVER_STAGED 32

#define RTC_MAX_DEPTH 8

/* Work counters of one render ("ray" = one World::intersect call, src/shape.rs:677). */
typedef struct rtc_stats {
    uint64_t rays;
    uint64_t primary_hits;
    uint64_t shaded_hits;
    uint64_t shadow_rays;
    uint64_t kernel_launches;
    double h2d_ms;
    double kernel_ms;
    double gather_ms;
    double d2h_ms;
    double total_ms; /* host wall clock of the whole call */
} rtc_stats;

typedef struct rtc_context rtc_context;

/* ABI version of the loaded library (== RTC_ABI_VERSION). */
int rtc_abi_version(void);

/* Message for the last non-RTC_OK status returned on this thread. */
const char* rtc_last_error(void);

/* Number of visible CUDA devices (0 when there is no driver/device). */
int rtc_device_count(void);

/* A context owns streams, the device scene, frame buffers and pinned staging for the GPUs in
 * device_ids (n_gpus >= 1; NULL device_ids = devices 0..n_gpus-1).  It amortises set-up over
 * frames (the reference re-renders per GIF frame: src/lua.rs:34-41).  One thread at a time
 * per context and ONE launch stream in flight per context: the persistent kernels share the context's tile
 * counter, so a render enqueued on a second stream (or rtc_render beside a user-stream rtc_render_device)
 * must be ordered after the previous one by the caller (an event or a synchronise).  Different contexts may be
 * used concurrently (Camera::render takes &self,
 * &World and both are Sync: src/shape.rs:17, src/material.rs:33).
 * The host threads that deliver staged frames (rtc_render) are created with the first render: one per hardware thread, or
 * the number in the environment variable RTC_B200_HOST_THREADS, read once at that point — the library's only environment
 * variable (a deployment knob like OMP_NUM_THREADS; nothing on a render call's path reads the environment). */
int rtc_context_create(const int* device_ids, int n_gpus, rtc_context** out_ctx);
void rtc_context_destroy(rtc_context* ctx);

/* Scene flattening pass: takes World.shapes in add_shape order (src/shape.rs:661-667) and
 * World.light, and packs them into structure-of-arrays device buffers on every GPU of the
 * context.  Replaces what World::intersect (src/shape.rs:677-683) reads per ray.  `light` and `shapes` are not referenced
 * after the call returns.  On a one-GPU context the arrays travel as one pinned blob and the call returns without
 * waiting for the copy: every later call of this library on the context is ordered behind it (on the context's own stream
 * by stream order, on a caller's stream by an event wait the library inserts).  Uploading a new scene while a render of
 * the old one is still in flight on another stream is the caller's to order, as it always was. */
int rtc_scene_upload(rtc_context* ctx, const rtc_light* light, const rtc_shape* shapes, size_t n_shapes);

/* Same flattening pass for a world with several point lights — the "FIXME -- multiple lights" of World::shade_hit
 * (src/shape.rs:686; the Lua loader already reads a list of lights and keeps the first, src/lua.rs:148-150).
 * The reference has no statement of this to mirror; the semantics are the book's shade_hit, built from the
 * reference's own pieces: surface = lighting(light_0, .., is_shadowed_by_light(over_point, light_0))
 * + lighting(light_1, .., is_shadowed_by_light(over_point, light_1)) + ... added in list order (each term carries the
 * material's ambient share, as Material::lighting does, src/material.rs:335-340), one shadow ray per light and
 * shaded hit.  n_lights = 1 is rtc_scene_upload exactly.  1 <= n_lights <= RTC_MAX_LIGHTS. */
int rtc_scene_upload_lights(rtc_context* ctx, const rtc_light* lights, size_t n_lights, const rtc_shape* shapes,
                            size_t n_shapes);

/* rtc_render_device without any allocation.  No counterpart in the reference; added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 * rtc_render_device may grow the context's tile-order buffer (cudaMalloc) on the first frame of a size; this variant of
 * the small-scene fast kernel keeps its scratch — the tile counter words and the tile hand-out order — in a static device
 * array of the library (4 MiB + 256 B, held by every GPU that loads the library until the process exits) and calls no
 * cudaMalloc, cudaFree or allocating library routine.  Inputs and outputs are the caller's, as in rtc_render_device
 * (same opts, shards and RTC_OPT_DST_FULL_FRAME included; the frame is bit-identical to rtc_render_device's).
 * Limits: RTC_PRECISION_F32_FAST on a scene that fits the small-scene form (<= 16 planes, <= 48 similarity-transformed
 * spheres, one light, light outside every sphere) on a one-GPU context, else RTC_ERR_UNSUPPORTED; at most
 * RTC_NOALLOC_MAX_TILES tiles of 8x4 pixels per call (7680x4320 fits), else RTC_ERR_UNSUPPORTED with the counts in
 * rtc_last_error(); `stream` is required and nothing is synchronised; no work counters.
 * Every call on a GPU shares the one array (whatever the context): calls that overlap on that GPU race — order them on
 * one stream or with events.  The array keeps the last call's contents; the call clears the words that must start at
 * zero on `stream` before the launch and rewrites every order word it reads. */
#define RTC_NOALLOC_MAX_TILES (1 << 20)
int rtc_render_device_noalloc(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts, void* d_out_rgb,
                              uint8_t* d_out_hit_mask, void* stream);

/* Long-lived caller frames (the Canvas a renderer reuses per GIF frame, src/lua.rs:34-41; a frame in memory shared by
 * one process per GPU): register the buffer once (cudaHostRegister: page-locked and visible to every GPU of the machine)
 * and rtc_render delivers finished rows into it by DMA — no staging copy and no pass by host threads; each GPU writes
 * its rows over its own PCIe link.  ptr / bytes must cover every out_rgb later passed from inside the range.
 * Unregister before freeing the memory.  Replaces nothing in the reference (Canvas.pixels is a Vec, src/canvas.rs:17-22). */
int rtc_register_host_buffer(rtc_context* ctx, void* ptr, size_t bytes);
int rtc_unregister_host_buffer(rtc_context* ctx, void* ptr);

/* What the last rtc_render on this context moved over PCIe and how: bytes copied device -> host for the frame (all GPUs
 * of the context), the row bands (kernel launches) it was rendered in, and how many of those went by DMA straight into the
 * caller's buffer (0: all staged; == bands: all direct; between: the mixed delivery of RTC_OPT_DELIVER_DIRECT's comment).
 * Any pointer may be null.  Measurement support for the bench line's d2h_bytes_per_step; no counterpart in the reference. */
int rtc_delivery_info(rtc_context* ctx, uint64_t* d2h_bytes, int* bands, int* bands_direct);

/* Camera::render / render_async (src/camera.rs:116-160): renders the uploaded scene into the
 * caller's HOST buffers.  out_rgb: hsize*vsize*3 elements of opts->out_format, row-major
 * y*hsize+x (Canvas.pixels, src/canvas.rs:43-46); pixels outside the rect are 0 (src/canvas.rs:37-41).
 * out_hit_mask (nullable): hsize*vsize bytes, bit k set when sub-sample k's primary ray hit
 * a shape (World::color_at's Some/None, src/shape.rs:704-709; the four fixed sub-samples only, not the
 * resample pass).  stats nullable. */
int rtc_render(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts, void* out_rgb,
               uint8_t* out_hit_mask, rtc_stats* stats);

/* Rows of the frame owned by a shard under opts' tiling (the compact buffer height). */
int rtc_shard_rows(const rtc_camera* cam, const rtc_render_opts* opts);

/* Same render, DEVICE buffers on the context's first GPU, enqueued on `stream`
 * (a cudaStream_t; NULL = the context's own stream, synchronised before returning).
 * A context with several GPUs shards the frame over them (interleaved row tiles of opts->tile_rows rows, tile k ->
 * GPU k % n_gpus; opts' shard fields are ignored): every GPU's kernel stores its pixels straight into the frame on the
 * first GPU through peer access (NVLink), ordered by events — `stream` continues when all of them are done.  The
 * destination is then always a full frame.  RTC_ERR_UNSUPPORTED when a GPU cannot reach the first one.
 * With shard_count > 1, d_out_rgb / d_out_hit_mask are the shard's compact buffers:
 * rtc_shard_rows() rows of hsize pixels, owned tiles stacked in frame order.
 * Counters and kernel_ms in `stats` are only filled when stream == NULL. */
int rtc_render_device(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts,
                      void* d_out_rgb, uint8_t* d_out_hit_mask, void* stream, rtc_stats* stats);

/* De-interleave: d_gathered holds shard_count compact shard buffers back to back, each padded
 * to `shard_stride_rows` rows (e.g. the output of an NCCL all-gather / gather of
 * rtc_render_device results); writes the assembled frame (hsize*vsize pixels of out_format)
 * to d_frame.  `elem_kind` is RTC_OUT_* or -1 for the 1-byte hit mask. */
int rtc_assemble_device(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts,
                        const void* d_gathered, int shard_stride_rows, int elem_kind, void* d_frame,
                        void* stream);

/* Cross-process peer access to a device buffer (one process per GPU, torchrun style).
 * rtc_ipc_export: fills `handle` (RTC_IPC_HANDLE_BYTES bytes, = cudaIpcMemHandle_t) for the
 * allocation containing d_ptr and the byte offset of d_ptr inside it; the bytes can be sent to
 * another process by any transport.  rtc_ipc_open maps that allocation into this process on the
 * context's first GPU (enabling peer access) and returns the device pointer corresponding to
 * d_ptr; rtc_ipc_close unmaps it.  Replaces nothing in the reference (it has no multi-device
 * path): this is the NVLink side of the multi-GPU framebuffer assembly. */
#define RTC_IPC_HANDLE_BYTES 64
int rtc_ipc_export(rtc_context* ctx, const void* d_ptr, void* handle, uint64_t* offset);
int rtc_ipc_open(rtc_context* ctx, const void* handle, uint64_t offset, void** out_d_ptr);
int rtc_ipc_close(rtc_context* ctx, void* d_ptr, uint64_t offset);

/* Completion flags for the peer-store frame assembly.  rtc_signal_device enqueues, after everything
 * already on `stream` (NULL = the context's stream), a store of `value` to the 32-bit word at d_flag —
 * which may be a peer-mapped word of another GPU (rtc_ipc_open) — with system-scope release: all
 * earlier stores of this stream (the rendered pixels) are visible before the flag.  rtc_wait_device
 * enqueues a kernel that returns once d_flags[i] >= value for all i in [0, n) (no host synchronisation either
 * way).  Rank 0 of a sharded render waits for the other ranks' flags instead of running a collective.
 * A wait gives up after the context's wait timeout (default 60 s — parity-mode frames of big scenes take seconds;
 * rtc_set_wait_timeout changes it): it then records the index of the first missing flag in the context and returns
 * — the stream goes on and the CUDA context stays usable; whatever the stream does next sees an incomplete frame.
 * rtc_wait_result synchronises `stream` and reports it: RTC_OK, or RTC_ERR_TIMEOUT with *first_missing = i (the rank
 * whose flag did not arrive; nullable) — call it before consuming a frame when a peer may have died.
 * The same flags give back-pressure between frames: the consumer signals "frame k consumed" after its last read and
 * the producers wait for it before they overwrite the frame (raytracer-challenge_b200/sharded.py). */
int rtc_signal_device(rtc_context* ctx, void* d_flag, uint32_t value, void* stream);
int rtc_wait_device(rtc_context* ctx, const void* d_flags, int n, uint32_t value, void* stream);
/* rtc_wait_device and, in the same kernel, rtc_signal_device(d_then_flag, then_value) once every flag has arrived (not after a
 * timeout): a rank 0 without a consumer of its own between frames releases the frame ("consumed") without a second launch.
 * 1 <= n <= 1024. */
int rtc_wait_signal_device(rtc_context* ctx, const void* d_flags, int n, uint32_t value, void* d_then_flag, uint32_t then_value,
                           void* stream);
int rtc_set_wait_timeout(rtc_context* ctx, double seconds);
int rtc_wait_result(rtc_context* ctx, void* stream, int* first_missing);

/* rtc_render_device with the flag protocol folded into the launch (one process per GPU, peer-store frame assembly):
 * the render starts only once *d_wait_flag >= wait_value (NULL: at once) — the consumer has released the destination
 * frame — and when the last pixel is out, done_value is release-stored to *d_done_flag at system scope (NULL: no
 * store).  Both flags may be peer-mapped words of another GPU.  The small-scene fast kernel does both inside its one
 * launch (every block polls the wait flag before its first store; the last block to retire stores the done flag after
 * system-scope fences), so a rank's frame is a single kernel; the other kernels get the flag kernels of
 * rtc_wait_device / rtc_signal_device enqueued around them.  A wait that times out is reported by rtc_wait_result.
 * One-GPU contexts; `stream` is required (nothing is synchronised). */
int rtc_render_device_synced(rtc_context* ctx, const rtc_camera* cam, const rtc_render_opts* opts, void* d_out_rgb,
                             uint8_t* d_out_hit_mask, void* stream, const void* d_wait_flag, uint32_t wait_value,
                             void* d_done_flag, uint32_t done_value);

/* Canvas delivery for the device-buffer entry points: copies a finished frame (hsize*vsize pixels of
 * `dev_format`, on the context's first GPU) into the caller's HOST buffer of `out_format`, after everything
 * already enqueued on `stream` (NULL = the context's stream).  Row bands cross PCIe on the copy stream while
 * the host delivers the previous band; dev_format F32_RGB with out_format F64_RGB widens on the host (exact),
 * which halves the PCIe bytes of a fast-mode frame.  This is the step between rank 0's assembled frame
 * (rtc_render_device + peer stores / rtc_assemble_device) and Canvas.pixels (src/canvas.rs:17-22).
 * Supported: dev_format == out_format, or F32 -> F64.  Returns when out_rgb is complete. */
int rtc_download_frame(rtc_context* ctx, const void* d_frame, int dev_format, uint32_t hsize, uint32_t vsize,
                       int out_format, void* out_rgb, void* stream);

/* Canvas::to_imgbuf (src/canvas.rs:61-79), the step behind the path when an image file is written: packs a
 * finished device frame (F64_RGB or F32_RGB, n_pixels pixels) into RGBA8 on the device — per channel
 * Color::scale(component.powf(1/gamma), 255) with the reference's truncate-then-clamp rule
 * (src/color.rs:55-65,100-114), alpha 255; `gamma` is Canvas.gamma (1.0 by default, src/canvas.rs:30).
 * 4 bytes per pixel instead of 24 then cross NVLink / PCIe.  Both buffers 16-byte aligned.
 * Enqueued on `stream` (NULL = the context's stream, synchronised before returning). */
int rtc_pack_rgba8_device(rtc_context* ctx, const void* d_frame, int frame_format, size_t n_pixels, float gamma,
                          void* d_out_rgba, void* stream);

/* Register-resident FMA micro-benchmark on the context's first GPU: measured non-tensor peak
 * of the FP32 (precision = RTC_PRECISION_F32_FAST) or FP64 pipe in TFLOP/s (FMA = 2 flops),
 * used as the roofline denominator beside the formula peak. */
int rtc_measure_peak_tflops(rtc_context* ctx, int precision, double* out_tflops);

/* Host-only self-check of the scene flattening pass (no device needed, nothing is uploaded): runs the pass on the
 * given world and verifies what the kernels rely on — every shape sits in exactly one intersect stream of each
 * precision (the streams of big scenes are reordered along a Morton curve), and every shape's bounding sphere lies
 * inside the bounding sphere of its group of eight, every group's inside its block's (the two upper levels of the
 * line-distance filter in front of the exact shape tests).  RTC_OK and *out filled when the pass ran; the caller
 * inspects the fields.  Replaces nothing in the reference: World::intersect is a plain loop (src/shape.rs:677-683). */
typedef struct rtc_selfcheck {
    int32_t n_planes, n_uniform_spheres, n_general_spheres, n_cubes;
    int32_t streams_are_permutations;   /* 1: each shape index appears exactly once per precision */
    int32_t morton_sorted;              /* 1: at least one stream was reordered (>= 64 entries) */
    int32_t n_groups_f32, n_blocks_f32, n_groups_f64, n_blocks_f64;   /* 0: no filter for that precision */
    int32_t small_scene_form;           /* 1: the scene fits the constant-bank f32 kernel */
    int32_t reserved;
    double worst_enclosure_f32;         /* max over entries of (distance of centres + r_child - r_parent) / r_parent; <= 0 = enclosed */
    double worst_enclosure_f64;
} rtc_selfcheck;
int rtc_scene_selfcheck(const rtc_light* lights, size_t n_lights, const rtc_shape* shapes, size_t n_shapes,
                        rtc_selfcheck* out);

#ifdef __cplusplus
}
#endif
#endif /* RTC_B200_H */
