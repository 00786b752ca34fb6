// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// C entry points (for ctypes) over the C++ restatement in orc_*.hpp.  Worlds are built
// from the same flat descriptors the product's C ABI takes (include/rtc.h), through the
// restated constructors' "stored matrices verbatim" path, so oracle and CUDA path see
// the same inputs.  Also exposes the restated transform builders (src/transform.rs:53-217)
// so tests can check the product's host-side mirror against them bit for bit.
#include <chrono>
#include <cstring>
#include <mutex>
#include <string>
#include "../include/rtc.h"
#include "orc_world.hpp"

using namespace orc;

namespace {
thread_local std::string g_err;

std::shared_ptr<Pattern> make_pattern(const rtc_shape& d) {
    Color a(d.pat_a[0], d.pat_a[1], d.pat_a[2]), b(d.pat_b[0], d.pat_b[1], d.pat_b[2]);
    std::shared_ptr<Pattern> p;
    switch (d.pattern_kind) {
        case RTC_PATTERN_NONE: return nullptr;
        case RTC_PATTERN_TEST: p = std::make_shared<TestPattern>(); break;
        case RTC_PATTERN_STRIPE: p = std::make_shared<StripePattern>(a, b); break;
        case RTC_PATTERN_GRADIENT: p = std::make_shared<GradientPattern>(a, b); break;
        case RTC_PATTERN_RING: p = std::make_shared<RingPattern>(a, b); break;
        case RTC_PATTERN_CHECKER: p = std::make_shared<CheckerPattern>(a, b); break;
        case RTC_PATTERN_GRID: p = std::make_shared<GridPattern>(a, b); break;
        default: throw std::runtime_error("bad pattern kind");
    }
    p->set_transform_inverse_raw(Matrix::from_row_major(d.pat_inv));
    return p;
}

Material make_material(const rtc_shape& d) {
    Material m;
    m.pattern = make_pattern(d);
    m.has_color = d.has_color != 0;
    m.color = Color(d.color[0], d.color[1], d.color[2]);
    m.ambient = d.ambient;
    m.diffuse = d.diffuse;
    m.specular = d.specular;
    m.shininess = d.shininess;
    m.reflectiveness = d.reflective;
    m.transparency = d.transparency;
    m.refractive_index = d.refractive_index;
    return m;
}
}  // namespace

struct orc_world {
    World world;
    explicit orc_world(const Light& l) : world(l) {}
};

struct orc_counters {
    uint64_t rays, shadow_rays, shaded_hits, sphere_tests, plane_tests, cube_tests;
    double seconds;
};

extern "C" {

const char* orc_last_error(void) { return g_err.c_str(); }

orc_world* orc_world_create(const rtc_light* light, const rtc_shape* shapes, size_t n) {
    try {
        Light l(Color(light->intensity[0], light->intensity[1], light->intensity[2]),
                Point(light->position[0], light->position[1], light->position[2]));
        orc_world* w = new orc_world(l);
        for (size_t i = 0; i < n; ++i) {
            const rtc_shape& d = shapes[i];
            Matrix inv = Matrix::from_row_major(d.inv), inv_t = Matrix::from_row_major(d.inv_t);
            Material m = make_material(d);
            std::unique_ptr<Shape> s;
            if (d.kind == RTC_SHAPE_SPHERE) s = std::make_unique<Sphere>(Sphere::Raw{}, inv, inv_t, m);
            else if (d.kind == RTC_SHAPE_PLANE) s = std::make_unique<Plane>(Plane::Raw{}, inv, inv_t, m);
            else if (d.kind == RTC_SHAPE_CUBE) s = std::make_unique<Cube>(Cube::Raw{}, inv, inv_t, m);
            else throw std::runtime_error("bad shape kind");
            w->world.add_shape(std::move(s));
            if ((int)w->world.shapes.back()->get_world_id() != d.world_id) {
                delete w;
                throw std::runtime_error("descriptor world_id differs from add_shape order");
            }
        }
        return w;
    } catch (const std::exception& e) {
        g_err = e.what();
        return nullptr;
    }
}

// Lights 1..n-1 of a multi-light world (extension, orc_world.hpp shade_hit); light 0 is orc_world_create's.
int orc_world_add_light(orc_world* w, const rtc_light* light) {
    w->world.extra_lights.emplace_back(Color(light->intensity[0], light->intensity[1], light->intensity[2]),
                                       Point(light->position[0], light->position[1], light->position[2]));
    return 0;
}

void orc_world_destroy(orc_world* w) { delete w; }

// Renders with the restated Camera::render_pixel.  rect: 0 = Camera::render's bounds
// (src/camera.rs:120-121), 1 = full (render_async, :149).  pixel_indices (nullable): render only
// these y*hsize+x indices, results compact in the same order (pixels outside the rect give 0).
// threads = 1 mirrors render; > 1 mirrors render_async (dynamic chunks over a thread pool).
int orc_render(const orc_world* w, const rtc_camera* cam, int rect, int max_depth, int threads,
               const uint32_t* pixel_indices, size_t n_pixels, double* out_rgb, uint8_t* out_hit,
               orc_counters* counters) {
    try {
        Camera c(cam->hsize, cam->vsize, 0.0);
        c.half_width = cam->half_width;
        c.half_height = cam->half_height;
        c.pixel_size = cam->pixel_size;
        c.view_transform_inv = Matrix::from_row_major(cam->view_inv);
        c.antialiasing_samples = (uint8_t)cam->samples;
        c.max_reflections = (uint8_t)max_depth;
        if (cam->samples > 255) throw std::runtime_error("oracle: antialiasing_samples is a u8 (src/camera.rs:26)");
        c.aa_seed = cam->aa_seed;
        const size_t total = pixel_indices ? n_pixels : (size_t)cam->hsize * cam->vsize;
        const uint32_t xlim = rect == 0 ? cam->hsize - 1 : cam->hsize;
        const uint32_t ylim = rect == 0 ? cam->vsize - 1 : cam->vsize;
        if (threads < 1) threads = 1;
        std::atomic<size_t> next{0};
        std::mutex mu;
        Counters sum;
        std::string err;
        const size_t chunk = 256;
        auto t0 = std::chrono::steady_clock::now();
        auto worker = [&]() {
            tls_counters() = Counters();
            try {
                for (;;) {
                    size_t begin = next.fetch_add(chunk);
                    if (begin >= total) break;
                    size_t end = std::min(total, begin + chunk);
                    for (size_t k = begin; k < end; ++k) {
                        size_t idx = pixel_indices ? pixel_indices[k] : k;
                        uint32_t x = (uint32_t)(idx % cam->hsize), y = (uint32_t)(idx / cam->hsize);
                        Color col(0., 0., 0.);
                        uint8_t bits = 0;
                        if (x < xlim && y < ylim) col = c.render_pixel(x, y, w->world, &bits);
                        out_rgb[3 * k + 0] = col.red;
                        out_rgb[3 * k + 1] = col.green;
                        out_rgb[3 * k + 2] = col.blue;
                        if (out_hit) out_hit[k] = bits;
                    }
                }
            } catch (const std::exception& e) {
                std::lock_guard<std::mutex> g(mu);
                err = e.what();
            }
            std::lock_guard<std::mutex> g(mu);
            sum.add(tls_counters());
        };
        if (threads == 1) {
            worker();
        } else {
            std::vector<std::thread> pool;
            for (int i = 0; i < threads; ++i) pool.emplace_back(worker);
            for (auto& t : pool) t.join();
        }
        auto t1 = std::chrono::steady_clock::now();
        if (!err.empty()) throw std::runtime_error(err);
        if (counters) {
            counters->rays = sum.rays;
            counters->shadow_rays = sum.shadow_rays;
            counters->shaded_hits = sum.shaded_hits;
            counters->sphere_tests = sum.sphere_tests;
            counters->plane_tests = sum.plane_tests;
            counters->cube_tests = sum.cube_tests;
            counters->seconds = std::chrono::duration<double>(t1 - t0).count();
        }
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// One sample of Camera::render_pixel: world.color_at(camera.ray_for_pixel_offset(x, xo, y, yo), max_depth)
// (src/camera.rs:64-76,96-105).  Lets the tests restate render_pixel / resample (:84-114) a second time, in
// Python, on top of single samples, and compare that with orc_render's statement of it.
int orc_sample(const orc_world* w, const rtc_camera* cam, uint32_t x, double xo, uint32_t y, double yo, int max_depth,
               double* out_rgb) {
    try {
        Camera c(cam->hsize, cam->vsize, 0.0);
        c.half_width = cam->half_width;
        c.half_height = cam->half_height;
        c.pixel_size = cam->pixel_size;
        c.view_transform_inv = Matrix::from_row_major(cam->view_inv);
        Color col = w->world.color_at(c.ray_for_pixel_offset(x, xo, y, yo), (uint8_t)max_depth);
        out_rgb[0] = col.red; out_rgb[1] = col.green; out_rgb[2] = col.blue;
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// The resample offset stream of include/rtc.h ("Resample offsets") as the oracle states it.
double orc_resample_offset(uint32_t aa_seed, uint32_t hsize, uint32_t x, uint32_t y, uint32_t k, uint32_t axis) {
    Camera c(hsize, 1, 0.0);
    c.aa_seed = aa_seed;
    return c.resample_offset(x, y, k, axis);
}

// ---- restated transform builders, for checking the product's host mirror -------------------
// ops: 0 translation(x,y,z) 1 scaling(x,y,z) 2 rotation_x(r) 3 rotation_y(r) 4 rotation_z(r)
//      5 shearing(xy,xz,yx,yz,zx,zy); args has 6 doubles per op.  Applied in listed order,
// each as `m = m.op(..)` on Matrix::identity() (src/transform.rs:53-105).
int orc_matrix_chain(const int* ops, const double* args, int n, double* out16) {
    try {
        Matrix m = Matrix::identity();
        for (int i = 0; i < n; ++i) {
            const double* a = args + 6 * i;
            switch (ops[i]) {
                case 0: m = m.translation(a[0], a[1], a[2]); break;
                case 1: m = m.scaling(a[0], a[1], a[2]); break;
                case 2: m = m.rotation_x(a[0]); break;
                case 3: m = m.rotation_y(a[0]); break;
                case 4: m = m.rotation_z(a[0]); break;
                case 5: m = m.shearing(a[0], a[1], a[2], a[3], a[4], a[5]); break;
                default: throw std::runtime_error("bad op");
            }
        }
        m.to_row_major(out16);
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// returns 0 and fills out16 when |det| > EPSILON (src/transform.rs:37), 2 when not invertible
int orc_matrix_inverse(const double* in16, double* out16, double* det) {
    Matrix m = Matrix::from_row_major(in16);
    if (det) *det = m.determinant_value;
    if (!m.is_invertable()) return 2;
    m.inverse().to_row_major(out16);
    return 0;
}

void orc_matrix_transpose(const double* in16, double* out16) {
    Matrix::from_row_major(in16).transpose().to_row_major(out16);
}

void orc_matrix_multiply(const double* a16, const double* b16, double* out16) {
    Matrix::from_row_major(a16).multiply(Matrix::from_row_major(b16)).to_row_major(out16);
}

void orc_view_transform(const double* from3, const double* to3, const double* up3, double* out16) {
    Matrix::make_view_transform(Point(from3[0], from3[1], from3[2]), Point(to3[0], to3[1], to3[2]),
                                Vector(up3[0], up3[1], up3[2]))
        .to_row_major(out16);
}

// Camera::new / new_with_transform (src/camera.rs:33-58); view16 nullable (identity).
int orc_camera_new(uint32_t hsize, uint32_t vsize, double fov, const double* view16, rtc_camera* out) {
    try {
        Camera c = view16 ? Camera(hsize, vsize, fov, Matrix::from_row_major(view16)) : Camera(hsize, vsize, fov);
        out->hsize = hsize;
        out->vsize = vsize;
        out->half_width = c.half_width;
        out->half_height = c.half_height;
        out->pixel_size = c.pixel_size;
        c.view_transform_inv.to_row_major(out->view_inv);
        out->samples = 1;
        out->aa_seed = 0;
        return 0;
    } catch (const std::exception& e) {
        g_err = e.what();
        return 1;
    }
}

// Color::scale (src/color.rs:100-114) over a buffer: the 8-bit rule used by the PPM writer.
void orc_quantize_u8(const double* rgb, size_t n, uint8_t* out) {
    for (size_t i = 0; i < n; ++i) out[i] = (uint8_t)Color::scale(rgb[i], 255);
}

}  // extern "C"
