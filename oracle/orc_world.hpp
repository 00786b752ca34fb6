// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement of Sphere / Plane / Cube, World (recursive shading), Camera and
// Canvas of joedane/raytracer-challenge `ch1`:
//   src/shape.rs:281-630 (shapes), :633-795 (World)
//   src/camera.rs:16-162, src/canvas.rs:16-51,86-109
// See orc_math.hpp for who may use this.
#pragma once
#include <algorithm>
#include <atomic>
#include <limits>
#include <thread>
#include "orc_scene.hpp"

namespace orc {

// src/shape.rs:281-394
struct Sphere : Shape {
    double r = 1.0;
    Matrix transform;            // NB: holds the INVERSE (:300)
    Matrix transform_transpose;  // inverse transposed (:301)
    Material material;
    uint8_t world_id = 0;

    Sphere() : Sphere(Matrix::identity()) {}                                          // :293
    explicit Sphere(const Matrix& m)                                                  // :297
        : transform(m.inverse()), transform_transpose(m.inverse().transpose()),
          material(Material::default_()) {}
    Sphere(const Matrix& m, const Material& mat)                                      // :308
        : transform(m.inverse()), transform_transpose(m.inverse().transpose()), material(mat) {}
    // flat-descriptor entry: stored matrices given verbatim
    struct Raw {};
    Sphere(Raw, const Matrix& inv, const Matrix& inv_t, const Material& mat)
        : transform(inv), transform_transpose(inv_t), material(mat) {}
    void set_transform(const Matrix& m) {                                             // :319
        transform = m.inverse();
        transform_transpose = transform.transpose();
    }
    static Sphere glass_sphere() {                                                    // :324
        Sphere s;
        s.material.transparency = 1.0;
        s.material.refractive_index = 1.5;
        return s;
    }
    void set_world_id(uint8_t id) override { world_id = id; }
    uint8_t get_world_id() const override { return world_id; }
    const Matrix& get_transform_inverse() const override { return transform; }
    const Matrix& get_transform_inverse_transpose() const override { return transform_transpose; }
    Intersections intersect_local(const Ray& ray) const override {                    // :361-375
        tls_counters().sphere_tests++;
        Vector sphere_to_ray = ray.origin.sub(Point(0., 0., 0.));
        double a = ray.direction.dot(ray.direction);
        double b = 2. * ray.direction.dot(sphere_to_ray);
        double c = sphere_to_ray.dot(sphere_to_ray) - 1.;
        double discriminant = (b * b) - 4. * a * c;
        if (discriminant < 0.) {
            return Intersections();
        } else {
            Intersections i(Intersection((-b - std::sqrt(discriminant)) / (2. * a), this));
            i.add(Intersection((-b + std::sqrt(discriminant)) / (2. * a), this));
            return i;
        }
    }
    Vector normal_at_local(const Point& p) const override { return p.sub(Point(0., 0., 0.)); } // :377
    const Material& get_material() const override { return material; }
    Material& get_material_mut() override { return material; }
};

// src/shape.rs:405-498
struct Plane : Shape {
    Matrix transform;
    Matrix transform_transpose;
    Material material;
    Vector local_normal = Vector(0., 1., 0.);
    uint8_t world_id = 0;

    Plane() : transform(Matrix::identity()), transform_transpose(Matrix::identity().transpose()),
              material(Material::default_()) {}                                       // :416
    explicit Plane(const Matrix& xf)                                                  // :426
        : transform(xf.inverse()), transform_transpose(xf.inverse().transpose()),
          material(Material::default_()) {}
    Plane(const Matrix& xf, const Material& mat)                                      // :436
        : transform(xf.inverse()), transform_transpose(xf.inverse().transpose()), material(mat) {}
    struct Raw {};
    Plane(Raw, const Matrix& inv, const Matrix& inv_t, const Material& mat)
        : transform(inv), transform_transpose(inv_t), material(mat) {}
    void set_transform(const Matrix& m) {                                             // :446-449 (bug-compatible)
        transform = m.inverse();
        transform_transpose = transform_transpose.transpose();
    }
    void set_world_id(uint8_t id) override { world_id = id; }
    uint8_t get_world_id() const override { return world_id; }
    Intersections intersect_local(const Ray& ray) const override {                    // :462-471
        tls_counters().plane_tests++;
        if (std::fabs(ray.direction.y) < EPSILON) {
            return Intersections::new_empty();
        } else {
            double t = -ray.origin.y / ray.direction.y;
            return Intersections(Intersection(t, this));
        }
    }
    const Matrix& get_transform_inverse() const override { return transform; }
    const Matrix& get_transform_inverse_transpose() const override { return transform_transpose; }
    Vector normal_at_local(const Point&) const override { return local_normal; }      // :481
    const Material& get_material() const override { return material; }
    Material& get_material_mut() override { return material; }
};

// src/shape.rs:500-630
struct Cube : Shape {
    Matrix transform;
    Matrix transform_transpose;
    Material material;
    uint8_t world_id = 0;

    Cube() : Cube(Matrix::identity()) {}
    explicit Cube(const Matrix& m)
        : transform(m.inverse()), transform_transpose(m.inverse().transpose()),
          material(Material::default_()) {}
    Cube(const Matrix& m, const Material& mat)
        : transform(m.inverse()), transform_transpose(m.inverse().transpose()), material(mat) {}
    struct Raw {};
    Cube(Raw, const Matrix& inv, const Matrix& inv_t, const Material& mat)
        : transform(inv), transform_transpose(inv_t), material(mat) {}
    void set_transform(const Matrix& m) {                                             // :535-538
        transform = m.inverse();
        transform_transpose = transform_transpose.transpose();
    }
    static std::pair<double, double> check_axis(double origin, double direction) {   // :540-564
        double tmin_numerator = -1.0 - origin;
        double tmax_numerator = 1. - origin;
        double tmin, tmax;
        if (std::fabs(direction) >= EPSILON) {
            tmin = tmin_numerator / direction;
            tmax = tmax_numerator / direction;
        } else {
            const double inf = std::numeric_limits<double>::infinity();
            tmin = tmin_numerator >= 0.0 ? inf : -inf;
            tmax = tmax_numerator >= 0.0 ? inf : -inf;
        }
        if (tmin > tmax) std::swap(tmin, tmax);
        return std::make_pair(tmin, tmax);
    }
    // Rust f64::max/min: if one argument is NaN the other is returned.
    static double rmax(double a, double b) { return std::fmax(a, b); }
    static double rmin(double a, double b) { return std::fmin(a, b); }
    void set_world_id(uint8_t id) override { world_id = id; }
    uint8_t get_world_id() const override { return world_id; }
    Intersections intersect_local(const Ray& ray) const override {                    // :577-591
        tls_counters().cube_tests++;
        auto xs = check_axis(ray.origin.x, ray.direction.x);
        auto ys = check_axis(ray.origin.y, ray.direction.y);
        auto zs = check_axis(ray.origin.z, ray.direction.z);
        double tmin = rmax(xs.first, rmax(ys.first, zs.first));
        double tmax = rmin(xs.second, rmin(ys.second, zs.second));
        Intersections is;
        if (tmin < tmax) {
            is.add(Intersection(tmin, this));
            is.add(Intersection(tmax, this));
        }
        return is;
    }
    const Matrix& get_transform_inverse() const override { return transform; }
    const Matrix& get_transform_inverse_transpose() const override { return transform_transpose; }
    Vector normal_at_local(const Point& p) const override {                           // :601-610
        double maxc = rmax(std::fabs(p.x), rmax(std::fabs(p.y), std::fabs(p.z)));
        if (maxc == std::fabs(p.x)) return Vector(p.x, 0., 0.);
        else if (maxc == std::fabs(p.y)) return Vector(0., p.y, 0.);
        else return Vector(0., 0., p.z);
    }
    const Material& get_material() const override { return material; }
    Material& get_material_mut() override { return material; }
    Vector normal_at(const Point& p) const override {                                 // :623-629
        const Matrix& xf = get_transform_inverse();
        Point local_point = p.transform(xf);
        Vector local_normal = normal_at_local(local_point);
        Vector world_normal = local_normal.transform(xf.transpose());
        return world_normal.normalize();
    }
};

// src/shape.rs:633-795
struct World {
    std::vector<std::unique_ptr<Shape>> shapes;
    Light light;
    std::vector<Light> extra_lights;   // extension, see shade_hit
    uint8_t last_world_id = 0;

    explicit World(const Light& l) : light(l) {}                                      // :642
    static World default_() {                                                         // :784-795
        World w{Light()};
        Material m = Material::solid_with_defaults(Color(0.8, 1.0, 0.6));
        m.diffuse = 0.7;
        m.specular = 0.2;
        w.add_shape(std::make_unique<Sphere>(Matrix::identity(), m));
        w.add_shape(std::make_unique<Sphere>(Matrix::identity().scaling(0.5, 0.5, 0.5)));
        return w;
    }
    Color background_color(const Ray&) const { return Color::BLACK(); }               // :652
    World& add_shape(std::unique_ptr<Shape> s) {                                      // :661-667
        uint8_t new_id = (uint8_t)(last_world_id + 1);  // release-build wrap
        s->set_world_id(new_id);
        last_world_id = new_id;
        shapes.push_back(std::move(s));
        return *this;
    }
    const Shape* get_shape(size_t i) const { return shapes[i].get(); }
    Shape* get_shape_mut(size_t i) { return shapes[i].get(); }

    Intersections intersect(const Ray& ray) const {                                   // :677-683
        tls_counters().rays++;
        Intersections is;
        for (const auto& s : shapes) is.merge(s->intersect(ray));
        return is;
    }
    Color shade_hit(const CachedVectors& comp, uint8_t reflections_remaining) const { // :685-700
        tls_counters().shaded_hits++;
        Color surface_color = comp.object->get_material().lighting(
            light, comp.object, comp.over_point, comp.eyev, comp.normal, is_shadowed(comp.over_point));
        // "FIXME -- multiple lights" (:686): the reference keeps one light and already passes it to
        // is_shadowed_by_light (:716).  EXTENSION (no reference statement to follow): further lights
        // contribute lighting(light_k, ..., is_shadowed_by_light(over_point, light_k)) each, added in list order — the
        // book's shade_hit.  With no extra light this is the reference's statement above, bit for bit.
        for (const Light& l : extra_lights)
            surface_color = surface_color.add(comp.object->get_material().lighting(
                l, comp.object, comp.over_point, comp.eyev, comp.normal, is_shadowed_by_light(comp.over_point, l)));
        Color reflected = reflected_color(comp, reflections_remaining);
        Color refracted = refracted_color(comp, reflections_remaining);
        const Material& material = comp.object->get_material();
        if (material.reflectiveness > 0.0 && material.transparency > 0.0) {
            double refl = reflectance(comp);
            return surface_color.add(reflected.mul_f64(refl).add(refracted.mul_f64(1.0 - refl)));
        } else {
            return surface_color.add(reflected).add(refracted);
        }
    }
    Color color_at(const Ray& r, uint8_t reflections_remaining, bool* hit_out = nullptr) const { // :702-710
        Intersections is = intersect(r);
        const Intersection* hit = is.get_hit();
        if (hit_out) *hit_out = hit != nullptr;
        if (hit) return shade_hit(hit->compute_vectors(r, is), reflections_remaining);
        return background_color(r);
    }
    bool is_shadowed(const Point& p) const { return is_shadowed_by_light(p, light); } // :712
    bool is_shadowed_by_light(const Point& p, const Light& l) const {                 // :716-727
        Vector v = l.position.sub(p);
        double distance = v.magnitude();
        Vector direction = v.normalize();
        Ray r(p, direction);
        tls_counters().shadow_rays++;
        Intersections is = intersect(r);
        const Intersection* h = is.get_hit();
        return h ? (h->t < distance) : false;
    }
    Color reflected_color(const CachedVectors& comps, uint8_t reflections_remaining) const { // :729-738
        if (reflections_remaining == 0 || comps.object->get_material().reflectiveness <= 0.) {
            return Color::BLACK();
        } else {
            Ray reflected_ray(comps.over_point, comps.reflectv);
            Color c = color_at(reflected_ray, (uint8_t)(reflections_remaining - 1));
            return c.mul_f64(comps.object->get_material().reflectiveness);
        }
    }
    Color refracted_color(const CachedVectors& comps, uint8_t refractions_remaining) const { // :751-766
        if (refractions_remaining == 0 || comps.object->get_material().transparency == 0.0)
            return Color::BLACK();
        double n_ratio = comps.n1 / comps.n2;
        double cos_i = comps.eyev.dot(comps.normal);
        double sin2_t = (n_ratio * n_ratio) * (1.0 - (cos_i * cos_i));
        if (sin2_t > 1.0) return Color::BLACK();
        double cos_t = std::sqrt(1.0 - sin2_t);
        Vector direction = comps.normal.mul(n_ratio * cos_i - cos_t).sub(comps.eyev.mul(n_ratio));
        Ray refract_ray(comps.under_point, direction);
        return color_at(refract_ray, (uint8_t)(refractions_remaining - 1))
            .mul_f64(comps.object->get_material().transparency);
    }
    static double powi5(double a) {  // f64::powi(5): a * ((a*a)*(a*a))
        double a2 = a * a;
        double a4 = a2 * a2;
        return a4 * a;
    }
    static double reflectance(const CachedVectors& comps) {                           // :768-781
        double cos = comps.eyev.dot(comps.normal);
        if (comps.n1 > comps.n2) {
            double n = comps.n1 / comps.n2;
            double sin2_t = (n * n) * (1.0 - (cos * cos));
            if (sin2_t > 1.0) return 1.0;
            double cos_t = std::sqrt(1.0 - sin2_t);
            cos = cos_t;
        }
        double q = (comps.n1 - comps.n2) / (comps.n1 + comps.n2);
        double r0 = q * q;
        return r0 + (1.0 - r0) * powi5(1.0 - cos);
    }
};

// src/canvas.rs:16-51
struct Canvas {
    uint32_t width, height;
    std::vector<Color> pixels;
    Canvas(uint32_t w, uint32_t h) : width(w), height(h), pixels((size_t)w * h, Color(0., 0., 0.)) {}
    void write_pixel(uint32_t x, uint32_t y, const Color& c) { pixels[(size_t)y * width + x] = c; }
    Color get_pixel(uint32_t x, uint32_t y) const { return pixels[(size_t)y * width + x]; }
};

// src/camera.rs:16-162
struct Camera {
    uint32_t hsize, vsize;
    double fov, half_height, half_width, pixel_size;
    uint8_t antialiasing_samples;
    Matrix view_transform_inv;
    static constexpr uint8_t MAX_REFLECTIONS = 5;                                     // :31
    uint8_t max_reflections = MAX_REFLECTIONS;  // test knob; the reference always uses 5

    Camera(uint32_t hsize_, uint32_t vsize_, double fov_)                             // :39-58
        : hsize(hsize_), vsize(vsize_), fov(fov_), antialiasing_samples(1),
          view_transform_inv(Matrix::identity()) {
        double half_view = std::tan(fov / 2.0);
        double aspect = (double)hsize / (double)vsize;
        if (aspect < 1.0) {
            half_height = half_view;
            half_width = half_view * aspect;
        } else {
            half_width = half_view;
            half_height = half_view / aspect;
        }
        pixel_size = (half_width * 2.) / (double)hsize;
    }
    Camera(uint32_t h, uint32_t v, double f, const Matrix& m) : Camera(h, v, f) {     // :33
        view_transform_inv = m.inverse();
    }
    void set_samples(uint8_t n) { antialiasing_samples = n; }
    Ray ray_for_pixel_offset(uint32_t x, double x_offset, uint32_t y, double y_offset) const { // :64-76
        double xoffset = ((double)x + x_offset) * pixel_size;
        double yoffset = ((double)y + y_offset) * pixel_size;
        double world_x = half_width - xoffset;
        double world_y = half_height - yoffset;
        const Matrix& vt = view_transform_inv;
        Point pixel = vt.transform_point(Point(world_x, world_y, -1.));
        Point origin = vt.transform_point(Point(0., 0., 0.));
        Vector direction = pixel.sub(origin).normalize();
        return Ray(origin, direction);
    }
    Ray ray_for_pixel(uint32_t x, uint32_t y) const { return ray_for_pixel_offset(x, 0.5, y, 0.5); }
    // Sub-pixel offsets of the resample pass.  The reference draws them from rand::thread_rng()
    // (src/camera.rs:85-89: x offset first, then y, per extra sample), which no run can reproduce; the
    // ABI (include/rtc.h, "Resample offsets") fixes a counter-based stream instead:
    //   u(pixel, k, axis) = splitmix64( splitmix64(aa_seed) ^ ((pixel_index << 9) | (k << 1) | axis) )
    //   offset = (u >> 11) * 2^-53   in [0, 1)
    // PARITY UNPINNED against the reference for samples >= 2 (it is non-deterministic there); everything
    // else in render_pixel / resample follows the reference statement by statement.
    uint32_t aa_seed = 0;
    static uint64_t splitmix64(uint64_t x) {
        x += 0x9E3779B97F4A7C15ull;
        x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
        x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
        return x ^ (x >> 31);
    }
    double resample_offset(uint32_t x, uint32_t y, unsigned k, unsigned axis) const {
        uint64_t pixel_index = (uint64_t)y * hsize + x;
        uint64_t u = splitmix64(splitmix64((uint64_t)aa_seed) ^ ((pixel_index << 9) | ((uint64_t)k << 1) | axis));
        return (double)(u >> 11) * (1.0 / 9007199254740992.0);
    }
    // :94-114 + resample :84-92.  `hit_bits`: bit k = fixed sub-sample k's primary ray hit something.
    Color render_pixel(uint32_t x, uint32_t y, const World& world, uint8_t* hit_bits = nullptr) const {
        bool h = false;
        if (antialiasing_samples == 1) {
            Color c = world.color_at(ray_for_pixel_offset(x, 0.5, y, 0.5), max_reflections, &h);
            if (hit_bits) *hit_bits = h ? 1 : 0;
            return c;
        } else {
            static const double ox[4] = {0.25, 0.75, 0.25, 0.75};
            static const double oy[4] = {0.25, 0.25, 0.75, 0.75};
            std::vector<Color> sample;
            uint8_t bits = 0;
            for (int k = 0; k < 4; ++k) {
                sample.push_back(world.color_at(ray_for_pixel_offset(x, ox[k], y, oy[k]), max_reflections, &h));
                if (h) bits |= (uint8_t)(1u << k);
            }
            if (hit_bits) *hit_bits = bits;
            Color average = Color::average_over(sample);
            bool any = false;
            for (const Color& c : sample)
                if (c.distance_from(average) > 0.01) any = true;
            if (any) {                                                                 // resample, :84-92
                for (unsigned k = 0; k < antialiasing_samples; ++k) {
                    double xo = resample_offset(x, y, k, 0);   // Rust evaluates arguments left to right:
                    double yo = resample_offset(x, y, k, 1);   // x's rng.gen() before y's
                    sample.push_back(world.color_at(ray_for_pixel_offset(x, xo, y, yo), max_reflections));
                }
                average = Color::average_over(sample);
            }
            return average;
        }
    }
    Canvas render(const World& w) const {                                             // :116-126 (off-by-one kept)
        Canvas canvas(hsize, vsize);
        for (uint32_t y = 0; y + 1 < vsize; ++y)
            for (uint32_t x = 0; x + 1 < hsize; ++x) canvas.write_pixel(x, y, render_pixel(x, y, w));
        return canvas;
    }
};

}  // namespace orc
