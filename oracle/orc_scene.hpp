// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement of the scene + shading layer of joedane/raytracer-challenge `ch1`,
// keeping the reference's data-structure shape (virtual dispatch per shape, a heap
// vector per per-shape intersection result, insertion-sorted intersection list,
// list-walking compute_refractive):
//   src/material.rs:10-46,48-242,244-369   Light, Pattern(s), Material::lighting
//   src/shape.rs:17-56,58-279,281-630,633-795   Shape, Intersection(s), Sphere/Plane/Cube, World
// See orc_math.hpp for who may use this.
#pragma once
#include <memory>
#include <string>
#include "orc_math.hpp"

namespace orc {

// Per-thread work counters (not in the reference; "ray" = one World::intersect call,
// src/shape.rs:677, callers :703 and :721).
struct Counters {
    uint64_t rays = 0;         // World::intersect invocations
    uint64_t shadow_rays = 0;  // of which from is_shadowed_by_light
    uint64_t shaded_hits = 0;  // shade_hit invocations
    uint64_t sphere_tests = 0;
    uint64_t plane_tests = 0;
    uint64_t cube_tests = 0;
    void add(const Counters& o) {
        rays += o.rays; shadow_rays += o.shadow_rays; shaded_hits += o.shaded_hits;
        sphere_tests += o.sphere_tests; plane_tests += o.plane_tests; cube_tests += o.cube_tests;
    }
};
inline Counters& tls_counters() {
    static thread_local Counters c;
    return c;
}

// src/material.rs:10-31
struct Light {
    Color intensity;
    Point position;
    Light() : intensity(Color::WHITE()), position(Point(-10., 10., -10.)) {}  // Default :26-31
    Light(const Color& i, const Point& p) : intensity(i), position(p) {}
};

struct Shape;

// src/material.rs:33-46
struct Pattern {
    Matrix xf_inv = Matrix::identity();
    virtual ~Pattern() {}
    void set_transform(const Matrix& xf) { xf_inv = xf.inverse(); }
    void set_transform_inverse_raw(const Matrix& inv) { xf_inv = inv; }  // flat-descriptor entry
    const Matrix& get_transform_inverse() const { return xf_inv; }
    virtual Color pattern_at(const Point& point) const = 0;
    inline Color pattern_at_shape(const Shape& shape, const Point& world_point) const;  // :41-45
};

struct TestPattern : Pattern {                                                       // :48-70
    Color pattern_at(const Point& p) const override { return Color(p.x, p.y, p.z); }
};
struct StripePattern : Pattern {                                                     // :72-104
    Color color_a, color_b;
    StripePattern(const Color& a, const Color& b) : color_a(a), color_b(b) {}
    Color pattern_at(const Point& p) const override {
        // Rust `%` on f64 is C fmod (sign of the dividend)
        if (std::fmod(std::floor(p.x), 2.0) == 0.) return color_a;
        return color_b;
    }
};
struct GradientPattern : Pattern {                                                   // :106-136
    Color color_a, color_b, _diff;
    GradientPattern(const Color& a, const Color& b) : color_a(a), color_b(b), _diff(b.sub(a)) {}
    Color pattern_at(const Point& p) const override {
        return color_a.add(_diff.mul_f64(p.x - std::floor(p.x)));
    }
};
struct RingPattern : Pattern {                                                       // :138-171
    Color color_a, color_b;
    RingPattern(const Color& a, const Color& b) : color_a(a), color_b(b) {}
    Color pattern_at(const Point& p) const override {
        if (std::fmod(std::floor(std::sqrt(p.x * p.x + p.y * p.y)), 2.0) == 0.0) return color_a;
        return color_b;
    }
};
struct CheckerPattern : Pattern {                                                    // :173-206
    Color color_a, color_b;
    CheckerPattern(const Color& a, const Color& b) : color_a(a), color_b(b) {}
    Color pattern_at(const Point& p) const override {
        if (std::fmod(std::floor(p.x) + std::floor(p.y) + std::floor(p.z), 2.0) == 0.0) return color_a;
        return color_b;
    }
};
struct GridPattern : Pattern {                                                       // :208-242
    Color color_base, color_grid;
    GridPattern(const Color& base, const Color& grid) : color_base(base), color_grid(grid) {}
    Color pattern_at(const Point& p) const override {
        if (std::fabs(p.x - std::floor(p.x)) < 0.01 || std::fabs(p.z - std::floor(p.z)) < 0.01)
            return color_grid;
        return color_base;
    }
};

// src/material.rs:244-369
struct Material {
    std::shared_ptr<Pattern> pattern;  // Option<Box<dyn Pattern>>
    bool has_color = true;             // Option<Color>
    Color color = Color::RED();
    double ambient = 0.1, diffuse = 0.9, specular = 0.9, shininess = 200.0;
    double reflectiveness = 0.0, transparency = 0.0, refractive_index = 1.0;

    static Material DEFAULT() { return Material(); }                                  // :273-283 (colour RED)
    static Material solid_with_defaults(const Color& c) {                             // :295
        Material m; m.color = c; m.has_color = true; return m;
    }
    static Material default_() { return solid_with_defaults(Color::WHITE()); }        // :364-369
    void set_pattern(std::shared_ptr<Pattern> p) { pattern = std::move(p); }          // :303
    const Pattern& get_pattern() const {                                              // :308
        if (!pattern) throw std::runtime_error("no pattern set");
        return *pattern;
    }
    inline Color lighting(const Light& light, const Shape* object, const Point& point,
                          const Vector& eye, const Vector& normal, bool in_shadow) const;  // :319
};

struct Intersections;

// src/shape.rs:17-49
struct Shape {
    virtual ~Shape() {}
    virtual void set_world_id(uint8_t id) = 0;
    virtual uint8_t get_world_id() const = 0;
    virtual Intersections intersect(const Ray& ray) const;                            // :23-26
    virtual Intersections intersect_local(const Ray& ray) const = 0;
    virtual const Matrix& get_transform_inverse() const = 0;
    virtual const Matrix& get_transform_inverse_transpose() const = 0;
    virtual Vector normal_at(const Point& p) const {                                  // :34-40
        const Matrix& xf = get_transform_inverse();
        Point local_point = p.transform(xf);
        Vector local_normal = normal_at_local(local_point);
        Vector world_normal = local_normal.transform(get_transform_inverse_transpose());
        return world_normal.normalize();
    }
    virtual Vector normal_at_local(const Point& p) const = 0;
    virtual const Material& get_material() const = 0;
    virtual Material& get_material_mut() = 0;
};

inline Color Pattern::pattern_at_shape(const Shape& shape, const Point& world_point) const {
    Point object_point = world_point.transform(shape.get_transform_inverse());
    Point pattern_point = object_point.transform(get_transform_inverse());
    return pattern_at(pattern_point);
}

inline Color Material::lighting(const Light& light, const Shape* object, const Point& point,
                                const Vector& eye, const Vector& normal, bool in_shadow) const {
    Color effective_color;
    if (pattern) {                                                                    // :326-333
        if (!object) throw std::runtime_error("must pass a shape when using a pattern");
        effective_color = pattern->pattern_at_shape(*object, point).mul(light.intensity);
    } else {
        if (!has_color)
            throw std::runtime_error("must specify a color on materials when no pattern is used");
        effective_color = color.mul(light.intensity);
    }
    Vector lightv = light.position.sub(point).normalize();                            // :335
    Color ambient_c = effective_color.mul_f64(ambient);
    if (in_shadow) return ambient_c;
    double light_dot_normal = lightv.dot(normal);
    Color diffuse_c, specular_c;
    if (light_dot_normal < 0.) {
        diffuse_c = Color::BLACK();
        specular_c = Color::BLACK();
    } else {
        diffuse_c = effective_color.mul_f64(diffuse * light_dot_normal);              // :349
        Vector reflectv = lightv.negate().reflect(normal);
        double reflect_dot_eye = reflectv.dot(eye);
        if (reflect_dot_eye <= 0.) {
            specular_c = Color::BLACK();
        } else {
            double factor = std::pow(reflect_dot_eye, shininess);                     // :355
            specular_c = light.intensity.mul_f64(specular * factor);
        }
    }
    return ambient_c.add(diffuse_c).add(specular_c);                                  // :359
}

// src/shape.rs:58-97
struct CachedVectors {
    double t;
    const Shape* object;
    Point point, over_point, under_point;
    Vector eyev, normal, reflectv;
    bool inside;
    double n1, n2;
    CachedVectors(double t_, const Shape* shape, const Ray& ray, const Point& point_,
                  const Vector& eyev_, const Vector& normal_, double n1_, double n2_) {
        inside = normal_.dot(eyev_) < 0.0;                                            // :77
        Vector nrm = inside ? normal_.negate() : normal_;
        over_point = point_.displace_by(nrm.mul(EPSILON));                            // :79
        under_point = point_.sub_vec(nrm.mul(EPSILON));                               // :80
        reflectv = ray.direction.reflect(nrm);                                        // :81
        t = t_; object = shape; point = point_; eyev = eyev_; normal = nrm; n1 = n1_; n2 = n2_;
    }
};

// src/shape.rs:99-164
struct Intersection {
    double t;
    const Shape* shape;
    Intersection(double t_, const Shape* s) : t(t_), shape(s) {}
    bool operator==(const Intersection& o) const { return t == o.t; }                 // :156-164 (t only)
    inline std::pair<double, double> compute_refractive(const Intersections& is) const;  // :115
    inline CachedVectors compute_vectors(const Ray& r, const Intersections& is) const;   // :144
};

// src/shape.rs:167-279
struct Intersections {
    std::vector<Intersection> _i;
    Intersections() {}
    explicit Intersections(const Intersection& i) { _i.push_back(i); }                // :181
    static Intersections new_empty() { return Intersections(); }
    void add(const Intersection& i) { _insert_sorted(i); }                            // :191
    void _insert_sorted(const Intersection& x) {                                      // :195-208
        if (_i.size() == 0) {
            _i.push_back(x);
        } else {
            for (size_t i = 0; i < _i.size(); ++i) {
                if (x.t < _i[i].t) {
                    _i.insert(_i.begin() + (long)i, x);
                    return;
                }
            }
            _i.push_back(x);
        }
    }
    size_t len() const { return _i.size(); }
    void merge(const Intersections& is) {                                             // :214
        for (const Intersection& i : is._i) add(i);
    }
    const Intersection* get_hit() const {                                             // :220-232
        if (len() == 0) return nullptr;
        for (size_t i = 0; i < len(); ++i)
            if (_i[i].t >= 0.0) return &_i[i];
        return nullptr;
    }
    const Intersection& operator[](size_t i) const { return _i[i]; }
};

inline Intersections Shape::intersect(const Ray& ray) const {
    Ray r = ray.transform(get_transform_inverse());
    return intersect_local(r);
}

inline std::pair<double, double> Intersection::compute_refractive(const Intersections& is) const {
    std::vector<const Intersection*> containers;                                      // :116
    double n1 = 1.0, n2 = 1.0;
    for (const Intersection& intersection : is._i) {
        if (intersection == *this) {                                                  // :121
            size_t n = containers.size();
            n1 = n == 0 ? 1.0 : containers[n - 1]->shape->get_material().refractive_index;
        }
        long pos = -1;                                                                // :127
        for (size_t k = 0; k < containers.size(); ++k)
            if (containers[k]->shape->get_world_id() == intersection.shape->get_world_id()) {
                pos = (long)k;
                break;
            }
        if (pos >= 0) containers.erase(containers.begin() + pos);
        else containers.push_back(&intersection);
        if (intersection == *this) {                                                  // :132
            size_t n = containers.size();
            n2 = n == 0 ? 1.0 : containers[n - 1]->shape->get_material().refractive_index;
            return std::make_pair(n1, n2);
        }
    }
    throw std::runtime_error("invalid refraction setup");                             // :140
}

inline CachedVectors Intersection::compute_vectors(const Ray& r, const Intersections& is) const {
    Point p = r.position(t);                                                          // :145
    Vector eyev = r.direction.negate();
    Vector normal = shape->normal_at(p);
    std::pair<double, double> n = compute_refractive(is);
    return CachedVectors(t, shape, r, p, eyev, normal, n.first, n.second);
}

}  // namespace orc
