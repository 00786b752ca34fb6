// ORACLE — TEST INFRASTRUCTURE ONLY.
//
// Known-answer tests that PIN the oracle: every unit test the reference holds for the
// `Camera::render` path, re-expressed against the C++ restatement
// with the reference's own expected values and tolerances (floats_equal = |a-b| < 1e-4,
// src/lib.rs:13-15, or exact equality where the reference uses assert_eq!).
// Each KAT names the reference test it mirrors (file:line).  Prints one line per test
// and exits non-zero if any fails.
#include <cstring>
#include <functional>
#include <iostream>
#include "orc_world.hpp"

using namespace orc;

static int g_failed = 0, g_run = 0;
static bool g_cur_ok = true;
static const double PI = 3.14159265358979323846264338327950288;  // std::f64::consts::PI

#define CHECK(cond)                                                        \
    do {                                                                   \
        if (!(cond)) {                                                     \
            g_cur_ok = false;                                              \
            std::printf("    check failed: %s (line %d)\n", #cond, __LINE__); \
        }                                                                  \
    } while (0)

static void kat(const char* name, const char* ref, const std::function<void()>& fn) {
    g_cur_ok = true;
    ++g_run;
    try {
        fn();
    } catch (const std::exception& e) {
        g_cur_ok = false;
        std::printf("    exception: %s\n", e.what());
    }
    std::printf("%s %s  [%s]\n", g_cur_ok ? "ok  " : "FAIL", name, ref);
    if (!g_cur_ok) ++g_failed;
}

static std::string f5(double v) {
    char b[64];
    std::snprintf(b, sizeof b, "%.5f", v);
    return b;
}

static Array2D arr(int n, std::initializer_list<double> v) {
    Array2D a(n, 0.0);
    int i = 0;
    for (double d : v) a.v[i++] = d;
    return a;
}

static void kats_vec() {
    kat("vec::negate/add/sub", "src/vec.rs:297-326", [] {
        Vector v(1.0, 2.0, 3.0), v2(10.0, 20.0, 30.0);
        CHECK(v.negate().x == -1.0);
        CHECK(v.add(v2).x == 11.0);
        CHECK(v.sub(v2).x == -9.0);
    });
    kat("vec::magnitude", "src/vec.rs:328-333", [] { CHECK(floats_equal(Vector(2.0, 5.0, 4.0).magnitude(), 6.708204)); });
    kat("vec::dot", "src/vec.rs:335-341", [] {
        Vector a(1.0, 2.0, 3.0), b(2.0, 3.0, 4.0);
        CHECK(floats_equal(a.dot(b), 20.0));
        CHECK(floats_equal(b.dot(a), 20.0));
    });
    kat("vec::cross", "src/vec.rs:343-352", [] {
        Vector a(1.0, 2.0, 3.0), b(2.0, 3.0, 4.0);
        CHECK(floats_equal(a.cross(b).x, -1.0));
        CHECK(floats_equal(b.cross(a).x, 1.0));
    });
    kat("vec::transform_point1", "src/vec.rs:354-362", [] {
        Point p(-4., 6., 8.);
        CHECK(p.transform(Matrix::identity().scaling(2., 3., 4.)) == Point(-8., 18., 32.));
    });
    kat("vec::translate_ray1", "src/vec.rs:364-374", [] {
        Ray r(Point(1., 2., 3.), Vector(0., 1., 0.));
        Ray rt = r.transform(Matrix::identity().translation(3., 4., 5.));
        CHECK(rt.origin == Point(4., 6., 8.));
        CHECK(rt.direction == Vector(0., 1., 0.));
    });
    kat("vec::translate_ray2", "src/vec.rs:376-386", [] {
        Ray r(Point(1., 2., 3.), Vector(0., 1., 0.));
        Ray rt = r.transform(Matrix::identity().scaling(2., 3., 4.));
        CHECK(rt.origin == Point(2., 6., 12.));
        CHECK(rt.direction == Vector(0., 3., 0.));
    });
    kat("vec::reflect1", "src/vec.rs:388-396", [] {
        Vector r = Vector(1., -1., 0.).reflect(Vector(0., 1., 0.));
        CHECK(floats_equal(r.x, 1.) && floats_equal(r.y, 1.) && floats_equal(r.z, 0.));
    });
    kat("vec::reflect2", "src/vec.rs:398-408", [] {
        double f = std::sqrt(2.0) / 2.;
        Vector r = Vector(0., -1., 0.).reflect(Vector(f, f, 0.));
        CHECK(floats_equal(r.x, 1.) && floats_equal(r.y, 0.) && floats_equal(r.z, 0.));
    });
}

static void kats_transform() {
    kat("transform::array_mul", "src/transform.rs:269-293", [] {
        Array2D m1 = arr(4, {1, 2, 3, 4, 5, 6, 7, 8, 9, 8, 7, 6, 5, 4, 3, 2});
        Array2D m2 = arr(4, {-2, 1, 2, 3, 3, 2, 1, -1, 4, 3, 6, 5, 1, 2, 7, 8});
        Array2D m3 = arr(4, {20, 22, 50, 48, 44, 54, 114, 108, 40, 58, 110, 102, 16, 26, 46, 42});
        CHECK(multiply_arrays(m1, m2).v == m3.v);
    });
    kat("transform::determinant1", "src/transform.rs:327-335", [] {
        Array2D m = arr(3, {1.0, 2.0, 6.0, -5.0, 8.0, -4.0, 2.0, 6.0, 4.0});
        CHECK(floats_equal(Matrix::determinant(m), -196.0));
    });
    kat("transform::invert", "src/transform.rs:343-360", [] {
        Matrix m = Matrix::rows4(-5.0, 2.0, 6.0, -8.0, 1.0, -5.0, 1.0, 8.0, 7.0, 7.0, -6.0, -7.0, 1.0, -3.0, 7.0, 4.0);
        const char* want[16] = {"0.21805", "0.45113", "0.24060", "-0.04511", "-0.80827", "-1.45677", "-0.44361", "0.52068",
                                "-0.07895", "-0.22368", "-0.05263", "0.19737", "-0.52256", "-0.81391", "-0.30075", "0.30639"};
        Matrix inv = m.inverse();
        for (int i = 0; i < 16; ++i) CHECK(f5(inv.matrix.v[i]) == want[i]);
    });
    kat("transform::transpose", "src/transform.rs:310-323", [] {
        Matrix m = Matrix::rows4(1, 2, 3, 4, 5, 6, 7, 8, 9, 10, 11, 12, 13, 14, 15, 16);
        Matrix t = Matrix::rows4(1, 5, 9, 13, 2, 6, 10, 14, 3, 7, 11, 15, 4, 8, 12, 16);
        CHECK(m.transpose() == t);
    });
    kat("transform::chained", "src/transform.rs:363-375", [] {
        Matrix m = Matrix::identity().rotation_x(PI / 2.).scaling(5., 5., 5.).translation(10., 5., 7.);
        Point tp = m.transform_point(Point(1., 0., 1.));
        CHECK(floats_equal(tp.x, 15.) && floats_equal(tp.y, 0.) && floats_equal(tp.z, 7.));
    });
    kat("transform::view1", "src/transform.rs:377-383", [] {
        CHECK(Matrix::make_view_transform(Point(0., 0., 0.), Point(0., 0., -1.), Vector(0., 1., 0.)) == Matrix::identity());
    });
    kat("transform::view2", "src/transform.rs:385-391", [] {
        CHECK(Matrix::make_view_transform(Point(0., 0., 0.), Point(0., 0., 1.), Vector(0., 1., 0.)) ==
              Matrix::identity().scaling(-1., 1., -1.));
    });
    kat("transform::view3", "src/transform.rs:393-401", [] {
        CHECK(Matrix::make_view_transform(Point(0., 0., 8.), Point(0., 0., 0.), Vector(0., 1., 0.)) ==
              Matrix::identity().translation(0., 0., -8.));
    });
    kat("transform::view4", "src/transform.rs:403-429", [] {
        Matrix m = Matrix::make_view_transform(Point(1., 3., 2.), Point(4., -2., 8.), Vector(1., 1., 0.));
        const char* want[16] = {"-0.50709", "0.50709", "0.67612", "-2.36643", "0.76772", "0.60609", "0.12122", "-2.82843",
                                "-0.35857", "0.59761", "-0.71714", "0.00000", "0.00000", "0.00000", "0.00000", "1.00000"};
        for (int i = 0; i < 16; ++i) {
            std::string got = f5(m.matrix.v[i]);
            if (got == "-0.00000") got = "0.00000";
            CHECK(got == want[i]);
        }
    });
    kat("transform::not_invertible_below_eps", "src/transform.rs:37,177", [] {
        CHECK(!Matrix::identity().scaling(0.002, 0.002, 0.002).is_invertable());  // det 8e-9 <= 1e-8
        CHECK(Matrix::identity().scaling(0.0022, 0.0022, 0.0022).is_invertable());
    });
}

static void kats_color() {
    kat("color::from_spec", "src/color.rs:147-151", [] {
        Color c;
        CHECK(Color::color_from_spec("#ffffff", &c) && c == Color::WHITE());
        CHECK(Color::color_from_spec("#ff0000", &c) && c == Color::RED());
    });
    kat("color::scale_truncates_and_clamps", "src/color.rs:100-114", [] {
        CHECK(Color::scale(0.999, 255) == 254);
        CHECK(Color::scale(1.0, 255) == 255);
        CHECK(Color::scale(1.9, 255) == 255);
        CHECK(Color::scale(-0.5, 255) == 0);
        CHECK(Color::scale(0.5, 255) == 127);
        CHECK(Color::scale(std::nan(""), 255) == 0);
    });
}

static void kats_camera() {
    kat("camera::pixel_size1", "src/camera.rs:174-178", [] { CHECK(floats_equal(Camera(200, 125, PI / 2.0).pixel_size, 0.01)); });
    kat("camera::pixel_size2", "src/camera.rs:180-185", [] { CHECK(floats_equal(Camera(125, 200, PI / 2.0).pixel_size, 0.01)); });
    kat("camera::camera1", "src/camera.rs:187-194", [] {
        Ray r = Camera(201, 101, PI / 2.0).ray_for_pixel(100, 50);
        CHECK(r.origin.approximately_equal(Point(0., 0., 0.)));
        CHECK(r.direction.approximately_equal(Vector(0., 0., -1.)));
    });
    kat("camera::camera2", "src/camera.rs:196-202", [] {
        Ray r = Camera(201, 101, PI / 2.0).ray_for_pixel(0, 0);
        CHECK(r.origin.approximately_equal(Point(0., 0., 0.)));
        CHECK(r.direction.approximately_equal(Vector(0.66519, 0.33259, -0.66851)));
    });
    kat("camera::camera3", "src/camera.rs:205-214", [] {
        Camera c(201, 101, PI / 2.0, Matrix::identity().translation(0., -2., 5.).rotation_y(PI / 4.0));
        Ray r = c.ray_for_pixel(100, 50);
        double n = std::sqrt(2.0) / 2.0;
        CHECK(r.origin.approximately_equal(Point(0., 2., -5.)));
        CHECK(r.direction.approximately_equal(Vector(n, 0., -n)));
    });
    kat("camera::render1", "src/camera.rs:216-231", [] {
        World world = World::default_();
        Camera camera(11, 11, PI / 2., Matrix::make_view_transform(Point(0., 0., -5.), Point(0., 0., 0.), Vector(0., 1., 0.)));
        Canvas c = camera.render(world);
        Color t = c.get_pixel(5, 5);
        CHECK(floats_equal(t.red, 0.38066) && floats_equal(t.green, 0.47583) && floats_equal(t.blue, 0.2855));
        // Camera::render's loop bounds leave the last row and column black (src/camera.rs:120-121)
        CHECK(c.get_pixel(10, 5) == Color(0., 0., 0.) && c.get_pixel(5, 10) == Color(0., 0., 0.));
    });
}

static void kats_material() {
    struct L { Point light; Vector eye; double want; const char* ref; const char* name; };
    double n = std::sqrt(2.0) / 2.;
    L cases[5] = {{Point(0., 0., -10.), Vector(0., 0., -1.), 1.9, "src/material.rs:379-391", "material::lighting1"},
                  {Point(0., 0., -10.), Vector(0., n, -n), 1.0, "src/material.rs:393-406", "material::lighting2"},
                  {Point(0., 10., -10.), Vector(0., 0., -1.), 0.7364, "src/material.rs:408-419", "material::lighting3"},
                  {Point(0., 10., -10.), Vector(0., -n, -n), 1.6364, "src/material.rs:421-433", "material::lighting4"},
                  {Point(0., 0., 10.), Vector(0., 0., -1.), 0.1, "src/material.rs:435-447", "material::lighting5"}};
    for (const L& c : cases)
        kat(c.name, c.ref, [&] {
            Material m = Material::default_();
            Light light(Color(1., 1., 1.), c.light);
            Color r = m.lighting(light, nullptr, Point(0., 0., 0.), c.eye, Vector(0., 0., -1.), false);
            CHECK(floats_equal(r.red, c.want) && floats_equal(r.green, c.want) && floats_equal(r.blue, c.want));
        });
    kat("material::shadow1", "src/material.rs:449-457", [] {
        Material m = Material::default_();
        Light light(Color::WHITE(), Point(0., 0., -10.));
        Color c = m.lighting(light, nullptr, Point(0., 0., 0.), Vector(0., 0., -1.), Vector(0., 0., -1.), true);
        CHECK(c == Color(0.1, 0.1, 0.1));
    });
    kat("material::shadow2-5", "src/material.rs:459-481", [] {
        World w = World::default_();
        CHECK(w.is_shadowed(Point(0., 10., 0.)) == false);
        CHECK(w.is_shadowed(Point(10., -10., 10.)) == true);
        CHECK(w.is_shadowed(Point(-20., 20., -20.)) == false);
        CHECK(w.is_shadowed(Point(-2., 2., -2.)) == false);
    });
    kat("material::pattern1", "src/material.rs:483-493", [] {
        Material material = Material::default_();
        material.set_pattern(std::make_shared<TestPattern>());
        Sphere shape(Matrix::identity().scaling(2., 2., 2.), material);
        CHECK(shape.get_material().get_pattern().pattern_at_shape(shape, Point(2., 3., 4.)) == Color(1., 1.5, 2.));
    });
    kat("material::pattern2", "src/material.rs:495-505", [] {
        Material material = Material::default_();
        auto p = std::make_shared<TestPattern>();
        p->set_transform(Matrix::identity().scaling(2., 2., 2.));
        material.set_pattern(p);
        Sphere shape;
        shape.material = material;
        CHECK(shape.get_material().get_pattern().pattern_at_shape(shape, Point(2., 3., 4.)) == Color(1., 1.5, 2.));
    });
    kat("material::pattern3", "src/material.rs:507-518", [] {
        Material material = Material::default_();
        auto p = std::make_shared<TestPattern>();
        p->set_transform(Matrix::identity().translation(0.5, 1., 1.5));
        material.set_pattern(p);
        Sphere shape(Matrix::identity().scaling(2., 2., 2.), material);
        CHECK(shape.get_material().get_pattern().pattern_at_shape(shape, Point(2.5, 3., 3.5)) == Color(0.75, 0.5, 0.25));
    });
    kat("material::stripe1-3", "src/material.rs:520-555", [] {
        {
            Material material = Material::default_();
            material.set_pattern(std::make_shared<StripePattern>(Color::WHITE(), Color::BLACK()));
            Sphere shape(Matrix::identity().scaling(2., 2., 2.), material);
            CHECK(shape.get_material().get_pattern().pattern_at_shape(shape, Point(1.5, 0., 0.)) == Color::WHITE());
        }
        {
            Material material = Material::default_();
            auto p = std::make_shared<StripePattern>(Color::WHITE(), Color::BLACK());
            p->set_transform(Matrix::identity().scaling(2., 2., 2.));
            material.set_pattern(p);
            Sphere shape;
            shape.material = material;
            CHECK(shape.get_material().get_pattern().pattern_at_shape(shape, Point(1.5, 0., 0.)) == Color::WHITE());
        }
        {
            Material material = Material::default_();
            auto p = std::make_shared<StripePattern>(Color::WHITE(), Color::BLACK());
            p->set_transform(Matrix::identity().translation(0.5, 0., 0.));
            material.set_pattern(p);
            Sphere shape(Matrix::identity().scaling(2., 2., 2.), material);
            CHECK(shape.get_material().get_pattern().pattern_at_shape(shape, Point(2.5, 0., 0.)) == Color::WHITE());
        }
    });
    kat("material::gradient", "src/material.rs:557-565", [] {
        GradientPattern p(Color::WHITE(), Color::BLACK());
        CHECK(p.pattern_at(Point(0., 0., 0.)) == Color::WHITE());
        CHECK(p.pattern_at(Point(0.25, 0., 0.)) == Color(0.75, 0.75, 0.75));
        CHECK(p.pattern_at(Point(0.5, 0., 0.)) == Color(0.5, 0.5, 0.5));
        CHECK(p.pattern_at(Point(0.75, 0., 0.)) == Color(0.25, 0.25, 0.25));
    });
    kat("material::stripe_negative_odd_floor_is_b", "src/material.rs:97-103 (Rust % = fmod)", [] {
        StripePattern p(Color::WHITE(), Color::BLACK());
        CHECK(p.pattern_at(Point(-0.5, 0., 0.)) == Color::BLACK());  // floor=-1, -1 % 2 = -1 != 0
        CHECK(p.pattern_at(Point(-1.5, 0., 0.)) == Color::WHITE());  // floor=-2
    });
}

void kats_shape();  // kat_shape.cpp

int main() {
    kats_vec();
    kats_transform();
    kats_color();
    kats_camera();
    kats_material();
    kats_shape();
    std::printf("KAT summary: %d run, %d failed\n", g_run, g_failed);
    return g_failed == 0 ? 0 : 1;
}

// shared with kat_shape.cpp
void kat_run(const char* name, const char* ref, const std::function<void()>& fn) { kat(name, ref, fn); }
void kat_fail(const char* expr, int line) {
    g_cur_ok = false;
    std::printf("    check failed: %s (line %d)\n", expr, line);
}
