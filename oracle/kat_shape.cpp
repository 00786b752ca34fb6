// ORACLE — TEST INFRASTRUCTURE ONLY.
// KATs mirroring the reference's shape/world unit tests (src/shape.rs:798-1539).
#include <functional>
#include "orc_world.hpp"

using namespace orc;

void kat_run(const char* name, const char* ref, const std::function<void()>& fn);
void kat_fail(const char* expr, int line);
#define CHECK(cond)                               \
    do {                                          \
        if (!(cond)) kat_fail(#cond, __LINE__);   \
    } while (0)

static bool either(double v, double a, double b) { return floats_equal(v, a) || floats_equal(v, b); }

static World world_with_floor(double reflect, double transp, double ri, bool use_set_transform) {
    // default world + a plane at y=-1 (+ optionally the red ball of test_refract_5/test_schlick_4)
    World w = World::default_();
    if (use_set_transform) {
        auto floor = std::make_unique<Plane>();
        floor->set_transform(Matrix::identity().translation(0., -1., 0.));
        floor->get_material_mut().reflectiveness = reflect;
        floor->get_material_mut().transparency = transp;
        floor->get_material_mut().refractive_index = ri;
        w.add_shape(std::move(floor));
        Material mball = Material::solid_with_defaults(Color(1.0, 0., 0.));
        mball.ambient = 0.5;
        w.add_shape(std::make_unique<Sphere>(Matrix::identity().translation(0., -3.5, -0.5), mball));
    } else {
        Material material = Material::default_();
        material.reflectiveness = reflect;
        w.add_shape(std::make_unique<Plane>(Matrix::identity().translation(0., -1., 0.), material));
    }
    return w;
}

void kats_shape() {
    kat_run("shape::intersect1-5", "src/shape.rs:807-865", [] {
        Sphere c;
        struct T { double oy, oz; int n; double a, b; } t[5] = {{0., -5., 2, 4., 6.}, {1., -5., 2, 5., 5.}, {2., -5., 0, 0., 0.},
                                                                 {0., 0., 2, -1., 1.}, {0., 5., 2, -4., -6.}};
        for (const T& k : t) {
            Intersections v = c.intersect(Ray(Point(0., k.oy, k.oz), Vector(0., 0., 1.)));
            CHECK((int)v.len() == k.n);
            if (k.n == 2) {
                CHECK(either(v[0].t, k.a, k.b));
                CHECK(either(v[1].t, k.a, k.b));
            }
        }
    });
    kat_run("shape::hits1", "src/shape.rs:884-893", [] {
        Sphere c;
        Intersections is(Intersection(1., &c));
        is.add(Intersection(2., &c));
        CHECK(is.get_hit() && is.get_hit()->t == 1.);
    });
    kat_run("shape::hits2", "src/shape.rs:895-904", [] {
        Sphere c;
        Intersections is(Intersection(-1., &c));
        is.add(Intersection(-2., &c));
        CHECK(is.get_hit() == nullptr);
    });
    kat_run("shape::shape_transform1", "src/shape.rs:906-916", [] {
        Sphere c(Matrix::identity().scaling(2., 2., 2.));
        Intersections is = c.intersect(Ray(Point(0., 0., -5.), Vector(0., 0., 1.)));
        CHECK(is.len() == 2);
        CHECK(either(is[0].t, 3., 7.) && either(is[1].t, 3., 7.));
    });
    kat_run("shape::shape_transform2", "src/shape.rs:918-926", [] {
        Sphere c(Matrix::identity().translation(5., 0., 0.));
        CHECK(c.intersect(Ray(Point(0., 0., -5.), Vector(0., 0., 1.))).len() == 0);
    });
    kat_run("shape::circle_normal1-5", "src/shape.rs:928-974", [] {
        Sphere c;
        CHECK(c.normal_at(Point(1., 0., 0.)).approximately_equal(Vector(1., 0., 0.)));
        CHECK(c.normal_at(Point(0., 1., 0.)).approximately_equal(Vector(0., 1., 0.)));
        CHECK(c.normal_at(Point(0., 0., 1.)).approximately_equal(Vector(0., 0., 1.)));
        double n = std::sqrt(3.0) / 3.0;
        Vector p = c.normal_at(Point(n, n, n));
        CHECK(p.approximately_equal(Vector(n, n, n)));
        CHECK(floats_equal(p.magnitude(), 1.));
    });
    kat_run("shape::translate_normal1", "src/shape.rs:976-986", [] {
        Sphere s(Matrix::identity().translation(0., 1., 0.));
        CHECK(s.normal_at(Point(0., 1.70711, -0.70711)).approximately_equal(Vector(0., 0.70711, -0.70711)));
    });
    kat_run("shape::translate_normal2", "src/shape.rs:988-1002", [] {
        const double PI = 3.14159265358979323846264338327950288;
        Sphere s(Matrix::identity().rotation_z(PI / 5.).scaling(1., 0.5, 1.));
        double n = std::sqrt(2.0) / 2.;
        CHECK(s.normal_at(Point(0., n, -n)).approximately_equal(Vector(0., 0.97014, -0.24254)));
    });
    kat_run("shape::world1/world2", "src/shape.rs:1004-1027", [] {
        World world = World::default_();
        Intersections is = world.intersect(Ray(Point(0., 0., -5.), Vector(0., 0., 1.)));
        CHECK(is.len() == 4);
        CHECK(is[0].t == 4. && is[1].t == 4.5 && is[2].t == 5.5 && is[3].t == 6.);
    });
    kat_run("shape::world3 (inside hit)", "src/shape.rs:1029-1039", [] {
        Ray r(Point(0., 0., 0.), Vector(0., 0., 1.));
        Sphere s;
        Intersections is = s.intersect(r);
        const Intersection* ii = nullptr;
        for (const Intersection& a : is._i)
            if (a.t == 1.0) ii = &a;
        CHECK(ii != nullptr);
        CachedVectors pc = ii->compute_vectors(r, is);
        CHECK(pc.inside == true);
        CHECK(pc.point == Point(0., 0., 1.));
        CHECK(pc.eyev == Vector(0., 0., -1.));
    });
    kat_run("shape::world4 (shade_hit outside)", "src/shape.rs:1041-1054", [] {
        World w = World::default_();
        Ray r(Point(0., 0., -5.), Vector(0., 0., 1.));
        Intersections is = w.get_shape(0)->intersect(r);
        const Intersection* ii = nullptr;
        for (const Intersection& a : is._i)
            if (a.t == 4.0) ii = &a;
        CHECK(ii != nullptr);
        Color c = w.shade_hit(ii->compute_vectors(r, is), 1);
        CHECK(floats_equal(c.red, 0.38066) && floats_equal(c.green, 0.47583) && floats_equal(c.blue, 0.2855));
    });
    kat_run("shape::world5 (shade_hit inside)", "src/shape.rs:1056-1071", [] {
        World w = World::default_();
        w.light = Light(Color::WHITE(), Point(0., 0.25, 0.));
        Ray r(Point(0., 0., 0.), Vector(0., 0., 1.));
        Intersections is = w.get_shape(1)->intersect(r);
        const Intersection* ii = nullptr;
        for (const Intersection& a : is._i)
            if (a.t == 0.5) ii = &a;
        CHECK(ii != nullptr);
        Color c = w.shade_hit(ii->compute_vectors(r, is), 1);
        CHECK(floats_equal(c.red, 0.90498) && floats_equal(c.green, 0.90498) && floats_equal(c.blue, 0.90498));
    });
    kat_run("shape::color_at1 (miss)", "src/shape.rs:1073-1082", [] {
        World w = World::default_();
        Color c = w.color_at(Ray(Point(0., 0., -5.), Vector(0., 1., 0.)), 1);
        CHECK(floats_equal(c.red, 0.) && floats_equal(c.green, 0.) && floats_equal(c.blue, 0.));
    });
    kat_run("shape::color_at2 (hit)", "src/shape.rs:1084-1093", [] {
        World w = World::default_();
        Color c = w.color_at(Ray(Point(0., 0., -5.), Vector(0., 0., 1.)), 1);
        CHECK(floats_equal(c.red, 0.38066) && floats_equal(c.green, 0.47583) && floats_equal(c.blue, 0.2855));
    });
    kat_run("shape::color_at3 (inner sphere)", "src/shape.rs:1095-1112", [] {
        World w = World::default_();
        w.get_shape_mut(0)->get_material_mut().ambient = 1.0;
        w.get_shape_mut(1)->get_material_mut().ambient = 1.0;
        Color cmp = w.get_shape(1)->get_material().color;
        Color c = w.color_at(Ray(Point(0., 0., 0.75), Vector(0., 0., -1.)), 1);
        CHECK(floats_equal(c.red, cmp.red) && floats_equal(c.green, cmp.green) && floats_equal(c.blue, cmp.blue));
    });
    kat_run("shape::color_at_jd1 (grazing hit not black)", "src/shape.rs:1114-1130", [] {
        const double PI = 3.14159265358979323846264338327950288;
        World w{Light()};
        auto c = std::make_unique<Sphere>(Matrix::identity().scaling(1.00001, 1.00001, 1.00001));
        c->get_material_mut().color = Color::RED();
        w.add_shape(std::move(c));
        double n = std::sin(PI * 3.0 / 4.0);
        Color col = w.color_at(Ray(Point(0., 0., -2.0 * n), Vector(0., 1., 1.).normalize()), 1);
        CHECK(!(col == Color::BLACK()));
    });
    kat_run("shape::shadow1 (shade_hit in shadow)", "src/shape.rs:1132-1144", [] {
        World w{Light(Color::WHITE(), Point(0., 0., -10.))};
        w.add_shape(std::make_unique<Sphere>());
        w.add_shape(std::make_unique<Sphere>(Matrix::identity().translation(0., 0., 10.)));
        Ray r(Point(0., 0., 5.), Vector(0., 0., 1.));
        Intersections is = w.intersect(r);
        int count = 0;
        const Intersection* ii = nullptr;
        for (const Intersection& a : is._i)
            if (a.t == 4.0) { ii = &a; ++count; }
        CHECK(count == 1);
        Color c = w.shade_hit(ii->compute_vectors(r, is), 1);
        CHECK(c == Color(0.1, 0.1, 0.1));
    });
    kat_run("shape::shadow2 (over_point)", "src/shape.rs:1146-1156", [] {
        Ray r(Point(0., 0., -5.), Vector(0., 0., 1.));
        Sphere shape(Matrix::identity().translation(0., 0., 1.));
        Intersections is = shape.intersect(r);
        const Intersection* ii = nullptr;
        for (const Intersection& a : is._i)
            if (a.t == 5.0) ii = &a;
        CHECK(ii != nullptr);
        CachedVectors v = ii->compute_vectors(r, is);
        CHECK(v.over_point.z < -EPSILON / 2.0);
        CHECK(v.point.z > v.over_point.z);
    });
    kat_run("shape::plane1-5", "src/shape.rs:1158-1203", [] {
        Plane p;
        CHECK(p.normal_at_local(Point(0., 0., 0.)) == Vector(0., 1., 0.));
        CHECK(p.normal_at_local(Point(10., 0., -10.)) == Vector(0., 1., 0.));
        CHECK(p.normal_at_local(Point(-5., 0., 150.)) == Vector(0., 1., 0.));
        CHECK(p.intersect(Ray(Point(0., 10., 0.), Vector(0., 0., 1.))).len() == 0);
        CHECK(p.intersect(Ray(Point(0., 0., 0.), Vector(0., 0., 1.))).len() == 0);
        Intersections a = p.intersect(Ray(Point(0., 1., 0.), Vector(0., -1., 0.)));
        CHECK(a.len() == 1 && a[0].t == 1.0);
        Intersections b = p.intersect(Ray(Point(0., -1., 0.), Vector(0., 1., 0.)));
        CHECK(b.len() == 1 && b[0].t == 1.0);
    });
    kat_run("shape::reflect1 (reflectv)", "src/shape.rs:1205-1213", [] {
        Plane shape;
        double n = std::sqrt(2.0) / 2.0;
        Ray ray(Point(0., 0., -1.), Vector(0., -n, n));
        Intersection i(n, &shape);
        CachedVectors comps = i.compute_vectors(ray, Intersections(i));
        CHECK(comps.reflectv == Vector(0., n, n));
    });
    kat_run("shape::reflect2 (non-reflective)", "src/shape.rs:1215-1229", [] {
        World world = World::default_();
        world.get_shape_mut(1)->get_material_mut().ambient = 1.0;
        Ray ray(Point(0., 0., 0.), Vector(0., 0., 1.));
        Intersection i(1.0, world.get_shape(1));
        CachedVectors comps = i.compute_vectors(ray, Intersections(i));
        CHECK(world.reflected_color(comps, 1) == Color::BLACK());
    });
    kat_run("shape::reflect3 (reflected_color)", "src/shape.rs:1231-1247", [] {
        World world = world_with_floor(0.5, 0.0, 1.0, false);
        double n = std::sqrt(2.0) / 2.0;
        Ray ray(Point(0., 0., -3.), Vector(0., -n, n));
        Intersection i(n * 2.0, world.get_shape(2));
        CachedVectors comps = i.compute_vectors(ray, Intersections(i));
        CHECK(world.reflected_color(comps, 1).approximately_equal(Color(0.19032, 0.2379, 0.14274)));
    });
    kat_run("shape::reflect4 (shade_hit reflective)", "src/shape.rs:1249-1266", [] {
        World world = world_with_floor(0.5, 0.0, 1.0, false);
        double n = std::sqrt(2.0) / 2.0;
        Ray ray(Point(0., 0., -3.), Vector(0., -n, n));
        Intersection i(n * 2.0, world.get_shape(2));
        CachedVectors comps = i.compute_vectors(ray, Intersections(i));
        CHECK(world.shade_hit(comps, 1).approximately_equal(Color(0.87677, 0.92436, 0.82918)));
    });
    kat_run("shape::reflect5 (n1/n2 table)", "src/shape.rs:1268-1304", [] {
        Sphere A = Sphere::glass_sphere();
        A.set_transform(Matrix::identity().scaling(2., 2., 2.));
        A.material.refractive_index = 1.5;
        A.world_id = 1;
        Sphere B = Sphere::glass_sphere();
        B.set_transform(Matrix::identity().translation(0., 0., -0.25));
        B.material.refractive_index = 2.0;
        B.world_id = 2;
        Sphere C = Sphere::glass_sphere();
        C.set_transform(Matrix::identity().translation(0., 0., 0.25));
        C.material.refractive_index = 2.5;
        C.world_id = 3;
        Intersections is(Intersection(2.0, &A));
        is.add(Intersection(2.75, &B));
        is.add(Intersection(3.25, &C));
        is.add(Intersection(4.75, &B));
        is.add(Intersection(5.25, &C));
        is.add(Intersection(6., &A));
        double want[6][2] = {{1.0, 1.5}, {1.5, 2.0}, {2.0, 2.5}, {2.5, 2.5}, {2.5, 1.5}, {1.5, 1.0}};
        Ray ray(Point(0., 0., -4.), Vector(0., 0., 1.));
        for (size_t i = 0; i < is.len(); ++i) {
            CachedVectors comps = is[i].compute_vectors(ray, is);
            CHECK(comps.n1 == want[i][0]);
            CHECK(comps.n2 == want[i][1]);
        }
    });
    kat_run("shape::under_point", "src/shape.rs:1306-1315", [] {
        Ray ray(Point(0., 0., -5.), Vector(0., 0., 1.));
        Sphere shape = Sphere::glass_sphere();
        shape.set_transform(Matrix::identity().translation(0., 0., 1.));
        Intersection i(5., &shape);
        CachedVectors comps = i.compute_vectors(ray, Intersections(i));
        CHECK(comps.under_point.z > EPSILON / 2.0);
        CHECK(comps.point.z < comps.under_point.z);
    });
    kat_run("shape::refract_1 (opaque)", "src/shape.rs:1317-1328", [] {
        World world = World::default_();
        Ray ray(Point(0., 0., -5.), Vector(0., 0., 1.));
        Intersection i(4.0, world.get_shape(0));
        Intersections is(i);
        is.add(Intersection(6.0, world.get_shape(0)));
        CHECK(world.refracted_color(i.compute_vectors(ray, is), 5) == Color::BLACK());
    });
    kat_run("shape::refract_2 (depth 0)", "src/shape.rs:1330-1345", [] {
        World world = World::default_();
        world.get_shape_mut(0)->get_material_mut().transparency = 1.0;
        world.get_shape_mut(0)->get_material_mut().refractive_index = 1.5;
        Ray ray(Point(0., 0., -5.), Vector(0., 0., 1.));
        Intersection i(4.0, world.get_shape(0));
        Intersections is(i);
        is.add(Intersection(6.0, world.get_shape(0)));
        CHECK(world.refracted_color(i.compute_vectors(ray, is), 0) == Color::BLACK());
    });
    kat_run("shape::refract_3 (total internal reflection)", "src/shape.rs:1347-1365", [] {
        World world = World::default_();
        world.get_shape_mut(0)->get_material_mut().transparency = 1.0;
        world.get_shape_mut(0)->get_material_mut().refractive_index = 1.5;
        double n = std::sqrt(2.0) / 2.0;
        Ray ray(Point(0., 0., n), Vector(0., 1., 0.));
        Intersections is(Intersection(-n, world.get_shape(0)));
        Intersection i(n, world.get_shape(0));
        is.add(i);
        CHECK(world.refracted_color(i.compute_vectors(ray, is), 5) == Color::BLACK());
    });
    kat_run("shape::refract_4 (through test pattern)", "src/shape.rs:1367-1395", [] {
        World world = World::default_();
        world.get_shape_mut(0)->get_material_mut().ambient = 1.0;
        world.get_shape_mut(0)->get_material_mut().set_pattern(std::make_shared<TestPattern>());
        world.get_shape_mut(1)->get_material_mut().transparency = 1.0;
        world.get_shape_mut(1)->get_material_mut().refractive_index = 1.5;
        const Shape* A = world.get_shape(0);
        const Shape* B = world.get_shape(1);
        Ray ray(Point(0., 0., 0.1), Vector(0., 1., 0.));
        Intersections is(Intersection(-0.9899, A));
        is.add(Intersection(-0.4899, B));
        is.add(Intersection(0.4899, B));
        is.add(Intersection(0.9899, A));
        Color color = world.refracted_color(is[2].compute_vectors(ray, is), 5);
        CHECK(color.approximately_equal(Color(0., 0.99888, 0.04725)));
    });
    kat_run("shape::refract_5 (shade_hit transparent floor)", "src/shape.rs:1397-1420", [] {
        World world = world_with_floor(0.0, 0.5, 1.5, true);
        double n = std::sqrt(2.0) / 2.0;
        Ray ray(Point(0., 0., -3.0), Vector(0., -n, n));
        Intersections is(Intersection(n * 2.0, world.get_shape(2)));
        Color color = world.shade_hit(is[0].compute_vectors(ray, is), 5);
        CHECK(color.approximately_equal(Color(0.93642, 0.68642, 0.68642)));
    });
    kat_run("shape::schlick_1 (TIR)", "src/shape.rs:1422-1431", [] {
        Sphere shape = Sphere::glass_sphere();
        double n = std::sqrt(2.0) / 2.0;
        Ray ray(Point(0., 0., n), Vector(0., 1., 0.));
        Intersections is(Intersection(-n, &shape));
        is.add(Intersection(n, &shape));
        CHECK(World::reflectance(is[1].compute_vectors(ray, is)) == 1.0);
    });
    kat_run("shape::schlick_2 (perpendicular)", "src/shape.rs:1433-1441", [] {
        Sphere shape = Sphere::glass_sphere();
        Ray ray(Point(0., 0., 0.), Vector(0., 1., 0.));
        Intersections is(Intersection(-1.0, &shape));
        is.add(Intersection(1.0, &shape));
        CHECK(floats_equal(World::reflectance(is[1].compute_vectors(ray, is)), 0.04));
    });
    kat_run("shape::schlick_3 (grazing)", "src/shape.rs:1443-1450", [] {
        Sphere shape = Sphere::glass_sphere();
        Ray ray(Point(0., 0.99, -2.), Vector(0., 0., 1.));
        Intersections is(Intersection(1.8589, &shape));
        CHECK(floats_equal(World::reflectance(is[0].compute_vectors(ray, is)), 0.48873));
    });
    kat_run("shape::schlick_4 (shade_hit reflective+transparent)", "src/shape.rs:1452-1476", [] {
        World world = world_with_floor(0.5, 0.5, 1.5, true);
        double n = std::sqrt(2.0) / 2.0;
        Ray ray(Point(0., 0., -3.0), Vector(0., -n, n));
        Intersections is(Intersection(n * 2.0, world.get_shape(2)));
        Color color = world.shade_hit(is[0].compute_vectors(ray, is), 5);
        CHECK(color.approximately_equal(Color(0.93391, 0.69643, 0.69243)));
    });
    kat_run("shape::cube1 (hits)", "src/shape.rs:1478-1499", [] {
        Cube cube;
        struct T { Point o; Vector d; double t1, t2; } tr[7] = {
            {Point(5., 0.5, 0.), Vector(-1., 0., 0.), 4., 6.}, {Point(-5., 0.5, 0.), Vector(1., 0., 0.), 4., 6.},
            {Point(0.5, 5., 0.), Vector(0., -1., 0.), 4., 6.}, {Point(0.5, -5., 0.), Vector(0., 1., 0.), 4., 6.},
            {Point(0.5, 0., 5.), Vector(0., 0., -1.), 4., 6.}, {Point(0.5, 0., -5.), Vector(0., 0., 1.), 4., 6.},
            {Point(0., 0.5, 0.), Vector(0., 0., 1.), -1., 1.}};
        for (const T& k : tr) {
            Intersections is = cube.intersect(Ray(k.o, k.d));
            CHECK(is.len() == 2);
            if (is.len() == 2) CHECK(floats_equal(is[0].t, k.t1) && floats_equal(is[1].t, k.t2));
        }
    });
    kat_run("shape::cube2 (misses)", "src/shape.rs:1501-1517", [] {
        Cube cube;
        struct T { Point o; Vector d; } tr[6] = {
            {Point(-2., 0., 0.), Vector(0.2673, 0.5345, 0.8018)}, {Point(0., -2., 0.), Vector(0.8018, 0.2673, 0.5345)},
            {Point(0., 0., -2.), Vector(0.5345, 0.8018, 0.2673)}, {Point(2., 0., 2.), Vector(0., 0., -1.0)},
            {Point(0., 2., 2.), Vector(0., -1., 0.)},             {Point(2., 2., 0.), Vector(-1., 0., 0.)}};
        for (const T& k : tr) CHECK(cube.intersect(Ray(k.o, k.d)).len() == 0);
    });
    kat_run("shape::cube_normals", "src/shape.rs:1519-1538", [] {
        Cube cube;
        struct T { Point p; Vector n; } tr[8] = {
            {Point(1., 0.5, -0.8), Vector(1., 0., 0.)},  {Point(-1., -0.2, 0.9), Vector(-1., 0., 0.)},
            {Point(-0.4, 1., -0.1), Vector(0., 1., 0.)}, {Point(0.3, -1., -0.7), Vector(0., -1., 0.)},
            {Point(-0.6, 0.3, 1.), Vector(0., 0., 1.)},  {Point(0.4, 0.4, -1.), Vector(0., 0., -1.)},
            {Point(1., 1., 1.), Vector(1., 0., 0.)},     {Point(-1., -1., -1.), Vector(-1., 0., 0.)}};
        for (const T& k : tr) CHECK(cube.normal_at_local(k.p) == k.n);
    });
}
