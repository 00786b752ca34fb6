// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product path.
//
// CPU restatement (C++17, build with -O2 -ffp-contract=off, no fast-math) of the
// math layer of joedane/raytracer-challenge `ch1`:
//   src/vec.rs:7-216      Vector / Point / Ray
//   src/transform.rs:8-217 Matrix (4x4, cofactor determinant/inverse, fluent builders)
//   src/color.rs:67-139   Color
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
// legs may build, link, load or execute this code.  The reference itself (Rust) cannot
// be compiled in this environment (no cargo/rustc), so there is no oracle/_ref; parity
// is pinned through the reference's own unit-test known answers (oracle/kat.cpp).
#pragma once
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <stdexcept>
#include <string>
#include <vector>

namespace orc {

// src/vec.rs:16
constexpr double EPSILON = 0.00000001;

// src/lib.rs:13-15
inline bool floats_equal(double a, double b) { return std::fabs(a - b) < 0.0001; }

struct Matrix;

// src/vec.rs:7-115
struct Vector {
    double x, y, z;
    Vector() : x(0), y(0), z(0) {}
    Vector(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
    Vector add(const Vector& o) const { return Vector(x + o.x, y + o.y, z + o.z); }       // :25
    Vector sub(const Vector& o) const { return Vector(x - o.x, y - o.y, z - o.z); }       // :33
    Vector negate() const { return Vector(-x, -y, -z); }                                  // :41
    Vector mul(double m) const { return Vector(x * m, y * m, z * m); }                    // :49
    Vector div(double d) const { return Vector(x / d, y / d, z / d); }                    // :57
    double magnitude() const { return std::sqrt(x * x + y * y + z * z); }                 // :65
    Vector normalize() const {                                                            // :69
        double m = magnitude();
        return Vector(x / m, y / m, z / m);
    }
    double dot(const Vector& o) const { return x * o.x + y * o.y + z * o.z; }             // :78
    Vector cross(const Vector& o) const {                                                 // :84
        return Vector(y * o.z - z * o.y, z * o.x - x * o.z, x * o.y - y * o.x);
    }
    inline Vector transform(const Matrix& m) const;                                       // :102
    Vector reflect(const Vector& n) const { return sub(n.mul(2. * dot(n))); }             // :106
    bool approximately_equal(const Vector& v) const {                                     // :110
        return floats_equal(x, v.x) && floats_equal(y, v.y) && floats_equal(z, v.z);
    }
    bool operator==(const Vector& o) const { return x == o.x && y == o.y && z == o.z; }
};

// src/vec.rs:145-190
struct Point {
    double x, y, z;
    Point() : x(0), y(0), z(0) {}
    Point(double x_, double y_, double z_) : x(x_), y(y_), z(z_) {}
    Point displace_by(const Vector& o) const { return Point(x + o.x, y + o.y, z + o.z); } // :162
    Vector sub(const Point& o) const { return Vector(x - o.x, y - o.y, z - o.z); }        // :168
    Point sub_vec(const Vector& o) const { return Point(x - o.x, y - o.y, z - o.z); }     // :174
    inline Point transform(const Matrix& m) const;                                        // :180
    bool approximately_equal(const Point& p) const {                                      // :184
        return floats_equal(x, p.x) && floats_equal(y, p.y) && floats_equal(z, p.z);
    }
    bool operator==(const Point& o) const { return x == o.x && y == o.y && z == o.z; }
};

// A square row-major array of run-time size, standing in for the `array2d` crate
// (storage and indexing only; src/transform.rs:2).
struct Array2D {
    int n;
    std::vector<double> v;
    Array2D(int n_, double fill) : n(n_), v((size_t)n_ * n_, fill) {}
    double& at(int r, int c) { return v[(size_t)r * n + c]; }
    double at(int r, int c) const { return v[(size_t)r * n + c]; }
};

// src/transform.rs:8-21
inline Array2D multiply_arrays(const Array2D& a1, const Array2D& a2) {
    Array2D out(4, 0.0);
    for (int row = 0; row < 4; ++row)
        for (int col = 0; col < 4; ++col) {
            double sum = 0.0;
            for (int i = 0; i < 4; ++i) {
                double v = a1.at(row, i) * a2.at(i, col);
                sum = sum + v;
            }
            out.at(row, col) = sum;
        }
    return out;
}

// src/transform.rs:23-226
struct Matrix {
    Array2D matrix;
    bool has_determinant;
    double determinant_value;

    explicit Matrix(const Array2D& m) : matrix(m) {                                       // :35-38
        double d = determinant(m);
        has_determinant = std::fabs(d) > EPSILON;
        determinant_value = d;
    }
    static Matrix from_row_major(const double* r16) {
        Array2D a(4, 0.0);
        for (int i = 0; i < 16; ++i) a.v[i] = r16[i];
        return Matrix(a);
    }
    static Matrix rows4(double a0, double a1, double a2, double a3, double b0, double b1, double b2,
                        double b3, double c0, double c1, double c2, double c3, double d0, double d1,
                        double d2, double d3) {
        double r[16] = {a0, a1, a2, a3, b0, b1, b2, b3, c0, c1, c2, c3, d0, d1, d2, d3};
        return from_row_major(r);
    }
    Matrix multiply(const Matrix& other) const {                                          // :31
        return Matrix(multiply_arrays(matrix, other.matrix));
    }
    static Matrix identity() {                                                            // :44
        return rows4(1., 0., 0., 0., 0., 1., 0., 0., 0., 0., 1., 0., 0., 0., 0., 1.);
    }
    // Fluent builders PRE-multiply: result = New * self (:59)
    Matrix translation(double x, double y, double z) const {                              // :53
        return rows4(1., 0., 0., x, 0., 1., 0., y, 0., 0., 1., z, 0., 0., 0., 1.).multiply(*this);
    }
    Matrix scaling(double x, double y, double z) const {                                  // :62
        return rows4(x, 0., 0., 0., 0., y, 0., 0., 0., 0., z, 0., 0., 0., 0., 1.).multiply(*this);
    }
    Matrix rotation_x(double r) const {                                                   // :71
        return rows4(1., 0., 0., 0., 0., std::cos(r), -std::sin(r), 0., 0., std::sin(r), std::cos(r),
                     0., 0., 0., 0., 1.)
            .multiply(*this);
    }
    Matrix rotation_y(double r) const {                                                   // :80
        return rows4(std::cos(r), 0., std::sin(r), 0., 0., 1., 0., 0., -std::sin(r), 0., std::cos(r),
                     0., 0., 0., 0., 1.)
            .multiply(*this);
    }
    Matrix rotation_z(double r) const {                                                   // :89
        return rows4(std::cos(r), -std::sin(r), 0., 0., std::sin(r), std::cos(r), 0., 0., 0., 0., 1.,
                     0., 0., 0., 0., 1.)
            .multiply(*this);
    }
    Matrix shearing(double xy, double xz, double yx, double yz, double zx, double zy) const { // :98
        return rows4(1., xy, xz, 0., yx, 1., yz, 0., zx, zy, 1., 0., 0., 0., 0., 1.).multiply(*this);
    }
    Vector transform_vector(const Vector& p) const {                                      // :107-120
        const Array2D& m = matrix;
        return Vector(m.at(0, 0) * p.x + m.at(0, 1) * p.y + m.at(0, 2) * p.z,
                      m.at(1, 0) * p.x + m.at(1, 1) * p.y + m.at(1, 2) * p.z,
                      m.at(2, 0) * p.x + m.at(2, 1) * p.y + m.at(2, 2) * p.z);
    }
    Point transform_point(const Point& p) const {                                         // :122-128
        const Array2D& m = matrix;
        return Point(m.at(0, 0) * p.x + m.at(0, 1) * p.y + m.at(0, 2) * p.z + m.at(0, 3),
                     m.at(1, 0) * p.x + m.at(1, 1) * p.y + m.at(1, 2) * p.z + m.at(1, 3),
                     m.at(2, 0) * p.x + m.at(2, 1) * p.y + m.at(2, 2) * p.z + m.at(2, 3));
    }
    static double determinant(const Array2D& m) {                                         // :130-143
        if (m.n == 2) {
            return m.at(0, 0) * m.at(1, 1) - m.at(0, 1) * m.at(1, 0);
        } else {
            double det = 0.0;
            for (int col = 0; col < m.n; ++col) det += m.at(0, col) * cofactor(m, 0, col);
            return det;
        }
    }
    static Array2D submatrix(const Array2D& m, int i, int j) {                            // :145-157
        if (m.n <= 2) throw std::runtime_error("Matrix is too small");
        int new_size = m.n - 1;
        Array2D out(new_size, 0.0);
        for (int row = 0; row < new_size; ++row)
            for (int col = 0; col < new_size; ++col)
                out.at(row, col) = m.at(row < i ? row : row + 1, col < j ? col : col + 1);
        return out;
    }
    static double minor_(const Array2D& m, int i, int j) {                                // :159
        return determinant(submatrix(m, i, j));
    }
    static double cofactor(const Array2D& m, int i, int j) {                              // :163-169
        double mn = minor_(m, i, j);
        if ((i + j) % 2 == 1) mn = -1.0 * mn;
        return mn;
    }
    bool is_invertable() const { return has_determinant; }                                // :171
    Matrix inverse() const {                                                              // :175-190
        if (!has_determinant) throw std::runtime_error("Matrix is not invertable");
        double d = determinant_value;
        int l = matrix.n;
        Array2D out(l, 0.0);
        for (int row = 0; row < l; ++row)
            for (int col = 0; col < l; ++col) {
                double c = cofactor(matrix, row, col);
                out.at(col, row) = c / d;
            }
        return Matrix(out);
    }
    Matrix transpose() const {                                                            // :192-202
        Array2D out(matrix.n, 0.0);
        for (int row = 0; row < out.n; ++row)
            for (int col = 0; col < out.n; ++col) out.at(row, col) = matrix.at(col, row);
        return Matrix(out);
    }
    static Matrix make_view_transform(const Point& from, const Point& to, const Vector& up) { // :204-217
        Vector forward = to.sub(from).normalize();
        Vector upn = up.normalize();
        Vector left = forward.cross(upn);
        Vector true_up = left.cross(forward);
        Matrix m = rows4(left.x, left.y, left.z, 0., true_up.x, true_up.y, true_up.z, 0., -(forward.x),
                         -(forward.y), -(forward.z), 0., 0., 0., 0., 1.);
        return Matrix(
            multiply_arrays(m.matrix, identity().translation(-from.x, -from.y, -from.z).matrix));
    }
    bool operator==(const Matrix& o) const { return matrix.v == o.matrix.v; }             // :220
    void to_row_major(double* out16) const {
        for (int i = 0; i < 16; ++i) out16[i] = matrix.v[i];
    }
};

inline Vector Vector::transform(const Matrix& m) const { return m.transform_vector(*this); }
inline Point Point::transform(const Matrix& m) const { return m.transform_point(*this); }

// src/vec.rs:192-216
struct Ray {
    Point origin;
    Vector direction;
    Ray(const Point& o, const Vector& d) : origin(o), direction(d) {}
    Point position(double t) const { return origin.displace_by(direction.mul(t)); }       // :207
    Ray transform(const Matrix& m) const {                                                // :211
        return Ray(m.transform_point(origin), m.transform_vector(direction));
    }
};

// src/color.rs:5-140
struct Color {
    double red, green, blue;
    Color() : red(0), green(0), blue(0) {}
    Color(double r, double g, double b) : red(r), green(g), blue(b) {}
    static Color BLACK() { return Color(0.0, 0.0, 0.0); }
    static Color WHITE() { return Color(1.0, 1.0, 1.0); }
    static Color RED() { return Color(1., 0., 0.); }
    Color add(const Color& o) const { return Color(red + o.red, green + o.green, blue + o.blue); } // :67
    Color sub(const Color& o) const { return Color(red - o.red, green - o.green, blue - o.blue); } // :75
    Color mul(const Color& o) const { return Color(red * o.red, green * o.green, blue * o.blue); } // :83
    Color mul_f64(double m) const { return Color(red * m, green * m, blue * m); }                  // :92
    static int32_t clamp(int32_t i, int32_t scale) { return i < 0 ? 0 : (i > scale ? scale : i); } // :100
    // Rust `as i32` from f64: truncation toward zero, saturating, NaN -> 0.              // :112
    static int32_t f64_as_i32(double v) {
        if (std::isnan(v)) return 0;
        if (v >= 2147483647.0) return INT32_MAX;
        if (v <= -2147483648.0) return INT32_MIN;
        return (int32_t)v;
    }
    static int32_t scale(double component, int32_t scale_) {
        return clamp(f64_as_i32(component * (double)scale_), scale_);
    }
    int32_t red_scaled(int32_t s) const { return scale(red, s); }                         // :43
    int32_t green_scaled(int32_t s) const { return scale(green, s); }
    int32_t blue_scaled(int32_t s) const { return scale(blue, s); }
    bool approximately_equal(const Color& o) const {                                      // :116
        return floats_equal(red, o.red) && floats_equal(green, o.green) && floats_equal(blue, o.blue);
    }
    double distance_from(const Color& o) const {                                          // :122
        double dr = red - o.red, dg = green - o.green, db = blue - o.blue;
        return std::sqrt(dr * dr + dg * dg + db * db);
    }
    static Color average_over(const std::vector<Color>& colors) {                         // :128
        double reds = 0., greens = 0., blues = 0.;
        for (size_t i = 0; i < colors.size(); ++i) {
            reds += colors[i].red;
            greens += colors[i].green;
            blues += colors[i].blue;
        }
        double l = (double)colors.size();
        return Color(reds / l, greens / l, blues / l);
    }
    bool operator==(const Color& o) const {
        return red == o.red && green == o.green && blue == o.blue;
    }
    // "#rrggbb" -> Color (src/color.rs:26-41)
    static bool color_from_spec(const char* s, Color* out) {
        if (!s || std::char_traits<char>::length(s) != 7 || s[0] != '#') return false;
        int v[3];
        for (int k = 0; k < 3; ++k) {
            unsigned int x = 0;
            char buf[3] = {s[1 + 2 * k], s[2 + 2 * k], 0};
            if (std::sscanf(buf, "%x", &x) != 1) return false;
            v[k] = (int)x;
        }
        *out = Color((double)v[0] / 255.0, (double)v[1] / 255.0, (double)v[2] / 255.0);
        return true;
    }
};

}  // namespace orc
