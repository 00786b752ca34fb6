"""Host-side mirror of the reference's ``Matrix`` (``src/transform.rs:23-226``).

Runs at scene-build time on the host, exactly as in the reference; the GPU path only ever
sees the already-inverted matrices (``Shape::get_transform_inverse``, ``src/shape.rs:30``).
Python floats are IEEE f64 and never contracted into FMAs, so by keeping the reference's
operation order (cofactor expansion along row 0, ``c / d`` per element, row-by-column
products summed left to right) every matrix built here is bit-identical to what the Rust
crate builds.  tests/test_host_mirror.py checks that against the oracle's restatement.
"""
import math

EPSILON = 0.00000001  # src/vec.rs:16


def _determinant(m):
    # src/transform.rs:130-143
    n = len(m)
    if n == 2:
        return m[0][0] * m[1][1] - m[0][1] * m[1][0]
    det = 0.0
    for col in range(n):
        det += m[0][col] * _cofactor(m, 0, col)
    return det


def _submatrix(m, i, j):
    # src/transform.rs:145-157
    n = len(m) - 1
    return [[m[r if r < i else r + 1][c if c < j else c + 1] for c in range(n)] for r in range(n)]


def _cofactor(m, i, j):
    # src/transform.rs:159-169
    minor = _determinant(_submatrix(m, i, j))
    if (i + j) % 2 == 1:
        minor = -1.0 * minor
    return minor


def _multiply(a, b):
    # src/transform.rs:8-21
    out = [[0.0] * 4 for _ in range(4)]
    for row in range(4):
        for col in range(4):
            s = 0.0
            for i in range(4):
                v = a[row][i] * b[i][col]
                s = s + v
            out[row][col] = s
    return out


class Matrix:
    """4x4 row-major f64 matrix with the determinant cached at construction (``:35-38``)."""

    __slots__ = ("m", "determinant")

    def __init__(self, rows):
        self.m = [[float(v) for v in row] for row in rows]
        d = _determinant(self.m)
        self.determinant = d if abs(d) > EPSILON else None

    @staticmethod
    def from_row_major(flat):
        return Matrix([flat[0:4], flat[4:8], flat[8:12], flat[12:16]])

    def to_row_major(self):
        return [v for row in self.m for v in row]

    @staticmethod
    def identity():  # :44
        return Matrix([[1., 0., 0., 0.], [0., 1., 0., 0.], [0., 0., 1., 0.], [0., 0., 0., 1.]])

    def multiply(self, other):  # :31
        return Matrix(_multiply(self.m, other.m))

    # The fluent builders PRE-multiply (New * self, :59) so chains read in application order.
    def translation(self, x, y, z):  # :53
        return Matrix([[1., 0., 0., x], [0., 1., 0., y], [0., 0., 1., z], [0., 0., 0., 1.]]).multiply(self)

    def scaling(self, x, y, z):  # :62
        return Matrix([[x, 0., 0., 0.], [0., y, 0., 0.], [0., 0., z, 0.], [0., 0., 0., 1.]]).multiply(self)

    def rotation_x(self, r):  # :71
        c, s = math.cos(r), math.sin(r)
        return Matrix([[1., 0., 0., 0.], [0., c, -s, 0.], [0., s, c, 0.], [0., 0., 0., 1.]]).multiply(self)

    def rotation_y(self, r):  # :80
        c, s = math.cos(r), math.sin(r)
        return Matrix([[c, 0., s, 0.], [0., 1., 0., 0.], [-s, 0., c, 0.], [0., 0., 0., 1.]]).multiply(self)

    def rotation_z(self, r):  # :89
        c, s = math.cos(r), math.sin(r)
        return Matrix([[c, -s, 0., 0.], [s, c, 0., 0.], [0., 0., 1., 0.], [0., 0., 0., 1.]]).multiply(self)

    def shearing(self, xy, xz, yx, yz, zx, zy):  # :98
        return Matrix([[1., xy, xz, 0.], [yx, 1., yz, 0.], [zx, zy, 1., 0.], [0., 0., 0., 1.]]).multiply(self)

    def transform_vector(self, v):  # :107-120
        m = self.m
        return (m[0][0] * v[0] + m[0][1] * v[1] + m[0][2] * v[2],
                m[1][0] * v[0] + m[1][1] * v[1] + m[1][2] * v[2],
                m[2][0] * v[0] + m[2][1] * v[1] + m[2][2] * v[2])

    def transform_point(self, p):  # :122-128
        m = self.m
        return (m[0][0] * p[0] + m[0][1] * p[1] + m[0][2] * p[2] + m[0][3],
                m[1][0] * p[0] + m[1][1] * p[1] + m[1][2] * p[2] + m[1][3],
                m[2][0] * p[0] + m[2][1] * p[1] + m[2][2] * p[2] + m[2][3])

    def is_invertable(self):  # :171
        return self.determinant is not None

    def inverse(self):  # :175-190
        if self.determinant is None:
            raise ValueError("Matrix is not invertable")  # the reference panics (:177)
        d = self.determinant
        out = [[0.0] * 4 for _ in range(4)]
        for row in range(4):
            for col in range(4):
                out[col][row] = _cofactor(self.m, row, col) / d
        return Matrix(out)

    def transpose(self):  # :192-202
        return Matrix([[self.m[c][r] for c in range(4)] for r in range(4)])

    @staticmethod
    def make_view_transform(frm, to, up):  # :204-217
        forward = _normalize((to[0] - frm[0], to[1] - frm[1], to[2] - frm[2]))
        upn = _normalize(up)
        left = _cross(forward, upn)
        true_up = _cross(left, forward)
        m = [[left[0], left[1], left[2], 0.], [true_up[0], true_up[1], true_up[2], 0.],
             [-(forward[0]), -(forward[1]), -(forward[2]), 0.], [0., 0., 0., 1.]]
        return Matrix(_multiply(m, Matrix.identity().translation(-frm[0], -frm[1], -frm[2]).m))

    def __eq__(self, other):  # :220
        return isinstance(other, Matrix) and self.m == other.m


def _normalize(v):  # src/vec.rs:65-76
    mag = math.sqrt(v[0] * v[0] + v[1] * v[1] + v[2] * v[2])
    return (v[0] / mag, v[1] / mag, v[2] / mag)


def _cross(a, b):  # src/vec.rs:84-90
    return (a[1] * b[2] - a[2] * b[1], a[2] * b[0] - a[0] * b[2], a[0] * b[1] - a[1] * b[0])
