// ORACLE — TEST INFRASTRUCTURE ONLY.  Not part of the product; nothing under rpt_b200/
// may include, link or call it.  CPU restatement (f64, -ffp-contract=off) of the vector
// maths the reference takes from nalgebra 0.24 / nalgebra-glm 0.10 (Cargo.toml:15-16;
// sources NOT in /root/reference, restated from their published algorithms).
// PARITY UNPINNED at the ulp level: the reference cannot be built here (no rustc) and
// ships no golden vectors for this path.
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace orc {

constexpr double PI = 3.14159265358979323846264338327950288;  // glm::pi::<f64>()
constexpr double INF = std::numeric_limits<double>::infinity();

struct Vec3 {
    double x = 0, y = 0, z = 0;
    Vec3() = default;
    Vec3(double a, double b, double c) : x(a), y(b), z(c) {}
    double operator[](int i) const { return i == 0 ? x : (i == 1 ? y : z); }
    double& operator[](int i) { return i == 0 ? x : (i == 1 ? y : z); }
};

inline Vec3 operator+(const Vec3& a, const Vec3& b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline Vec3 operator-(const Vec3& a, const Vec3& b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec3 operator-(const Vec3& a) { return {-a.x, -a.y, -a.z}; }
inline Vec3 operator*(const Vec3& a, double s) { return {a.x * s, a.y * s, a.z * s}; }
inline Vec3 operator*(double s, const Vec3& a) { return {s * a.x, s * a.y, s * a.z}; }
inline Vec3 operator/(const Vec3& a, double s) { return {a.x / s, a.y / s, a.z / s}; }
inline Vec3& operator+=(Vec3& a, const Vec3& b) { a = a + b; return a; }
inline Vec3& operator*=(Vec3& a, double s) { a = a * s; return a; }
inline Vec3& operator/=(Vec3& a, double s) { a = a / s; return a; }

// nalgebra Matrix::dot for a 3-vector: unrolled a0*b0 + a1*b1 + a2*b2, left to right.
inline double dot(const Vec3& a, const Vec3& b) { return a.x * b.x + a.y * b.y + a.z * b.z; }
inline Vec3 cross(const Vec3& a, const Vec3& b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline Vec3 cmul(const Vec3& a, const Vec3& b) { return {a.x * b.x, a.y * b.y, a.z * b.z}; }  // component_mul
inline double length2(const Vec3& a) { return dot(a, a); }           // glm::length2 / norm_squared
inline double length(const Vec3& a) { return std::sqrt(dot(a, a)); } // glm::length / norm / magnitude
// nalgebra normalize(): self / self.norm()  (component-wise division)
inline Vec3 normalize(const Vec3& a) { return a / length(a); }
// glm::reflect_vec(i, n) = i - n * (n.dot(i) * 2)
inline Vec3 reflect_vec(const Vec3& i, const Vec3& n) { return i - n * (dot(n, i) * 2.0); }
inline Vec3 vmin(const Vec3& a, const Vec3& b) { return {std::fmin(a.x, b.x), std::fmin(a.y, b.y), std::fmin(a.z, b.z)}; }
inline Vec3 vmax(const Vec3& a, const Vec3& b) { return {std::fmax(a.x, b.x), std::fmax(a.y, b.y), std::fmax(a.z, b.z)}; }
inline double clampd(double v, double lo, double hi) { return v < lo ? lo : (v > hi ? hi : v); }  // f64::clamp
inline double signum(double v) { return std::isnan(v) ? v : (std::signbit(v) ? -1.0 : 1.0); }     // f64::signum

struct Mat3 {
    double m[3][3];  // m[row][col]
};
struct Mat4 {
    double m[4][4];
};

// nalgebra matrix * vector (gemv, axpy over columns): y = col0*x0; y += col1*x1; ...
inline Vec3 mul(const Mat3& a, const Vec3& v) {
    Vec3 r;
    r.x = a.m[0][0] * v.x + a.m[0][1] * v.y + a.m[0][2] * v.z;
    r.y = a.m[1][0] * v.x + a.m[1][1] * v.y + a.m[1][2] * v.z;
    r.z = a.m[2][0] * v.x + a.m[2][1] * v.y + a.m[2][2] * v.z;
    return r;
}
// (M * (x,y,z,w)).xyz
inline Vec3 mul_h(const Mat4& a, const Vec3& v, double w) {
    Vec3 r;
    r.x = a.m[0][0] * v.x + a.m[0][1] * v.y + a.m[0][2] * v.z + a.m[0][3] * w;
    r.y = a.m[1][0] * v.x + a.m[1][1] * v.y + a.m[1][2] * v.z + a.m[1][3] * w;
    r.z = a.m[2][0] * v.x + a.m[2][1] * v.y + a.m[2][2] * v.z + a.m[2][3] * w;
    return r;
}

inline Mat3 mat4_to_mat3(const Mat4& a) {
    Mat3 r;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) r.m[i][j] = a.m[i][j];
    return r;
}

// nalgebra 3x3 determinant (minors along the first row)
inline double determinant(const Mat3& a) {
    const double m11 = a.m[0][0], m12 = a.m[0][1], m13 = a.m[0][2];
    const double m21 = a.m[1][0], m22 = a.m[1][1], m23 = a.m[1][2];
    const double m31 = a.m[2][0], m32 = a.m[2][1], m33 = a.m[2][2];
    const double minor_m12_m23 = m22 * m33 - m32 * m23;
    const double minor_m11_m23 = m21 * m33 - m31 * m23;
    const double minor_m11_m22 = m21 * m32 - m31 * m22;
    return m11 * minor_m12_m23 - m12 * minor_m11_m23 + m13 * minor_m11_m22;
}

// glm::inverse_transpose(m3) = m.try_inverse().transpose()  (nalgebra 3x3 closed form)
inline Mat3 inverse_transpose(const Mat3& a) {
    const double m11 = a.m[0][0], m12 = a.m[0][1], m13 = a.m[0][2];
    const double m21 = a.m[1][0], m22 = a.m[1][1], m23 = a.m[1][2];
    const double m31 = a.m[2][0], m32 = a.m[2][1], m33 = a.m[2][2];
    const double minor_m12_m23 = m22 * m33 - m32 * m23;
    const double minor_m11_m23 = m21 * m33 - m31 * m23;
    const double minor_m11_m22 = m21 * m32 - m31 * m22;
    const double det = m11 * minor_m12_m23 - m12 * minor_m11_m23 + m13 * minor_m11_m22;
    Mat3 inv;
    inv.m[0][0] = minor_m12_m23 / det;
    inv.m[0][1] = (m13 * m32 - m33 * m12) / det;
    inv.m[0][2] = (m12 * m23 - m22 * m13) / det;
    inv.m[1][0] = -minor_m11_m23 / det;
    inv.m[1][1] = (m11 * m33 - m31 * m13) / det;
    inv.m[1][2] = (m13 * m21 - m23 * m11) / det;
    inv.m[2][0] = minor_m11_m22 / det;
    inv.m[2][1] = (m12 * m31 - m32 * m11) / det;
    inv.m[2][2] = (m11 * m22 - m21 * m12) / det;
    Mat3 t;
    for (int i = 0; i < 3; i++)
        for (int j = 0; j < 3; j++) t.m[i][j] = inv.m[j][i];
    return t;
}

// glm::inverse(m4): nalgebra do_inverse4 (cofactor expansion on the column-major array)
inline Mat4 inverse(const Mat4& a) {
    double m[16], inv[16];
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++) m[c * 4 + r] = a.m[r][c];
    inv[0] = m[5] * m[10] * m[15] - m[5] * m[11] * m[14] - m[9] * m[6] * m[15] + m[9] * m[7] * m[14] + m[13] * m[6] * m[11] - m[13] * m[7] * m[10];
    inv[4] = -m[4] * m[10] * m[15] + m[4] * m[11] * m[14] + m[8] * m[6] * m[15] - m[8] * m[7] * m[14] - m[12] * m[6] * m[11] + m[12] * m[7] * m[10];
    inv[8] = m[4] * m[9] * m[15] - m[4] * m[11] * m[13] - m[8] * m[5] * m[15] + m[8] * m[7] * m[13] + m[12] * m[5] * m[11] - m[12] * m[7] * m[9];
    inv[12] = -m[4] * m[9] * m[14] + m[4] * m[10] * m[13] + m[8] * m[5] * m[14] - m[8] * m[6] * m[13] - m[12] * m[5] * m[10] + m[12] * m[6] * m[9];
    inv[1] = -m[1] * m[10] * m[15] + m[1] * m[11] * m[14] + m[9] * m[2] * m[15] - m[9] * m[3] * m[14] - m[13] * m[2] * m[11] + m[13] * m[3] * m[10];
    inv[5] = m[0] * m[10] * m[15] - m[0] * m[11] * m[14] - m[8] * m[2] * m[15] + m[8] * m[3] * m[14] + m[12] * m[2] * m[11] - m[12] * m[3] * m[10];
    inv[9] = -m[0] * m[9] * m[15] + m[0] * m[11] * m[13] + m[8] * m[1] * m[15] - m[8] * m[3] * m[13] - m[12] * m[1] * m[11] + m[12] * m[3] * m[9];
    inv[13] = m[0] * m[9] * m[14] - m[0] * m[10] * m[13] - m[8] * m[1] * m[14] + m[8] * m[2] * m[13] + m[12] * m[1] * m[10] - m[12] * m[2] * m[9];
    inv[2] = m[1] * m[6] * m[15] - m[1] * m[7] * m[14] - m[5] * m[2] * m[15] + m[5] * m[3] * m[14] + m[13] * m[2] * m[7] - m[13] * m[3] * m[6];
    inv[6] = -m[0] * m[6] * m[15] + m[0] * m[7] * m[14] + m[4] * m[2] * m[15] - m[4] * m[3] * m[14] - m[12] * m[2] * m[7] + m[12] * m[3] * m[6];
    inv[10] = m[0] * m[5] * m[15] - m[0] * m[7] * m[13] - m[4] * m[1] * m[15] + m[4] * m[3] * m[13] + m[12] * m[1] * m[7] - m[12] * m[3] * m[5];
    inv[14] = -m[0] * m[5] * m[14] + m[0] * m[6] * m[13] + m[4] * m[1] * m[14] - m[4] * m[2] * m[13] - m[12] * m[1] * m[6] + m[12] * m[2] * m[5];
    inv[3] = -m[1] * m[6] * m[11] + m[1] * m[7] * m[10] + m[5] * m[2] * m[11] - m[5] * m[3] * m[10] - m[9] * m[2] * m[7] + m[9] * m[3] * m[6];
    inv[7] = m[0] * m[6] * m[11] - m[0] * m[7] * m[10] - m[4] * m[2] * m[11] + m[4] * m[3] * m[10] + m[8] * m[2] * m[7] - m[8] * m[3] * m[6];
    inv[11] = -m[0] * m[5] * m[11] + m[0] * m[7] * m[9] + m[4] * m[1] * m[11] - m[4] * m[3] * m[9] - m[8] * m[1] * m[7] + m[8] * m[3] * m[5];
    inv[15] = m[0] * m[5] * m[10] - m[0] * m[6] * m[9] - m[4] * m[1] * m[10] + m[4] * m[2] * m[9] + m[8] * m[1] * m[6] - m[8] * m[2] * m[5];
    const double det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    Mat4 out;
    if (det == 0.0) {  // glm::inverse falls back to zeros
        for (int r = 0; r < 4; r++)
            for (int c = 0; c < 4; c++) out.m[r][c] = 0.0;
        return out;
    }
    const double inv_det = 1.0 / det;
    for (int c = 0; c < 4; c++)
        for (int r = 0; r < 4; r++) out.m[r][c] = inv[c * 4 + r] * inv_det;
    return out;
}

// nalgebra Rotation3::from_axis_angle (unit axis), Rodrigues form
inline Mat3 rot_from_axis_angle(const Vec3& u, double angle) {
    Mat3 r;
    if (angle == 0.0) {
        for (int i = 0; i < 3; i++)
            for (int j = 0; j < 3; j++) r.m[i][j] = (i == j) ? 1.0 : 0.0;
        return r;
    }
    const double ux = u.x, uy = u.y, uz = u.z;
    const double sqx = ux * ux, sqy = uy * uy, sqz = uz * uz;
    const double s = std::sin(angle), c = std::cos(angle);
    const double one_m_cos = 1.0 - c;
    r.m[0][0] = sqx + (1.0 - sqx) * c;
    r.m[0][1] = ux * uy * one_m_cos - uz * s;
    r.m[0][2] = ux * uz * one_m_cos + uy * s;
    r.m[1][0] = ux * uy * one_m_cos + uz * s;
    r.m[1][1] = sqy + (1.0 - sqy) * c;
    r.m[1][2] = uy * uz * one_m_cos - ux * s;
    r.m[2][0] = ux * uz * one_m_cos - uy * s;
    r.m[2][1] = uy * uz * one_m_cos + ux * s;
    r.m[2][2] = sqz + (1.0 - sqz) * c;
    return r;
}

// nalgebra Rotation3::rotation_between(a, b) (= scaled_rotation_between(a, b, 1)).
// Returns false for antiparallel inputs (the reference then retries with (0,1,1e-8),
// src/material.rs:186-194).
inline bool rotation_between(const Vec3& a, const Vec3& b, Mat3* out) {
    const double la = length(a), lb = length(b);
    if (la > 0.0 && lb > 0.0) {
        const Vec3 na = a / la, nb = b / lb;
        const Vec3 c = cross(na, nb);
        const double sq = length2(c);
        const double eps = std::numeric_limits<double>::epsilon();
        if (sq > eps * eps) {
            const Vec3 axis = c / std::sqrt(sq);
            *out = rot_from_axis_angle(axis, std::acos(dot(na, nb)) * 1.0);
            return true;
        }
        if (dot(na, nb) < 0.0) return false;
    }
    *out = rot_from_axis_angle(Vec3(1, 0, 0), 0.0);
    return true;
}

struct Quat {
    double w = 1;
    Vec3 v;
};

// glm::quat_rotation(orig, dest) = UnitQuaternion::rotation_between(..).unwrap_or(identity)
inline Quat quat_rotation(const Vec3& a, const Vec3& b) {
    Quat id;
    const double la = length(a), lb = length(b);
    if (!(la > 0.0) || !(lb > 0.0)) return id;
    const Vec3 na = a / la, nb = b / lb;
    const Vec3 c = cross(na, nb);
    const double sq = length2(c);
    const double eps = std::numeric_limits<double>::epsilon();
    if (sq > eps * eps) {
        const Vec3 axis = c / std::sqrt(sq);
        const double cs = dot(na, nb);
        if (cs <= -1.0) return id;  // None -> unwrap_or(identity)
        if (cs >= 1.0) return id;
        const double half = std::acos(cs) * 1.0 / 2.0;
        Quat q;
        q.w = std::cos(half);
        q.v = axis * std::sin(half);
        return q;
    }
    return id;  // parallel (identity) or antiparallel (None -> identity)
}

// glm::quat_rotate_vec3(q, v) = UnitQuaternion * v
inline Vec3 quat_rotate_vec3(const Quat& q, const Vec3& v) {
    const Vec3 t = cross(q.v, v) * 2.0;
    const Vec3 c = cross(q.v, t);
    return t * q.w + c + v;
}

}  // namespace orc
