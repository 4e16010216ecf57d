// oracle/na32.h — TEST INFRASTRUCTURE ONLY (never linked into the product).
//
// f32 linear algebra restated in the *operation order* of nalgebra 0.31.4 (+ simba 0.7.3), the crate the
// reference (litzendraht/tiny_renderer, Cargo.lock) does all of its arithmetic with.  nalgebra's sources
// are NOT under /root/reference (un-vendored dependency), so the orders below are a restatement of that
// crate's published algorithms:
//   * static-size Matrix*Vector / Matrix*Matrix: column-axpy gemv (base/blas_uninit.rs: gemm_uninit ->
//     gemv_uninit -> axcpy): y = A[:,0]*x0 ; y = A[:,j]*xj + y  for j = 1..
//   * dot: special cases for dim 2/3/4 (base/blas.rs dotx): 2: a+b ; 3: (a+b)+c ; 4: (a+c)+(b+d)
//   * cross (base/ops / matrix.rs): (ay*bz - az*by, az*bx - ax*bz, ax*by - ay*bx)
//   * norm = sqrt(dot(v,v)); normalize = component / norm (true division, `unscale`)
//   * try_inverse: 3x3 by minors with per-entry division by det; 4x4 the Mesa gluInvertMatrix cofactor
//     expansion on the column-major array, then every entry *= 1/det (linalg/inverse.rs)
//   * Rotation3::rotation_between / from_axis_angle (geometry/rotation_specialization.rs)
// PARITY STATUS vs the real Rust binary: UNPINNED (no Rust toolchain, no goldens in the reference).
// Compile with -ffp-contract=off and never with -ffast-math: every * and + below is one IEEE-754 op.
//
// Sensitivity switches (tools/oracle_sensitivity.py only; never defined in the oracle the tests use): because the
// orders above are a restatement, the tool rebuilds the oracle with a different but equally plausible order for one
// primitive at a time and counts how many output pixels move, which bounds what "unpinned" can cost.
//   NA32_ALT_GEMV_PAIRWISE   4-dim mat*vec rows summed (t0 + t2) + (t1 + t3) (the dot4 association) instead of axpy order
//   NA32_ALT_DOT3_RIGHT      3-dim dot as p + (q + r)
//   NA32_ALT_NORMALIZE_RECIP normalize multiplies by 1/norm instead of dividing each component
//   NA32_ALT_INV4_DIVIDE     4x4 inverse divides every cofactor by det instead of multiplying by 1/det
#pragma once
#include <cmath>
#include <cstdint>
#include <limits>

namespace na32 {

struct V2 { float x, y; };
struct V3 { float x, y, z; };
struct V4 { float x, y, z, w; };
// Column-major storage like nalgebra: m[c*R + r].
struct M3 { float m[9];  float& at(int r, int c) { return m[c * 3 + r]; } float at(int r, int c) const { return m[c * 3 + r]; } };
struct M4 { float m[16]; float& at(int r, int c) { return m[c * 4 + r]; } float at(int r, int c) const { return m[c * 4 + r]; } };
struct M23 { float m[6]; float& at(int r, int c) { return m[c * 2 + r]; } float at(int r, int c) const { return m[c * 2 + r]; } };

inline V3 sub(V3 a, V3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
inline V3 add(V3 a, V3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
inline V3 scale(V3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
inline V3 unscale(V3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }

inline float dot(V3 a, V3 b) {
    float p = a.x * b.x, q = a.y * b.y, r = a.z * b.z;
#ifdef NA32_ALT_DOT3_RIGHT
    return p + (q + r);
#else
    return (p + q) + r;
#endif
}
inline float dot(V4 a, V4 b) {
    float p = a.x * b.x, q = a.y * b.y, r = a.z * b.z, s = a.w * b.w;
    p += r; q += s;
    return p + q;
}
inline V3 cross(V3 a, V3 b) {
    return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x};
}
inline float norm_squared(V3 a) { return 0.0f + dot(a, a); }
inline float norm(V3 a) { return std::sqrt(norm_squared(a)); }
#ifdef NA32_ALT_NORMALIZE_RECIP
inline V3 normalize(V3 a) { return scale(a, 1.0f / norm(a)); }
#else
inline V3 normalize(V3 a) { return unscale(a, norm(a)); }
#endif

// gemv, column-axpy order.
inline V4 mul(const M4& a, V4 x) {
    float xs[4] = {x.x, x.y, x.z, x.w};
    float y[4];
#ifdef NA32_ALT_GEMV_PAIRWISE
    for (int i = 0; i < 4; ++i) {
        float t0 = a.at(i, 0) * xs[0], t1 = a.at(i, 1) * xs[1], t2 = a.at(i, 2) * xs[2], t3 = a.at(i, 3) * xs[3];
        y[i] = (t0 + t2) + (t1 + t3);
    }
    return {y[0], y[1], y[2], y[3]};
#endif
    for (int i = 0; i < 4; ++i) y[i] = a.at(i, 0) * xs[0];
    for (int j = 1; j < 4; ++j)
        for (int i = 0; i < 4; ++i) y[i] = a.at(i, j) * xs[j] + y[i];
    return {y[0], y[1], y[2], y[3]};
}
inline V3 mul(const M3& a, V3 x) {
    float xs[3] = {x.x, x.y, x.z};
    float y[3];
    for (int i = 0; i < 3; ++i) y[i] = a.at(i, 0) * xs[0];
    for (int j = 1; j < 3; ++j)
        for (int i = 0; i < 3; ++i) y[i] = a.at(i, j) * xs[j] + y[i];
    return {y[0], y[1], y[2]};
}
inline V2 mul(const M23& a, V3 x) {
    float xs[3] = {x.x, x.y, x.z};
    float y[2];
    for (int i = 0; i < 2; ++i) y[i] = a.at(i, 0) * xs[0];
    for (int j = 1; j < 3; ++j)
        for (int i = 0; i < 2; ++i) y[i] = a.at(i, j) * xs[j] + y[i];
    return {y[0], y[1]};
}
// gemm = one gemv per output column.
inline M4 mul(const M4& a, const M4& b) {
    M4 r;
    for (int j = 0; j < 4; ++j) {
        V4 col = mul(a, V4{b.at(0, j), b.at(1, j), b.at(2, j), b.at(3, j)});
        r.at(0, j) = col.x; r.at(1, j) = col.y; r.at(2, j) = col.z; r.at(3, j) = col.w;
    }
    return r;
}
inline M4 transpose(const M4& a) {
    M4 r;
    for (int i = 0; i < 4; ++i) for (int j = 0; j < 4; ++j) r.at(i, j) = a.at(j, i);
    return r;
}
// Row-major literal, like nalgebra's matrix![] macro.
inline M4 m4_rows(float a00, float a01, float a02, float a03, float a10, float a11, float a12, float a13,
                  float a20, float a21, float a22, float a23, float a30, float a31, float a32, float a33) {
    M4 r;
    r.at(0,0)=a00; r.at(0,1)=a01; r.at(0,2)=a02; r.at(0,3)=a03;
    r.at(1,0)=a10; r.at(1,1)=a11; r.at(1,2)=a12; r.at(1,3)=a13;
    r.at(2,0)=a20; r.at(2,1)=a21; r.at(2,2)=a22; r.at(2,3)=a23;
    r.at(3,0)=a30; r.at(3,1)=a31; r.at(3,2)=a32; r.at(3,3)=a33;
    return r;
}

// Point3::to_homogeneous appends 1, Vector3::to_homogeneous appends 0.
inline V4 point_h(V3 p) { return {p.x, p.y, p.z, 1.0f}; }
inline V4 vector_h(V3 v) { return {v.x, v.y, v.z, 0.0f}; }
// Point3::from_homogeneous: None iff w == 0, else xyz / w (division).
inline bool point_from_h(V4 v, V3* out) {
    if (v.w == 0.0f) return false;
    *out = {v.x / v.w, v.y / v.w, v.z / v.w};
    return true;
}
// Vector3::from_homogeneous: Some(xyz) iff w == 0.
inline bool vector_from_h(V4 v, V3* out) {
    if (!(v.w == 0.0f)) return false;
    *out = {v.x, v.y, v.z};
    return true;
}

// linalg/inverse.rs, dimension 3.
inline bool try_inverse(const M3& a, M3* out) {
    float m11 = a.at(0,0), m12 = a.at(0,1), m13 = a.at(0,2);
    float m21 = a.at(1,0), m22 = a.at(1,1), m23 = a.at(1,2);
    float m31 = a.at(2,0), m32 = a.at(2,1), m33 = a.at(2,2);
    float minor_m12_m23 = m22 * m33 - m32 * m23;
    float minor_m11_m23 = m21 * m33 - m31 * m23;
    float minor_m11_m22 = m21 * m32 - m31 * m22;
    float det = m11 * minor_m12_m23 - m12 * minor_m11_m23 + m13 * minor_m11_m22;
    if (det == 0.0f) return false;
    out->at(0,0) = minor_m12_m23 / det;
    out->at(0,1) = (m13 * m32 - m33 * m12) / det;
    out->at(0,2) = (m12 * m23 - m22 * m13) / det;
    out->at(1,0) = -minor_m11_m23 / det;
    out->at(1,1) = (m11 * m33 - m31 * m13) / det;
    out->at(1,2) = (m13 * m21 - m23 * m11) / det;
    out->at(2,0) = minor_m11_m22 / det;
    out->at(2,1) = (m12 * m31 - m32 * m11) / det;
    out->at(2,2) = (m11 * m22 - m21 * m12) / det;
    return true;
}

// linalg/inverse.rs do_inverse4: Mesa's gluInvertMatrix on the linear (column-major) array.
inline bool try_inverse(const M4& a, M4* out) {
    const float* m = a.m;
    float inv[16];
    inv[0]  =  m[5]*m[10]*m[15] - m[5]*m[11]*m[14] - m[9]*m[6]*m[15] + m[9]*m[7]*m[14] + m[13]*m[6]*m[11] - m[13]*m[7]*m[10];
    inv[4]  = -m[4]*m[10]*m[15] + m[4]*m[11]*m[14] + m[8]*m[6]*m[15] - m[8]*m[7]*m[14] - m[12]*m[6]*m[11] + m[12]*m[7]*m[10];
    inv[8]  =  m[4]*m[9]*m[15]  - m[4]*m[11]*m[13] - m[8]*m[5]*m[15] + m[8]*m[7]*m[13] + m[12]*m[5]*m[11] - m[12]*m[7]*m[9];
    inv[12] = -m[4]*m[9]*m[14]  + m[4]*m[10]*m[13] + m[8]*m[5]*m[14] - m[8]*m[6]*m[13] - m[12]*m[5]*m[10] + m[12]*m[6]*m[9];
    inv[1]  = -m[1]*m[10]*m[15] + m[1]*m[11]*m[14] + m[9]*m[2]*m[15] - m[9]*m[3]*m[14] - m[13]*m[2]*m[11] + m[13]*m[3]*m[10];
    inv[5]  =  m[0]*m[10]*m[15] - m[0]*m[11]*m[14] - m[8]*m[2]*m[15] + m[8]*m[3]*m[14] + m[12]*m[2]*m[11] - m[12]*m[3]*m[10];
    inv[9]  = -m[0]*m[9]*m[15]  + m[0]*m[11]*m[13] + m[8]*m[1]*m[15] - m[8]*m[3]*m[13] - m[12]*m[1]*m[11] + m[12]*m[3]*m[9];
    inv[13] =  m[0]*m[9]*m[14]  - m[0]*m[10]*m[13] - m[8]*m[1]*m[14] + m[8]*m[2]*m[13] + m[12]*m[1]*m[10] - m[12]*m[2]*m[9];
    inv[2]  =  m[1]*m[6]*m[15]  - m[1]*m[7]*m[14]  - m[5]*m[2]*m[15] + m[5]*m[3]*m[14] + m[13]*m[2]*m[7]  - m[13]*m[3]*m[6];
    inv[6]  = -m[0]*m[6]*m[15]  + m[0]*m[7]*m[14]  + m[4]*m[2]*m[15] - m[4]*m[3]*m[14] - m[12]*m[2]*m[7]  + m[12]*m[3]*m[6];
    inv[10] =  m[0]*m[5]*m[15]  - m[0]*m[7]*m[13]  - m[4]*m[1]*m[15] + m[4]*m[3]*m[13] + m[12]*m[1]*m[7]  - m[12]*m[3]*m[5];
    inv[14] = -m[0]*m[5]*m[14]  + m[0]*m[6]*m[13]  + m[4]*m[1]*m[14] - m[4]*m[2]*m[13] - m[12]*m[1]*m[6]  + m[12]*m[2]*m[5];
    inv[3]  = -m[1]*m[6]*m[11]  + m[1]*m[7]*m[10]  + m[5]*m[2]*m[11] - m[5]*m[3]*m[10] - m[9]*m[2]*m[7]   + m[9]*m[3]*m[6];
    inv[7]  =  m[0]*m[6]*m[11]  - m[0]*m[7]*m[10]  - m[4]*m[2]*m[11] + m[4]*m[3]*m[10] + m[8]*m[2]*m[7]   - m[8]*m[3]*m[6];
    inv[11] = -m[0]*m[5]*m[11]  + m[0]*m[7]*m[9]   + m[4]*m[1]*m[11] - m[4]*m[3]*m[9]  - m[8]*m[1]*m[7]   + m[8]*m[3]*m[5];
    inv[15] =  m[0]*m[5]*m[10]  - m[0]*m[6]*m[9]   - m[4]*m[1]*m[10] + m[4]*m[2]*m[9]  + m[8]*m[1]*m[6]   - m[8]*m[2]*m[5];
    float det = m[0]*inv[0] + m[1]*inv[4] + m[2]*inv[8] + m[3]*inv[12];
    if (det == 0.0f) return false;
#ifdef NA32_ALT_INV4_DIVIDE
    for (int k = 0; k < 16; ++k) out->m[k] = inv[k] / det;
    return true;
#endif
    float inv_det = 1.0f / det;
    for (int k = 0; k < 16; ++k) out->m[k] = inv[k] * inv_det;
    return true;
}

// Rotation3::rotation_between(a, b) (scaled_rotation_between with n = 1).  Returns false where nalgebra
// returns None (antiparallel vectors), which the reference unwrap()s into a panic (shader.rs:921).
inline bool rotation_between(V3 a, V3 b, M3* out) {
    M3 id{}; id.at(0,0) = 1.0f; id.at(1,1) = 1.0f; id.at(2,2) = 1.0f;
    float an = norm(a), bn = norm(b);
    if (!(an <= 0.0f) && !(bn <= 0.0f)) {          // try_normalize(0): None iff n <= 0
        V3 na = unscale(a, an), nb = unscale(b, bn);
        V3 c = cross(na, nb);
        float sq = norm_squared(c);
        const float eps = std::numeric_limits<float>::epsilon();
        if (sq > eps * eps) {                        // Unit::try_new(c, eps)
            V3 u = unscale(c, std::sqrt(sq));
            float angle = std::acos(dot(na, nb)) * 1.0f;
            if (angle != 0.0f) {                     // from_axis_angle
                float sqx = u.x * u.x, sqy = u.y * u.y, sqz = u.z * u.z;
                float s = std::sin(angle), co = std::cos(angle);
                float omc = 1.0f - co;
                M3 r;
                r.at(0,0) = sqx + (1.0f - sqx) * co;
                r.at(0,1) = u.x * u.y * omc - u.z * s;
                r.at(0,2) = u.x * u.z * omc + u.y * s;
                r.at(1,0) = u.x * u.y * omc + u.z * s;
                r.at(1,1) = sqy + (1.0f - sqy) * co;
                r.at(1,2) = u.y * u.z * omc - u.x * s;
                r.at(2,0) = u.x * u.z * omc - u.y * s;
                r.at(2,1) = u.y * u.z * omc + u.x * s;
                r.at(2,2) = sqz + (1.0f - sqz) * co;
                *out = r;
            } else {
                *out = id;
            }
            return true;
        }
        if (dot(na, nb) < 0.0f) return false;        // angle PI: axis undefined
    }
    *out = id;
    return true;
}

// Rust scalar cast semantics: `as` from float saturates, truncates toward zero, NaN -> 0.
inline int32_t as_i32(float v) {
    if (v != v) return 0;
    if (v >= 2147483648.0f) return INT32_MAX;
    if (v <= -2147483648.0f) return INT32_MIN;
    return (int32_t)v;
}
inline uint32_t as_u32(float v) {
    if (v != v) return 0;
    if (v <= 0.0f) return 0;
    if (v >= 4294967296.0f) return UINT32_MAX;
    return (uint32_t)v;
}
inline uint8_t as_u8(float v) {
    if (v != v) return 0;
    if (v <= 0.0f) return 0;
    if (v >= 255.0f) return 255;
    return (uint8_t)v;
}

}  // namespace na32
