// tr_math.cuh — f32 vector/matrix primitives for host (K0 prepare) and device (K1..K6).
//
// Every function is written so that each source-level * / + - is exactly one IEEE-754 binary32 operation
// in the order the reference's arithmetic (nalgebra 0.31.4 behind /root/reference/src/scene/shader.rs)
// performs it: mat*vec accumulates column by column ((a0*x0 + a1*x1) + a2*x2) + a3*x3, dot products use the
// crate's small-dimension forms, normalisation divides by the norm.  The translation unit MUST be compiled
// with nvcc -fmad=false and host -ffp-contract=off; tests/test_fma_canary.py fails otherwise.
// Matrices here are ROW-major float[16] / float[9] (m[r*4+c]).
#pragma once
#include <cmath>
#include <cstdint>

#if defined(__CUDACC__)
#define TR_HD __host__ __device__ __forceinline__
#else
#define TR_HD inline
#endif

namespace tr {

struct f3 { float x, y, z; };
struct f4 { float x, y, z, w; };
struct f2 { float x, y; };
struct mat4 { float m[16]; };
struct mat3 { float m[9]; };

TR_HD f3 operator-(f3 a, f3 b) { return {a.x - b.x, a.y - b.y, a.z - b.z}; }
TR_HD f3 operator+(f3 a, f3 b) { return {a.x + b.x, a.y + b.y, a.z + b.z}; }
TR_HD f3 operator*(f3 a, float s) { return {a.x * s, a.y * s, a.z * s}; }
TR_HD f3 operator/(f3 a, float s) { return {a.x / s, a.y / s, a.z / s}; }

TR_HD float dot3(f3 a, f3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
TR_HD f3 cross3(f3 a, f3 b) { return {a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x}; }
TR_HD float length3(f3 a) { return sqrtf(dot3(a, a)); }
TR_HD f3 unit3(f3 a) { return a / length3(a); }

// y = M * x, accumulated one column of M at a time.
TR_HD f4 xform(const mat4& M, f4 x) {
    const float* m = M.m;
    f4 y{m[0] * x.x, m[4] * x.x, m[8] * x.x, m[12] * x.x};
    y = {m[1] * x.y + y.x, m[5] * x.y + y.y, m[9] * x.y + y.z, m[13] * x.y + y.w};
    y = {m[2] * x.z + y.x, m[6] * x.z + y.y, m[10] * x.z + y.z, m[14] * x.z + y.w};
    y = {m[3] * x.w + y.x, m[7] * x.w + y.y, m[11] * x.w + y.z, m[15] * x.w + y.w};
    return y;
}
TR_HD f3 xform(const mat3& M, f3 x) {
    const float* m = M.m;
    f3 y{m[0] * x.x, m[3] * x.x, m[6] * x.x};
    y = {m[1] * x.y + y.x, m[4] * x.y + y.y, m[7] * x.y + y.z};
    y = {m[2] * x.z + y.x, m[5] * x.z + y.y, m[8] * x.z + y.z};
    return y;
}
TR_HD mat4 matmul(const mat4& A, const mat4& B) {
    mat4 R;
    for (int c = 0; c < 4; ++c) {
        f4 col = xform(A, f4{B.m[c], B.m[4 + c], B.m[8 + c], B.m[12 + c]});
        R.m[c] = col.x; R.m[4 + c] = col.y; R.m[8 + c] = col.z; R.m[12 + c] = col.w;
    }
    return R;
}
TR_HD mat4 transposed(const mat4& A) {
    mat4 R;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) R.m[r * 4 + c] = A.m[c * 4 + r];
    return R;
}

// Homogeneous divide of a point (true division per component); ok=false when w == 0.
TR_HD f3 dehomog(f4 v, bool& ok) {
    ok = !(v.w == 0.0f);
    return {v.x / v.w, v.y / v.w, v.z / v.w};
}

// Rust `as` casts from f32: truncate toward zero, saturate, NaN -> 0.
TR_HD int32_t cast_i32(float v) {
#if defined(__CUDA_ARCH__)
    return __float2int_rz(v);
#else
    if (v != v) return 0;
    if (v >= 2147483648.0f) return INT32_MAX;
    if (v <= -2147483648.0f) return INT32_MIN;
    return (int32_t)v;
#endif
}
TR_HD uint32_t cast_u32(float v) {
#if defined(__CUDA_ARCH__)
    return __float2uint_rz(v);
#else
    if (v != v || v <= 0.0f) return 0;
    if (v >= 4294967296.0f) return UINT32_MAX;
    return (uint32_t)v;
#endif
}
TR_HD uint32_t cast_u8(float v) {
#if defined(__CUDA_ARCH__)
    return min(__float2uint_rz(v), 255u);
#else
    if (v != v || v <= 0.0f) return 0;
    if (v >= 255.0f) return 255;
    return (uint32_t)v;
#endif
}

// Monotone map f32 -> u32 (larger float => larger key) and its inverse.
TR_HD uint32_t float_order(float z) {
    union { float f; uint32_t u; } c; c.f = z;
    return (c.u & 0x80000000u) ? ~c.u : (c.u | 0x80000000u);
}
TR_HD float float_unorder(uint32_t k) {
    union { float f; uint32_t u; } c;
    c.u = (k & 0x80000000u) ? (k & 0x7FFFFFFFu) : ~k;
    return c.f;
}

}  // namespace tr
