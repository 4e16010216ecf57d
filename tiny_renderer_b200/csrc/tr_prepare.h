// tr_prepare.h — K0: per-pass uniform set-up on the host (bit-exact f32, no FMA), uploaded as a kernel
// parameter block.  Replaces default_prepare / shadow_pass_prepare_1 / shadow_pass_prepare_2
// (/root/reference/src/scene/shader.rs:183-279) and hoists the per-fragment constants of the shadow and
// occlusion fragment shaders (shader.rs:763-764, 882-885, 899-900, 920-929).
#pragma once
#include "tr_types.cuh"

namespace tr {

// Cofactor expansion of a 4x4 on its column-major linear array, term order of the widely used Mesa
// gluInvertMatrix (which nalgebra 0.31's 4x4 try_inverse follows); {sign, a, b, c} -> sign * m[a]*m[b]*m[c].
static const signed char kCof[16][6][4] = {
    {{ 1, 5,10,15}, {-1, 5,11,14}, {-1, 9, 6,15}, { 1, 9, 7,14}, { 1,13, 6,11}, {-1,13, 7,10}},
    {{-1, 1,10,15}, { 1, 1,11,14}, { 1, 9, 2,15}, {-1, 9, 3,14}, {-1,13, 2,11}, { 1,13, 3,10}},
    {{ 1, 1, 6,15}, {-1, 1, 7,14}, {-1, 5, 2,15}, { 1, 5, 3,14}, { 1,13, 2, 7}, {-1,13, 3, 6}},
    {{-1, 1, 6,11}, { 1, 1, 7,10}, { 1, 5, 2,11}, {-1, 5, 3,10}, {-1, 9, 2, 7}, { 1, 9, 3, 6}},
    {{-1, 4,10,15}, { 1, 4,11,14}, { 1, 8, 6,15}, {-1, 8, 7,14}, {-1,12, 6,11}, { 1,12, 7,10}},
    {{ 1, 0,10,15}, {-1, 0,11,14}, {-1, 8, 2,15}, { 1, 8, 3,14}, { 1,12, 2,11}, {-1,12, 3,10}},
    {{-1, 0, 6,15}, { 1, 0, 7,14}, { 1, 4, 2,15}, {-1, 4, 3,14}, {-1,12, 2, 7}, { 1,12, 3, 6}},
    {{ 1, 0, 6,11}, {-1, 0, 7,10}, {-1, 4, 2,11}, { 1, 4, 3,10}, { 1, 8, 2, 7}, {-1, 8, 3, 6}},
    {{ 1, 4, 9,15}, {-1, 4,11,13}, {-1, 8, 5,15}, { 1, 8, 7,13}, { 1,12, 5,11}, {-1,12, 7, 9}},
    {{-1, 0, 9,15}, { 1, 0,11,13}, { 1, 8, 1,15}, {-1, 8, 3,13}, {-1,12, 1,11}, { 1,12, 3, 9}},
    {{ 1, 0, 5,15}, {-1, 0, 7,13}, {-1, 4, 1,15}, { 1, 4, 3,13}, { 1,12, 1, 7}, {-1,12, 3, 5}},
    {{-1, 0, 5,11}, { 1, 0, 7, 9}, { 1, 4, 1,11}, {-1, 4, 3, 9}, {-1, 8, 1, 7}, { 1, 8, 3, 5}},
    {{-1, 4, 9,14}, { 1, 4,10,13}, { 1, 8, 5,14}, {-1, 8, 6,13}, {-1,12, 5,10}, { 1,12, 6, 9}},
    {{ 1, 0, 9,14}, {-1, 0,10,13}, {-1, 8, 1,14}, { 1, 8, 2,13}, { 1,12, 1,10}, {-1,12, 2, 9}},
    {{-1, 0, 5,14}, { 1, 0, 6,13}, { 1, 4, 1,14}, {-1, 4, 2,13}, {-1,12, 1, 6}, { 1,12, 2, 5}},
    {{ 1, 0, 5,10}, {-1, 0, 6, 9}, {-1, 4, 1,10}, { 1, 4, 2, 9}, { 1, 8, 1, 6}, {-1, 8, 2, 5}},
};

inline bool invert4(const mat4& A, mat4* out) {
    float m[16], inv[16];
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) m[c * 4 + r] = A.m[r * 4 + c];   // to column-major
    for (int k = 0; k < 16; ++k) {
        float acc = 0.0f;
        for (int t = 0; t < 6; ++t) {
            const signed char* q = kCof[k][t];
            float prod = m[q[1]] * m[q[2]] * m[q[3]];
            if (t == 0) acc = q[0] < 0 ? -prod : prod;
            else acc = q[0] < 0 ? acc - prod : acc + prod;
        }
        inv[k] = acc;
    }
    float det = m[0] * inv[0] + m[1] * inv[4] + m[2] * inv[8] + m[3] * inv[12];
    if (det == 0.0f) return false;
    float inv_det = 1.0f / det;
    for (int r = 0; r < 4; ++r) for (int c = 0; c < 4; ++c) out->m[r * 4 + c] = inv[c * 4 + r] * inv_det;
    return true;
}

// Rotation taking +z onto `to` (nalgebra Rotation3::rotation_between(+z, to), shader.rs:921).
// Returns false where the reference panics (to antiparallel to +z).
inline bool rotation_from_z(f3 to, mat3* R) {
    const mat3 ident = {{1, 0, 0, 0, 1, 0, 0, 0, 1}};
    const f3 zaxis{0.0f, 0.0f, 1.0f};
    float la = length3(zaxis), lb = length3(to);
    if (!(la <= 0.0f) && !(lb <= 0.0f)) {
        f3 a = zaxis / la, b = to / lb;
        f3 c = cross3(a, b);
        float sq = 0.0f + dot3(c, c);
        const float eps = 1.1920929e-7f;
        if (sq > eps * eps) {
            f3 u = c / sqrtf(sq);
            float ang = acosf(dot3(a, b)) * 1.0f;
            if (ang == 0.0f) { *R = ident; return true; }
            float s = sinf(ang), co = cosf(ang), k = 1.0f - co;
            float xx = u.x * u.x, yy = u.y * u.y, zz = u.z * u.z;
            *R = {{xx + (1.0f - xx) * co,      u.x * u.y * k - u.z * s, u.x * u.z * k + u.y * s,
                   u.x * u.y * k + u.z * s,    yy + (1.0f - yy) * co,   u.y * u.z * k - u.x * s,
                   u.x * u.z * k - u.y * s,    u.y * u.z * k + u.x * s, zz + (1.0f - zz) * co}};
            return true;
        }
        if (dot3(a, b) < 0.0f) return false;
    }
    *R = ident;
    return true;
}

enum PrepareError { PREP_OK = 0, PREP_SINGULAR = 1, PREP_LIGHT_W = 2, PREP_ROTATION = 3 };

// default_prepare, shader.rs:183-230.
inline int prepare_camera(Uniforms& u, uint32_t width, uint32_t height, f3 light, f3 from, f3 at, f3 up) {
    f3 bz = unit3(from - at);
    f3 by = unit3(up - bz * dot3(bz, up));
    f3 bx = unit3(cross3(by, bz));
    mat4 model = {{bx.x, bx.y, bx.z, 0.0f, by.x, by.y, by.z, 0.0f, bz.x, bz.y, bz.z, 0.0f, 0.0f, 0.0f, 0.0f, 1.0f}};
    mat4 view = {{1.0f, 0.0f, 0.0f, -from.x, 0.0f, 1.0f, 0.0f, -from.y, 0.0f, 0.0f, 1.0f, -from.z, 0.0f, 0.0f, 0.0f, 1.0f}};
    float coef = -1.0f / 5.0f;
    mat4 proj = {{1.0f, 0.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, 0.0f, 1.0f, 0.0f, 0.0f, 0.0f, coef, 1.0f}};
    float w = (float)(width - 1), h = (float)(height - 1), d = 255.0f;
    mat4 viewport = {{w / 2.0f, 0.0f, 0.0f, w / 2.0f, 0.0f, h / 2.0f, 0.0f, h / 2.0f, 0.0f, 0.0f, d / 2.0f, d / 2.0f, 0.0f, 0.0f, 0.0f, 1.0f}};
    u.vpmv = matmul(matmul(matmul(viewport, proj), model), view);
    u.m = model;
    if (!invert4(transposed(model), &u.it_m)) return PREP_SINGULAR;
    u.camera_direction = bz;
    f4 tl = xform(u.m, f4{light.x, light.y, light.z, 0.0f});
    if (!(tl.w == 0.0f)) return PREP_LIGHT_W;
    u.t_light = unit3(f3{tl.x, tl.y, tl.z});
    u.width = width; u.height = height;
    return PREP_OK;
}

// shadow_pass_prepare_1 (shader.rs:234-255): the camera sits at the light direction; returns the light vpmv.
inline int prepare_light(Uniforms& u, uint32_t width, uint32_t height, f3 light, f3 at, f3 up) {
    int rc = prepare_camera(u, width, height, light, light, at, up);
    u.shadow = u.vpmv;
    return rc;
}

// shadow_pass_prepare_2 (shader.rs:259-279) + the hoisted per-fragment constants.  `u.shadow` must already
// hold the light pass matrix.  want_ao: also build the 16 occlusion sample offsets (shader.rs:882-885, 916-929).
inline int prepare_camera_with_shadow(Uniforms& u, uint32_t width, uint32_t height, f3 light, f3 from, f3 at, f3 up, bool want_ao) {
    int rc = prepare_camera(u, width, height, light, from, at, up);
    if (rc) return rc;
    if (!invert4(u.vpmv, &u.i_vpmv)) return PREP_SINGULAR;
    if (!invert4(u.m, &u.i_m)) return PREP_SINGULAR;
    u.shadow_from_screen = matmul(u.shadow, u.i_vpmv);
    u.ao_w_rowsum = (std::fabs(u.shadow.m[12]) + std::fabs(u.shadow.m[13])) + std::fabs(u.shadow.m[14]);
    u.ao_x_rowsum = (std::fabs(u.shadow.m[0]) + std::fabs(u.shadow.m[1])) + std::fabs(u.shadow.m[2]);
    u.ao_y_rowsum = (std::fabs(u.shadow.m[4]) + std::fabs(u.shadow.m[5])) + std::fabs(u.shadow.m[6]);
    if (want_ao) {
        f4 l4 = xform(u.i_m, f4{u.t_light.x, u.t_light.y, u.t_light.z, 0.0f});
        if (!(l4.w == 0.0f)) return PREP_LIGHT_W;
        mat3 rot;
        if (!rotation_from_z(f3{l4.x, l4.y, l4.z}, &rot)) return PREP_ROTATION;
        const float step = 0.02f;
        const float angle_coef = (2.0f * 3.14159274101257324f) / 16.0f;
        for (int k = 0; k < 16; ++k) {
            float a = angle_coef * (float)k;
            u.ao_step[k] = xform(rot, f3{sinf(a), 0.0f, cosf(a)}) * step;
        }
    }
    return PREP_OK;
}

}  // namespace tr
