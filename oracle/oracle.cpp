// oracle/oracle.cpp — TEST INFRASTRUCTURE ONLY.  CPU restatement of the reference render path
// (litzendraht/tiny_renderer: src/scene.rs, src/scene/shader.rs, src/scene/util.rs), single-threaded
// scalar f32 exactly like the reference (its ThreadPool is never used, scene.rs:72-74).
//
// Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load
// this library; the product (tiny_renderer_b200/) never does.
//
// PARITY STATUS: "parity unpinned" against the real Rust binary — the reference ships no tests or
// golden vectors, no Rust toolchain exists on this box, and the arithmetic lives in nalgebra 0.31.4
// which is not vendored (see na32.h).  What IS pinned: asset decode (sha256 KATs vs PIL) and
// GPU<->oracle agreement.  Build: g++ -O2 -ffp-contract=off (never -ffast-math).
#include "na32.h"

#include <algorithm>
#include <cstdio>
#include <cstring>
#include <string>
#include <vector>

using namespace na32;

namespace {

constexpr uint32_t kNoTriangle = 0xFFFFFFFFu;

struct Image {           // image::RgbImage: RGB8, row 0 = top
    std::vector<uint8_t> px;
    uint32_t w = 0, h = 0;
};

struct Model {           // util.rs:16-22
    std::vector<V3> positions;
    std::vector<V2> tex_coords;   // (u, v); the 3rd obj component is ignored (shader.rs:144)
    std::vector<V3> normals;
    std::vector<uint32_t> tris;   // 9 per polygon: p0 t0 n0 p1 t1 n1 p2 t2 n2 (first 3 corners only, scene.rs:224-226)
    Image texture, normal_map, normal_map_tangent, specular_map;
};

struct Diag {
    uint64_t shadow_index_oob = 0;   // shader.rs:774-778, 909-912, 932-935 would panic (index >= len)
    uint64_t shadow_x_wrap = 0;      // round(x) >= width: silently lands on the next row in the reference
    uint64_t texel_oob = 0;          // util.rs get_pixel would panic
    uint64_t nan_depth = 0;
};

struct Buffer {          // shader.rs:14-38
    uint32_t width = 0, height = 0;
    std::vector<float> z_buffer, shadow_buffer;
    V3 camera_direction{};
    V3 t_light_direction{};
    M4 vpmv_matrix{}, i_vpmv_matrix{}, m_matrix{}, i_m_matrix{}, it_m_matrix{}, shadow_matrix{};
    V3 vertex_intensities{};
    M3 vertex_t_positions{};
    M3 vertex_t_normals{};
    M23 vertex_uvs{};
    int32_t raster[2][3] = {};      // vertex_t_raster: [row][col]
    V3 vertex_z_values{};
    uint8_t fragment_color[3] = {};
    Diag diag;
    const char* panic = nullptr;     // set where the reference would panic
};

// ---- util.rs ---------------------------------------------------------------------------------------

// util.rs:7-13
void color_blend(const uint8_t c1[3], const uint8_t c2[3], float t, uint8_t out[3]) {
    for (int k = 0; k < 3; ++k) out[k] = as_u8(t * (float)c1[k] + (1.0f - t) * (float)c2[k]);
}

const uint8_t* get_pixel(Buffer& b, const Image& img, uint32_t x, uint32_t y) {
    if (x >= img.w || y >= img.h) {   // image::get_pixel panics; counted and clamped here
        b.diag.texel_oob++;
        x = std::min(x, img.w - 1);
        y = std::min(y, img.h - 1);
    }
    return &img.px[3 * ((size_t)y * img.w + x)];
}

// util.rs:34-41
const uint8_t* get_color_at_uv(Buffer& b, const Model& m, V2 uv) {
    uint32_t x = as_u32(uv.x * (float)m.texture.w), y = as_u32(uv.y * (float)m.texture.h);
    return get_pixel(b, m.texture, x, y);
}
// util.rs:44-57
V3 get_normal_at_uv(Buffer& b, const Model& m, V2 uv) {
    uint32_t x = as_u32(uv.x * (float)m.normal_map.w), y = as_u32(uv.y * (float)m.normal_map.h);
    const uint8_t* p = get_pixel(b, m.normal_map, x, y);
    return normalize(V3{(float)p[0] / 255.0f - 0.5f, (float)p[1] / 255.0f - 0.5f, (float)p[2] / 255.0f - 0.5f});
}
// util.rs:60-73 — coordinates sized by normal_map, texel read from normal_map_tangent (bug-compatible).
V3 get_normal_tangent_at_uv(Buffer& b, const Model& m, V2 uv) {
    uint32_t x = as_u32(uv.x * (float)m.normal_map.w), y = as_u32(uv.y * (float)m.normal_map.h);
    const uint8_t* p = get_pixel(b, m.normal_map_tangent, x, y);
    return normalize(V3{(float)p[0] / 255.0f - 0.5f, (float)p[1] / 255.0f - 0.5f, (float)p[2] / 255.0f - 0.5f});
}
// util.rs:76-83 — raw channel 0 in [0,255].
float get_specular_value_at_uv(Buffer& b, const Model& m, V2 uv) {
    uint32_t x = as_u32(uv.x * (float)m.specular_map.w), y = as_u32(uv.y * (float)m.specular_map.h);
    return (float)get_pixel(b, m.specular_map, x, y)[0];
}

// ---- shader.rs helpers ------------------------------------------------------------------------------

struct Idx3 { uint32_t a, b, c; uint32_t operator[](int i) const { return i == 0 ? a : (i == 1 ? b : c); } };

// shader.rs:116-124
bool should_cull_face(const V3 p[3], V3 camera_direction) {
    V3 face_normal = cross(sub(p[1], p[0]), sub(p[2], p[0]));
    return dot(camera_direction, face_normal) <= 0.0f;
}
// shader.rs:127-133
void get_vertex_positions(const Model& m, Idx3 idx, V3 out[3]) {
    for (int i = 0; i < 3; ++i) out[i] = m.positions[idx[i]];
}
// shader.rs:136-147
void store_vertex_uvs(M23& uvs, const Model& m, Idx3 idx) {
    for (int i = 0; i < 3; ++i) {
        uvs.at(0, i) = m.tex_coords[idx[i]].x;
        uvs.at(1, i) = 1.0f - m.tex_coords[idx[i]].y;
    }
}
// shader.rs:150-165
void store_vertex_transformation_results(Buffer& b, const V3 p[3], const M4& mat) {
    float z[3];
    for (int i = 0; i < 3; ++i) {
        V3 t;
        if (!point_from_h(mul(mat, point_h(p[i])), &t)) { b.panic = "from_homogeneous: w == 0 (shader.rs:158)"; t = {0, 0, 0}; }
        b.raster[0][i] = as_i32(t.x);
        b.raster[1][i] = as_i32(t.y);
        z[i] = t.z;
    }
    b.vertex_z_values = {z[0], z[1], z[2]};
}
// shader.rs:169-180
bool process_z_value(Buffer& b, V3 bar, uint32_t x, uint32_t y) {
    size_t index = (size_t)x + (size_t)(y * b.width);
    float z = dot(bar, b.vertex_z_values);
    if (z != z) b.diag.nan_depth++;
    if (z <= b.z_buffer[index]) return false;
    b.z_buffer[index] = z;
    return true;
}
V3 transform_normal(Buffer& b, V3 n) {   // Vector3::from_homogeneous(it_m * n.to_homogeneous()).unwrap().normalize()
    V3 t;
    if (!vector_from_h(mul(b.it_m_matrix, vector_h(n)), &t)) { b.panic = "Vector3::from_homogeneous: w != 0"; t = {0, 0, 1}; }
    return normalize(t);
}

// ---- prepares (shader.rs:183-279) --------------------------------------------------------------------

void default_prepare(Buffer& b, uint32_t width, uint32_t height, V3 light, V3 look_from, V3 look_at, V3 up) {
    V3 new_z = normalize(sub(look_from, look_at));
    V3 new_y = normalize(sub(up, scale(new_z, dot(new_z, up))));
    V3 new_x = normalize(cross(new_y, new_z));
    M4 model = m4_rows(new_x.x, new_x.y, new_x.z, 0.0f,
                       new_y.x, new_y.y, new_y.z, 0.0f,
                       new_z.x, new_z.y, new_z.z, 0.0f,
                       0.0f, 0.0f, 0.0f, 1.0f);
    M4 view = m4_rows(1.0f, 0.0f, 0.0f, -look_from.x,
                      0.0f, 1.0f, 0.0f, -look_from.y,
                      0.0f, 0.0f, 1.0f, -look_from.z,
                      0.0f, 0.0f, 0.0f, 1.0f);
    float coef = -1.0f / 5.0f;
    M4 projection = m4_rows(1.0f, 0.0f, 0.0f, 0.0f,
                            0.0f, 1.0f, 0.0f, 0.0f,
                            0.0f, 0.0f, 1.0f, 0.0f,
                            0.0f, 0.0f, coef, 1.0f);
    float w = (float)(width - 1), h = (float)(height - 1), d = 255.0f;
    M4 viewport = m4_rows(w / 2.0f, 0.0f, 0.0f, w / 2.0f,
                          0.0f, h / 2.0f, 0.0f, h / 2.0f,
                          0.0f, 0.0f, d / 2.0f, d / 2.0f,
                          0.0f, 0.0f, 0.0f, 1.0f);
    b.vpmv_matrix = mul(mul(mul(viewport, projection), model), view);   // left-assoc, shader.rs:221
    b.m_matrix = model;
    if (!try_inverse(transpose(model), &b.it_m_matrix)) b.panic = "it_m: singular (shader.rs:224)";
    b.camera_direction = new_z;
    V3 tl;
    if (!vector_from_h(mul(b.m_matrix, vector_h(light)), &tl)) { b.panic = "t_light: w != 0 (shader.rs:227)"; tl = {0, 0, 1}; }
    b.t_light_direction = normalize(tl);
}
void shadow_pass_prepare_1(Buffer& b, uint32_t w, uint32_t h, V3 light, V3, V3 look_at, V3 up) {
    default_prepare(b, w, h, light, light, look_at, up);
    b.shadow_matrix = b.vpmv_matrix;
}
void shadow_pass_prepare_2(Buffer& b, uint32_t w, uint32_t h, V3 light, V3 look_from, V3 look_at, V3 up) {
    default_prepare(b, w, h, light, look_from, look_at, up);
    if (!try_inverse(b.vpmv_matrix, &b.i_vpmv_matrix)) b.panic = "i_vpmv: singular (shader.rs:277)";
    if (!try_inverse(b.m_matrix, &b.i_m_matrix)) b.panic = "i_m: singular (shader.rs:278)";
}

typedef void (*PrepareFn)(Buffer&, uint32_t, uint32_t, V3, V3, V3, V3);
typedef bool (*VertexFn)(Buffer&, const Model&, Idx3, Idx3, Idx3);
typedef bool (*FragmentFn)(Buffer&, const Model&, uint32_t, uint32_t, V3);
struct ShaderPass { PrepareFn prepare; VertexFn vertex; FragmentFn fragment; };   // shader.rs:84-88

const uint8_t kBlack[3] = {0, 0, 0};

// ---- vertex shaders -----------------------------------------------------------------------------------

// default: shader.rs:285-316
bool vs_default(Buffer& b, const Model& m, Idx3 pi, Idx3 ti, Idx3) {
    V3 p[3]; get_vertex_positions(m, pi, p);
    if (should_cull_face(p, b.camera_direction)) return false;
    V3 face_normal = cross(sub(p[1], p[0]), sub(p[2], p[0]));
    float diff = dot(b.t_light_direction, transform_normal(b, face_normal));
    b.vertex_intensities = {diff, diff, diff};
    store_vertex_transformation_results(b, p, b.vpmv_matrix);
    store_vertex_uvs(b.vertex_uvs, m, ti);
    return true;
}
// phong: shader.rs:349-384 ; shadow pass 2: shader.rs:711-747
bool vs_phong(Buffer& b, const Model& m, Idx3 pi, Idx3 ti, Idx3 ni) {
    V3 p[3]; get_vertex_positions(m, pi, p);
    if (should_cull_face(p, b.camera_direction)) return false;
    float in[3];
    for (int i = 0; i < 3; ++i) in[i] = dot(b.t_light_direction, transform_normal(b, m.normals[ni[i]]));
    b.vertex_intensities = {in[0], in[1], in[2]};
    store_vertex_transformation_results(b, p, b.vpmv_matrix);
    store_vertex_uvs(b.vertex_uvs, m, ti);
    return true;
}
// normal_map / specular / occlusion pass 2: shader.rs:416-437, 475-496, 849-870
bool vs_plain(Buffer& b, const Model& m, Idx3 pi, Idx3 ti, Idx3) {
    V3 p[3]; get_vertex_positions(m, pi, p);
    if (should_cull_face(p, b.camera_direction)) return false;
    store_vertex_transformation_results(b, p, b.vpmv_matrix);
    store_vertex_uvs(b.vertex_uvs, m, ti);
    return true;
}
// darboux: shader.rs:549-595
bool vs_darboux(Buffer& b, const Model& m, Idx3 pi, Idx3 ti, Idx3 ni) {
    V3 p[3]; get_vertex_positions(m, pi, p);
    if (should_cull_face(p, b.camera_direction)) return false;
    for (int i = 0; i < 3; ++i) {
        V3 t;
        if (!point_from_h(mul(b.m_matrix, point_h(p[i])), &t)) { b.panic = "m*p: w == 0"; t = {0, 0, 0}; }
        b.vertex_t_positions.at(0, i) = t.x; b.vertex_t_positions.at(1, i) = t.y; b.vertex_t_positions.at(2, i) = t.z;
    }
    for (int i = 0; i < 3; ++i) {
        V3 n = transform_normal(b, m.normals[ni[i]]);
        b.vertex_t_normals.at(0, i) = n.x; b.vertex_t_normals.at(1, i) = n.y; b.vertex_t_normals.at(2, i) = n.z;
    }
    store_vertex_transformation_results(b, p, b.vpmv_matrix);
    store_vertex_uvs(b.vertex_uvs, m, ti);
    return true;
}
// shadow / occlusion pass 1: shader.rs:671-692, 809-830 (no cull, light-space transform)
bool vs_light(Buffer& b, const Model& m, Idx3 pi, Idx3 ti, Idx3) {
    V3 p[3]; get_vertex_positions(m, pi, p);
    store_vertex_transformation_results(b, p, b.shadow_matrix);
    store_vertex_uvs(b.vertex_uvs, m, ti);
    return true;
}

// ---- fragment shaders -----------------------------------------------------------------------------------

void set_color(Buffer& b, const uint8_t c[3]) { b.fragment_color[0] = c[0]; b.fragment_color[1] = c[1]; b.fragment_color[2] = c[2]; }

// default: shader.rs:318-333
bool fs_default(Buffer& b, const Model& m, uint32_t x, uint32_t y, V3 bar) {
    if (!process_z_value(b, bar, x, y)) return false;
    V2 uv = mul(b.vertex_uvs, bar);
    const uint8_t* color = get_color_at_uv(b, m, uv);
    uint8_t out[3]; color_blend(color, kBlack, b.vertex_intensities.x, out); set_color(b, out);
    return true;
}
// phong: shader.rs:386-401
bool fs_phong(Buffer& b, const Model& m, uint32_t x, uint32_t y, V3 bar) {
    if (!process_z_value(b, bar, x, y)) return false;
    V2 uv = mul(b.vertex_uvs, bar);
    const uint8_t* color = get_color_at_uv(b, m, uv);
    float diff = dot(bar, b.vertex_intensities);
    uint8_t out[3]; color_blend(color, kBlack, diff, out); set_color(b, out);
    return true;
}
// normal_map: shader.rs:439-459
bool fs_normal_map(Buffer& b, const Model& m, uint32_t x, uint32_t y, V3 bar) {
    if (!process_z_value(b, bar, x, y)) return false;
    V2 uv = mul(b.vertex_uvs, bar);
    const uint8_t* color = get_color_at_uv(b, m, uv);
    V3 n = transform_normal(b, get_normal_at_uv(b, m, uv));
    float diff = dot(b.t_light_direction, n);
    uint8_t out[3]; color_blend(color, kBlack, diff, out); set_color(b, out);
    return true;
}
// specular: shader.rs:498-534
bool fs_specular(Buffer& b, const Model& m, uint32_t x, uint32_t y, V3 bar) {
    if (!process_z_value(b, bar, x, y)) return false;
    V2 uv = mul(b.vertex_uvs, bar);
    const uint8_t* color = get_color_at_uv(b, m, uv);
    V3 n = transform_normal(b, get_normal_at_uv(b, m, uv));
    V3 l = b.t_light_direction;
    V3 r = normalize(sub(scale(scale(n, dot(l, n)), 2.0f), l));
    float diff = dot(l, n);
    float rz = r.z;
    float base = (rz != rz) ? 0.0f : (rz > 0.0f ? rz : 0.0f);   // f32::max(NaN, 0) = 0
    float spec = 0.6f * std::pow(base, get_specular_value_at_uv(b, m, uv));
    uint8_t out[3];
    for (int k = 0; k < 3; ++k) {
        float v = (diff + spec) * (float)color[k];
        float mn = (v != v) ? 255.0f : (v < 255.0f ? v : 255.0f);   // f32::min(NaN, 255) = 255
        out[k] = as_u8(mn);
    }
    set_color(b, out);
    return true;
}
// darboux: shader.rs:597-655
bool fs_darboux(Buffer& b, const Model& m, uint32_t x, uint32_t y, V3 bar) {
    if (!process_z_value(b, bar, x, y)) return false;
    V2 uv = mul(b.vertex_uvs, bar);
    const uint8_t* color = get_color_at_uv(b, m, uv);
    V3 nt = get_normal_tangent_at_uv(b, m, uv);
    V3 local_z = mul(b.vertex_t_normals, bar);
    V3 r0 = normalize(mul(b.vertex_t_positions, V3{-1.0f, 1.0f, 0.0f}));
    V3 r1 = normalize(mul(b.vertex_t_positions, V3{-1.0f, 0.0f, 1.0f}));
    V3 r2 = normalize(mul(b.vertex_t_normals, bar));
    M3 a;
    a.at(0,0) = r0.x; a.at(0,1) = r0.y; a.at(0,2) = r0.z;
    a.at(1,0) = r1.x; a.at(1,1) = r1.y; a.at(1,2) = r1.z;
    a.at(2,0) = r2.x; a.at(2,1) = r2.y; a.at(2,2) = r2.z;
    M3 ia;
    if (!try_inverse(a, &ia)) { b.panic = "darboux: singular local basis (shader.rs:631)"; return true; }
    const M23& u = b.vertex_uvs;
    V3 lx = mul(ia, V3{u.at(0,1) - u.at(0,0), u.at(0,2) - u.at(0,0), 0.0f});
    V3 ly = mul(ia, V3{u.at(1,1) - u.at(1,0), u.at(1,2) - u.at(1,0), 0.0f});
    V3 c0 = normalize(lx), c1 = normalize(ly), c2 = normalize(local_z);
    M3 t;
    t.at(0,0) = c0.x; t.at(1,0) = c0.y; t.at(2,0) = c0.z;
    t.at(0,1) = c1.x; t.at(1,1) = c1.y; t.at(2,1) = c1.z;
    t.at(0,2) = c2.x; t.at(1,2) = c2.y; t.at(2,2) = c2.z;
    V3 n = normalize(mul(t, nt));
    float diff = dot(b.t_light_direction, n);
    uint8_t out[3]; color_blend(color, kBlack, diff, out); set_color(b, out);
    return true;
}
// shadow / occlusion pass 1: shader.rs:694-709, 832-847
bool fs_light(Buffer& b, const Model&, uint32_t x, uint32_t y, V3 bar) {
    size_t index = (size_t)(x + y * b.width);
    float z = dot(bar, b.vertex_z_values);
    if (z >= b.shadow_buffer[index]) b.shadow_buffer[index] = z;
    return false;
}
// Shadow-buffer lookup: `(x.round() as u32 + (y.round() as u32) * width) as usize` then indexing
// (shader.rs:774-778).  The reference panics when the index is out of range and silently reads the next
// row when only x overflows; both are counted.  Out-of-range reads return f32::MIN (defined behaviour
// shared with the GPU backend).
float shadow_fetch(Buffer& b, V3 s) {
    uint32_t sx = as_u32(std::round(s.x)), sy = as_u32(std::round(s.y));
    if (sx >= b.width) b.diag.shadow_x_wrap++;
    uint32_t idx = sx + sy * b.width;   // wrapping u32, like a release build
    if ((size_t)idx >= b.shadow_buffer.size()) { b.diag.shadow_index_oob++; return std::numeric_limits<float>::lowest(); }
    return b.shadow_buffer[idx];
}
// shadow pass 2: shader.rs:749-788
bool fs_shadow(Buffer& b, const Model& m, uint32_t x, uint32_t y, V3 bar) {
    if (!process_z_value(b, bar, x, y)) return false;
    V3 s;
    M4 sm = mul(b.shadow_matrix, b.i_vpmv_matrix);
    if (!point_from_h(mul(sm, point_h(V3{(float)x, (float)y, dot(bar, b.vertex_z_values)})), &s)) { b.panic = "shadow coord: w == 0"; return true; }
    float coef = 1.0f;
    if (s.z + 1.0f < shadow_fetch(b, s)) coef = 0.3f;
    V2 uv = mul(b.vertex_uvs, bar);
    const uint8_t* color = get_color_at_uv(b, m, uv);
    float diff = dot(bar, b.vertex_intensities);
    uint8_t out[3]; color_blend(color, kBlack, diff * coef, out); set_color(b, out);
    return true;
}
// occlusion pass 2: shader.rs:872-947
bool fs_occlusion(Buffer& b, const Model&, uint32_t x, uint32_t y, V3 bar) {
    if (!process_z_value(b, bar, x, y)) return false;
    V3 light;
    if (!vector_from_h(mul(b.i_m_matrix, vector_h(b.t_light_direction)), &light)) { b.panic = "light: w != 0 (shader.rs:885)"; return true; }
    V4 frag = point_h(V3{(float)x, (float)y, dot(bar, b.vertex_z_values)});
    V3 wp, fs;
    if (!point_from_h(mul(b.i_vpmv_matrix, frag), &wp)) { b.panic = "world pos: w == 0"; return true; }
    if (!point_from_h(mul(mul(b.shadow_matrix, b.i_vpmv_matrix), frag), &fs)) { b.panic = "shadow coord: w == 0"; return true; }
    float fsv = shadow_fetch(b, fs);
    const float threshold = 1.0f;
    float occ = 1.0f;
    const int n_samples = 16;
    const float step = 0.02f;
    const float angle_coef = (2.0f * 3.14159265358979323846f) / (float)n_samples;
    M3 rot;
    if (!rotation_between(V3{0.0f, 0.0f, 1.0f}, light, &rot)) { b.panic = "rotation_between: None (shader.rs:921)"; return true; }
    for (int i = 0; i < n_samples; ++i) {
        float a = angle_coef * (float)i;
        V3 dir = mul(rot, V3{std::sin(a), 0.0f, std::cos(a)});
        V3 sample = add(wp, scale(dir, step));
        V3 sc;
        if (!point_from_h(mul(b.shadow_matrix, point_h(sample)), &sc)) { b.panic = "sample coord: w == 0"; return true; }
        float sv = shadow_fetch(b, sc);
        if (sv - threshold > fsv) {
            float strength = (sv - fsv) / 20.0f;
            strength = (strength != strength) ? 1.0f : (strength < 1.0f ? strength : 1.0f);
            occ -= (1.0f / (float)n_samples) * strength;
        }
    }
    const uint8_t white[3] = {255, 255, 255};
    uint8_t out[3]; color_blend(white, kBlack, occ, out); set_color(b, out);
    return true;
}

// ---- pipeline registry (shader.rs:97-112) -----------------------------------------------------------------

bool make_pipeline(const std::string& name, std::vector<ShaderPass>* passes) {
    passes->clear();
    if (name == "default") passes->push_back({default_prepare, vs_default, fs_default});
    else if (name == "phong") passes->push_back({default_prepare, vs_phong, fs_phong});
    else if (name == "normal_map" || name == "true_normal") passes->push_back({default_prepare, vs_plain, fs_normal_map});
    else if (name == "specular") passes->push_back({default_prepare, vs_plain, fs_specular});
    else if (name == "darboux") passes->push_back({default_prepare, vs_darboux, fs_darboux});
    else if (name == "shadow") {
        passes->push_back({shadow_pass_prepare_1, vs_light, fs_light});
        passes->push_back({shadow_pass_prepare_2, vs_phong, fs_shadow});
    } else if (name == "occlusion") {
        passes->push_back({shadow_pass_prepare_1, vs_light, fs_light});
        passes->push_back({shadow_pass_prepare_2, vs_plain, fs_occlusion});
    } else return false;   // reference: panic!("Provided pipeline name is not supported!")
    return true;
}

// ---- Scene (scene.rs) -----------------------------------------------------------------------------------

struct Scene {
    uint32_t width, height;
    Model model;
    Buffer buffer;
    std::vector<ShaderPass> passes;
    V3 light_direction{0.0f, 0.0f, -1.0f}, look_from{0.0f, 0.0f, 1.0f}, look_at{0.0f, 0.0f, 0.0f}, up{0.0f, 1.0f, 0.0f};   // scene.rs:66-69
    std::vector<uint8_t> frame_buffer;     // bottom-up: pixel (0,0) is bottom-left, index i + j*w
    std::vector<uint32_t> visibility;      // oracle extra: polygon index of the last fragment that passed
    uint64_t bbox_tests = 0, fragments = 0;
};

// scene.rs:174-197
V3 to_barycentric_coord(int32_t px, int32_t py, const int32_t r[2][3]) {
    V3 u{(float)(r[0][1] - r[0][0]), (float)(r[0][2] - r[0][0]), (float)(r[0][0] - px)};
    V3 v{(float)(r[1][1] - r[1][0]), (float)(r[1][2] - r[1][0]), (float)(r[1][0] - py)};
    V3 c = cross(u, v);
    if (std::fabs(c.z) < 1.0f) return {-1.0f, 1.0f, 1.0f};
    return {1.0f - (c.x + c.y) / c.z, c.x / c.z, c.y / c.z};
}

// scene.rs:128-137
void scene_clear(Scene& s) {
    const float fmin = std::numeric_limits<float>::lowest();   // f32::MIN
    size_t n = (size_t)s.width * s.height;
    for (size_t i = 0; i < n; ++i) {
        s.buffer.z_buffer[i] = fmin;
        s.buffer.shadow_buffer[i] = fmin;
        s.frame_buffer[3 * i + 0] = 0; s.frame_buffer[3 * i + 1] = 0; s.frame_buffer[3 * i + 2] = 0;
        s.visibility[i] = kNoTriangle;
    }
}

// scene.rs:151-268
void scene_render(Scene& s) {
    Buffer& b = s.buffer;
    size_t n_tri = s.model.tris.size() / 9;
    for (const ShaderPass& pass : s.passes) {
        pass.prepare(b, s.width, s.height, s.light_direction, s.look_from, s.look_at, s.up);
        for (size_t t = 0; t < n_tri; ++t) {
            const uint32_t* ix = &s.model.tris[9 * t];
            if (!pass.vertex(b, s.model, Idx3{ix[0], ix[3], ix[6]}, Idx3{ix[1], ix[4], ix[7]}, Idx3{ix[2], ix[5], ix[8]})) continue;
            const int32_t (*r)[3] = b.raster;
            int32_t llx = std::min(std::min(r[0][0], r[0][1]), r[0][2]), lly = std::min(std::min(r[1][0], r[1][1]), r[1][2]);
            int32_t urx = std::max(std::max(r[0][0], r[0][1]), r[0][2]), ury = std::max(std::max(r[1][0], r[1][1]), r[1][2]);
            int32_t x_min = std::max(0, llx), x_max = std::min(urx, (int32_t)(s.width - 1));
            int32_t y_min = std::max(0, lly), y_max = std::min(ury, (int32_t)(s.height - 1));
            for (int32_t i = x_min; i <= x_max; ++i) {
                for (int32_t j = y_min; j <= y_max; ++j) {
#ifndef ORC_TIMING      // (the timed build leaves out what the reference does not do: two counters and the visibility store)
                    s.bbox_tests++;
#endif
                    V3 bar = to_barycentric_coord(i, j, r);
                    if (bar.x < 0.0f || bar.y < 0.0f || bar.z < 0.0f) continue;
#ifndef ORC_TIMING
                    s.fragments++;
#endif
                    if (!pass.fragment(b, s.model, (uint32_t)i, (uint32_t)j, bar)) continue;
                    size_t pixel_index = (size_t)(i + j * (int32_t)s.width);
                    s.frame_buffer[3 * pixel_index + 0] = b.fragment_color[0];
                    s.frame_buffer[3 * pixel_index + 1] = b.fragment_color[1];
                    s.frame_buffer[3 * pixel_index + 2] = b.fragment_color[2];
#ifndef ORC_TIMING
                    s.visibility[pixel_index] = (uint32_t)t;
#endif
                }
            }
        }
    }
}

}  // namespace

// ---- C API (ctypes) -----------------------------------------------------------------------------------------
extern "C" {

struct orc_image { const uint8_t* rgb8_topdown; uint32_t w, h; };

// Mirrors Scene::new (scene.rs:47-88).  tri_idx: n_tri x 9 (p0 t0 n0 p1 t1 n1 p2 t2 n2), 0-based.
// Returns nullptr for an unknown pipeline name (reference: panic, shader.rs:108) or bad indices.
void* orc_scene_new(uint32_t width, uint32_t height, const float* positions, uint32_t n_pos, const float* uvs,
                    uint32_t n_uv, const float* normals, uint32_t n_nrm, const uint32_t* tri_idx, uint32_t n_tri,
                    const orc_image* maps /*[4]: texture, normal_map, normal_map_tangent, specular_map*/,
                    const char* pipeline_name) {
    Scene* s = new Scene();
    s->width = width; s->height = height;
    if (!make_pipeline(pipeline_name, &s->passes)) { delete s; return nullptr; }
    for (uint32_t i = 0; i < n_pos; ++i) s->model.positions.push_back({positions[3 * i], positions[3 * i + 1], positions[3 * i + 2]});
    for (uint32_t i = 0; i < n_uv; ++i) s->model.tex_coords.push_back({uvs[2 * i], uvs[2 * i + 1]});
    for (uint32_t i = 0; i < n_nrm; ++i) s->model.normals.push_back({normals[3 * i], normals[3 * i + 1], normals[3 * i + 2]});
    for (uint32_t t = 0; t < n_tri; ++t)
        for (int k = 0; k < 9; ++k) {
            uint32_t v = tri_idx[9 * t + k];
            uint32_t lim = (k % 3 == 0) ? n_pos : (k % 3 == 1 ? n_uv : n_nrm);
            if (v >= lim) { delete s; return nullptr; }
            s->model.tris.push_back(v);
        }
    Image* dst[4] = {&s->model.texture, &s->model.normal_map, &s->model.normal_map_tangent, &s->model.specular_map};
    for (int k = 0; k < 4; ++k) {
        dst[k]->w = maps[k].w; dst[k]->h = maps[k].h;
        dst[k]->px.assign(maps[k].rgb8_topdown, maps[k].rgb8_topdown + (size_t)3 * maps[k].w * maps[k].h);
    }
    size_t n = (size_t)width * height;
    s->buffer.width = width; s->buffer.height = height;
    s->buffer.z_buffer.assign(n, 0.0f);         // Buffer::new: 0.0, not f32::MIN (shader.rs:46-47)
    s->buffer.shadow_buffer.assign(n, 0.0f);
    s->frame_buffer.assign(3 * n, 0);
    s->visibility.assign(n, kNoTriangle);
    return s;
}
void orc_scene_free(void* h) { delete (Scene*)h; }
void orc_scene_clear(void* h) { scene_clear(*(Scene*)h); }
void orc_scene_set_light_direction(void* h, const float* d) { ((Scene*)h)->light_direction = {d[0], d[1], d[2]}; }
void orc_scene_set_camera(void* h, const float* from, const float* at, const float* up) {
    Scene* s = (Scene*)h;
    s->look_from = {from[0], from[1], from[2]}; s->look_at = {at[0], at[1], at[2]}; s->up = {up[0], up[1], up[2]};
}
// Returns 0, or 1 if the reference would have panicked somewhere (message via orc_scene_panic).
int orc_scene_render(void* h) { Scene* s = (Scene*)h; s->buffer.panic = nullptr; scene_render(*s); return s->buffer.panic ? 1 : 0; }
const char* orc_scene_panic(void* h) { const char* p = ((Scene*)h)->buffer.panic; return p ? p : ""; }

// get_frame_buffer (scene.rs:92-97): clone + vertical flip -> top-down RGB8.
void orc_scene_get_frame_buffer(void* h, uint8_t* dst) {
    Scene* s = (Scene*)h;
    size_t row = (size_t)3 * s->width;
    for (uint32_t j = 0; j < s->height; ++j) std::memcpy(dst + (size_t)j * row, &s->frame_buffer[(size_t)(s->height - 1 - j) * row], row);
}
// get_z_buffer / get_shadow_buffer (scene.rs:101-125): saturating f32 -> u8 gray, flipped.
static void gray_view(Scene* s, const std::vector<float>& src, uint8_t* dst) {
    for (uint32_t j = 0; j < s->height; ++j)
        for (uint32_t i = 0; i < s->width; ++i) {
            uint8_t g = as_u8(src[(size_t)(s->height - 1 - j) * s->width + i]);
            uint8_t* p = dst + 3 * ((size_t)j * s->width + i);
            p[0] = g; p[1] = g; p[2] = g;
        }
}
void orc_scene_get_z_buffer(void* h, uint8_t* dst) { Scene* s = (Scene*)h; gray_view(s, s->buffer.z_buffer, dst); }
void orc_scene_get_shadow_buffer(void* h, uint8_t* dst) { Scene* s = (Scene*)h; gray_view(s, s->buffer.shadow_buffer, dst); }
// Raw, bottom-up (index i + j*w) views of the internal state.
void orc_scene_raw_frame(void* h, uint8_t* dst) { Scene* s = (Scene*)h; std::memcpy(dst, s->frame_buffer.data(), s->frame_buffer.size()); }
void orc_scene_raw_z(void* h, float* dst) { Scene* s = (Scene*)h; std::memcpy(dst, s->buffer.z_buffer.data(), 4 * s->buffer.z_buffer.size()); }
void orc_scene_raw_shadow(void* h, float* dst) { Scene* s = (Scene*)h; std::memcpy(dst, s->buffer.shadow_buffer.data(), 4 * s->buffer.shadow_buffer.size()); }
void orc_scene_raw_visibility(void* h, uint32_t* dst) { Scene* s = (Scene*)h; std::memcpy(dst, s->visibility.data(), 4 * s->visibility.size()); }
// out[0..5]: shadow_index_oob, shadow_x_wrap, texel_oob, nan_depth, bbox_tests, fragments
void orc_scene_diagnostics(void* h, uint64_t* out) {
    Scene* s = (Scene*)h; const Diag& d = s->buffer.diag;
    out[0] = d.shadow_index_oob; out[1] = d.shadow_x_wrap; out[2] = d.texel_oob; out[3] = d.nan_depth;
    out[4] = s->bbox_tests; out[5] = s->fragments;
}
// The uniforms the last prepare left in the Buffer, column-major 4x4 each: vpmv, i_vpmv, m, i_m, it_m, shadow
// then camera_direction(3), t_light_direction(3): 102 floats.
void orc_scene_uniforms(void* h, float* out) {
    Buffer& b = ((Scene*)h)->buffer;
    const M4* ms[6] = {&b.vpmv_matrix, &b.i_vpmv_matrix, &b.m_matrix, &b.i_m_matrix, &b.it_m_matrix, &b.shadow_matrix};
    for (int k = 0; k < 6; ++k) std::memcpy(out + 16 * k, ms[k]->m, 64);
    out[96] = b.camera_direction.x; out[97] = b.camera_direction.y; out[98] = b.camera_direction.z;
    out[99] = b.t_light_direction.x; out[100] = b.t_light_direction.y; out[101] = b.t_light_direction.z;
}

// The per-pass constants of the occlusion fragment (shader.rs:882-885, 899-900, 916-929) as the reference recomputes them
// for every fragment: out[0..48) = the 16 sample offsets dir_k * 0.02 (xyz each), out[48..64) = shadow_matrix * i_vpmv
// (column-major).  Returns 1 where the reference would panic (light w != 0, rotation_between None).
int orc_scene_ao_uniforms(void* h, float* out) {
    Buffer& b = ((Scene*)h)->buffer;
    V3 light;
    if (!vector_from_h(mul(b.i_m_matrix, vector_h(b.t_light_direction)), &light)) return 1;
    M3 rot;
    if (!rotation_between(V3{0.0f, 0.0f, 1.0f}, light, &rot)) return 1;
    const float angle_coef = (2.0f * 3.14159265358979323846f) / 16.0f;
    for (int i = 0; i < 16; ++i) {
        float a = angle_coef * (float)i;
        V3 d = scale(mul(rot, V3{std::sin(a), 0.0f, std::cos(a)}), 0.02f);
        out[3 * i] = d.x; out[3 * i + 1] = d.y; out[3 * i + 2] = d.z;
    }
    M4 sm = mul(b.shadow_matrix, b.i_vpmv_matrix);
    std::memcpy(out + 48, sm.m, 64);
    return 0;
}

// Many-light ambient-occlusion batch (BASELINE.json config 4; semantics SURVEY.md §8(d)): for each light
// `set_light_direction; clear; render` with the scene's (occlusion) pipeline, accum[p] += gray (u32), where
// p is the bottom-up pixel index and gray is channel 0 of the frame the reference would have produced.
int orc_scene_render_ao(void* h, const float* light_dirs, uint32_t n_lights, uint32_t* accum) {
    Scene* s = (Scene*)h;
    size_t n = (size_t)s->width * s->height;
    int rc = 0;
    for (uint32_t k = 0; k < n_lights; ++k) {
        s->light_direction = {light_dirs[3 * k], light_dirs[3 * k + 1], light_dirs[3 * k + 2]};
        scene_clear(*s);
        s->buffer.panic = nullptr;
        scene_render(*s);
        if (s->buffer.panic) rc = 1;
        for (size_t p = 0; p < n; ++p) accum[p] += s->frame_buffer[3 * p];
    }
    return rc;
}

}  // extern "C"
