// pipeline.cuh — the reference's three-role shader surface (/root/reference/src/scene/shader.rs:55-88):
//
//     pub struct ShaderPass { prepare: Prepare, vertex: VertexShader, fragment: FragmentShader }      shader.rs:84-88
//
// rebuilt as compile-time roles.  A pipeline is ONE struct in ONE header under pipelines/ with
//
//     static int  prepare(Uniforms&, w, h, light, from, at, up)          host   (K0; `Prepare`,        shader.rs:55-63)
//     static void vertex(const Uniforms&, const VertexIn&, float (&)[kVarWords], diag)   device (K1; `VertexShader`,   shader.rs:66-72)
//     static u32  fragment(const ShadeArgs&, const Uniforms&, const FragmentIn&)         device (K5; `FragmentShader`, shader.rs:75-80)
//
// plus a few constants (name, optional alias, whether a light-depth pass runs first, how many varying words the vertex
// role writes, whose vertex role it runs when it shares one).  pipeline_table.cuh lists the structs once; the name lookup (ShaderPipeline::new, shader.rs:97-112), the
// kernel instantiations (k_vertex_setup<PIPE>, k_raster<PIPE, THREADS>) and the host dispatch are all generated from
// that list, so adding a pipeline is one new header plus one line of the table.
//
// What the reference's closures share through helper functions stays shared here, outside the roles: face culling and
// the vertex transform (shader.rs:116-165) in k_vertex_setup, the depth test (process_z_value, shader.rs:169-180) in
// k_raster, the samplers and color_blend (util.rs) in k_shade.cuh.  The first pass of the two-pass pipelines (light
// depth, shader.rs:671-709 / 809-847: no cull, transform by the light matrix, `>=` max) is the same for both and is
// the PassLightDepth tag below.
#pragma once
#include "k_shade.cuh"

namespace tr {

// What a vertex closure reads from the Model for one polygon (object space, shader.rs:127-133 + the obj normals).
struct VertexIn {
    f3 p[3];            // positions
    f3 n[3];            // vertex normals
    f3 face_normal;     // (p1 - p0) x (p2 - p0), already formed by the cull test (shader.rs:117-119)
};

// What a fragment closure reads for one covered pixel: the winning polygon, the pixel, its depth and barycentrics as
// the reference computes them (scene.rs:178-196), and the polygon's uv rows / varyings written by the vertex stage.
struct FragmentIn {
    uint32_t tri;
    int px, py;
    float z, b0, b1, b2;
    float4 r2, r3;              // setup record words 8..15: (z2, flags, u0, u1) (u2, v0, v1, v2)
    const float4* var;          // this polygon's varyings (kVarWords)
};

// uv = vertex_uvs * bar (2x3 mat-vec, column order; shader.rs:327)
__device__ __forceinline__ f2 fragment_uv(const FragmentIn& f) {
    return f2{(f.r2.z * f.b0 + f.r2.w * f.b1) + f.r3.x * f.b2, (f.r3.y * f.b0 + f.r3.z * f.b1) + f.r3.w * f.b2};
}
// color_blend(texel, black, diff) on all three channels (shader.rs:330 and its siblings)
__device__ __forceinline__ uint32_t blend_texel(uchar4 c, float diff) {
    return pack_rgb(blend_black(c.x, diff), blend_black(c.y, diff), blend_black(c.z, diff));
}
// it_m * (n, 0) -> must have w == 0 -> normalised   (shader.rs:300-303, 368-371)
__device__ __forceinline__ f3 normal_to_camera(const Uniforms& u, f3 n, unsigned long long* diag) {
    f4 t = xform(u.it_m, f4{n.x, n.y, n.z, 0.0f});
    if (!(t.w == 0.0f)) atomicAdd(&diag[D_W_ZERO], 1ull);
    return unit3(f3{t.x, t.y, t.z});
}

// Kernel modes that are not pipelines: the shared first pass of shadow / occlusion, and visibility alone (camera pass of
// the many-light AO batch).  They carry the same constants so that k_raster<PIPE, THREADS> takes them like a pipeline.
constexpr int kModeDepthOnly = -2;     // K4: shadow buffer, `>=` max, no culling upstream
constexpr int kModeVisibility = -1;    // K3 alone: keys only
struct PassLightDepth {
    static constexpr int id = kModeDepthOnly;
    static constexpr bool heavy_fragment = false;
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs&, const Uniforms&, const FragmentIn&) { return 0u; }
};
struct PassVisibility {
    static constexpr int id = kModeVisibility;
    static constexpr bool heavy_fragment = false;
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs&, const Uniforms&, const FragmentIn&) { return 0u; }
};

}  // namespace tr
