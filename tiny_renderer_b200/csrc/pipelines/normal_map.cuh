// pipelines/normal_map.cuh — `normal_map` (README spelling `true_normal`): normals from the object-space normal map
// (shader.rs:413-468).
#pragma once
#include "../pipeline.cuh"

namespace tr {

struct PipelineNormalMap {
    static constexpr int id = kPipeNormalMap;
    static constexpr const char* name = "normal_map";
    static constexpr const char* alias = "true_normal";        // README.md:18 names it so; the code matches "normal_map" (shader.rs:103)
    static constexpr bool light_pass = false;
    static constexpr int vertex_role = kPipeNormalMap;
    static constexpr int varying_words = 0;
    static constexpr bool heavy_fragment = false;

    static int prepare(Uniforms& u, uint32_t w, uint32_t h, f3 light, f3 from, f3 at, f3 up) { return prepare_camera(u, w, h, light, from, at, up); }

    // vertex (shader.rs:416-437): nothing beyond the shared transform / uv store
    static __device__ __forceinline__ void vertex(const Uniforms&, const VertexIn&, float (&)[kVarWords], unsigned long long*) {}

    // n = normalize(it_m * (normal_map(uv), 0))   (shader.rs:451-454)
    static __device__ __forceinline__ f3 mapped_normal(const ShadeArgs& a, const Uniforms& u, f2 uv) {
        const f3 nm = decode_normal(fetch_texel(a.maps, 1, 1, uv, a.diag));
        const f4 t = xform(u.it_m, f4{nm.x, nm.y, nm.z, 0.0f});
        return unit3_shared(f3{t.x, t.y, t.z});
    }

    // fragment (shader.rs:439-459)
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs& a, const Uniforms& u, const FragmentIn& f) {
        const f2 uv = fragment_uv(f);
        const uchar4 c = fetch_texel(a.maps, 0, 0, uv, a.diag);
        return blend_texel(c, dot3(u.t_light, mapped_normal(a, u, uv)));
    }
};

}  // namespace tr
