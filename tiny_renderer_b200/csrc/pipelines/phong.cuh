// pipelines/phong.cuh — `phong`: per-vertex intensities interpolated over the polygon (shader.rs:346-410).
#pragma once
#include "../pipeline.cuh"

namespace tr {

struct PipelinePhong {
    static constexpr int id = kPipePhong;
    static constexpr const char* name = "phong";
    static constexpr const char* alias = nullptr;
    static constexpr bool light_pass = false;
    static constexpr int vertex_role = kPipePhong;
    static constexpr int varying_words = 3;
    static constexpr bool heavy_fragment = false;

    static int prepare(Uniforms& u, uint32_t w, uint32_t h, f3 light, f3 from, f3 at, f3 up) { return prepare_camera(u, w, h, light, from, at, up); }

    // vertex (shader.rs:362-373): one intensity per corner from the obj normals
    static __device__ __forceinline__ void vertex(const Uniforms& u, const VertexIn& in, float (&v)[kVarWords], unsigned long long* diag) {
#pragma unroll
        for (int i = 0; i < 3; ++i) v[i] = dot3(u.t_light, normal_to_camera(u, in.n[i], diag));
    }

    // bar . vertex_intensities (shader.rs:397)
    static __device__ __forceinline__ float intensity(const FragmentIn& f) {
        const float4 in = __ldg(&f.var[0]);
        return (f.b0 * in.x + f.b1 * in.y) + f.b2 * in.z;
    }

    // fragment (shader.rs:386-401)
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs& a, const Uniforms&, const FragmentIn& f) {
        const uchar4 c = fetch_texel(a.maps, 0, 0, fragment_uv(f), a.diag);
        return blend_texel(c, intensity(f));
    }
};

}  // namespace tr
