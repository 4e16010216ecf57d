// pipelines/default.cuh — `default`: flat shading, one intensity per polygon (shader.rs:282-342).
#pragma once
#include "../pipeline.cuh"

namespace tr {

struct PipelineDefault {
    static constexpr int id = kPipeDefault;
    static constexpr const char* name = "default";
    static constexpr const char* alias = nullptr;
    static constexpr bool light_pass = false;
    static constexpr int vertex_role = kPipeDefault;
    static constexpr int varying_words = 3;
    static constexpr bool heavy_fragment = false;

    // prepare: default_prepare (shader.rs:183-230)
    static int prepare(Uniforms& u, uint32_t w, uint32_t h, f3 light, f3 from, f3 at, f3 up) { return prepare_camera(u, w, h, light, from, at, up); }

    // vertex (shader.rs:297-305): face normal -> it_m -> normalised -> . t_light, the same value for the three corners
    static __device__ __forceinline__ void vertex(const Uniforms& u, const VertexIn& in, float (&v)[kVarWords], unsigned long long* diag) {
        const float d = dot3(u.t_light, normal_to_camera(u, in.face_normal, diag));
        v[0] = d; v[1] = d; v[2] = d;
    }

    // fragment (shader.rs:318-333)
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs& a, const Uniforms&, const FragmentIn& f) {
        const uchar4 c = fetch_texel(a.maps, 0, 0, fragment_uv(f), a.diag);
        const float diff = __ldg(&f.var[0]).x;                                   // shader.rs:329
        return blend_texel(c, diff);
    }
};

}  // namespace tr
