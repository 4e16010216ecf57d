// pipelines/specular.cuh — `specular`: normal map + specular highlight from the specular map (shader.rs:472-543).
#pragma once
#include "normal_map.cuh"

namespace tr {

struct PipelineSpecular {
    static constexpr int id = kPipeSpecular;
    static constexpr const char* name = "specular";
    static constexpr const char* alias = nullptr;
    static constexpr bool light_pass = false;
    static constexpr int vertex_role = kPipeSpecular;
    static constexpr int varying_words = 0;
    static constexpr bool heavy_fragment = false;

    static int prepare(Uniforms& u, uint32_t w, uint32_t h, f3 light, f3 from, f3 at, f3 up) { return prepare_camera(u, w, h, light, from, at, up); }

    // vertex (shader.rs:475-496): nothing beyond the shared transform / uv store
    static __device__ __forceinline__ void vertex(const Uniforms&, const VertexIn&, float (&)[kVarWords], unsigned long long*) {}

    // fragment (shader.rs:498-534).  powf: CUDA's and glibc's differ by a few ulp — inside north_star's 1-LSB budget.
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs& a, const Uniforms& u, const FragmentIn& f) {
        const f2 uv = fragment_uv(f);
        const uchar4 c = fetch_texel(a.maps, 0, 0, uv, a.diag);
        const f3 n = PipelineNormalMap::mapped_normal(a, u, uv);
        const float diff = dot3(u.t_light, n);
        const f3 refl = unit3_shared((n * dot3(u.t_light, n)) * 2.0f - u.t_light);         // shader.rs:515-518
        const float e = (float)fetch_texel(a.maps, 3, 3, uv, a.diag).x;                    // util.rs:82: raw 0..255
        const float spec = 0.6f * powf(fmaxf(refl.z, 0.0f), e);
        const float k = diff + spec;
        return pack_rgb(cast_u8(fminf(k * (float)c.x, 255.0f)), cast_u8(fminf(k * (float)c.y, 255.0f)),
                        cast_u8(fminf(k * (float)c.z, 255.0f)));                           // shader.rs:526-530
    }
};

}  // namespace tr
