// pipelines/shadow.cuh — `shadow`: light-depth pass (PassLightDepth), then phong shading dimmed where the fragment lies
// behind the light's nearest surface (shader.rs:668-802).
#pragma once
#include "phong.cuh"

namespace tr {

struct PipelineShadow {
    static constexpr int id = kPipeShadow;
    static constexpr const char* name = "shadow";
    static constexpr const char* alias = nullptr;
    static constexpr bool light_pass = true;            // pass 1: shadow_pass_prepare_1 + no-cull vertex + depth-max fragment (shader.rs:671-709)
    static constexpr int vertex_role = kPipePhong;        // pass 2 shares the phong vertex closure (shader.rs:711-747 == :349-384)
    static constexpr int varying_words = 3;
    static constexpr bool heavy_fragment = false;

    // prepare of pass 2: shadow_pass_prepare_2 (shader.rs:259-279); u.shadow already holds the light matrix
    static int prepare(Uniforms& u, uint32_t w, uint32_t h, f3 light, f3 from, f3 at, f3 up) {
        return prepare_camera_with_shadow(u, w, h, light, from, at, up, false);
    }

    // vertex of pass 2 (shader.rs:711-747) = the phong vertex closure
    static __device__ __forceinline__ void vertex(const Uniforms& u, const VertexIn& in, float (&v)[kVarWords], unsigned long long* diag) {
        PipelinePhong::vertex(u, in, v, diag);
    }

    // fragment of pass 2 (shader.rs:749-788)
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs& a, const Uniforms& u, const FragmentIn& f) {
        const uchar4 c = fetch_texel(a.maps, 0, 0, fragment_uv(f), a.diag);
        float diff = PipelinePhong::intensity(f);                                        // shader.rs:784
        bool ok;
        DiagLocal dl;
        const f3 s = dehomog_shared(xform(u.shadow_from_screen, f4{(float)f.px, (float)f.py, f.z, 1.0f}), ok);
        dl.w_zero += ok ? 0u : 1u;
        const float sv = shadow_lookup(a.shadow, s, u.width, u.width * u.height, dl);
        dl.flush(a.diag);
        const float coef = (s.z + 1.0f < sv) ? 0.3f : 1.0f;                              // shader.rs:776-780
        diff = diff * coef;
        return blend_texel(c, diff);
    }
};

}  // namespace tr
