// pipelines/occlusion.cuh — `occlusion`: light-depth pass (PassLightDepth), then a 16-sample ring of shadow-buffer
// lookups around the fragment's world position (shader.rs:806-963).  The sample loop itself (occlusion_gray) lives in
// k_shade.cuh because the many-light batch (K6 k_ao_accumulate) runs it too.
#pragma once
#include "../pipeline.cuh"

namespace tr {

struct PipelineOcclusion {
    static constexpr int id = kPipeOcclusion;
    static constexpr const char* name = "occlusion";
    static constexpr const char* alias = nullptr;
    static constexpr bool light_pass = true;            // pass 1: shader.rs:809-847
    static constexpr int vertex_role = kPipeOcclusion;
    static constexpr int varying_words = 0;
    static constexpr bool heavy_fragment = false;

    // prepare of pass 2: shadow_pass_prepare_2 + the per-pass constants the fragment closure recomputes per fragment
    // (rotation_between, the 16 step vectors; shader.rs:882-885, 916-929)
    static int prepare(Uniforms& u, uint32_t w, uint32_t h, f3 light, f3 from, f3 at, f3 up) {
        return prepare_camera_with_shadow(u, w, h, light, from, at, up, true);
    }

    // vertex of pass 2 (shader.rs:849-870): nothing beyond the shared transform / uv store
    static __device__ __forceinline__ void vertex(const Uniforms&, const VertexIn&, float (&)[kVarWords], unsigned long long*) {}

    // fragment of pass 2 (shader.rs:872-947)
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs& a, const Uniforms& u, const FragmentIn& f) {
        DiagLocal dl;
        bool ok;
        const f3 wp = dehomog(xform(u.i_vpmv, f4{(float)f.px, (float)f.py, f.z, 1.0f}), ok);
        dl.w_zero += ok ? 0u : 1u;
        const uint32_t g = occlusion_gray(a.shadow, u, f.px, f.py, f.z, wp, dl);
        dl.flush(a.diag);
        return pack_rgb(g, g, g);
    }
};

}  // namespace tr
