// pipelines/darboux.cuh — `darboux`: tangent-space normal map in a per-fragment Darboux frame (shader.rs:546-664).
#pragma once
#include "../pipeline.cuh"

namespace tr {

struct PipelineDarboux {
    static constexpr int id = kPipeDarboux;
    static constexpr const char* name = "darboux";
    static constexpr const char* alias = nullptr;
    static constexpr bool light_pass = false;
    static constexpr int vertex_role = kPipeDarboux;
    static constexpr int varying_words = 21;
    static constexpr bool heavy_fragment = true;        // 62 registers: the tile kernel keeps fewer CTAs per SM (k_raster.cuh)

    static int prepare(Uniforms& u, uint32_t w, uint32_t h, f3 light, f3 from, f3 at, f3 up) { return prepare_camera(u, w, h, light, from, at, up); }

    // vertex (shader.rs:562-584): vertex_t_positions = m * p and vertex_t_normals = normalize(it_m * n).  Rows 0 and 1 of
    // the fragment's local basis (shader.rs:612-623) only depend on the polygon — unit(P*(-1,1,0)), unit(P*(-1,0,1)) with
    // P = columns m*p_i — and are formed here, once, in the fragment's operation order.
    static __device__ __forceinline__ void vertex(const Uniforms& u, const VertexIn& in, float (&v)[kVarWords], unsigned long long* diag) {
        f3 tp[3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            bool ok;
            tp[i] = dehomog(xform(u.m, f4{in.p[i].x, in.p[i].y, in.p[i].z, 1.0f}), ok);
        }
        const mat3 P = {{tp[0].x, tp[1].x, tp[2].x, tp[0].y, tp[1].y, tp[2].y, tp[0].z, tp[1].z, tp[2].z}};
        const f3 r0 = unit3(xform(P, f3{-1.0f, 1.0f, 0.0f}));
        const f3 r1 = unit3(xform(P, f3{-1.0f, 0.0f, 1.0f}));
        v[3] = r0.x; v[4] = r0.y; v[5] = r0.z; v[6] = r1.x; v[7] = r1.y; v[8] = r1.z;
#pragma unroll
        for (int i = 0; i < 3; ++i) {
            const f3 n = normal_to_camera(u, in.n[i], diag);
            v[12 + 3 * i] = n.x; v[13 + 3 * i] = n.y; v[14 + 3 * i] = n.z;   // column i of vertex_t_normals
        }
    }

    // fragment (shader.rs:597-655)
    static __device__ __forceinline__ uint32_t fragment(const ShadeArgs& a, const Uniforms& u, const FragmentIn& f) {
        const f2 uv = fragment_uv(f);
        const uchar4 c = fetch_texel(a.maps, 0, 0, uv, a.diag);
        const f3 nt = decode_normal(fetch_texel(a.maps, 2, 1, uv, a.diag));             // util.rs:61-63: sized by normal_map
        const float4* v = f.var;
        const float4 v0 = __ldg(&v[0]), v1 = __ldg(&v[1]), v2 = __ldg(&v[2]), v3 = __ldg(&v[3]), v4 = __ldg(&v[4]), v5 = __ldg(&v[5]);
        const mat3 N = {{v3.x, v3.w, v4.z, v3.y, v4.x, v4.w, v3.z, v4.y, v5.x}};         // vertex_t_normals, columns n0 n1 n2
        const f3 local_z = xform(N, f3{f.b0, f.b1, f.b2});
        const f3 row2 = unit3_shared(local_z);
        // A rows: (v0.w v1.x v1.y), (v1.z v1.w v2.x), row2 — inverse by minors, each entry divided by det
        const float m11 = v0.w, m12 = v1.x, m13 = v1.y, m21 = v1.z, m22 = v1.w, m23 = v2.x, m31 = row2.x, m32 = row2.y, m33 = row2.z;
        const float mi1 = m22 * m33 - m32 * m23, mi2 = m21 * m33 - m31 * m23, mi3 = m21 * m32 - m31 * m22;
        const float det = m11 * mi1 - m12 * mi2 + m13 * mi3;
        if (det == 0.0f) { atomicAdd(&a.diag[D_SINGULAR_BASIS], 1ull); return 0u; }
        // every entry of the adjugate divided by det (nalgebra 3x3 try_inverse), one shared divisor
        const float c12 = m13 * m32 - m33 * m12, c22 = m11 * m33 - m31 * m13, c32 = m12 * m31 - m32 * m11;
        // uvs: r2.z r2.w r3.x = u0 u1 u2 ; r3.y r3.z r3.w = v0 v1 v2
        const f3 du{f.r2.w - f.r2.z, f.r3.x - f.r2.z, 0.0f}, dv{f.r3.z - f.r3.y, f.r3.w - f.r3.y, 0.0f};
        f3 lx, ly;
        if (in_fast_division_range(det)) {
            // The third column of the inverse only ever meets the zero component of du / dv.  On this branch det is
            // finite and normal, hence so are the cofactors (products of unit-vector components): 0 * I(.,2) is a zero
            // whose only possible trace is the sign of a zero sum, which no later step can turn into a different
            // colour (it reaches the result through x / len, products and sums only).  So the column is not formed.
            const SharedDivisor d = shared_divisor(det);
            const float i00 = quotient(mi1, d), i01 = quotient(c12, d), i10 = quotient(-mi2, d), i11 = quotient(c22, d),
                        i20 = quotient(mi3, d), i21 = quotient(c32, d);
            lx = unit3_shared(f3{i01 * du.y + i00 * du.x, i11 * du.y + i10 * du.x, i21 * du.y + i20 * du.x});
            ly = unit3_shared(f3{i01 * dv.y + i00 * dv.x, i11 * dv.y + i10 * dv.x, i21 * dv.y + i20 * dv.x});
        } else {
            const float c13 = m12 * m23 - m22 * m13, c23 = m13 * m21 - m23 * m11, c33 = m11 * m22 - m21 * m12;
            const mat3 I = {{mi1 / det, c12 / det, c13 / det, -mi2 / det, c22 / det, c23 / det, mi3 / det, c32 / det, c33 / det}};
            lx = unit3_shared(xform(I, du));
            ly = unit3_shared(xform(I, dv));
        }
        const f3 lz = unit3_shared(local_z);
        const mat3 B = {{lx.x, ly.x, lz.x, lx.y, ly.y, lz.y, lx.z, ly.z, lz.z}};
        const f3 n = unit3_shared(xform(B, nt));
        return blend_texel(c, dot3(u.t_light, n));
    }
};

}  // namespace tr
