// tr_types.cuh — data layout shared by the host driver and the sm_100a kernels.
#pragma once
#include "tr_math.cuh"

namespace tr {

// (the pipeline ids are generated from the pipeline list: pipeline_table.cuh)

enum DiagSlot : int {
    D_SHADOW_OOB = 0,   // shadow index >= w*h           (reference: index panic)
    D_SHADOW_XWRAP,     // round(x) >= w                 (reference: reads next row)
    D_TEXEL_OOB,        // texel outside the map         (reference: get_pixel panic)
    D_NAN_DEPTH,
    D_W_ZERO,           // homogeneous w == 0            (reference: unwrap panic)
    D_SINGULAR_BASIS,   // darboux local basis singular  (reference: unwrap panic)
    D_BINNED,
    D_RESERVED,
    D_COUNT
};

// Per-pass constants = the uniform part of the reference's `Buffer` (shader.rs:21-28) plus values the
// reference recomputes per fragment although they are constant over a pass (hoisted to K0, same op order).
struct Uniforms {
    mat4 vpmv, i_vpmv, m, i_m, it_m, shadow;
    mat4 shadow_from_screen;      // shadow * i_vpmv          (shader.rs:763-764, 899-900)
    f3 camera_direction;
    f3 t_light;
    f3 ao_step[16];               // (rot * (sin a_k, 0, cos a_k)) * 0.02   (shader.rs:920-929)
    float ao_w_rowsum;            // |s30| + |s31| + |s32| of `shadow` (bounds how far a sample's w moves; k_shade.cuh)
    float ao_x_rowsum, ao_y_rowsum;   // the same sums over rows 0 and 1 (bound the samples' shadow-buffer coordinates)
    uint32_t debug_flags;         // bit 0: occlusion samples always take the range-tested division
    uint32_t width, height;
    uint32_t n_triangles;
    uint32_t epoch;               // renders since the last clear (see Key below)
};

// ---- visibility key ------------------------------------------------------------------------------------
// One u64 per pixel: hi = ~float_order(z + 0.0f), lo = epoch << 24 | triangle id.  A smaller key is a nearer
// fragment (the reference keeps the LARGER z: shader.rs:175), ties go to the smaller lo = earlier render,
// then lower polygon index — exactly what the reference's in-order `z <= zbuf -> reject` loop leaves behind.
// lo == 0xFFFFFFFF marks "no triangle"; such an entry only yields to a strictly larger z.
constexpr uint32_t kNoTri = 0xFFFFFFFFu;
constexpr uint32_t kTriMask = 0x00FFFFFFu;
constexpr uint32_t kMaxTriangles = 0x00FFFFFEu;
constexpr uint32_t kMaxEpoch = 0xFEu;

TR_HD uint64_t make_key(float z, uint32_t lo) { return ((uint64_t)(~float_order(z + 0.0f)) << 32) | lo; }
TR_HD float key_depth(uint64_t k) { return float_unorder(~(uint32_t)(k >> 32)); }
TR_HD uint32_t key_lo(uint64_t k) { return (uint32_t)k; }
// true when candidate (z, lo) replaces cur under the reference's rule.
TR_HD bool key_beats(uint64_t cand, uint64_t cur) {
    uint32_t ch = (uint32_t)(cand >> 32), kh = (uint32_t)(cur >> 32);
    return ch < kh || (ch == kh && (uint32_t)cur != kNoTri && (uint32_t)cand < (uint32_t)cur);
}

// ---- K1 output: one 64-byte setup record per polygon (+ varyings) -------------------------------------------
// rec[0..5]  raster x0 y0 x1 y1 x2 y2 (i32 bit patterns)        vertex_t_raster   shader.rs:34
// rec[6..8]  z0 z1 z2                                            vertex_z_values   shader.rs:35
// rec[9]     flags (bit0 = drawn i.e. not culled)
// rec[10..15] u0 u1 u2 v0 v1 v2 (v already flipped 1-v)          vertex_uvs        shader.rs:33
// var[0..2]  vertex_intensities                                  shader.rs:30
// var[3..11] darboux: unit(p1-p0), unit(p2-p0) rows (6) then 3 pad; var[12..20] vertex_t_normals (col-major 3x3)
constexpr int kRecWords = 16;
constexpr int kVarWords = 24;
constexpr uint32_t kFlagDrawn = 1u;

}  // namespace tr
