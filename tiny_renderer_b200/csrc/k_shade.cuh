// k_shade.cuh — what the fragment closures share (samplers, color_blend, shadow-buffer lookups, the occlusion sample
// loop) and K6 ambient-occlusion accumulation (deferred, many-light batch).  The fragment closures themselves are the
// `fragment` roles of the pipeline structs under pipelines/ (pipeline.cuh), called from the fused tile kernel in
// k_raster.cuh for the winner of every pixel.
//
// Texels are point-sampled through texture objects (the texture path).  Replaces the fragment closures of
// /root/reference/src/scene/shader.rs (:318-333 default, :386-401 phong, :439-459 normal_map, :498-534 specular,
// :597-655 darboux, :749-788 shadow, :872-947 occlusion), the samplers of /root/reference/src/scene/util.rs:34-83
// and color_blend (util.rs:7-13).
#pragma once
#include <type_traits>
#include <cuda_runtime.h>
#include "k_tile.cuh"

namespace tr {

struct Maps {
    cudaTextureObject_t tex[4];     // texture, normal_map, normal_map_tangent, specular_map (RGBA8, point, unnormalised)
    uint32_t w[4], h[4];
};

struct ShadeArgs {
    const unsigned long long* keys;
    const uint32_t* tile_state;     // per raster tile: 1 = cleared, keys of the tile are stale (k_raster.cuh)
    uint32_t tiles_x;
    const float4* rec;
    const float4* var;
    const float* shadow;
    uint8_t* rgb;                   // top-down: row (height-1-j)
    uint32_t* ao_accum;             // K6: bottom-up u32 sums (index i + j*w)
    unsigned long long* diag;
    Maps maps;
    int fresh;                      // frame was cleared: pixels not shaded now are written as (0,0,0)
    // per raster tile: nonzero = the tile's RGB in memory is all zeros already (a cleared frame need not rewrite it)
    uint32_t* rgb_zero;
    int rgb_tma;                    // the frame layout allows the TMA tile store (row pitch a multiple of 16 bytes)
};

// util.rs:34-41 addressing: (uv * size) as u32, image::get_pixel panics when outside -> counted, clamped.
__device__ __forceinline__ uchar4 fetch_texel(const Maps& mp, int which, int size_of, f2 uv, unsigned long long* diag) {
    uint32_t x = cast_u32(uv.x * (float)mp.w[size_of]), y = cast_u32(uv.y * (float)mp.h[size_of]);
    if (x >= mp.w[which] || y >= mp.h[which]) {
        atomicAdd(&diag[D_TEXEL_OOB], 1ull);
        x = min(x, mp.w[which] - 1); y = min(y, mp.h[which] - 1);
    }
    return tex2D<uchar4>(mp.tex[which], (float)x + 0.5f, (float)y + 0.5f);
}
// util.rs:51-56 / 67-72
__device__ __forceinline__ f3 decode_normal(uchar4 c) {
    // c / 255.0 with integer c in 0..255: exact_quotient gives the IEEE quotient (integer operands far below 2^21)
    const float r255 = 1.0f / 255.0f;
    return unit3_shared(f3{exact_quotient((float)c.x, 255.0f, r255) - 0.5f, exact_quotient((float)c.y, 255.0f, r255) - 0.5f,
                    exact_quotient((float)c.z, 255.0f, r255) - 0.5f});
}
// color_blend(c, (0,0,0), t), util.rs:7-13
__device__ __forceinline__ uint32_t blend_black(uint32_t c, float t) { return cast_u8(t * (float)c + (1.0f - t) * 0.0f); }
__device__ __forceinline__ uint32_t pack_rgb(uint32_t r, uint32_t g, uint32_t b) { return r | (g << 8) | (b << 16); }

// Events on which the reference panics or misreads, counted in registers and flushed once per thread.
struct DiagLocal {
    uint32_t w_zero = 0, xwrap = 0, oob = 0;
    __device__ __forceinline__ void flush(unsigned long long* diag) const {
        if (w_zero | xwrap | oob) {
            if (w_zero) atomicAdd(&diag[D_W_ZERO], (unsigned long long)w_zero);
            if (xwrap) atomicAdd(&diag[D_SHADOW_XWRAP], (unsigned long long)xwrap);
            if (oob) atomicAdd(&diag[D_SHADOW_OOB], (unsigned long long)oob);
        }
    }
};

// counter += cond as one predicated add (the compiler's own rendering of `if (c) ++n` is add + predicated undo).
__device__ __forceinline__ void count_if_ge(uint32_t& n, float a, float b) {
    asm("{ .reg .pred p; setp.ge.f32 p, %1, %2; @p add.u32 %0, %0, 1; }" : "+r"(n) : "f"(a), "f"(b));
}
__device__ __forceinline__ void count_if_geu(uint32_t& n, uint32_t a, uint32_t b) {
    asm("{ .reg .pred p; setp.hs.u32 p, %1, %2; @p add.u32 %0, %0, 1; }" : "+r"(n) : "r"(a), "r"(b));
}
__device__ __forceinline__ void count_if_zero(uint32_t& n, float w) {       // w == 0.0 (either sign; NaN does not count)
    asm("{ .reg .pred p; setp.eq.f32 p, %1, 0f00000000; @p add.u32 %0, %0, 1; }" : "+r"(n) : "f"(w));
}

// `f32::round() as u32` (shader.rs:774-775): round half away from zero, then Rust's saturating cast (negative and
// NaN -> 0, >= 2^32 -> u32::MAX).  For x >= 0 that is floor(x + 0.5); the sum is formed rounding toward zero, so it
// never reaches the next integer n unless x + 0.5 >= n exactly (n < 2^24 is representable; above 2^23 x is an integer
// and the sum truncates back to x), and the truncating, saturating conversion finishes the job.  For x < 0 both sides
// give 0 (x + 0.5 < 0.5 truncates to 0 or saturates).  tr_selftest_round_u32 checks all 2^32 bit patterns on device.
__device__ __forceinline__ uint32_t round_u32(float x) { return __float2uint_rz(__fadd_rz(x, 0.5f)); }
// The literal form, kept for the self-test.
__device__ __forceinline__ uint32_t round_u32_literal(float x) {
    const float r = roundf(x);
    if (r != r || r <= 0.0f) return 0u;
    return r >= 4294967296.0f ? 0xffffffffu : (uint32_t)r;
}

// Shadow-buffer lookup of shader.rs:774-778 / 909-912 / 932-935: round, saturating cast, wrapping u32 index.
__device__ __forceinline__ float shadow_lookup(const float* shadow, f3 s, uint32_t width, uint32_t n_px, DiagLocal& dl) {
    uint32_t sx = round_u32(s.x), sy = round_u32(s.y);
    count_if_geu(dl.xwrap, sx, width);
    uint32_t idx = sx + sy * width;
    count_if_geu(dl.oob, idx, n_px);
    return __ldg(&shadow[min(idx, n_px)]);      // shadow[n_px] is a sentinel holding the cleared value (f32::MIN)
}

// A note on what was tried here in r2 and measured not to pay (profiles/README.md "AO index predictor"): of every sample
// the reference keeps only two rounded integers, so the sample's rows were predicted as (rows at wp) + (rows . step_k from
// K0), divided with one MUFU.RCP, and the nearest integer accepted only when farther than a proven per-pixel bound eps from
// the rounding boundary, recomputing the rest the reference's way: 22 instead of 45 instructions per accepted sample.  But
// the reference's own rounding noise (5 roundings per term of a row whose terms reach 4,000 at 4K, times x / w up to 3,840)
// already puts eps at 6.5e-3 px in x and 3.7e-3 in y, so 2 % of the samples are undecided — with 32 lanes and 8 samples per
// group practically every warp runs the 65-instruction recomputation loop 1-2 times per group, which eats the saving:
// k_ao_accumulate 50.1 us against 49.3, AO batch 17.64 k against 17.87 k passes/s.  Bit-identical (all AO parity tests
// passed with it), not faster; removed.
#ifndef TR_AO_GROUP
#define TR_AO_GROUP 8
#endif
constexpr int kAoGroup = TR_AO_GROUP;
static_assert(16 % kAoGroup == 0, "sample groups");
// Occlusion fragment (shader.rs:886-947) for a pixel whose depth z and world position wp = from_homogeneous(i_vpmv *
// (px, py, z, 1)) are known; returns the gray level.
__device__ __forceinline__ uint32_t occlusion_gray(const float* shadow, const Uniforms& u, int px, int py, float z, f3 wp, DiagLocal& dl) {
    const uint32_t n_px = u.width * u.height;
    bool ok;
    const f4 fh = xform(u.shadow_from_screen, f4{(float)px, (float)py, z, 1.0f});
    f3 fs = dehomog_shared(fh, ok);
    count_if_zero(dl.w_zero, fh.w);
    float fsv = shadow_lookup(shadow, fs, u.width, n_px, dl);
    float occ = 1.0f;
    const SharedDivisor d20 = shared_divisor(20.0f);        // (sv - fsv) / 20.0 below: numerators in (1, ~300)
    // The 16 sample points lie within 0.02 of wp, so their homogeneous w = W(p) (fourth row of u.shadow) lies within
    // `reach` of W(wp): |row . (p - wp)| <= rowsum * 0.0203 (0.02-long steps, plus the rounding of wp + step for
    // |wp| <= 1024) and each computed W carries at most 2^-21 * T of rounding error (T bounds the terms' magnitudes).
    // When reach <= |W(wp)| / 2 and |W(wp)| is in [2^-60, 2^60], every sample's w is in [2^-61, 2^61): non-zero and inside
    // the shared divisor's range, so the loop needs neither the range test nor the w == 0 count per sample.  The
    // test only selects between two forms of the same arithmetic (u.debug_flags bit 0 forces the tested form).
    const float* sm = u.shadow.m;
    const float reach_a = fmaxf(fmaxf(fabsf(wp.x), fabsf(wp.y)), fabsf(wp.z));
    const float w_base = fabsf(((sm[12] * wp.x + sm[13] * wp.y) + sm[14] * wp.z) + sm[15]);
    const float span = reach_a + 0.03f;
    const float reach = 0.0204f * u.ao_w_rowsum + 0x1p-20f * (u.ao_w_rowsum * span + fabsf(sm[15]));
    // The same argument bounds the samples' homogeneous x and y around X(wp), Y(wp) (rows 0 and 1), hence their
    // shadow-buffer coordinates: |x / w| <= (|X(wp)| + reach_x) / (|W(wp)| - reach).  When that stays below width - 1
    // (height - 1 for y) the rounded, saturated index parts are at most width - 1 and height - 1 — negative
    // coordinates saturate to 0, as in the reference — so no sample can wrap into the next row or leave the buffer,
    // and the loop needs neither the two diagnostics counts nor the clamp to the sentinel element.
    const float x_base = fabsf(((sm[0] * wp.x + sm[1] * wp.y) + sm[2] * wp.z) + sm[3]);
    const float y_base = fabsf(((sm[4] * wp.x + sm[5] * wp.y) + sm[6] * wp.z) + sm[7]);
    const float reach_x = 0.0204f * u.ao_x_rowsum + 0x1p-20f * (u.ao_x_rowsum * span + fabsf(sm[3]));
    const float reach_y = 0.0204f * u.ao_y_rowsum + 0x1p-20f * (u.ao_y_rowsum * span + fabsf(sm[7]));
    const float w_least = w_base - reach;
    const bool untested = reach_a <= 1024.0f && w_base >= 0x1p-60f && w_base <= 0x1p60f && reach <= 0.5f * w_base &&
                          x_base + reach_x <= (float)(u.width - 1u) * w_least && y_base + reach_y <= (float)(u.height - 1u) * w_least &&
                          (u.debug_flags & 1u) == 0u;
    auto samples = [&](auto untested_c) {
        constexpr bool UNTESTED = decltype(untested_c)::value;
        // Groups of kAoGroup samples: all shadow-buffer gathers of a group are issued before the first one is used.
        // The untested form is unrolled completely, so the 16 steps are constant-bank operands of the additions.
#pragma unroll (UNTESTED ? 16 / kAoGroup : 1)
        for (int k0 = 0; k0 < 16; k0 += kAoGroup) {
            float sv[kAoGroup];
#pragma unroll
            for (int j = 0; j < kAoGroup; ++j) {
                f3 sp = wp + u.ao_step[k0 + j];
                const f4 sh = xform(u.shadow, f4{sp.x, sp.y, sp.z, 1.0f});
                if (UNTESTED) {
                    const SharedDivisor d = shared_divisor(sh.w);
                    // z of the sample is not used
                    sv[j] = __ldg(&shadow[round_u32(quotient(sh.x, d)) + round_u32(quotient(sh.y, d)) * u.width]);
                } else {
                    f3 sc = dehomog_shared(sh, ok);
                    count_if_zero(dl.w_zero, sh.w);
                    sv[j] = shadow_lookup(shadow, sc, u.width, n_px, dl);
                }
            }
#pragma unroll
            for (int j = 0; j < kAoGroup; ++j) {
                if (sv[j] - 1.0f > fsv) {
                    float strength = fminf(quotient(sv[j] - fsv, d20), 1.0f);
                    occ = __fmaf_rn(-0.0625f, strength, occ);       // occ -= 0.0625 * strength: the product is exact
                }
            }
        }
    };
    if (untested) samples(std::true_type{}); else samples(std::false_type{});
    return cast_u8(occ * 255.0f + (1.0f - occ) * 0.0f);
}

// K6 — many-light ambient occlusion (BASELINE config 4).  The camera's visibility does not depend on the light, so
// the batch first compacts the covered pixels once (k_ao_compact: pixel coordinates, depth and world position of the
// winning fragment — everything of shader.rs:886-905 that is light-independent) and then runs one k_ao_accumulate per
// light over that list only: fully populated warps, no key / tile-state traffic, 20 bytes read per covered pixel.
constexpr int kShadeThreads = 256;

struct AoList {
    float4* frag;          // wp.x wp.y wp.z z
    uint32_t* pix;         // px | py << 16
    uint32_t* count;       // device counter (entries appended)
};

__global__ void __launch_bounds__(kShadeThreads) k_ao_compact(const ShadeArgs a, const Uniforms u, AoList out) {
    const int px = (int)(blockIdx.x * kShadeThreads + threadIdx.x), py = (int)blockIdx.y;
    bool covered = false;
    float4 frag = make_float4(0.0f, 0.0f, 0.0f, 0.0f);
    // A cleared tile has no winners and its keys are stale.  (A warp's 32 pixels lie in one raster tile.)
    if (px < (int)u.width && a.tile_state[(uint32_t)(py / kTileH) * a.tiles_x + (uint32_t)(px / kTileW)] == 0u) {
        const uint32_t lo = key_lo(__ldg(&a.keys[(size_t)py * u.width + px]));
        if (lo != kNoTri && (lo >> 24) == 0u) {
            const float4* r = a.rec + (size_t)(lo & kTriMask) * 4;
            float4 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
            float z, b0, b1, b2;
            fragment_depth(__float_as_int(r0.x), __float_as_int(r0.y), __float_as_int(r0.z), __float_as_int(r0.w),
                           __float_as_int(r1.x), __float_as_int(r1.y), r1.z, r1.w, r2.x, px, py, z, b0, b1, b2);
            bool ok;
            f3 wp = dehomog(xform(u.i_vpmv, f4{(float)px, (float)py, z, 1.0f}), ok);
            if (!ok) atomicAdd(&a.diag[D_W_ZERO], 1ull);      // from_homogeneous(i_vpmv * ..) with w == 0: shader.rs:890 panics
            frag = make_float4(wp.x, wp.y, wp.z, z);
            covered = true;
        }
    }
    const unsigned m = __ballot_sync(0xffffffffu, covered);
    if (m == 0u) return;
    const int lane = threadIdx.x & 31;
    uint32_t base = 0;
    if (lane == 0) base = atomicAdd(out.count, (uint32_t)__popc(m));
    base = __shfl_sync(0xffffffffu, base, 0);
    if (covered) {
        const uint32_t i = base + __popc(m & ((1u << lane) - 1u));
        out.frag[i] = frag;
        out.pix[i] = (uint32_t)px | ((uint32_t)py << 16);
    }
}

#ifndef TR_AO_MIN_CTAS
#define TR_AO_MIN_CTAS 6      // 40 registers; measured best with groups of 8 (tools/tune_ao.sh)
#endif
// One thread per entry of the compacted list.  (A fixed grid walking the list with a stride — which needs no count on
// the host — was measured in r2: it spills at 40 registers, and at 5 CTAs / 48 registers the kernel alone takes 58.1 us
// against 49.6; the host now learns the count behind the first light passes' depth work instead, tinyr_batch.inc.)
__global__ void __launch_bounds__(kShadeThreads, TR_AO_MIN_CTAS) k_ao_accumulate(const ShadeArgs a, const Uniforms u, const AoList in, uint32_t n) {
    const uint32_t i = blockIdx.x * kShadeThreads + threadIdx.x;
    if (i >= n) return;
    const float4 frag = __ldg(&in.frag[i]);
    const uint32_t pix = __ldg(&in.pix[i]);
    const int px = (int)(pix & 0xffffu), py = (int)(pix >> 16);
    DiagLocal dl;
    const uint32_t g = occlusion_gray(a.shadow, u, px, py, frag.w, f3{frag.x, frag.y, frag.z}, dl);
    dl.flush(a.diag);
    // integer add: the order of the light passes (which may overlap on the device) cannot change the sum
    atomicAdd(&a.ao_accum[(size_t)py * u.width + px], g);
}

}  // namespace tr
