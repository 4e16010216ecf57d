// k_raster.cuh — K3 raster_visibility / K4 shadow_depth: sort-middle tile rasteriser.
//
// One CTA owns one TILE_W x TILE_H screen tile.  (1) The CTA scans the polygon bounding boxes (K1 output) and
// stages the records of the polygons touching its tile in shared memory, in polygon order.  (2) Each warp owns
// 8x4-pixel blocks of the tile, one pixel per lane: 32 lanes test 32 staged bounding boxes against the block,
// a ballot yields the candidate mask, and every lane evaluates the reference's barycentric / inside / depth
// arithmetic for its own pixel with the operands broadcast from shared memory.  Visibility is resolved in
// registers as a 64-bit min over packed (depth, epoch|polygon) keys — the same order the global 64-bit
// atomicMin formulation gives, without the L2 round trips — and written once, coalesced.
//
// Replaces /root/reference/src/scene.rs:232-258 (bbox walk, to_barycentric_coord :174-197, inside test :245)
// and process_z_value (/root/reference/src/scene/shader.rs:169-180); in DEPTH_ONLY mode the pass-1 fragment
// shaders of shadow / occlusion (shader.rs:694-709, 832-847).
#pragma once
#include "tr_types.cuh"

namespace tr {

constexpr int kTileW = 64, kTileH = 32;
constexpr int kRasterThreads = 256;
constexpr int kListCap = 768;          // staged polygons per round; longer lists are processed in rounds
constexpr int kStageWords = 12;        // x0 y0 x1 y1 x2 y2 z0 z1 z2 id xmin|ymin xmax|ymax  (bbox packed 16:16, tile-relative)

struct RasterArgs {
    const float4* rec;
    const int4* bbox;
    uint32_t n_triangles;
    uint32_t width, height;
    uint32_t epoch;
    int fresh;                 // 1: target was cleared (init from constants, scene.rs:128-137), 0: merge with memory
    unsigned long long* keys;  // visibility mode
    float* depth;              // DEPTH_ONLY mode (shadow buffer)
    unsigned long long* diag;
};

// The reference's per-pixel arithmetic, literally (scene.rs:178-196, shader.rs:174).  Returns false when the
// pixel is outside the polygon or the polygon is degenerate.
__device__ __forceinline__ bool fragment_depth(int x0, int y0, int x1, int y1, int x2, int y2, float z0, float z1, float z2,
                                               int px, int py, float& z, float& b0, float& b1, float& b2) {
    float ux = (float)(x1 - x0), uy = (float)(x2 - x0), uz = (float)(x0 - px);
    float vx = (float)(y1 - y0), vy = (float)(y2 - y0), vz = (float)(y0 - py);
    float cx = uy * vz - uz * vy;
    float cy = uz * vx - ux * vz;
    float cz = ux * vy - uy * vx;
    if (fabsf(cz) < 1.0f) return false;
    b0 = 1.0f - (cx + cy) / cz;
    b1 = cx / cz;
    b2 = cy / cz;
    if (b0 < 0.0f || b1 < 0.0f || b2 < 0.0f) return false;
    z = (b0 * z0 + b1 * z1) + b2 * z2;
    return true;
}

template <bool DEPTH_ONLY>
__global__ void __launch_bounds__(kRasterThreads) k_raster(const RasterArgs a) {
    __shared__ float s_tri[kListCap * kStageWords];
    __shared__ int s_warp_count[kRasterThreads / 32];
    __shared__ int s_count;
    __shared__ int s_next;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx0 = blockIdx.x * kTileW, ty0 = blockIdx.y * kTileH;
    const int tx1 = min(tx0 + kTileW, (int)a.width) - 1, ty1 = min(ty0 + kTileH, (int)a.height) - 1;
    const float kFloatMin = -3.402823466e38f;

    // Per-lane register state for the 8 blocks this warp owns: block b of the tile is (b & 7, b >> 3) in units
    // of 8x4 pixels; warp w owns row w of blocks (8 blocks across the 64-pixel tile width).
    constexpr int kBlocksPerWarp = (kTileW / 8) * (kTileH / 4) / (kRasterThreads / 32);
    unsigned long long best[kBlocksPerWarp];
    float bestz[kBlocksPerWarp];
#pragma unroll
    for (int b = 0; b < kBlocksPerWarp; ++b) {
        int blk = warp * kBlocksPerWarp + b;
        int px = tx0 + (blk % (kTileW / 8)) * 8 + (lane & 7), py = ty0 + (blk / (kTileW / 8)) * 4 + (lane >> 3);
        bool in = px < (int)a.width && py < (int)a.height;
        if (DEPTH_ONLY) bestz[b] = (a.fresh || !in) ? kFloatMin : a.depth[(size_t)py * a.width + px];
        else best[b] = (a.fresh || !in) ? make_key(kFloatMin, kNoTri) : a.keys[(size_t)py * a.width + px];
    }

    uint32_t scan_from = 0;
    while (scan_from < a.n_triangles) {
        // ---- (1) stage the next <= kListCap polygons whose bbox touches the tile, in polygon order ----
        if (tid == 0) { s_count = 0; s_next = (int)a.n_triangles; }
        __syncthreads();
        for (uint32_t base = scan_from; base < a.n_triangles; base += kRasterThreads) {
            uint32_t t = base + tid;
            bool hit = false;
            int4 bb = make_int4(1, 0, 0, 0);
            if (t < a.n_triangles) {
                bb = __ldg(&a.bbox[t]);
                hit = bb.x <= bb.z && bb.x <= tx1 && bb.z >= tx0 && bb.y <= ty1 && bb.w >= ty0;
            }
            unsigned m = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) s_warp_count[warp] = __popc(m);
            __syncthreads();
            int before = s_count, total = 0;
            for (int w = 0; w < kRasterThreads / 32; ++w) { int c = s_warp_count[w]; if (w < warp) before += c; total += c; }
            int slot = before + __popc(m & ((1u << lane) - 1u));
            bool overflow = s_count + total > kListCap;
            if (hit && slot < kListCap) {
                const float4* r = a.rec + (size_t)t * 4;
                float4 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
                float* d = s_tri + slot * kStageWords;
                d[0] = r0.x; d[1] = r0.y; d[2] = r0.z; d[3] = r0.w; d[4] = r1.x; d[5] = r1.y;
                d[6] = r1.z; d[7] = r1.w; d[8] = r2.x;
                d[9] = __int_as_float((int)t);
                d[10] = __int_as_float(((max(bb.x, tx0) - tx0) << 16) | (max(bb.y, ty0) - ty0));
                d[11] = __int_as_float(((min(bb.z, tx1) - tx0) << 16) | (min(bb.w, ty1) - ty0));
            }
            if (hit && slot >= kListCap) atomicMin(&s_next, (int)t);   // first polygon that did not fit
            __syncthreads();
            if (tid == 0) s_count = min(s_count + total, kListCap);
            __syncthreads();
            if (overflow) break;
        }
        const int n_list = s_count;
        scan_from = (uint32_t)s_next;   // == n_triangles when everything was staged

        // ---- (2) per-warp blocks: ballot the staged bboxes against the block, evaluate per pixel ----
#pragma unroll
        for (int b = 0; b < kBlocksPerWarp; ++b) {
            int blk = warp * kBlocksPerWarp + b;
            int bx0 = (blk % (kTileW / 8)) * 8, by0 = (blk / (kTileW / 8)) * 4;   // tile-relative
            int px = tx0 + bx0 + (lane & 7), py = ty0 + by0 + (lane >> 3);
            for (int base = 0; base < n_list; base += 32) {
                int i = base + lane;
                bool cand = false;
                if (i < n_list) {
                    int lo = __float_as_int(s_tri[i * kStageWords + 10]), hi = __float_as_int(s_tri[i * kStageWords + 11]);
                    cand = (lo >> 16) <= bx0 + 7 && (hi >> 16) >= bx0 && (lo & 0xffff) <= by0 + 3 && (hi & 0xffff) >= by0;
                }
                unsigned m = __ballot_sync(0xffffffffu, cand);
                while (m) {
                    int j = base + __ffs(m) - 1;
                    m &= m - 1;
                    const float* d = s_tri + j * kStageWords;
                    float z, b0, b1, b2;
                    if (fragment_depth(__float_as_int(d[0]), __float_as_int(d[1]), __float_as_int(d[2]), __float_as_int(d[3]),
                                       __float_as_int(d[4]), __float_as_int(d[5]), d[6], d[7], d[8], px, py, z, b0, b1, b2)) {
                        if (z != z) { atomicAdd(&a.diag[D_NAN_DEPTH], 1ull); continue; }
                        if (DEPTH_ONLY) {
                            if (z >= bestz[b]) bestz[b] = z;                         // shader.rs:703
                        } else {
                            unsigned long long k = make_key(z, (a.epoch << 24) | (uint32_t)__float_as_int(d[9]));
                            if (key_beats(k, best[b])) best[b] = k;                  // shader.rs:175
                        }
                    }
                }
            }
        }
        __syncthreads();
    }

    // ---- (3) write the tile once ----
#pragma unroll
    for (int b = 0; b < kBlocksPerWarp; ++b) {
        int blk = warp * kBlocksPerWarp + b;
        int px = tx0 + (blk % (kTileW / 8)) * 8 + (lane & 7), py = ty0 + (blk / (kTileW / 8)) * 4 + (lane >> 3);
        if (px < (int)a.width && py < (int)a.height) {
            if (DEPTH_ONLY) a.depth[(size_t)py * a.width + px] = bestz[b];
            else a.keys[(size_t)py * a.width + px] = best[b];
        }
    }
}

}  // namespace tr
