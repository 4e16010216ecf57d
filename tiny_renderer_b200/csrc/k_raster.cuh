// k_raster.cuh — K3 raster_visibility / K4 shadow_depth: sort-middle tile rasteriser.
//
// One CTA owns one kTileW x kTileH screen tile.  (1) It takes the tile's polygon list from the bins K1+K2 filled
// (falling back to a scan of all bounding boxes when a bin overflowed) and stages per-polygon edge equations in
// shared memory.  (2) Each warp owns 8x4-pixel blocks of the tile, one pixel per lane: 32 lanes test 32 staged
// bounding boxes against the block, a ballot yields the candidate mask, and every lane evaluates the inside test
// for its own pixel with operands broadcast from shared memory.  Visibility is resolved in registers as a 64-bit
// min over packed (depth, epoch|polygon) keys — the order a global 64-bit atomicMin would give, without the L2
// round trips — and each tile is written once, coalesced.  Tiles no polygon touches only write the cleared value.
//
// Arithmetic: the reference evaluates, per pixel, c = u x v in f32 and bar = (1-(cx+cy)/cz, cx/cz, cy/cz)
// (/root/reference/src/scene.rs:174-197), rejects bar.* < 0 (:245) and takes z = bar . z (shader.rs:174).  When
// the polygon's unclamped bounding box is at most kExactExtent pixels wide and high every product and sum in c
// is an integer below 2^24, hence exact in f32, and `bar.* >= 0` is equivalent to the integer edge tests used
// here (proof in DESIGN.md §K3); the divisions and z are then done in f32 exactly as the reference does, for
// inside pixels only.  Larger polygons take the literal f32 path (fragment_depth).
//
// Replaces scene.rs:232-258 and process_z_value (/root/reference/src/scene/shader.rs:169-180); in DEPTH_ONLY
// mode the pass-1 fragment shaders of shadow / occlusion (shader.rs:694-709, 832-847).
#pragma once
#include "tr_types.cuh"

namespace tr {

constexpr int kTileW = 64, kTileH = 32;
constexpr int kRasterThreads = 256;
constexpr int kListCap = 512;          // staged polygons per round; longer lists are processed in rounds
constexpr int kBinCap = 256;           // ids per tile bin (K2); a fuller bin makes the tile scan all polygons
constexpr int kStageWords = 12;        // A1 B1 C1 A2 B2 C2 |cz| id|flag z0 z1 z2 pad   (or raw x0 y0 x1 y1 x2 y2 when inexact)
constexpr int kExactExtent = 2048;
constexpr uint32_t kInexactFlag = 0x80000000u;

struct Bins {
    uint32_t* count;                   // per tile; reset to 0 by the raster kernel that consumes it
    uint32_t* list;                    // per tile kBinCap polygon ids (unordered)
    uint32_t tiles_x, tiles_y;
};

struct RasterArgs {
    const float4* rec;
    const int4* bbox;
    Bins bins;
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

// Stage polygon t into slot: edge equations relative to the tile origin, or raw coordinates when inexact.
__device__ __forceinline__ void stage_polygon(const RasterArgs& a, uint32_t t, int4 bb, int slot, int tx0, int ty0, int tx1, int ty1,
                                              float* s_tri, int* s_lo, int* s_hi) {
    const float4* r = a.rec + (size_t)t * 4;
    float4 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
    int x0 = __float_as_int(r0.x), y0 = __float_as_int(r0.y), x1 = __float_as_int(r0.z), y1 = __float_as_int(r0.w);
    int x2 = __float_as_int(r1.x), y2 = __float_as_int(r1.y);
    long long ex = (long long)max(max(x0, x1), x2) - min(min(x0, x1), x2), ey = (long long)max(max(y0, y1), y2) - min(min(y0, y1), y2);
    int4* d = reinterpret_cast<int4*>(s_tri + slot * kStageWords);
    int lo = ((max(bb.x, tx0) - tx0) << 16) | (max(bb.y, ty0) - ty0);
    int hi = ((min(bb.z, tx1) - tx0) << 16) | (min(bb.w, ty1) - ty0);
    uint32_t id = t;
    if (ex <= kExactExtent && ey <= kExactExtent) {
        int ux = x1 - x0, uy = x2 - x0, vx = y1 - y0, vy = y2 - y0;
        int cz = ux * vy - uy * vx;
        int s = cz < 0 ? -1 : 1;
        int rx = x0 - tx0, ry = y0 - ty0;
        // cx(lx,ly) = uy*(y0-py) - (x0-px)*vy ,  cy(lx,ly) = (x0-px)*vx - ux*(y0-py)   with px = tx0+lx, py = ty0+ly
        d[0] = make_int4(s * (uy * ry - rx * vy), s * vy, -s * uy, s * (rx * vx - ux * ry));
        d[1] = make_int4(-s * vx, s * ux, s * cz, (int)id);
        if (cz == 0) { lo = 0x7fff7fff; hi = 0; }       // |cz| < 1: degenerate, never produces a fragment (scene.rs:188)
    } else {
        d[0] = make_int4(x0, y0, x1, y1);
        d[1] = make_int4(x2, y2, 0, (int)(id | kInexactFlag));
    }
    d[2] = make_int4(__float_as_int(r1.z), __float_as_int(r1.w), __float_as_int(r2.x), 0);
    s_lo[slot] = lo;
    s_hi[slot] = hi;
}

template <bool DEPTH_ONLY>
__global__ void __launch_bounds__(kRasterThreads) k_raster(const RasterArgs a) {
    __shared__ __align__(16) float s_tri[kListCap * kStageWords];
    __shared__ int s_lo[kListCap], s_hi[kListCap];
    __shared__ int s_warp_count[kRasterThreads / 32];
    __shared__ int s_count;
    __shared__ int s_next;

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int tx0 = blockIdx.x * kTileW, ty0 = blockIdx.y * kTileH;
    const int tx1 = min(tx0 + kTileW, (int)a.width) - 1, ty1 = min(ty0 + kTileH, (int)a.height) - 1;
    const float kFloatMin = -3.402823466e38f;
    constexpr int kBlocksX = kTileW / 8;
    constexpr int kBlocksPerWarp = kBlocksX * (kTileH / 4) / (kRasterThreads / 32);

    const uint32_t tile = blockIdx.y * a.bins.tiles_x + blockIdx.x;
    const uint32_t n_bin = a.bins.count[tile];
    __syncthreads();
    if (tid == 0) a.bins.count[tile] = 0;                 // self-cleaning: ready for the next K1+K2

    if (n_bin == 0) {
        // Nothing touches this tile: a cleared target gets the cleared value, an uncleared one keeps its contents.
        if (!a.fresh) return;
#pragma unroll
        for (int b = 0; b < kBlocksPerWarp; ++b) {
            int blk = warp * kBlocksPerWarp + b;
            int px = tx0 + (blk % kBlocksX) * 8 + (lane & 7), py = ty0 + (blk / kBlocksX) * 4 + (lane >> 3);
            if (px < (int)a.width && py < (int)a.height) {
                if (DEPTH_ONLY) a.depth[(size_t)py * a.width + px] = kFloatMin;
                else a.keys[(size_t)py * a.width + px] = make_key(kFloatMin, kNoTri);
            }
        }
        return;
    }

    unsigned long long best[kBlocksPerWarp];
    float bestz[kBlocksPerWarp];
#pragma unroll
    for (int b = 0; b < kBlocksPerWarp; ++b) {
        int blk = warp * kBlocksPerWarp + b;
        int px = tx0 + (blk % kBlocksX) * 8 + (lane & 7), py = ty0 + (blk / kBlocksX) * 4 + (lane >> 3);
        bool in = px < (int)a.width && py < (int)a.height;
        if (DEPTH_ONLY) bestz[b] = (a.fresh || !in) ? kFloatMin : a.depth[(size_t)py * a.width + px];
        else best[b] = (a.fresh || !in) ? make_key(kFloatMin, kNoTri) : a.keys[(size_t)py * a.width + px];
    }

    const bool use_bin = n_bin <= (uint32_t)kBinCap;
    uint32_t scan_from = 0;
    bool more = true;
    while (more) {
        int n_list;
        if (use_bin) {
            // ---- (1a) the bin holds the tile's polygons ----
            for (int i = tid; i < (int)n_bin; i += kRasterThreads) {
                uint32_t t = a.bins.list[(size_t)tile * kBinCap + i];
                stage_polygon(a, t, __ldg(&a.bbox[t]), i, tx0, ty0, tx1, ty1, s_tri, s_lo, s_hi);
            }
            n_list = (int)n_bin;
            more = false;
            __syncthreads();
        } else {
            // ---- (1b) overflowed bin: scan all bounding boxes, kListCap polygons per round ----
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
                if (hit && slot < kListCap) stage_polygon(a, t, bb, slot, tx0, ty0, tx1, ty1, s_tri, s_lo, s_hi);
                if (hit && slot >= kListCap) atomicMin(&s_next, (int)t);   // first polygon that did not fit
                __syncthreads();
                if (tid == 0) s_count = min(s_count + total, kListCap);
                __syncthreads();
                if (overflow) break;
            }
            n_list = s_count;
            scan_from = (uint32_t)s_next;   // == n_triangles when everything was staged
            more = scan_from < a.n_triangles;
        }

        // ---- (2) per-warp blocks: ballot the staged bboxes against the block, evaluate per pixel ----
#pragma unroll
        for (int b = 0; b < kBlocksPerWarp; ++b) {
            int blk = warp * kBlocksPerWarp + b;
            int bx0 = (blk % kBlocksX) * 8, by0 = (blk / kBlocksX) * 4;   // tile-relative
            int lx = bx0 + (lane & 7), ly = by0 + (lane >> 3);
            for (int base = 0; base < n_list; base += 32) {
                int i = base + lane;
                bool cand = false;
                if (i < n_list) {
                    int lo = s_lo[i], hi = s_hi[i];
                    cand = (lo >> 16) <= bx0 + 7 && (hi >> 16) >= bx0 && (lo & 0xffff) <= by0 + 3 && (hi & 0xffff) >= by0;
                }
                unsigned m = __ballot_sync(0xffffffffu, cand);
                while (m) {
                    int j = base + __ffs(m) - 1;
                    m &= m - 1;
                    const int4* d = reinterpret_cast<const int4*>(s_tri + j * kStageWords);
                    int4 e = d[0], f = d[1];
                    float z;
                    if (!((uint32_t)f.w & kInexactFlag)) {
                        int e1 = e.x + lx * e.y + ly * e.z;          // cx * sign(cz)
                        int e2 = e.w + lx * f.x + ly * f.y;          // cy * sign(cz)
                        int e0 = f.z - e1 - e2;
                        if ((e0 | e1 | e2) < 0) continue;
                        int4 g = d[2];
                        float cx = (float)e1, cy = (float)e2, cz = (float)f.z;
                        float b0 = 1.0f - (cx + cy) / cz, b1 = cx / cz, b2 = cy / cz;
                        z = (b0 * __int_as_float(g.x) + b1 * __int_as_float(g.y)) + b2 * __int_as_float(g.z);
                    } else {
                        int4 g = d[2];
                        float b0, b1, b2;
                        if (!fragment_depth(e.x, e.y, e.z, e.w, f.x, f.y, __int_as_float(g.x), __int_as_float(g.y), __int_as_float(g.z),
                                            tx0 + lx, ty0 + ly, z, b0, b1, b2)) continue;
                    }
                    if (z != z) { atomicAdd(&a.diag[D_NAN_DEPTH], 1ull); continue; }
                    if (DEPTH_ONLY) {
                        if (z >= bestz[b]) bestz[b] = z;                         // shader.rs:703
                    } else {
                        unsigned long long k = make_key(z, (a.epoch << 24) | ((uint32_t)f.w & kTriMask));
                        if (key_beats(k, best[b])) best[b] = k;                  // shader.rs:175
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
        int px = tx0 + (blk % kBlocksX) * 8 + (lane & 7), py = ty0 + (blk / kBlocksX) * 4 + (lane >> 3);
        if (px < (int)a.width && py < (int)a.height) {
            if (DEPTH_ONLY) a.depth[(size_t)py * a.width + px] = bestz[b];
            else a.keys[(size_t)py * a.width + px] = best[b];
        }
    }
}

}  // namespace tr
