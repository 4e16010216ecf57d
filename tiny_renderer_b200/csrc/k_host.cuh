// k_host.cuh — K10 frame_to_host: sparse, damage-tracked delivery of a finished RGB8 frame into HOST memory.
//
// Replaces the full-frame copy behind Scene::get_frame_buffer (/root/reference/src/scene.rs:92-97: clone + flip of
// 3*W*H bytes per frame) for destinations the library has allocated itself (tr_host_alloc: pinned, mapped, zeroed,
// written by nobody else).  82 % of a 4K diablo frame is black and the device knows where: per raster tile the
// rgb_zero flag of k_raster, per tile row a ballot over the 192 bytes of the row.  The kernel stores, straight over
// PCIe into the destination frame,
//     * every tile row that holds a non-zero byte now, and
//     * zeros into every tile row that held one after the previous delivery to this destination but is black now,
// and nothing else.  What the destination held before is known exactly: a 32-bit row mask per tile, kept on the device
// per destination frame (host-frame registry in tinyr_batch.inc), updated here.  The caller therefore always finds
// the exact top-down frame get_frame_buffer returns; only the bytes that changed or are non-zero cross the link.
// No staging buffer, no host-side scatter: the host CPU does not touch a pixel.
//
// Rows of a tile are 192 bytes (64 px) at a 64-byte aligned offset, written as 12 x 16-byte stores by 12 adjacent lanes.
// (A form with one warp per strip of four adjacent tiles — store instructions covering up to 512 contiguous bytes, so that
// everything leaves as full 128-byte lines — was measured in r2 and is no faster: 62.8 / 64.9 / 66.8 Gpx/s end to end with
// 8 / 16 / 32 CTAs against 65.3 / 63.6 / 65.0 for this form on the same box; SM-issued stores to host memory top out near
// 47 GB/s here whatever their shape, a DMA copy reaches 54-57.)
#pragma once
#include <cstdint>
#include <cuda_runtime.h>
#include "k_tile.cuh"

namespace tr {

constexpr int kHostThreads = 128;

struct HostStoreArgs {
    const uint8_t* rgb;          // device frame, top-down rows, pitch = width * 3 (a multiple of 16)
    uint8_t* dst;                // host frame (device-visible address of pinned memory), same layout
    uint32_t* row_mask;          // per raster tile: bit r = tile row r (bottom-up inside the tile) of dst holds a non-zero byte
    const uint32_t* rgb_zero;    // per raster tile: nonzero = the tile of rgb is all zeros (k_raster.cuh)
    uint32_t width, height, tiles_x, n_tiles;
    unsigned long long* stats;   // [0] bytes stored to the host, [1] tile rows stored
};

__device__ __forceinline__ void st_host_v4(uint4* p, uint4 v) {
    // streaming store towards the link; the line is never read again on the device
    asm volatile("st.global.cs.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

static_assert(kTileW * 3 == 12 * 16, "a tile row is 12 16-byte columns");
// One warp per tile (grid-stride over the tiles); lane l < 24 handles 16-byte column l % 12 of row 2*i + l / 12 in
// step i.  All loads of a lane are issued before the first ballot.
__global__ void __launch_bounds__(kHostThreads) k_frame_to_host(const HostStoreArgs a) {
    const int lane = threadIdx.x & 31;
    const uint32_t warp = (blockIdx.x * kHostThreads + threadIdx.x) >> 5, n_warps = gridDim.x * kHostThreads / 32;
    const size_t pitch = (size_t)a.width * 3;
    const int sub = lane / 12, col = lane % 12;                 // lanes 24..31 idle in the row steps
    unsigned long long rows_stored = 0, bytes_stored = 0;
    for (uint32_t tile = warp; tile < a.n_tiles; tile += n_warps) {
        const uint32_t prev = a.row_mask[tile];
        const bool zero_tile = a.rgb_zero[tile] != 0u;
        if (zero_tile && prev == 0u) continue;
        const uint32_t tx = tile % a.tiles_x, ty = tile / a.tiles_x;
        const int tx0 = (int)tx * kTileW, ty0 = (int)ty * kTileH;
        const int rows = min(kTileH, (int)a.height - ty0);
        const int seg16 = (min(kTileW, (int)a.width - tx0) * 3) / 16;          // 16-byte columns of a row segment (<= 12)
        const bool active = sub < 2 && col < seg16;
        // tile row r (bottom-up) = frame row height-1-(ty0+r) (top-down)
        uint4 v[kTileH / 2];
        uint32_t now = 0u;
        if (!zero_tile) {
#pragma unroll
            for (int i = 0; i < kTileH / 2; ++i) {
                const int r = 2 * i + sub;
                v[i] = make_uint4(0u, 0u, 0u, 0u);
                if (active && r < rows)
                    v[i] = __ldg(reinterpret_cast<const uint4*>(a.rgb + (size_t)(a.height - 1 - (ty0 + r)) * pitch + (size_t)tx0 * 3) + col);
            }
#pragma unroll
            for (int i = 0; i < kTileH / 2; ++i) {
                const unsigned b = __ballot_sync(0xffffffffu, (v[i].x | v[i].y | v[i].z | v[i].w) != 0u);
                if (b & 0x00000fffu) now |= 1u << (2 * i);
                if (b & 0x00fff000u) now |= 1u << (2 * i + 1);
            }
        }
        const uint32_t write = now | prev;          // rows to store: non-zero now, or non-zero before (then zeros)
#pragma unroll
        for (int i = 0; i < kTileH / 2; ++i) {
            const int r = 2 * i + sub;
            if (active && r < rows && ((write >> r) & 1u)) {
                const uint4 out = zero_tile ? make_uint4(0u, 0u, 0u, 0u) : v[i];
                st_host_v4(reinterpret_cast<uint4*>(a.dst + (size_t)(a.height - 1 - (ty0 + r)) * pitch + (size_t)tx0 * 3) + col, out);
            }
        }
        if (lane == 0) {
            if (now != prev) a.row_mask[tile] = now;
            const int n = __popc(write & (rows >= 32 ? 0xffffffffu : ((1u << rows) - 1u)));
            rows_stored += n;
            bytes_stored += (unsigned long long)n * seg16 * 16;
        }
    }
    if (lane == 0 && rows_stored) {
        atomicAdd(&a.stats[0], bytes_stored);
        atomicAdd(&a.stats[1], rows_stored);
    }
}

}  // namespace tr
