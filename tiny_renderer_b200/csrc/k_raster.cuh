// k_raster.cuh — K3 raster_visibility (+ K5 shading, fused) / K4 shadow_depth: sort-middle tile rasteriser.
//
// One CTA owns one kTileW x kTileH screen tile.  (1) It filters the polygon list of the binning cell the tile lies
// in (K1+K2; a scan of all bounding boxes when the cell's bin overflowed) against the tile and stages per-polygon
// edge equations in shared memory, together with a 32-bit mask of the 8x8-pixel blocks of the tile the triangle
// can reach (bounding box, narrowed by the three edge functions' maxima per block).  (2) Warps pull blocks off the
// tile, two pixels per lane: 32 lanes test the block's bit of 32 masks, a ballot yields the candidates, and every lane evaluates the
// inside test for its own pixel with operands broadcast from shared memory.  Visibility is resolved in registers —
// (depth, epoch|polygon) compared the way a 64-bit atomicMin on the packed key would order them, without the L2
// round trips — and the key of every pixel of a touched tile is written once.  (3) In the shading modes the lane
// then runs the pipeline's fragment math for its winner (k_shade.cuh) with the barycentrics it already holds,
// the tile's RGB is collected in shared memory and stored with 16-byte writes.  Tiles no polygon touches cost one
// word of tile state (see RasterArgs::tile_state) plus, when shading, their zeroed RGB.
//
// Arithmetic: the reference evaluates, per pixel, c = u x v in f32 and bar = (1-(cx+cy)/cz, cx/cz, cy/cz)
// (/root/reference/src/scene.rs:174-197), rejects bar.* < 0 (:245) and takes z = bar . z (shader.rs:174).  When
// the polygon's unclamped bounding box is at most kExactExtent pixels wide and high every product and sum in c
// is an integer below 2^24, hence exact in f32, and `bar.* >= 0` is equivalent to the integer edge tests used
// here (proof in DESIGN.md §4); the three quotients are formed by exact_quotient() (k_tile.cuh: bit-identical to the
// IEEE division for these integer operands) and z in f32 exactly as the reference does, for inside pixels only.  Larger polygons take the literal f32 path (fragment_depth).
//
// Replaces scene.rs:232-258 and process_z_value (/root/reference/src/scene/shader.rs:169-180) plus the fragment
// closures (k_shade.cuh); in depth-only mode the pass-1 fragment shaders of shadow / occlusion
// (shader.rs:694-709, 832-847).
#pragma once
#include <type_traits>
#include <cuda.h>               // CUtensorMap
#include "pipeline.cuh"

namespace tr {

// Two CTA shapes of the same kernel (template parameter THREADS, measured in profiles/README.md):
//   * 256 threads, 512-polygon lists, 6 CTAs per SM (40 registers, no spills; also the shared-memory limit; pipelines
//     with heavy_fragment set — darboux, whose fragment code spills below 64 registers: 4) — a lone frame: the heavy tiles set the frame's duration, and
//     eight warps finish one sooner than four;
//   * 128 threads, 192-polygon lists, 12 CTAs per SM (darboux: 8) — overlapped batches (view batches, the many-light
//     AO batch): the same number of resident warps in smaller barrier groups, and a tile's ragged tail idles four
//     warps instead of eight; the longer tiles no longer matter because the next frame's work fills in behind them.
#ifndef TR_MIN_CTAS_LIGHT
#define TR_MIN_CTAS_LIGHT 6
#endif
#ifndef TR_MIN_CTAS_DARBOUX
#define TR_MIN_CTAS_DARBOUX 4
#endif
#ifndef TR_RASTER_THREADS_BATCH
#define TR_RASTER_THREADS_BATCH 128
#endif
#ifndef TR_LIST_CAP_BATCH
#define TR_LIST_CAP_BATCH 192
#endif
#ifndef TR_MIN_CTAS_LIGHT_BATCH
#define TR_MIN_CTAS_LIGHT_BATCH 12
#endif
#ifndef TR_MIN_CTAS_DARBOUX_BATCH
#define TR_MIN_CTAS_DARBOUX_BATCH 8
#endif
constexpr int kRasterThreadsBatch = TR_RASTER_THREADS_BATCH;
__host__ __device__ constexpr int raster_list_cap(int threads) { return threads == kRasterThreads ? kListCap : TR_LIST_CAP_BATCH; }
__host__ __device__ constexpr int raster_min_ctas(int threads, bool darboux) {
    return threads == kRasterThreads ? (darboux ? TR_MIN_CTAS_DARBOUX : TR_MIN_CTAS_LIGHT)
                                     : (darboux ? TR_MIN_CTAS_DARBOUX_BATCH : TR_MIN_CTAS_LIGHT_BATCH);
}
#define TR_RASTER_BOUNDS __launch_bounds__(THREADS, raster_min_ctas(THREADS, PIPE::heavy_fragment))
#ifndef TR_PIXELS_PER_LANE
#define TR_PIXELS_PER_LANE 2
#endif

// Kernel modes: PIPE is a pipeline struct (pipelines/*.cuh: visibility + its `fragment` role) or one of the two tags of
// pipeline.cuh, PassLightDepth (K4) / PassVisibility (K3 alone).

struct RasterArgs {
    const float4* rec;
    const int4* bbox;
    Bins bins;
    uint32_t n_triangles;
    uint32_t width, height;
    uint32_t tiles_x;
    uint32_t epoch;
    int fresh;                 // 1: target was cleared (init from constants, scene.rs:128-137), 0: merge with memory
    unsigned long long* keys;  // visibility modes
    // Visibility modes, per tile: 1 = every pixel of the tile holds the cleared value and the key memory of the tile
    // is stale.  Scene::clear therefore costs nothing for tiles no polygon touches; K6 and the getters never read
    // their keys.
    uint32_t* tile_state;
    const uint32_t* tile_order;   // launch position -> tile coordinates, y << 16 | x
    float* depth;              // depth-only mode (shadow buffer)
    uint32_t* depth_clear;     // depth-only mode, per tile: nonzero = every depth of the tile holds the cleared value already
    unsigned long long* diag;
    uint32_t n_tiles;
    // Lone-frame launches: per tile (under its position in the static centre-out order), the SM cycles this launch spent on it.  The vertex
    // set-up kernel of the NEXT lone frame sorts the launch order by it, longest first (k_setup.cuh: sort_tiles).  Non-null
    // also means that tile_order was written by the set-up kernel in front of this launch, i.e. is read after the wait.
    // Lone frames: the diagnostics block of the set-up kernel run whose output this launch reads (possibly a run of an
    // earlier frame of the same view, k_setup.cuh SetupOut::diag_zero); folded into diag once per launch.  Null: none.
    const unsigned long long* setup_diag;
    const unsigned long long* light_diag;  // the same for the light pass (set-up + depth raster) whose shadow buffer this launch reads
    uint32_t* tile_cost;
    const uint32_t* tile_perm; // launch position -> position in the static centre-out order (the index tile_cost is kept under); null: the same
};

// The tile is cut into kBlocksX x kBlocksY blocks of 8 x kBlockH pixels (one warp pass each); bit (by * kBlocksX + bx)
// of a block mask stands for block (bx, by).
constexpr int kPix = TR_PIXELS_PER_LANE;                // pixels per lane in the block loop (1 or 2), 4 rows apart
constexpr int kBlockH = 4 * kPix;
constexpr int kBlocksX = kTileW / 8;
constexpr int kBlocksY = kTileH / kBlockH;
constexpr int kBlocks = kBlocksX * kBlocksY;
static_assert(kPix == 1 || kPix == 2, "the fragment stage selects between at most two pixels per lane");
static_assert(kTileH % kBlockH == 0 && kTileW % 8 == 0, "tiles must hold whole blocks");
static_assert(kBlocks <= 32, "one 32-bit block mask per staged polygon");

// Blocks a tile-relative, clamped, inclusive bounding box overlaps.
__device__ __forceinline__ uint32_t bbox_block_mask(int lx0, int ly0, int lx1, int ly1) {
    const int bx0 = lx0 >> 3, bx1 = lx1 >> 3, by0 = ly0 / kBlockH, by1 = ly1 / kBlockH;
    uint32_t rep = 0;
#pragma unroll
    for (int r = 0; r < kBlocksY; ++r) rep |= 1u << (r * kBlocksX);
    const uint32_t cols = (2u << bx1) - (1u << bx0);
    const int lo = kBlocksX * by0, hi = kBlocksX * (by1 + 1) - 1;
    return (cols * rep) & (0xffffffffu >> (31 - hi)) & (0xffffffffu << lo);
}

// Blocks that can hold an inside pixel of an exact-path polygon: with e1, e2 the staged edge functions and
// e0 = cz - e1 - e2, a pixel is inside iff all three are >= 0 (see the sweep below); each is linear in the pixel, so
// its largest value over a block sits at one corner, and a block where any of the three maxima is negative holds no
// inside pixel.  Integer arithmetic throughout (|values| < 2^23), hence conservative AND exact.
__device__ __forceinline__ uint32_t edge_block_mask(int4 e, int4 f) {
    const int a1 = e.y, b1 = e.z, a2 = f.x, b2 = f.y, a0 = -(a1 + a2), b0 = -(b1 + b2);
    const int m0 = (f.z - e.x - e.w) + 7 * max(a0, 0) + (kBlockH - 1) * max(b0, 0);
    const int m1 = e.x + 7 * max(a1, 0) + (kBlockH - 1) * max(b1, 0);
    const int m2 = e.w + 7 * max(a2, 0) + (kBlockH - 1) * max(b2, 0);
    uint32_t outside = 0;                   // one bit per block pushed in from the right: block 0 ends up leftmost
#pragma unroll
    for (int by = 0; by < kBlocksY; ++by) {
        int v0 = m0 + by * kBlockH * b0, v1 = m1 + by * kBlockH * b1, v2 = m2 + by * kBlockH * b2;
#pragma unroll
        for (int bx = 0; bx < kBlocksX; ++bx) {
            outside = __funnelshift_l((uint32_t)(v0 | v1 | v2), outside, 1);      // sign bit: some maximum is negative
            v0 += 8 * a0; v1 += 8 * a1; v2 += 8 * a2;
        }
    }
    return ~(__brev(outside) >> (32 - kBlocks));
}

// Stage polygon t into slot: edge equations relative to the tile origin, or raw coordinates when inexact.  s_mask /
// s_maskx receive the blocks its bounding box overlaps (exact path / literal path; the other one is 0).
__device__ __forceinline__ void stage_polygon(uint32_t t, int4 bb, float4 r0, float4 r1, float4 r2, int slot, int tx0, int ty0, int tx1, int ty1,
                                              float* s_tri, uint32_t* s_mask, uint32_t* s_maskx, int* s_inexact) {
    int x0 = __float_as_int(r0.x), y0 = __float_as_int(r0.y), x1 = __float_as_int(r0.z), y1 = __float_as_int(r0.w);
    int x2 = __float_as_int(r1.x), y2 = __float_as_int(r1.y);
    long long ex = (long long)max(max(x0, x1), x2) - min(min(x0, x1), x2), ey = (long long)max(max(y0, y1), y2) - min(min(y0, y1), y2);
    int4* d = reinterpret_cast<int4*>(s_tri + slot * kStageWords);
    uint32_t blocks = bbox_block_mask(max(bb.x, tx0) - tx0, max(bb.y, ty0) - ty0, min(bb.z, tx1) - tx0, min(bb.w, ty1) - ty0);
    uint32_t id = t;
    bool exact = true;
    // (non-finite vertex depths also take the literal path: the exact sweep then needs no NaN test per fragment,
    // because barycentrics in [0, 1] times finite depths cannot produce one)
    const bool finite_z = fabsf(r1.z) <= 3.402823466e38f && fabsf(r1.w) <= 3.402823466e38f && fabsf(r2.x) <= 3.402823466e38f;
    if (ex <= kExactExtent && ey <= kExactExtent && finite_z) {
        int ux = x1 - x0, uy = x2 - x0, vx = y1 - y0, vy = y2 - y0;
        int cz = ux * vy - uy * vx;
        int s = cz < 0 ? -1 : 1;
        int rx = x0 - tx0, ry = y0 - ty0;
        // cx(lx,ly) = uy*(y0-py) - (x0-px)*vy ,  cy(lx,ly) = (x0-px)*vx - ux*(y0-py)   with px = tx0+lx, py = ty0+ly
        d[0] = make_int4(s * (uy * ry - rx * vy), s * vy, -s * uy, s * (rx * vx - ux * ry));
        d[1] = make_int4(-s * vx, s * ux, s * cz, (int)id);
        if (cz == 0) blocks = 0u;                       // |cz| < 1: degenerate, never produces a fragment (scene.rs:188)
    } else {
        d[0] = make_int4(x0, y0, x1, y1);
        d[1] = make_int4(x2, y2, 0, (int)(id | kInexactFlag));
        exact = false;                                  // visible to the literal-path sweep only
        atomicAdd(s_inexact, 1);
    }
    d[2] = make_int4(__float_as_int(r1.z), __float_as_int(r1.w), __float_as_int(r2.x), __float_as_int(1.0f / fabsf((float)d[1].z)));
    s_mask[slot] = exact ? blocks : 0u;
    s_maskx[slot] = exact ? 0u : blocks;
}

#ifdef TR_TIMELINE
// Diagnostic build only (make EXTRA=-DTR_TIMELINE, tools/lone_timeline.py): per CTA of the last k_raster launch, words
// [0] entry, [1] past griddepcontrol.wait, [2] first filter round done, [3] exit (globaltimer, ns), [4] SM id,
// [5] staged polygons of the first round.  Read back with tr_debug_timeline (tinyr.cu, same macro).
constexpr int kTimelineWords = 8, kTimelineCtas = 8192;
__device__ unsigned long long g_timeline[kTimelineCtas * kTimelineWords];
__device__ __forceinline__ unsigned long long timeline_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
__device__ __forceinline__ void timeline_mark(uint32_t pos, int word, unsigned long long v) {
    if (threadIdx.x == 0 && pos < (uint32_t)kTimelineCtas) g_timeline[pos * kTimelineWords + word] = v;
}
#define TR_TIMELINE_MARK(pos, word, v) timeline_mark(pos, word, v)
#else
#define TR_TIMELINE_MARK(pos, word, v) ((void)0)
#endif

// Winner state of a lane's pixels as left in memory by earlier work on the target (a render without clear, or an
// earlier round of a long polygon list).  Out of line on purpose: inlined, the compiler predicates these loads and
// conversions into every block's start-up, where a cleared frame — the common case — never needs them.
struct BlockState { float z[kPix]; uint32_t lo[kPix]; };
template <bool DEPTH_ONLY>
__device__ __noinline__ BlockState load_block_state(const unsigned long long* keys, const float* depth, int width, int height, int px, int py) {
    BlockState st;
#pragma unroll
    for (int k = 0; k < kPix; ++k) {
        st.z[k] = -3.402823466e38f; st.lo[k] = kNoTri;
        if (px < width && py + 4 * k < height) {
            const size_t at = (size_t)(py + 4 * k) * width + px;
            if (DEPTH_ONLY) st.z[k] = depth[at];
            else { unsigned long long key = keys[at]; st.z[k] = key_depth(key); st.lo[k] = key_lo(key); }
        }
    }
    return st;
}

// The tile kernel's body: one CTA, one tile.  launch_pos selects the tile through a.tile_order; the tensor maps must
// live in parameter (or constant / global) memory.
template <class PIPE, int THREADS>
__device__ __forceinline__ void raster_tile(const RasterArgs& a, const ShadeArgs& sa, const Uniforms& u, const CUtensorMap* rgb_map,
                                            const CUtensorMap* rgb_map_top, uint32_t launch_pos) {
    constexpr int MODE = PIPE::id;
    constexpr bool DEPTH_ONLY = MODE == kModeDepthOnly;
    constexpr bool SHADE = MODE >= 0;
    constexpr int kRgbRow = kTileW * 3;
    constexpr int kThreads = THREADS, kCap = raster_list_cap(THREADS);
    __shared__ __align__(16) float s_tri[kCap * kStageWords];
    __shared__ uint32_t s_mask[kCap], s_maskx[kCap];   // per staged polygon: blocks to visit (exact / literal path)
    // the tile's RGB in frame order (top-down: tile row r at s_rgb row rows-1-r, rows = the tile's on-screen rows), the
    // box of the TMA store
    __shared__ __align__(128) uint8_t s_rgb[SHADE ? kTileH * kRgbRow : 16];
    __shared__ int s_count;
    __shared__ int s_next;
    __shared__ int s_block;                                 // next block of the tile to hand to a warp
    __shared__ int s_inexact;                               // staged polygons that need the literal f32 path

    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // CTAs are handed out in launch order; for a lone frame tile_order lists the tiles from the screen centre outwards
    // so that the tiles that usually carry the geometry start first and the cheap border tiles fill the tail of the grid.
    const uint32_t tile_xy = __ldg(&a.tile_order[launch_pos]);
    const uint32_t tile = (tile_xy >> 16) * a.tiles_x + (tile_xy & 0xffffu);
    const int tx0 = (int)(tile_xy & 0xffffu) * kTileW, ty0 = (int)(tile_xy >> 16) * kTileH;
    const int tx1 = min(tx0 + kTileW, (int)a.width) - 1, ty1 = min(ty0 + kTileH, (int)a.height) - 1;
    const float kFloatMin = -3.402823466e38f;

    const uint32_t cell = (uint32_t)(ty0 / kCellH) * a.bins.cells_x + (uint32_t)(tx0 / kCellW);
    // Everything above only reads launch-time constants (the tile order table is written at scene creation); from here
    // on the kernel reads what the vertex set-up kernel in front of it wrote (tinyr.cu: launch_raster).
    asm volatile("griddepcontrol.wait;" ::: "memory");
    TR_TIMELINE_MARK(launch_pos, 1, timeline_now());
    TR_TIMELINE_MARK(launch_pos, 2, 0ull);
    TR_TIMELINE_MARK(launch_pos, 5, 0ull);
    const uint32_t n_bin = a.bins.count[a.bins.parity * a.bins.n_cells + cell];
    // The filter below is a chain of dependent loads (bin count -> list entry -> bounding box -> setup record) that
    // the whole CTA waits for.  The first list entry of every thread is fetched alongside the count (the slot exists
    // whatever the count says; it is only used when it turns out to be valid), and a polygon's bounding box and
    // record travel together, so the chain is two loads deep instead of four.
#ifndef TR_NO_FILTER_PREFETCH
    const uint32_t t_first = a.bins.list[(size_t)cell * kBinCap + tid];
#endif
    const bool cleared = a.fresh || (!DEPTH_ONLY && a.tile_state[tile] != 0u);
    // With a cleared frame the tile's RGB is rebuilt completely (zeros where nothing is drawn) and stored in bulk.
    const bool rgb_bulk = SHADE && sa.fresh;
    const int row_bytes = (tx1 - tx0 + 1) * 3;

    const uint32_t epoch_hi = a.epoch << 24;
    const uint32_t s_block_at = (uint32_t)__cvta_generic_to_shared(&s_block);
    const bool use_bin = n_bin <= (uint32_t)kBinCap;
    const uint32_t n_src = use_bin ? n_bin : a.n_triangles;
    bool from_memory = !cleared;                          // later rounds of a long list merge with the earlier ones
    bool touched = false;
    uint32_t scan_from = 0;
    bool more = n_src > 0;
    while (more) {
        // ---- (1) filter the source list (cell bin, or all polygons) against the tile, kCap polygons per round ----
        if (tid == 0) { s_count = 0; s_next = (int)n_src; s_block = 0; s_inexact = 0; }
        __syncthreads();
        // Warps walk the list independently (no barrier per step): a warp ballots its 32 entries, reserves slots for
        // the hits with one shared atomic and stages them.  The order of the staged list therefore varies from run to
        // run; the result does not (winners are ordered by depth and polygon id, never by list position).  When the
        // list overflows, s_next ends up at the smallest source index that was not staged and the next round restarts
        // there; entries beyond it that did get a slot are staged again, which changes nothing (an equal key never
        // beats itself).
        for (uint32_t base = scan_from + warp * 32; base < n_src; base += kThreads) {
            if (*(volatile int*)&s_count >= kCap) { if (lane == 0) atomicMin(&s_next, (int)base); break; }
            uint32_t i = base + lane, t = 0;
            bool hit = false;
            int4 bb = make_int4(1, 0, 0, 0);
            float4 r0 = make_float4(0.0f, 0.0f, 0.0f, 0.0f), r1 = r0, r2 = r0;
            if (i < n_src) {
#ifndef TR_NO_FILTER_PREFETCH
                t = !use_bin ? i : (i == (uint32_t)tid ? t_first : a.bins.list[(size_t)cell * kBinCap + i]);
#else
                t = use_bin ? a.bins.list[(size_t)cell * kBinCap + i] : i;
#endif
                bb = __ldg(&a.bbox[t]);
#ifndef TR_NO_FILTER_PREFETCH
                const float4* r = a.rec + (size_t)t * 4;
                r0 = __ldg(r); r1 = __ldg(r + 1); r2 = __ldg(r + 2);
#endif
                hit = bb.x <= bb.z && bb.x <= tx1 && bb.z >= tx0 && bb.y <= ty1 && bb.w >= ty0;
            }
#ifdef TR_NO_FILTER_PREFETCH
            if (hit) { const float4* r = a.rec + (size_t)t * 4; r0 = __ldg(r); r1 = __ldg(r + 1); r2 = __ldg(r + 2); }
#endif
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (m == 0u) continue;
            int first = 0;
            if (lane == 0) first = atomicAdd(&s_count, __popc(m));
            first = __shfl_sync(0xffffffffu, first, 0);
            const int slot = first + __popc(m & ((1u << lane) - 1u));
            if (hit && slot < kCap) stage_polygon(t, bb, r0, r1, r2, slot, tx0, ty0, tx1, ty1, s_tri, s_mask, s_maskx, &s_inexact);
            if (hit && slot >= kCap) atomicMin(&s_next, (int)i);       // a source entry that did not fit
        }
        __syncthreads();
        const int n_list = min(s_count, kCap);
        const int n_inexact = s_inexact;
#ifdef TR_TIMELINE
        if (!touched) { TR_TIMELINE_MARK(launch_pos, 2, timeline_now()); TR_TIMELINE_MARK(launch_pos, 5, (unsigned long long)n_list); }
#endif
        scan_from = (uint32_t)s_next;       // == n_src when everything was staged
        more = scan_from < n_src;
        // A round that staged nothing reached the end of the source list (more is false).  In the shading modes it still
        // has to run when earlier rounds left winners behind: the fragment stage only runs in the last round, and the
        // round before this one ended with `more` set (it filled the list exactly, or a warp saw the list full before it
        // had looked at its remaining entries), so nothing has been shaded yet.  The block loop below then runs with an
        // empty list: winners come from memory and their barycentrics are re-derived.
        if (n_list == 0 && !(SHADE && touched)) break;
        touched = true;
        // Narrow the bounding-box masks of the exact-path polygons to the blocks the triangle itself reaches (one lane
        // per staged polygon), and pad both mask arrays with zeros to a whole warp for the sweeps below.
        for (int i = tid; i < ((n_list + 31) & ~31); i += kThreads) {
            if (i >= n_list) { s_mask[i] = 0u; s_maskx[i] = 0u; continue; }
            const uint32_t m = s_mask[i];
            if (m) {
                const int4* d = reinterpret_cast<const int4*>(s_tri + i * kStageWords);
                s_mask[i] = m & edge_block_mask(d[0], d[1]);
            }
        }
        __syncthreads();

        // ---- (2) warps pull 8 x (4*kPix) blocks off the tile (a lane owns kPix pixels, 4 rows apart); ballot the staged
        //      bboxes against the block, evaluate per pixel ----
#ifdef TR_STATIC_BLOCKS
        for (int blk = warp; blk < kBlocks; blk += kThreads / 32) {
#else
        for (;;) {
            // (the assembler wraps an atomic add in a ~15-instruction warp-aggregation sequence even for one lane; the
            // wrapping increment stays a single ATOMS.INC)
            int blk = 0;
            if (lane == 0) asm volatile("atom.shared.inc.u32 %0, [%1], 0x7fffffff;" : "=r"(blk) : "r"(s_block_at) : "memory");
            blk = __shfl_sync(0xffffffffu, blk, 0);
            if (blk >= kBlocks) break;
#endif
            const int bx0 = (blk % kBlocksX) * 8, by0 = (blk / kBlocksX) * kBlockH;   // tile-relative
            const int lx = bx0 + (lane & 7), ly = by0 + (lane >> 3);                  // pixel k: (lx, ly + 4k)
            const int px = tx0 + lx, py = ty0 + ly;
            // running winners: depth, low key word (epoch | polygon; kNoTri = nothing drawn yet) and barycentrics
            float bz[kPix], w0[kPix], w1[kPix], w2[kPix];
            uint32_t blo[kPix];
            bool fresh_winner[kPix];                // the winner was found in this round (w0..w2 are its barycentrics)
#pragma unroll
            for (int k = 0; k < kPix; ++k) { bz[k] = kFloatMin; blo[k] = kNoTri; w0[k] = w1[k] = w2[k] = 0.0f; fresh_winner[k] = false; }
            if (from_memory) {
                const BlockState st = load_block_state<DEPTH_ONLY>(a.keys, a.depth, (int)a.width, (int)a.height, px, py);
#pragma unroll
                for (int k = 0; k < kPix; ++k) { bz[k] = st.z[k]; blo[k] = st.lo[k]; }
            }
            // One sweep over the staged list: ballot the block's bit of the polygons' block masks, then every lane
            // evaluates the candidates for its pixels.  Polygons too large for the integer path (kInexactFlag; rare) carry
            // their mask in s_maskx, so the exact sweep never sees them and a second, separate sweep runs the literal f32
            // path — two loop bodies, so the compiler cannot hoist the slow path's set-up into the hot one.
            auto sweep = [&](auto exact_c) {
                constexpr bool EXACT = decltype(exact_c)::value;
                const uint32_t* masks = EXACT ? s_mask : s_maskx;
                for (int base = 0; base < n_list; base += 32) {
                    unsigned m = __ballot_sync(0xffffffffu, (masks[base + lane] >> blk) & 1u);
                    while (m) {
                        int j = base + __ffs(m) - 1;
                        m &= m - 1;
                        const int4* d = reinterpret_cast<const int4*>(s_tri + j * kStageWords);
                        const int4 e = d[0], f = d[1];
                        int e1[kPix], e2[kPix];
                        bool inside[kPix], any = !EXACT;
                        if (EXACT) {
                            e1[0] = e.x + lx * e.y + ly * e.z;           // cx * sign(cz)
                            e2[0] = e.w + lx * f.x + ly * f.y;           // cy * sign(cz)
#pragma unroll
                            for (int k = 0; k < kPix; ++k) {
                                if (k) { e1[k] = e1[k - 1] + 4 * e.z; e2[k] = e2[k - 1] + 4 * f.y; }
                                inside[k] = ((f.z - e1[k] - e2[k]) | e1[k] | e2[k]) >= 0;
                                any = any || inside[k];
                            }
                        }
                        if (!any) continue;
                        const int4 g = d[2];
                        const uint32_t lo = epoch_hi | ((uint32_t)f.w & kTriMask);
#pragma unroll
                        for (int k = 0; k < kPix; ++k) {
                            float z, b0, b1, b2;
                            if (EXACT) {
                                if (!inside[k]) continue;
                                // (cx+cy)/cz, cx/cz, cy/cz of scene.rs:189-193: integer operands, 0 <= numerator <= |cz| <= 2^21
                                const float cx = (float)e1[k], cy = (float)e2[k], cz = (float)f.z, rcz = __int_as_float(g.w);
                                b0 = 1.0f - exact_quotient(cx + cy, cz, rcz); b1 = exact_quotient(cx, cz, rcz); b2 = exact_quotient(cy, cz, rcz);
                                z = (b0 * __int_as_float(g.x) + b1 * __int_as_float(g.y)) + b2 * __int_as_float(g.z);
                            } else {
                                if (!fragment_depth(e.x, e.y, e.z, e.w, f.x, f.y, __int_as_float(g.x), __int_as_float(g.y), __int_as_float(g.z),
                                                    px, py + 4 * k, z, b0, b1, b2)) continue;
                            }
                            if (!EXACT && z != z) { atomicAdd(&a.diag[D_NAN_DEPTH], 1ull); continue; }
                            if (DEPTH_ONLY) {
                                if (z >= bz[k]) bz[k] = z;                               // shader.rs:703
                            } else if (z > bz[k] || (z == bz[k] && lo < blo[k] && blo[k] != kNoTri)) {
                                // shader.rs:175 `z <= zbuf -> reject`, applied in polygon order == key_beats() on packed keys
                                bz[k] = z; blo[k] = lo;
                                if (SHADE) { w0[k] = b0; w1[k] = b1; w2[k] = b2; fresh_winner[k] = true; }
                            }
                        }
                    }
                }
            };
            sweep(std::true_type{});
            if (n_inexact) sweep(std::false_type{});
#pragma unroll
            for (int k = 0; k < kPix; ++k) {
                if (px < (int)a.width && py + 4 * k < (int)a.height) {
                    const size_t at = (size_t)(py + 4 * k) * a.width + px;
                    if (DEPTH_ONLY) a.depth[at] = bz[k];
                    else a.keys[at] = make_key(bz[k], blo[k]);
                } else {
                    blo[k] = kNoTri;                // off-screen lanes of a border tile: nothing to shade below
                }
            }
            if (SHADE && !more) {
                // ---- (3) fragment stage for the winners that passed the depth test in THIS render; one copy of the
                //      fragment code (the pixel's state is selected into scalars) keeps the kernel in the instruction cache
                uint8_t* const tile_px = s_rgb + (ty1 - ty0 - ly) * kRgbRow + lx * 3;
#pragma unroll 1
                for (int k = 0; k < kPix; ++k) {
                    const bool second = k != 0;
                    float z = bz[0], v0 = w0[0], v1 = w1[0], v2 = w2[0];
                    uint32_t lo = blo[0];
                    bool fresh = fresh_winner[0];
                    if (kPix > 1 && second) { z = bz[kPix - 1]; v0 = w0[kPix - 1]; v1 = w1[kPix - 1]; v2 = w2[kPix - 1]; lo = blo[kPix - 1]; fresh = fresh_winner[kPix - 1]; }
                    const int qy = py + 4 * k;
                    uint32_t out = 0u;
                    const bool shaded = lo != kNoTri && (lo >> 24) == a.epoch;
                    if (shaded) {
                        if (!fresh) {           // won in an earlier round of a long list: re-derive its barycentrics
                            const float4* r = a.rec + (size_t)(lo & kTriMask) * 4;
                            float4 r0 = __ldg(r), r1 = __ldg(r + 1), r2 = __ldg(r + 2);
                            float zz;
                            fragment_depth(__float_as_int(r0.x), __float_as_int(r0.y), __float_as_int(r0.z), __float_as_int(r0.w),
                                           __float_as_int(r1.x), __float_as_int(r1.y), r1.z, r1.w, r2.x, px, qy, zz, v0, v1, v2);
                        }
                        const uint32_t tri = lo & kTriMask;
                        const float4* rec = sa.rec + (size_t)tri * 4;
                        out = PIPE::fragment(sa, u, FragmentIn{tri, px, qy, z, v0, v1, v2, __ldg(rec + 2), __ldg(rec + 3), sa.var + (size_t)tri * 6});
                    }
                    if (rgb_bulk) {
                        if (qy <= ty1) {                // (rows of a partial top tile beyond the frame have no place in s_rgb)
                            uint8_t* p = tile_px - k * (4 * kRgbRow);
                            p[0] = (uint8_t)out; p[1] = (uint8_t)(out >> 8); p[2] = (uint8_t)(out >> 16);
                        }
                    } else if (shaded) {
                        uint8_t* p = sa.rgb + ((size_t)(a.height - 1 - qy) * a.width + px) * 3;
                        p[0] = (uint8_t)out; p[1] = (uint8_t)(out >> 8); p[2] = (uint8_t)(out >> 16);
                    }
                }
            }
        }
        from_memory = true;
        if (SHADE) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");     // s_rgb -> the TMA store below
        __syncthreads();
    }

    if (!touched) {
        // Nothing touches this tile: an uncleared target keeps its contents, a cleared one its cleared state.
        if (cleared) {
            if (DEPTH_ONLY) {
                if (a.depth_clear[tile] != 0u) return;      // still cleared from an earlier pass: nothing to write
                __syncthreads();                            // (every warp has read the flag before it changes)
                if (tid == 0) a.depth_clear[tile] = 1u;
                if ((a.width & 3u) == 0u) {                 // rows of the tile are 16-byte aligned and whole float4s
                    const int quads = (tx1 - tx0 + 1) >> 2;
                    const float4 v = make_float4(kFloatMin, kFloatMin, kFloatMin, kFloatMin);
                    const int c = tid % (kTileW / 4);
                    if (c < quads)
                        for (int r = tid / (kTileW / 4); r <= ty1 - ty0; r += kThreads / (kTileW / 4))
                            reinterpret_cast<float4*>(a.depth + (size_t)(ty0 + r) * a.width + tx0)[c] = v;
                } else {
                    for (int i = tid; i < kTileW * kTileH; i += kThreads) {
                        int px = tx0 + i % kTileW, py = ty0 + i / kTileW;
                        if (px < (int)a.width && py < (int)a.height) a.depth[(size_t)py * a.width + px] = kFloatMin;
                    }
                }
            } else if (tid == 0) {
                a.tile_state[tile] = 1u;
            }
        }
    } else if (tid == 0) {
        if (DEPTH_ONLY) a.depth_clear[tile] = 0u;
        else a.tile_state[tile] = 0u;
    }

    if (rgb_bulk) {
        // ---- (4) the tile's RGB: one TMA tensor store of the kTileW x kTileH box (clipped at the frame border by the
        //      hardware).  A tile nothing touches is all zeros; when it already was, nothing is written at all.
        if (!touched && sa.rgb_zero[tile] != 0u) return;
        if (!touched) {
            for (int i = tid; i < kTileH * kRgbRow / 16; i += kThreads) reinterpret_cast<uint4*>(s_rgb)[i] = make_uint4(0u, 0u, 0u, 0u);
            asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            __syncthreads();
        }
        const int rows = ty1 - ty0 + 1;
        if (sa.rgb_tma) {
            if (tid == 0) {
                // Box coordinates must not be negative (the store faults), so the partial tile row at the top of the
                // frame has its own tensor map whose box is as high as that row; clipping on the high side (the right
                // border) is done by the hardware.
                const uint32_t src = (uint32_t)__cvta_generic_to_shared(s_rgb);
                const int x = tx0 * 3, y = (int)a.height - ty0 - rows;
                const uint64_t map = reinterpret_cast<uint64_t>(rows == kTileH ? rgb_map : rgb_map_top);
                asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                             :: "l"(map), "r"(src), "r"(x), "r"(y) : "memory");
                asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
            }
        } else {
            for (int i = tid; i < rows * row_bytes; i += kThreads) {
                int r = i / row_bytes, c = i % row_bytes;
                sa.rgb[((size_t)(a.height - 1 - (ty0 + r)) * a.width + tx0) * 3 + c] = s_rgb[(rows - 1 - r) * kRgbRow + c];
            }
        }
        if (tid == 0) sa.rgb_zero[tile] = touched ? 0u : 1u;
    } else if (SHADE && touched && tid == 0) {
        sa.rgb_zero[tile] = 0u;
    }
}

// One launch = one frame, one CTA per tile.  The lone shape also records what each tile cost (tile_cost), from which the
// set-up kernel of the next lone frame sorts the launch order.  (A form that launched only as many CTAs as are resident
// and let each walk the order through a device counter was measured with it: 1 us slower per 4K frame than this one
// with the sorted order, no faster without — profiles/r2_probes.md — and removed.)
template <class PIPE, int THREADS>
__global__ void TR_RASTER_BOUNDS k_raster(const RasterArgs a, const ShadeArgs sa, const Uniforms u, const __grid_constant__ CUtensorMap rgb_map,
                                         const __grid_constant__ CUtensorMap rgb_map_top) {
    constexpr bool LONE = THREADS == kRasterThreads;
    // (shared, not registers: the tile code's register budget is exhausted.  The tile's index into tile_cost arrives by an
    // asynchronous copy, so that no thread waits for it: a CTA of an empty tile lives for half a microsecond, and
    // two dependent loads at its end would treble that for 2,900 of the 4,080 CTAs of a 4K frame)
    __shared__ uint32_t s_t0, s_rank;
    if (LONE && a.tile_cost != nullptr) {
        asm volatile("griddepcontrol.wait;" ::: "memory");     // the launch order is the set-up kernel's output
        if (threadIdx.x == 0) {
            s_t0 = (uint32_t)clock();
            if (a.tile_perm != nullptr)
                asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" ::"r"((uint32_t)__cvta_generic_to_shared(&s_rank)), "l"(a.tile_perm + blockIdx.x) : "memory");
            else
                s_rank = blockIdx.x;
        }
    }
    TR_TIMELINE_MARK(blockIdx.x, 0, timeline_now());
    raster_tile<PIPE, THREADS>(a, sa, u, &rgb_map, &rgb_map_top, blockIdx.x);
#ifdef TR_TIMELINE
    unsigned sm;
    asm volatile("mov.u32 %0, %%smid;" : "=r"(sm));
    TR_TIMELINE_MARK(blockIdx.x, 3, timeline_now());
    TR_TIMELINE_MARK(blockIdx.x, 4, (unsigned long long)sm);
#endif
    if (LONE && blockIdx.x == 0 && threadIdx.x == 0) {       // (behind raster_tile's wait for the kernels in front)
        for (int which = 0; which < 2; ++which) {
            const unsigned long long* from = which ? a.light_diag : a.setup_diag;
            if (from == nullptr) continue;
            for (int k = 0; k < D_COUNT; ++k) {
                const unsigned long long v = from[k];
                if (v) atomicAdd(&a.diag[k], v);
            }
        }
    }
    if (LONE && a.tile_cost != nullptr && threadIdx.x == 0) {
        asm volatile("cp.async.wait_all;" ::: "memory");
        a.tile_cost[s_rank] = (uint32_t)clock() - s_t0;
    }
}

// Several views in ONE launch (multi-view-per-launch batching at small resolutions): blockIdx.y
// selects the view, i.e. one frame slot's buffers and one camera's uniforms out of the parameter block.  At 800x800 — the
// reference's own resolution (main.rs:6-7) — a frame is 325 tiles for 148 SMs and two launches per frame leave the device
// launch- and latency-bound; a batch of V views becomes two launches of V x 325 CTAs.  Every access to the arguments is
// an indexed constant-bank load here instead of an immediate operand, which the 4K paths (k_raster above) do not pay.
constexpr int kMaxViewsPerLaunch = 8;
struct ViewArgs {
    RasterArgs a;
    ShadeArgs sa;
    Uniforms u;
    CUtensorMap rgb_map, rgb_map_top;
};
struct ViewBatch { ViewArgs v[kMaxViewsPerLaunch]; };
static_assert(sizeof(ViewBatch) <= 32000, "kernel parameters are limited to 32,764 bytes");
template <class PIPE, int THREADS>
__global__ void TR_RASTER_BOUNDS k_raster_views(const __grid_constant__ ViewBatch b) {
    const ViewArgs& v = b.v[blockIdx.y];
    raster_tile<PIPE, THREADS>(v.a, v.sa, v.u, &v.rgb_map, &v.rgb_map_top, blockIdx.x);
}

}  // namespace tr
