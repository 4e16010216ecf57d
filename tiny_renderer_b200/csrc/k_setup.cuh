// k_setup.cuh — K1 vertex_setup (+ K2 bin, fused): one thread per polygon, coalesced float4 loads of the corner-expanded
// static geometry, SoA setup records out.  Replaces the reference's vertex closures
// (/root/reference/src/scene/shader.rs:285-316, 349-384, 416-437, 475-496, 549-595, 671-692, 711-747,
// 809-830, 849-870), their helpers (shader.rs:116-165) and the bounding box of scene.rs:232-239.
#pragma once
#include "pipeline_table.cuh"

namespace tr {

// Static per-polygon geometry, written once at scene creation (24 floats = 6 x float4):
//   [0..8]  p0 p1 p2 (xyz)   [9..14] u0 u1 u2 v0' v1' v2' with v' = 1 - v   [15..23] n0 n1 n2 (xyz)
constexpr int kStaticWords = 24;

// Launch order of the NEXT tile kernel, longest tile first (lone frames only; null cost = off).  A lone 4K frame spends a
// third of its tile kernel in a tail of late-started heavy tiles (per-CTA timeline: tools/lone_timeline.py); launching
// the tiles in the order of what they cost in the previous lone frame — views change little from frame to frame — cuts
// a darboux frame's tile kernel from 57 to 46 us and a phong frame's from 36 to 27 (tools/lone_order_probe.py).
struct TileSort {
    const uint32_t* cost;     // per centre-out position: SM cycles of that tile in the previous lone launch of this pass (k_raster.cuh)
    const uint32_t* base;     // n_tiles launch positions -> y << 16 | x, centre-out (ties keep this order)
    uint32_t* order;          // out: n_tiles entries, a permutation of base
    uint32_t* perm;           // out: the centre-out position of every entry of order (where the tile kernel files its cost)
    uint32_t n_tiles;
};
constexpr int kSortBuckets = 16, kSortShift = 13;        // 8,192 cycles (4.2 us) per bucket: as good as an exact sort (probe)
constexpr int kSortMaxTiles = 16384;                     // one byte of shared memory per tile (8K: 16,200 tiles)
static_assert(kSortBuckets * 4 == 64, "the prefix below handles two entries per lane of one warp");

#ifdef TR_TIMELINE
// Diagnostic build only: per block of the last set-up launch, [0] entry and [1] exit (globaltimer, ns); block 255 = the sort
// block; g_timeline_phase: per block, thread 0's time after the geometry loads, the transforms, the binning, the record stores.
__device__ unsigned long long g_timeline_setup[256 * 2];
__device__ unsigned long long g_timeline_phase[256 * 4];
#define TR_SORT_MARK(k) do { if (threadIdx.x == 0) { unsigned long long t_; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_) :: "memory"); g_timeline_phase[255 * 4 + (k)] = t_; } } while (0)
#define TR_SETUP_MARK(k, dep) do { if (threadIdx.x == 0 && blockIdx.x < 255) { unsigned long long t_; \
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_) : "r"(__float_as_int(dep)) : "memory"); g_timeline_phase[blockIdx.x * 4 + (k)] = t_; } } while (0)
#else
#define TR_SETUP_MARK(k, dep) ((void)0)
#define TR_SORT_MARK(k) ((void)0)
#endif

// One CTA of the set-up kernel (its last block, 128 threads): counting sort of the centre-out positions by cost bucket,
// longest first.  It has to finish inside the ~6 us the other blocks take, so nothing in it is a chain of dependent
// memory operations: costs and table entries are fetched at once (32 + 32 coalesced loads in flight per thread), counted
// and placed with shared-memory atomics.  Warp w owns the w-th quarter of the centre-out order and writes its members of a bucket behind
// those of warp w - 1, so a bucket keeps the centre-out order up to the order inside one warp's quarter — which only
// matters to speed, never to the image: every position is written exactly once (the output is a permutation of base).
// Measured (per-block timeline, profiles/r2_probes.md): 8.3 us against 4.0 us for the slowest vertex block — 3.2 us counting
// and 3.3 us scattering, because three quarters of the tiles fall into the last bucket, i.e. onto four addresses.  A form
// without atomics (a private counter column per thread over contiguous positions, shuffle scans) was slower still, 13.7 us:
// its loads no longer coalesce; serving the lanes of the last bucket (or, with match.any, of every bucket) with one atomic
// per step was slower too, 10.6 and 10.9 us.  The host therefore sorts twice per view and re-uses the order while the
// view stands still.
__device__ __forceinline__ uint32_t sort_bucket(uint32_t cycles) {        // bucket 0 = longest
    return (uint32_t)(kSortBuckets - 1) - min((uint32_t)(kSortBuckets - 1), cycles >> kSortShift);
}
__device__ __forceinline__ void sort_tiles(const TileSort& q) {
    __shared__ uint8_t s_bucket[kSortMaxTiles];
    __shared__ uint32_t s_run[kSortBuckets * 4];         // entry b * 4 + w: warp w's count of bucket b, then its next output slot
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const uint32_t n = q.n_tiles;
    const uint32_t per = ((n + 3u) / 4u + 31u) & ~31u;   // positions per warp
    const uint32_t w0 = min(n, (uint32_t)warp * per), w1 = min(n, w0 + per);
    // the table entries this lane will place first: in flight while the costs are fetched and counted
    uint32_t xy[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) { const uint32_t p = w0 + 32u * j + lane; xy[j] = p < w1 ? __ldg(&q.base[p]) : 0u; }
    if (tid < kSortBuckets * 4) s_run[tid] = 0u;
    __syncthreads();
    TR_SORT_MARK(0);
    for (uint32_t r0 = 0; r0 < n; r0 += 4096u) {
        uint32_t c[32];
#pragma unroll
        for (int j = 0; j < 32; ++j) { const uint32_t p = r0 + 128u * j + tid; c[j] = p < n ? q.cost[p] : 0u; }
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const uint32_t p = r0 + 128u * j + tid;
            if (p < n) {
                const uint32_t b = sort_bucket(c[j]), owner = (p >= per ? 1u : 0u) + (p >= 2u * per ? 1u : 0u) + (p >= 3u * per ? 1u : 0u);
                s_bucket[p] = (uint8_t)b;
                atomicAdd(&s_run[b * 4 + owner], 1u);
            }
        }
    }
    __syncthreads();
    TR_SORT_MARK(1);
    if (warp == 0) {            // exclusive prefix over the 64 counts in (bucket, warp) order, two entries per lane
        const uint32_t a = s_run[2 * lane], b = s_run[2 * lane + 1];
        uint32_t incl = a + b;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) { const uint32_t v = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += v; }
        s_run[2 * lane] = incl - a - b;
        s_run[2 * lane + 1] = incl - b;
    }
    __syncthreads();
    TR_SORT_MARK(2);
    for (uint32_t i0 = w0; i0 < w1; i0 += 1024u) {
        if (i0 != w0) {
#pragma unroll
            for (int j = 0; j < 32; ++j) { const uint32_t p = i0 + 32u * j + lane; xy[j] = p < w1 ? __ldg(&q.base[p]) : 0u; }
        }
#pragma unroll
        for (int j = 0; j < 32; ++j) {
            const uint32_t p = i0 + 32u * j + lane;
            if (p < w1) {
                const uint32_t slot = atomicAdd(&s_run[s_bucket[p] * 4 + warp], 1u);
                q.order[slot] = xy[j];
                q.perm[slot] = p;
            }
        }
    }
    TR_SORT_MARK(3);
}

struct SetupOut {
    float4* rec;      // n_triangles x 4 float4 (kRecWords)
    float4* var;      // n_triangles x 6 float4 (kVarWords), may be null when the pipeline needs none
    int4* bbox;       // n_triangles: xmin ymin xmax ymax clamped to the screen; xmin > xmax when not drawn
    Bins bins;        // K2: per-cell polygon lists for the raster kernel that follows
    TileSort sort;    // cost != null: the grid has one extra block that writes the tile kernel's launch order
    // Lone camera passes whose output the host may re-use for the next frames of the same view (tinyr.cu: run_setup) count
    // their diagnostics into a block of their own, which the tile kernel of every frame that uses this output folds
    // into the scene's: `diag` then points at that block and diag_zero at the block of the NEXT run, cleared here.
    unsigned long long* diag_zero;
};

// LIGHT_PASS: shadow/occlusion pass 1 — no culling, transform by u.shadow, no varyings.  Otherwise `pipeline` selects the
// pipeline whose `vertex` role (pipeline.cuh) writes the varyings.  The camera pass is ONE kernel for all pipelines, with a
// switch over the vertex roles generated from the pipeline list, not one instantiation per pipeline: a kernel per pipeline
// was measured in r2 (smaller kernels, 792-1,552 instructions instead of 2,080) and cost the phong + darboux view batch
// 5 % (266 vs 279 Gpx/s; each pipeline alone unchanged) — two scenes with different pipelines then keep two set-up
// kernels resident beside their tile kernels, and the SMs' 32 KB L1.5 instruction cache is already oversubscribed.
template <bool LIGHT_PASS, bool SORT = false>
__device__ __forceinline__ void vertex_setup(const float4* __restrict__ geom, const SetupOut& out, const Uniforms& u, int pipeline,
                                             unsigned long long* diag) {
    // The tile kernel behind this one is launched with programmatic stream serialization (tinyr.cu: launch_raster): let
    // it start now; it waits for this grid's completion before it touches anything written here.
    asm volatile("griddepcontrol.launch_dependents;");
    if (out.diag_zero != nullptr && blockIdx.x == 0 && threadIdx.x < D_COUNT) out.diag_zero[threadIdx.x] = 0ull;
    if (SORT && out.sort.cost != nullptr && blockIdx.x == gridDim.x - 1) {
        // the extra block: its share of the counter zeroing below, then the launch order
        uint32_t* other = out.bins.count + (out.bins.parity ^ 1u) * out.bins.n_cells;
        for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < out.bins.n_cells; c += gridDim.x * blockDim.x) other[c] = 0u;
        sort_tiles(out.sort);
        return;
    }
    uint32_t t = blockIdx.x * blockDim.x + threadIdx.x;
    const bool valid = t < u.n_triangles;        // no early return: the whole warp takes part in the binning
    if (!valid) t = u.n_triangles - 1;
    float g[kStaticWords];
#pragma unroll
    for (int k = 0; k < 6; ++k) {
        float4 v = __ldg(&geom[(size_t)t * 6 + k]);
        g[4 * k] = v.x; g[4 * k + 1] = v.y; g[4 * k + 2] = v.z; g[4 * k + 3] = v.w;
    }
    f3 p[3] = {{g[0], g[1], g[2]}, {g[3], g[4], g[5]}, {g[6], g[7], g[8]}};
    TR_SETUP_MARK(0, g[0] + g[23]);
    bool drawn = true;
    f3 face_normal{0.0f, 0.0f, 0.0f};
    if (!LIGHT_PASS) {
        // should_cull_face, shader.rs:116-124 (object-space normal against camera_direction)
        face_normal = cross3(p[1] - p[0], p[2] - p[0]);
        drawn = !(dot3(u.camera_direction, face_normal) <= 0.0f);
    }
    // store_vertex_transformation_results, shader.rs:150-165
    const mat4& M = LIGHT_PASS ? u.shadow : u.vpmv;
    int32_t rx[3], ry[3];
    float rz[3];
#pragma unroll
    for (int i = 0; i < 3; ++i) {
        bool ok;
        f3 s = dehomog(xform(M, f4{p[i].x, p[i].y, p[i].z, 1.0f}), ok);
        if (!ok && drawn && valid) atomicAdd(&diag[D_W_ZERO], 1ull);
        rx[i] = cast_i32(s.x);
        ry[i] = cast_i32(s.y);
        rz[i] = s.z;
    }
    TR_SETUP_MARK(1, rz[0] + rz[1] + rz[2]);
    // scene.rs:232-239
    int32_t xmin = max(0, min(min(rx[0], rx[1]), rx[2])), xmax = min(max(max(rx[0], rx[1]), rx[2]), (int32_t)u.width - 1);
    int32_t ymin = max(0, min(min(ry[0], ry[1]), ry[2])), ymax = min(max(max(ry[0], ry[1]), ry[2]), (int32_t)u.height - 1);
    if (!drawn || ymin > ymax) { xmin = 1; xmax = 0; }
    if (valid) out.bbox[t] = make_int4(xmin, ymin, xmax, ymax);
    // K2 bin (fused): append the polygon to every binning cell its clamped bounding box touches.  The order inside
    // a bin is whatever the atomics give; visibility does not depend on it (the packed-key min is commutative).
    // A lone frame waits for this kernel's slowest block (per-block timeline, profiles/r2_probes.md: median block 3.0 us,
    // the block with the large polygons 6.7 us, all of the difference in here), so no thread waits for one counter's
    // round trip at a time: a polygon of at most kSmallCells cells issues all its appends back to back and writes its
    // list entries at the very END of the kernel, behind the record and varying stores; larger polygons are walked by the
    // whole warp, lanes spread over the polygon's cells, kLargeBatch polygons' appends in flight together.
    constexpr int kSmallCells = 6, kLargeBatch = 8;
    uint32_t tl[kSmallCells], sl[kSmallCells];
#pragma unroll
    for (int i = 0; i < kSmallCells; ++i) { tl[i] = 0u; sl[i] = kBinCap; }
    {
        // zero the counter set the NEXT pass will fill (its last reader finished before this kernel started)
        uint32_t* other = out.bins.count + (out.bins.parity ^ 1u) * out.bins.n_cells;
        for (uint32_t c = blockIdx.x * blockDim.x + threadIdx.x; c < out.bins.n_cells; c += gridDim.x * blockDim.x) other[c] = 0u;
        uint32_t* count = out.bins.count + out.bins.parity * out.bins.n_cells;
        const bool any = valid && xmin <= xmax;
        const int bx0 = any ? xmin / kCellW : 0, by0 = any ? ymin / kCellH : 0;
        const int nx = any ? xmax / kCellW - bx0 + 1 : 0, ny = any ? ymax / kCellH - by0 + 1 : 0;
        const int lane = threadIdx.x & 31;
        const unsigned peers = 0xffffffffu;
        const bool small = any && nx * ny <= kSmallCells;
        if (small) {
#pragma unroll
            for (int i = 0; i < kSmallCells; ++i) {
                tl[i] = (uint32_t)(by0 + i / nx) * out.bins.cells_x + (uint32_t)(bx0 + i % nx);
                if (i < nx * ny) sl[i] = atomicAdd(&count[tl[i]], 1u);
            }
        }
        unsigned todo = __ballot_sync(peers, any && !small);
        while (todo) {
            uint32_t lc[kLargeBatch], ls[kLargeBatch], lt[kLargeBatch];
#pragma unroll
            for (int k = 0; k < kLargeBatch; ++k) {
                lc[k] = 0u; ls[k] = kBinCap; lt[k] = 0u;
                if (todo == 0u) continue;           // (uniform: todo is the same in every lane)
                const int src = __ffs(todo) - 1;
                todo &= todo - 1;
                const int sbx0 = __shfl_sync(peers, bx0, src), sby0 = __shfl_sync(peers, by0, src);
                const int snx = __shfl_sync(peers, nx, src), sny = __shfl_sync(peers, ny, src);
                lt[k] = __shfl_sync(peers, t, src);
                if (lane < snx * sny) {             // the first 32 cells: one per lane, the append's slot is collected below
                    lc[k] = (uint32_t)(sby0 + lane / snx) * out.bins.cells_x + (uint32_t)(sbx0 + lane % snx);
                    ls[k] = atomicAdd(&count[lc[k]], 1u);
                }
                for (int i = lane + 32; i < snx * sny; i += 32) {       // (a polygon of more than 32 cells: rare)
                    const uint32_t cell = (uint32_t)(sby0 + i / snx) * out.bins.cells_x + (uint32_t)(sbx0 + i % snx);
                    const uint32_t slot = atomicAdd(&count[cell], 1u);
                    if (slot < (uint32_t)kBinCap) out.bins.list[(size_t)cell * kBinCap + slot] = lt[k];
                }
            }
#pragma unroll
            for (int k = 0; k < kLargeBatch; ++k)
                if (ls[k] < (uint32_t)kBinCap) out.bins.list[(size_t)lc[k] * kBinCap + ls[k]] = lt[k];
        }
    }

    TR_SETUP_MARK(2, 0.0f);
    if (!valid) return;                     // (no appends pending: an invalid lane binned nothing)
    float4* rec = out.rec + (size_t)t * 4;
    rec[0] = make_float4(__int_as_float(rx[0]), __int_as_float(ry[0]), __int_as_float(rx[1]), __int_as_float(ry[1]));
    rec[1] = make_float4(__int_as_float(rx[2]), __int_as_float(ry[2]), rz[0], rz[1]);
    rec[2] = make_float4(rz[2], __int_as_float(drawn ? kFlagDrawn : 0u), g[9], g[10]);
    rec[3] = make_float4(g[11], g[12], g[13], g[14]);
    TR_SETUP_MARK(3, 0.0f);
    if constexpr (!LIGHT_PASS) {
        int words = 0, role = pipeline;
        switch (pipeline) {                 // which vertex role this pipeline runs (pipelines may share one: shadow runs phong's)
#define TR_X(N) case kPipe##N: role = Pipeline##N::vertex_role; words = Pipeline##N::varying_words; break;
            TR_PIPELINES(TR_X)
#undef TR_X
        }
        if (drawn && out.var != nullptr && words != 0) {      // (words == 0: nothing beyond the shared transform / uv store)
            float v[kVarWords];
#pragma unroll
            for (int k = 0; k < kVarWords; ++k) v[k] = 0.0f;
            VertexIn in;
#pragma unroll
            for (int i = 0; i < 3; ++i) { in.p[i] = p[i]; in.n[i] = f3{g[15 + 3 * i], g[16 + 3 * i], g[17 + 3 * i]}; }
            in.face_normal = face_normal;
            switch (role) {                     // the vertex role of every pipeline in the table that has one of its own
#define TR_X(N) case kPipe##N: if constexpr (Pipeline##N::vertex_role == kPipe##N && Pipeline##N::varying_words > 0) Pipeline##N::vertex(u, in, v, diag); break;
                TR_PIPELINES(TR_X)
#undef TR_X
            }
            float4* var = out.var + (size_t)t * 6;
#pragma unroll
            for (int k = 0; k < kVarWords / 4; ++k)
                if (4 * k < words) var[k] = make_float4(v[4 * k], v[4 * k + 1], v[4 * k + 2], v[4 * k + 3]);
        }
    }
    // the list entries of a small polygon, whose slots have long arrived
#pragma unroll
    for (int i = 0; i < kSmallCells; ++i)
        if (sl[i] < (uint32_t)kBinCap) out.bins.list[(size_t)tl[i] * kBinCap + sl[i]] = t;
}

// SORT: the lone-frame form, whose last block writes the tile kernel's launch order (sort_tiles; 25 KB of shared memory
// that the batch form, resident beside other frames' tile kernels, does not carry).
template <bool LIGHT_PASS, bool SORT = false>
__global__ void __launch_bounds__(128) k_vertex_setup(const float4* __restrict__ geom, const SetupOut out, const Uniforms u, int pipeline,
                                                     unsigned long long* diag) {
#ifdef TR_TIMELINE
    unsigned long long t_in;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_in));
#endif
    vertex_setup<LIGHT_PASS, SORT>(geom, out, u, pipeline, diag);
#ifdef TR_TIMELINE
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t_out;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t_out));
        const int slot = (SORT && out.sort.cost != nullptr && blockIdx.x == gridDim.x - 1) ? 255 : min((int)blockIdx.x, 254);
        g_timeline_setup[2 * slot] = t_in; g_timeline_setup[2 * slot + 1] = t_out;
    }
#endif
}

// Several views in one launch (see k_raster_views): blockIdx.y selects the view's uniforms and frame slot.
constexpr int kMaxSetupViews = 8;
struct SetupViewArgs { SetupOut out; Uniforms u; };
struct SetupViewBatch { SetupViewArgs v[kMaxSetupViews]; };
template <bool LIGHT_PASS>
__global__ void __launch_bounds__(128) k_vertex_setup_views(const float4* __restrict__ geom, const __grid_constant__ SetupViewBatch b, int pipeline,
                                                           unsigned long long* diag) {
    const SetupViewArgs& v = b.v[blockIdx.y];
    vertex_setup<LIGHT_PASS>(geom, v.out, v.u, pipeline, diag);
}

}  // namespace tr
