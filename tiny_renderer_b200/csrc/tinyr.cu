// tinyr.cu — host driver behind the C ABI of include/tinyr.h: owns the device state of a scene and issues
// K0 (host prepare) -> K1 vertex_setup -> K3 raster_visibility [K4 shadow_depth first for two-pass pipelines]
// -> K5 shade on one CUDA stream.  Mirrors the reference's Scene (/root/reference/src/scene.rs:25-151).
// There is no CPU fallback: every entry point fails with TR_ERR_NO_DEVICE / TR_ERR_CUDA without a GPU.
#include "../../include/tinyr.h"
#include "tr_error.h"
#include "pipeline_table.cuh"
#include "k_setup.cuh"
#include "k_raster.cuh"
#include "k_misc.cuh"
#include "k_peer.cuh"
#include "k_host.cuh"

#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstring>
#include <string>
#include <algorithm>
#include <map>
#include <mutex>
#include <vector>

using namespace tr;

#define CU(call)                                                                                         \
    do {                                                                                                 \
        cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) return fail(TR_ERR_CUDA, "%s -> %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
    } while (0)

// u32 words of the AO allocation: the sums, zero padding to whole uint4, the barrier flags of K9 (k_peer.cuh)
static inline size_t ao_vec_count(size_t n_px) { return (n_px + 3) / 4; }
static inline size_t ao_alloc_words(size_t n_px) { return ao_vec_count(n_px) * 4 + kPeerFlagWords; }

enum KernelSlot { KS_SETUP = 0, KS_BIN, KS_RASTER, KS_SHADOW, KS_SHADE, KS_AO, KS_CLEAR, KS_CHECKSUM, KS_COUNT };

struct tr_scene {
    int device = 0;
    uint32_t width = 0, height = 0, n_tri = 0, ring = 1, cur = 0;
    size_t n_px = 0;
    int pipeline = kPipeDefault;
    // reference Scene state (scene.rs:32-35, defaults :66-69)
    f3 light{0.0f, 0.0f, -1.0f}, from{0.0f, 0.0f, 1.0f}, at{0.0f, 0.0f, 0.0f}, up{0.0f, 1.0f, 0.0f};
    Uniforms uni{};            // last camera-pass uniforms (for tr_scene_get_uniforms)
    uint32_t epoch = 0;        // renders since the last clear
    bool fresh_keys = false, fresh_rgb = false, fresh_shadow = false;   // a clear is pending (folded into the next render)
    // device
    cudaStream_t own_stream = nullptr, stream = nullptr;
    float4* geom = nullptr;
    // One slot = everything a frame in flight owns, so that consecutive frames of a view batch can overlap on the
    // GPU (tr_scene_render_views rotates through the slots, each on its own stream).  The single-frame API works on
    // slot `cur` and the scene stream.
    struct Slot {
        unsigned long long* keys = nullptr;
        uint32_t* tile_state = nullptr;    // per raster tile: 1 = cleared, key memory stale
        uint8_t* rgb = nullptr;
        float* shadow = nullptr;
        // per raster tile, nonzero = the tile of rgb / shadow holds its cleared value in memory (k_raster skips the store)
        uint32_t* rgb_zero = nullptr;
        uint32_t* shadow_clear = nullptr;
        CUtensorMap rgb_map{};             // rgb as a (width*3 bytes) x height u8 tensor, box = one raster tile
        CUtensorMap rgb_map_top{};         // same tensor, box as high as the partial tile row at the top of the frame
        bool rgb_tma = false;
        SetupOut cam{}, lit{};
        // Re-use of a lone camera pass's set-up output (run_setup): what it was computed for, the diagnostics blocks of
        // its runs (k_setup.cuh SetupOut::diag_zero) and how many runs there were
        struct SetupKey { f3 from{0, 0, 0}, at{0, 0, 0}, up{0, 0, 0}, light{0, 0, 0}; int pipeline = -1; bool valid = false; } setup_key;
        unsigned long long* setup_diag = nullptr;      // 2 x D_COUNT
        uint32_t setup_runs = 0;
        bool setup_reused = false;                     // the frame being queued re-uses the output (no set-up launch)
        // Re-use of a lone frame's light pass (render_impl): the shadow buffer holds "cleared + the light pass of
        // shadow_key" (light, target, up) and nothing has written to it since; its diagnostics blocks and run count
        SetupKey shadow_key;
        unsigned long long* light_diag = nullptr;      // 2 x D_COUNT
        uint32_t light_runs = 0;
        unsigned long long* light_diag_cur = nullptr;  // non-null while render_impl queues a light pass that counts into its own block
        Bins bins{};
        cudaStream_t stream = nullptr;
        cudaEvent_t done = nullptr;
    };
    std::vector<Slot> slots;
    uint32_t tiles_x = 0, tiles_y = 0;
    // launch position -> tile coordinates (y << 16 | x), two orders of n_tiles entries each: [0] tiles sorted by
    // distance from the screen centre, [1] plain row order
    uint32_t* tile_order = nullptr;
    // lone frames: per pass (0 light depth, 1 camera) the tiles' cost in the last lone launch and the launch order the
    // set-up kernel sorts from it (k_setup.cuh: sort_tiles); order_sorted[p]: the set-up kernel in front wrote order p
    uint32_t* tile_cost = nullptr;       // [pass][centre-out position]
    uint32_t* tile_perm = nullptr;       // [pass][launch position of the sorted order] -> centre-out position
    uint32_t* tile_order_sorted = nullptr;
    bool order_sorted[2] = {false, false};
    // the view (eye or light, target, up) whose lone frame left tile_cost[p] behind; the sorted order is only used for a
    // view close to it (view_close)
    struct CostView { f3 eye{0, 0, 0}, at{0, 0, 0}, up{0, 0, 0}; int pipeline = -1; bool valid = false; } cost_view[2];
    CostView order_view[2];            // the view the sorted order of pass p was last built for, and how often in a row:
    uint32_t order_builds[2] = {0, 0}; // a view that stands still is sorted twice (the second time from a sorted frame's costs), then re-uses the order
    uint32_t order_uses[2] = {0, 0};   // frames since the order was last built (it is rebuilt every 256 frames: costs recorded under a disturbance do not stick)
    CostView last_view[2];             // the view of the previous lone frame: costs are only recorded once the view has settled
    bool overlapping = false;            // a batch is running frames / light passes concurrently on the slot streams
    uint32_t debug_flags = 0;            // tr_scene_set_debug_flags -> Uniforms::debug_flags
    uint32_t* ao = nullptr;            // n_px sums, zero padding to a multiple of 4, then the kPeerFlagWords barrier flags of K9
    // K9 (k_peer.cuh): the accumulators of all ranks, mapped through CUDA IPC by tr_scene_ao_open_peers
    uint32_t* peer_buf[kMaxPeers] = {};
    uint32_t peer_world = 0, peer_rank = 0, peer_value = 1;
    unsigned long long* nvls_mc = nullptr;   // multicast mapping of all ranks' accumulators (tr_multi_*, NVLS): K9n instead of K9
    bool ao_vmm = false;               // `ao` is a VMM mapping owned by a tr_multi (not cudaFree'd here)
    bool peer_ipc = true;              // peer_buf entries were opened from CUDA IPC handles (one process per GPU); false: plain peer pointers (tr_multi_*)
    int* peer_status = nullptr;        // mapped host word the kernel sets when a barrier times out
    AoList ao_list{};                  // covered pixels of the camera view (many-light AO batch), allocated on first use
    uint32_t* ao_count_host = nullptr; // pinned: length of ao_list, copied behind k_ao_compact
    cudaEvent_t ev_ao_count = nullptr;
    unsigned long long* diag = nullptr;
    // K10 (k_host.cuh): [0] bytes / [1] tile rows stored into host frames by the sparse delivery; host-side counters beside them
    unsigned long long* host_stats = nullptr;
    uint64_t frames_sparse = 0, frames_copied = 0, bytes_copied = 0;
    unsigned long long* d_checksums = nullptr; size_t checksum_cap = 0;
    uint8_t* scratch = nullptr; size_t scratch_bytes = 0;
    cudaArray_t arrays[4] = {nullptr, nullptr, nullptr, nullptr};
    Maps maps{};
    // timing
    cudaEvent_t ev_begin = nullptr, ev_end = nullptr, ev_fork = nullptr;
    bool profiling = false;
    struct Span { int slot; cudaEvent_t a, b; };
    std::vector<Span> spans;
    std::vector<cudaEvent_t> event_pool;
    float kernel_ms[KS_COUNT] = {0};
    uint64_t kernel_launches[KS_COUNT] = {0};
};

namespace {

inline tr_scene::Slot& S(tr_scene* s) { return s->slots[s->cur]; }

// ---- the pipeline registry (ShaderPipeline::new, shader.rs:97-112), generated from TR_PIPELINES ------------------
// One entry per pipeline struct: its names, whether a light-depth pass runs first, its host `prepare` role and the
// launchers of the kernels instantiated with its `vertex` / `fragment` roles.
struct PipelineEntry {
    const char* name;
    const char* alias;
    bool light_pass;
    int (*prepare)(Uniforms&, uint32_t, uint32_t, f3, f3, f3, f3);
    void (*raster[2])(tr_scene*, dim3, const RasterArgs&, const ShadeArgs&, const Uniforms&);      // [0] lone-frame CTA shape, [1] batch shape
    void (*raster_views)(tr_scene*, dim3, const ViewBatch&);                                       // several views per launch (small frames)
};
template <class PIPE, int T> void launch_raster(tr_scene* s, dim3 grid, const RasterArgs& a, const ShadeArgs& sa, const Uniforms& u);
template <class PIPE> void launch_raster_views(tr_scene* s, dim3 grid, const ViewBatch& b);
template <class PIPE> constexpr PipelineEntry pipeline_entry() {
    return {PIPE::name, PIPE::alias, PIPE::light_pass, &PIPE::prepare,
            {&launch_raster<PIPE, kRasterThreads>, &launch_raster<PIPE, kRasterThreadsBatch>}, &launch_raster_views<PIPE>};
}
const PipelineEntry kPipelines[kPipeCount] = {
#define TR_X(N) pipeline_entry<Pipeline##N>(),
    TR_PIPELINES(TR_X)
#undef TR_X
};
// the two kernel modes that are not pipelines (pipeline.cuh)
void (*const kRasterLightDepth[2])(tr_scene*, dim3, const RasterArgs&, const ShadeArgs&, const Uniforms&) =
    {&launch_raster<PassLightDepth, kRasterThreads>, &launch_raster<PassLightDepth, kRasterThreadsBatch>};
void (*const kRasterVisibility[2])(tr_scene*, dim3, const RasterArgs&, const ShadeArgs&, const Uniforms&) =
    {&launch_raster<PassVisibility, kRasterThreads>, &launch_raster<PassVisibility, kRasterThreadsBatch>};

int pipeline_from_name(const char* name) {
    if (!name) return -1;
    for (int p = 0; p < kPipeCount; ++p)
        if (strcmp(kPipelines[p].name, name) == 0 || (kPipelines[p].alias && strcmp(kPipelines[p].alias, name) == 0)) return p;
    return -1;
}

int ensure_scratch(tr_scene* s, size_t bytes) {
    if (s->scratch_bytes >= bytes) return TR_OK;
    if (s->scratch) cudaFree(s->scratch);
    s->scratch = nullptr; s->scratch_bytes = 0;
    CU(cudaMalloc(&s->scratch, bytes));
    s->scratch_bytes = bytes;
    return TR_OK;
}

// Profiling spans: CUDA events around a launch while tr_scene_set_profiling is on.
struct Scoped {
    tr_scene* s; int slot; cudaEvent_t a = nullptr, b = nullptr;
    cudaEvent_t grab() {
        cudaEvent_t e;
        if (!s->event_pool.empty()) { e = s->event_pool.back(); s->event_pool.pop_back(); }
        else cudaEventCreate(&e);
        return e;
    }
    Scoped(tr_scene* s_, int slot_) : s(s_), slot(slot_) {
        s->kernel_launches[slot]++;
        if (s->profiling) { a = grab(); b = grab(); cudaEventRecord(a, s->stream); }
    }
    ~Scoped() {
        if (s->profiling) { cudaEventRecord(b, s->stream); s->spans.push_back({slot, a, b}); }
    }
};

int launch_ok(const char* what) {
    cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) return fail(TR_ERR_CUDA, "launch of %s failed: %s", what, cudaGetErrorString(e));
    return TR_OK;
}

// The slot's RGB frame as a 2-D u8 tensor (row = width*3 bytes) whose box is one raster tile: the tile kernel stores
// its shared-memory tile with one TMA instruction and the hardware clips it at the frame border.  Needs a row pitch
// that is a multiple of 16 bytes; other widths keep the byte-wise store.  cuTensorMapEncodeTiled comes from the
// driver through the runtime's entry-point query, so libcuda is not linked.
int make_rgb_map(tr_scene* s, tr_scene::Slot* sl) {
    sl->rgb_tma = false;
    if ((s->width * 3) % 16 != 0) return TR_OK;
    using EncodeTiled = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                     const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                     CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult found;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &found) != cudaSuccess || found != cudaDriverEntryPointSuccess || !fn)
        return fail(TR_ERR_CUDA, "cuTensorMapEncodeTiled is not available from this driver");
    const cuuint64_t dims[2] = {(cuuint64_t)s->width * 3, (cuuint64_t)s->height};
    const cuuint64_t pitch[1] = {(cuuint64_t)s->width * 3};
    const cuuint32_t step[2] = {1, 1};
    const uint32_t top_rows = s->height % kTileH ? s->height % kTileH : kTileH;
    CUtensorMap* maps[2] = {&sl->rgb_map, &sl->rgb_map_top};
    const uint32_t box_rows[2] = {std::min<uint32_t>(kTileH, s->height), top_rows};
    for (int k = 0; k < 2; ++k) {
        const cuuint32_t box[2] = {(cuuint32_t)kTileW * 3, (cuuint32_t)box_rows[k]};
        CUresult r = reinterpret_cast<EncodeTiled>(fn)(maps[k], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, sl->rgb, dims, pitch, box, step,
                                                       CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE,
                                                       CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) return fail(TR_ERR_CUDA, "cuTensorMapEncodeTiled failed (%d) for a %ux%u frame", (int)r, s->width, s->height);
    }
    sl->rgb_tma = true;
    return TR_OK;
}

int fill_state(tr_scene* s, float z_init, bool do_keys, bool do_shadow, bool do_rgb) {
    if (!do_keys && !do_shadow && !do_rgb) return TR_OK;
    Scoped sp(s, KS_CLEAR);
    k_fill_state<<<148 * 8, 256, 0, s->stream>>>(S(s).keys, S(s).shadow, reinterpret_cast<uint32_t*>(S(s).rgb), s->n_px,
                                                 (s->n_px * 3 + 3) / 4, z_init, do_keys, do_shadow, do_rgb);
    const size_t tile_bytes = (size_t)s->tiles_x * s->tiles_y * 4;
    if (do_keys) CU(cudaMemsetAsync(S(s).tile_state, 0, tile_bytes, s->stream));
    if (do_rgb) CU(cudaMemsetAsync(S(s).rgb_zero, 1, tile_bytes, s->stream));
    if (do_shadow) S(s).shadow_key.valid = false;       // (the shadow buffer no longer holds a light pass)
    if (do_shadow) CU(cudaMemsetAsync(S(s).shadow_clear, z_init == -3.402823466e38f ? 1 : 0, tile_bytes, s->stream));
    return launch_ok("k_fill_state");
}

// Turn pending (lazy) clears into memory contents, for getters that read state a render has not rewritten.
int materialise(tr_scene* s) {
    int rc = fill_state(s, -3.402823466e38f, s->fresh_keys, s->fresh_shadow, s->fresh_rgb);
    if (rc == TR_OK && !s->fresh_keys) {
        // tiles still in the tile-level cleared state: write their keys out
        Scoped sp(s, KS_CLEAR);
        k_materialise_tiles<<<dim3(s->tiles_x, s->tiles_y), 256, 0, s->stream>>>(S(s).keys, S(s).tile_state, s->width, s->height);
        rc = launch_ok("k_materialise_tiles");
    }
    s->fresh_keys = s->fresh_shadow = s->fresh_rgb = false;
    return rc;
}

int prep_error(int rc) {
    switch (rc) {
        case PREP_SINGULAR: return fail(TR_ERR_REFERENCE_UNDEFINED, "singular transform: the reference panics on try_inverse().unwrap() (shader.rs:224,277,278)");
        case PREP_LIGHT_W: return fail(TR_ERR_REFERENCE_UNDEFINED, "Vector3::from_homogeneous on w != 0: the reference panics (shader.rs:227,885)");
        case PREP_ROTATION: return fail(TR_ERR_REFERENCE_UNDEFINED, "light direction antiparallel to +z: Rotation3::rotation_between is None and the reference panics (shader.rs:921)");
        default: return TR_OK;
    }
}

int upload_map(tr_scene* s, int k, const tr_image& img) {
    if (!img.rgb8_topdown || img.width == 0 || img.height == 0) return fail(TR_ERR_INVALID_ARGUMENT, "map %d is empty", k);
    std::vector<uchar4> rgba((size_t)img.width * img.height);
    for (size_t i = 0; i < rgba.size(); ++i)
        rgba[i] = make_uchar4(img.rgb8_topdown[3 * i], img.rgb8_topdown[3 * i + 1], img.rgb8_topdown[3 * i + 2], 255);
    cudaChannelFormatDesc fmt = cudaCreateChannelDesc<uchar4>();
    CU(cudaMallocArray(&s->arrays[k], &fmt, img.width, img.height));
    CU(cudaMemcpy2DToArray(s->arrays[k], 0, 0, rgba.data(), (size_t)img.width * 4, (size_t)img.width * 4, img.height, cudaMemcpyHostToDevice));
    cudaResourceDesc res{};
    res.resType = cudaResourceTypeArray;
    res.res.array.array = s->arrays[k];
    cudaTextureDesc td{};
    td.addressMode[0] = td.addressMode[1] = cudaAddressModeClamp;
    td.filterMode = cudaFilterModePoint;       // nearest texel, like util.rs:34-41
    td.readMode = cudaReadModeElementType;
    td.normalizedCoords = 0;
    CU(cudaCreateTextureObject(&s->maps.tex[k], &res, &td, nullptr));
    s->maps.w[k] = img.width; s->maps.h[k] = img.height;
    return TR_OK;
}

int alloc_setup(SetupOut* o, uint32_t n_tri, bool with_var) {
    size_t n = n_tri ? n_tri : 1;
    CU(cudaMalloc(&o->rec, n * kRecWords * 4));
    CU(cudaMalloc(&o->bbox, n * sizeof(int4)));
    o->var = nullptr;
    if (with_var) CU(cudaMalloc(&o->var, n * kVarWords * 4));
    return TR_OK;
}

// Buffer::new (shader.rs:41-50): z and shadow buffers start at 0.0 (NOT f32::MIN), frame at 0.
int reset_to_new(tr_scene* s) {
    for (uint32_t r = 0; r < s->ring; ++r) {
        s->cur = r;
        int rc = fill_state(s, 0.0f, true, true, true);
        if (rc) return rc;
    }
    s->cur = 0;
    s->epoch = 0;
    s->fresh_keys = s->fresh_rgb = s->fresh_shadow = false;
    return TR_OK;
}

int create_impl(tr_scene* s, const tr_scene_desc* d) {
    int n_dev = 0;
    if (cudaGetDeviceCount(&n_dev) != cudaSuccess || n_dev == 0) { cudaGetLastError(); return fail(TR_ERR_NO_DEVICE, "no CUDA device (this backend has no CPU fallback)"); }
    if (d->device < 0 || d->device >= n_dev) return fail(TR_ERR_INVALID_ARGUMENT, "device %d out of range (have %d)", d->device, n_dev);
    s->device = d->device;
    CU(cudaSetDevice(s->device));
    s->pipeline = pipeline_from_name(d->pipeline_name);
    if (s->pipeline < 0) return fail(TR_ERR_UNKNOWN_PIPELINE, "Provided pipeline name is not supported! (%s)", d->pipeline_name ? d->pipeline_name : "null");
    if (d->width == 0 || d->height == 0 || d->width > 32768 || d->height > 32768) return fail(TR_ERR_INVALID_ARGUMENT, "bad frame size %ux%u", d->width, d->height);
    if (d->n_triangles > kMaxTriangles) return fail(TR_ERR_UNSUPPORTED, "more than %u polygons", kMaxTriangles);
    if ((d->n_triangles && (!d->tri_idx || !d->positions || !d->uvs || !d->normals))) return fail(TR_ERR_INVALID_ARGUMENT, "null geometry array");
    s->width = d->width; s->height = d->height; s->n_tri = d->n_triangles;
    s->n_px = (size_t)d->width * d->height;
    s->ring = d->ring ? d->ring : 1;

    // Corner-expanded static geometry: 24 floats per polygon (k_setup.cuh), v flipped here (shader.rs:144).
    std::vector<float> g((size_t)(s->n_tri ? s->n_tri : 1) * kStaticWords, 0.0f);
    for (uint32_t t = 0; t < s->n_tri; ++t) {
        const uint32_t* ix = d->tri_idx + (size_t)9 * t;
        float* o = &g[(size_t)t * kStaticWords];
        for (int c = 0; c < 3; ++c) {
            uint32_t p = ix[3 * c], tx = ix[3 * c + 1], n = ix[3 * c + 2];
            if (p >= d->n_positions || tx >= d->n_uvs || n >= d->n_normals) return fail(TR_ERR_BAD_POLYGON, "polygon %u: index out of range", t);
            o[3 * c] = d->positions[3 * p]; o[3 * c + 1] = d->positions[3 * p + 1]; o[3 * c + 2] = d->positions[3 * p + 2];
            o[9 + c] = d->uvs[2 * tx];
            o[12 + c] = 1.0f - d->uvs[2 * tx + 1];
            o[15 + 3 * c] = d->normals[3 * n]; o[16 + 3 * c] = d->normals[3 * n + 1]; o[17 + 3 * c] = d->normals[3 * n + 2];
        }
    }
    CU(cudaStreamCreateWithFlags(&s->own_stream, cudaStreamNonBlocking));
    s->stream = s->own_stream;
    CU(cudaMalloc(&s->geom, g.size() * 4));
    CU(cudaMemcpy(s->geom, g.data(), g.size() * 4, cudaMemcpyHostToDevice));
    int rc;
    s->tiles_x = (s->width + kTileW - 1) / kTileW;
    s->tiles_y = (s->height + kTileH - 1) / kTileH;
    const size_t n_tiles = (size_t)s->tiles_x * s->tiles_y;
    {
        std::vector<uint32_t> order(n_tiles);
        for (size_t i = 0; i < n_tiles; ++i) order[i] = (uint32_t)i;
        const double cx = 0.5 * (s->tiles_x - 1), cy = 0.5 * (s->tiles_y - 1);
        // distance in pixels (tiles are kTileW x kTileH), ties by index: a deterministic order
        auto dist = [&](uint32_t t) {
            double dx = ((double)(t % s->tiles_x) - cx) * kTileW, dy = ((double)(t / s->tiles_x) - cy) * kTileH;
            return dx * dx + dy * dy;
        };
        std::stable_sort(order.begin(), order.end(), [&](uint32_t p, uint32_t q) { return dist(p) < dist(q); });
        std::vector<uint32_t> packed(2 * n_tiles);
        for (size_t i = 0; i < n_tiles; ++i) {
            packed[i] = ((order[i] / s->tiles_x) << 16) | (order[i] % s->tiles_x);
            packed[n_tiles + i] = ((uint32_t)(i / s->tiles_x) << 16) | (uint32_t)(i % s->tiles_x);
        }
#ifdef TR_TIMELINE
        // Diagnostic build only: TINYR_ORDER_PERM = file of n_tiles u32, entry i = the centre-out position that launches i-th
        // (tools/lone_timeline.py: what a cost-sorted launch order would buy).
        if (const char* f = getenv("TINYR_ORDER_PERM")) {
            std::vector<uint32_t> perm(n_tiles), was(packed.begin(), packed.begin() + n_tiles);
            FILE* fp = fopen(f, "rb");
            if (fp && fread(perm.data(), 4, n_tiles, fp) == n_tiles)
                for (size_t i = 0; i < n_tiles; ++i) packed[i] = was[perm[i] % n_tiles];
            if (fp) fclose(fp);
        }
#endif
        CU(cudaMalloc(&s->tile_order, packed.size() * 4));
        CU(cudaMemcpy(s->tile_order, packed.data(), packed.size() * 4, cudaMemcpyHostToDevice));
        CU(cudaMalloc(&s->tile_perm, 2 * n_tiles * 4));
        CU(cudaMemset(s->tile_perm, 0, 2 * n_tiles * 4));
        CU(cudaMalloc(&s->tile_cost, 2 * n_tiles * 4));
        CU(cudaMemset(s->tile_cost, 0, 2 * n_tiles * 4));
        CU(cudaMalloc(&s->tile_order_sorted, 2 * n_tiles * 4));
        CU(cudaMemcpy(s->tile_order_sorted, packed.data(), n_tiles * 4, cudaMemcpyHostToDevice));
        CU(cudaMemcpy(s->tile_order_sorted + n_tiles, packed.data(), n_tiles * 4, cudaMemcpyHostToDevice));
    }
    s->slots.assign(s->ring, tr_scene::Slot{});
    for (uint32_t r = 0; r < s->ring; ++r) {
        tr_scene::Slot& sl = s->slots[r];
        if ((rc = alloc_setup(&sl.cam, s->n_tri, true))) return rc;
        if ((rc = alloc_setup(&sl.lit, s->n_tri, false))) return rc;
        sl.bins.cells_x = (s->width + kCellW - 1) / kCellW;
        sl.bins.cells_y = (s->height + kCellH - 1) / kCellH;
        sl.bins.n_cells = sl.bins.cells_x * sl.bins.cells_y;
        sl.bins.parity = 0;
        CU(cudaMalloc(&sl.bins.count, (size_t)sl.bins.n_cells * 2 * 4));
        CU(cudaMemset(sl.bins.count, 0, (size_t)sl.bins.n_cells * 2 * 4));
        CU(cudaMalloc(&sl.bins.list, (size_t)sl.bins.n_cells * kBinCap * 4));
        CU(cudaMalloc(&sl.tile_state, n_tiles * 4));
        CU(cudaMemset(sl.tile_state, 0, n_tiles * 4));
        CU(cudaMalloc(&sl.keys, s->n_px * 8));
        CU(cudaMalloc(&sl.rgb, s->n_px * 3 + 16));
        CU(cudaMalloc(&sl.shadow, (s->n_px + 1) * 4));            // + the sentinel out-of-range lookups read (k_shade.cuh)
        {
            const float cleared = -3.402823466e38f;
            CU(cudaMemcpy(sl.shadow + s->n_px, &cleared, 4, cudaMemcpyHostToDevice));
        }
        CU(cudaMalloc(&sl.rgb_zero, n_tiles * 4));
        CU(cudaMemset(sl.rgb_zero, 0, n_tiles * 4));
        CU(cudaMalloc(&sl.shadow_clear, n_tiles * 4));
        CU(cudaMemset(sl.shadow_clear, 0, n_tiles * 4));
        CU(cudaMalloc(&sl.setup_diag, 2 * D_COUNT * 8));
        CU(cudaMemset(sl.setup_diag, 0, 2 * D_COUNT * 8));
        CU(cudaMalloc(&sl.light_diag, 2 * D_COUNT * 8));
        CU(cudaMemset(sl.light_diag, 0, 2 * D_COUNT * 8));
        if ((rc = make_rgb_map(s, &sl))) return rc;
        CU(cudaStreamCreateWithFlags(&sl.stream, cudaStreamNonBlocking));
        CU(cudaEventCreateWithFlags(&sl.done, cudaEventDisableTiming));
    }
    CU(cudaMalloc(&s->ao, ao_alloc_words(s->n_px) * 4));
    CU(cudaMemset(s->ao, 0, ao_alloc_words(s->n_px) * 4));
    CU(cudaMalloc(&s->diag, D_COUNT * 8));
    CU(cudaMemset(s->diag, 0, D_COUNT * 8));
    CU(cudaMalloc(&s->host_stats, 16));
    CU(cudaMemset(s->host_stats, 0, 16));
    const tr_image* imgs[4] = {&d->texture, &d->normal_map, &d->normal_map_tangent, &d->specular_map};
    for (int k = 0; k < 4; ++k) if ((rc = upload_map(s, k, *imgs[k]))) return rc;
    CU(cudaEventCreate(&s->ev_begin));
    CU(cudaEventCreate(&s->ev_end));
    CU(cudaEventCreateWithFlags(&s->ev_fork, cudaEventDisableTiming));
    if ((rc = reset_to_new(s))) return rc;
    CU(cudaStreamSynchronize(s->stream));
    return TR_OK;
}

// CUDA loads kernels lazily, and loading one may wait for the device to go idle: a rank's exchange kernel spinning at its
// barrier while another rank's FIRST launch of the same kernel is still being loaded would never be released (seen with
// two scenes on one GPU).  Every variant is therefore loaded when the peers are opened.
void preload_peer_kernels() {
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, k_ao_reduce_peers<0>);
    cudaFuncGetAttributes(&fa, k_ao_reduce_peers<2>);
    cudaFuncGetAttributes(&fa, k_ao_reduce_peers<4>);
    cudaFuncGetAttributes(&fa, k_ao_reduce_peers<8>);
    cudaFuncGetAttributes(&fa, k_ao_reduce_nvls);
    cudaGetLastError();
}

void close_peers(tr_scene* s) {
    for (uint32_t r = 0; r < s->peer_world; ++r)
        if (s->peer_ipc && r != s->peer_rank && s->peer_buf[r]) cudaIpcCloseMemHandle(s->peer_buf[r]);
    for (auto& p : s->peer_buf) p = nullptr;
    s->peer_world = 0;
}

void destroy_impl(tr_scene* s) {
    if (!s) return;
    cudaSetDevice(s->device);
    if (s->stream) cudaStreamSynchronize(s->stream);
    close_peers(s);
    if (s->peer_status) cudaFreeHost(s->peer_status);
    if (s->ao_count_host) cudaFreeHost(s->ao_count_host);
    if (s->ev_ao_count) cudaEventDestroy(s->ev_ao_count);
    for (auto& sp : s->spans) { cudaEventDestroy(sp.a); cudaEventDestroy(sp.b); }
    for (auto e : s->event_pool) cudaEventDestroy(e);
    if (s->ev_begin) cudaEventDestroy(s->ev_begin);
    if (s->ev_end) cudaEventDestroy(s->ev_end);
    if (s->ev_fork) cudaEventDestroy(s->ev_fork);
    for (int k = 0; k < 4; ++k) {
        if (s->maps.tex[k]) cudaDestroyTextureObject(s->maps.tex[k]);
        if (s->arrays[k]) cudaFreeArray(s->arrays[k]);
    }
    for (auto& sl : s->slots) {
        if (sl.stream) { cudaStreamSynchronize(sl.stream); cudaStreamDestroy(sl.stream); }
        if (sl.done) cudaEventDestroy(sl.done);
        cudaFree(sl.keys); cudaFree(sl.tile_state); cudaFree(sl.rgb); cudaFree(sl.shadow); cudaFree(sl.rgb_zero); cudaFree(sl.shadow_clear); cudaFree(sl.setup_diag); cudaFree(sl.light_diag);
        cudaFree(sl.cam.rec); cudaFree(sl.cam.var); cudaFree(sl.cam.bbox); cudaFree(sl.lit.rec); cudaFree(sl.lit.bbox);
        cudaFree(sl.bins.count); cudaFree(sl.bins.list);
    }
    cudaFree(s->geom); cudaFree(s->tile_order); cudaFree(s->tile_cost); cudaFree(s->tile_perm); cudaFree(s->tile_order_sorted); if (!s->ao_vmm) cudaFree(s->ao); cudaFree(s->ao_list.frag); cudaFree(s->ao_list.pix); cudaFree(s->ao_list.count); cudaFree(s->diag); cudaFree(s->host_stats); cudaFree(s->d_checksums); cudaFree(s->scratch);
    if (s->own_stream) cudaStreamDestroy(s->own_stream);
    delete s;
}

// ---- pass issue ------------------------------------------------------------------------------------------

// A lone frame (not part of an overlapped batch) runs the 256-thread tile kernel (k_raster.cuh); debug bit 1 forces the
// batch shape.  TINYR_LONE_SORT=0: lone frames keep the static centre-out launch order (A/B in profiles/r2_probes.md).
static const bool g_lone_sort = !(getenv("TINYR_LONE_SORT") && atoi(getenv("TINYR_LONE_SORT")) == 0);
static const uint32_t g_order_builds = getenv("TINYR_ORDER_BUILDS") ? (uint32_t)atoi(getenv("TINYR_ORDER_BUILDS")) : 2u;
bool lone_shape(const tr_scene* s) { return !s->overlapping && (s->debug_flags & 2u) == 0u; }
// TINYR_SETUP_CACHE=0: every lone frame runs the set-up kernel (A/B in profiles/r2_probes.md).
static const bool g_setup_cache = !(getenv("TINYR_SETUP_CACHE") && atoi(getenv("TINYR_SETUP_CACHE")) == 0);
bool setup_key_same(const tr_scene::Slot::SetupKey& a, const tr_scene::Slot::SetupKey& b) {
    auto eq = [](f3 p, f3 q) { return memcmp(&p, &q, sizeof(f3)) == 0; };      // bit for bit: -0.0 is not 0.0 here
    return a.valid && b.valid && a.pipeline == b.pipeline && eq(a.from, b.from) && eq(a.at, b.at) && eq(a.up, b.up) && eq(a.light, b.light);
}
// lone frames of at least two waves of tiles (800x800: 325 tiles, one wave whatever the order) record tile costs
bool lone_costs(const tr_scene* s) {
    const uint32_t n_tiles = s->tiles_x * s->tiles_y;
    return lone_shape(s) && g_lone_sort && (s->debug_flags & 4u) == 0u && n_tiles > 1024u && n_tiles <= (uint32_t)kSortMaxTiles;
}
tr_scene::CostView pass_view(const tr_scene* s, int pass) {
    tr_scene::CostView v;
    v.eye = pass == 0 ? f3{s->at.x + s->light.x, s->at.y + s->light.y, s->at.z + s->light.z} : s->from;
    v.at = s->at; v.up = s->up; v.pipeline = s->pipeline; v.valid = true;
    return v;
}
bool view_same(const tr_scene::CostView& a, const tr_scene::CostView& b) {
    auto eq = [](f3 p, f3 q) { return memcmp(&p, &q, sizeof(f3)) == 0; };
    return a.valid && b.valid && a.pipeline == b.pipeline && eq(a.eye, b.eye) && eq(a.at, b.at) && eq(a.up, b.up);
}
// The previous frame's costs predict this frame's when the view moved little: measured on the 4K turntable, 1 degree per
// frame keeps most of the gain of a still view, 5.6 degrees per frame has none left (profiles/r2_probes.md); the
// reference's app loop turns the camera by 3 rad/s x frame time (app.rs:12,178), i.e. well under 0.1 degree per frame.
bool view_close(const tr_scene::CostView& a, const tr_scene::CostView& b) {
    if (!a.valid || !b.valid) return false;
    const double kCos = 0.99905;           // cos(2.5 degrees)
    auto sub = [](f3 p, f3 q) { return f3{p.x - q.x, p.y - q.y, p.z - q.z}; };
    auto dot = [](f3 p, f3 q) { return (double)p.x * q.x + (double)p.y * q.y + (double)p.z * q.z; };
    const f3 da = sub(a.eye, a.at), db = sub(b.eye, b.at), dt = sub(a.at, b.at);
    const double la = std::sqrt(dot(da, da)), lb = std::sqrt(dot(db, db)), ua = std::sqrt(dot(a.up, a.up)), ub = std::sqrt(dot(b.up, b.up));
    if (!(la > 0.0 && lb > 0.0 && ua > 0.0 && ub > 0.0)) return false;
    return dot(da, db) >= kCos * la * lb && dot(a.up, b.up) >= kCos * ua * ub && std::fabs(la / lb - 1.0) <= 0.03 &&
           std::sqrt(dot(dt, dt)) <= 0.03 * lb;
}

int run_setup(tr_scene* s, const Uniforms& u, bool light_pass) {
    if (s->n_tri == 0) return TR_OK;
    const int pass = light_pass ? 0 : 1;
    const uint32_t n_tiles = s->tiles_x * s->tiles_y;
    s->order_sorted[pass] = false;
    S(s).setup_reused = false;
    const tr_scene::CostView now = pass_view(s, pass);
    // what this lone frame does about the tile kernel's launch order: re-use the one built for this very view, build one, or neither
    const bool costs = lone_costs(s) && view_close(s->cost_view[pass], now);
    const bool reuse_order = costs && view_same(s->order_view[pass], now) && s->order_builds[pass] >= g_order_builds && ++s->order_uses[pass] < 256u;
    const bool build_order = costs && !reuse_order;
    // The set-up output — records, varyings, bounding boxes, bins — is a pure function of model, pipeline, camera and light.
    // A lone frame of a single-pass pipeline whose inputs equal, bit for bit, those of the set-up launch that last wrote
    // this slot re-uses that output: no launch (the app loop with no key down re-renders one view; SURVEY.md 8(d) config 2
    // times 100 frames of one view).  The two passes of shadow / occlusion share the slot's bin lists: a light pass that
    // runs invalidates the camera pass's output (below); a light pass that is itself re-used (render_impl) leaves it alone.
    tr_scene::Slot::SetupKey key;
    key.from = s->from; key.at = s->at; key.up = s->up; key.light = s->light; key.pipeline = s->pipeline; key.valid = true;
    const bool cacheable = g_setup_cache && (s->debug_flags & 4u) == 0u && lone_shape(s) && !light_pass;
    if (cacheable && !build_order && setup_key_same(S(s).setup_key, key)) {
        S(s).setup_reused = true;
        s->order_sorted[pass] = reuse_order;
        return TR_OK;
    }
    Scoped sp(s, KS_SETUP);
    dim3 grid((s->n_tri + 127) / 128);
    S(s).bins.parity ^= 1u;                 // this pass fills the other counter set; the kernel zeroes the one just used
    S(s).cam.bins = S(s).bins;
    S(s).lit.bins = S(s).bins;
    // (the vertex roles of all pipelines live in one kernel: k_setup.cuh)
    SetupOut out = light_pass ? S(s).lit : S(s).cam;
    unsigned long long* diag = s->diag;
    S(s).setup_key.valid = false;
    if (light_pass && S(s).light_diag_cur != nullptr) {
        diag = S(s).light_diag_cur;
        out.diag_zero = S(s).light_diag + (size_t)((S(s).light_runs + 1u) & 1u) * D_COUNT;
    }
    if (cacheable) {
        S(s).setup_runs++;
        diag = S(s).setup_diag + (size_t)(S(s).setup_runs & 1u) * D_COUNT;
        out.diag_zero = S(s).setup_diag + (size_t)((S(s).setup_runs + 1u) & 1u) * D_COUNT;
        S(s).setup_key = key;
    }
    if (reuse_order) {
        // a view that stands still (the app loop with no key down): the order built for it stands too.  The sort block
        // ends 4 us after the vertex blocks (k_setup.cuh), which a frame that needs no new order does not have to wait for.
        s->order_sorted[pass] = true;
    } else if (build_order) {
        // one more block sorts the tile kernel's launch order by the tiles' cost in the previous lone frame
        s->order_builds[pass] = view_same(s->order_view[pass], now) ? s->order_builds[pass] + 1 : 1;
        s->order_view[pass] = now;
        s->order_uses[pass] = 0;
        out.sort.cost = s->tile_cost + (size_t)pass * n_tiles;
        out.sort.base = s->tile_order;
        out.sort.order = s->tile_order_sorted + (size_t)pass * n_tiles;
        out.sort.perm = s->tile_perm + (size_t)pass * n_tiles;
        out.sort.n_tiles = n_tiles;
        grid.x += 1;
        if (light_pass) k_vertex_setup<true, true><<<grid, 128, 0, s->stream>>>(s->geom, out, u, s->pipeline, diag);
        else k_vertex_setup<false, true><<<grid, 128, 0, s->stream>>>(s->geom, out, u, s->pipeline, diag);
        s->order_sorted[pass] = true;
        return launch_ok("k_vertex_setup");
    } else {
        s->order_builds[pass] = 0;          // (a frame with the static order in between: whatever order exists is stale)
    }
    if (light_pass) k_vertex_setup<true><<<grid, 128, 0, s->stream>>>(s->geom, out, u, s->pipeline, diag);
    else k_vertex_setup<false><<<grid, 128, 0, s->stream>>>(s->geom, out, u, s->pipeline, diag);
    return launch_ok("k_vertex_setup");
}

ShadeArgs shade_args(tr_scene* s, bool fresh_rgb) {
    ShadeArgs a{};
    a.keys = S(s).keys; a.tile_state = S(s).tile_state; a.tiles_x = s->tiles_x;
    a.rec = S(s).cam.rec; a.var = S(s).cam.var; a.shadow = S(s).shadow;
    a.rgb = S(s).rgb; a.ao_accum = s->ao; a.diag = s->diag; a.maps = s->maps; a.fresh = fresh_rgb ? 1 : 0;
    a.rgb_zero = S(s).rgb_zero; a.rgb_tma = S(s).rgb_tma ? 1 : 0;
    return a;
}

// The tile kernel is launched with programmatic stream serialization: it may start while the vertex set-up kernel in
// front of it (which signals griddepcontrol.launch_dependents on entry) is still running, does everything that does
// not depend on that kernel's output — tile coordinates, shared-memory set-up — and then waits (griddepcontrol.wait)
// for its completion and memory flush.  For a lone frame this hides the tile kernel's launch latency behind K1.
// TINYR_NO_PDL=1 in the environment restores plain launches (A/B in profiles/README.md).
static const bool g_pdl = getenv("TINYR_NO_PDL") == nullptr;
template <class PIPE, int T>
void launch_raster(tr_scene* s, dim3 grid, const RasterArgs& a, const ShadeArgs& sa, const Uniforms& u) {
    if (!g_pdl) {
        k_raster<PIPE, T><<<grid, T, 0, s->stream>>>(a, sa, u, S(s).rgb_map, S(s).rgb_map_top);
        return;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = dim3(T); cfg.dynamicSmemBytes = 0; cfg.stream = s->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, k_raster<PIPE, T>, a, sa, u, S(s).rgb_map, S(s).rgb_map_top);
}
template <class PIPE>
void launch_raster_views(tr_scene* s, dim3 grid, const ViewBatch& b) {
    constexpr int T = kRasterThreadsBatch;
    if (!g_pdl) {
        k_raster_views<PIPE, T><<<grid, T, 0, s->stream>>>(b);
        return;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = dim3(T); cfg.dynamicSmemBytes = 0; cfg.stream = s->stream;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at; cfg.numAttrs = 1;
    cudaLaunchKernelEx(&cfg, k_raster_views<PIPE, T>, b);
}

// Arguments of the tile kernel for the current slot.
RasterArgs raster_args(tr_scene* s, bool depth_only, uint32_t epoch, bool fresh, bool row_order) {
    RasterArgs a{};
    a.rec = depth_only ? S(s).lit.rec : S(s).cam.rec;
    a.bbox = depth_only ? S(s).lit.bbox : S(s).cam.bbox;
    a.bins = S(s).bins;
    a.n_triangles = s->n_tri; a.width = s->width; a.height = s->height; a.tiles_x = s->tiles_x;
    a.epoch = epoch; a.fresh = fresh ? 1 : 0;
    a.keys = S(s).keys; a.tile_state = S(s).tile_state; a.depth = S(s).shadow; a.depth_clear = S(s).shadow_clear; a.diag = s->diag;
    a.tile_order = s->tile_order + (row_order ? (size_t)s->tiles_x * s->tiles_y : 0);
    a.n_tiles = s->tiles_x * s->tiles_y;
    return a;
}

// Several cleared views in one launch per stage (k_raster_views / k_vertex_setup_views): view j renders into slot
// slots[j] with uniforms cam[j] (and lit[j] for the light pass of the two-pass pipelines).  Everything on s->stream.
int run_views_multi(tr_scene* s, const uint32_t* slots, const Uniforms* lit, const Uniforms* cam, uint32_t m) {
    const PipelineEntry& pipe = kPipelines[s->pipeline];
    const uint32_t keep = s->cur;
    const dim3 sgrid((s->n_tri + 127) / 128, m), rgrid(s->tiles_x * s->tiles_y, m);
    int rc = TR_OK;
    for (int pass = pipe.light_pass ? 0 : 1; pass < 2 && rc == TR_OK; ++pass) {
        const bool light = pass == 0;
        static SetupViewBatch sb;           // (parameter blocks are copied by the launch; static: 10 KB off the stack)
        static ViewBatch vb;
        static std::mutex mu;
        std::lock_guard<std::mutex> lock(mu);
        for (uint32_t j = 0; j < m; ++j) {
            s->cur = slots[j];
            const Uniforms& u = light ? lit[j] : cam[j];
            S(s).bins.parity ^= 1u;
            S(s).cam.bins = S(s).bins;
            S(s).lit.bins = S(s).bins;
            S(s).setup_key.valid = false;       // (the slot's set-up output is overwritten)
            if (light) S(s).shadow_key.valid = false;
            sb.v[j].out = light ? S(s).lit : S(s).cam;
            sb.v[j].u = u;
            vb.v[j].a = raster_args(s, light, 0, true, true);
            vb.v[j].sa = shade_args(s, true);
            vb.v[j].u = u;
            vb.v[j].rgb_map = S(s).rgb_map;
            vb.v[j].rgb_map_top = S(s).rgb_map_top;
        }
        if (s->n_tri) {
            s->kernel_launches[KS_SETUP]++;
            if (light) k_vertex_setup_views<true><<<sgrid, 128, 0, s->stream>>>(s->geom, sb, s->pipeline, s->diag);
            else k_vertex_setup_views<false><<<sgrid, 128, 0, s->stream>>>(s->geom, sb, s->pipeline, s->diag);
            if ((rc = launch_ok("k_vertex_setup_views"))) break;
        }
        s->kernel_launches[light ? KS_SHADOW : KS_RASTER]++;
        if (light) launch_raster_views<PassLightDepth>(s, rgrid, vb);
        else { pipe.raster_views(s, rgrid, vb); s->kernel_launches[KS_SHADE]++; }
        rc = launch_ok("k_raster_views");
    }
    s->cur = keep;
    return rc;
}

// mode: kModeDepthOnly (K4), kModeVisibility (K3 alone) or a pipeline id (K3 + K5 fused).
int run_raster(tr_scene* s, int mode, uint32_t epoch, bool fresh, const Uniforms& u, bool fresh_rgb) {
    const bool depth_only = mode == kModeDepthOnly;
    RasterArgs a = raster_args(s, depth_only, epoch, fresh, s->overlapping);
    ShadeArgs sa = shade_args(s, fresh_rgb);
    // (tile order: centre-out shortens a lone frame's tail; overlapped batches do better in plain row order, where one
    // frame's cheap border tiles run beside the previous frame's heavy ones — measured: profiles/README.md)
    dim3 grid(s->tiles_x * s->tiles_y);
    Scoped sp(s, depth_only ? KS_SHADOW : KS_RASTER);
    // CTA shape: 256 threads for a lone frame, 128 for overlapped batches (k_raster.cuh)
    const int shape = lone_shape(s) ? 0 : 1;      // (debug bit 1: time the batch shape on a lone frame)
    const int pass = depth_only ? 0 : 1;
    if (S(s).setup_key.valid && !depth_only)      // (valid: the slot's set-up output comes from a lone camera pass that counted into its own block)
        a.setup_diag = S(s).setup_diag + (size_t)(S(s).setup_runs & 1u) * D_COUNT;
    if (depth_only) {
        S(s).shadow_key.valid = false;              // (render_impl sets it again when this is a light pass it may re-use)
        if (S(s).light_diag_cur != nullptr) a.diag = S(s).light_diag_cur;
    } else if (S(s).shadow_key.valid && kPipelines[s->pipeline].light_pass && mode >= 0) {
        a.light_diag = S(s).light_diag + (size_t)(S(s).light_runs & 1u) * D_COUNT;
    }
    if (lone_costs(s)) {
        if (s->order_sorted[pass]) {
            a.tile_order = s->tile_order_sorted + (size_t)pass * a.n_tiles;
            a.tile_perm = s->tile_perm + (size_t)pass * a.n_tiles;
        }
        // (a view that moves fast — 5.6 degrees per frame on the bench's turntable — records nothing and pays nothing:
        // the recording costs such a frame about 1 us; a view that has settled is sorted from its third frame on)
        const tr_scene::CostView now = pass_view(s, pass);
        if (s->order_sorted[pass] || view_close(s->last_view[pass], now)) {
            a.tile_cost = s->tile_cost + (size_t)pass * a.n_tiles;
            s->cost_view[pass] = now;
        }
        s->last_view[pass] = now;
    }
    s->order_sorted[pass] = false;
    if (mode == kModeDepthOnly) kRasterLightDepth[shape](s, grid, a, sa, u);
    else if (mode == kModeVisibility) kRasterVisibility[shape](s, grid, a, sa, u);
    else kPipelines[mode].raster[shape](s, grid, a, sa, u);
    if (!depth_only && mode != kModeVisibility) s->kernel_launches[KS_SHADE]++;     // K5 runs inside this launch
    return launch_ok("k_raster");
}

// K6: deferred occlusion pass of the many-light AO batch.
// K6 part 1: compact the covered pixels of the camera slot (depth + world position), once per AO batch.  The count
// travels to the host asynchronously (mapped word + event): tr_scene_render_ao only waits for it after it has queued the
// depth passes of the first lights, so the device never idles on the round trip.
int run_ao_compact(tr_scene* s, const Uniforms& u) {
    if (!s->ao_list.frag) {
        CU(cudaMalloc(&s->ao_list.frag, s->n_px * sizeof(float4)));
        CU(cudaMalloc(&s->ao_list.pix, s->n_px * 4));
        CU(cudaMalloc(&s->ao_list.count, 4));
        CU(cudaHostAlloc((void**)&s->ao_count_host, 4, cudaHostAllocDefault));
        CU(cudaEventCreateWithFlags(&s->ev_ao_count, cudaEventDisableTiming));
    }
    CU(cudaMemsetAsync(s->ao_list.count, 0, 4, s->stream));
    ShadeArgs a = shade_args(s, false);
    dim3 grid((s->width + kShadeThreads - 1) / kShadeThreads, s->height);
    {
        Scoped sp(s, KS_AO);
        k_ao_compact<<<grid, kShadeThreads, 0, s->stream>>>(a, u, s->ao_list);
    }
    int rc = launch_ok("k_ao_compact");
    if (rc) return rc;
    CU(cudaMemcpyAsync(s->ao_count_host, s->ao_list.count, 4, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaEventRecord(s->ev_ao_count, s->stream));
    return TR_OK;
}

// K6 part 2: one light's occlusion term for every covered pixel, added into the accumulator.
int run_ao_accumulate(tr_scene* s, const Uniforms& u, uint32_t n_covered) {
    if (n_covered == 0) return TR_OK;
    ShadeArgs a = shade_args(s, false);                  // shadow buffer of the current (light) slot
    dim3 grid((n_covered + kShadeThreads - 1) / kShadeThreads);
    Scoped sp(s, KS_AO);
    k_ao_accumulate<<<grid, kShadeThreads, 0, s->stream>>>(a, u, s->ao_list, n_covered);
    return launch_ok("k_ao_accumulate");
}

// Scene::render (scene.rs:151-268): every pass = prepare, all polygons through the vertex stage, raster.
int render_impl(tr_scene* s) {
    CU(cudaSetDevice(s->device));
    if (s->epoch > kMaxEpoch) {
        // 255 renders without a clear: fold the epochs of the stored winners into epoch 0 and go on (k_misc.cuh)
        Scoped sp(s, KS_CLEAR);
        k_collapse_epochs<<<dim3(s->tiles_x, s->tiles_y), 256, 0, s->stream>>>(S(s).keys, S(s).tile_state, s->width, s->height);
        int rc0 = launch_ok("k_collapse_epochs");
        if (rc0) return rc0;
        s->epoch = 1;
    }
    int rc;
    Uniforms u{};
    u.n_triangles = s->n_tri; u.epoch = s->epoch;
    // every pass = its prepare role, all polygons through its vertex role, raster + its fragment role (scene.rs:200-266)
    const PipelineEntry& pipe = kPipelines[s->pipeline];
    if (pipe.light_pass) {
        if ((rc = prep_error(prepare_light(u, s->width, s->height, s->light, s->at, s->up)))) return rc;
        u.n_triangles = s->n_tri; u.epoch = s->epoch;
        // The shadow buffer of a lone frame is a pure function of model, light, target and up when the pass starts from a
        // cleared buffer — and merging the same pass into it again (a render without clear) changes nothing.  While those
        // inputs stay bit-identical and nothing else has written to the buffer, later frames re-use it: no set-up launch,
        // no depth raster (a camera that moves under a fixed light; the app loop with no key down).
        tr_scene::Slot& sl = S(s);
        tr_scene::Slot::SetupKey lk;
        lk.light = s->light; lk.at = s->at; lk.up = s->up; lk.pipeline = -2; lk.valid = true;
        const bool may_cache = g_setup_cache && (s->debug_flags & 4u) == 0u && lone_shape(s) && s->n_tri != 0;
        if (!(may_cache && setup_key_same(sl.shadow_key, lk))) {
            const bool keep = may_cache && s->fresh_shadow;     // "cleared + this pass" is what the buffer will hold
            if (keep) { sl.light_runs++; sl.light_diag_cur = sl.light_diag + (size_t)(sl.light_runs & 1u) * D_COUNT; }
            rc = run_setup(s, u, true);
            if (rc == TR_OK) rc = run_raster(s, kModeDepthOnly, s->epoch, s->fresh_shadow, u, false);
            sl.light_diag_cur = nullptr;
            if (rc) return rc;
            if (keep) sl.shadow_key = lk;
        }
        s->fresh_shadow = false;
    }
    if ((rc = prep_error(pipe.prepare(u, s->width, s->height, s->light, s->from, s->at, s->up)))) return rc;
    u.debug_flags = s->debug_flags;
    if ((rc = run_setup(s, u, false))) return rc;
    if ((rc = run_raster(s, s->pipeline, s->epoch, s->fresh_keys, u, s->fresh_rgb))) return rc;
    s->fresh_keys = false;
    s->fresh_rgb = false;
    s->uni = u;
    s->epoch++;
    return TR_OK;
}

int checksum_into(tr_scene* s, const uint8_t* img, size_t n_bytes, unsigned long long* d_out) {
    CU(cudaMemsetAsync(d_out, 0, 8, s->stream));
    Scoped sp(s, KS_CHECKSUM);
    k_checksum<<<148 * 4, 256, 0, s->stream>>>(reinterpret_cast<const uint32_t*>(img), n_bytes / 4, d_out);
    return launch_ok("k_checksum");
}

int deliver_frame(tr_scene* s, uint8_t* dst);      // tinyr_batch.inc: sparse store into tr_host_alloc memory, else a plain copy

}  // namespace

// ============================================ C ABI ==========================================================
#define NEED(s) do { if (!(s)) return fail(TR_ERR_INVALID_ARGUMENT, "%s: null scene", __func__); clear_error(); } while (0)

extern "C" {

const char* tr_last_error(void) { return last_error(); }
const char* tr_version(void) { return TR_VERSION_STRING; }

int tr_device_count(int* n) {
    if (!n) return fail(TR_ERR_INVALID_ARGUMENT, "null argument");
    *n = 0;
    if (cudaGetDeviceCount(n) != cudaSuccess) { cudaGetLastError(); *n = 0; return fail(TR_ERR_NO_DEVICE, "no CUDA device"); }
    return TR_OK;
}

int tr_scene_create(tr_scene** out, const tr_scene_desc* desc) {
    if (!out || !desc) return fail(TR_ERR_INVALID_ARGUMENT, "tr_scene_create: null argument");
    *out = nullptr;
    clear_error();
    tr_scene* s = new tr_scene();
    int rc = create_impl(s, desc);
    if (rc) { std::string keep = last_error(); destroy_impl(s); fail(rc, "%s", keep.c_str()); return rc; }
    *out = s;
    return TR_OK;
}

void tr_scene_destroy(tr_scene* s) { destroy_impl(s); }

// Scene::clear (scene.rs:128-137).  Lazy: the z/shadow/frame fills are folded into the tile initialisation of
// the next render; getters called before that materialise them.
int tr_scene_clear(tr_scene* s) {
    NEED(s);
    s->fresh_keys = s->fresh_rgb = s->fresh_shadow = true;
    s->epoch = 0;
    return TR_OK;
}

int tr_scene_set_light_direction(tr_scene* s, const float d[3]) {
    NEED(s);
    if (!d) return fail(TR_ERR_INVALID_ARGUMENT, "null direction");
    s->light = {d[0], d[1], d[2]};
    return TR_OK;
}

int tr_scene_set_camera(tr_scene* s, const float from[3], const float at[3], const float up[3]) {
    NEED(s);
    if (!from || !at || !up) return fail(TR_ERR_INVALID_ARGUMENT, "null camera vector");
    s->from = {from[0], from[1], from[2]}; s->at = {at[0], at[1], at[2]}; s->up = {up[0], up[1], up[2]};
    return TR_OK;
}

int tr_scene_render(tr_scene* s) { NEED(s); return render_impl(s); }

int tr_scene_synchronize(tr_scene* s) {
    NEED(s);
    CU(cudaSetDevice(s->device));
    CU(cudaStreamSynchronize(s->stream));
    return TR_OK;
}

int tr_scene_set_pipeline(tr_scene* s, const char* name) {
    NEED(s);
    int p = pipeline_from_name(name);
    if (p < 0) return fail(TR_ERR_UNKNOWN_PIPELINE, "Provided pipeline name is not supported! (%s)", name ? name : "null");
    CU(cudaSetDevice(s->device));
    s->pipeline = p;
    return reset_to_new(s);
}

int tr_scene_set_stream(tr_scene* s, void* stream) {
    NEED(s);
    CU(cudaSetDevice(s->device));
    CU(cudaStreamSynchronize(s->stream));
    s->stream = stream ? (cudaStream_t)stream : s->own_stream;
    return TR_OK;
}

// get_frame_buffer (scene.rs:92-97): frames are stored top-down on the device, so this is one D2H copy.
int tr_scene_get_frame_buffer(tr_scene* s, uint8_t* dst) {
    NEED(s);
    if (!dst) return fail(TR_ERR_INVALID_ARGUMENT, "null destination");
    CU(cudaSetDevice(s->device));
    int rc = materialise(s);
    if (rc) return rc;
    if ((rc = deliver_frame(s, dst))) return rc;
    CU(cudaStreamSynchronize(s->stream));
    return TR_OK;
}

int tr_scene_read_frame_raw(tr_scene* s, uint8_t* dst) {
    NEED(s);
    if (!dst) return fail(TR_ERR_INVALID_ARGUMENT, "null destination");
    CU(cudaSetDevice(s->device));
    int rc = materialise(s);
    if (rc) return rc;
    if ((rc = ensure_scratch(s, s->n_px * 3))) return rc;
    size_t n = s->n_px * 3;
    k_flip_rows<<<(unsigned)((n + 255) / 256), 256, 0, s->stream>>>(S(s).rgb, s->scratch, s->width * 3, s->height);
    if ((rc = launch_ok("k_flip_rows"))) return rc;
    CU(cudaMemcpyAsync(dst, s->scratch, n, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return TR_OK;
}

static int gray_view(tr_scene* s, uint8_t* dst, bool shadow) {
    if (!dst) return fail(TR_ERR_INVALID_ARGUMENT, "null destination");
    CU(cudaSetDevice(s->device));
    int rc = materialise(s);
    if (rc) return rc;
    if ((rc = ensure_scratch(s, s->n_px * 3))) return rc;
    k_gray_view<<<(unsigned)((s->n_px + 255) / 256), 256, 0, s->stream>>>(shadow ? nullptr : S(s).keys, S(s).shadow, s->scratch, s->width, s->height);
    if ((rc = launch_ok("k_gray_view"))) return rc;
    CU(cudaMemcpyAsync(dst, s->scratch, s->n_px * 3, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return TR_OK;
}
int tr_scene_get_z_buffer(tr_scene* s, uint8_t* dst) { NEED(s); return gray_view(s, dst, false); }
int tr_scene_get_shadow_buffer(tr_scene* s, uint8_t* dst) { NEED(s); return gray_view(s, dst, true); }

static int read_unpacked(tr_scene* s, float* depth, uint32_t* tri) {
    CU(cudaSetDevice(s->device));
    int rc = materialise(s);
    if (rc) return rc;
    if ((rc = ensure_scratch(s, s->n_px * 4))) return rc;
    k_unpack_keys<<<(unsigned)((s->n_px + 255) / 256), 256, 0, s->stream>>>(S(s).keys, depth ? (float*)s->scratch : nullptr,
                                                                         tri ? (uint32_t*)s->scratch : nullptr, s->n_px);
    if ((rc = launch_ok("k_unpack_keys"))) return rc;
    CU(cudaMemcpyAsync(depth ? (void*)depth : (void*)tri, s->scratch, s->n_px * 4, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return TR_OK;
}
int tr_scene_read_depth(tr_scene* s, float* dst) { NEED(s); if (!dst) return fail(TR_ERR_INVALID_ARGUMENT, "null destination"); return read_unpacked(s, dst, nullptr); }
int tr_scene_read_visibility(tr_scene* s, uint32_t* dst) { NEED(s); if (!dst) return fail(TR_ERR_INVALID_ARGUMENT, "null destination"); return read_unpacked(s, nullptr, dst); }

int tr_scene_read_shadow(tr_scene* s, float* dst) {
    NEED(s);
    if (!dst) return fail(TR_ERR_INVALID_ARGUMENT, "null destination");
    CU(cudaSetDevice(s->device));
    int rc = materialise(s);
    if (rc) return rc;
    CU(cudaMemcpyAsync(dst, S(s).shadow, s->n_px * 4, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return TR_OK;
}

int tr_scene_diagnostics(tr_scene* s, uint64_t out[8]) {
    NEED(s);
    if (!out) return fail(TR_ERR_INVALID_ARGUMENT, "null destination");
    CU(cudaSetDevice(s->device));
    CU(cudaMemcpyAsync(out, s->diag, D_COUNT * 8, cudaMemcpyDeviceToHost, s->stream));
    CU(cudaStreamSynchronize(s->stream));
    return TR_OK;
}

int tr_scene_get_uniforms(tr_scene* s, float out[102]) {
    NEED(s);
    if (!out) return fail(TR_ERR_INVALID_ARGUMENT, "null destination");
    const mat4* ms[6] = {&s->uni.vpmv, &s->uni.i_vpmv, &s->uni.m, &s->uni.i_m, &s->uni.it_m, &s->uni.shadow};
    for (int k = 0; k < 6; ++k) memcpy(out + 16 * k, ms[k]->m, 64);
    out[96] = s->uni.camera_direction.x; out[97] = s->uni.camera_direction.y; out[98] = s->uni.camera_direction.z;
    out[99] = s->uni.t_light.x; out[100] = s->uni.t_light.y; out[101] = s->uni.t_light.z;
    return TR_OK;
}

// The hoisted per-fragment constants of the shadow / occlusion fragments (tr_prepare.h), for parity tests against the
// values the reference recomputes per fragment.
int tr_scene_get_ao_uniforms(tr_scene* s, float out[64]) {
    NEED(s);
    if (!out) return fail(TR_ERR_INVALID_ARGUMENT, "null destination");
    for (int k = 0; k < 16; ++k) { out[3 * k] = s->uni.ao_step[k].x; out[3 * k + 1] = s->uni.ao_step[k].y; out[3 * k + 2] = s->uni.ao_step[k].z; }
    memcpy(out + 48, s->uni.shadow_from_screen.m, 64);
    return TR_OK;
}

int tr_scene_set_debug_flags(tr_scene* s, uint32_t flags) {
    NEED(s);
    s->debug_flags = flags;
    return TR_OK;
}

int tr_scene_device_frame(tr_scene* s, void** p) {
    NEED(s);
    if (!p) return fail(TR_ERR_INVALID_ARGUMENT, "null destination");
    CU(cudaSetDevice(s->device));
    int rc = materialise(s);
    *p = S(s).rgb;
    return rc;
}

}  // extern "C"

#include "tinyr_batch.inc"
#include "tinyr_multi.inc"

#ifdef TR_TIMELINE
// Diagnostic build only (see k_raster.cuh): copies the per-CTA timeline of the last tile-kernel launch on the current device.
extern "C" int tr_debug_timeline(uint64_t* out, uint32_t n_ctas) {
    using namespace tr;
    if (!out || n_ctas > (uint32_t)kTimelineCtas) return fail(TR_ERR_INVALID_ARGUMENT, "tr_debug_timeline: bad argument");
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpyFromSymbol(out, g_timeline, (size_t)n_ctas * kTimelineWords * 8));
    return TR_OK;
}
extern "C" int tr_debug_timeline_setup(uint64_t* out512) {
    using namespace tr;
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpyFromSymbol(out512, g_timeline_setup, 512 * 8));
    return TR_OK;
}
extern "C" int tr_debug_timeline_phase(uint64_t* out1024) {
    using namespace tr;
    CU(cudaDeviceSynchronize());
    CU(cudaMemcpyFromSymbol(out1024, g_timeline_phase, 1024 * 8));
    return TR_OK;
}
#endif
