/* tinyr.h — C ABI of the B200-native rasterization backend for litzendraht/tiny_renderer.
 *
 * This is the drop-in boundary.  The reference has no FFI; what it has is the `Scene` public API
 * (/root/reference/src/scene.rs) driven by app::run (/root/reference/src/app.rs:137-146,170,208-213).
 * Every entry point below names the reference item it replaces.  Plain pointers and sizes only; no C++,
 * CUDA or torch types.  All functions return a tr_status (0 = OK, negative = error) unless noted, never
 * abort the process, and leave a message retrievable with tr_last_error() (thread-local).
 *
 * Threading (scene.rs takes &mut self everywhere): a tr_scene is NOT thread-safe; distinct scenes may be
 * driven from distinct threads.  All device work of a scene is issued on one CUDA stream; tr_scene_render
 * is asynchronous, every getter synchronises.
 *
 * Pixel conventions: the reference's frame_buffer has pixel (0,0) at the bottom-left, index i + j*w
 * (scene.rs:24,260); Scene::get_frame_buffer returns the vertically flipped (top-down) image
 * (scene.rs:92-97).  "topdown" getters mirror the reference getters, "raw" getters use i + j*w.
 */
#ifndef TINYR_H
#define TINYR_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define TR_VERSION_STRING "tinyr 0.1 (sm_100a)"
#define TR_NO_TRIANGLE 0xFFFFFFFFu

typedef enum tr_status {
    TR_OK = 0,
    TR_ERR_INVALID_ARGUMENT = -1,
    TR_ERR_UNKNOWN_PIPELINE = -2,     /* reference: panic!("Provided pipeline name is not supported!") shader.rs:108 */
    TR_ERR_BAD_POLYGON = -3,          /* reference: panic!("Encountered some garbage...") scene.rs:218 */
    TR_ERR_IO = -4,                   /* reference: `?` on File::open / image::open, app.rs:94,99 */
    TR_ERR_FORMAT = -5,               /* malformed OBJ / TGA */
    TR_ERR_CUDA = -6,
    TR_ERR_NO_DEVICE = -7,            /* no CUDA device: there is NO CPU fallback */
    TR_ERR_REFERENCE_UNDEFINED = -8,  /* inputs for which the reference panics (singular basis, light = -z in occlusion, ...) */
    TR_ERR_RESERVED_9 = -9,           /* (was TR_ERR_NCCL: the library links no NCCL; the sum over GPUs is its own peer-memory kernel) */
    TR_ERR_UNSUPPORTED = -10
} tr_status;

typedef struct tr_scene tr_scene;

/* image::RgbImage as produced by image::open(..).into_rgb8() (app.rs:99): RGB8, row 0 = top. */
typedef struct tr_image {
    const uint8_t* rgb8_topdown;
    uint32_t width, height;
} tr_image;

/* Arguments of Scene::new (scene.rs:47-56).  obj::raw::RawObj is flattened to arrays:
 *   positions  n_positions x 3 (x,y,z; obj `w` is ignored by the reference, util.rs:25-31)
 *   uvs        n_uvs x 2       (u,v;   3rd component ignored, shader.rs:144)
 *   normals    n_normals x 3
 *   tri_idx    n_triangles x 9 (p0 t0 n0 p1 t1 n1 p2 t2 n2), 0-based; only the first three corners of a
 *              polygon are used by the reference (scene.rs:224-226)
 * The caller keeps ownership; create copies everything to the device (Model is immutable afterwards). */
typedef struct tr_scene_desc {
    uint32_t width, height;
    const float* positions;   uint32_t n_positions;
    const float* uvs;         uint32_t n_uvs;
    const float* normals;     uint32_t n_normals;
    const uint32_t* tri_idx;  uint32_t n_triangles;
    tr_image texture, normal_map, normal_map_tangent, specular_map;
    const char* pipeline_name;   /* default phong normal_map (alias true_normal) specular darboux shadow occlusion */
    int32_t device;              /* CUDA device ordinal */
    uint32_t ring;               /* number of frame/visibility buffer sets to rotate through in batch calls (0 -> 1) */
} tr_scene_desc;

/* ---- Scene surface (scene.rs) ------------------------------------------------------------------------- */
int tr_scene_create(tr_scene** out, const tr_scene_desc* desc);                  /* Scene::new            scene.rs:47  */
void tr_scene_destroy(tr_scene* s);                                              /* Drop                                */
int tr_scene_clear(tr_scene* s);                                                 /* Scene::clear          scene.rs:128 */
int tr_scene_set_light_direction(tr_scene* s, const float dir[3]);               /* set_light_direction   scene.rs:140 */
int tr_scene_set_camera(tr_scene* s, const float look_from[3], const float look_at[3], const float up[3]); /* scene.rs:145 */
int tr_scene_render(tr_scene* s);                                                /* Scene::render         scene.rs:151 */
/* (tr_scene_render computes what Scene::render computes; passes whose inputs are bit-identical to the previous frame's —
 * the vertex stage of an unchanged view, the light pass of an unchanged light — are not run again, their output is
 * re-used; buffers, getters and diagnostics are the same as if they had run.  TINYR_SETUP_CACHE=0 disables that.) */
int tr_scene_get_frame_buffer(tr_scene* s, uint8_t* dst_rgb8_topdown);           /* get_frame_buffer      scene.rs:92  */
int tr_scene_get_z_buffer(tr_scene* s, uint8_t* dst_rgb8_topdown);               /* get_z_buffer          scene.rs:101 */
int tr_scene_get_shadow_buffer(tr_scene* s, uint8_t* dst_rgb8_topdown);          /* get_shadow_buffer     scene.rs:115 */

/* ---- additions: raw state, visibility, diagnostics ---------------------------------------------------- */
int tr_scene_read_frame_raw(tr_scene* s, uint8_t* dst_rgb8_bottomup);            /* Scene.frame_buffer as stored, i + j*w */
int tr_scene_read_depth(tr_scene* s, float* dst);                                /* Buffer.z_buffer       shader.rs:18 */
int tr_scene_read_shadow(tr_scene* s, float* dst);                               /* Buffer.shadow_buffer  shader.rs:19 */
int tr_scene_read_visibility(tr_scene* s, uint32_t* dst_tri_id);                 /* winning polygon index per pixel, TR_NO_TRIANGLE if none */
int tr_scene_set_pipeline(tr_scene* s, const char* pipeline_name);               /* ShaderPipeline::new   shader.rs:97 (buffers re-zeroed) */
int tr_scene_synchronize(tr_scene* s);
/* out[0] shadow index >= w*h (reference panics, shader.rs:778,912,935)   out[1] round(x) >= w (reference reads the next row)
 * out[2] texel out of range (reference panics, util.rs:40)                out[3] NaN depth
 * out[4] homogeneous w == 0 (reference unwrap panic, shader.rs:158)       out[5] singular darboux basis (shader.rs:631)
 * out[6] triangles binned  out[7] reserved */
int tr_scene_diagnostics(tr_scene* s, uint64_t out[8]);
/* The uniforms of the last prepare (shader.rs:183-279), row-major 4x4: vpmv, i_vpmv, m, i_m, it_m, shadow,
 * then camera_direction[3], t_light_direction[3] — 102 floats. */
int tr_scene_get_uniforms(tr_scene* s, float out[102]);
/* The constants K0 hoists out of the shadow / occlusion fragments (shader.rs:763-764, 899-900, 916-929): out[0..48) the 16
 * occlusion sample offsets rot * (sin a_k, 0, cos a_k) * 0.02 (xyz each; occlusion pipeline only), out[48..64)
 * shadow_matrix * i_vpmv, row-major. */
int tr_scene_get_ao_uniforms(tr_scene* s, float out[64]);

/* ---- batch modes (BASELINE.json configs 4 and 5) -------------------------------------------------------- */
typedef struct tr_view {
    float look_from[3], look_at[3], up[3], light_direction[3];
} tr_view;
/* For each view: set_light_direction, set_camera, clear, render (app.rs:170,208-210).  checksums (optional,
 * host, n) receives a 64-bit checksum of each top-down RGB8 frame computed on the device (see
 * tr_checksum_rgb8); frames (optional, host, n*3*w*h) receives every frame.  With ring > 1 (tr_scene_desc) consecutive
 * frames overlap on the device.  Returns after the host outputs have landed; with neither output requested the call
 * is asynchronous on the scene stream like tr_scene_render. */
int tr_scene_render_views(tr_scene* s, const tr_view* views, uint32_t n_views, uint64_t* checksums, uint8_t* frames_rgb8_topdown);
/* Many-light ambient occlusion: for each light `set_light_direction; clear; render` with the occlusion
 * pipeline (shader.rs:806-963) and accum[i + j*w] += gray (u32).  Camera visibility is computed once.
 * The sums stay on the device (tr_scene_ao_device_buffer); accum_host (optional) receives a copy.
 * reset != 0 zeroes the accumulator first.  Afterwards the shadow-buffer getters show the depth pass of ONE of the
 * lights (the last one with ring == 1; with ring > 1 whichever pass ran last on the camera's slot, or the earlier
 * contents when none did): they are a debug view, not part of the batch's result. */
int tr_scene_render_ao(tr_scene* s, const float* light_dirs_xyz, uint32_t n_lights, int reset, uint32_t* accum_host);
int tr_scene_ao_device_buffer(tr_scene* s, void** device_ptr_u32, size_t* n_elements);
int tr_scene_read_ao(tr_scene* s, uint32_t* accum_host);
/* Multi-GPU sum of the accumulators, one process per GPU (the reference renders its
 * light passes in one loop on one device, shader.rs:806-963, so there is no reference interface to cite: this is the
 * exchange step of the sharded batch).  Lights are split over the ranks, every rank runs tr_scene_render_ao on its
 * share, then ONE kernel per rank adds the accumulators over NVLink peer memory (k_peer.cuh): no NCCL call, no host
 * staging; afterwards every rank's accumulator holds the total (u32 sums: bit-identical for any rank count).
 *   tr_scene_ao_ipc_handle    64-byte CUDA IPC handle of this scene's accumulator, to be passed to the other ranks
 *   tr_scene_ao_open_peers    handles = world x 64 bytes in rank order (entry `rank` is ignored); world <= 16; world == 1
 *                             needs no handles.  Maps the peers' accumulators; call again after the set of ranks changes
 *   tr_scene_ao_reduce_peers  collective, asynchronous on the scene stream: every rank calls it once per reduction.  A
 *                             rank that never arrives makes the others time out after 10 s: tr_scene_read_ao then fails
 *                             and keeps failing until tr_scene_ao_open_peers is called again on every rank (it resets the
 *                             barrier flags; the ranks must synchronise between it and the next reduction).  The kernel's
 *                             barrier needs the 148 CTAs of every rank's launch resident at the same time: true for one
 *                             rank per GPU (CTAs are dispatched in index order and a CTA only waits for its own index on
 *                             the other ranks), and for at most two scenes sharing one GPU
 *   tr_scene_ao_close_peers   unmaps (also done by tr_scene_destroy) */
int tr_scene_ao_ipc_handle(tr_scene* s, void* handle64);
int tr_scene_ao_open_peers(tr_scene* s, const void* handles, uint32_t world, uint32_t rank);
int tr_scene_ao_reduce_peers(tr_scene* s);
int tr_scene_ao_close_peers(tr_scene* s);
/* ---- native multi-GPU driver: one host process, one scene per device (tinyr_multi.inc) -------------------------------
 * The reference's process entry (main.rs:12-36) drives one Scene; these drive the two batch modes that shard naturally
 * over the devices of one box without Python, torch.distributed or NCCL.  `devices` lists CUDA ordinals (a device may
 * appear twice — not more often — : two scenes on one GPU, which lets a single-GPU box exercise the sharded paths).
 *   tr_multi_render_views  device i renders views [i n / G, (i + 1) n / G); checksums / frames (optional, host, n entries
 *                          in VIEW order) are written by the devices directly; no data-path exchange
 *   tr_multi_render_ao     occlusion pipeline: device i accumulates lights [i n / G, (i + 1) n / G), then the accumulators
 *                          are summed over the devices by k_ao_reduce_peers over NVLink peer memory (cudaDeviceEnablePeerAccess);
 *                          accum_host (optional, w*h u32) receives the total; bit-identical for any G
 *   tr_multi_set_camera    the same camera on every device's scene (lights / views are per call)
 *   tr_multi_scene         the i-th device's scene, for everything else (getters, diagnostics, replicas of single frames) */
typedef struct tr_multi tr_multi;
int tr_multi_create(tr_multi** out, const tr_scene_desc* desc, const int32_t* devices, uint32_t n_devices);     /* desc->device is ignored */
int tr_multi_create_from_dir(tr_multi** out, const char* asset_dir, const char* pipeline_name, uint32_t width, uint32_t height,
                             const int32_t* devices, uint32_t n_devices, uint32_t ring);
void tr_multi_destroy(tr_multi* m);
uint32_t tr_multi_device_count(const tr_multi* m);
tr_scene* tr_multi_scene(tr_multi* m, uint32_t i);
int tr_multi_set_camera(tr_multi* m, const float look_from[3], const float look_at[3], const float up[3]);
int tr_multi_render_views(tr_multi* m, const tr_view* views, uint32_t n_views, uint64_t* checksums, uint8_t* frames_rgb8_topdown);
int tr_multi_render_ao(tr_multi* m, const float* light_dirs_xyz, uint32_t n_lights, uint32_t* accum_host);
/* "nvls": the accumulators live in memory bound to one NVSwitch multicast object and are summed INSIDE the switch
 * (multimem.ld_reduce / multimem.st, k_peer.cuh k_ao_reduce_nvls) — only with TINYR_NVLS=1 in the environment, distinct
 * devices and multicast support; otherwise "peer: <reason>": the peer-memory kernel, which is the default because it is
 * the faster of the two at 2, 4 and 8 B200s (profiles/r2_probes.md).  Results are identical. */
const char* tr_multi_exchange_kind(tr_multi* m);
/* Benchmark hook: reps back-to-back exchanges on all devices at once; the slowest device's CUDA-event time / reps, in us. */
int tr_multi_time_exchange(tr_multi* m, uint32_t reps, double* us_per_exchange);
uint64_t tr_checksum_rgb8(const uint8_t* rgb8, size_t n_bytes);                  /* host twin of the device checksum */
/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  The device checksum (K8, the one
 * tr_scene_render_views returns per view) of a caller's frame `rgb8` (host or any UVA pointer) on `device`, with ALL device
 * scratch — the accumulator and the device copy of the frame — in a file-scope static __device__ array
 * (csrc/tinyr_static.cu) that every GPU using this entry point holds until the process exits.  Allocates nothing.
 * n_bytes > TR_STATIC_CHECKSUM_MAX_BYTES -> TR_ERR_INVALID_ARGUMENT.  Synchronous (legacy default stream).  Every call
 * on a device shares the array: calls that overlap on the GPU would race; inside one process a mutex prevents that. */
#define TR_STATIC_CHECKSUM_MAX_BYTES (8u << 20)   /* 8 MiB: a 1920x1080 RGB8 frame is 6,220,800 bytes */
int tr_checksum_rgb8_static(int device, const uint8_t* rgb8, size_t n_bytes, uint64_t* checksum);
/* Deterministic light set for tr_scene_render_ao (host only): SplitMix64(seed) -> uniform sphere directions, computed
 * in double and rounded to f32, rejecting directions on which the reference's occlusion pipeline panics or reads
 * its shadow buffer on the wrong row: light parallel to `up` (shader.rs:194), light == -z (shader.rs:921), any model
 * vertex outside |ndc| <= ndc_margin in the light's view (shader.rs:774-778, 909-912, 932-935; use 0.97).
 * n_drawn (optional) receives the number of candidates drawn. */
int tr_ao_generate_lights(const float* positions_xyz, uint32_t n_positions, const float look_at[3], const float up[3],
                          uint64_t seed, uint32_t n_lights, float ndc_margin, float* out_dirs_xyz, uint64_t* n_drawn);

/* ---- device plumbing -------------------------------------------------------------------------------- */
int tr_device_count(int* n);
int tr_scene_set_stream(tr_scene* s, void* cuda_stream);                          /* run on the caller's cudaStream_t (NULL = scene's own) */
int tr_scene_device_frame(tr_scene* s, void** device_ptr_rgb8_topdown);
/* Test hooks: bit 0 makes the occlusion samples always take the range-tested division (k_shade.cuh occlusion_gray picks
 * between two forms of the same arithmetic per pixel; results must not depend on the choice); bit 1 makes single-frame
 * renders launch the tile kernel in its batch CTA shape (128 threads), so that shape can be timed on a lone frame; bit 2
 * makes every lone frame run all its passes in the static launch order (no re-use of unchanged passes, no cost-sorted
 * order), so both forms can be timed and compared in one process. */
int tr_scene_set_debug_flags(tr_scene* s, uint32_t flags);           /* zero-copy access for tests / interop */
int tr_scene_timer_begin(tr_scene* s);                                            /* CUDA event on the scene stream */
int tr_scene_timer_end(tr_scene* s, float* elapsed_ms);                           /* records + synchronises */
/* Per-kernel device times (CUDA events around every launch while enabled).  names: K1 vertex_setup, K2 bin,
 * K3 raster_visibility, K4 shadow_depth, K5 shade, K6 ao_accumulate, K7 clear/pack, K8 checksum. */
int tr_scene_set_profiling(tr_scene* s, int enabled);
int tr_scene_kernel_times(tr_scene* s, float total_ms[8], uint64_t launches[8]);
/* Device self-test of the rasteriser's 3-instruction exact division (k_tile.cuh exact_quotient) against IEEE f32
 * division over n_pairs pseudo-random integer pairs 0 <= a <= b <= 2^21; *mismatches must come back 0. */
int tr_selftest_exact_quotient(int device, uint64_t n_pairs, uint64_t seed, uint64_t* mismatches);
/* Device self-test of the fragment stage's shared-divisor quotients (k_tile.cuh shared_divisor/quotient: nvcc's own
 * division sequence with the divisor-only part hoisted) against `a / b`; *mismatches must come back 0. */
int tr_selftest_shared_divisor(int device, uint64_t n_pairs, uint64_t seed, uint64_t* mismatches);
/* Device self-test of the shadow-lookup rounding (k_shade.cuh round_u32: `f32::round() as u32`, shader.rs:774-775)
 * against its literal form on all 2^32 f32 bit patterns; *mismatches must come back 0. */
int tr_selftest_round_u32(int device, uint64_t* mismatches);
/* Device self-test of the normalisations' square root (k_tile.cuh sqrt_fast_path: nvcc's own IEEE sqrt fast path,
 * spelled out to share its range test with the division that follows) against sqrtf() on every f32 bit pattern it
 * accepts; *mismatches must come back 0, *accepted is the number of patterns tested (about 1.9 * 2^30). */
int tr_selftest_sqrt(int device, uint64_t* mismatches, uint64_t* accepted);
/* Micro-benchmark: rate of global 64-bit atomicMin on a width x height key plane, distinct
 * addresses, warp = 8x4 pixel block, `rounds` updating atomics per key; *atomics_per_second is the best of `repeats`
 * timed launches (CUDA events).  Not on the render path: k_raster resolves visibility in registers. */
int tr_bench_atomic_min(int device, uint32_t width, uint32_t height, uint32_t rounds, uint32_t repeats, double* atomics_per_second);
/* Host frames.  tr_host_alloc returns pinned, device-mapped, ZEROED host memory.  Frames delivered into it
 * (tr_scene_render_views' `frames`, tr_scene_get_frame_buffer's dst — the clone of scene.rs:92-97) travel SPARSELY
 * (k_host.cuh): the device stores, straight over PCIe into the destination frame, only the 64-pixel tile rows that hold
 * a non-zero byte now and zeros into those that held one after the previous delivery to the same address; a 4K diablo
 * frame is ~5 MB instead of 24.9 MB on the link.  The caller always finds the exact top-down RGB8 frame.  Contract: between
 * deliveries nobody but the library writes to the memory; after writing to it yourself call tr_host_invalidate on the
 * range (the next delivery then rewrites whole frames).  Any other destination pointer (malloc, numpy, foreign pinned
 * memory) takes a plain full-frame copy.  TINYR_NO_SPARSE_D2H=1 in the environment forces the full copy everywhere. */
void* tr_host_alloc(size_t bytes);
void tr_host_free(void* p);
int tr_host_invalidate(void* p, size_t bytes);
/* out[0] bytes stored into host frames by the sparse delivery, out[1] tile rows stored, out[2] frames delivered sparsely,
 * out[3] frames delivered by a full copy, out[4] bytes of those copies; reset != 0 zeroes the counters. */
int tr_scene_transfer_stats(tr_scene* s, uint64_t out[5], int reset);

/* ---- asset loaders kept from the reference (app.rs:87-131) ------------------------------------------- */
typedef struct tr_mesh {
    float* positions;   uint32_t n_positions;
    float* uvs;         uint32_t n_uvs;
    float* normals;     uint32_t n_normals;
    uint32_t* tri_idx;  uint32_t n_triangles;
} tr_mesh;
typedef struct tr_pixels {
    uint8_t* rgb8_topdown;
    uint32_t width, height;
} tr_pixels;
int tr_load_obj(const char* path, tr_mesh* out);        /* obj::raw::parse_obj, app.rs:94 (PTN polygons only, scene.rs:216-219) */
void tr_mesh_free(tr_mesh* m);
int tr_load_tga(const char* path, tr_pixels* out);      /* image::open(path)?.into_rgb8(), app.rs:99 */
void tr_pixels_free(tr_pixels* p);
int tr_write_tga(const char* path, const uint8_t* rgb8_topdown, uint32_t width, uint32_t height, int rle);  /* headless output */
/* app::run's load sequence (app.rs:87-146): <dir>/model.obj, texture.tga, normal_map.tga, normal_map_tangent.tga, specular_map.tga */
int tr_scene_create_from_dir(tr_scene** out, const char* asset_dir, const char* pipeline_name, uint32_t width,
                             uint32_t height, int32_t device, uint32_t ring);

const char* tr_last_error(void);
const char* tr_version(void);

#ifdef __cplusplus
}
#endif
#endif /* TINYR_H */
