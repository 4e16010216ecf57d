// k_misc.cuh — K7 clear / views / flips and K8 frame checksum.
#pragma once
#include "k_tile.cuh"

namespace tr {

// Scene::clear (/root/reference/src/scene.rs:128-137) materialised, or Buffer::new's zero state
// (/root/reference/src/scene/shader.rs:46-47) when z_init = 0.  Grid-stride, 16-byte stores where possible.
__global__ void __launch_bounds__(256) k_fill_state(unsigned long long* keys, float* shadow, uint32_t* rgb_words, size_t n_px,
                                                    size_t n_rgb_words, float z_init, int do_keys, int do_shadow, int do_rgb) {
    size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    const unsigned long long k = make_key(z_init, kNoTri);
    if (do_keys) for (size_t i = i0; i < n_px; i += stride) keys[i] = k;
    if (do_shadow) for (size_t i = i0; i < n_px; i += stride) shadow[i] = z_init;
    if (do_rgb) for (size_t i = i0; i < n_rgb_words; i += stride) rgb_words[i] = 0u;
}

// Writes the cleared key into every tile still in the tile-level cleared state (k_raster.cuh) and marks it as
// materialised; getters that read key memory call this first.  One CTA per raster tile.
__global__ void __launch_bounds__(256) k_materialise_tiles(unsigned long long* keys, uint32_t* tile_state, uint32_t width, uint32_t height) {
    const uint32_t tile = blockIdx.y * gridDim.x + blockIdx.x;
    if (tile_state[tile] == 0u) return;
    const unsigned long long k = make_key(-3.402823466e38f, kNoTri);
    for (int i = threadIdx.x; i < kTileW * kTileH; i += 256) {
        uint32_t px = blockIdx.x * kTileW + i % kTileW, py = blockIdx.y * kTileH + i / kTileW;
        if (px < width && py < height) keys[(size_t)py * width + px] = k;
    }
    __syncthreads();
    if (threadIdx.x == 0) tile_state[tile] = 0u;
}

// The epoch field of a key (renders since the last clear) has 8 bits.  The reference allows any number of renders without
// a clear (scene.rs:151 never clears), so when the field would overflow every stored winner is moved to epoch 0 — all
// that matters about an old winner is that it orders before whatever the NEXT render draws at the same depth — and the
// scene continues at epoch 1.  One CTA per raster tile; tiles in the tile-level cleared state hold no keys.
__global__ void __launch_bounds__(256) k_collapse_epochs(unsigned long long* keys, const uint32_t* tile_state, uint32_t width, uint32_t height) {
    const uint32_t tile = blockIdx.y * gridDim.x + blockIdx.x;
    if (tile_state[tile] != 0u) return;
    for (int i = threadIdx.x; i < kTileW * kTileH; i += 256) {
        uint32_t px = blockIdx.x * kTileW + i % kTileW, py = blockIdx.y * kTileH + i / kTileW;
        if (px < width && py < height) {
            unsigned long long k = keys[(size_t)py * width + px];
            if (key_lo(k) != kNoTri) keys[(size_t)py * width + px] = (k & 0xFFFFFFFF00000000ull) | (key_lo(k) & kTriMask);
        }
    }
}

// Raw bottom-up views of the state the reference keeps in Buffer.z_buffer / visibility (derived from keys).
__global__ void __launch_bounds__(256) k_unpack_keys(const unsigned long long* keys, float* depth, uint32_t* tri, size_t n_px) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_px) return;
    unsigned long long k = keys[i];
    if (depth) depth[i] = key_depth(k);
    if (tri) tri[i] = key_lo(k) == kNoTri ? kNoTri : (key_lo(k) & kTriMask);
}

// get_z_buffer / get_shadow_buffer (scene.rs:101-125): saturating `f32 as u8` gray, vertically flipped.
__global__ void __launch_bounds__(256) k_gray_view(const unsigned long long* keys, const float* plane, uint8_t* dst, uint32_t w, uint32_t h) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)w * h) return;
    uint32_t x = (uint32_t)(i % w), y = (uint32_t)(i / w);
    float v = keys ? key_depth(keys[i]) : plane[i];
    uint8_t g = (uint8_t)cast_u8(v);
    uint8_t* p = dst + ((size_t)(h - 1 - y) * w + x) * 3;
    p[0] = g; p[1] = g; p[2] = g;
}

// Vertical flip of an RGB8 image (top-down storage -> the reference's bottom-up frame_buffer order).
__global__ void __launch_bounds__(256) k_flip_rows(const uint8_t* src, uint8_t* dst, uint32_t row_bytes, uint32_t h) {
    size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= (size_t)row_bytes * h) return;
    uint32_t y = (uint32_t)(i / row_bytes), b = (uint32_t)(i % row_bytes);
    dst[(size_t)(h - 1 - y) * row_bytes + b] = src[i];
}

// Self-test of exact_quotient() (k_tile.cuh) against the IEEE division it replaces: random and adversarial integer
// pairs 0 <= a <= b <= 2^21.  Counts mismatching bit patterns.
__global__ void __launch_bounds__(256) k_selftest_quotient(unsigned long long n, unsigned long long seed, unsigned long long* mismatches) {
    unsigned long long i0 = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long bad = 0;
    for (unsigned long long i = i0; i < n; i += stride) {
        unsigned long long z = seed + (i + 1ull) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        uint32_t b = (uint32_t)(z & 0x1FFFFFu) + 1u;                       // 1 .. 2^21
        if ((i & 7ull) == 0ull) b = 0x200000u - (uint32_t)((z >> 40) & 0xFFu);   // near the top of the range
        if ((i & 7ull) == 1ull) b = (1u << ((z >> 50) % 22u)) + (uint32_t)((z >> 56) & 3u);   // near powers of two
        if (b > 0x200000u) b = 0x200000u;
        uint32_t a = (uint32_t)((z >> 21) % (unsigned long long)(b + 1u));      // 0 .. b
        if ((i & 15ull) == 2ull) a = b - (uint32_t)((z >> 60) & 3u) > b ? b : b - (uint32_t)((z >> 60) & 3u);
        float fa = (float)a, fb = (float)b;
        float q_ref = fa / fb;
        float q = exact_quotient(fa, fb, 1.0f / fb);
        bad += __float_as_uint(q) != __float_as_uint(q_ref) ? 1ull : 0ull;
    }
    if (bad) atomicAdd(mismatches, bad);
}

// Self-test of sqrt_fast_path() (k_tile.cuh) against sqrtf() on every f32 bit pattern it accepts.
__global__ void __launch_bounds__(256) k_selftest_sqrt(unsigned long long* mismatches, unsigned long long* accepted) {
    const unsigned long long i0 = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long bad = 0, n_ok = 0;
    for (unsigned long long i = i0; i < (1ull << 32); i += stride) {
        const float x = __uint_as_float((uint32_t)i);
        bool ok;
        const float s = sqrt_fast_path(x, ok);
        if (!ok) continue;
        ++n_ok;
        bad += __float_as_uint(s) == __float_as_uint(sqrtf(x)) ? 0ull : 1ull;
    }
    if (bad) atomicAdd(mismatches, bad);
    atomicAdd(accepted, n_ok);
}

// Self-test of round_u32() (k_shade.cuh) against the literal `f32::round() as u32` on every f32 bit pattern.
__global__ void __launch_bounds__(256) k_selftest_round_u32(unsigned long long* mismatches) {
    const unsigned long long i0 = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long bad = 0;
    for (unsigned long long i = i0; i < (1ull << 32); i += stride) {
        const float x = __uint_as_float((uint32_t)i);
        bad += round_u32(x) == round_u32_literal(x) ? 0ull : 1ull;
    }
    if (bad) atomicAdd(mismatches, bad);
}

// Micro-benchmark for SURVEY.md §8(d): the rate of 64-bit atomicMin on distinct addresses in the pattern a rasteriser
// that resolves visibility in global memory would produce — each warp owns an 8x4 pixel block of a row-major key
// plane (8 consecutive keys on 4 rows) and every lane hits its own key `rounds` times with a decreasing value, so every
// atomic updates.  k_raster resolves winners in registers instead; this number is what that design is measured against.
__global__ void __launch_bounds__(256) k_bench_atomic_min(unsigned long long* keys, uint32_t width, uint32_t height, uint32_t rounds) {
    const uint32_t blocks_x = width / 8, blocks_y = height / 4;
    const uint32_t n_blocks = blocks_x * blocks_y;
    const uint32_t lane = threadIdx.x & 31u, warps = gridDim.x * (blockDim.x >> 5);
    for (uint32_t b = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5); b < n_blocks; b += warps) {
        const uint32_t px = (b % blocks_x) * 8 + (lane & 7u), py = (b / blocks_x) * 4 + (lane >> 3);
        unsigned long long* at = keys + (size_t)py * width + px;
        for (uint32_t r = 0; r < rounds; ++r) atomicMin(at, ((unsigned long long)(0x7fffffffu - r) << 32) | b);
    }
}

// Self-test of shared_divisor()/quotient() (k_tile.cuh) against `a / b`: divisors of moderate magnitude, numerators
// from vanishing to 2^12 times the divisor (what normalisations and homogeneous divides feed it).  Counts results whose
// bits differ from the division's, ignoring the sign of a zero result.
__global__ void __launch_bounds__(256) k_selftest_shared_divisor(unsigned long long n, unsigned long long seed, unsigned long long* mismatches) {
    unsigned long long i0 = (unsigned long long)blockIdx.x * blockDim.x + threadIdx.x, stride = (unsigned long long)gridDim.x * blockDim.x;
    unsigned long long bad = 0;
    for (unsigned long long i = i0; i < n; i += stride) {
        unsigned long long z = seed + (i + 1ull) * 0x9E3779B97F4A7C15ull;
        z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
        z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
        z ^= z >> 31;
        // b: random mantissa, exponent in [-24, 24], random sign
        uint32_t eb = 127u - 24u + (uint32_t)((z >> 23) % 49ull);
        float b = __uint_as_float((uint32_t)(z & 0x7FFFFFu) | (eb << 23) | ((uint32_t)(z >> 63) << 31));
        // a: random mantissa, exponent = exponent(b) + [-40, 12]
        unsigned long long w = z * 0xD6E8FEB86659FD93ull + 0x2545F4914F6CDD1Dull;
        w ^= w >> 32;
        uint32_t ea = eb - 40u + (uint32_t)((w >> 23) % 53ull);
        float a = __uint_as_float((uint32_t)(w & 0x7FFFFFu) | (ea << 23) | ((uint32_t)(w >> 63) << 31));
        if ((i & 31ull) == 0ull) a = b;                                   // quotient exactly 1
        if ((i & 31ull) == 1ull) a = 0.0f;
        float q_ref = a / b;
        float q = in_fast_division_range(b) ? quotient(a, shared_divisor(b)) : a / b;
        bool same = __float_as_uint(q) == __float_as_uint(q_ref) || (q == 0.0f && q_ref == 0.0f);
        bad += same ? 0ull : 1ull;
    }
    if (bad) atomicAdd(mismatches, bad);
}

// K8: position-weighted 64-bit checksum of a byte image viewed as little-endian u32 words:
// sum_i word_i * (2 i + 1) mod 2^64 (tr_checksum_rgb8 is the host twin).  n_words = bytes / 4 (bytes % 4 == 0).
__global__ void __launch_bounds__(256) k_checksum(const uint32_t* words, size_t n_words, unsigned long long* out) {
    size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    unsigned long long acc = 0;
    for (size_t i = i0; i < n_words; i += stride) acc += (unsigned long long)__ldg(&words[i]) * (2ull * i + 1ull);
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    __shared__ unsigned long long s[8];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < 8; ++w) t += s[w];
        atomicAdd(out, t);
    }
}

}  // namespace tr
