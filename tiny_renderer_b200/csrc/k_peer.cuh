// K9: sum of the ranks' ambient-occlusion accumulators over NVLink peer memory.
//
// The many-light AO batch (tr_scene_render_ao) leaves one u32 sum per pixel on every rank.  The reference has no
// multi-GPU path; what this replaces is "all-reduce(sum) of W*H u32" after the last light pass.  One launch per rank,
// on the stream the light passes ran on, does the whole exchange:
//
//   barrier A   every rank's light passes are complete (a CTA of this kernel only runs once they are)
//   reduce      rank r owns elements [r n / R, (r + 1) n / R): it loads that slice from all R accumulators (its own and
//               R - 1 peer-mapped ones: NVLink reads), adds, and stores the sum into all R accumulators (NVLink writes)
//   barrier B   every rank's stores have landed, so when the kernel ends on a rank its accumulator holds the total
//
// u32 addition is associative and commutative, so the total is bit-identical for any rank count and any order.
// Per rank 2 (R - 1) / R of the buffer crosses NVLink (4K: 29 MB in, 29 MB out at R = 8); nothing goes through the host.
//
// The barriers are flags in the peers' allocations (behind the accumulator): CTA b of rank r writes `value` into word
// [b][r] of every rank's flag area and waits until words [b][0..R) of its own area have reached it.  Values only grow
// (two per call), so the flags are never reset.  Every rank launches the same grid, CTAs start in index order, and a CTA
// only waits for the CTA of its own index on the other ranks, so the wait cannot cycle; it is still bounded by a
// timeout (status word in mapped host memory) so that a rank that never arrives cannot hang the GPU.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

namespace tr {

constexpr uint32_t kMaxPeers = 16;
constexpr uint32_t kPeerBlocks = 148;              // one CTA per SM
constexpr uint32_t kPeerThreads = 512;
constexpr uint32_t kPeerFlagWords = kPeerBlocks * kMaxPeers;

struct PeerArgs {
    uint32_t* buf[kMaxPeers];        // accumulator of every rank (index = rank; own entry = local pointer)
    uint32_t world, rank;
    uint32_t n_vec;                  // accumulator length in uint4 (the allocation is padded with zeros)
    uint32_t flag_word;              // u32 offset of the flag area in every allocation
    uint32_t value;                  // barrier A waits for `value`, barrier B for `value + 1`
    unsigned long long timeout_ns;
    int* status;                     // mapped host memory; 1 = a barrier timed out (the sums are then invalid)
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    return t;
}
// Peer memory is read exactly once per call and must not be served from a line cached by an earlier call.
__device__ __forceinline__ uint4 ld_peer(const uint4* p) {
    uint4 v;
    asm volatile("ld.relaxed.sys.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
    return v;
}
__device__ __forceinline__ void st_peer(uint4* p, uint4 v) {
    asm volatile("st.relaxed.sys.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}

// All threads of the CTA call it.  Returns false (for every thread) when the wait timed out.
__device__ __forceinline__ bool peer_barrier(const PeerArgs& a, uint32_t value) {
    __threadfence_system();          // this thread's stores, before the CTA's flag is raised
    __syncthreads();
    bool ok = true;
    if (threadIdx.x < a.world) {
        const uint32_t peer = threadIdx.x;
        st_release_sys(a.buf[peer] + a.flag_word + blockIdx.x * kMaxPeers + a.rank, value);
        const uint32_t* mine = a.buf[a.rank] + a.flag_word + blockIdx.x * kMaxPeers + peer;
        const unsigned long long t0 = global_timer_ns();
        while ((int32_t)(ld_acquire_sys(mine) - value) < 0) {
            if (global_timer_ns() - t0 > a.timeout_ns) {
                *(volatile int*)a.status = 1;
                ok = false;
                break;
            }
        }
    }
    return __syncthreads_and(ok) != 0;
}

// WORLD > 0: rank count known at compile time (all loads of an element in flight together); WORLD == 0: any count.
template <int WORLD>
__global__ void __launch_bounds__(kPeerThreads, 1) k_ao_reduce_peers(const PeerArgs a) {
    const uint32_t world = WORLD > 0 ? (uint32_t)WORLD : a.world;
    if (!peer_barrier(a, a.value)) return;
    const uint32_t lo = (uint32_t)((unsigned long long)a.n_vec * a.rank / world);
    const uint32_t hi = (uint32_t)((unsigned long long)a.n_vec * (a.rank + 1) / world);
    for (uint32_t i = lo + blockIdx.x * kPeerThreads + threadIdx.x; i < hi; i += kPeerBlocks * kPeerThreads) {
        uint4 sum = make_uint4(0u, 0u, 0u, 0u);
        if (WORLD > 0) {
            uint4 part[WORLD > 0 ? WORLD : 1];
#pragma unroll
            for (int k = 0; k < WORLD; ++k) {
                // start at the own copy and walk the ranks upwards, so that at any moment the R readers pull from R different peers
                uint32_t p = a.rank + k;
                p = p >= world ? p - world : p;
                part[k] = ld_peer(reinterpret_cast<const uint4*>(a.buf[p]) + i);
            }
#pragma unroll
            for (int k = 0; k < WORLD; ++k) { sum.x += part[k].x; sum.y += part[k].y; sum.z += part[k].z; sum.w += part[k].w; }
        } else {
            for (uint32_t k = 0; k < world; ++k) {
                uint32_t p = a.rank + k;
                p = p >= world ? p - world : p;
                const uint4 v = ld_peer(reinterpret_cast<const uint4*>(a.buf[p]) + i);
                sum.x += v.x; sum.y += v.y; sum.z += v.z; sum.w += v.w;
            }
        }
#pragma unroll
        for (int k = 0; k < (WORLD > 0 ? WORLD : 1); ++k) {
            if (WORLD > 0) {
                uint32_t p = a.rank + k;
                p = p >= world ? p - world : p;
                st_peer(reinterpret_cast<uint4*>(a.buf[p]) + i, sum);
            }
        }
        if (WORLD == 0) {
            for (uint32_t k = 0; k < world; ++k) {
                uint32_t p = a.rank + k;
                p = p >= world ? p - world : p;
                st_peer(reinterpret_cast<uint4*>(a.buf[p]) + i, sum);
            }
        }
    }
    peer_barrier(a, a.value + 1u);
}

// K9n: the same sum through the NVSwitch (NVLS).  When the accumulators of all ranks are bound to one multicast object
// (tinyr_multi.inc: one process, cuMulticast* + cuMemCreate instead of cudaMalloc), `multimem.ld_reduce` on the multicast
// address returns the sum of the R copies formed INSIDE the switch, and `multimem.st` writes a value to all R copies:
// rank r reads its slice once (one element in, instead of R - 1 peer loads) and writes it once (instead of R - 1 peer
// stores), so per rank 1 / R of the buffer crosses its NVLink in each direction, whatever R is.  The integer multimem
// operations are scalar; the u32 sums travel as packed u64 pairs — exact, because a 32-bit sum never carries (255 per
// light and pixel: 16 M lights would be needed).  Same barriers as k_ao_reduce_peers (flags through the unicast mappings).
__device__ __forceinline__ unsigned long long mm_ld_reduce_add_u64(const unsigned long long* mc) {
    unsigned long long v;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(v) : "l"(mc) : "memory");
    return v;
}
__device__ __forceinline__ void mm_st_u64(unsigned long long* mc, unsigned long long v) {
    asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" :: "l"(mc), "l"(v) : "memory");
}
constexpr int kNvlsUnroll = 8;      // in-switch reductions in flight per thread
__global__ void __launch_bounds__(kPeerThreads, 1) k_ao_reduce_nvls(const PeerArgs a, unsigned long long* mc) {
    if (!peer_barrier(a, a.value)) return;
    const unsigned long long n64 = 2ull * a.n_vec;
    const unsigned long long lo = n64 * a.rank / a.world, hi = n64 * (a.rank + 1) / a.world;
    const unsigned long long stride = (unsigned long long)kPeerBlocks * kPeerThreads;
    for (unsigned long long i = lo + blockIdx.x * kPeerThreads + threadIdx.x; i < hi; i += stride * kNvlsUnroll) {
        unsigned long long v[kNvlsUnroll];
#pragma unroll
        for (int k = 0; k < kNvlsUnroll; ++k) if (i + k * stride < hi) v[k] = mm_ld_reduce_add_u64(mc + i + k * stride);
#pragma unroll
        for (int k = 0; k < kNvlsUnroll; ++k) if (i + k * stride < hi) mm_st_u64(mc + i + k * stride, v[k]);
    }
    peer_barrier(a, a.value + 1u);
}

}  // namespace tr
