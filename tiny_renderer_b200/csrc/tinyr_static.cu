// tinyr_static.cu — tr_checksum_rgb8_static: the K8 frame checksum for callers that must not allocate.
//
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops, harnesses that forbid it).  The existing device checksum
// (k_checksum, k_misc.cuh, reached through tr_scene_render_views) accumulates into cudaMalloc'ed memory of a scene.
// This variant keeps ALL of its device scratch — the 64-bit accumulator and the device copy of the caller's frame — in one
// file-scope `static __device__` array of kStaticScratchBytes; every GPU that loads this module holds that array until
// the process exits.  The entry point calls no cudaMalloc / cudaFree / cudaHostAlloc and no allocating library routine.
//
// The array is shared by every call on a device and keeps the previous call's contents: the accumulator is zeroed by the
// call itself before the kernel runs, the frame copy is overwritten up to n_bytes and never read beyond.  Calls on one
// device that overlapped on the GPU would race on it; inside this process they cannot (the call is synchronous and holds
// a mutex), callers that share a GPU between processes each have their own copy of the array.
#include <cstdint>
#include <cstdio>
#include <mutex>
#include <cuda_runtime.h>
#include "../../include/tinyr.h"
#include "tr_error.h"

namespace tr {
namespace {

constexpr size_t kStaticFrameOffset = 256;                                  // [0, 8): accumulator; frame copy from byte 256
constexpr size_t kStaticScratchBytes = kStaticFrameOffset + (size_t)TR_STATIC_CHECKSUM_MAX_BYTES;
constexpr int kStaticMaxDevices = 64;

// In host code this name is a host-side stand-in for the symbol, NOT a device address: the address a kernel can use
// comes from cudaGetSymbolAddress, per device (static_scratch() below).
static __device__ __align__(256) unsigned char g_static_scratch[kStaticScratchBytes];

std::mutex g_static_mu;
unsigned char* g_static_ptr[kStaticMaxDevices];                            // device address of g_static_scratch, per device

// The device address of the array on `device` (which must be current), resolved once per device.
int static_scratch(int device, unsigned char** out) {
    if (!g_static_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            cudaGetLastError();
            return fail(TR_ERR_CUDA, "tr_checksum_rgb8_static: cudaGetSymbolAddress on device %d: %s", device, cudaGetErrorString(e));
        }
        g_static_ptr[device] = static_cast<unsigned char*>(p);
    }
    *out = g_static_ptr[device];
    return TR_OK;
}

// K8 as in k_misc.cuh (sum_i word_i * (2 i + 1) mod 2^64 over little-endian u32 words), with the words and the accumulator
// both inside the static array: acc = scratch, words = scratch + kStaticFrameOffset.
__global__ void __launch_bounds__(256) k_checksum_static(unsigned char* scratch, size_t n_words) {
    const uint32_t* words = reinterpret_cast<const uint32_t*>(scratch + kStaticFrameOffset);
    unsigned long long* out = reinterpret_cast<unsigned long long*>(scratch);
    const size_t i0 = (size_t)blockIdx.x * blockDim.x + threadIdx.x, stride = (size_t)gridDim.x * blockDim.x;
    unsigned long long acc = 0;
    for (size_t i = i0; i < n_words; i += stride) acc += (unsigned long long)__ldg(&words[i]) * (2ull * i + 1ull);
    for (int o = 16; o > 0; o >>= 1) acc += __shfl_down_sync(0xffffffffu, acc, o);
    __shared__ unsigned long long s[8];
    if ((threadIdx.x & 31) == 0) s[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        unsigned long long t = 0;
        for (int w = 0; w < 8; ++w) t += s[w];
        if (t) atomicAdd(out, t);
    }
}

#define CUS(call)                                                                                              \
    do {                                                                                                       \
        const cudaError_t e_ = (call);                                                                         \
        if (e_ != cudaSuccess) {                                                                               \
            cudaGetLastError();                                                                                \
            return fail(TR_ERR_CUDA, "tr_checksum_rgb8_static: %s: %s", #call, cudaGetErrorString(e_));        \
        }                                                                                                      \
    } while (0)

}  // namespace
}  // namespace tr

extern "C" int tr_checksum_rgb8_static(int device, const uint8_t* rgb8, size_t n_bytes, uint64_t* checksum) {
    using namespace tr;
    if (!checksum || (!rgb8 && n_bytes)) return fail(TR_ERR_INVALID_ARGUMENT, "tr_checksum_rgb8_static: null argument");
    if (n_bytes > (size_t)TR_STATIC_CHECKSUM_MAX_BYTES)
        return fail(TR_ERR_INVALID_ARGUMENT, "tr_checksum_rgb8_static: %zu bytes exceed the static scratch of this entry point (at most %zu bytes per call)",
                    n_bytes, (size_t)TR_STATIC_CHECKSUM_MAX_BYTES);
    if (device < 0 || device >= kStaticMaxDevices) return fail(TR_ERR_INVALID_ARGUMENT, "tr_checksum_rgb8_static: device %d out of range", device);
    clear_error();
    std::lock_guard<std::mutex> lock(g_static_mu);
    CUS(cudaSetDevice(device));
    unsigned char* scratch = nullptr;
    const int rc = static_scratch(device, &scratch);
    if (rc != TR_OK) return rc;
    const size_t n_words = n_bytes / 4;                    // trailing bytes (n % 4) are not covered, like k_checksum
    CUS(cudaMemsetAsync(scratch, 0, 8, 0));                // the accumulator holds the previous call's sum
    if (n_words) {
        CUS(cudaMemcpyAsync(scratch + kStaticFrameOffset, rgb8, n_words * 4, cudaMemcpyDefault, 0));
        k_checksum_static<<<148 * 4, 256, 0, 0>>>(scratch, n_words);
        CUS(cudaGetLastError());
    }
    unsigned long long h = 0;
    CUS(cudaMemcpyAsync(&h, scratch, 8, cudaMemcpyDeviceToHost, 0));
    CUS(cudaStreamSynchronize(0));
    *checksum = h;
    return TR_OK;
}
