// Probe: NVLS (NVSwitch multicast) from ONE process over G devices — multicast object, per-device physical memory bound to it,
// unicast + multicast mappings, then multimem.ld_reduce.add.u64 (in-switch sum of the G copies) and multimem.st (broadcast)
// from a kernel on device 0.  u32 sums travel as packed u64 pairs: exact as long as no 32-bit sum carries (the AO sums stay
// below 2^19).  Prints what works and how long a 33 MB reduction takes.
//   nvcc -gencode arch=compute_100a,code=sm_100a -o mc_probe mc_probe.cu && ./mc_probe [G]
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <vector>

#define DRV(name) static decltype(&name) p_##name = nullptr
DRV(cuDeviceGetAttribute); DRV(cuMulticastCreate); DRV(cuMulticastAddDevice); DRV(cuMulticastBindMem); DRV(cuMulticastGetGranularity);
DRV(cuMemCreate); DRV(cuMemAddressReserve); DRV(cuMemMap); DRV(cuMemSetAccess); DRV(cuMemGetAllocationGranularity); DRV(cuDeviceGet);
template <class F> static bool load(F& fn, const char* name) {
    void* p = nullptr; cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint(name, &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess || !p) { printf("no driver entry point %s\n", name); return false; }
    fn = reinterpret_cast<F>(p); return true;
}
#define CK(call) do { CUresult r_ = (call); if (r_ != CUDA_SUCCESS) { printf("%s -> CUresult %d (line %d)\n", #call, (int)r_, __LINE__); return 1; } } while (0)

// rank r of R reduces slice [r n / R, (r + 1) n / R) of the u64-packed accumulators in the switch and broadcasts the sums
__global__ void k_mc_reduce(unsigned long long* mc, size_t lo, size_t hi) {
    for (size_t i = lo + blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < hi; i += (size_t)gridDim.x * blockDim.x) {
        unsigned long long v;
        asm volatile("multimem.ld_reduce.relaxed.sys.global.add.u64 %0, [%1];" : "=l"(v) : "l"(mc + i) : "memory");
        asm volatile("multimem.st.relaxed.sys.global.u64 [%0], %1;" :: "l"(mc + i), "l"(v) : "memory");
    }
}
__global__ void k_fill(uint32_t* p, size_t n, uint32_t base) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) p[i] = base + (uint32_t)(i % 1000);
}

int main(int argc, char** argv) {
    int G = argc > 1 ? atoi(argv[1]) : 2, n_dev = 0;
    cudaGetDeviceCount(&n_dev);
    if (n_dev < G) { printf("need %d devices, have %d\n", G, n_dev); return 0; }
    cudaFree(0);
    if (!load(p_cuDeviceGetAttribute, "cuDeviceGetAttribute") || !load(p_cuMulticastCreate, "cuMulticastCreate") || !load(p_cuMulticastAddDevice, "cuMulticastAddDevice") ||
        !load(p_cuMulticastBindMem, "cuMulticastBindMem") || !load(p_cuMulticastGetGranularity, "cuMulticastGetGranularity") || !load(p_cuMemCreate, "cuMemCreate") ||
        !load(p_cuMemAddressReserve, "cuMemAddressReserve") || !load(p_cuMemMap, "cuMemMap") || !load(p_cuMemSetAccess, "cuMemSetAccess") ||
        !load(p_cuMemGetAllocationGranularity, "cuMemGetAllocationGranularity") || !load(p_cuDeviceGet, "cuDeviceGet")) return 1;
    for (int d = 0; d < G; ++d) {
        int ok = 0; CUdevice dev; CK(p_cuDeviceGet(&dev, d));
        CK(p_cuDeviceGetAttribute(&ok, CU_DEVICE_ATTRIBUTE_MULTICAST_SUPPORTED, dev));
        printf("device %d multicast supported: %d\n", d, ok);
        if (!ok) return 0;
        cudaSetDevice(d); cudaFree(0);
        for (int e = 0; e < G; ++e) if (e != d) cudaDeviceEnablePeerAccess(e, 0);
        cudaGetLastError();
    }
    const size_t n_px = 3840ull * 2160ull, want = n_px * 4;
    CUmulticastObjectProp mp{}; mp.numDevices = G; mp.handleTypes = 0; mp.flags = 0;
    size_t gran = 0;
    mp.size = want;
    CK(p_cuMulticastGetGranularity(&gran, &mp, CU_MULTICAST_GRANULARITY_RECOMMENDED));
    const size_t size = (want + gran - 1) / gran * gran;
    mp.size = size;
    printf("granularity %zu, size %zu\n", gran, size);
    CUmemGenericAllocationHandle mc;
    CK(p_cuMulticastCreate(&mc, &mp));
    for (int d = 0; d < G; ++d) { CUdevice dev; CK(p_cuDeviceGet(&dev, d)); CK(p_cuMulticastAddDevice(mc, dev)); }
    std::vector<CUmemAccessDesc> acc(G);
    for (int d = 0; d < G; ++d) { acc[d].location.type = CU_MEM_LOCATION_TYPE_DEVICE; acc[d].location.id = d; acc[d].flags = CU_MEM_ACCESS_FLAGS_PROT_READWRITE; }
    std::vector<CUdeviceptr> uc(G);
    for (int d = 0; d < G; ++d) {
        CUmemAllocationProp ap{}; ap.type = CU_MEM_ALLOCATION_TYPE_PINNED; ap.location.type = CU_MEM_LOCATION_TYPE_DEVICE; ap.location.id = d;
        CUmemGenericAllocationHandle h;
        CK(p_cuMemCreate(&h, size, &ap, 0));
        CK(p_cuMulticastBindMem(mc, 0, h, 0, size, 0));
        CK(p_cuMemAddressReserve(&uc[d], size, gran, 0, 0));
        CK(p_cuMemMap(uc[d], size, 0, h, 0));
        CK(p_cuMemSetAccess(uc[d], size, acc.data(), G));
    }
    CUdeviceptr mcva;
    CK(p_cuMemAddressReserve(&mcva, size, gran, 0, 0));
    CK(p_cuMemMap(mcva, size, 0, mc, 0));
    CK(p_cuMemSetAccess(mcva, size, acc.data(), G));
    printf("mapped: multicast va %p\n", (void*)mcva);
    for (int d = 0; d < G; ++d) { cudaSetDevice(d); k_fill<<<296, 256>>>((uint32_t*)uc[d], n_px, 7u * (d + 1)); cudaDeviceSynchronize(); }
    // every device reduces its slice (one stream each); timed on device 0 .. G-1 together with events on device 0's view
    const size_t n64 = n_px / 2;
    std::vector<cudaStream_t> st(G);
    std::vector<cudaEvent_t> e0(G), e1(G);
    for (int d = 0; d < G; ++d) { cudaSetDevice(d); cudaStreamCreate(&st[d]); cudaEventCreate(&e0[d]); cudaEventCreate(&e1[d]); }
    for (int rep = 0; rep < 3; ++rep) {
        if (rep) for (int d = 0; d < G; ++d) { cudaSetDevice(d); k_fill<<<296, 256>>>((uint32_t*)uc[d], n_px, 7u * (d + 1)); cudaDeviceSynchronize(); }
        for (int d = 0; d < G; ++d) {
            cudaSetDevice(d);
            cudaEventRecord(e0[d], st[d]);
            k_mc_reduce<<<148, 512, 0, st[d]>>>((unsigned long long*)mcva, n64 * d / G, n64 * (d + 1) / G);
            cudaEventRecord(e1[d], st[d]);
        }
        float worst = 0;
        for (int d = 0; d < G; ++d) {
            cudaSetDevice(d);
            cudaError_t e = cudaStreamSynchronize(st[d]);
            if (e != cudaSuccess) { printf("kernel on device %d failed: %s\n", d, cudaGetErrorString(e)); return 1; }
            float ms = 0; cudaEventElapsedTime(&ms, e0[d], e1[d]); worst = ms > worst ? ms : worst;
        }
        printf("rep %d: in-switch reduce + broadcast of %zu MB over %d devices: %.1f us (slowest device's own slice)\n", rep, want >> 20, G, worst * 1e3);
    }
    // check: every copy holds the sum of the G fills
    std::vector<uint32_t> h(n_px);
    size_t bad = 0;
    for (int d = 0; d < G; ++d) {
        cudaSetDevice(d);
        cudaMemcpy(h.data(), (void*)uc[d], want, cudaMemcpyDeviceToHost);
        for (size_t i = 0; i < n_px; ++i) { uint32_t w = 0; for (int e = 0; e < G; ++e) w += 7u * (e + 1) + (uint32_t)(i % 1000); if (h[i] != w) ++bad; }
    }
    printf("mismatching words over all copies: %zu\n", bad);
    return 0;
}
