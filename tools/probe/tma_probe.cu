// Probe: does a 2-D tiled TMA store (u8, box 192 x 32) accept boxes that stick out of the tensor (negative y, x past
// the row end)?  nvcc -gencode arch=compute_100a,code=sm_100a -o tma_probe tma_probe.cu && ./tma_probe X Y
#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
#include <cstdlib>

__global__ void k_store(const __grid_constant__ CUtensorMap map, int x, int y) {
    __shared__ __align__(128) uint8_t tile[32 * 192];
    for (int i = threadIdx.x; i < 32 * 192; i += blockDim.x) tile[i] = (uint8_t)(1 + i / 192);   // row r holds r+1
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t src = (uint32_t)__cvta_generic_to_shared(tile);
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%2, %3}], [%1];"
                     :: "l"(reinterpret_cast<uint64_t>(&map)), "r"(src), "r"(x), "r"(y) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
}

int main(int argc, char** argv) {
    const int W = 336 * 3, H = 250;
    uint8_t* d;
    cudaMalloc(&d, (size_t)W * H);
    using Encode = CUresult (*)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
    void* fn = nullptr;
    cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    CUtensorMap map;
    cuuint64_t dims[2] = {(cuuint64_t)W, (cuuint64_t)H}, pitch[1] = {(cuuint64_t)W};
    cuuint32_t box[2] = {192, 32}, step[2] = {1, 1};
    CUresult r = ((Encode)fn)(&map, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, d, dims, pitch, box, step, CU_TENSOR_MAP_INTERLEAVE_NONE,
                              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    printf("encode %d\n", (int)r);
    int cases[1][2] = {{argc > 2 ? atoi(argv[1]) : 0, argc > 2 ? atoi(argv[2]) : 0}};
    for (auto& c : cases) {
        cudaMemset(d, 0, (size_t)W * H);
        k_store<<<1, 256>>>(map, c[0], c[1]);
        cudaError_t e = cudaDeviceSynchronize();
        std::vector<uint8_t> h((size_t)W * H);
        if (e == cudaSuccess) cudaMemcpy(h.data(), d, h.size(), cudaMemcpyDeviceToHost);
        long nz = 0; for (auto v : h) nz += v != 0;
        // expected: rows max(y,0)..min(y+31,H-1), bytes max(x,0)..min(x+191,W-1)
        int y0 = c[1] < 0 ? 0 : c[1], y1 = c[1] + 31 >= H ? H - 1 : c[1] + 31, x0 = c[0] < 0 ? 0 : c[0], x1 = c[0] + 191 >= W ? W - 1 : c[0] + 191;
        long expect = (long)(y1 - y0 + 1) * (x1 - x0 + 1);
        bool rows_ok = true;
        if (e == cudaSuccess) for (int yy = y0; yy <= y1; ++yy) if (h[(size_t)yy * W + x0] != (uint8_t)(1 + yy - c[1])) rows_ok = false;
        printf("x %4d y %4d : %s  nonzero %ld expected %ld rows %s\n", c[0], c[1], cudaGetErrorString(e), nz, expect, rows_ok ? "ok" : "WRONG");
        if (e != cudaSuccess) { printf("context lost\n"); return 1; }
    }
    return 0;
}
