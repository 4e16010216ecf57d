// Throughput probe for the packed f32x2 FMA (FFMA2) of sm_100a against scalar FMUL / FADD / FFMA.
// FFMA2 with an opaque -0 addend is an exact packed multiply, with an opaque 1.0 multiplier an exact packed add
// (ptxas contracts mul.rn.f32x2 + add.rn.f32x2 into one fused FFMA2, so the constants must not be visible to it).
// build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -fmad=false -o ffma2_probe ffma2_probe.cu ; run on the GPU box.
#include <cstdio>
#include <cuda_runtime.h>
typedef unsigned long long u64;
struct K { u64 neg0, one; float b, c; };
__device__ __forceinline__ u64 pk(float a, float b) { u64 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ u64 fma2(u64 a, u64 b, u64 c) { u64 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
constexpr int kAcc = 8, kIter = 2048;
// MODE 0: scalar mul then add (2 instr per value)   1: packed mul then add (2 FFMA2 per 2 values)
// MODE 2: scalar fused FFMA                          3: packed fused FFMA2
template <int MODE>
__global__ void __launch_bounds__(256) k(float* out, const K p) {
    float a[2 * kAcc];
    u64 v[kAcc];
#pragma unroll
    for (int i = 0; i < 2 * kAcc; ++i) a[i] = (float)(threadIdx.x + i) * 1e-3f;
#pragma unroll
    for (int i = 0; i < kAcc; ++i) v[i] = pk(a[2 * i], a[2 * i + 1]);
    const u64 b2 = pk(p.b, p.b), c2 = pk(p.c, p.c);
    for (int it = 0; it < kIter; ++it) {
#pragma unroll
        for (int i = 0; i < kAcc; ++i) {
            if (MODE == 0) { a[2 * i] = a[2 * i] * p.b + p.c; a[2 * i + 1] = a[2 * i + 1] * p.b + p.c; }
            if (MODE == 1) { v[i] = fma2(fma2(v[i], b2, p.neg0), p.one, c2); }
            if (MODE == 2) { a[2 * i] = __fmaf_rn(a[2 * i], p.b, p.c); a[2 * i + 1] = __fmaf_rn(a[2 * i + 1], p.b, p.c); }
            if (MODE == 3) { v[i] = fma2(v[i], b2, c2); }
        }
    }
    float s = 0.0f;
#pragma unroll
    for (int i = 0; i < kAcc; ++i) {
        float x, y;
        asm("mov.b64 {%0, %1}, %2;" : "=f"(x), "=f"(y) : "l"(v[i]));
        s += x + y + a[2 * i] + a[2 * i + 1];
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}
template <int MODE>
void run(const char* name, float* out, K p, int values_per_instr_pair) {
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    const int grid = 148 * 8;
    k<MODE><<<grid, 256>>>(out, p);
    cudaEventRecord(e0);
    k<MODE><<<grid, 256>>>(out, p);
    cudaEventRecord(e1);
    cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double values = (double)grid * 256 * 2 * kAcc * kIter;            // (a*b+c) results
    printf("%-28s %8.3f ms  %7.2f G results/s  (%d instr per 2 results)\n", name, ms, values / ms * 1e-6, values_per_instr_pair);
}
int main() {
    float* out; cudaMalloc(&out, 148 * 8 * 256 * 4);
    K p{0x8000000080000000ull, 0x3f8000003f800000ull, 0.999f, 1e-3f};
    run<0>("scalar FMUL + FADD", out, p, 4);
    run<1>("packed mul2 + add2 (2 FFMA2)", out, p, 2);
    run<2>("scalar FFMA", out, p, 2);
    run<3>("packed FFMA2", out, p, 1);
    float h[4]; cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost);
    printf("check %g %s\n", h[0], cudaGetErrorString(cudaGetLastError()));
    return 0;
}
