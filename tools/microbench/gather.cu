// Microbenchmark: divergent 8-byte gathers from a 28.7 MB table (the Mo kernel's access pattern), through
// LDG (ld.global.nc), LDG with L1 bypass (ld.global.cg), a 1-D texture object on linear memory, and shared memory.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather gather.cu ; prints ns per 32-lane gather and cycles per lane-gather per SM.
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)
__device__ __forceinline__ uint32_t lcg(uint32_t &s) { s = s * 1664525u + 1013904223u; return s; }
template <int MODE, int ILP>
__global__ void k_gather(const float2 *__restrict__ tab, cudaTextureObject_t tex, uint32_t n, int iters, float *out) {
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    float acc = 0.f;
    for (int it = 0; it < iters; ++it) {
        uint32_t idx[ILP]; float2 v[ILP];
#pragma unroll
        for (int k = 0; k < ILP; ++k) idx[k] = (uint32_t)(((uint64_t)lcg(s) * n) >> 32);
#pragma unroll
        for (int k = 0; k < ILP; ++k) {
            if (MODE == 0) v[k] = __ldg(&tab[idx[k]]);
            else if (MODE == 1) v[k] = __ldcg(&tab[idx[k]]);
            else if (MODE == 2) v[k] = tex1Dfetch<float2>(tex, (int)idx[k]);
            else { float a = __ldg(reinterpret_cast<const float *>(tab) + 2 * idx[k]), b = __ldg(reinterpret_cast<const float *>(tab) + 2 * idx[k] + 1); v[k] = make_float2(a, b); }
        }
#pragma unroll
        for (int k = 0; k < ILP; ++k) acc += v[k].x * 0.5f + v[k].y;
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
template <int ILP>
__global__ void k_gather_smem(const float2 *__restrict__ tab, uint32_t n, int iters, float *out) {
    extern __shared__ float2 sm[];
    for (uint32_t i = threadIdx.x; i < n; i += blockDim.x) sm[i] = tab[i];
    __syncthreads();
    uint32_t s = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    float acc = 0.f;
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < ILP; ++k) { uint32_t i = (uint32_t)(((uint64_t)lcg(s) * n) >> 32); float2 v = sm[i]; acc += v.x * 0.5f + v.y; }
    }
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
}
int main() {
    const uint32_t n = 30u * 119766u;
    float2 *tab; float *out;
    CK(cudaMalloc(&tab, n * sizeof(float2))); CK(cudaMemset(tab, 0, n * sizeof(float2)));
    const int blocks = 148 * 8, threads = 256, iters = 256;
    CK(cudaMalloc(&out, blocks * threads * sizeof(float)));
    cudaResourceDesc rd = {}; rd.resType = cudaResourceTypeLinear; rd.res.linear.devPtr = tab; rd.res.linear.desc = cudaCreateChannelDesc<float2>(); rd.res.linear.sizeInBytes = n * sizeof(float2);
    cudaTextureDesc td = {}; td.readMode = cudaReadModeElementType;
    cudaTextureObject_t tex; CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto report = [&](const char *name, int ilp, float ms) {
        double lanes = (double)blocks * threads * iters * ilp;
        printf("%-28s ILP %2d  %8.3f ms  %6.2f G lane-gathers/s  %5.2f SM-cycles per lane-gather (1.965 GHz, 148 SMs)\n", name, ilp, ms, lanes / ms / 1e6,
               ms * 1e-3 * 1.965e9 * 148 / lanes);
    };
#define RUN(MODE, ILP, NAME) { k_gather<MODE, ILP><<<blocks, threads>>>(tab, tex, n, 8, out); cudaEventRecord(e0); k_gather<MODE, ILP><<<blocks, threads>>>(tab, tex, n, iters, out); cudaEventRecord(e1); CK(cudaEventSynchronize(e1)); float ms; cudaEventElapsedTime(&ms, e0, e1); report(NAME, ILP, ms); }
    RUN(0, 1, "ldg.nc float2") RUN(0, 4, "ldg.nc float2") RUN(0, 8, "ldg.nc float2") RUN(0, 16, "ldg.nc float2")
    RUN(1, 4, "ld.cg float2") RUN(1, 16, "ld.cg float2")
    RUN(2, 1, "tex1Dfetch float2") RUN(2, 4, "tex1Dfetch float2") RUN(2, 8, "tex1Dfetch float2") RUN(2, 16, "tex1Dfetch float2")
    RUN(3, 8, "2 x ldg.nc float")
    {   // shared-memory gathers from a 16 K-entry (128 KB) table for scale
        const uint32_t ns = 16384; CK(cudaFuncSetAttribute(k_gather_smem<8>, cudaFuncAttributeMaxDynamicSharedMemorySize, ns * 8));
        k_gather_smem<8><<<148, 1024, ns * 8>>>(tab, ns, 8, out); cudaEventRecord(e0); k_gather_smem<8><<<148, 1024, ns * 8>>>(tab, ns, 1024, out); cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
        float ms; cudaEventElapsedTime(&ms, e0, e1); double lanes = 148.0 * 1024 * 1024 * 8;
        printf("%-28s         %8.3f ms  %6.2f G lane-gathers/s  %5.2f SM-cycles per lane-gather\n", "smem float2 (128 KB table)", ms, lanes / ms / 1e6, ms * 1e-3 * 1.965e9 * 148 / lanes);
    }
    return 0;
}
