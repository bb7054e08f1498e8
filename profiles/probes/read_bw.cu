// Read-only streaming ceiling with the decoders' access pattern: persistent CTAs pull 12 KB maps
// (two tensors) into shared memory with 16-byte cp.async, two stages in flight, and touch nothing.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o read_bw read_bw.cu && ./read_bw
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
__global__ void __launch_bounds__(96, 5) k(const float *a, const float *b, long long maps, float *sink, int two)
{
    extern __shared__ __align__(16) float sm[];  // 2 stages x 2 tensors x 3072 floats
    const unsigned base = (unsigned)__cvta_generic_to_shared(sm);
    float acc = 0.f;
    int st = 0;
    auto pre = [&](long long m, int s) {
        const float *pa = a + m * 3072, *pb = b + m * 3072;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const int q = threadIdx.x + j * 96;
            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(base + (s * 6144 + q * 4) * 4), "l"(pa + q * 4) : "memory");
            if (two) asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(base + (s * 6144 + 3072 + q * 4) * 4), "l"(pb + q * 4) : "memory");
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    long long m = blockIdx.x;
    if (m < maps) pre(m, 0);
    for (; m < maps; m += gridDim.x, st ^= 1) {
        const long long nx = m + gridDim.x;
        if (nx < maps) pre(nx, st ^ 1);
        else asm volatile("cp.async.commit_group;" ::: "memory");
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncthreads();
        acc += sm[st * 6144 + threadIdx.x];
        __syncthreads();
    }
    if (acc == 123.456f) *sink = acc;
}
int main()
{
    const long long maps = 16384LL * 17;
    float *a, *b, *sink;
    cudaMalloc(&a, maps * 3072 * 4); cudaMalloc(&b, maps * 3072 * 4); cudaMalloc(&sink, 4);
    cudaMemset(a, 0, maps * 3072 * 4); cudaMemset(b, 0, maps * 3072 * 4);
    cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 49152);
    for (int two = 0; two < 2; ++two)
        for (int per_sm : {2, 3, 4, 5}) {
            cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
            k<<<148 * per_sm, 96, 49152>>>(a, b, maps, sink, two);
            cudaEventRecord(e0);
            for (int r = 0; r < 5; ++r) k<<<148 * per_sm, 96, 49152>>>(a, b, maps, sink, two);
            cudaEventRecord(e1); cudaEventSynchronize(e1);
            float ms; cudaEventElapsedTime(&ms, e0, e1); ms /= 5;
            printf("tensors %d CTAs/SM %d: %.3f ms  %.0f GB/s  (%s)\n", two + 1, per_sm, ms, maps * 3072.0 * 4 * (two + 1) / ms / 1e6,
                   cudaGetErrorString(cudaGetLastError()));
        }
    return 0;
}
