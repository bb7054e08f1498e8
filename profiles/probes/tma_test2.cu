#include <cuda.h>
#include <cuda_runtime.h>
#include <cstdio>
#include <cstdint>
#include <vector>
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned parity)
{
    asm volatile("{\n.reg .pred p;\nW:\nmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n@p bra D;\nbra W;\nD:\n}\n" ::"r"(mbar), "r"(parity) : "memory");
}
__global__ void k(const __grid_constant__ CUtensorMap tm, const CUtensorMap *gtm, int use_g, int c0, int c1, uint8_t *out, int nbytes)
{
    extern __shared__ __align__(128) unsigned char fp[];
    __shared__ __align__(8) unsigned long long bar;
    const unsigned fpa = (unsigned)__cvta_generic_to_shared(fp), mbar = (unsigned)__cvta_generic_to_shared(&bar);
    const int lane = threadIdx.x;
    unsigned long long gdesc = use_g ? (unsigned long long)(uintptr_t)gtm : (unsigned long long)(uintptr_t)&tm;
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    if (lane == 0) {
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(mbar), "r"(nbytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(fpa), "l"(gdesc), "r"(c0), "r"(c1), "r"(mbar) : "memory");
    }
    mbar_wait(mbar, 0);
    for (int i = lane; i < nbytes; i += 32) out[i] = fp[i];
}
typedef CUresult (*encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *, const cuuint32_t *,
                              const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int main()
{
    void *fn = nullptr; cudaDriverEntryPointQueryResult q;
    cudaFree(0);
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q);
    const int sh = 480, sw = 640, ss = 1920;
    std::vector<uint8_t> img(sh * ss);
    for (size_t i = 0; i < img.size(); ++i) img[i] = (uint8_t)((i * 2654435761u) >> 13);
    uint8_t *dimg, *dout; CUtensorMap *dtm;
    cudaMalloc(&dimg, img.size()); cudaMalloc(&dout, 65536); cudaMalloc(&dtm, 128);
    cudaMemcpy(dimg, img.data(), img.size(), cudaMemcpyHostToDevice);
    const cuuint32_t es[2] = {1, 1};
    for (int variant = 0; variant < 8; ++variant) {
        const int boxw = variant < 2 ? 64 : (variant < 4 ? 160 : 256), boxh = variant < 2 ? 32 : (variant < 4 ? 50 : 90);
        const int c0s[8] = {16, -16, -32, 1904, 1888, -160, 7, -7}; const int c1s[8] = {8, -3, 470, -20, 479, 0, 3, -3}; const int c0 = c0s[variant], c1 = c1s[variant];
        CUtensorMap want;
        const cuuint64_t gdim2[2] = {3 * sw, sh}, gstride2[1] = {ss};
        const cuuint32_t box2[2] = {(cuuint32_t)boxw, (cuuint32_t)boxh};
        int rc = (int)((encode_fn)fn)(&want, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, dimg, gdim2, gstride2, box2, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                      CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        cudaMemcpy(dtm, &want, 128, cudaMemcpyHostToDevice);
        for (int use_g = 0; use_g < 2; ++use_g) {
            cudaMemset(dout, 0xcd, 65536);
            k<<<1, 32, boxw * boxh>>>(want, dtm, use_g, c0, c1, dout, boxw * boxh);
            cudaError_t e = cudaDeviceSynchronize();
            printf("variant %d box %dx%d at (%d,%d) use_g %d encode %d: %s\n", variant, boxw, boxh, c0, c1, use_g, rc, cudaGetErrorString(e));
            if (e != cudaSuccess) return 1;
            std::vector<uint8_t> out(65536);
            cudaMemcpy(out.data(), dout, 65536, cudaMemcpyDeviceToHost);
            int bad = 0;
            for (int r = 0; r < boxh; ++r) for (int c = 0; c < boxw; ++c) {
                int x = c0 + c, y = c1 + r;
                uint8_t w = (x >= 0 && x < 3 * sw && y >= 0 && y < sh) ? img[(size_t)y * ss + x] : 0;
                if (out[r * boxw + c] != w) ++bad;
            }
            printf("  bad %d of %d\n", bad, boxw * boxh);
        }
    }
    return 0;
}
