// tma_host.cuh -- host side of the TMA paths: the driver's tensor-map encoder, fetched through the
// runtime (cudaGetDriverEntryPoint: the library does not link libcuda), and a 2-D tiled map maker.
#pragma once
#include <cuda.h>  // CUtensorMap and its enums (types only)
#include <cuda_runtime.h>

namespace udp {

typedef CUresult (*tma_encode_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *,
                                  const cuuint64_t *, const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

// nullptr when the driver does not offer the encoder (the callers then keep to their non-TMA kernels).
inline tma_encode_fn tma_encoder()
{
    static tma_encode_fn fn = nullptr;
    static int state = 0;  // 0 not tried, 1 ok, -1 unavailable (a benign race: every thread computes the same value)
    if (state == 0) {
        void *p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess && p &&
            q == cudaDriverEntryPointSuccess) {
            fn = (tma_encode_fn)p;
            state = 1;
        } else {
            (void)cudaGetLastError();
            state = -1;
        }
    }
    return state == 1 ? fn : nullptr;
}

// Row-major 2-D tensor [dim1][dim0] of `type`, rows `stride_bytes` apart, box [box1][box0]; no swizzle,
// zeros outside the tensor.  false if the encoder is missing or rejects the shape.
inline bool tma_make_2d(CUtensorMap *tm, CUtensorMapDataType type, const void *base, unsigned long long dim0,
                        unsigned long long dim1, unsigned long long stride_bytes, unsigned box0, unsigned box1)
{
    tma_encode_fn fn = tma_encoder();
    if (!fn) return false;
    const cuuint64_t gdim[2] = {dim0, dim1}, gstride[1] = {stride_bytes};
    const cuuint32_t box[2] = {box0, box1}, estride[2] = {1, 1};
    return fn(tm, type, 2, const_cast<void *>(base), gdim, gstride, box, estride, CU_TENSOR_MAP_INTERLEAVE_NONE,
              CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

}  // namespace udp
