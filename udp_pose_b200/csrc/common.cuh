// common.cuh -- shared helpers for the sm_100a UDP codec kernels.
//
// Compile flags that are part of the numerical contract (see __graft_entry__.build()):
//   -fmad=false   : no implicit FMA contraction; fused ops appear only where fmaf()/fma() is
//                   written, mirroring OpenCV's FMA order and numpy's
//                   separately rounded multiplies and adds;
//   no -use_fast_math: IEEE division / sqrt, denormals kept.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdlib.h>

#include "../../include/udp_codec.h"

#define UDP_LAUNCH_CHECK()                     \
    do {                                       \
        cudaError_t e__ = cudaGetLastError();  \
        if (e__ != cudaSuccess) return (int)e__; \
    } while (0)

// UDP_FORCE_GENERIC=1 in the environment routes the decoders through the shape-generic kernels
// (used by the tests to cover both paths).
inline bool udp_force_generic()
{
    const char *e = getenv("UDP_FORCE_GENERIC");
    return e && e[0] == '1';
}

namespace udp {

// Number of SMs of the current device (148 on B200), queried once per device.
inline int sm_count()
{
    static int cached[64] = {0};
    int dev = 0;
    cudaGetDevice(&dev);
    if (dev < 0 || dev >= 64) dev = 0;
    if (!cached[dev]) {
        int n = 0;
        if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || n < 1) n = 148;
        cached[dev] = n;
    }
    return cached[dev];
}

// Opt a persistent kernel into `smem` bytes of dynamic shared memory and return how many of its
// CTAs fit on one SM.  0 on success, otherwise a cudaError_t.
inline int persistent_grid(const void *kern, int threads, size_t smem, int *per_sm)
{
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    int n = 0;
    e = cudaOccupancyMaxActiveBlocksPerMultiprocessor(&n, kern, threads, smem);
    if (e != cudaSuccess) return (int)e;
    *per_sm = n < 1 ? 1 : n;
    return 0;
}

constexpr double kPi = 3.141592653589793;  // math.pi

// 16-byte streaming store: outputs are written once and never re-read by this library.
__device__ __forceinline__ void st_stream_f4(float *p, float4 v) { __stcs(reinterpret_cast<float4 *>(p), v); }
// 16-byte read-once load (keeps L1 for the gathers that do reuse lines).
__device__ __forceinline__ float4 ld_stream_f4(const float *p) { return __ldcs(reinterpret_cast<const float4 *>(p)); }

// Packed fp32x2 values live in 64-bit registers from the moment they are formed, so ptxas keeps
// them as aligned register pairs (holding the halves separately cost two MOVs per FFMA2 in the
// first build of this file -- 31 % of all executed instructions were moves).
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk(float lo, float hi)
{
    f32x2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ float lo_of(f32x2 v)
{
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return a;
}
__device__ __forceinline__ float hi_of(f32x2 v)
{
    float a, b;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(a), "=f"(b) : "l"(v));
    return b;
}
__device__ __forceinline__ f32x2 ffma2p(f32x2 a, float k, f32x2 c)
{
    f32x2 d;
    asm("{.reg .b64 kk;\n mov.b64 kk, {%2, %2};\n fma.rn.f32x2 %0, %1, kk, %3;}" : "=l"(d) : "l"(a), "f"(k), "l"(c));
    return d;
}
__device__ __forceinline__ f32x2 fmul2p(f32x2 a, float k)
{
    f32x2 d;
    asm("{.reg .b64 kk;\n mov.b64 kk, {%2, %2};\n mul.rn.f32x2 %0, %1, kk;}" : "=l"(d) : "l"(a), "f"(k));
    return d;
}
__device__ __forceinline__ f32x2 fadd2p(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 fsub2p(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 fmul2pp(f32x2 a, f32x2 b)
{
    f32x2 d;
    asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b));
    return d;
}
__device__ __forceinline__ f32x2 ffma2pp(f32x2 a, f32x2 b, f32x2 c)
{
    f32x2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// Butterfly reductions in blocks whose size need not be a multiple of 32 (the last warp is then partial): shuffles
// name the lanes that exist, and a partner that does not exist contributes the lane's own value.
__device__ __forceinline__ unsigned warp_lanes()
{
    const unsigned n = blockDim.x - (threadIdx.x & ~31u);
    return n >= 32u ? 0xffffffffu : ((1u << n) - 1u);
}
template <class T>
__device__ __forceinline__ T shfl_xor_or_self(unsigned lanes, T v, int o)
{
    const T u = __shfl_xor_sync(lanes, v, o);
    return ((lanes >> ((threadIdx.x & 31u) ^ (unsigned)o)) & 1u) ? u : v;
}

// Wait until the mbarrier at shared address `mbar` completes the phase of parity `parity`
// (try_wait suspends the thread in hardware between polls).
__device__ __forceinline__ void mbar_wait(unsigned mbar, unsigned parity)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "UDP_MBAR_WAIT:\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
        "@p bra UDP_MBAR_DONE;\n"
        "bra UDP_MBAR_WAIT;\n"
        "UDP_MBAR_DONE:\n"
        "}\n" ::"r"(mbar), "r"(parity) : "memory");
}

// center/scale loader honouring the caller's dtype.
struct CS {
    double c0, c1;   // center as double (exact widening when float32)
    double st0, st1; // scale*200 rounded in the caller's dtype, widened
    float c0f, c1f, st0f, st1f;
};

__device__ __forceinline__ CS load_cs(const void *center, const void *scale, int cs_f64, int n)
{
    CS r;
    if (cs_f64) {
        const double *c = reinterpret_cast<const double *>(center) + 2 * (size_t)n;
        const double *s = reinterpret_cast<const double *>(scale) + 2 * (size_t)n;
        r.c0 = c[0]; r.c1 = c[1];
        r.st0 = __dmul_rn(s[0], 200.0); r.st1 = __dmul_rn(s[1], 200.0);
        r.c0f = r.c1f = r.st0f = r.st1f = 0.f;
    } else {
        const float *c = reinterpret_cast<const float *>(center) + 2 * (size_t)n;
        const float *s = reinterpret_cast<const float *>(scale) + 2 * (size_t)n;
        r.c0f = c[0]; r.c1f = c[1];
        r.st0f = __fmul_rn(s[0], 200.0f); r.st1f = __fmul_rn(s[1], 200.0f);
        r.c0 = r.c0f; r.c1 = r.c1f; r.st0 = r.st0f; r.st1 = r.st1f;
    }
    return r;
}

}  // namespace udp
