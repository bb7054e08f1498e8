// crop.cu -- the UDP person crop: bit-exact restatement of OpenCV's fixed-point bilinear
// warpAffine (u8, 3 channels, WARP_INVERSE_MAP | INTER_LINEAR, constant-0 border).
//
// Reference call sites: deep_hrnet/lib/dataset/JointsDataset.py:227,
// RSN/dataset/JointsDataset.py:197-198.  OpenCV arithmetic:
//   X = (rint((M01*y+M02)*1024) + 16 + rint(M00*x*1024)) >> 5      (Q5 source coordinate)
//   sx = X>>5, fx = X&31 ; likewise Y ; weights (32-fy)(32-fx)*32 ... sum to 2^15
//   dst = (sum w*p + 2^14) >> 15
// The four weights factor exactly (integers, no intermediate rounding):
//   sum w*p = 32 * [ (32-fy)*((32-fx)*p00 + fx*p01) + fy*((32-fx)*p10 + fx*p11) ]
// so dst = (S + 512) >> 10 with S the bracket -- one dp4a per source row pair.
//
// Layout: one CTA = one person x ROWS output rows.  The per-column deltas
// rint(M00*x*1024), rint(M10*x*1024) and the per-row bases are float64 and are computed
// once per CTA into shared memory; the per-pixel work is integer only.  The finished
// ROWS x W x 3 byte tile is staged in shared memory and leaves as 16-byte streaming stores.
#include "common.cuh"

namespace udp {

constexpr int CROP_ROWS = 16;

__device__ __forceinline__ uint32_t ldg_u8(const uint8_t *p) { return (uint32_t)__ldg(p); }

// Border pixel (some of the 2x2 taps fall outside the image): per-tap bounds checks, byte loads.
// Kept out of line so the interior path stays branch- and predicate-free.
__device__ __noinline__ uint32_t crop_border_pixel(const uint8_t *src, int sh, int sw, int ss, int Xq10, int Yq10)
{
    const int X = Xq10 >> 5, Y = Yq10 >> 5;
    const int sx = X >> 5, sy = Y >> 5;
    if (sx < -1 || sx >= sw || sy < -1 || sy >= sh) return 0u;  // all four taps outside
    const uint32_t fx = X & 31, fy = Y & 31, wy0 = 32u - fy;
    const bool x0 = sx >= 0, x1 = sx + 1 < sw, y0 = sy >= 0, y1 = sy + 1 < sh;
    uint32_t c[4] = {0u, 0u, 0u, 0u};  // r | g<<8 | b<<16 of the taps 00 01 10 11
    if (y0 && x0) { const uint8_t *q = src + (size_t)sy * ss + sx * 3; c[0] = ldg_u8(q) | ldg_u8(q + 1) << 8 | ldg_u8(q + 2) << 16; }
    if (y0 && x1) { const uint8_t *q = src + (size_t)sy * ss + (sx + 1) * 3; c[1] = ldg_u8(q) | ldg_u8(q + 1) << 8 | ldg_u8(q + 2) << 16; }
    if (y1 && x0) { const uint8_t *q = src + (size_t)(sy + 1) * ss + sx * 3; c[2] = ldg_u8(q) | ldg_u8(q + 1) << 8 | ldg_u8(q + 2) << 16; }
    if (y1 && x1) { const uint8_t *q = src + (size_t)(sy + 1) * ss + (sx + 1) * 3; c[3] = ldg_u8(q) | ldg_u8(q + 1) << 8 | ldg_u8(q + 2) << 16; }
    uint32_t o = 0;
#pragma unroll
    for (int ch = 0; ch < 3; ++ch) {
        const uint32_t t0 = (c[0] >> (8 * ch)) & 255u, t1 = (c[1] >> (8 * ch)) & 255u;
        const uint32_t t2 = (c[2] >> (8 * ch)) & 255u, t3 = (c[3] >> (8 * ch)) & 255u;
        const uint32_t S = wy0 * ((32u - fx) * t0 + fx * t1) + fy * ((32u - fx) * t2 + fx * t3);
        o |= ((S + 512u) >> 10) << (8 * ch);
    }
    return o;
}

__global__ void __launch_bounds__(512)
warp_affine_u8_kernel(const uint8_t *const *src_ptrs, const int32_t *src_h, const int32_t *src_w,
                      const int32_t *src_stride, const float *M, uint8_t *dst, int W, int H)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    int32_t *ad = reinterpret_cast<int32_t *>(smem_raw);  // [W]
    int32_t *bd = ad + W;                                 // [W]
    int32_t *X0 = bd + W;                                 // [CROP_ROWS]
    int32_t *Y0 = X0 + CROP_ROWS;                         // [CROP_ROWS]
    uint8_t *tile = reinterpret_cast<uint8_t *>(Y0 + CROP_ROWS);  // [CROP_ROWS][W*3], 16B aligned (W even)

    const int n = blockIdx.y;
    const int y_base = blockIdx.x * CROP_ROWS;
    const int rows = min(CROP_ROWS, H - y_base);
    const float *m = M + 6 * (size_t)n;
    const double m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5];

    for (int x = threadIdx.x; x < W; x += (int)blockDim.x) {
        ad[x] = __double2int_rn(__dmul_rn(__dmul_rn(m00, (double)x), 1024.0));
        bd[x] = __double2int_rn(__dmul_rn(__dmul_rn(m10, (double)x), 1024.0));
    }
    if (threadIdx.x < rows) {
        double y = (double)(y_base + threadIdx.x);
        X0[threadIdx.x] = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(m01, y), m02), 1024.0)) + 16;
        Y0[threadIdx.x] = __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(m11, y), m12), 1024.0)) + 16;
    }
    __syncthreads();

    const uint8_t *src = src_ptrs[n];
    const int sh = src_h[n], sw = src_w[n], ss = src_stride[n];
    const int row_bytes = W * 3;

    // Thread = output column x (consecutive lanes = consecutive x, so a warp's loads touch one
    // or two 128-byte lines), looping over the CTA's rows with the column deltas in registers.
    // ncu on the earlier 4-pixels-per-thread layout: LSU wavefronts at 75 % of peak with five
    // wavefronts per request; on the 1-pixel-per-iteration layout ~98 instructions per pixel,
    // half of them loop/index overhead -- hence this shape.
    const uint8_t *src_al = reinterpret_cast<const uint8_t *>((uintptr_t)src & ~(uintptr_t)3);
    const int mis = (int)((uintptr_t)src & 3);
    const uint32_t sw5 = sw > 5 ? (uint32_t)(sw - 5) : 0u, sh1 = sh > 1 ? (uint32_t)(sh - 1) : 0u;
    for (int x = threadIdx.x; x < W; x += (int)blockDim.x) {
        const int adx = ad[x], bdx = bd[x];
        uint8_t *o = tile + x * 3;
#pragma unroll 4
        for (int r = 0; r < rows; ++r, o += row_bytes) {
            const int X = X0[r] + adx, Y = Y0[r] + bdx;  // Q10
            const int sx = X >> 10, sy = Y >> 10;
            uint32_t px;  // byte 0,1,2 = r,g,b
            // interior: 1 <= sx, sx + 4 < sw, 0 <= sy, sy + 1 < sh (keeps the 12-byte reads inside the image)
            if ((uint32_t)(sx - 1) < sw5 && (uint32_t)sy < sh1) {
                const uint32_t fx = ((uint32_t)X >> 5) & 31u, fy6 = ((uint32_t)Y << 1) & 0x7c0u;  // fy * 64
                const uint32_t wx = (32u - fx) | (fx << 8), wxh = wx << 16;
                const int o0 = sy * ss + (sx * 3 + mis), o1 = o0 + ss;
                const uint32_t *q0 = reinterpret_cast<const uint32_t *>(src_al + (o0 & ~3));
                const uint32_t *q1 = reinterpret_cast<const uint32_t *>(src_al + (o1 & ~3));
                const uint32_t u0 = __ldg(q0), u1 = __ldg(q0 + 1), u2 = __ldg(q0 + 2);
                const uint32_t v0 = __ldg(q1), v1 = __ldg(q1 + 1), v2 = __ldg(q1 + 2);
                const uint32_t s0 = (uint32_t)o0 << 3, s1 = (uint32_t)o1 << 3;  // funnel shift uses bits 0..4
                const uint32_t lo0 = __funnelshift_r(u0, u1, s0), hi0 = __funnelshift_r(u1, u2, s0);
                const uint32_t lo1 = __funnelshift_r(v0, v1, s1), hi1 = __funnelshift_r(v1, v2, s1);
                const uint32_t a0 = __byte_perm(lo0, lo1, 0x7430);  // [r00 r01 r10 r11]
                const uint32_t g0 = __byte_perm(lo0, hi0, 0x5241);  // [g00 g01 b00 b01]
                const uint32_t g1 = __byte_perm(lo1, hi1, 0x5241);  // [g10 g11 b10 b11]
                // (S + 512) >> 10 == byte 2 of (64*S + 32768): no shifts, PRMT picks the byte
                const uint32_t wy6 = 2048u - fy6;
                const uint32_t Sr = wy6 * __dp4a(a0, wx, 0u) + (fy6 * __dp4a(a0, wxh, 0u) + 32768u);
                const uint32_t Sg = wy6 * __dp4a(g0, wx, 0u) + (fy6 * __dp4a(g1, wx, 0u) + 32768u);
                const uint32_t Sb = wy6 * __dp4a(g0, wxh, 0u) + (fy6 * __dp4a(g1, wxh, 0u) + 32768u);
                px = __byte_perm(__byte_perm(Sr, Sg, 0x4462), Sb, 0x4610);
            } else {
                px = crop_border_pixel(src, sh, sw, ss, X, Y);
            }
            o[0] = (uint8_t)px; o[1] = (uint8_t)(px >> 8); o[2] = (uint8_t)(px >> 16);
        }
    }
    __syncthreads();

    // tile -> global, 16-byte streaming stores (tile rows are contiguous in dst as well).
    uint8_t *out = dst + ((size_t)n * H + y_base) * row_bytes;
    const int total = rows * row_bytes;
    if ((row_bytes & 15) == 0) {
        const int4 *t4 = reinterpret_cast<const int4 *>(tile);
        int4 *o4 = reinterpret_cast<int4 *>(out);
        for (int i = threadIdx.x; i < total / 16; i += (int)blockDim.x) __stcs(o4 + i, t4[i]);
    } else {
        for (int i = threadIdx.x; i < total; i += (int)blockDim.x) out[i] = tile[i];
    }
}

}  // namespace udp

extern "C" int udp_warp_affine_u8(const uint8_t *const *src_ptrs, const int32_t *src_h,
                                  const int32_t *src_w, const int32_t *src_stride, const float *M,
                                  uint8_t *dst, int W, int H, int N, void *stream)
{
    if (N < 0 || W < 1 || H < 1 || (W & 1)) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!src_ptrs || !src_h || !src_w || !src_stride || !M || !dst) return UDP_EINVAL;
    size_t smem = sizeof(int32_t) * (2 * (size_t)W + 2 * udp::CROP_ROWS) + (size_t)udp::CROP_ROWS * W * 3;
    if (smem > 200 * 1024) return UDP_EINVAL;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(udp::warp_affine_u8_kernel,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
    }
    const int threads = W <= 96 ? 96 : (W <= 512 ? ((W + 31) / 32) * 32 : 512);
    for (int n0 = 0; n0 < N; n0 += 65535) {  // gridDim.y limit
        int nn = N - n0 < 65535 ? N - n0 : 65535;
        dim3 grid((H + udp::CROP_ROWS - 1) / udp::CROP_ROWS, nn);
        udp::warp_affine_u8_kernel<<<grid, threads, smem, (cudaStream_t)stream>>>(
            src_ptrs + n0, src_h + n0, src_w + n0, src_stride + n0, M + 6 * (size_t)n0,
            dst + (size_t)n0 * H * W * 3, W, H);
        UDP_LAUNCH_CHECK();
    }
    return 0;
}
