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
#include <cmath>
#include <cstring>
#include <mutex>
#include <type_traits>

#include "decode_pipe.cuh"
#include "tma_host.cuh"

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


// ---------------------------------------------------------------------------------------
// Second-generation crop: 32 x 32 output tiles gathered from a shared-memory copy of the
// tile's source footprint.
//
// ncu on the kernel above (profiles/r01_crop_v6_ncu.txt): the L1 data pipe is the limiter --
// 81 % of peak wavefronts, 4.3 wavefronts per global load request, because 60 % of the
// training-time boxes are rotated and a warp's 32 consecutive output pixels then walk
// diagonally through up to 32 source rows (one 128-byte line each, six loads per pixel), and
// every border pixel diverges into the out-of-line path.  Here:
//   * the footprint (axis-aligned bounding box of the tile's taps, exact because the Q10
//     column deltas and row bases are monotonic) is copied with 16-byte cp.async chunks that
//     keep the global 16-byte phase, so any image alignment / stride works; rows and chunks
//     outside the image buffer are zero-filled, and for tiles that cross the left/right image
//     edge a fix-up pass zeroes the out-of-image columns: the gather itself is branch-free and
//     OpenCV's constant-0 border falls out of the data;
//   * the gather reads shared memory, where only bank conflicts matter, not 128-byte lines;
//   * the bilinear weights go through two IDP.2A per channel (16-bit weight pairs wy*wx
//     against byte pairs) instead of dp4a + two IMADs;
//   * a warp's 32 pixels (96 bytes) are repacked with two shuffles and leave as 24 aligned
//     4-byte stores -- no output tile in shared memory, no second barrier.
// Arithmetic is the same integer formula as above: bit-exact with cv2.warpAffine.
// ---------------------------------------------------------------------------------------
constexpr int CT = 32;                       // tile edge (output pixels)
constexpr int CT_THREADS = 128;              // 4 warps, 8 tile rows each
#ifndef UDP_CROP_SMEM_KB
#define UDP_CROP_SMEM_KB 24
#endif
// footprint buffer (bytes), + 16 bytes slack; a template parameter of the tiled kernels.  34 KB (6 CTAs per SM, 0.5 %
// of the SURVEY 8d boxes miss the TMA box budget instead of 6 %) is 3.7 % faster up to 256x256 crops, 24 KB
// (8 CTAs per SM) 1 % faster at 384x288 (measured)
constexpr int CT_SMEM_LARGE_CROP = UDP_CROP_SMEM_KB * 1024;
#ifndef UDP_CROP_SMEM_SMALL_KB
#define UDP_CROP_SMEM_SMALL_KB 34
#endif
constexpr int CT_SMEM_SMALL_CROP = UDP_CROP_SMEM_SMALL_KB * 1024;

__device__ __forceinline__ void zero_bytes(uint8_t *base, int lo, int hi)
{
    while (lo < hi && (lo & 3)) base[lo++] = 0;
    while (lo + 4 <= hi) { *reinterpret_cast<uint32_t *>(base + lo) = 0u; lo += 4; }
    while (lo < hi) base[lo++] = 0;
}

struct TileGeom {
    int xa, ya, ncols, nrows, maxc;
};

__device__ __forceinline__ uint32_t lds_u32(unsigned addr)
{
    uint32_t v;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(addr));
    return v;
}

// One output pixel from the footprint copy.  K0 folds the buffer's shared-space address,
// shift0 - ya*pitch - 3*xa, so a tap address is sy*pitch + 3*sx + K0.
template <bool UNIFORM_PHASE>
__device__ __forceinline__ uint32_t crop_tile_pixel(int X, int Y, int pitch, int K0, int shift0, int ya, int ssm)
{
    const int sx = X >> 10, sy = Y >> 10;
    const uint32_t fx = ((uint32_t)X >> 5) & 31u, fy = ((uint32_t)Y >> 5) & 31u;
    int a0, a1;
    if (UNIFORM_PHASE) {  // stride % 16 == 0: every row has the phase shift0
        a0 = sy * pitch + (3 * sx + K0);
        a1 = a0 + pitch;
    } else {
        const int ry = sy - ya;
        const int s0 = (shift0 + ry * ssm) & 15, s1 = (s0 + ssm) & 15;
        a0 = sy * pitch + (3 * sx + K0) + (s0 - shift0);
        a1 = a0 + pitch + (s1 - s0);
    }
    // the buffer is 16-byte aligned in shared space, so (a & 3) is the byte phase of the tap
    const unsigned q0 = (unsigned)a0 & ~3u, q1 = (unsigned)a1 & ~3u;
    const uint32_t u0 = lds_u32(q0), u1 = lds_u32(q0 + 4), u2 = lds_u32(q0 + 8);
    const uint32_t v0 = lds_u32(q1), v1 = lds_u32(q1 + 4), v2 = lds_u32(q1 + 8);
    const uint32_t h0 = (uint32_t)a0 << 3, h1 = (uint32_t)a1 << 3;  // funnel shift uses bits 0..4
    const uint32_t lo0 = __funnelshift_r(u0, u1, h0), hi0 = __funnelshift_r(u1, u2, h0);
    const uint32_t lo1 = __funnelshift_r(v0, v1, h1), hi1 = __funnelshift_r(v1, v2, h1);
    const uint32_t c_r = __byte_perm(lo0, lo1, 0x7430);  // [r00 r01 r10 r11]
    const uint32_t c_t = __byte_perm(lo0, hi0, 0x5241);  // [g00 g01 b00 b01]
    const uint32_t c_b = __byte_perm(lo1, hi1, 0x5241);  // [g10 g11 b10 b11]
    const uint32_t wxp = fx * 0xffffu + 32u;             // (32 - fx) | fx << 16
    const uint32_t W0 = (32u - fy) * wxp, W1 = fy * wxp; // (wy*wx0 | wy*wx1 << 16), each <= 1024
    const uint32_t Sr = __dp2a_hi(W1, c_r, __dp2a_lo(W0, c_r, 512u));
    const uint32_t Sg = __dp2a_lo(W1, c_b, __dp2a_lo(W0, c_t, 512u));
    const uint32_t Sb = __dp2a_hi(W1, c_b, __dp2a_hi(W0, c_t, 512u));
    // (S + 512) >> 10 is byte 2 of S << 6
    return __byte_perm(__byte_perm(Sr << 6, Sg << 6, 0x4462), Sb << 6, 0x4610);
}

// The same pixel in two halves, so that a caller can put the loads of the next row in front of
// the arithmetic of the current one (uniform 16-byte phase only: pitch % 16 == 0).
#ifndef UDP_CROP_PRED3
#define UDP_CROP_PRED3 1
#endif
#ifndef UDP_CROP_AXIS
#define UDP_CROP_AXIS 1
#endif
#ifndef UDP_CROP_F32_FEWER
#define UDP_CROP_F32_FEWER 1
#endif
#ifndef UDP_CROP_LARGE_MINB
#define UDP_CROP_LARGE_MINB 6
#endif
#ifndef UDP_CROP_PITCHSEL
#define UDP_CROP_PITCHSEL 1
#endif
struct CropTaps {
    uint32_t u0, u1, u2, v0, v1, v2, h0, W0, W1;
};
__device__ __forceinline__ CropTaps crop_tile_load(int X, int Y, int pitch, int K0)
{
    CropTaps tp;
    const int sx = X >> 10, sy = Y >> 10;
    const uint32_t fx = ((uint32_t)X >> 5) & 31u, fy = ((uint32_t)Y >> 5) & 31u;
    const uint32_t wxp = fx * 0xffffu + 32u;  // (32 - fx) | fx << 16
    tp.W1 = fy * wxp;                         // (wy * wx0 | wy * wx1 << 16), each <= 1024
    tp.W0 = wxp * 32u - tp.W1;
    const int a0 = sy * pitch + (3 * sx + K0);
    const unsigned q0 = (unsigned)a0 & ~3u, q1 = q0 + (unsigned)pitch;  // pitch % 4 == 0: both rows share the byte phase
    tp.u0 = lds_u32(q0); tp.u1 = lds_u32(q0 + 4);
    tp.v0 = lds_u32(q1); tp.v1 = lds_u32(q1 + 4);
#if UDP_CROP_PRED3
    // the six bytes of a row reach into the third word only at byte phase 3 (a quarter of the lanes): the
    // other lanes stay out of that load, which then costs about one wavefront instead of two or three.
    // (Their u2 / v2 keep whatever the registers held: only bytes the funnel shift never selects.)
    asm volatile("{\n.reg .pred p;\nsetp.eq.u32 p, %2, 3;\n@p ld.shared.u32 %0, [%3+8];\n@p ld.shared.u32 %1, [%4+8];\n}"
                 : "+r"(tp.u2), "+r"(tp.v2) : "r"((unsigned)a0 & 3u), "r"(q0), "r"(q1));
#else
    tp.u2 = lds_u32(q0 + 8); tp.v2 = lds_u32(q1 + 8);
#endif
    tp.h0 = (uint32_t)a0 << 3;
    return tp;
}
__device__ __forceinline__ uint32_t crop_tile_finish(const CropTaps &tp)
{
    const uint32_t lo0 = __funnelshift_r(tp.u0, tp.u1, tp.h0), hi0 = __funnelshift_r(tp.u1, tp.u2, tp.h0);
    const uint32_t lo1 = __funnelshift_r(tp.v0, tp.v1, tp.h0), hi1 = __funnelshift_r(tp.v1, tp.v2, tp.h0);
    const uint32_t c_r = __byte_perm(lo0, lo1, 0x7430);  // [r00 r01 r10 r11]
    const uint32_t c_t = __byte_perm(lo0, hi0, 0x5241);  // [g00 g01 b00 b01]
    const uint32_t c_b = __byte_perm(lo1, hi1, 0x5241);  // [g10 g11 b10 b11]
    const uint32_t Sr = __dp2a_hi(tp.W1, c_r, __dp2a_lo(tp.W0, c_r, 512u));
    const uint32_t Sg = __dp2a_lo(tp.W1, c_b, __dp2a_lo(tp.W0, c_t, 512u));
    const uint32_t Sb = __dp2a_hi(tp.W1, c_b, __dp2a_hi(tp.W0, c_t, 512u));
    return __byte_perm(__byte_perm(Sr << 6, Sg << 6, 0x4462), Sb << 6, 0x4610);
}

// Can the footprints of this person's tiles be fetched by TMA (warp_affine_u8_tma_kernel below)?
// Tiles are 32 columns x tr rows (tr = 32, 16 or 8: the tallest whose box fits half the buffer, so
// two boxes can be in flight).  The box is the same for every tile of the person -- a bound on
// the footprint of any tile.  The Q10 column deltas are rint(m00*x*1024) (products exact in
// float64) and the row bases rint((m01*y+m02)*1024)+16, both monotonic, so over a tile
//   span(X) <= |m00|*31*1024 + |m01|*(tr-1)*1024 + 2 (+ 2^-20 from the one rounded addition)
// and the tap columns number (Xmax>>10) + 2 - (Xmin>>10) <= (span >> 10) + 3; rows likewise.
#ifndef UDP_CROP_NBOX
#define UDP_CROP_NBOX 2
#endif
constexpr int CT_NBOX = UDP_CROP_NBOX;                   // TMA boxes in flight per CTA
__host__ __device__ constexpr int ct_box(int smem) { return (smem / CT_NBOX) & ~127; }  // one box
// bytes between the boxes and colc[]: 128 (descriptor) + 8 * CT_NBOX (mbarriers) + 8 (pitch) + 8 * CT_NBOX (cring), 16-aligned
constexpr int CT_TAIL = (128 + 8 * CT_NBOX + 8 + 8 * CT_NBOX + 15) & ~15;
struct CropPlan {
    int lead, boxw, boxh, tr, boxw_min;
    bool tma, axis;
};
// axis: no rotation (every evaluation-time box, 40 % of the training-time ones): the source column depends
// on the output column only and the source row on the output row only -- see crop_gather_axis.
__device__ __forceinline__ bool crop_axis_aligned(const float *m) { return m[1] == 0.0f && m[3] == 0.0f; }
__device__ __forceinline__ CropPlan crop_plan(const uint8_t *src, int sh, int sw, int ss, const float *m, int box_bytes)
{
    CropPlan p;
    p.lead = (int)((uintptr_t)src & 15);
    p.boxw = p.boxh = p.tr = p.boxw_min = 0;
    p.tma = false;
    p.axis = crop_axis_aligned(m);
    const double a00 = fabs((double)m[0]) * 1024.0, a01 = fabs((double)m[1]) * 1024.0;
    const double a10 = fabs((double)m[3]) * 1024.0, a11 = fabs((double)m[4]) * 1024.0;
    // (also NaN) tame coefficients only: Q10 coordinates stay far from int32 overflow, boxes below 256
    if (!(a00 < 8192.0 && a01 < 8192.0 && a10 < 8192.0 && a11 < 8192.0 && fabsf(m[2]) < 1048576.0f && fabsf(m[5]) < 1048576.0f))
        return p;
    if (!(sh > 0 && sw > 0 && ss >= 16 && (ss & 15) == 0)) return p;
#pragma unroll
    for (int tr = CT; tr >= 8; tr >>= 1) {
        const int ex = (int)(a00 * 31.0 + a01 * (double)(tr - 1)), ey = (int)(a10 * 31.0 + a11 * (double)(tr - 1));
        const int bmin = (3 * (((ex + 4) >> 10) + 3) + 15 + 15) & ~15;  // + 15: the box starts on a 16-byte boundary
        const int boxh = ((ey + 4) >> 10) + 3;
        if (bmin <= 512 && boxh <= 256 && bmin * boxh <= box_bytes) {  // 256 two-byte elements per box row at most
            int boxw = bmin;
            // rotated gathers walk across the rows of the box: an odd number of 16-byte chunks per row spreads
            // them over the banks (the default; the TMA kernel may pick another width, crop_pick_pitch)
            if (!p.axis && !(boxw & 16) && boxw + 16 <= 512 && (boxw + 16) * boxh <= box_bytes) boxw += 16;
            p.boxw = boxw; p.boxw_min = bmin; p.boxh = boxh; p.tr = tr; p.tma = true;
            break;
        }
    }
    return p;
}

__device__ __forceinline__ int q10_col(double m, int x) { return __double2int_rn(__dmul_rn(__dmul_rn(m, (double)x), 1024.0)); }
__device__ __forceinline__ int q10_row(double m, double c, int y)
{
    return __double2int_rn(__dmul_rn(__dadd_rn(__dmul_rn(m, (double)y), c), 1024.0)) + 16;
}

// ---------------------------------------------------------------------------------------
// Where a finished row of 32 pixels goes.  Both tiled kernels are templated on one of these:
//   CropOutU8  -- the crop itself: uint8 HWC (a2), a warp's 96 bytes repacked into 24 word stores;
//   CropOutF32 -- SURVEY 8(f) rank 1: ToTensor + Normalize fused into the crop
//                 (tools/train.py:127-143, JointsDataset.py:237-238): float32 CHW
//                 ((x / 255) - mean) / std, plus (optionally) the x-mirrored copy that the flip
//                 test feeds the model (np.flip(input, 3), core/function.py:154-155).  Three
//                 (six) fully coalesced 128-byte stores per warp row; the uint8 crop never
//                 exists in memory.
// Each lane owns one pixel of the row; a warp's next row is 4 below (CT_THREADS / 32).
// ---------------------------------------------------------------------------------------
#ifndef UDP_CROP_U8_ROWMAJOR
#define UDP_CROP_U8_ROWMAJOR false
#endif
struct CropU8Args {
    uint8_t *dst;
};
struct CropOutU8 {
    using Args = CropU8Args;
    static constexpr bool ROW_MAJOR = UDP_CROP_U8_ROWMAJOR;
    static constexpr bool IS_F32 = false;
    uint8_t *base, *p;
    size_t row_bytes, step;
    uint32_t osel;
    bool act;
    __device__ __forceinline__ void init(const Args &a, int n, int W, int H, int lane)
    {
        // A warp's row of 32 pixels is 96 bytes = 24 aligned words.  Word j starts inside pixel floor(4j/3):
        // pixels 0, 1, 2 of every group of four hold the first byte of one word each (at byte 0, 1, 2 of the
        // pixel), pixel 3 of none.  So lane l (l % 4 != 3) fetches its right neighbour's pixel with ONE
        // shuffle, cuts word j = l - l/4 out of the two and stores it: 24 lanes, 96 contiguous bytes.
        osel = (lane % 4 == 0) ? 0x4210u : ((lane % 4 == 1) ? 0x5421u : 0x6542u);
        asm volatile("" : "+r"(osel));  // keep it in a register (ptxas rematerialised it per pixel)
        row_bytes = (size_t)W * 3;
        step = (size_t)(CT_THREADS / 32) * row_bytes;
        base = a.dst + (size_t)n * H * row_bytes + 4 * (lane - lane / 4);
        p = base;
        act = lane % 4 != 3;
    }
    __device__ __forceinline__ void seek(int y, int x0)
    {
        p = base + (unsigned)(y * (int)row_bytes + x0 * 3);
        asm volatile("" : "+l"(p));  // a register pair for the tile (ptxas re-added the base from the constant bank per row: 2 %)
    }
    __device__ __forceinline__ void emit(uint32_t px)
    {
        const uint32_t nx = __shfl_down_sync(0xffffffffu, px, 1);
        if (act) *reinterpret_cast<uint32_t *>(p) = __byte_perm(px, nx, osel);
        p += step;
    }
    // row i * ROWSTEP below the row of the last seek(); with the crop width known at compile time (WC) the
    // offset is an immediate of the store (the running 64-bit pointer cost four instructions per row)
    template <int ROWSTEP, int WC>
    __device__ __forceinline__ void emit_at(uint32_t px, int i)
    {
        const uint32_t nx = __shfl_down_sync(0xffffffffu, px, 1);
        const size_t off = WC ? (size_t)(i * ROWSTEP * WC * 3) : (size_t)(i * ROWSTEP) * row_bytes;
        if (act) *reinterpret_cast<uint32_t *>(p + off) = __byte_perm(px, nx, osel);
    }
};

struct CropF32Args {
    float *dst, *dst_flipped;  // [N,3,H,W]; dst_flipped may be null
    float mean[3], sd[3], inv[3];
    int exact_div;             // 1: the reciprocal + one correction was checked on the host for all 256 inputs
};
// ((x / 255) - mean) / sd in float32, each step rounded as torch does (true divisions).
__device__ __forceinline__ float crop_norm(uint32_t px, int c, float mean, float sd, float inv, bool fast)
{
    const float x = __fsub_rn(__uint_as_float(__byte_perm(px, 0x4B000000u, 0x7440u | (uint32_t)c)), 8388608.0f);  // byte c, exact
    const float y = 0x1.010102p-8f;
    const float q0 = __fmul_rn(x, y);
    const float q = fmaf(fmaf(-q0, 255.0f, x), y, q0);  // x / 255 correctly rounded (div255; x >= 0)
    const float t = __fsub_rn(q, mean);
    if (!fast) return __fdiv_rn(t, sd);
    const float u = __fmul_rn(t, inv);
    return fmaf(fmaf(-u, sd, t), inv, u);
}
struct CropOutF32 {
    using Args = CropF32Args;
    static constexpr bool ROW_MAJOR = false;
    static constexpr bool IS_F32 = true;
    float *base, *basef, *p, *pf;
    size_t plane, step;
    int W_;
    f32x2 nmean01, nsd01, inv01;  // channels 0 and 1 ride the packed FP32 pipe
    float mean2, sd2, inv2;
    bool fast;
    __device__ __forceinline__ void init(const Args &a, int n, int W, int H, int lane)
    {
        plane = (size_t)H * W;
        step = (size_t)(CT_THREADS / 32) * W;
        W_ = W;
        base = a.dst + (size_t)n * 3 * plane + lane;
        basef = a.dst_flipped ? a.dst_flipped + (size_t)n * 3 * plane + (W - 1 - lane) : nullptr;
        p = base;
        pf = basef;
        nmean01 = pk(-a.mean[0], -a.mean[1]);
        nsd01 = pk(-a.sd[0], -a.sd[1]);
        inv01 = pk(a.inv[0], a.inv[1]);
        mean2 = a.mean[2]; sd2 = a.sd[2]; inv2 = a.inv[2];
        fast = a.exact_div != 0;
    }
    __device__ __forceinline__ void seek(int y, int x0)
    {
        p = base + (y * W_ + x0);
        if (basef) pf = basef + (y * W_ - x0);
    }
    __device__ __forceinline__ void norm3(uint32_t px, float &v0, float &v1, float &v2) const
    {
        v2 = crop_norm(px, 2, mean2, sd2, inv2, fast);
        if (fast) {  // the same operations as crop_norm, two channels per instruction
            const float y = 0x1.010102p-8f;
            const f32x2 x = fsub2p(pk(__uint_as_float(__byte_perm(px, 0x4B000000u, 0x7440u)),
                                      __uint_as_float(__byte_perm(px, 0x4B000000u, 0x7441u))), pk(8388608.0f, 8388608.0f));
            const f32x2 q0 = fmul2p(x, y);
            const f32x2 q = ffma2p(ffma2p(q0, -255.0f, x), y, q0);
            const f32x2 t = fadd2p(q, nmean01);
            const f32x2 u = fmul2pp(t, inv01);
            const f32x2 o = ffma2pp(ffma2pp(u, nsd01, t), inv01, u);
            v0 = lo_of(o);
            v1 = hi_of(o);
        } else {
            v0 = crop_norm(px, 0, -lo_of(nmean01), -lo_of(nsd01), lo_of(inv01), false);
            v1 = crop_norm(px, 1, -hi_of(nmean01), -hi_of(nsd01), hi_of(inv01), false);
        }
    }
    __device__ __forceinline__ void emit(uint32_t px)
    {
        float v0, v1, v2;
        norm3(px, v0, v1, v2);
        __stcs(p, v0); __stcs(p + plane, v1); __stcs(p + 2 * plane, v2);
        p += step;
        if (basef) {
            __stcs(pf, v0); __stcs(pf + plane, v1); __stcs(pf + 2 * plane, v2);
            pf += step;
        }
    }
    template <int ROWSTEP, int WC>
    __device__ __forceinline__ void emit_at(uint32_t px, int i)
    {
        float v0, v1, v2;
        norm3(px, v0, v1, v2);
        const size_t off = WC ? (size_t)(i * ROWSTEP * WC) : (size_t)(i * ROWSTEP) * W_;
        __stcs(p + off, v0); __stcs(p + off + plane, v1); __stcs(p + off + 2 * plane, v2);
        if (basef) { __stcs(pf + off, v0); __stcs(pf + off + plane, v1); __stcs(pf + off + 2 * plane, v2); }
    }
};

__host__ __device__ inline void crop_args_advance(CropU8Args &a, size_t elems) { a.dst += elems; }
__host__ __device__ inline void crop_args_advance(CropF32Args &a, size_t elems)
{
    a.dst += elems;
    if (a.dst_flipped) a.dst_flipped += elems;
}

// grid = (tile columns, persons); a CTA walks down its tile column, so everything that depends
// only on the person and the column (matrix, Q10 column deltas, image geometry) is computed once.
template <class OUT, int CT_SMEM>
__global__ void __launch_bounds__(CT_THREADS)
warp_affine_u8_tile_kernel(const uint8_t *const *src_ptrs, const int32_t *src_h, const int32_t *src_w,
                           const int32_t *src_stride, const float *M, typename OUT::Args oargs, int W, int H, int skip_tma)
{
    extern __shared__ __align__(16) unsigned char fp[];  // footprint copy
    const unsigned fpa = (unsigned)__cvta_generic_to_shared(fp);

    const int n = blockIdx.y, t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const int x0 = blockIdx.x * CT;
    const float *m = M + 6 * (size_t)n;
    const double m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5];
    const int adx = q10_col(m00, x0 + lane), bdx = q10_col(m10, x0 + lane);  // Q10 deltas of this lane's column
    const int A0 = __shfl_sync(0xffffffffu, adx, 0), A1 = __shfl_sync(0xffffffffu, adx, CT - 1);
    const int B0 = __shfl_sync(0xffffffffu, bdx, 0), B1 = __shfl_sync(0xffffffffu, bdx, CT - 1);
    const int Amin = min(A0, A1), Amax = max(A0, A1), Bmin = min(B0, B1), Bmax = max(B0, B1);

    const uint8_t *src = src_ptrs[n];
    const int sh = src_h[n], sw = src_w[n], ss = src_stride[n];
    if (skip_tma && crop_plan(src, sh, sw, ss, m, ct_box(CT_SMEM)).tma) return;  // the TMA kernel does this person
    const int ssm = ss & 15;
    // the image buffer as 16-byte chunks: [buf, buf + lim + 16) may be read (images are < 2 GiB)
    const char *buf = reinterpret_cast<const char *>((uintptr_t)src & ~(uintptr_t)15);
    const int lead = (int)((uintptr_t)src & 15);
    const bool img_ok = sh > 0 && sw > 0;
    const unsigned lim = img_ok ? (unsigned)((((long long)(sh - 1) * ss + 3LL * sw + lead + 15) & ~15LL) - 16) : 0u;
    OUT out;
    out.init(oargs, n, W, H, lane);

    // Q10 row bases of the whole tile column, once per CTA (float64 products, as OpenCV rounds them)
    int2 *rowq = reinterpret_cast<int2 *>(fp + CT_SMEM + 16);
    for (int y = t; y < H; y += CT_THREADS) rowq[y] = make_int2(q10_row(m01, m02, y), q10_row(m11, m12, y));
    __syncthreads();

    auto bbox = [&](int2 qa, int2 qb) {  // footprint of the tile rows whose first / last bases are given
        TileGeom g;
        g.xa = (min(qa.x, qb.x) + Amin) >> 10;
        g.ya = (min(qa.y, qb.y) + Bmin) >> 10;
        g.ncols = ((max(qa.x, qb.x) + Amax) >> 10) + 2 - g.xa;
        g.nrows = ((max(qa.y, qb.y) + Bmax) >> 10) + 2 - g.ya;
        g.maxc = ((30 + 3 * g.ncols) >> 4) | 1;  // 16-byte chunks per row at worst-case phase; odd count
        return g;
    };

    for (int y0 = 0; y0 < H; y0 += CT) {
        auto bbox_rows = [&](int ry0, int rows) { return bbox(rowq[y0 + ry0], rowq[y0 + ry0 + rows - 1]); };
        TileGeom g = bbox_rows(0, CT);
        // rows per pass: the whole tile unless its footprint does not fit the buffer (large scale)
        int nr = CT;
        if (g.nrows * (g.maxc << 4) > CT_SMEM) {
            for (nr = CT / 2; nr > 1; nr >>= 1) {
                bool fits = true;
                for (int ry0 = 0; ry0 < CT; ry0 += nr) {
                    const TileGeom h = bbox_rows(ry0, nr);
                    if ((long long)h.nrows * (h.maxc << 4) > CT_SMEM) { fits = false; break; }
                }
                if (fits) break;
            }
        }

        for (int ry0 = 0; ry0 < CT; ry0 += nr) {
            if (nr != CT) g = bbox_rows(ry0, nr);
            const int pitch = g.maxc << 4;
            const bool direct = (long long)g.nrows * pitch > CT_SMEM;  // absurd scale: per-pixel global path
            // 16-byte phase of pixel (xa, ya); row r of the footprint has phase (shift0 + r*ssm) & 15
            const int base_off = lead + 3 * g.xa;  // byte offset of pixel (xa, row 0) from buf
            const int shift0 = (g.ya * ssm + base_off) & 15;
            if (!direct) {
                // ---- stage the footprint with 16-byte cp.async chunks that keep the global phase.
                // nch chunks per row are copied (exact when every row has the same phase); a chunk is
                // copied iff it lies inside the image buffer, else zero-filled.
                const int nch = ((ssm == 0 ? shift0 : 15) + 3 * g.ncols + 15) >> 4;
                // a row goes to a group of G = 2^lg lanes (one chunk per lane), 32/G rows per warp per
                // turn: no index division, and every lane's pointers advance by constants
                const int lg = nch > 16 ? 5 : 32 - __clz(nch - 1);  // smallest G >= nch, G <= 32
                const int c0 = lane & ((1 << lg) - 1), r0 = (warp << (5 - lg)) + (lane >> lg);
                const int rstep = (CT_THREADS / 32) << (5 - lg);
                const bool rows_in = g.ya >= 0 && g.ya + g.nrows <= sh;
                const int off0 = g.ya * ss + base_off;  // pixel (xa, ya); fits: images are < 2 GiB
                if (ssm == 0 && rows_in && off0 - 15 >= 0 && (unsigned)(off0 + (g.nrows - 1) * ss + pitch - 16) <= lim) {
                    // every chunk is in the buffer and every row has phase shift0: pure pointer stepping
                    for (int c = c0; c < nch; c += 32) {
                        const char *gp = buf + (off0 + r0 * ss - shift0 + (c << 4));
                        unsigned d = fpa + r0 * pitch + (c << 4);
                        const int gstep = rstep * ss, dstep = rstep * pitch;
                        for (int r = r0; r < g.nrows; r += rstep, gp += gstep, d += dstep)
                            asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gp) : "memory");
                    }
                } else {
                    for (int c = c0; c < nch; c += 32) {
                        for (int r = r0; r < g.nrows; r += rstep) {
                            const int y = g.ya + r;
                            const int off = y * ss + base_off - ((shift0 + r * ssm) & 15) + (c << 4);
                            const unsigned d = fpa + r * pitch + (c << 4);
                            if ((unsigned)y < (unsigned)sh && (unsigned)off <= lim && img_ok)
                                asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(buf + off) : "memory");
                            else
                                asm volatile("st.shared.v4.b32 [%0], {%1, %1, %1, %1};\n" ::"r"(d), "r"(0) : "memory");
                        }
                    }
                }
                cp_async_commit();
                cp_async_wait_all();
                __syncthreads();
                if (g.xa < 0 || g.xa + g.ncols > sw) {  // zero the columns left / right of the image
                    for (int rr = t; rr < g.nrows; rr += CT_THREADS) {
                        const int y = g.ya + rr;
                        if (y < 0 || y >= sh) continue;
                        const int shr = (shift0 + rr * ssm) & 15;
                        uint8_t *base = fp + rr * pitch;
                        if (g.xa < 0) zero_bytes(base, shr, shr + 3 * min(g.ncols, -g.xa));
                        if (g.xa + g.ncols > sw) zero_bytes(base, shr + 3 * max(0, sw - g.xa), shr + 3 * g.ncols);
                    }
                    __syncthreads();
                }
            }
            const int K0 = (int)fpa + shift0 - g.ya * pitch - 3 * g.xa;
            out.seek(y0 + ry0 + warp, x0);
            // rows of this pass owned by this warp: ry0 + warp + 4*i; their Q10 bases come from rowq
#define UDP_CROP_ROWS(PIXEL_EXPR)                                                                     \
            for (int r = ry0 + warp; r < ry0 + nr; r += CT_THREADS / 32) {                            \
                const int2 q = rowq[y0 + r];                                                          \
                const int X = q.x + adx, Y = q.y + bdx;                                               \
                out.emit(PIXEL_EXPR);                                                                 \
            }
            if (direct) {  // absurd scale: per-pixel global path
                UDP_CROP_ROWS(crop_border_pixel(src, sh, sw, ss, X, Y))
            } else if (ssm == 0) {
#pragma unroll 2
                UDP_CROP_ROWS(crop_tile_pixel<true>(X, Y, pitch, K0, shift0, g.ya, 0))
            } else {
#pragma unroll 2
                UDP_CROP_ROWS(crop_tile_pixel<false>(X, Y, pitch, K0, shift0, g.ya, ssm))
            }
#undef UDP_CROP_ROWS
            __syncthreads();  // the next pass / tile overwrites the footprint buffer
        }
    }
}

// ---------------------------------------------------------------------------------------
// Third generation: the footprint of a tile arrives with ONE TMA tensor load.
//
// The image is described to the TMA unit as a 2-D tensor of two-byte elements (a box row may then be
// 512 bytes: scales up to 5 source pixels per output pixel) of 3*sw + lead bytes x sh rows, row
// pitch = stride, based at the 16-byte-aligned address below the first pixel; the box is the
// person's footprint bound (crop_plan).  Box coordinates are arbitrary byte / row positions,
// positions (the byte start rounded down to 16: the TMA unit faults on anything else), and
// whatever falls outside the tensor is filled with zeros by the hardware -- OpenCV's
// constant-0 border with no bounds logic at all.  One CTA = one person: it patches a template
// descriptor in shared memory (tensormap.replace: address, extents, pitch, box), publishes it
// to a slot of a device-global ring (tensormap.cp_fenceproxy) and walks its tiles: thread 0
// issues the load, everybody waits on the mbarrier, gathers (same integer pixel code as the
// cp.async kernel, uniform phase 0) and meets at one barrier per tile.
// Needs stride % 16 == 0 and a box that fits; other persons are left to the cp.async kernel.
// ---------------------------------------------------------------------------------------
// Pool of descriptors in global memory (the library allocates nothing at run time).  A CTA claims
// a free slot with one compare-and-swap, starting at a ticket so that claims rarely collide, and
// releases it after its last load has landed; 8192 slots against at most a few thousand resident
// CTAs (all streams together), so the first probe almost always succeeds.
constexpr int TM_SLOTS = 8192;
__device__ __align__(128) unsigned char g_tmap_ring[TM_SLOTS][128];
__device__ unsigned int g_tmap_next;
__device__ unsigned int g_tmap_busy[TM_SLOTS];

// ROWS rows (every fourth of a tile) of one warp: the taps of row i+1 are loaded before the
// arithmetic, shuffles and store of row i.
template <int ROWS, int WC, class OUT>
__device__ __forceinline__ void crop_gather_rows(const int2 *rq, int adx, int bdx, int pitch, int K0, OUT &out)
{
    // The float32 policy's emit is ~100 instructions: unrolled into every row of every tile shape it overflowed the
    // instruction cache (`no_instruction` stalls at 5.8 per issue, 3.0 -> 6.0 ms): its row loops stay rolled.
    if constexpr (OUT::IS_F32) {
        int2 q0 = rq[0];
        CropTaps cur = crop_tile_load(q0.x + adx, q0.y + bdx, pitch, K0);
#pragma unroll 1
        for (int i = 0; i < ROWS; ++i) {
            CropTaps nxt = cur;
            if (i + 1 < ROWS) {
                const int2 q = rq[4 * (i + 1)];
                nxt = crop_tile_load(q.x + adx, q.y + bdx, pitch, K0);
            }
            out.template emit_at<CT_THREADS / 32, WC>(crop_tile_finish(cur), i);
            cur = nxt;
        }
    } else {
        int2 q[ROWS];
#pragma unroll
        for (int i = 0; i < ROWS; ++i) q[i] = rq[4 * i];
        CropTaps cur = crop_tile_load(q[0].x + adx, q[0].y + bdx, pitch, K0);
#pragma unroll
        for (int i = 0; i < ROWS; ++i) {
            CropTaps nxt;
            if (i + 1 < ROWS) nxt = crop_tile_load(q[i + 1].x + adx, q[i + 1].y + bdx, pitch, K0);
            out.template emit_at<CT_THREADS / 32, WC>(crop_tile_finish(cur), i);
            if (i + 1 < ROWS) cur = nxt;
        }
    }
}

// ---------------------------------------------------------------------------------------
// Axis-aligned persons (rotation 0: m01 == m10 == 0).  The four weights of a pixel factor into a column
// part and a row part, S = wy0 * h(sy, x) + wy1 * h(sy + 1, x) with h(r, x) = wx0 * p[r][sx] + wx1 * p[r][sx + 1]
// (exact integers, so the result is the same bit for bit): sx, wx and with them the byte offset, the byte
// phase and every selector are constants of the lane, sy and wy are the same for the whole warp.  A warp
// walks DOWN its rows and keeps h of the last source row: a source row is loaded and blended horizontally
// once (3 shared loads, the third one only on the lanes at byte phase 3, 2 funnel shifts, 2 PRMT, 3 IDP.2A)
// however many output rows use it -- 1.7 source rows per output row at the mean scale of the SURVEY 8d
// boxes instead of 2, and half as many instructions per pixel as the general gather.
// ---------------------------------------------------------------------------------------
struct CropHRow {
    uint32_t r, g, b;
};
__device__ __forceinline__ CropHRow crop_axis_hrow(unsigned addr, bool phi3, uint32_t sh8, uint32_t wxp)
{
    const uint32_t u0 = lds_u32(addr), u1 = lds_u32(addr + 4);
    uint32_t u2;  // deliberately not initialised: lanes not at byte phase 3 never select a byte of it
    asm volatile("{\n.reg .pred p;\nsetp.ne.u32 p, %1, 0;\n@p ld.shared.u32 %0, [%2+8];\n}" : "+r"(u2) : "r"((unsigned)phi3), "r"(addr));
    const uint32_t lo = __funnelshift_r(u0, u1, sh8), hi = __funnelshift_r(u1, u2, sh8);  // [r0 g0 b0 r1] [g1 b1 . .]
    const uint32_t c_rg = __byte_perm(lo, hi, 0x4130);  // [r0 r1 g0 g1]
    const uint32_t c_b = __byte_perm(lo, hi, 0x5252);   // [b0 b1 b0 b1]
    CropHRow h;
    h.r = __dp2a_lo(wxp, c_rg, 0u);
    h.g = __dp2a_hi(wxp, c_rg, 0u);
    h.b = __dp2a_lo(wxp, c_b, 0u);
    return h;
}
// RW consecutive output rows of one warp.  rq: their Q10 row bases (y component; bdx is 0 for every lane);
// base: shared address of this lane's first word in source row 0 (the box's rows are `pitch` apart).
// rq[i] = (byte offset of the top source row from the box origin's row 0, i.e. sy * pitch; Q10 row coordinate).
template <int RW, int WC, class OUT>
__device__ __forceinline__ void crop_gather_axis(const int2 *rq, unsigned base, int pitch, bool phi3, uint32_t sh8,
                                                 uint32_t wxp, OUT &out)
{
    int2 qn = rq[0];
    int cur = qn.x - pitch;  // offset of the source row held in hb
    CropHRow ht = {0u, 0u, 0u}, hb = {0u, 0u, 0u};
    constexpr int UNR = OUT::IS_F32 ? 1 : RW;
#pragma unroll UNR
    for (int i = 0; i < RW; ++i) {
        const int2 qc = qn;
        if (i + 1 < RW) qn = rq[i + 1];  // one row ahead (all of them at once cost registers: spills in this loop)
        const int top = qc.x, bot = top + pitch;
        if (bot != cur) {  // (equal: enlargement, both rows are already here)
            if (top == cur) ht = hb;
            else ht = crop_axis_hrow(base + (unsigned)top, phi3, sh8, wxp);
            hb = crop_axis_hrow(base + (unsigned)bot, phi3, sh8, wxp);
            cur = bot;
        }
        // (S + 512) >> 10 is byte 2 of 64 * S + 32768, S = (32 - fy) * ht + fy * hb <= 32 * 8160
        const uint32_t w1 = ((uint32_t)qc.y << 1) & 0x7c0u, w0 = 2048u - w1;  // fy * 64
        const uint32_t Sr = w1 * hb.r + (w0 * ht.r + 32768u);
        const uint32_t Sg = w1 * hb.g + (w0 * ht.g + 32768u);
        const uint32_t Sb = w1 * hb.b + (w0 * ht.b + 32768u);
        out.template emit_at<1, WC>(__byte_perm(__byte_perm(Sr, Sg, 0x4462), Sb, 0x4610), i);
    }
}

// Rotated persons: a warp's 32 taps lie on a slanted line through the box, lane i at (row, byte) =
// i * (m10, 3 * m00), so which banks they hit depends on the row pitch.  The pitch may be any multiple of 16
// bytes that still fits the buffer: warp 0 tries them on four sample rows (wavefronts of one load = the
// largest number of distinct words in one bank; both tap rows and all three words of a row follow the same
// pattern) and keeps the best.  Offline model over the SURVEY 8d boxes: 2.7 wavefronts per load with the
// "odd number of chunks" rule, 2.1 with the pick.
__device__ __forceinline__ int crop_pick_pitch(const CropPlan &pl, const float *m, int box_bytes, int lane)
{
    const double m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5];
    const int adx = q10_col(m00, lane), bdx = q10_col(m10, lane);
    int sx[4], sy[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        const int y = 9 * k;
        sx[k] = (q10_row(m01, m02, y) + adx) >> 10;
        sy[k] = (q10_row(m11, m12, y) + bdx) >> 10;
    }
    int best = pl.boxw, best_wf = 1 << 30;
    for (int pw = pl.boxw_min; pw <= 512 && pw * pl.boxh <= box_bytes && pw < pl.boxw_min + 128; pw += 16) {
        int wf = 0;
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            const int word = (sy[k] * pw + 3 * sx[k]) >> 2, bank = word & 31;
            const unsigned mb = __match_any_sync(0xffffffffu, bank), mw = __match_any_sync(0xffffffffu, word);
            const unsigned leaders = __ballot_sync(0xffffffffu, lane == __ffs(mw) - 1);
            wf += __reduce_max_sync(0xffffffffu, __popc(mb & leaders));
        }
        if (wf < best_wf) { best_wf = wf; best = pw; }
    }
    return best;
}

template <class OUT, int CT_SMEM, int WC>
__global__ void __launch_bounds__(CT_THREADS, ((CT_SMEM > 30 * 1024) ? 6 : UDP_CROP_LARGE_MINB) - (OUT::IS_F32 ? UDP_CROP_F32_FEWER : 0))  // CTAs per SM: footprint buffers, registers
warp_affine_u8_tma_kernel(const __grid_constant__ CUtensorMap tmpl, const uint8_t *const *src_ptrs,
                          const int32_t *src_h, const int32_t *src_w, const int32_t *src_stride,
                          const float *M, typename OUT::Args oargs, int W_, int H)
{
    const int W = WC ? WC : W_;  // WC: the crop width as a compile-time constant (0: any width)
    // CT_NBOX boxes of CT_BOX bytes | 128 B descriptor | mbarriers | pitch | int2 cring[CT_NBOX] | int4 colc[W/32] | int2 rowq[H]
    // dynamic shared memory is only guaranteed 16-byte alignment; TMA destinations need 128
    extern __shared__ __align__(16) unsigned char fp_raw[];
    unsigned char *fp = fp_raw + ((128u - ((unsigned)__cvta_generic_to_shared(fp_raw) & 127u)) & 127u);  // still a shared-space pointer
    const unsigned fpa = (unsigned)__cvta_generic_to_shared(fp);
    const unsigned sdesc = fpa + CT_SMEM, mbar = sdesc + 128;
    const int tiles_x = W / CT;
    int4 *colc = reinterpret_cast<int4 *>(fp + CT_SMEM + CT_TAIL);  // per tile column: min/max of the Q10 column deltas
    int2 *rowq = reinterpret_cast<int2 *>(fp + CT_SMEM + CT_TAIL + 16 * tiles_x);

    const int n = blockIdx.x, t = threadIdx.x, lane = t & 31, warp = t >> 5;
    const float *m = M + 6 * (size_t)n;
    const uint8_t *src = src_ptrs[n];
    const int sh = src_h[n], sw = src_w[n], ss = src_stride[n];
    constexpr int CT_BOX = ct_box(CT_SMEM);
    const CropPlan pl = crop_plan(src, sh, sw, ss, m, CT_BOX);
    if (!pl.tma) return;
    const double m00 = m[0], m01 = m[1], m02 = m[2], m10 = m[3], m11 = m[4], m12 = m[5];
    const int lead = pl.lead, TR = pl.tr;
    const bool axis = UDP_CROP_AXIS && pl.axis;
    int pitch = pl.boxw;
    volatile int *pitch_sh = reinterpret_cast<volatile int *>(fp + CT_SMEM + 128 + 8 * CT_NBOX);  // after the mbarriers
    // corners (first tap column / row) of the tiles whose boxes are in flight, one per box: written by warp 0
    // when it issues the load, read by everybody once the tile's turn has come
    int2 *cring = reinterpret_cast<int2 *>(fp + CT_SMEM + 128 + 8 * CT_NBOX + 8);

    // Q10 row bases; axis-aligned persons (whose X base is the same for every row: Xc) keep the byte offset
    // of the row's top source row there instead
    const int Xc = q10_row(m01, m02, 0);
    for (int y = t; y < H; y += CT_THREADS) {
        const int qy = q10_row(m11, m12, y);
        rowq[y] = make_int2(axis ? (qy >> 10) * pitch : q10_row(m01, m02, y), qy);
    }
    for (int c = t; c < tiles_x; c += CT_THREADS) {
        const int A0 = q10_col(m00, c * CT), A1 = q10_col(m00, c * CT + CT - 1);
        const int B0 = q10_col(m10, c * CT), B1 = q10_col(m10, c * CT + CT - 1);
        colc[c] = make_int4(min(A0, A1), max(A0, A1), min(B0, B1), max(B0, B1));
    }
    unsigned long long gdesc = 0;
    if (warp == 0) {
        reinterpret_cast<uint32_t *>(fp + CT_SMEM)[lane] = reinterpret_cast<const uint32_t *>(&tmpl)[lane];
#if UDP_CROP_PITCHSEL
        if (!pl.axis) pitch = crop_pick_pitch(pl, m, CT_BOX, lane);
#endif
        if (lane == 0) *pitch_sh = pitch;
        __syncwarp();
        unsigned slot = 0;
        if (lane == 0) {
            const unsigned long long base = (unsigned long long)(uintptr_t)src - (unsigned)lead;
            asm volatile("tensormap.replace.tile.global_address.shared::cta.b1024.b64 [%0], %1;" ::"r"(sdesc), "l"(base) : "memory");
            asm volatile("tensormap.replace.tile.global_dim.shared::cta.b1024.b32 [%0], 0, %1;" ::"r"(sdesc), "r"((3 * sw + lead + 1) >> 1) : "memory");
            asm volatile("tensormap.replace.tile.global_dim.shared::cta.b1024.b32 [%0], 1, %1;" ::"r"(sdesc), "r"(sh) : "memory");
            asm volatile("tensormap.replace.tile.global_stride.shared::cta.b1024.b64 [%0], 0, %1;" ::"r"(sdesc), "l"((unsigned long long)ss) : "memory");
            asm volatile("tensormap.replace.tile.box_dim.shared::cta.b1024.b32 [%0], 0, %1;" ::"r"(sdesc), "r"(pitch >> 1) : "memory");
            asm volatile("tensormap.replace.tile.box_dim.shared::cta.b1024.b32 [%0], 1, %1;" ::"r"(sdesc), "r"(pl.boxh) : "memory");
            slot = atomicAdd(&g_tmap_next, 1u) & (TM_SLOTS - 1);
            while (atomicCAS(&g_tmap_busy[slot], 0u, 1u) != 0u) slot = (slot + 1) & (TM_SLOTS - 1);
#pragma unroll
            for (int i = 0; i < CT_NBOX; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(mbar + 8 * i) : "memory");
        }
        slot = __shfl_sync(0xffffffffu, slot, 0);
        gdesc = (unsigned long long)(uintptr_t)g_tmap_ring[slot];
        asm volatile("tensormap.cp_fenceproxy.global.shared::cta.tensormap::generic.release.gpu.sync.aligned [%0], [%1], 128;"
                     ::"l"(gdesc), "r"(sdesc) : "memory");
        if (lane == 0) {
            asm volatile("fence.proxy.tensormap::generic.acquire.gpu [%0], 128;" ::"l"(gdesc) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
    }
    // tile order: down a tile column, then the next column; tile k lives in box k % CT_NBOX and the load cursor
    // (warp 0 only) runs CT_NBOX - 1 tiles ahead of the gather cursor.
    __syncthreads();  // rowq, colc, the descriptor, the barriers and the pitch are in place
    pitch = *pitch_sh;
    const unsigned tx_bytes = (unsigned)(pitch * pl.boxh);
    const int ntiles = tiles_x * (H / TR);

    // tile (cx, y0): first tap column / row of its footprint; the box starts at the 16-byte boundary below
    auto corner = [&](int cx_, int y0_) {
        const int2 qa = rowq[y0_], qb = rowq[y0_ + TR - 1];
        const int4 cc = colc[cx_];
        return make_int2(((axis ? Xc : min(qa.x, qb.x)) + cc.x) >> 10, (min(qa.y, qb.y) + cc.z) >> 10);
    };
    auto issue = [&](int2 g_, int b_) {  // one thread: publish the corner, arm the barrier of box b_, start the load
        const unsigned bar = mbar + 8 * b_;
        cring[b_] = g_;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(tx_bytes) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(fpa + b_ * CT_BOX), "l"(gdesc), "r"(((3 * g_.x + lead) & ~15) >> 1), "r"(g_.y), "r"(bar) : "memory");
    };
    int lcx = 0, ly0 = 0, lbox = 0;  // load cursor (warp 0)
    if (warp == 0) {
#pragma unroll
        for (int i = 0; i < CT_NBOX - 1; ++i)
            if (i < ntiles) {
                const int2 g_ = corner(lcx, ly0);
                if (lane == 0) issue(g_, lbox);
                ly0 += TR;
                if (ly0 >= H) { ly0 = 0; ++lcx; }
                lbox = lbox + 1 == CT_NBOX ? 0 : lbox + 1;
            }
    }
    __syncthreads();  // cring[0 .. CT_NBOX - 2]

    OUT out;
    out.init(oargs, n, W, H, lane);
    // not zero-filled by the hardware: bytes [0, lead) of a row (they precede the image) and, when the row has
    // an odd number of bytes, the second half of its last two-byte element -- neither exists for 16-byte
    // aligned images of even width (every decoded image of the reference's datasets)
    const int rowb = 3 * sw + lead;
    const bool may_fix = lead != 0 || (rowb & 1);
    int cx = 0, y0 = 0, adx = 0, bdx = 0, b = 0;
    int ax_col = 0;          // axis-aligned persons: 3 * sx + lead of this lane's column ...
    uint32_t ax_wxp = 0;     // ... and its weight pair (32 - fx) | fx << 16
    unsigned par = 0;        // parity of box b's current fill
    for (int k = 0; k < ntiles; ++k) {
        if (warp == 0 && k + CT_NBOX - 1 < ntiles) {  // the box being refilled was released by the barrier that ended tile k-1
            const int2 g_ = corner(lcx, ly0);
            if (lane == 0) issue(g_, lbox);
            ly0 += TR;
            if (ly0 >= H) { ly0 = 0; ++lcx; }
            lbox = lbox + 1 == CT_NBOX ? 0 : lbox + 1;
        }
        const int2 g = cring[b];
        if (y0 == 0) {  // new tile column: this lane's Q10 column deltas
            adx = q10_col(m00, cx * CT + lane);
            bdx = q10_col(m10, cx * CT + lane);
            if (axis) {
                const int X = Xc + adx;  // m01 == 0: the same for every row
                ax_col = 3 * (X >> 10) + lead;
                ax_wxp = (((uint32_t)X >> 5) & 31u) * 0xffffu + 32u;
            }
        }
        const int xb = 3 * g.x + lead, c0 = xb & ~15;
        const unsigned box = fpa + b * CT_BOX;
        mbar_wait(mbar + 8 * b, par);
        bool fix = false;
        if (may_fix) {
            const bool fix_l = lead != 0 && c0 <= 0, fix_r = (rowb & 1) && c0 <= rowb && rowb < c0 + pitch;
            fix = fix_l || fix_r;
            if (fix) {
                for (int rr = t; rr < pl.boxh; rr += CT_THREADS)
                    if ((unsigned)(g.y + rr) < (unsigned)sh) {
                        if (fix_l) zero_bytes(fp + b * CT_BOX + rr * pitch, min(pitch, -c0), min(pitch, lead - c0));
                        if (fix_r) fp[b * CT_BOX + rr * pitch + (rowb - c0)] = 0;
                    }
                __syncthreads();
            }
        }
        if (axis) {
            const int colb = ax_col - c0;  // this lane's byte offset in a box row
            const unsigned base = box + (unsigned)(colb & ~3) - (unsigned)(g.y * pitch);
            const bool phi3 = (colb & 3) == 3;
            const uint32_t sh8 = (uint32_t)colb << 3;
            const int yw = y0 + warp * (TR >> 2);
            out.seek(yw, cx * CT);
            if (TR == CT) crop_gather_axis<CT / 4, WC>(rowq + yw, base, pitch, phi3, sh8, ax_wxp, out);
            else if (TR == CT / 2) crop_gather_axis<CT / 8, WC>(rowq + yw, base, pitch, phi3, sh8, ax_wxp, out);
            else crop_gather_axis<CT / 16, WC>(rowq + yw, base, pitch, phi3, sh8, ax_wxp, out);
        } else {
            const int K0 = (int)box + (xb - c0) - g.y * pitch - 3 * g.x;
            out.seek(y0 + warp, cx * CT);
            if (TR == CT) crop_gather_rows<CT / 4, WC>(rowq + y0 + warp, adx, bdx, pitch, K0, out);
            else if (TR == CT / 2) crop_gather_rows<CT / 8, WC>(rowq + y0 + warp, adx, bdx, pitch, K0, out);
            else crop_gather_rows<CT / 16, WC>(rowq + y0 + warp, adx, bdx, pitch, K0, out);
        }
        if (fix) asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // generic writes before the next TMA write
        __syncthreads();  // this box may be refilled
        y0 += TR;
        if (y0 >= H) { y0 = 0; ++cx; }
        if (++b == CT_NBOX) { b = 0; par ^= 1u; }
    }
    // every load of this CTA has landed (each tile waited on its barrier): hand the descriptor slot back
    if (t == 0) atomicExch(&g_tmap_busy[(gdesc - (unsigned long long)(uintptr_t)g_tmap_ring[0]) >> 7], 0u);
}

}  // namespace udp

// The template descriptor: any valid 2-D uint8 tiled map; every field that matters is replaced on
// the device.  cuTensorMapEncodeTiled comes from the driver through the runtime (no -lcuda).
static int crop_tma_template(CUtensorMap *out)
{
    static CUtensorMap tm;
    static int state = 0;  // 1 ready, -1 unavailable
    static std::mutex mu;
    std::lock_guard<std::mutex> lock(mu);
    if (state == 0) {
        state = -1;
        void *ring = nullptr;
        if (cudaGetSymbolAddress(&ring, udp::g_tmap_ring) == cudaSuccess &&
            udp::tma_make_2d(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, ring, 4096, 4096, 4096, 64, 32))
            state = 1;
        (void)cudaGetLastError();
    }
    if (state == 1) *out = tm;
    return state;
}

static bool crop_tiles_ok(int W, int H)
{
    return !udp_force_generic() && W % udp::CT == 0 && H % udp::CT == 0 && H <= 1024 && W <= 2048;
}
static bool crop_tma_armed(CUtensorMap *tmpl)
{
    const char *no_tma = getenv("UDP_CROP_NO_TMA");
    return !(no_tma && no_tma[0] == '1') && crop_tma_template(tmpl) == 1;
}

extern "C" int udp_warp_affine_launches(int W, int H)
{
    CUtensorMap tmpl;
    return crop_tiles_ok(W, H) && crop_tma_armed(&tmpl) ? 2 : 1;
}

struct CropFork {
    cudaStream_t side = nullptr;
    cudaEvent_t fork = nullptr, join = nullptr;
    std::mutex mu;
    bool ok = false, tried = false;
};
static bool crop_fork_enabled()
{
    const char *e = getenv("UDP_CROP_FORK");  // UDP_CROP_FORK=0: both launches on the caller's stream
    return !(e && e[0] == '0');
}
static CropFork *crop_fork_get()
{
    static CropFork forks[64];
    static std::mutex init_mu;
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 64) return nullptr;
    CropFork &f = forks[dev];
    std::lock_guard<std::mutex> lock(init_mu);
    if (!f.tried) {
        f.tried = true;
        f.ok = cudaStreamCreateWithFlags(&f.side, cudaStreamNonBlocking) == cudaSuccess &&
               cudaEventCreateWithFlags(&f.fork, cudaEventDisableTiming) == cudaSuccess &&
               cudaEventCreateWithFlags(&f.join, cudaEventDisableTiming) == cudaSuccess;
    }
    return f.ok ? &f : nullptr;
}

template <class OUT, int SM, int WC>
static int crop_launch_tiled_sm(const uint8_t *const *src_ptrs, const int32_t *src_h, const int32_t *src_w,
                                const int32_t *src_stride, const float *M, const typename OUT::Args &oargs,
                                size_t person_elems, int W, int H, int N, cudaStream_t stream);

// Footprint buffer by crop shape (measured), and -- for the uint8 crop -- the TMA kernel with the
// crop width of the reference's configurations (192, 256, 288) as a compile-time constant.
template <class OUT>
static int crop_launch_tiled(const uint8_t *const *src_ptrs, const int32_t *src_h, const int32_t *src_w,
                             const int32_t *src_stride, const float *M, const typename OUT::Args &oargs,
                             size_t person_elems, int W, int H, int N, cudaStream_t stream)
{
    constexpr bool U8 = std::is_same<OUT, udp::CropOutU8>::value;
#define UDP_CROP_GO(SMB, WCV) \
    return crop_launch_tiled_sm<OUT, SMB, WCV>(src_ptrs, src_h, src_w, src_stride, M, oargs, person_elems, W, H, N, stream)
    if ((long long)W * H <= 256 * 256) {
        if (U8 && W == 192) UDP_CROP_GO(udp::CT_SMEM_SMALL_CROP, U8 ? 192 : 0);
        if (U8 && W == 256) UDP_CROP_GO(udp::CT_SMEM_SMALL_CROP, U8 ? 256 : 0);
        UDP_CROP_GO(udp::CT_SMEM_SMALL_CROP, 0);
    }
    if (U8 && W == 288) UDP_CROP_GO(udp::CT_SMEM_LARGE_CROP, U8 ? 288 : 0);
    UDP_CROP_GO(udp::CT_SMEM_LARGE_CROP, 0);
#undef UDP_CROP_GO
}

template <class OUT, int SM, int WC>
static int crop_launch_tiled_sm(const uint8_t *const *src_ptrs, const int32_t *src_h, const int32_t *src_w,
                                const int32_t *src_stride, const float *M, const typename OUT::Args &oargs,
                                size_t person_elems, int W, int H, int N, cudaStream_t stream)
{
    const int tiles_x = W / udp::CT;
    CUtensorMap tmpl;
    const int tma = crop_tma_armed(&tmpl);
    // The TMA kernel takes the persons whose images / boxes suit the TMA unit, the cp.async kernel the
    // others (its CTAs for TMA persons exit at once).  The two touch disjoint persons, so the cp.async
    // launch goes to a side stream forked from and joined back into the caller's stream (one side
    // stream and event pair per device, created on first use; 3 % faster than back to back:
    // 2.28 -> 2.21 ms per 16 384 persons).  UDP_CROP_FORK=0 keeps both on the caller's stream.
    cudaStream_t tile_stream = stream;
    CropFork *fk = nullptr;
    if (tma && crop_fork_enabled() && (fk = crop_fork_get()) != nullptr) {
        // No fork while the caller's stream is being captured: the shared side stream would join that capture,
        // and a concurrent plain call from another thread would then be illegal.  If the fork cannot be set up,
        // both kernels simply run on the caller's stream.
        cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
        std::lock_guard<std::mutex> lock(fk->mu);  // record + wait must pair up per call
        if (cudaStreamIsCapturing(stream, &cap) == cudaSuccess && cap == cudaStreamCaptureStatusNone &&
            cudaEventRecord(fk->fork, stream) == cudaSuccess && cudaStreamWaitEvent(fk->side, fk->fork, 0) == cudaSuccess) {
            tile_stream = fk->side;
        } else {
            (void)cudaGetLastError();
            fk = nullptr;
        }
    }
    for (int n0 = 0; n0 < N; n0 += 65535) {  // gridDim.y limit
        const int nn = N - n0 < 65535 ? N - n0 : 65535;
        typename OUT::Args a = oargs;
        udp::crop_args_advance(a, (size_t)n0 * person_elems);
        udp::warp_affine_u8_tile_kernel<OUT, SM><<<dim3(tiles_x, nn), udp::CT_THREADS, SM + 16 + 8 * H, tile_stream>>>(
            src_ptrs + n0, src_h + n0, src_w + n0, src_stride + n0, M + 6 * (size_t)n0, a, W, H, tma);
        UDP_LAUNCH_CHECK();
    }
    if (tma) {
        udp::warp_affine_u8_tma_kernel<OUT, SM, WC><<<N, udp::CT_THREADS, SM + udp::CT_TAIL + 16 * tiles_x + 8 * H + 128, stream>>>(
            tmpl, src_ptrs, src_h, src_w, src_stride, M, oargs, W, H);
        UDP_LAUNCH_CHECK();
    }
    if (fk) {  // the caller's stream must not run ahead of the side stream's kernel: a failed join is an error
        std::lock_guard<std::mutex> lock(fk->mu);
        cudaError_t e = cudaEventRecord(fk->join, fk->side);
        if (e == cudaSuccess) e = cudaStreamWaitEvent(stream, fk->join, 0);
        if (e != cudaSuccess) {
            cudaStreamSynchronize(fk->side);  // last resort: order by blocking
            return (int)e;
        }
    }
    return 0;
}

extern "C" int udp_warp_affine_u8(const uint8_t *const *src_ptrs, const int32_t *src_h,
                                  const int32_t *src_w, const int32_t *src_stride, const float *M,
                                  uint8_t *dst, int W, int H, int N, void *stream)
{
    if (N < 0 || W < 1 || H < 1 || (W & 1)) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!src_ptrs || !src_h || !src_w || !src_stride || !M || !dst) return UDP_EINVAL;
    if (crop_tiles_ok(W, H) && ((uintptr_t)dst & 3) == 0) {
        udp::CropU8Args a{dst};
        return crop_launch_tiled<udp::CropOutU8>(src_ptrs, src_h, src_w, src_stride, M, a, (size_t)H * W * 3, W, H, N,
                                                 (cudaStream_t)stream);
    }
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

// float32 x / y for the 256 inputs of one channel: does reciprocal + one FMA correction give
// the correctly rounded quotient everywhere?  (It does for the ImageNet constants; any other
// constants are checked here, on the host, and fall back to a true division if not.)
static bool crop_norm_fast_ok(const float mean[3], const float sd[3], float inv[3])
{
    bool ok = true;
    for (int c = 0; c < 3; ++c) {
        inv[c] = 1.0f / sd[c];
        for (int b = 0; b < 256 && ok; ++b) {
            const float t = (float)b / 255.0f - mean[c];
            const float u = t * inv[c];
            const float q = std::fmaf(std::fmaf(-u, sd[c], t), inv[c], u);
            const float want = t / sd[c];
            ok = std::memcmp(&q, &want, 4) == 0;
        }
    }
    return ok;
}

extern "C" int udp_warp_affine_normalize_f32(const uint8_t *const *src_ptrs, const int32_t *src_h,
                                             const int32_t *src_w, const int32_t *src_stride, const float *M,
                                             const float *mean, const float *std_, float *dst, float *dst_flipped,
                                             int W, int H, int N, void *stream)
{
    if (N < 0 || W < 1 || H < 1 || !mean || !std_) return UDP_EINVAL;
    if (!crop_tiles_ok(W, H)) return UDP_EINVAL;  // crop sizes in multiples of 32 (every reference config)
    if (N == 0) return 0;
    if (!src_ptrs || !src_h || !src_w || !src_stride || !M || !dst || dst == dst_flipped) return UDP_EINVAL;
    if ((((uintptr_t)dst | (uintptr_t)dst_flipped) & 3) != 0) return UDP_EINVAL;
    udp::CropF32Args a;
    a.dst = dst;
    a.dst_flipped = dst_flipped;
    for (int c = 0; c < 3; ++c) {
        if (!(std_[c] != 0.0f)) return UDP_EINVAL;
        a.mean[c] = mean[c];
        a.sd[c] = std_[c];
    }
    a.exact_div = crop_norm_fast_ok(a.mean, a.sd, a.inv) ? 1 : 0;
    return crop_launch_tiled<udp::CropOutF32>(src_ptrs, src_h, src_w, src_stride, M, a, (size_t)3 * H * W, W, H, N,
                                              (cudaStream_t)stream);
}
