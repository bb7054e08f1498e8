// decode_offset_prune.cuh -- third generation of the fused UDPv1 offset decoder: an EXACT
// branch-and-bound arg-max of the 15x15-blurred classification map (inference.py:165,168).
//
// The reference blurs the whole map and takes np.argmax of the result.  The second-generation kernel
// (decode_offset_pipe_kernel) did the same on the packed FP32 pipe: 3 300 warp-instructions per 64x48
// map, 1 440 of them the FMAs of the separable blur, issue-bound at 12 warps per SM
// (profiles/r01_decode_offset_v9_ncu.txt: 0.59 of the HBM roofline).  Most of that blur cannot matter:
//
//   every tap is positive and the taps sum to 1 (+- 1e-7), so a blurred value is bounded by the largest
//   input value under its 15x15 window:   b(y, x) <= max(window) + EPS * max|map|      (rounding slack)
//   and any blurred value that HAS been computed, L, is a lower bound of the maximum.
//
// So: (1) the flip-averaged map A is formed once (as before) together with the maxima of its 8x8 blocks;
// (2) one warp computes L = b at the raw arg-max, exactly as the full pass would; (3) a block can hold the
// arg-max only if the maximum over its 3x3 block neighbourhood (which covers every 15x15 window of its
// pixels, 7 < 8) reaches L - EPS * max|A|; (4) the row pass runs only for the 8-row groups that touch a
// candidate block's windows, the column pass only for candidate blocks, both with the unchanged FMA order
// (row: left-to-right chain; column: centre, then symmetric pairs), so every value that is computed has the
// bits the reference computes, and every pixel that could reach the maximum IS computed: arg-max index,
// first-index tie rule and the blurred maximum (`maxvals`) stay bit-exact.  A peaked map (what a network
// emits) keeps 9-16 of 48 blocks; a map of pure noise keeps all of them and costs what the full pass costs.
//
// Shared memory per CTA: two stages x (P, Q).  P receives out_i by TMA and is overwritten in place by
// A2 (row pairs (8g+a, 8g+a+4) interleaved as float2, the packed operand of the row pass); Q receives
// flipped_i and is overwritten by T (row-pass results, row-major).  The loads of map i+1 are issued at the
// top of map i into the other stage.
#pragma once
#include "decode_pipe.cuh"

namespace udp {

template <int W, int H, int NT>
struct PruneGeom {
    static_assert(W % 8 == 0 && H % 8 == 0 && W >= 16 && H >= 16, "map shape");
    static constexpr int R = 7;
    static constexpr int GX = W / 8, GY = H / 8, NBLK = GX * GY;  // 8x8 blocks
    static constexpr int NQ = H / 2;                              // row pairs q = 4g + a <-> rows (8g+a, 8g+a+4)
    static constexpr int SR = (W / 4) % 2 == 1 ? W : W + 4;       // raw row stride (floats), SR/4 odd: TMA box [H][SR]
    static constexpr int SA = ((2 * W) / 4) % 2 == 1 ? 2 * W : 2 * W + 4;  // A2 row-pair stride (floats), SA/4 odd
    static constexpr int ST = SR;                                 // T row stride (floats), a multiple of 4
    static constexpr int RAW = H * SR;
    static_assert(NQ * SA <= RAW && H * ST <= RAW, "A2 and T must fit the raw buffers");
    static constexpr int BUF = ((RAW + 31) / 32) * 32;            // floats, a multiple of 128 bytes
    static constexpr size_t SMEM = (size_t)4 * BUF * sizeof(float);
    static constexpr unsigned BYTES = (unsigned)(RAW * 4);        // one tensor load
    static constexpr int CELLS = NQ * (W / 4);                    // phase A: (row pair, 4 columns)
    static_assert(CELLS % 32 == 0, "whole warps of cells");
    static constexpr int CPT = (CELLS + NT - 1) / NT;             // cells per thread
    static constexpr int NWARP = NT / 32;
};

// rounding slack of the bound, relative to max|A|: 30 roundings of 2^-24 on terms that sum to <= max|A|,
// |sum(taps)^2 - 1| < 1e-6, doubled for the rounding of the threshold itself
constexpr float PRUNE_EPS = 2.0e-5f;

// four mbarriers (stage s: P at 2s, Q at 2s+1); every thread tracks the parities, one thread issues
template <int W, int H, int NT>
struct PruneFeed {
    using G = PruneGeom<W, H, NT>;
    unsigned smem0, bars, par;
    __device__ __forceinline__ void init(const float *smem, unsigned long long *bar_storage)
    {
        smem0 = (unsigned)__cvta_generic_to_shared(smem);
        bars = (unsigned)__cvta_generic_to_shared(bar_storage);
        par = 0u;
        if (threadIdx.x == 0) {
#pragma unroll
            for (int i = 0; i < 4; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8 * i) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
    }
    __device__ __forceinline__ void issue(int buf, const CUtensorMap *tm, int row) const
    {
        const unsigned bar = bars + 8 * buf;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(G::BYTES) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem0 + (unsigned)buf * (unsigned)(G::BUF * 4)), "l"((unsigned long long)(uintptr_t)tm), "r"(0), "r"(row), "r"(bar)
                     : "memory");
    }
    __device__ __forceinline__ void wait(int buf)
    {
        mbar_wait(bars + 8 * buf, (par >> buf) & 1u);
        par ^= 1u << buf;
    }
    static __device__ __forceinline__ void release() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
};

// cell c -> (block, row-in-group a, chunk): the 8 cells of a block are 8 consecutive lanes
template <int W, int H, int NT>
__device__ __forceinline__ void prune_cell(int c, int &g, int &j, int &a, int &cx)
{
    using G = PruneGeom<W, H, NT>;
    const int blk = c >> 3, sub = c & 7;
    g = blk / G::GX;
    j = blk - g * G::GX;
    a = sub >> 1;
    cx = 2 * j + (sub & 1);
}

// A(r, c) out of the row-pair interleaved copy
template <int W, int H, int NT>
__device__ __forceinline__ float prune_a_at(const float *A2, int r, int c)
{
    using G = PruneGeom<W, H, NT>;
    return A2[(4 * (r >> 3) + (r & 3)) * G::SA + 2 * c + ((r >> 2) & 1)];
}

// One blurred value, the arithmetic of the full pass: 15 row chains (lane i: row reflect(y - 7 + i)),
// then the column combine (centre, symmetric pairs).  Executed by one full warp; result in every lane.
template <int W, int H, int NT>
__device__ __forceinline__ float prune_blur_at(const float *A2, int y, int x, int lane)
{
    using G = PruneGeom<W, H, NT>;
    constexpr int R = 7;
    const int l15 = lane < 15 ? lane : 14;
    const int r = reflect101(y - R + l15, H);
    const int cidx = 2 * reflect101(x - R + l15, W);  // lane t: float offset of the column of tap t, shared by shuffle
    const float *rowp = A2 + (4 * (r >> 3) + (r & 3)) * G::SA + ((r >> 2) & 1);
    float s = __fmul_rn(rowp[__shfl_sync(0xffffffffu, cidx, 0)], Taps<R>::k(0));
#pragma unroll
    for (int t = 1; t < 2 * R + 1; ++t) s = fmaf(rowp[__shfl_sync(0xffffffffu, cidx, t)], Taps<R>::k(t), s);
    float b = fmaf(__shfl_sync(0xffffffffu, s, R), Taps<R>::k(R), 0.0f);
#pragma unroll
    for (int i = 1; i <= R; ++i)
        b = fmaf(__fadd_rn(__shfl_sync(0xffffffffu, s, R + i), __shfl_sync(0xffffffffu, s, R - i)), Taps<R>::k(R + i), b);
    return b;
}

// Row-pass item: row pair q = 4g + a, the 8 columns of block column j.  T rows 8g+a and 8g+a+4.
template <int W, int H, int NT>
__device__ __forceinline__ void prune_row_item(const float *A2, float *T, int g, int j, int a)
{
    using G = PruneGeom<W, H, NT>;
    constexpr int R = 7;
    const int x0 = 8 * j;
    const float *row = A2 + (4 * g + a) * G::SA;
    f32x2 v[24];  // v[k] = columns x0 - 8 + k of the two rows
#pragma unroll
    for (int m = 0; m < 12; ++m) {
        int xc = x0 - 8 + 2 * m;
        xc = xc < 0 ? 0 : (xc > W - 2 ? W - 2 : xc);  // out-of-map pairs are replaced below
        const float4 t = *reinterpret_cast<const float4 *>(row + 2 * xc);
        v[2 * m] = pk(t.x, t.y);
        v[2 * m + 1] = pk(t.z, t.w);
    }
    if (j == 0) {  // reflect-101: column -1-i <- column 1+i
#pragma unroll
        for (int i = 0; i < R; ++i) v[7 - i] = v[9 + i];
    }
    if (j == G::GX - 1) {  // column W+i <- column W-2-i
#pragma unroll
        for (int i = 0; i < R; ++i) v[16 + i] = v[14 - i];
    }
    float *t0 = T + (8 * g + a) * G::ST + x0, *t1 = t0 + 4 * G::ST;
#pragma unroll
    for (int o = 0; o < 8; o += 2) {  // two outputs at a time: their row-0 halves and row-1 halves leave as 8-byte stores
        f32x2 s0 = fmul2p(v[o + 1], Taps<R>::k(0)), s1 = fmul2p(v[o + 2], Taps<R>::k(0));
#pragma unroll
        for (int t = 1; t < 2 * R + 1; ++t) {
            s0 = ffma2p(v[o + 1 + t], Taps<R>::k(t), s0);
            s1 = ffma2p(v[o + 2 + t], Taps<R>::k(t), s1);
        }
        *reinterpret_cast<f32x2 *>(t0 + o) = pk(lo_of(s0), lo_of(s1));
        *reinterpret_cast<f32x2 *>(t1 + o) = pk(hi_of(s0), hi_of(s1));
    }
}

// Column-pass item: rows 8g .. 8g+7 of the column pair (2c, 2c+1); updates the thread's running arg-max
// (np.argmax order within the item: rows ascending, strict >, so the first row of a tie stays).
template <int W, int H, int NT>
__device__ __forceinline__ void prune_col_item(const float *T, int g, int c, float &bv, int &bi)
{
    using G = PruneGeom<W, H, NT>;
    constexpr int R = 7;
    const float *col = T + 2 * c;
    f32x2 t[8 + 2 * R];
#pragma unroll
    for (int k = 0; k < 8 + 2 * R; ++k) {
        int m = 8 * g - R + k;
        if (k < R) m = m < 0 ? -m : m;                          // reflect-101 above row 0
        if (k >= 8 + R) m = m > H - 1 ? 2 * (H - 1) - m : m;    // and below row H-1
        t[k] = *reinterpret_cast<const f32x2 *>(col + m * G::ST);
    }
    float v0 = -INFINITY, v1 = -INFINITY;
    int y0 = 0, y1 = 0;
#pragma unroll
    for (int o = 0; o < 8; ++o) {
        f32x2 s = ffma2p(t[o + R], Taps<R>::k(R), pk(0.f, 0.f));
#pragma unroll
        for (int i = 1; i <= R; ++i) s = ffma2p(fadd2p(t[o + R + i], t[o + R - i]), Taps<R>::k(R + i), s);
        const float sa = lo_of(s), sb = hi_of(s);
        if (sa > v0) { v0 = sa; y0 = 8 * g + o; }
        if (sb > v1) { v1 = sb; y1 = 8 * g + o; }
    }
    const int i0 = y0 * W + 2 * c, i1 = y1 * W + 2 * c + 1;
    if (v0 > bv || (v0 == bv && i0 < bi)) { bv = v0; bi = i0; }
    if (v1 > bv || (v1 == bv && i1 < bi)) { bv = v1; bi = i1; }
}

}  // namespace udp
