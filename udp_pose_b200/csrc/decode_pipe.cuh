// decode_pipe.cuh -- the streaming skeleton shared by both fused decoders (second generation).
//
// What the first-generation kernels got wrong (ncu, profiles/r01_decode_gauss_v1_ncu.txt and
// profiles/r01_decode_offset_fast_ncu.txt): one CTA per map with load -> sync -> blur -> sync
// phases left HBM idle while a CTA computed and the SM idle while it loaded (17-27 % occupancy,
// 36-48 % issue-active, 20-30 % of DRAM peak), and ~8000 warp-instructions per 64x48 map.
//
// This version:
//  * persistent CTAs (a few per SM) walk the (person, joint) maps with a TMA pipeline (third
//    generation; the second used 16-byte cp.async copies, see PipeFeed below): while map i is
//    being decoded, the tensor loads of map i+1 (and of its mirrored partner map) are in flight.
//    Three buffers rotate: P holds out_i and,
//    once the row pass has consumed it, T_i; Q holds flipped_i and is refilled with
//    flipped_{i+1} as soon as the row pass has read it; N receives out_{i+1} and becomes the
//    next P.  The loads place each row at a padded stride (SR/4 odd) so that the later per-row
//    LDS.128 reads are bank-conflict free; no thread holds HBM data in registers while waiting.
//  * the flip test ((out + mirror(partner)) * 0.5, utils/transforms.py:15-47 +
//    core/function.py:171) is fused into the ROW pass: a thread owns two rows (p, p + H/2) and a
//    column segment, averages its window straight into row-paired registers and runs cv2's
//    left-to-right FMA chain with packed fma.rn.f32x2 -- the padded copy A2 of the first
//    generation, its store/load round trip and one barrier are gone.
//  * T (the row-pass result, with reflected halo rows) aliases the buffer out_i came from, so
//    shared memory is three map-sized buffers (four in the first build of this file, which
//    limited the 96x72 maps to one CTA per SM).
//  * the column pass pairs two COLUMNS (x, x+1), cv2's centre-then-symmetric-pairs order.
//  * the raw arg-max needs only FMNMX per element: block max first, then the (rare) threads
//    whose local max equals it rescan their registers for the first index.
// Arithmetic and summation order are unchanged: results stay bit-identical.
// Maps holding a non-finite value are handed to the generic, NaN-exact code path.
#pragma once
#include "blur_fast.cuh"
#include "tma_host.cuh"

namespace udp {

__device__ __forceinline__ void cp_async16(void *smem_dst, const void *gsrc)
{
    const unsigned d = (unsigned)__cvta_generic_to_shared(smem_dst);
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;\n" ::"r"(d), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;\n" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;\n" ::: "memory"); }

__device__ __forceinline__ float fmax3(float a, float b, float c)
{
    float d;
    asm("max.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}
__device__ __forceinline__ float fmin3(float a, float b, float c)
{
    float d;
    asm("min.f32 %0, %1, %2, %3;" : "=f"(d) : "f"(a), "f"(b), "f"(c));
    return d;
}

// x / 255.0f, correctly rounded, without the division sequence: q = x*y, one FMA residual,
// one FMA correction (y = RN(1/255)).  Checked exhaustively against x / 255.0f for every
// finite float32 (only -0.0 needed the select).
__device__ __forceinline__ float div255(float x)
{
    const float y = 0x1.010102p-8f;
    const float q = __fmul_rn(x, y);
    const float r = fmaf(-q, 255.0f, x);
    const float q2 = fmaf(r, y, q);
    return x == 0.0f ? x : q2;
}

template <int W, int H, int R, int NT>
struct PipeGeom {
    static_assert(W % 4 == 0 && H % 2 == 0, "map shape");
    static constexpr int PAD = (R <= 4) ? 4 : 8;          // halo columns loaded per side (float4 units)
    static constexpr int HP = H / 2;                      // thread rows: (p, p + HP)
    static_assert(NT % HP == 0, "NT must be a multiple of H/2");
    static constexpr int XS = NT / HP;                    // column segments
    static_assert(W % XS == 0, "row-pass split");
    static constexpr int SEG = W / XS;
    static_assert(SEG % 4 == 0 && SEG >= R + 1, "segment");
    static constexpr int NCOL = SEG + 2 * PAD;            // window columns per row
    static constexpr int NV4 = NCOL / 4;
    static constexpr int CP = W / 2;                      // column pairs
    static_assert(NT % CP == 0, "NT must be a multiple of W/2");
    static constexpr int YS = NT / CP;
    static_assert(H % YS == 0, "column-pass split");
    static constexpr int RS = H / YS;
    static constexpr int SR = (W / 4) % 2 == 1 ? W : W + 4;  // raw row stride, SR/4 odd
    static constexpr int ST = W + 2;                      // T row stride
    static constexpr int TROWS = H;                       // no halo rows: the column pass reflects its row index (4 % faster than
                                                          // writing R reflected rows above and below, and T fits the raw buffer)
    static constexpr int TOFF = 0;
    static constexpr int RAW = H * SR;                    // floats per raw tensor
    static constexpr int BUF0 = (RAW > TROWS * ST) ? RAW : TROWS * ST;
    static constexpr int BUF = ((BUF0 + 31) / 32) * 32;   // one buffer, a multiple of 128 bytes (TMA destination)
    static constexpr size_t SMEM = (size_t)3 * BUF * sizeof(float);
    static constexpr int NWARP = (NT + 31) / 32;
};

// The TMA feed of the pipeline: one tensor load per map instead of H*W/4 cp.async per tensor (which
// cost the SM an LDGSTS + three filler issue slots each and a quarter of the kernels' shared-memory
// wavefronts).  The network output is described as a 2-D float tensor [maps*H][W]; the box is
// [H][SR]: the SR - W columns past the row end lie outside the tensor and are zero-filled, so a map
// lands directly at the padded row stride the row pass wants.  One mbarrier per buffer; every
// thread tracks the parities, thread 0 issues.
template <int W, int H, int R, int NT>
struct PipeFeed {
    using G = PipeGeom<W, H, R, NT>;
    static constexpr unsigned BYTES = (unsigned)(G::SR * H * 4);
    unsigned smem0, bars, par;
    __device__ __forceinline__ void init(const float *smem, unsigned long long *bar_storage)
    {
        smem0 = (unsigned)__cvta_generic_to_shared(smem);
        bars = (unsigned)__cvta_generic_to_shared(bar_storage);
        par = 0u;
        if (threadIdx.x == 0) {
#pragma unroll
            for (int i = 0; i < 3; ++i) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bars + 8 * i) : "memory");
            asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
        }
        __syncthreads();
    }
    // one thread: arm the barrier of `buf` and start the load of the map whose first row is `row`
    __device__ __forceinline__ void issue(int buf, const CUtensorMap *tm, int row) const
    {
        const unsigned bar = bars + 8 * buf;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(BYTES) : "memory");
        asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];"
                     ::"r"(smem0 + (unsigned)buf * (unsigned)(G::BUF * 4)), "l"((unsigned long long)(uintptr_t)tm), "r"(0), "r"(row), "r"(bar)
                     : "memory");
    }
    // every thread: the load into `buf` has landed
    __device__ __forceinline__ void wait(int buf)
    {
        mbar_wait(bars + 8 * buf, (par >> buf) & 1u);
        par ^= 1u << buf;
    }
    // every thread, before the barrier that precedes a refill: order this thread's shared-memory
    // writes (T) before the asynchronous-proxy write of the next load
    static __device__ __forceinline__ void release() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
};

inline bool pipe_make_map(CUtensorMap *tm, const float *base, int W, int H, int SR, long long maps)
{
    if (maps * H >= 0x7fffffffLL) return false;  // row coordinates are int32
    return tma_make_2d(tm, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, base, (unsigned long long)W, (unsigned long long)(maps * H),
                       (unsigned long long)W * 4, (unsigned)SR, (unsigned)H);
}

// (person, joint) of a map index that advances by a fixed step, without 64-bit divisions.
struct MapWalk {
    int n, k, sn, sk, K;
    __device__ __forceinline__ void init(long long map, int step, int K_)
    {
        K = K_;
        n = (int)(map / K_); k = (int)(map - (long long)n * K_);
        sn = step / K_; sk = step - sn * K_;
    }
    __device__ __forceinline__ void advance()
    {
        n += sn; k += sk;
        if (k >= K) { k -= K; ++n; }
    }
};

// The averaged window of one thread: v[j] = (row p, row p + HP) at column x0 - PAD + j.
// mode: 0 = no flip, 1 = flip (+), 2 = flip with the partner negated (x-offset channels).
template <int W, int H, int R, int NT, bool FLIP, bool DIV255>
__device__ __forceinline__ void pipe_window(const float *raw_out, const float *raw_fl, int p, int xs,
                                            f32x2 (&v)[PipeGeom<W, H, R, NT>::NCOL])
{
    using G = PipeGeom<W, H, R, NT>;
    constexpr int PAD = G::PAD, SEG = G::SEG;
    const int x0 = xs * SEG;
    const float *a0 = raw_out + p * G::SR, *a1 = a0 + G::HP * G::SR;
    const float *f0 = raw_fl + p * G::SR, *f1 = f0 + G::HP * G::SR;
    // the chunks of the segment itself are always inside the map: one base each (x0 folded in) and constant offsets;
    // only the halo chunks need the clamp (-40 instructions per thread and map; 4 % on the 96x72 decoder)
    const float *ax = a0 + x0, *fx = f0 + (W - 4 - x0);
#pragma unroll
    for (int q = 0; q < G::NV4; ++q) {
        int xc = x0 - PAD + 4 * q;
        xc = xc < 0 ? 0 : (xc > W - 4 ? W - 4 : xc);  // out-of-map chunks are replaced below
        const bool inner = 4 * q >= PAD && 4 * q < PAD + SEG;
        const float4 r0 = *reinterpret_cast<const float4 *>(inner ? ax + (4 * q - PAD) : a0 + xc);
        const float4 r1 = *reinterpret_cast<const float4 *>(inner ? ax + (4 * q - PAD) + G::HP * G::SR : a1 + xc);
        float e[8] = {r0.x, r1.x, r0.y, r1.y, r0.z, r1.z, r0.w, r1.w};
        if (FLIP) {  // the two sums of a row pair land in a register pair, so the halving is one packed multiply
            const float4 g0 = *reinterpret_cast<const float4 *>(inner ? fx - (4 * q - PAD) : f0 + (W - 4 - xc));
            const float4 g1 = *reinterpret_cast<const float4 *>(inner ? fx - (4 * q - PAD) + G::HP * G::SR : f1 + (W - 4 - xc));
            e[0] = __fadd_rn(r0.x, g0.w); e[1] = __fadd_rn(r1.x, g1.w);
            e[2] = __fadd_rn(r0.y, g0.z); e[3] = __fadd_rn(r1.y, g1.z);
            e[4] = __fadd_rn(r0.z, g0.y); e[5] = __fadd_rn(r1.z, g1.y);
            e[6] = __fadd_rn(r0.w, g0.x); e[7] = __fadd_rn(r1.w, g1.x);
        }
        if (DIV255) {
#pragma unroll
            for (int i = 0; i < 8; ++i) e[i] = div255(FLIP ? __fmul_rn(e[i], 0.5f) : e[i]);
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const f32x2 s = pk(e[2 * i], e[2 * i + 1]);
            v[4 * q + i] = (FLIP && !DIV255) ? fmul2p(s, 0.5f) : s;
        }
    }
    // reflect-101 columns at the map borders: column -1-i <- column 1+i, column W+i <- column W-2-i
    if (xs == 0) {
#pragma unroll
        for (int i = 0; i < R; ++i) v[PAD - 1 - i] = v[PAD + 1 + i];
    }
    if (xs == G::XS - 1) {
#pragma unroll
        for (int i = 0; i < R; ++i) v[PAD + SEG + i] = v[PAD + SEG - 2 - i];
    }
}

// Row filter from the register window into T (stride ST, H rows).
template <int W, int H, int R, int NT>
__device__ __forceinline__ void pipe_row_filter(const f32x2 (&v)[PipeGeom<W, H, R, NT>::NCOL],
                                                float *T, int p, int xs)
{
    using G = PipeGeom<W, H, R, NT>;
    constexpr int PAD = G::PAD, SEG = G::SEG;
    const int x0 = xs * SEG, y0 = p, y1 = p + G::HP;
    float *t0 = T + (y0 + G::TOFF) * G::ST + x0, *t1 = T + (y1 + G::TOFF) * G::ST + x0;
#pragma unroll
    for (int o = 0; o < SEG; o += 2) {
        f32x2 s0 = fmul2p(v[PAD - R + o], Taps<R>::k(0)), s1 = fmul2p(v[PAD - R + o + 1], Taps<R>::k(0));
#pragma unroll
        for (int i = 1; i < 2 * R + 1; ++i) {
            s0 = ffma2p(v[PAD - R + o + i], Taps<R>::k(i), s0);
            s1 = ffma2p(v[PAD - R + o + 1 + i], Taps<R>::k(i), s1);
        }
        const f32x2 ra = pk(lo_of(s0), lo_of(s1)), rb = pk(hi_of(s0), hi_of(s1));
        *reinterpret_cast<f32x2 *>(t0 + o) = ra;
        *reinterpret_cast<f32x2 *>(t1 + o) = rb;
    }
}

// Column pass: thread = (column pair c, y segment); emit(y, x, a, b) gets the blurred values a = (y,x), b = (y,x+1).
template <int W, int H, int R, int NT, typename Emit>
__device__ __forceinline__ void pipe_col_pass(const float *T, Emit emit)
{
    using G = PipeGeom<W, H, R, NT>;
    constexpr int RS = G::RS;
    const int c = threadIdx.x % G::CP, ys = threadIdx.x / G::CP;
    f32x2 t[RS + 2 * R];
    const float *col = T + 2 * c;
    const int m0 = ys * RS - R;  // map row of t[0]; only the first / last segment reaches outside the map
#pragma unroll
    for (int i = 0; i < RS + 2 * R; ++i) {
        int m = m0 + i;
        if (i < R) m = m < 0 ? -m : m;                                   // reflect-101 above row 0
        if (i >= RS + R) m = m > H - 1 ? 2 * (H - 1) - m : m;            // and below row H-1
        t[i] = *reinterpret_cast<const f32x2 *>(col + m * G::ST);
    }
#pragma unroll
    for (int o = 0; o < RS; ++o) {
        f32x2 s = ffma2p(t[o + R], Taps<R>::k(R), pk(0.f, 0.f));
#pragma unroll
        for (int i = 1; i <= R; ++i) s = ffma2p(fadd2p(t[o + R + i], t[o + R - i]), Taps<R>::k(R + i), s);
        emit(ys * RS + o, 2 * c, lo_of(s), hi_of(s));
    }
}

// One blurred value from T (same order as the column pass).
template <int W, int H, int R, int NT>
__device__ __forceinline__ float pipe_col_at(const float *T, int y, int x)
{
    using G = PipeGeom<W, H, R, NT>;
    const float *q = T + x;
    float s = fmaf(q[y * G::ST], Taps<R>::k(R), 0.0f);
#pragma unroll
    for (int i = 1; i <= R; ++i) {
        const int a = y + i > H - 1 ? 2 * (H - 1) - (y + i) : y + i, b = y - i < 0 ? i - y : y - i;
        s = fmaf(__fadd_rn(q[a * G::ST], q[b * G::ST]), Taps<R>::k(R + i), s);
    }
    return s;
}

// Raw first-index arg-max over the thread's own SEG columns of its two rows.
// Step 1: local maxima (FMNMX only).  Step 2 (after the block max M is known): first index == M.
template <int W, int H, int R, int NT>
__device__ __forceinline__ float pipe_local_max(const f32x2 (&v)[PipeGeom<W, H, R, NT>::NCOL], float &fsum)
{
    using G = PipeGeom<W, H, R, NT>;
    float m0 = -INFINITY, m1 = -INFINITY;
    f32x2 acc = pk(0.f, 0.f);
#pragma unroll
    for (int j = G::PAD; j < G::PAD + G::SEG; j += 2) {
        m0 = fmax3(m0, lo_of(v[j]), lo_of(v[j + 1]));
        m1 = fmax3(m1, hi_of(v[j]), hi_of(v[j + 1]));
        acc = fadd2p(acc, fadd2p(v[j], v[j + 1]));
    }
    fsum = lo_of(acc) + hi_of(acc);
    return fmaxf(m0, m1);
}

template <int W, int H, int R, int NT>
__device__ __forceinline__ int pipe_first_index(const f32x2 (&v)[PipeGeom<W, H, R, NT>::NCOL], float M,
                                                int p, int xs)
{
    using G = PipeGeom<W, H, R, NT>;
    int best = 0x7fffffff;
    const int b0 = p * W + xs * G::SEG, b1 = b0 + G::HP * W;
#pragma unroll
    for (int j = G::SEG - 1; j >= 0; --j) {  // descending, so the lowest index is written last
        if (hi_of(v[G::PAD + j]) == M) best = b1 + j;
    }
#pragma unroll
    for (int j = G::SEG - 1; j >= 0; --j) {
        if (lo_of(v[G::PAD + j]) == M) best = b0 + j;
    }
    return best;
}

}  // namespace udp
