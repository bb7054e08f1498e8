// blur_fast.cuh -- compile-time-shaped, packed-FP32 (FFMA2/FADD2) version of the bit-exact
// separable Gaussian blur for the decoders.  Same arithmetic and summation order as blur.cuh
// (= cv2.GaussianBlur); what changes is how the work is laid out so that the
// kernel stops being instruction-issue bound (ncu, profiles/r01_decode_offset_v1.txt: 42 %
// issue-active at 23 % occupancy, ~9400 warp-instructions per 64x48 map).
//
//  * Blackwell's packed f32x2 pipe: one FFMA2/FADD2 does two IEEE fp32 operations.  The row
//    pass pairs two ROWS (y, y+1) at the same column, the column pass pairs two COLUMNS
//    (x, x+1) at the same row, so every packed operand is an aligned 8-byte value in shared
//    memory and no register shuffling is needed.
//  * Each thread slides along its row pair / column pair with the whole window unrolled in
//    registers: 0.5 shared-memory loads per output instead of 2.75.
//  * One small CTA (64 or 128 threads) per map, several CTAs per SM: no cross-map sync, HBM
//    latency is hidden by the other resident CTAs.
//
// Shared memory for one H x W map:
//   A2 : [H/2][SP] float2, element (p, x) = (map[2p][x], map[2p+1][x]); data starts at column
//        PADL, reflect-101 pad columns on both sides; SP/2 is odd -> conflict-free LDS.128
//        when consecutive lanes own consecutive row pairs.
//   T  : [H + 2R][ST] float, the row-pass result with R reflected halo rows above and below.
#pragma once
#include "blur.cuh"

namespace udp {

__device__ __forceinline__ float2 ffma2(float2 a, float k, float2 c)
{
    float2 d;
    asm("{.reg .b64 ra, rb, rc, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n mov.b64 rc, {%5,%6};\n"
        " fma.rn.f32x2 rd, ra, rb, rc;\n mov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(k), "f"(c.x), "f"(c.y));
    return d;
}
__device__ __forceinline__ float2 fmul2(float2 a, float k)
{
    float2 d;
    asm("{.reg .b64 ra, rb, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%4};\n"
        " mul.rn.f32x2 rd, ra, rb;\n mov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(k));
    return d;
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b)
{
    float2 d;
    asm("{.reg .b64 ra, rb, rd;\n mov.b64 ra, {%2,%3};\n mov.b64 rb, {%4,%5};\n"
        " add.rn.f32x2 rd, ra, rb;\n mov.b64 {%0,%1}, rd;}"
        : "=f"(d.x), "=f"(d.y) : "f"(a.x), "f"(a.y), "f"(b.x), "f"(b.y));
    return d;
}

template <int W, int H, int R, int NT>
struct FastGeom {
    static_assert(W % 4 == 0 && H % 2 == 0, "map shape");
    static constexpr int PADL = (R <= 4) ? 4 : 8;
    static constexpr int PADR = (R <= 4) ? 6 : 10;
    static constexpr int SP = W + PADL + PADR;  // float2 per row pair; SP/2 odd
    static_assert((SP / 2) % 2 == 1, "SP/2 must be odd");
    static constexpr int ST = W + 2;            // floats per T row
    static constexpr int TROWS = H + 2 * R;
    static constexpr int HP = H / 2;
    static constexpr int XS = NT / HP;          // row pass: x segments per row pair
    static_assert(XS >= 1 && W % XS == 0, "row-pass split");
    static constexpr int SEG = W / XS;
    static_assert(SEG % 2 == 0, "segment");
    static constexpr int CP = W / 2;            // column pass: column pairs
    static constexpr int YS = NT / CP;          // y segments
    static_assert(YS >= 1 && H % YS == 0, "column-pass split");
    static constexpr int RS = H / YS;
    static constexpr size_t A2_BYTES = (size_t)HP * SP * 8;
    static constexpr size_t T_BYTES = (size_t)TROWS * ST * 4;
    static constexpr size_t SMEM = A2_BYTES + ((T_BYTES + 15) / 16) * 16;
};

// Stage 1: global -> A2, fusing the flip test ((out + mirrored partner) * 0.5) and the RSN
// pre-division.  Tracks the raw first-index argmax and a running sum used only to detect
// non-finite inputs (those maps take the generic, NaN-exact path).
template <int W, int H, int R, int NT>
__device__ __forceinline__ void fast_load(const float *__restrict__ src, const float *__restrict__ fl,
                                          float pre_div, float2 *A2, float &best_v, int &best_i,
                                          float &fsum)
{
    using G = FastGeom<W, H, R, NT>;
    constexpr int W4 = W / 4, ITEMS = G::HP * W4;
    best_v = -INFINITY;
    best_i = 0x7fffffff;
    float2 acc = make_float2(0.f, 0.f);
#pragma unroll
    for (int it0 = 0; it0 < ITEMS; it0 += NT) {
        const int it = it0 + threadIdx.x;
        if (ITEMS % NT == 0 || it < ITEMS) {
            const int p = it / W4, x = (it - p * W4) << 2;
            const float *r0 = src + (2 * p) * W + x;
            float4 a = ld_stream_f4(r0), b = ld_stream_f4(r0 + W);
            if (fl) {
                const float *f0 = fl + (2 * p) * W + (W - 4 - x);
                const float4 fa = ld_stream_f4(f0), fb = ld_stream_f4(f0 + W);
                a.x = __fmul_rn(__fadd_rn(a.x, fa.w), 0.5f); a.y = __fmul_rn(__fadd_rn(a.y, fa.z), 0.5f);
                a.z = __fmul_rn(__fadd_rn(a.z, fa.y), 0.5f); a.w = __fmul_rn(__fadd_rn(a.w, fa.x), 0.5f);
                b.x = __fmul_rn(__fadd_rn(b.x, fb.w), 0.5f); b.y = __fmul_rn(__fadd_rn(b.y, fb.z), 0.5f);
                b.z = __fmul_rn(__fadd_rn(b.z, fb.y), 0.5f); b.w = __fmul_rn(__fadd_rn(b.w, fb.x), 0.5f);
            }
            if (pre_div != 0.0f) {
                a.x = __fdiv_rn(a.x, pre_div); a.y = __fdiv_rn(a.y, pre_div);
                a.z = __fdiv_rn(a.z, pre_div); a.w = __fdiv_rn(a.w, pre_div);
                b.x = __fdiv_rn(b.x, pre_div); b.y = __fdiv_rn(b.y, pre_div);
                b.z = __fdiv_rn(b.z, pre_div); b.w = __fdiv_rn(b.w, pre_div);
            }
            float4 *dst = reinterpret_cast<float4 *>(A2 + p * G::SP + G::PADL + x);
            dst[0] = make_float4(a.x, b.x, a.y, b.y);
            dst[1] = make_float4(a.z, b.z, a.w, b.w);
            acc = fadd2(acc, fadd2(fadd2(make_float2(a.x, b.x), make_float2(a.y, b.y)),
                                   fadd2(make_float2(a.z, b.z), make_float2(a.w, b.w))));
            const int i0 = (2 * p) * W + x, i1 = i0 + W;
            // row 2p first (lower flat indices), then row 2p+1; ties keep the lower index
            if (a.x > best_v || (a.x == best_v && i0 < best_i)) { best_v = a.x; best_i = i0; }
            if (a.y > best_v) { best_v = a.y; best_i = i0 + 1; }
            if (a.z > best_v) { best_v = a.z; best_i = i0 + 2; }
            if (a.w > best_v) { best_v = a.w; best_i = i0 + 3; }
            if (b.x > best_v) { best_v = b.x; best_i = i1; }
            if (b.y > best_v) { best_v = b.y; best_i = i1 + 1; }
            if (b.z > best_v) { best_v = b.z; best_i = i1 + 2; }
            if (b.w > best_v) { best_v = b.w; best_i = i1 + 3; }
        }
    }
    fsum = acc.x + acc.y;
}

template <int W, int H, int R, int NT>
__device__ __forceinline__ void fast_pads(float2 *A2)
{
    using G = FastGeom<W, H, R, NT>;
    for (int t = threadIdx.x; t < G::HP * 2 * R; t += NT) {
        const int p = t / (2 * R), q = t - p * 2 * R;
        float2 *row = A2 + p * G::SP + G::PADL;
        if (q < R) row[-1 - q] = row[1 + q];
        else { const int i = q - R; row[W + i] = row[W - 2 - i]; }
    }
}

// Stage 2: row pass.  Thread = (row pair p, x segment); the whole SEG + 2R window is unrolled.
template <int W, int H, int R, int NT>
__device__ __forceinline__ void fast_row_pass(const float2 *A2, float *T)
{
    using G = FastGeom<W, H, R, NT>;
    constexpr int SEG = G::SEG, NV = (SEG + 2 * R + 2) / 2;
    const int p = threadIdx.x % G::HP, xs = threadIdx.x / G::HP;
    if (xs >= G::XS) return;
    const int x0 = xs * SEG;
    // v[j] = A2[p][PADL + x0 - R - 1 + j]; output o uses v[o + 1 + i], i = 0..2R
    const float4 *src = reinterpret_cast<const float4 *>(A2 + p * G::SP + G::PADL + x0 - R - 1);
    float2 v[2 * NV];
#pragma unroll
    for (int q = 0; q < NV; ++q) {
        const float4 t = src[q];
        v[2 * q] = make_float2(t.x, t.y);
        v[2 * q + 1] = make_float2(t.z, t.w);
    }
    const int y0 = 2 * p, y1 = y0 + 1;
    float *t0 = T + (y0 + R) * G::ST + x0, *t1 = t0 + G::ST;
    // reflected halo rows: row -r = row r (1..R), row H-1+r = row H-1-r
    const int d0 = (y0 >= 1 && y0 <= R) ? -y0 : ((y0 >= H - 1 - R && y0 <= H - 2) ? 2 * (H - 1) - y0 : -1000);
    const int d1 = (y1 >= 1 && y1 <= R) ? -y1 : ((y1 >= H - 1 - R && y1 <= H - 2) ? 2 * (H - 1) - y1 : -1000);
    float *h0 = T + (d0 + R) * G::ST + x0, *h1 = T + (d1 + R) * G::ST + x0;
#pragma unroll
    for (int o = 0; o < SEG; o += 2) {
        float2 s0 = fmul2(v[o + 1], Taps<R>::k(0)), s1 = fmul2(v[o + 2], Taps<R>::k(0));
#pragma unroll
        for (int i = 1; i < 2 * R + 1; ++i) {
            s0 = ffma2(v[o + 1 + i], Taps<R>::k(i), s0);
            s1 = ffma2(v[o + 2 + i], Taps<R>::k(i), s1);
        }
        const float2 ra = make_float2(s0.x, s1.x), rb = make_float2(s0.y, s1.y);
        *reinterpret_cast<float2 *>(t0 + o) = ra;
        *reinterpret_cast<float2 *>(t1 + o) = rb;
        if (d0 != -1000) *reinterpret_cast<float2 *>(h0 + o) = ra;
        if (d1 != -1000) *reinterpret_cast<float2 *>(h1 + o) = rb;
    }
}

// Stage 3: column pass.  Thread = (column pair c, y segment); emit(y, x, float2) gets the two
// blurred values (y, x) and (y, x+1).
template <int W, int H, int R, int NT, typename Emit>
__device__ __forceinline__ void fast_col_pass(const float *T, Emit emit)
{
    using G = FastGeom<W, H, R, NT>;
    constexpr int RS = G::RS;
    const int c = threadIdx.x % G::CP, ys = threadIdx.x / G::CP;
    if (ys >= G::YS) return;
    const float *base = T + (ys * RS) * G::ST + 2 * c;  // T row index = y + R; first needed y = ys*RS - R
    float2 t[RS + 2 * R];
#pragma unroll
    for (int i = 0; i < RS + 2 * R; ++i) t[i] = *reinterpret_cast<const float2 *>(base + i * G::ST);
#pragma unroll
    for (int o = 0; o < RS; ++o) {
        float2 s = ffma2(t[o + R], Taps<R>::k(R), make_float2(0.f, 0.f));
#pragma unroll
        for (int i = 1; i <= R; ++i) s = ffma2(fadd2(t[o + R + i], t[o + R - i]), Taps<R>::k(R + i), s);
        emit(ys * RS + o, 2 * c, s);
    }
}

// One blurred value recomputed from T with the same column-pass order (used for the 3x3
// neighbourhood of the Taylor step; T carries the reflected halo rows).
template <int W, int H, int R, int NT>
__device__ __forceinline__ float fast_col_at(const float *T, int y, int x)
{
    using G = FastGeom<W, H, R, NT>;
    const float *q = T + (y + R) * G::ST + x;
    float s = fmaf(q[0], Taps<R>::k(R), 0.0f);
#pragma unroll
    for (int i = 1; i <= R; ++i) s = fmaf(__fadd_rn(q[i * G::ST], q[-i * G::ST]), Taps<R>::k(R + i), s);
    return s;
}

}  // namespace udp
