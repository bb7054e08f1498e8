// decode.cu -- flip_back(+average) and the fused 'gaussian' decoder:
//   flip-back + average  ->  raw argmax  ->  7x7 blur + renormalise  ->  clip/log  ->
//   Taylor (DARK-style) refinement  ->  UDP inverse affine,
// one CTA per (person, joint) map, the map read from HBM exactly once.
//
// Reference (paths relative to the reference root):
//   flip_back / flip_back_offset   deep_hrnet/lib/utils/transforms.py:15-47
//   average                        deep_hrnet/lib/core/function.py:171
//   get_max_preds                  deep_hrnet/lib/core/inference.py:30-58
//   post                           deep_hrnet/lib/core/inference.py:60-145
//   get_final_preds (gaussian)     deep_hrnet/lib/core/inference.py:149-155,176-186
//   transform_preds                deep_hrnet/lib/core/inference.py:20-27
//   RSN prm glue (/255.0)          RSN/exps/RSN18.coco.e1.se.36x8x132000_prm/test.py:183-196
#include "decode_pipe.cuh"

namespace udp {

// ---------------------------------------------------------------------------------------
// a9/a10/a11 as a standalone elementwise kernel (the decoders fuse it; this one serves the
// reference's own call sites, which want the flipped-back tensor itself).
// ---------------------------------------------------------------------------------------
__global__ void flip_back_kernel(const float *__restrict__ out, const float *__restrict__ fl,
                                 const int32_t *__restrict__ perm, int K, int w, int h,
                                 int triplets, float *__restrict__ dst, long long total4)
{
    const int w4 = w >> 2, per = triplets ? 3 : 1;
    const int per_map = w4 * h;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < total4;
         i += (long long)gridDim.x * blockDim.x) {
        const long long map = i / per_map;  // index into [N, K*per]
        const int j = (int)(i - map * per_map);
        const int y = j / w4, x = (j - y * w4) << 2;
        const int C = K * per;
        const long long n = map / C;
        const int c = (int)(map - n * C);
        const int joint = c / per, t = c - joint * per;
        const long long smap = n * C + (long long)perm[joint] * per + t;
        float4 f = ld_stream_f4(fl + smap * (long long)(h * w) + (size_t)y * w + (w - 4 - x));
        const float sg = (triplets && t == 1) ? -1.0f : 1.0f;
        float4 r = make_float4(sg * f.w, sg * f.z, sg * f.y, sg * f.x);
        if (out) {
            const float4 a = ld_stream_f4(out + i * 4);
            r.x = __fmul_rn(__fadd_rn(a.x, r.x), 0.5f); r.y = __fmul_rn(__fadd_rn(a.y, r.y), 0.5f);
            r.z = __fmul_rn(__fadd_rn(a.z, r.z), 0.5f); r.w = __fmul_rn(__fadd_rn(a.w, r.w), 0.5f);
        }
        st_stream_f4(dst + i * 4, r);
    }
}

// np.clip(x, 0.001, 50) then np.log, float32 (inference.py:81-82).  The log is the correctly
// rounded float32 logarithm (evaluated in float64, rounded once).
__device__ __forceinline__ double clip_log(float v)
{
    if (v == v) v = v < 0.001f ? 0.001f : (v > 50.0f ? 50.0f : v);
    return (double)(float)log((double)v);
}

// 2x2 inverse as np.linalg.inv computes it (LAPACK gesv: LU, partial pivoting); an exactly
// zero pivot is the LinAlgError branch -> zero matrix (inference.py:126-132).
__device__ __forceinline__ void inv2x2(double a, double b, double c, double d, double *inv)
{
    double r00 = a, r01 = b, r10 = c, r11 = d;
    double e00 = 1, e01 = 0, e10 = 0, e11 = 1;
    if (fabs(c) > fabs(a)) { r00 = c; r01 = d; r10 = a; r11 = b; e00 = 0; e01 = 1; e10 = 1; e11 = 0; }
    inv[0] = inv[1] = inv[2] = inv[3] = 0.0;
    if (r00 == 0.0) return;
    const double l = __ddiv_rn(r10, r00);
    const double u11 = __dsub_rn(r11, __dmul_rn(l, r01));
    if (u11 == 0.0) return;
    {
        const double y1 = __dsub_rn(e10, __dmul_rn(l, e00));
        const double x1 = __ddiv_rn(y1, u11);
        inv[0] = __ddiv_rn(__dsub_rn(e00, __dmul_rn(r01, x1)), r00);
        inv[2] = x1;
    }
    {
        const double y1 = __dsub_rn(e11, __dmul_rn(l, e01));
        const double x1 = __ddiv_rn(y1, u11);
        inv[1] = __ddiv_rn(__dsub_rn(e01, __dmul_rn(r01, x1)), r00);
        inv[3] = x1;
    }
}

struct DecodeGaussArgs {
    const float *hm, *hm_flipped;
    const int32_t *flip_perm;
    const void *center, *scale;
    int cs_f64;
    float pre_div;
    int K, w, h;
    void *preds;
    float *maxvals;
    void *preds_in;
    int32_t *argmax_idx;
    float *packed;
    float *blurred;
};

// Final stores of one map (thread 0): preds_in (:177-179), transform_preds, maxvals, index, packed.
template <bool POST>
__device__ __forceinline__ void gauss_store(const DecodeGaussArgs &p, long long map, int n, double cx,
                                            double cy, float bv, int bi)
{
    const int w = p.w, h = p.h;
    double pin_x, pin_y, ox = cx, oy = cy;
    if (POST) {  // :178-179 on a float64 array
        pin_x = __dmul_rn(__ddiv_rn(cx, w - 1.0), 4 * w - 1.0);
        pin_y = __dmul_rn(__ddiv_rn(cy, h - 1.0), 4 * h - 1.0);
    } else {     // float32 array / python float
        pin_x = (double)__fmul_rn(__fdiv_rn((float)cx, (float)(w - 1.0)), (float)(4 * w - 1.0));
        pin_y = (double)__fmul_rn(__fdiv_rn((float)cy, (float)(h - 1.0)), (float)(4 * h - 1.0));
    }
    if (p.center) {
        const CS cs = load_cs(p.center, p.scale, p.cs_f64, n);
        transform_pred(cx, cy, cs, p.cs_f64, POST, w, h, ox, oy);
    }
    if (POST) {
        double *pr = reinterpret_cast<double *>(p.preds) + map * 2;
        pr[0] = ox; pr[1] = oy;
        if (p.preds_in) { double *q = reinterpret_cast<double *>(p.preds_in) + map * 2; q[0] = pin_x; q[1] = pin_y; }
    } else {
        float *pr = reinterpret_cast<float *>(p.preds) + map * 2;
        pr[0] = (float)ox; pr[1] = (float)oy;
        if (p.preds_in) { float *q = reinterpret_cast<float *>(p.preds_in) + map * 2; q[0] = (float)pin_x; q[1] = (float)pin_y; }
    }
    p.maxvals[map] = bv;
    if (p.argmax_idx) p.argmax_idx[map] = bi;
    if (p.packed) { float *q = p.packed + map * 3; q[0] = (float)ox; q[1] = (float)oy; q[2] = bv; }
}

// The 7-tap Taylor (DARK-style) step of post() (:83-144), executed by one full warp: lane l < 7
// supplies the blurred value of its tap through tapval(yy, xx) (coordinates already
// edge-replicated).  Returns the refined coordinates in every lane.
template <typename TapVal>
__device__ __forceinline__ void taylor_step(int ix, int iy, int w, int h, float mn, float range,
                                            float maxori, TapVal tapval, double &cx, double &cy)
{
    const int lane = threadIdx.x & 31;
    const int tdy[7] = {0, 0, 0, 1, -1, 1, -1};
    const int tdx[7] = {0, 1, -1, 0, 0, 1, -1};
    double tap = 0.0;
    if (lane < 7) {
        const int yy = min(max(iy + tdy[lane], 0), h - 1), xx = min(max(ix + tdx[lane], 0), w - 1);
        const float raw = tapval(yy, xx);
        const float nv = __fmul_rn(__fdiv_rn(__fsub_rn(raw, mn), range), maxori);  // :79
        tap = clip_log(nv);
    }
    const double I = __shfl_sync(0xffffffffu, tap, 0), Ix1 = __shfl_sync(0xffffffffu, tap, 1);
    const double Ix1_ = __shfl_sync(0xffffffffu, tap, 2), Iy1 = __shfl_sync(0xffffffffu, tap, 3);
    const double Iy1_ = __shfl_sync(0xffffffffu, tap, 4), Ix1y1 = __shfl_sync(0xffffffffu, tap, 5);
    const double Ix1_y1_ = __shfl_sync(0xffffffffu, tap, 6);
    const double dx = __dmul_rn(0.5, __dsub_rn(Ix1, Ix1_));
    const double dy = __dmul_rn(0.5, __dsub_rn(Iy1, Iy1_));
    const double dxx = __dadd_rn(__dsub_rn(Ix1, __dmul_rn(2.0, I)), Ix1_);
    const double dyy = __dadd_rn(__dsub_rn(Iy1, __dmul_rn(2.0, I)), Iy1_);
    double t = __dsub_rn(Ix1y1, Ix1);
    t = __dsub_rn(t, Iy1); t = __dadd_rn(t, I); t = __dadd_rn(t, I);
    t = __dsub_rn(t, Ix1_); t = __dsub_rn(t, Iy1_); t = __dadd_rn(t, Ix1_y1_);
    const double dxy = __dmul_rn(0.5, t);
    double inv[4];
    inv2x2(dxx, dxy, dxy, dyy, inv);
    cx = __dsub_rn((double)ix, __dadd_rn(__dmul_rn(inv[0], dx), __dmul_rn(inv[1], dy)));
    cy = __dsub_rn((double)iy, __dadd_rn(__dmul_rn(inv[2], dx), __dmul_rn(inv[3], dy)));
}

// The same step done by ONE thread from the seven blurred tap values (order: centre, x+1, x-1,
// y+1, y-1, (x+1,y+1), (x-1,y-1)); used by the batched tail of the streaming kernel.
__device__ __forceinline__ void taylor_scalar(const float *tap, int ix, int iy, float mn, float range,
                                              float maxori, double &cx, double &cy)
{
    double I[7];
#pragma unroll
    for (int t = 0; t < 7; ++t)
        I[t] = clip_log(__fmul_rn(__fdiv_rn(__fsub_rn(tap[t], mn), range), maxori));  // :79, :81-82
    const double dx = __dmul_rn(0.5, __dsub_rn(I[1], I[2]));
    const double dy = __dmul_rn(0.5, __dsub_rn(I[3], I[4]));
    const double dxx = __dadd_rn(__dsub_rn(I[1], __dmul_rn(2.0, I[0])), I[2]);
    const double dyy = __dadd_rn(__dsub_rn(I[3], __dmul_rn(2.0, I[0])), I[4]);
    double t = __dsub_rn(I[5], I[1]);
    t = __dsub_rn(t, I[3]); t = __dadd_rn(t, I[0]); t = __dadd_rn(t, I[0]);
    t = __dsub_rn(t, I[2]); t = __dsub_rn(t, I[4]); t = __dadd_rn(t, I[6]);
    const double dxy = __dmul_rn(0.5, t);
    double inv[4];
    inv2x2(dxx, dxy, dxy, dyy, inv);
    cx = __dsub_rn((double)ix, __dadd_rn(__dmul_rn(inv[0], dx), __dmul_rn(inv[1], dy)));
    cy = __dsub_rn((double)iy, __dadd_rn(__dmul_rn(inv[2], dx), __dmul_rn(inv[3], dy)));
}

// Generic path for one map: any (w, h) with w % 4 == 0, NaN-exact, optional blurred write-back.
// Called by all threads of the CTA; smem must hold BlurGeom<3>::smem_floats(w, h) floats.
template <bool POST>
__device__ __noinline__ void decode_gauss_generic_map(const DecodeGaussArgs &p, float *smem, long long map)
{
    constexpr int R = 3, L = 8;
    __shared__ float red_v[32];
    __shared__ int red_i[32];
    __shared__ float red_mn[32], red_mx[32];
    __shared__ int red_nan[32];
    __shared__ float nb[9];

    const int w = p.w, h = p.h, hw = h * w;
    const int n = (int)(map / p.K), k = (int)(map - (long long)n * p.K);
    float *A = smem;
    float *T = A + (size_t)h * BlurGeom<R>::stride_a(w);

    const float *src = p.hm + map * (long long)hw;
    const float *fl = p.hm_flipped ? p.hm_flipped + ((long long)n * p.K + p.flip_perm[k]) * (long long)hw : nullptr;

    float bv;
    int bi;
    load_tile<R>(src, fl, 1.0f, p.pre_div, w, h, A, bv, bi);
    __syncthreads();
    if (POST) fill_pads<R>(A, w, h);
    block_argmax(bv, bi, red_v, red_i);  // contains the __syncthreads that publishes the pads

    // get_max_preds: x = idx % w, y = floor(idx / w); coords zeroed unless maxval > 0
    const bool mask = bv > 0.0f;
    const int ix = mask ? bi % w : 0, iy = mask ? bi / w : 0;
    double cx = (double)ix, cy = (double)iy;

    if (POST) {
        row_pass<R>(A, T, w, h);
        __syncthreads();
        float mn = INFINITY, mx = -INFINITY;
        int has_nan = 0;
        float *B = A;  // A is dead after the row pass; reused for the blurred map if requested
        const bool keep = p.blurred != nullptr;
        col_pass<R, L>(T, w, h, [&](int y, int x, float4 s) {
            mn = fminf(fminf(mn, fminf(s.x, s.y)), fminf(s.z, s.w));
            mx = fmaxf(fmaxf(mx, fmaxf(s.x, s.y)), fmaxf(s.z, s.w));
            has_nan |= (s.x != s.x) | (s.y != s.y) | (s.z != s.z) | (s.w != s.w);
            const int dy = y - iy + 1;
            if (dy >= 0 && dy <= 2) {
                const float vv[4] = {s.x, s.y, s.z, s.w};
#pragma unroll
                for (int c = 0; c < 4; ++c) {
                    const int dx = x + c - ix + 1;
                    if (dx >= 0 && dx <= 2) nb[dy * 3 + dx] = vv[c];
                }
            }
            if (keep) *reinterpret_cast<float4 *>(B + y * w + x) = s;
        });
        // block min / max (np.min / np.max propagate NaN)
        const unsigned lanes = warp_lanes();
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            mn = fminf(mn, shfl_xor_or_self(lanes, mn, o));
            mx = fmaxf(mx, shfl_xor_or_self(lanes, mx, o));
            has_nan |= shfl_xor_or_self(lanes, has_nan, o);
        }
        const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
        if (lane == 0) { red_mn[warp] = mn; red_mx[warp] = mx; red_nan[warp] = has_nan; }
        __syncthreads();
        mn = red_mn[0]; mx = red_mx[0]; has_nan = red_nan[0];
        for (int q = 1; q < nwarps; ++q) {
            mn = fminf(mn, red_mn[q]); mx = fmaxf(mx, red_mx[q]); has_nan |= red_nan[q];
        }
        if (has_nan) mn = mx = __int_as_float(0x7fc00000);
        const float range = __fsub_rn(mx, mn);
        const float maxori = bv;  // np.max of the raw map (:74)

        if (keep) {  // post() writes the renormalised blurred map back (:79-80)
            float *dst = p.blurred + map * (long long)hw;
            for (int j = threadIdx.x; j < hw / 4; j += blockDim.x) {
                float4 s = *reinterpret_cast<const float4 *>(B + j * 4);
                s.x = __fmul_rn(__fdiv_rn(__fsub_rn(s.x, mn), range), maxori);
                s.y = __fmul_rn(__fdiv_rn(__fsub_rn(s.y, mn), range), maxori);
                s.z = __fmul_rn(__fdiv_rn(__fsub_rn(s.z, mn), range), maxori);
                s.w = __fmul_rn(__fdiv_rn(__fsub_rn(s.w, mn), range), maxori);
                st_stream_f4(dst + j * 4, s);
            }
        }

        if (threadIdx.x < 32)
            taylor_step(ix, iy, w, h, mn, range, maxori,
                        [&](int yy, int xx) { return nb[(yy - iy + 1) * 3 + (xx - ix + 1)]; }, cx, cy);
    }

    if (threadIdx.x == 0) gauss_store<POST>(p, map, n, cx, cy, bv, bi);
    __syncthreads();  // the static scratch above may be reused by the caller's next map
}

template <bool POST>
__global__ void decode_gaussian_kernel(DecodeGaussArgs p)
{
    extern __shared__ __align__(16) float smem[];
    decode_gauss_generic_map<POST>(p, smem, blockIdx.x);
}

// ---------------------------------------------------------------------------------------
// Streaming fast path (decode_pipe.cuh): persistent CTAs, 2-stage cp.async pipeline.
// ---------------------------------------------------------------------------------------
// Per map only two barriers remain: (B) raw stage consumed / warp maxima published, (C) T
// complete (a landed map is read after a per-thread mbarrier wait, no CTA barrier).  Everything that is serial per map -- min/max combine, the float64
// log + Taylor + 2x2 solve, the inverse affine and the result stores -- is deferred: each map
// leaves 7 tap values, per-warp min/max, maxval and index in a shared-memory mailbox, and every 32
// maps the 32 lanes of one warp finish one mailbox entry each (first build: that tail ran on 7
// lanes per map while the other warps waited at the next barrier -- 37 % of stall samples were
// barrier stalls, 19 % of all instructions were the tail).
// NBUF == 3: out_i / T_i, out_{i+1}, flipped.  NBUF == 2 (large maps): no buffer for out_{i+1} -- it is fetched into the
// only `out` buffer once the column pass has left it; one more CTA per SM (three of 58 KB at 96x72 instead of two of 88 KB)
// covers the exposed load.
template <int W, int H, int NT, int MINB, bool POST, bool FLIP, bool DIV255, int NBUF>
__global__ void __launch_bounds__(NT, MINB)
decode_gaussian_pipe_kernel(const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ CUtensorMap tm_fl,
                            DecodeGaussArgs p, long long maps)
{
    constexpr int R = 3;
    using G = PipeGeom<W, H, R, NT>;
    constexpr int NW = G::NWARP;
    constexpr int B_MN = 7, B_MX = 7 + NW, B_BV = 7 + 2 * NW, B_BI = B_BV + 1, B_MAP = B_BV + 2;
    constexpr int BOXW = ((B_MAP + 1) | 1);  // odd stride: conflict-free entry-per-thread reads
    // dynamic shared memory is only guaranteed 16-byte alignment; TMA destinations need 128
    extern __shared__ __align__(16) float smem_raw[];
    float *smem = smem_raw + ((128u - ((unsigned)__cvta_generic_to_shared(smem_raw) & 127u)) & 127u) / 4u;  // still a shared-space pointer
    __shared__ __align__(8) unsigned long long s_bars[3];
    __shared__ float s_wmax[NW];
    __shared__ int s_idx[2];  // first index of the maximum, one slot per raw buffer
    constexpr int NBOX = 32;
    __shared__ float s_box[NBOX * BOXW];
    constexpr int hw = H * W;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pr = threadIdx.x % G::HP, xs = threadIdx.x / G::HP;
    constexpr bool has_fl = FLIP;
    // NT need not be a multiple of 32 (144 at 96x72): the last warp is then partial.  Shuffles name the lanes that
    // exist, and a butterfly partner that does not exist contributes the lane's own value.
    const unsigned wmask = (NT % 32 != 0 && warp == NW - 1) ? ((1u << (NT % 32)) - 1u) : 0xffffffffu;
    auto xor_or_self = [&](float v, int o) {
        const float u = __shfl_xor_sync(wmask, v, o);
        return (NT % 32 == 0 || ((wmask >> (lane ^ o)) & 1u)) ? u : v;
    };

    PipeFeed<W, H, R, NT> feed;
    feed.init(smem, s_bars);
    MapWalk nxt;  // (person, joint) of the map being prefetched
    nxt.init(blockIdx.x, (int)gridDim.x, p.K);
    auto row_of = [&](long long m) { return (int)(m * H); };  // first tensor row of map m
    auto fl_row = [&]() {  // partner map of the map `nxt` points at
        return (int)(((long long)nxt.n * p.K + __ldg(p.flip_perm + nxt.k)) * H);
    };
    auto flush = [&](int count) {  // executed by one full warp
        if (lane < count) {
            const float *b = s_box + lane * BOXW;
            const float bv = b[B_BV];
            const int bi = __float_as_int(b[B_BI]);
            const long long map = __float_as_int(b[B_MAP]);
            const bool mask = bv > 0.0f;
            const int ix = mask ? bi % W : 0, iy = mask ? bi / W : 0;
            double cx = (double)ix, cy = (double)iy;
            if (POST) {
                float mn = b[B_MN], mx = b[B_MX];
#pragma unroll
                for (int q = 1; q < NW; ++q) { mn = fminf(mn, b[B_MN + q]); mx = fmaxf(mx, b[B_MX + q]); }
                taylor_scalar(b, ix, iy, mn, __fsub_rn(mx, mn), bv, cx, cy);
            }
            gauss_store<POST>(p, map, (int)(map / p.K), cx, cy, bv, bi);
        }
    };

    // buffer roles: bP holds out_i (then T_i), bN receives out_{i+1}; buffer QB always holds the flipped map
    constexpr int QB = NBUF - 1;
    long long map = blockIdx.x;
    int bP = 0, bN = NBUF == 3 ? 1 : 0, slot = 0;
    if (map < maps && threadIdx.x == 0) {
        feed.issue(bP, &tm_out, row_of(map));
        if (has_fl) feed.issue(QB, &tm_fl, fl_row());
        s_idx[bP] = 0x7fffffff;  // visible to the first atomicMin through barrier (B)
    }
    for (; map < maps; map += gridDim.x, bP ^= (NBUF == 3 ? 1 : 0), bN ^= (NBUF == 3 ? 1 : 0)) {
        float *bufP = smem + bP * G::BUF, *bufQ = smem + QB * G::BUF;
        feed.wait(bP);  // per-thread wait: no CTA barrier is needed to read a landed map
        if (has_fl) feed.wait(QB);
        const long long next = map + gridDim.x;
        nxt.advance();
        // mailbox full: one warp finishes the 32 queued maps (one barrier so that the last entry is
        // visible).  The next mailbox write comes after barrier (C) of this iteration, which this
        // warp has to pass too.
        if (slot == NBOX) {
            __syncthreads();
            if (warp == 0) flush(NBOX);
            slot = 0;
        }

        f32x2 v[G::NCOL];
        pipe_window<W, H, R, NT, FLIP, DIV255>(bufP, bufQ, pr, xs, v);
        float fs;
        const float lm = pipe_local_max<W, H, R, NT>(v, fs);
        float wm = lm;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) wm = fmaxf(wm, xor_or_self(wm, o));
        if (lane == 0) s_wmax[warp] = wm;
        const bool bad = !(fabsf(fs) < INFINITY);
        feed.release();  // this thread's T_{i-1} stores (in bN) are ordered before the bulk load below
        // (B) both raw buffers are consumed and everyone has left the previous column pass:
        // out_{i+1} may land in bN (it held T_{i-1}), T_i may overwrite out_i, flipped_{i+1} may land in Q
        const bool anybad = __syncthreads_or(bad);
        if (NBUF == 3 && threadIdx.x == 0) {
            if (next < maps) feed.issue(bN, &tm_out, row_of(next));
            s_idx[bN] = 0x7fffffff;  // the next map's slot; its last reader passed (B) with us
        }
        if (anybad) {  // a non-finite value: NaN-exact generic path for this map
            if (NBUF == 3 && next < maps) feed.wait(bN);  // the load in flight must land before the buffers are reused
            __syncthreads();
            decode_gauss_generic_map<POST>(p, smem, map);  // clobbers all the buffers
            feed.release();
            __syncthreads();
            if (threadIdx.x == 0) {
                if (NBUF == 2) s_idx[0] = 0x7fffffff;
                if (next < maps) {
                    feed.issue(bN, &tm_out, row_of(next));
                    if (has_fl) feed.issue(QB, &tm_fl, fl_row());
                }
            }
            continue;
        }
        if (has_fl && next < maps && threadIdx.x == 0) feed.issue(QB, &tm_fl, fl_row());
        float M = s_wmax[0];
#pragma unroll
        for (int q = 1; q < NW; ++q) M = fmaxf(M, s_wmax[q]);
        if (lm == M) atomicMin(&s_idx[bP], pipe_first_index<W, H, R, NT>(v, M, pr, xs));
        float *T = bufP;
        if (POST) pipe_row_filter<W, H, R, NT>(v, T, pr, xs);
        __syncthreads();  // (C) T complete, s_idx final
        float *box = s_box + slot * BOXW;
        const int bi = s_idx[bP];
        if (POST) {
            if (warp == NW - 1 && lane < 7) {  // the seven taps of the Taylor step (:100-109), edge-replicated
                const int tdy[7] = {0, 0, 0, 1, -1, 1, -1};
                const int tdx[7] = {0, 1, -1, 0, 0, 1, -1};
                const bool mask = M > 0.0f;
                const int ix = mask ? bi % W : 0, iy = mask ? bi / W : 0;
                const int yy = min(max(iy + tdy[lane], 0), H - 1), xx = min(max(ix + tdx[lane], 0), W - 1);
                box[lane] = pipe_col_at<W, H, R, NT>(T, yy, xx);
            }
            float mn = INFINITY, mx = -INFINITY;
            pipe_col_pass<W, H, R, NT>(T, [&](int, int, float a, float b) {
                mn = fmin3(mn, a, b);
                mx = fmax3(mx, a, b);
            });
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                mn = fminf(mn, xor_or_self(mn, o));
                mx = fmaxf(mx, xor_or_self(mx, o));
            }
            if (lane == 0) { box[B_MN + warp] = mn; box[B_MX + warp] = mx; }
        }
        if (threadIdx.x == 0) {
            box[B_BV] = M;
            box[B_BI] = __int_as_float(bi);
            box[B_MAP] = __int_as_float((int)map);
        }
        ++slot;
        if (NBUF == 2) {  // (D) everybody has left the column pass: out_{i+1} may land where T_i was
            feed.release();
            __syncthreads();
            if (threadIdx.x == 0) {
                if (next < maps) feed.issue(0, &tm_out, row_of(next));
                s_idx[0] = 0x7fffffff;
            }
        }
    }
    __syncthreads();
    if (slot && warp == 0) flush(slot);
}

template <int W, int H, int NT, int MINB, int NBUF>
static void (*pick_gauss_pipe(bool post, bool flip, bool div))(CUtensorMap, CUtensorMap, DecodeGaussArgs, long long)
{
#define UDP_PICK(P, F, D) if (post == P && flip == F && div == D) return decode_gaussian_pipe_kernel<W, H, NT, MINB, P, F, D, NBUF>;
    UDP_PICK(true, true, false) UDP_PICK(true, false, false) UDP_PICK(true, true, true) UDP_PICK(true, false, true)
    UDP_PICK(false, true, false) UDP_PICK(false, false, false) UDP_PICK(false, true, true)
#undef UDP_PICK
    return decode_gaussian_pipe_kernel<W, H, NT, MINB, false, false, true, NBUF>;
}

}  // namespace udp

extern "C" int udp_flip_back(const float *out, const float *flipped, const int32_t *flip_perm,
                             int K, int w, int h, int triplets, float *dst, int N, void *stream)
{
    if (N < 0 || K < 1 || w < 4 || (w & 3) || h < 1) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!flipped || !flip_perm || !dst || dst == flipped) return UDP_EINVAL;
    if ((((uintptr_t)out | (uintptr_t)flipped | (uintptr_t)dst) & 15) != 0) return UDP_EINVAL;
    long long total4 = (long long)N * K * (triplets ? 3 : 1) * h * (w >> 2);
    long long blocks = (total4 + 255) / 256;
    if (blocks > 148 * 32) blocks = 148 * 32;
    udp::flip_back_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(
        out, flipped, flip_perm, K, w, h, triplets, dst, total4);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_decode_gaussian(const float *hm, const float *hm_flipped,
                                   const int32_t *flip_perm, const void *center, const void *scale,
                                   int cs_f64, float pre_div, int post_process, int K, int w, int h,
                                   void *preds, float *maxvals, void *preds_in, int32_t *argmax_idx,
                                   float *packed, float *blurred, int N, void *stream)
{
    if (N < 0 || K < 1 || w < 8 || (w & 3) || h < 8) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!hm || !preds || !maxvals || (hm_flipped && !flip_perm) || ((center == nullptr) != (scale == nullptr)))
        return UDP_EINVAL;
    if (blurred && !post_process) return UDP_EINVAL;
    if ((((uintptr_t)hm | (uintptr_t)hm_flipped | (uintptr_t)blurred) & 15) != 0) return UDP_EINVAL;  // 16-byte vector access
    udp::DecodeGaussArgs a{hm, hm_flipped, flip_perm, center, scale, cs_f64, pre_div, K, w, h,
                           preds, maxvals, preds_in, argmax_idx, packed, blurred};
    long long maps = (long long)N * K;
    if (maps > 0x7fffffffLL) return UDP_EINVAL;
    size_t gsmem = udp::BlurGeom<3>::smem_floats(w, h) * sizeof(float);
    if (gsmem > 200 * 1024) return UDP_EINVAL;

    // Streaming fast path: the reference's map shapes, plain or /255 (RSN) pre-scaling, no
    // blurred write-back.  Everything else (and every map holding a NaN/inf) runs the generic code.
    if (!udp_force_generic() && !blurred && (pre_div == 0.0f || pre_div == 255.0f)) {
#define UDP_GAUSS_PIPE(WW, HH, NT, MINB, NBUF)                                                          \
        if (w == WW && h == HH) do {                                                             \
            using G = udp::PipeGeom<WW, HH, 3, NT>;                                              \
            const size_t need = (size_t)NBUF * G::BUF * sizeof(float);                           \
            const size_t smem = (need > gsmem ? need : gsmem) + 128;  /* + alignment slack */    \
            void (*kern)(CUtensorMap, CUtensorMap, udp::DecodeGaussArgs, long long) =                   \
                udp::pick_gauss_pipe<WW, HH, NT, MINB, NBUF>(post_process != 0, hm_flipped != nullptr, pre_div != 0.0f); \
            CUtensorMap tm_out, tm_fl;                                                           \
            const bool tma_ok = udp::pipe_make_map(&tm_out, hm, WW, HH, G::SR, maps) &&          \
                                udp::pipe_make_map(&tm_fl, hm_flipped ? hm_flipped : hm, WW, HH, G::SR, maps); \
            if (!tma_ok) break;  /* no tensor-map encoder: the generic kernel below */          \
            int per_sm = 0;                                                                      \
            int rc = udp::persistent_grid((const void *)kern, NT, smem, &per_sm);                \
            if (rc) return rc;                                                                   \
            long long grid = (long long)per_sm * udp::sm_count();                                \
            if (grid > maps) grid = maps;                                                        \
            kern<<<(unsigned)grid, NT, smem, (cudaStream_t)stream>>>(tm_out, tm_fl, a, maps);    \
            UDP_LAUNCH_CHECK();                                                                  \
            return 0;                                                                            \
        } while (0);
#ifndef UDP_GAUSS_MINB
#define UDP_GAUSS_MINB 4
#endif
#ifndef UDP_GAUSS_BIG_NBUF
#define UDP_GAUSS_BIG_NBUF 2
#define UDP_GAUSS_BIG_MINB 3
#endif
#ifndef UDP_GAUSS_NBUF
#define UDP_GAUSS_NBUF 3
#endif
#ifndef UDP_GAUSS_SQ_MINB
#define UDP_GAUSS_SQ_MINB 3
#endif
        UDP_GAUSS_PIPE(48, 64, 96, UDP_GAUSS_MINB, UDP_GAUSS_NBUF)
        UDP_GAUSS_PIPE(64, 64, 128, UDP_GAUSS_SQ_MINB, UDP_GAUSS_NBUF)
#ifndef UDP_GAUSS_BIG_NT
#define UDP_GAUSS_BIG_NT 144   // 144 threads x 3 CTAs x 2 buffers: 2.99 ms per 16 384 persons; 288 x 2 x 3: 3.55; 288 x 3 x 2 (spills): 3.89
#endif
        UDP_GAUSS_PIPE(72, 96, UDP_GAUSS_BIG_NT, UDP_GAUSS_BIG_MINB, UDP_GAUSS_BIG_NBUF)
#undef UDP_GAUSS_PIPE
    }
    auto kern = post_process ? udp::decode_gaussian_kernel<true> : udp::decode_gaussian_kernel<false>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsmem);
    if (e != cudaSuccess) return (int)e;
    const int threads = (h * w <= 4096) ? 128 : 256;
    kern<<<(unsigned)maps, threads, gsmem, (cudaStream_t)stream>>>(a);
    UDP_LAUNCH_CHECK();
    return 0;
}
