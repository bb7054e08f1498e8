// decode_offset.cu -- fused UDPv1 'offset' decoder (classification map + x/y offset voting):
//   flip-back(+negate x offsets, swap joint triplets) + average  ->  15x15 blur of the
//   classification map  ->  first-index argmax on the BLURRED map  ->  7x7-blurred x/y offsets
//   read at the argmax  ->  UDP inverse affine.
//
// Reference (paths relative to the reference root):
//   flip_back_offset             deep_hrnet/lib/utils/transforms.py:31-47
//   average                      deep_hrnet/lib/core/function.py:171
//   get_final_preds (offset)     deep_hrnet/lib/core/inference.py:156-174,176-186
//   get_max_preds                deep_hrnet/lib/core/inference.py:30-58
//   transform_preds              deep_hrnet/lib/core/inference.py:20-27
//
// The reference blurs the two offset maps in full (:166-167) but reads them at ONE pixel
// (:173-174).  A 7x7 blur at one pixel depends on a 7x7 window only, so this kernel gathers
// the two windows (and their flipped counterparts) straight from global memory after the
// argmax is known, and runs the same row-then-column FMA order on them: identical bits,
// two thirds of the input tensor never leave HBM.
#include "decode_pipe.cuh"
#include "decode_offset_prune.cuh"

namespace udp {

struct DecodeOffsetArgs {
    const float *out, *out_flipped;
    // where the 7x7 offset windows are gathered from: the same tensors, or (udp_decode_offset_split) other copies of
    // them with the same [N,3K,h,w] layout -- e.g. the caller's pinned host buffers, mapped into the device's address
    // space, when only the classification channels were uploaded
    const float *off_out, *off_flipped;
    const int32_t *flip_perm;
    float kpd;
    const void *center, *scale;
    int cs_f64;
    int K, w, h;
    float *preds, *maxvals, *preds_in;
    int32_t *argmax_idx;
    float *packed;
};

// Everything after the blurred argmax: gather the two 7x7 offset windows, blur them at the
// argmax, vote, inverse affine, store.  Called by all threads of the CTA.
__device__ __forceinline__ void offset_finish(const DecodeOffsetArgs &p, long long map, int n,
                                              const float *src, const float *fl, float bv, int bi)
{
    __shared__ float rowres[2][7];
    __shared__ float offs[2];
    const int w = p.w, h = p.h, hw = h * w;
    const bool mask = bv > 0.0f;
    const int ix = mask ? bi % w : 0, iy = mask ? bi / w : 0;

    // 7x7 blur of (offset * kpd) at (iy, ix): threads 0..13 = (channel, window row)
    if (threadIdx.x < 14) {
        const int ch = threadIdx.x / 7, i = threadIdx.x - ch * 7;
        const int yy = reflect101(iy - 3 + i, h);
        const float *a = p.off_out + (src - p.out) + (size_t)(1 + ch) * hw + (size_t)yy * w;
        const float *f = fl ? p.off_flipped + (fl - p.out_flipped) + (size_t)(1 + ch) * hw + (size_t)yy * w : nullptr;
        const float fs = ch == 0 ? -1.0f : 1.0f;  // x offsets change sign under the mirror (:40)
        float v[7];
#pragma unroll
        for (int t = 0; t < 7; ++t) {
            const int xx = reflect101(ix - 3 + t, w);
            v[t] = __ldg(a + xx);
            if (f) v[t] = __fmul_rn(__fadd_rn(v[t], fs * __ldg(f + (w - 1 - xx))), 0.5f);
            v[t] = __fmul_rn(v[t], p.kpd);  // :161-162
        }
        float s = __fmul_rn(v[0], Taps<3>::k(0));
#pragma unroll
        for (int t = 1; t < 7; ++t) s = fmaf(v[t], Taps<3>::k(t), s);
        rowres[ch][i] = s;
    }
    __syncthreads();
    if (threadIdx.x < 2) {
        const float *t = rowres[threadIdx.x];
        float s = fmaf(t[3], Taps<3>::k(3), 0.0f);
        s = fmaf(__fadd_rn(t[4], t[2]), Taps<3>::k(4), s);
        s = fmaf(__fadd_rn(t[5], t[1]), Taps<3>::k(5), s);
        s = fmaf(__fadd_rn(t[6], t[0]), Taps<3>::k(6), s);
        offs[threadIdx.x] = s;
    }
    __syncthreads();

    if (threadIdx.x == 0) {
        const float px = __fadd_rn((float)ix, offs[0]);  // :173-174
        const float py = __fadd_rn((float)iy, offs[1]);
        double ox = px, oy = py;
        if (p.center) {
            const CS cs = load_cs(p.center, p.scale, p.cs_f64, n);
            transform_pred((double)px, (double)py, cs, p.cs_f64, false, w, h, ox, oy);
        }
        p.preds[map * 2] = (float)ox;
        p.preds[map * 2 + 1] = (float)oy;
        if (p.preds_in) {
            p.preds_in[map * 2] = __fmul_rn(__fdiv_rn(px, (float)(w - 1.0)), (float)(4 * w - 1.0));
            p.preds_in[map * 2 + 1] = __fmul_rn(__fdiv_rn(py, (float)(h - 1.0)), (float)(4 * h - 1.0));
        }
        p.maxvals[map] = bv;
        if (p.argmax_idx) p.argmax_idx[map] = bi;
        if (p.packed) { float *q = p.packed + map * 3; q[0] = (float)ox; q[1] = (float)oy; q[2] = bv; }
    }
}

// Generic path: any (w, h) with w % 4 == 0, NaN-exact argmax.
__device__ __noinline__ void decode_offset_generic(const DecodeOffsetArgs &p, float *smem,
                                                      long long map, int n, const float *src,
                                                      const float *fl)
{
    constexpr int R = 7, L = 8;
    __shared__ float red_v[32];
    __shared__ int red_i[32];
    const int w = p.w, h = p.h;
    float *A = smem;
    float *T = A + (size_t)h * BlurGeom<R>::stride_a(w);
    float rv;
    int ri;
    load_tile<R>(src, fl, 1.0f, 0.0f, w, h, A, rv, ri);  // raw argmax unused in this mode
    __syncthreads();
    fill_pads<R>(A, w, h);
    __syncthreads();
    row_pass<R>(A, T, w, h);
    __syncthreads();
    float bv = -INFINITY;
    int bi = 0x7fffffff;
    col_pass<R, L>(T, w, h, [&](int y, int x, float4 s) {
        const int base = y * w + x;
        if (arg_better(s.x, base, bv, bi)) { bv = s.x; bi = base; }
        if (arg_better(s.y, base + 1, bv, bi)) { bv = s.y; bi = base + 1; }
        if (arg_better(s.z, base + 2, bv, bi)) { bv = s.z; bi = base + 2; }
        if (arg_better(s.w, base + 3, bv, bi)) { bv = s.w; bi = base + 3; }
    });
    block_argmax(bv, bi, red_v, red_i);
    offset_finish(p, map, n, src, fl, bv, bi);
}

__global__ void decode_offset_kernel(DecodeOffsetArgs p)
{
    extern __shared__ __align__(16) float smem[];
    const int hw = p.h * p.w;
    const long long map = blockIdx.x;  // n*K + k
    const int n = (int)(map / p.K), k = (int)(map - (long long)n * p.K);
    const float *src = p.out + (map * 3) * (long long)hw;  // channel (k, cls)
    const int kp = p.out_flipped ? p.flip_perm[k] : k;
    const float *fl = p.out_flipped ? p.out_flipped + (((long long)n * p.K + kp) * 3) * (long long)hw : nullptr;
    decode_offset_generic(p, smem, map, n, src, fl);
}

// The vote (:160-174) of the streaming kernel's batched tail, split in two so that every thread
// of the CTA shares the scattered loads:
//   offset_row_item  one (map, channel, window row): gather 7 (+7 flipped) offsets straight from
//                    global memory, average, scale by KPD, cv2's left-to-right row FMA chain;
//   offset_final     one map: symmetric column combine of the 2 x 7 row results, add to the
//                    integer peak, inverse affine, stores.
__device__ __forceinline__ float offset_row_item(const DecodeOffsetArgs &p, long long map, float bv, int bi,
                                                 int ch, int i)
{
    const int w = p.w, h = p.h, hw = h * w;
    const int n = (int)(map / p.K), k = (int)(map - (long long)n * p.K);
    const bool mask = bv > 0.0f;
    const int ix = mask ? bi % w : 0, iy = mask ? bi / w : 0;
    const int yo = reflect101(iy - 3 + i, h) * w;
    const float *a = p.off_out + (map * 3 + 1 + ch) * (long long)hw + yo;
    const float *f = p.out_flipped
                         ? p.off_flipped + (((long long)n * p.K + p.flip_perm[k]) * 3 + 1 + ch) * (long long)hw + yo
                         : nullptr;
    const float fs = ch == 0 ? -1.0f : 1.0f;  // x offsets change sign under the mirror (:40)
    float v[7];
#pragma unroll
    for (int t = 0; t < 7; ++t) {
        const int xx = reflect101(ix - 3 + t, w);
        v[t] = __ldg(a + xx);
        if (f) v[t] = __fmul_rn(__fadd_rn(v[t], fs * __ldg(f + (w - 1 - xx))), 0.5f);
        v[t] = __fmul_rn(v[t], p.kpd);  // :161-162
    }
    float sacc = __fmul_rn(v[0], Taps<3>::k(0));
#pragma unroll
    for (int t = 1; t < 7; ++t) sacc = fmaf(v[t], Taps<3>::k(t), sacc);
    return sacc;
}

__device__ __forceinline__ void offset_final(const DecodeOffsetArgs &p, long long map, float bv, int bi,
                                             const float *rows /* [2][7] */)
{
    const int w = p.w, h = p.h;
    const int n = (int)(map / p.K);
    const bool mask = bv > 0.0f;
    const int ix = mask ? bi % w : 0, iy = mask ? bi / w : 0;
    float offs[2];
#pragma unroll
    for (int ch = 0; ch < 2; ++ch) {
        const float *t = rows + ch * 7;
        float sacc = fmaf(t[3], Taps<3>::k(3), 0.0f);
        sacc = fmaf(__fadd_rn(t[4], t[2]), Taps<3>::k(4), sacc);
        sacc = fmaf(__fadd_rn(t[5], t[1]), Taps<3>::k(5), sacc);
        sacc = fmaf(__fadd_rn(t[6], t[0]), Taps<3>::k(6), sacc);
        offs[ch] = sacc;
    }
    const float px = __fadd_rn((float)ix, offs[0]);  // :173-174
    const float py = __fadd_rn((float)iy, offs[1]);
    double ox = px, oy = py;
    if (p.center) {
        const CS cs = load_cs(p.center, p.scale, p.cs_f64, n);
        transform_pred((double)px, (double)py, cs, p.cs_f64, false, w, h, ox, oy);
    }
    p.preds[map * 2] = (float)ox;
    p.preds[map * 2 + 1] = (float)oy;
    if (p.preds_in) {
        p.preds_in[map * 2] = __fmul_rn(__fdiv_rn(px, (float)(w - 1.0)), (float)(4 * w - 1.0));
        p.preds_in[map * 2 + 1] = __fmul_rn(__fdiv_rn(py, (float)(h - 1.0)), (float)(4 * h - 1.0));
    }
    p.maxvals[map] = bv;
    if (p.argmax_idx) p.argmax_idx[map] = bi;
    if (p.packed) { float *q = p.packed + map * 3; q[0] = (float)ox; q[1] = (float)oy; q[2] = bv; }
}

// Streaming fast path (decode_pipe.cuh): persistent CTAs, 2-stage cp.async pipeline over the
// classification maps, 15x15 packed-FP32 blur, arg-max on the blurred map.  Two barriers per
// map; the per-warp arg-max candidates go to a 32-entry mailbox.  Every 32 maps all threads
// share the 32 x 14 offset-window rows (scattered global loads, all in flight together), and
// after the next barrier one warp finishes the 32 maps, one per lane.  Maps holding a
// non-finite value are re-done by the generic path above (np.argmax's NaN rule).
template <int W, int H, int NT, int MINB, bool FLIP>
__global__ void __launch_bounds__(NT, MINB)
decode_offset_pipe_kernel(const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ CUtensorMap tm_fl,
                          DecodeOffsetArgs p, long long maps)
{
    constexpr int R = 7;
    using G = PipeGeom<W, H, R, NT>;
    constexpr int NW = G::NWARP, NBOX = 32;
    constexpr int B_V = 0, B_I = NW, B_MAP = 2 * NW, BOXW = 2 * NW + 1;
    // dynamic shared memory is only guaranteed 16-byte alignment; TMA destinations need 128
    extern __shared__ __align__(16) float smem_raw[];
    float *smem = smem_raw + ((128u - ((unsigned)__cvta_generic_to_shared(smem_raw) & 127u)) & 127u) / 4u;  // still a shared-space pointer
    __shared__ __align__(8) unsigned long long s_bars[3];
    __shared__ float s_box[NBOX * BOXW];
    __shared__ float s_rows[NBOX * 14];
    constexpr int hw = H * W;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int pr = threadIdx.x % G::HP, xs = threadIdx.x / G::HP;

    PipeFeed<W, H, R, NT> feed;
    feed.init(smem, s_bars);
    MapWalk nxt;  // (person, joint) of the map being prefetched
    nxt.init(blockIdx.x, (int)gridDim.x, p.K);
    auto src_of = [&](long long m) { return p.out + (m * 3) * (long long)hw; };
    auto row_of = [&](long long m) { return (int)(m * 3 * H); };  // first tensor row of the classification map of joint map m
    auto fl_row = [&]() {  // partner map of the map `nxt` points at
        return (int)((((long long)nxt.n * p.K + __ldg(p.flip_perm + nxt.k)) * 3) * H);
    };
    auto entry = [&](int e, float &bv, int &bi) -> long long {  // combine the per-warp candidates
        const float *b = s_box + e * BOXW;
        bv = b[B_V];
        bi = __float_as_int(b[B_I]);
#pragma unroll
        for (int q = 1; q < NW; ++q) {
            const float v = b[B_V + q];
            const int i = __float_as_int(b[B_I + q]);
            if (arg_better(v, i, bv, bi)) { bv = v; bi = i; }
        }
        return (long long)__float_as_int(b[B_MAP]);
    };
    auto flush_rows = [&](int count) {  // all threads: (map, channel, row) items
        for (int item = threadIdx.x; item < count * 14; item += NT) {
            const int e = item / 14, r = item - e * 14;
            float bv;
            int bi;
            const long long m = entry(e, bv, bi);
            s_rows[item] = offset_row_item(p, m, bv, bi, r / 7, r - (r / 7) * 7);
        }
    };
    auto flush_final = [&](int count) {  // one full warp, one map per lane
        if (lane < count) {
            float bv;
            int bi;
            const long long m = entry(lane, bv, bi);
            offset_final(p, m, bv, bi, s_rows + lane * 14);
        }
    };

    // buffer roles: bP holds out_i (then T_i), bN receives out_{i+1}; buffer 2 always holds the flipped map
    long long map = blockIdx.x;
    int bP = 0, bN = 1, slot = 0;
    if (map < maps && threadIdx.x == 0) {
        feed.issue(bP, &tm_out, row_of(map));
        if (FLIP) feed.issue(2, &tm_fl, fl_row());
    }
    for (; map < maps; map += gridDim.x, bP ^= 1, bN ^= 1) {
        float *bufP = smem + bP * G::BUF, *bufQ = smem + 2 * G::BUF;
        feed.wait(bP);  // per-thread wait: no CTA barrier is needed to read a landed map
        if (FLIP) feed.wait(2);
        const long long next = map + gridDim.x;
        nxt.advance();
        // mailbox full: everyone gathers offset rows now (one barrier so that the last entry is
        // visible); warp 0 finishes the maps after barrier (B).  The next mailbox / s_rows write
        // comes after (C), resp. the next flush.
        const bool finish = slot == NBOX;
        if (finish) { __syncthreads(); flush_rows(NBOX); slot = 0; }

        f32x2 v[G::NCOL];
        pipe_window<W, H, R, NT, FLIP, false>(bufP, bufQ, pr, xs, v);
        f32x2 acc = pk(0.f, 0.f);
#pragma unroll
        for (int j = G::PAD; j < G::PAD + G::SEG; j += 2) acc = fadd2p(acc, fadd2p(v[j], v[j + 1]));
        const bool bad = !(fabsf(lo_of(acc) + hi_of(acc)) < INFINITY);
        feed.release();  // this thread's T_{i-1} stores (in bN) are ordered before the bulk load below
        // (B) both raw buffers are consumed and everyone has left the previous column pass:
        // out_{i+1} may land in bN (it held T_{i-1}), T_i may overwrite out_i, flipped_{i+1} may land in Q
        const bool anybad = __syncthreads_or(bad);
        if (next < maps && threadIdx.x == 0) feed.issue(bN, &tm_out, row_of(next));
        if (finish && warp == 0) flush_final(NBOX);
        if (anybad) {  // NaN / inf somewhere in the map
            if (next < maps) feed.wait(bN);  // the load in flight must land before the buffers are reused
            __syncthreads();
            const int gn = (int)(map / p.K), gk = (int)(map - (long long)gn * p.K);
            const float *gfl = FLIP ? p.out_flipped + (((long long)gn * p.K + p.flip_perm[gk]) * 3) * (long long)hw
                                    : nullptr;
            decode_offset_generic(p, smem, map, gn, src_of(map), gfl);  // clobbers all three buffers
            feed.release();
            __syncthreads();
            if (next < maps && threadIdx.x == 0) {
                feed.issue(bN, &tm_out, row_of(next));
                if (FLIP) feed.issue(2, &tm_fl, fl_row());
            }
            continue;
        }
        if (FLIP && next < maps && threadIdx.x == 0) feed.issue(2, &tm_fl, fl_row());
        float *T = bufP;
        pipe_row_filter<W, H, R, NT>(v, T, pr, xs);
        __syncthreads();  // (C) T complete
        float v0 = -INFINITY, v1 = -INFINITY;
        int y0 = 0, y1 = 0;
        pipe_col_pass<W, H, R, NT>(T, [&](int y, int, float a, float b) {
            if (a > v0) { v0 = a; y0 = y; }
            if (b > v1) { v1 = b; y1 = y; }
        });
        const int xc = 2 * (threadIdx.x % G::CP);
        float bv = v0;
        int bi = y0 * W + xc;
        if (arg_better(v1, y1 * W + xc + 1, bv, bi)) { bv = v1; bi = y1 * W + xc + 1; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        float *box = s_box + slot * BOXW;
        if (lane == 0) { box[B_V + warp] = bv; box[B_I + warp] = __int_as_float(bi); }
        if (threadIdx.x == 0) box[B_MAP] = __int_as_float((int)map);
        ++slot;
    }
    __syncthreads();
    if (slot) {
        flush_rows(slot);
        __syncthreads();
        if (warp == 0) flush_final(slot);
    }
}

// Third generation (decode_offset_prune.cuh): exact branch-and-bound arg-max.  Persistent CTAs, two-stage TMA
// feed of (out_i, flipped_i), four barriers per map, the same 32-entry mailbox and batched vote as above.
template <int W, int H, int NT, int MINB, int STAGES, bool FLIP>
__global__ void __launch_bounds__(NT, MINB)
decode_offset_prune_kernel(const __grid_constant__ CUtensorMap tm_out, const __grid_constant__ CUtensorMap tm_fl,
                           DecodeOffsetArgs p, long long maps)
{
    using G = PruneGeom<W, H, NT>;
    constexpr int NW = G::NWARP, NBOX = 16, NBLK = G::NBLK, GX = G::GX, GY = G::GY;
    constexpr int PX = GX + 2, NPAD = (GY + 2) * PX;   // block tables carry a one-block border: no bounds checks
    constexpr int NSLOT = (NBLK + 31) / 32;             // blocks per lane in the warp-0 phase
    constexpr int B_V = 0, B_I = NW, B_MAP = 2 * NW, BOXW = 2 * NW + 1;
    extern __shared__ __align__(16) float smem_raw[];
    float *smem = smem_raw + ((128u - ((unsigned)__cvta_generic_to_shared(smem_raw) & 127u)) & 127u) / 4u;
    __shared__ __align__(8) unsigned long long s_bars[4];
    __shared__ float s_box[NBOX * BOXW];
    __shared__ float s_rows[NBOX * 14];
    __shared__ float s_bmax[NPAD];
    __shared__ float s_abs[NW];
    // every warp derives the lower bound, the candidate blocks and the two work lists for itself (same inputs, same
    // result): 150 redundant instructions per warp instead of three warps waiting at a barrier for the fourth
    __shared__ unsigned char s_cand_w[NW][NPAD];
    __shared__ unsigned char s_titems_w[NW][NBLK], s_citems_w[NW][NBLK];  // one entry (g << 4 | j) per block, four items each
    constexpr int hw = H * W;
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;

    unsigned char *s_cand = s_cand_w[warp];
    unsigned char *s_titems = s_titems_w[warp], *s_citems = s_citems_w[warp];
    for (int i = tid; i < NPAD; i += NT) s_bmax[i] = -INFINITY;
    for (int i = lane; i < NPAD; i += 32) s_cand[i] = 0;
    PruneFeed<W, H, NT> feed;
    feed.init(smem, s_bars);  // (ends with a barrier: the borders above are in place)
    MapWalk nxt;  // (person, joint) of the map being prefetched
    nxt.init(blockIdx.x, (int)gridDim.x, p.K);
    auto src_of = [&](long long m) { return p.out + (m * 3) * (long long)hw; };
    auto fl_row = [&]() { return (int)((((long long)nxt.n * p.K + __ldg(p.flip_perm + nxt.k)) * 3) * H); };
    auto entry = [&](int e, float &bv, int &bi) -> long long {
        const float *b = s_box + e * BOXW;
        bv = b[B_V];
        bi = __float_as_int(b[B_I]);
#pragma unroll
        for (int q = 1; q < NW; ++q) {
            const float v = b[B_V + q];
            const int i = __float_as_int(b[B_I + q]);
            if (arg_better(v, i, bv, bi)) { bv = v; bi = i; }
        }
        return (long long)__float_as_int(b[B_MAP]);
    };
    auto flush_rows = [&](int count) {
        for (int item = tid; item < count * 14; item += NT) {
            const int e = item / 14, r = item - e * 14;
            float bv;
            int bi;
            const long long m = entry(e, bv, bi);
            s_rows[item] = offset_row_item(p, m, bv, bi, r / 7, r - (r / 7) * 7);
        }
    };
    auto flush_final = [&](int count) {
        if (lane < count) {
            float bv;
            int bi;
            const long long m = entry(lane, bv, bi);
            offset_final(p, m, bv, bi, s_rows + lane * 14);
        }
    };
    auto issue_pair = [&](int st) {  // one thread: the two tensor loads of the map `nxt` points at, into stage st
        feed.issue(2 * st, &tm_out, (int)(((long long)nxt.n * p.K + nxt.k) * 3 * H));
        if (FLIP) feed.issue(2 * st + 1, &tm_fl, fl_row());
    };

    // this thread's cells (map-independent): float offsets into the raw out / raw flipped / A2 buffers of a stage
    int offP[G::CPT], offQ[G::CPT], offA[G::CPT], bix[G::CPT];
#pragma unroll
    for (int k = 0; k < G::CPT; ++k) {
        int g, j, a, cx;
        prune_cell<W, H, NT>(k * NT + tid, g, j, a, cx);
        offP[k] = (8 * g + a) * G::SR + 4 * cx;
        offQ[k] = (8 * g + a) * G::SR + (W - 4 - 4 * cx);
        offA[k] = (4 * g + a) * G::SA + 8 * cx;
        bix[k] = (g + 1) * PX + j + 1;
    }
    // the blocks of this lane (phase L), as indices into the bordered tables, and as (g << 4 | j) list codes
    int pidx[NSLOT], code[NSLOT];
    bool vld[NSLOT];
#pragma unroll
    for (int s = 0; s < NSLOT; ++s) {
        const int b = s * 32 + lane, g = b / GX, j = b - g * GX;
        vld[s] = b < NBLK;
        pidx[s] = vld[s] ? (g + 1) * PX + j + 1 : PX + 1;   // (lanes past the last block look at block 0 and are masked)
        code[s] = (g << 4) | j;
    }

    long long map = blockIdx.x;
    int st = 0, slot = 0;
    if (map < maps && tid == 0) issue_pair(0);
    // STAGES == 2: the loads of map i+1 are in flight while map i is decoded (53 KB per CTA).  STAGES == 1: no prefetch,
    // 27 KB per CTA: twice as many, smaller CTAs per SM hide the load latency and the short serial phases of each other
    for (; map < maps; map += gridDim.x, st ^= (STAGES == 2 ? 1 : 0)) {
        const int ns = STAGES == 2 ? (st ^ 1) : 0;  // stage of the next map
        float *P = smem + (2 * st) * G::BUF, *Q = P + G::BUF;
        const long long next = map + gridDim.x;
        nxt.advance();
        feed.wait(2 * st);
        if (FLIP) feed.wait(2 * st + 1);
        const bool finish = slot == NBOX;
        if (finish) { __syncthreads(); flush_rows(NBOX); slot = 0; }

        // ---- phase A: flip average of this thread's cells (registers), block maxima, max |A|, finiteness
        float e[G::CPT][8];
        float amax = 0.0f;
        f32x2 fsum = pk(0.f, 0.f);
#pragma unroll
        for (int k = 0; k < G::CPT; ++k) {
            if (G::CELLS % NT == 0 || k * NT + tid < G::CELLS) {  // (whole warps: CELLS % 32 == 0)
                const float4 r0 = *reinterpret_cast<const float4 *>(P + offP[k]);
                const float4 r1 = *reinterpret_cast<const float4 *>(P + offP[k] + 4 * G::SR);
                float *o = e[k];
                if (FLIP) {
                    const float4 g0 = *reinterpret_cast<const float4 *>(Q + offQ[k]);
                    const float4 g1 = *reinterpret_cast<const float4 *>(Q + offQ[k] + 4 * G::SR);
                    o[0] = __fmul_rn(__fadd_rn(r0.x, g0.w), 0.5f); o[1] = __fmul_rn(__fadd_rn(r1.x, g1.w), 0.5f);
                    o[2] = __fmul_rn(__fadd_rn(r0.y, g0.z), 0.5f); o[3] = __fmul_rn(__fadd_rn(r1.y, g1.z), 0.5f);
                    o[4] = __fmul_rn(__fadd_rn(r0.z, g0.y), 0.5f); o[5] = __fmul_rn(__fadd_rn(r1.z, g1.y), 0.5f);
                    o[6] = __fmul_rn(__fadd_rn(r0.w, g0.x), 0.5f); o[7] = __fmul_rn(__fadd_rn(r1.w, g1.x), 0.5f);
                } else {
                    o[0] = r0.x; o[1] = r1.x; o[2] = r0.y; o[3] = r1.y; o[4] = r0.z; o[5] = r1.z; o[6] = r0.w; o[7] = r1.w;
                }
                float m = fmax3(fmax3(o[0], o[1], o[2]), fmax3(o[3], o[4], o[5]), fmaxf(o[6], o[7]));
                amax = fmax3(amax, fmax3(fabsf(o[0]), fabsf(o[1]), fabsf(o[2])), fmax3(fabsf(o[3]), fabsf(o[4]), fabsf(o[5])));
                amax = fmax3(amax, fabsf(o[6]), fabsf(o[7]));
                fsum = fadd2p(fsum, fadd2p(fadd2p(pk(o[0], o[1]), pk(o[2], o[3])), fadd2p(pk(o[4], o[5]), pk(o[6], o[7]))));
                m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 1));
                m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 2));
                m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, 4));
                if ((lane & 7) == 0) s_bmax[bix[k]] = m;
            }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        if (lane == 0) s_abs[warp] = amax;
        const bool bad = !(fabsf(lo_of(fsum) + hi_of(fsum)) < INFINITY);
        // (1) every raw value has been read: A2 may overwrite out_i, T may overwrite flipped_i ...
        const bool anybad = __syncthreads_or(bad);
        // ... and everybody has left the column pass of map i-1, the last reader of the other stage (whose
        // generic-proxy stores were fenced before that map's barrier (4)): prefetch map i+1 into it
        if (STAGES == 2 && !anybad && next < maps && tid == 0) issue_pair(ns);
        if (finish && warp == 0) flush_final(NBOX);
        if (anybad) {  // NaN / inf somewhere in the map: the generic, NaN-exact path (clobbers all buffers; no load in flight)
            const int gn = (int)(map / p.K), gk = (int)(map - (long long)gn * p.K);
            const float *gfl = FLIP ? p.out_flipped + (((long long)gn * p.K + p.flip_perm[gk]) * 3) * (long long)hw : nullptr;
            decode_offset_generic(p, smem, map, gn, src_of(map), gfl);
            feed.release();
            __syncthreads();
            if (next < maps && tid == 0) issue_pair(ns);
            continue;
        }
#pragma unroll
        for (int k = 0; k < G::CPT; ++k) {
            if (G::CELLS % NT == 0 || k * NT + tid < G::CELLS) {
                float4 *d = reinterpret_cast<float4 *>(P + offA[k]);
                d[0] = make_float4(e[k][0], e[k][1], e[k][2], e[k][3]);
                d[1] = make_float4(e[k][4], e[k][5], e[k][6], e[k][7]);
            }
        }
        __syncthreads();  // (2) A2 complete
        // ---- every warp: lower bound L, candidate blocks, work lists
        int nT = 0, nC = 0;
        {
            float bm = -INFINITY, bmv[NSLOT];
            int bp = 0;
#pragma unroll
            for (int s = 0; s < NSLOT; ++s) {
                bmv[s] = s_bmax[pidx[s]];
                if (vld[s] && bmv[s] > bm) { bm = bmv[s]; bp = pidx[s]; }
            }
#pragma unroll
            for (int o = 16; o > 0; o >>= 1) {
                const float ov = __shfl_xor_sync(0xffffffffu, bm, o);
                const int ob = __shfl_xor_sync(0xffffffffu, bp, o);
                if (ov > bm || (ov == bm && ob < bp)) { bm = ov; bp = ob; }
            }
            const int g0 = bp / PX - 1, j0 = bp - (g0 + 1) * PX - 1;
            // the pixel of that block holding the maximum: lane -> (row pair a, column), two rows each
            const float2 pv = *reinterpret_cast<const float2 *>(P + (4 * g0 + (lane >> 3)) * G::SA + 2 * (8 * j0 + (lane & 7)));
            const float pm = fmaxf(pv.x, pv.y);
            const unsigned hit = __ballot_sync(0xffffffffu, pm == bm);   // (bm IS one of these 64 values)
            const int src = __ffs(hit) - 1;
            const int py = 8 * g0 + (src >> 3) + (__shfl_sync(0xffffffffu, (pv.y > pv.x) ? 4 : 0, src)), px = 8 * j0 + (src & 7);
            const float L = prune_blur_at<W, H, NT>(P, py, px, lane);
            float aabs = s_abs[0];
#pragma unroll
            for (int q = 1; q < NW; ++q) aabs = fmaxf(aabs, s_abs[q]);
            const float thr = L - PRUNE_EPS * aabs;
            bool cd[NSLOT];
#pragma unroll
            for (int s = 0; s < NSLOT; ++s) {  // candidate: the 3x3 block neighbourhood reaches the threshold
                const float *b = s_bmax + pidx[s];
                const float u = fmax3(fmax3(b[-PX - 1], b[-PX], b[-PX + 1]), fmax3(b[-1], bmv[s], b[1]), fmax3(b[PX - 1], b[PX], b[PX + 1]));
                cd[s] = vld[s] && u >= thr;
                if (vld[s]) s_cand[pidx[s]] = cd[s] ? 1 : 0;
            }
            __syncwarp();
            const unsigned lt = (1u << lane) - 1u;
#pragma unroll
            for (int s = 0; s < NSLOT; ++s) {
                const bool tn = vld[s] && (cd[s] || s_cand[pidx[s] - PX] || s_cand[pidx[s] + PX]);
                const unsigned mc = __ballot_sync(0xffffffffu, cd[s]), mt = __ballot_sync(0xffffffffu, tn);
                if (cd[s]) s_citems[nC + __popc(mc & lt)] = (unsigned char)code[s];
                if (tn) s_titems[nT + __popc(mt & lt)] = (unsigned char)code[s];
                nC += __popc(mc);
                nT += __popc(mt);
            }
            nT *= 4;
            nC *= 4;
            __syncwarp();
        }
        float *T = Q;
        for (int it = tid; it < nT; it += NT) {
            const int v = s_titems[it >> 2];
            prune_row_item<W, H, NT>(P, T, v >> 4, v & 15, it & 3);
        }
        feed.release();  // this thread's A2 / T stores are ordered before the bulk loads that will reuse the stage
        __syncthreads();  // (4) T complete
        float bv = -INFINITY;
        int bi = 0x7fffffff;
        for (int it = tid; it < nC; it += NT) {
            const int v = s_citems[it >> 2];
            prune_col_item<W, H, NT>(T, v >> 4, 4 * (v & 15) + (it & 3), bv, bi);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const float ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (ov > bv || (ov == bv && oi < bi)) { bv = ov; bi = oi; }
        }
        float *box = s_box + slot * BOXW;
        if (lane == 0) { box[B_V + warp] = bv; box[B_I + warp] = __int_as_float(bi); }
        if (tid == 0) box[B_MAP] = __int_as_float((int)map);
        ++slot;
        if (STAGES == 1) {  // everybody has left the column pass: the only stage may be refilled
            __syncthreads();
            if (next < maps && tid == 0) issue_pair(0);
        }
    }
    __syncthreads();
    if (slot) {
        flush_rows(slot);
        __syncthreads();
        if (warp == 0) flush_final(slot);
    }
}

}  // namespace udp

#include "decode_offset_warp.cuh"

static int decode_offset_launch(const float *out, const float *out_flipped, const float *off_out, const float *off_flipped,
                                const int32_t *flip_perm, double kpd, const void *center,
                                const void *scale, int cs_f64, int K, int w, int h, float *preds,
                                float *maxvals, float *preds_in, int32_t *argmax_idx, float *packed,
                                int N, void *stream)
{
    if (N < 0 || K < 1 || w < 8 || (w & 3) || h < 8) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!out || !preds || !maxvals || (out_flipped && !flip_perm) || ((center == nullptr) != (scale == nullptr)))
        return UDP_EINVAL;
    if ((((uintptr_t)out | (uintptr_t)out_flipped) & 15) != 0) return UDP_EINVAL;  // 16-byte vector access
    if (!off_out || (out_flipped && !off_flipped)) return UDP_EINVAL;
    udp::DecodeOffsetArgs a{out, out_flipped, off_out, off_flipped, flip_perm, (float)kpd, center, scale, cs_f64, K, w, h,
                            preds, maxvals, preds_in, argmax_idx, packed};
    long long maps = (long long)N * K;
    if (maps > 0x7fffffffLL) return UDP_EINVAL;
    size_t gsmem = udp::BlurGeom<7>::smem_floats(w, h) * sizeof(float);
    if (gsmem > 200 * 1024) return UDP_EINVAL;
    if (!udp_force_generic()) {
#define UDP_OFF_PIPE(WW, HH, NT, MINB)                                                               \
        if (w == WW && h == HH) do {                                                             \
            using G = udp::PipeGeom<WW, HH, 7, NT>;                                              \
            const size_t smem = (G::SMEM > gsmem ? G::SMEM : gsmem) + 128;  /* + alignment slack */ \
            void (*kern)(CUtensorMap, CUtensorMap, udp::DecodeOffsetArgs, long long) =           \
                out_flipped ? udp::decode_offset_pipe_kernel<WW, HH, NT, MINB, true>             \
                            : udp::decode_offset_pipe_kernel<WW, HH, NT, MINB, false>;           \
            CUtensorMap tm_out, tm_fl;  /* [3K maps per person][H][W]; the kernel loads every third map */ \
            const bool tma_ok = udp::pipe_make_map(&tm_out, out, WW, HH, G::SR, 3 * maps) &&     \
                                udp::pipe_make_map(&tm_fl, out_flipped ? out_flipped : out, WW, HH, G::SR, 3 * maps); \
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
        // third generation (exact branch-and-bound arg-max), opt-in with UDP_OFFSET_PRUNE=1: bit-exact (the decode tests
        // pass with it) and a third fewer instructions on peaked maps, but its five short phases per map leave the SM
        // idle more than the three long ones of the full-blur pipe kernel: 2.15 ms against 1.84 ms per 16 384 persons
#define UDP_OFF_PRUNE(WW, HH, NT, MINB, STG)                                                                  \
        if (w == WW && h == HH && prune_on) do {                                                 \
            using G = udp::PruneGeom<WW, HH, NT>;                                                \
            const size_t need = (size_t)2 * STG * G::BUF * sizeof(float);                        \
            const size_t smem = (need > gsmem ? need : gsmem) + 128;                             \
            void (*kern)(CUtensorMap, CUtensorMap, udp::DecodeOffsetArgs, long long) =           \
                out_flipped ? udp::decode_offset_prune_kernel<WW, HH, NT, MINB, STG, true>       \
                            : udp::decode_offset_prune_kernel<WW, HH, NT, MINB, STG, false>;     \
            CUtensorMap tm_out, tm_fl;                                                           \
            const bool tma_ok = udp::pipe_make_map(&tm_out, out, WW, HH, G::SR, 3 * maps) &&     \
                                udp::pipe_make_map(&tm_fl, out_flipped ? out_flipped : out, WW, HH, G::SR, 3 * maps); \
            if (!tma_ok) break;                                                                  \
            int per_sm = 0;                                                                      \
            int rc = udp::persistent_grid((const void *)kern, NT, smem, &per_sm);                \
            if (rc) return rc;                                                                   \
            long long grid = (long long)per_sm * udp::sm_count();                                \
            if (grid > maps) grid = maps;                                                        \
            kern<<<(unsigned)grid, NT, smem, (cudaStream_t)stream>>>(tm_out, tm_fl, a, maps);    \
            UDP_LAUNCH_CHECK();                                                                  \
            return 0;                                                                            \
        } while (0);
        const char *pe = getenv("UDP_OFFSET_PRUNE");
        const bool prune_on = pe && pe[0] == '1';
#ifndef UDP_PRUNE_NT
#define UDP_PRUNE_NT 64
#define UDP_PRUNE_MINB 6
#define UDP_PRUNE_STAGES 1
#endif
        UDP_OFF_PRUNE(48, 64, UDP_PRUNE_NT, UDP_PRUNE_MINB, UDP_PRUNE_STAGES)
        UDP_OFF_PRUNE(64, 64, UDP_PRUNE_NT, UDP_PRUNE_MINB, UDP_PRUNE_STAGES)
#undef UDP_OFF_PRUNE
        // the default for 64-row maps (decode_offset_warp.cuh): a warp owns a 16-column band through both passes; one CTA
        // per SM, NTEAM teams sharing a ring of NP raw slots.  UDP_OFFSET_PIPE=1 selects the pipe kernel below instead.
#define UDP_OFF_TEAM(WW, HH, TNT, NTEAM, NP)                                                     \
        if (w == WW && h == HH && !pipe_on) do {                                                 \
            using G = udp::PipeGeom<WW, HH, 7, TNT>;                                             \
            using WG = udp::WarpGeom<WW, HH, TNT>;                                               \
            const size_t smem = (size_t)(NP * (out_flipped ? 2 : 1) * G::BUF + NTEAM * WG::TFLOATS) * sizeof(float) + 128; \
            void (*kern)(CUtensorMap, CUtensorMap, udp::DecodeOffsetArgs, long long) =           \
                out_flipped ? udp::decode_offset_team_kernel<WW, HH, TNT, NTEAM, NP, true>       \
                            : udp::decode_offset_team_kernel<WW, HH, TNT, NTEAM, NP, false>;     \
            CUtensorMap tm_out, tm_fl;                                                           \
            const bool tma_ok = udp::pipe_make_map(&tm_out, out, WW, HH, G::SR, 3 * maps) &&     \
                                udp::pipe_make_map(&tm_fl, out_flipped ? out_flipped : out, WW, HH, G::SR, 3 * maps); \
            if (!tma_ok) break;                                                                  \
            int per_sm = 0;                                                                      \
            int rc = udp::persistent_grid((const void *)kern, TNT * NTEAM, smem, &per_sm);       \
            if (rc) return rc;                                                                   \
            long long grid = (long long)per_sm * udp::sm_count();                                \
            if (grid * NTEAM > maps) grid = (maps + NTEAM - 1) / NTEAM;                          \
            kern<<<(unsigned)grid, TNT * NTEAM, smem, (cudaStream_t)stream>>>(tm_out, tm_fl, a, maps); \
            UDP_LAUNCH_CHECK();                                                                  \
            return 0;                                                                            \
        } while (0);
        const char *pp = getenv("UDP_OFFSET_PIPE");
        const bool pipe_on = pp && pp[0] == '1';
#ifndef UDP_OFF_NTEAM
#define UDP_OFF_NTEAM 5
#define UDP_OFF_NP 4
#endif
        UDP_OFF_TEAM(48, 64, 96, UDP_OFF_NTEAM, UDP_OFF_NP)
        UDP_OFF_TEAM(64, 64, 128, 4, 3)
#undef UDP_OFF_TEAM
#ifndef UDP_OFF_NT
#define UDP_OFF_NT 96
#define UDP_OFF_MINB 4
#endif
        UDP_OFF_PIPE(48, 64, UDP_OFF_NT, UDP_OFF_MINB)
        UDP_OFF_PIPE(64, 64, 128, 3)
        UDP_OFF_PIPE(72, 96, 288, 2)
#undef UDP_OFF_PIPE
    }
    cudaError_t e = cudaFuncSetAttribute(udp::decode_offset_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsmem);
    if (e != cudaSuccess) return (int)e;
    const int threads = (h * w <= 4096) ? 128 : 256;
    udp::decode_offset_kernel<<<(unsigned)maps, threads, gsmem, (cudaStream_t)stream>>>(a);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_decode_offset(const float *out, const float *out_flipped,
                                 const int32_t *flip_perm, double kpd, const void *center,
                                 const void *scale, int cs_f64, int K, int w, int h, float *preds,
                                 float *maxvals, float *preds_in, int32_t *argmax_idx, float *packed,
                                 int N, void *stream)
{
    return decode_offset_launch(out, out_flipped, out, out_flipped, flip_perm, kpd, center, scale, cs_f64, K, w, h, preds,
                                maxvals, preds_in, argmax_idx, packed, N, stream);
}

extern "C" int udp_decode_offset_split(const float *cls_out, const float *cls_flipped, const float *off_out,
                                       const float *off_flipped, const int32_t *flip_perm, double kpd,
                                       const void *center, const void *scale, int cs_f64, int K, int w, int h,
                                       float *preds, float *maxvals, float *preds_in, int32_t *argmax_idx,
                                       float *packed, int N, void *stream)
{
    return decode_offset_launch(cls_out, cls_flipped, off_out, off_flipped, flip_perm, kpd, center, scale, cs_f64, K, w, h,
                                preds, maxvals, preds_in, argmax_idx, packed, N, stream);
}

// Host plumbing for the split call (no kernels): a strided 2-D asynchronous copy (cudaMemcpy2DAsync; kind: 1 = host to
// device, 2 = device to host) and the device-side address of a pinned, mapped host buffer.
extern "C" int udp_copy2d_async(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t height,
                                int kind, void *stream)
{
    if (!dst || !src || width > dpitch || width > spitch || (kind != 1 && kind != 2)) return UDP_EINVAL;
    if (width == 0 || height == 0) return 0;
    return (int)cudaMemcpy2DAsync(dst, dpitch, src, spitch, width, height,
                                  kind == 1 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToHost, (cudaStream_t)stream);
}

extern "C" int udp_host_device_pointer(void *host_ptr, void **device_ptr)
{
    if (!host_ptr || !device_ptr) return UDP_EINVAL;
    return (int)cudaHostGetDevicePointer(device_ptr, host_ptr, 0);
}
