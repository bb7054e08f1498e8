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
#include "blur_fast.cuh"

namespace udp {

struct DecodeOffsetArgs {
    const float *out, *out_flipped;
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
        const float *a = src + (size_t)(1 + ch) * hw + (size_t)yy * w;
        const float *f = fl ? fl + (size_t)(1 + ch) * hw + (size_t)yy * w : nullptr;
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
__device__ __forceinline__ void decode_offset_generic(const DecodeOffsetArgs &p, float *smem,
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

// Fast path: compile-time map shape, packed-FP32 blur (blur_fast.cuh).  Maps holding a
// non-finite value are re-done by the generic path above (np.argmax's NaN rule).
template <int W, int H, int NT>
__global__ void __launch_bounds__(NT) decode_offset_fast_kernel(DecodeOffsetArgs p)
{
    constexpr int R = 7;
    using G = FastGeom<W, H, R, NT>;
    extern __shared__ __align__(16) float smem[];
    __shared__ float red_v[32];
    __shared__ int red_i[32];
    constexpr int hw = H * W;
    const long long map = blockIdx.x;
    const int n = (int)(map / p.K), k = (int)(map - (long long)n * p.K);
    const float *src = p.out + (map * 3) * (long long)hw;
    const int kp = p.out_flipped ? p.flip_perm[k] : k;
    const float *fl = p.out_flipped ? p.out_flipped + (((long long)n * p.K + kp) * 3) * (long long)hw : nullptr;

    float2 *A2 = reinterpret_cast<float2 *>(smem);
    float *T = smem + G::A2_BYTES / 4;
    float rv, fsum;
    int ri;
    fast_load<W, H, R, NT>(src, fl, 0.0f, A2, rv, ri, fsum);
    if (__syncthreads_or(!(fabsf(fsum) < INFINITY))) {  // NaN / inf somewhere in the map
        decode_offset_generic(p, smem, map, n, src, fl);
        return;
    }
    fast_pads<W, H, R, NT>(A2);
    __syncthreads();
    fast_row_pass<W, H, R, NT>(A2, T);
    __syncthreads();
    float v0 = -INFINITY, v1 = -INFINITY;
    int y0 = 0, y1 = 0, xc = 0;
    fast_col_pass<W, H, R, NT>(T, [&](int y, int x, float2 s) {
        xc = x;
        if (s.x > v0) { v0 = s.x; y0 = y; }
        if (s.y > v1) { v1 = s.y; y1 = y; }
    });
    float bv = v0;
    int bi = y0 * W + xc;
    if (arg_better(v1, y1 * W + xc + 1, bv, bi)) { bv = v1; bi = y1 * W + xc + 1; }
    if (threadIdx.x >= G::CP * G::YS) { bv = -INFINITY; bi = 0x7fffffff; }
    block_argmax(bv, bi, red_v, red_i);
    offset_finish(p, map, n, src, fl, bv, bi);
}

}  // namespace udp

extern "C" int udp_decode_offset(const float *out, const float *out_flipped,
                                 const int32_t *flip_perm, double kpd, const void *center,
                                 const void *scale, int cs_f64, int K, int w, int h, float *preds,
                                 float *maxvals, float *preds_in, int32_t *argmax_idx, float *packed,
                                 int N, void *stream)
{
    if (N < 0 || K < 1 || w < 8 || (w & 3) || h < 8) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!out || !preds || !maxvals || (out_flipped && !flip_perm) || ((center == nullptr) != (scale == nullptr)))
        return UDP_EINVAL;
    udp::DecodeOffsetArgs a{out, out_flipped, flip_perm, (float)kpd, center, scale, cs_f64, K, w, h,
                            preds, maxvals, preds_in, argmax_idx, packed};
    long long maps = (long long)N * K;
    if (maps > 0x7fffffffLL) return UDP_EINVAL;
    size_t gsmem = udp::BlurGeom<7>::smem_floats(w, h) * sizeof(float);
    if (gsmem > 200 * 1024) return UDP_EINVAL;
    if (!udp_force_generic()) {
#define UDP_OFF_FAST(WW, HH, NT)                                                                 \
        if (w == WW && h == HH) {                                                                \
            using G = udp::FastGeom<WW, HH, 7, NT>;                                                \
            size_t smem = G::SMEM > gsmem ? G::SMEM : gsmem;                                       \
            auto kern = udp::decode_offset_fast_kernel<WW, HH, NT>;                                \
            cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); \
            if (e != cudaSuccess) return (int)e;                                                   \
            kern<<<(unsigned)maps, NT, smem, (cudaStream_t)stream>>>(a);                           \
            UDP_LAUNCH_CHECK();                                                                    \
            return 0;                                                                              \
        }
        UDP_OFF_FAST(48, 64, 64)
        UDP_OFF_FAST(64, 64, 64)
        UDP_OFF_FAST(72, 96, 128)
#undef UDP_OFF_FAST
    }
    cudaError_t e = cudaFuncSetAttribute(udp::decode_offset_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)gsmem);
    if (e != cudaSuccess) return (int)e;
    const int threads = (h * w <= 4096) ? 128 : 256;
    udp::decode_offset_kernel<<<(unsigned)maps, threads, gsmem, (cudaStream_t)stream>>>(a);
    UDP_LAUNCH_CHECK();
    return 0;
}
