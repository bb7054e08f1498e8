// nms.cu -- SURVEY 8(f)-3: COCO OKS rescoring and oks_nms / soft_oks_nms, batched over images.
//
// Reference: deep_hrnet/lib/dataset/coco.py:321-356 (per image: rescore every person, then NMS) and
// deep_hrnet/lib/nms/nms.py:75-94 (oks_iou), :97-124 (oks_nms), :127-136 (rescore), :139-180
// (soft_oks_nms) -- numpy loops over persons on the host, the step right after get_final_preds in
// COCO evaluation.
//
// coco_rescore_kernel: one thread per person.  Mean of the joint scores above in_vis_thre,
//   accumulated in float32 in joint order (the reference's scalar loop runs on a float32 record),
//   times the float64 box score.
// oks_nms_kernel: one CTA per image, persons of an image contiguous (CSR offsets).  Both NMS
//   flavours are the same loop: pick the best-scored live person (block arg-max; the reference's
//   argsort()[::-1] only matters through its head), record it, then every thread scores live
//   candidates against it: hard mode kills candidates with oks > thresh, soft mode decays their
//   score by exp(-oks^2 / thresh) and stops after max_dets picks.  oks_iou is reproduced operation
//   by operation: float32 differences and squares, float64 from "/ vars" on, numpy's pairwise
//   summation order (8 partial sums, then the tail) over the K or the visibility-masked joints.
//   exp() is CUDA's float64 exp (<= 1 ulp; numpy's SIMD exp is not bit-identical), so OKS values
//   agree to ~1e-16 and the kept lists agree unless an OKS lands within that of the threshold.
//   Equal scores: numpy's default argsort leaves their order unspecified; here the higher row wins.
#include <math.h>

#include "common.cuh"

namespace udp {

constexpr int NMS_THREADS = 256;
constexpr int NMS_MAXK = 32;

struct NmsParams {
    double var[NMS_MAXK];  // (sigmas * 2) ** 2
    double thresh;
    float vis;             // oks_iou's in_vis_thre (float32 compare: float32 array > python float)
    int masked;
    int soft;
    int max_dets;
    int K;
    int n_max;             // persons per image the dynamic shared memory was sized for
};

__global__ void coco_rescore_kernel(const float *__restrict__ preds, const double *__restrict__ box_score, int K, int N,
                                    float vis, double *__restrict__ score)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    const float *p = preds + (long long)n * K * 3;
    float acc = 0.0f;
    int cnt = 0;
    for (int k = 0; k < K; ++k) {
        const float t = p[3 * k + 2];
        if (t > vis) { acc = __fadd_rn(acc, t); ++cnt; }
    }
    const double mean = cnt ? (double)__fdiv_rn(acc, (float)cnt) : 0.0;
    score[n] = __dmul_rn(mean, box_score[n]);
}

// np.add.reduce over m float64 values: 0 + pairwise_sum (numpy/core/src/umath/loops_utils.h: plain
// loop below 8 elements, else 8 running sums, a fixed combine tree, then the tail).
__device__ __forceinline__ double np_sum(const double *a, int m)
{
    double res;
    if (m < 8) {
        res = -0.0;
        for (int i = 0; i < m; ++i) res = __dadd_rn(res, a[i]);
    } else {
        double r[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) r[j] = a[j];
        int i = 8;
        for (; i < m - (m & 7); i += 8) {
#pragma unroll
            for (int j = 0; j < 8; ++j) r[j] = __dadd_rn(r[j], a[i + j]);
        }
        res = __dadd_rn(__dadd_rn(__dadd_rn(r[0], r[1]), __dadd_rn(r[2], r[3])),
                        __dadd_rn(__dadd_rn(r[4], r[5]), __dadd_rn(r[6], r[7])));
        for (; i < m; ++i) res = __dadd_rn(res, a[i]);
    }
    return __dadd_rn(0.0, res);
}

__device__ __forceinline__ bool pick_better(double bv, int bi, double av, int ai)
{   // is (bv, bi) a better pick than (av, ai)?  -1 = nothing yet
    if (bi < 0) return false;
    if (ai < 0) return true;
    return bv > av || (bv == av && bi > ai);
}

__global__ void __launch_bounds__(NMS_THREADS)
oks_nms_kernel(const float *__restrict__ preds, const double *__restrict__ area, const double *__restrict__ score,
               const int32_t *__restrict__ img_off, const __grid_constant__ NmsParams P, int32_t *__restrict__ keep,
               int32_t *__restrict__ keep_cnt, double *__restrict__ oks_top)
{
    extern __shared__ __align__(16) unsigned char nms_smem[];
    __shared__ float s_g[NMS_MAXK * 3];
    __shared__ double s_rv[NMS_THREADS / 32];
    __shared__ int s_ri[NMS_THREADS / 32];
    __shared__ int s_pick;
    const int b = blockIdx.x, tid = threadIdx.x, lane = tid & 31, wid = tid >> 5;
    const int o0 = img_off[b], n = img_off[b + 1] - o0;
    const int K = P.K;
    if (n > P.n_max || n < 0) {  // offsets disagree with max_per_image: refuse, visibly
        if (tid == 0) keep_cnt[b] = -1;
        return;
    }
    double *s_score = reinterpret_cast<double *>(nms_smem);
    unsigned char *s_alive = nms_smem + (size_t)((n + 1) & ~1) * sizeof(double);
    for (int j = tid; j < n; j += NMS_THREADS) {
        s_score[j] = score[o0 + j];
        s_alive[j] = 1;
    }
    __syncthreads();
    const double eps = 2.220446049250313e-16;  // np.spacing(1)
    int cnt = 0;
    for (int it = 0; it < n; ++it) {
        if (P.soft && it >= P.max_dets) break;
        // ---- block arg-max over the live persons
        double bv = 0.0;
        int bi = -1;
        for (int j = tid; j < n; j += NMS_THREADS)
            if (s_alive[j] && (bi < 0 || s_score[j] >= bv)) { bv = s_score[j]; bi = j; }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
            if (pick_better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
        }
        if (lane == 0) { s_rv[wid] = bv; s_ri[wid] = bi; }
        __syncthreads();
        if (wid == 0) {
            bv = lane < NMS_THREADS / 32 ? s_rv[lane] : 0.0;
            bi = lane < NMS_THREADS / 32 ? s_ri[lane] : -1;
#pragma unroll
            for (int o = 4; o > 0; o >>= 1) {
                const double ov = __shfl_xor_sync(0xffffffffu, bv, o);
                const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
                if (pick_better(ov, oi, bv, bi)) { bv = ov; bi = oi; }
            }
            if (lane == 0) {
                s_pick = bi;
                if (bi >= 0) { keep[o0 + cnt] = o0 + bi; s_alive[bi] = 0; }
            }
        }
        __syncthreads();
        const int pick = s_pick;
        if (pick < 0) break;
        ++cnt;
        if (tid < 3 * K) s_g[tid] = preds[(long long)(o0 + pick) * K * 3 + tid];
        const double a_g = area[o0 + pick];
        __syncthreads();
        // ---- score the live candidates against the pick
        for (int j = tid; j < n; j += NMS_THREADS) {
            if (!s_alive[j]) continue;
            const float *d = preds + (long long)(o0 + j) * K * 3;
            const double denom = __dadd_rn(__ddiv_rn(__dadd_rn(a_g, area[o0 + j]), 2.0), eps);
            double v[NMS_MAXK];
            int m = 0;
            for (int k = 0; k < K; ++k) {
                const float dx = __fsub_rn(d[3 * k], s_g[3 * k]), dy = __fsub_rn(d[3 * k + 1], s_g[3 * k + 1]);
                const float q = __fadd_rn(__fmul_rn(dx, dx), __fmul_rn(dy, dy));
                if (P.masked && !(d[3 * k + 2] > P.vis)) continue;
                const double e = __ddiv_rn(__ddiv_rn(__ddiv_rn((double)q, P.var[k]), denom), 2.0);
                v[m++] = exp(-e);
            }
            const double ovr = m ? __ddiv_rn(np_sum(v, m), (double)m) : 0.0;
            if (it == 0 && oks_top) oks_top[o0 + j] = ovr;
            if (P.soft)
                s_score[j] = __dmul_rn(s_score[j], exp(__ddiv_rn(-__dmul_rn(ovr, ovr), P.thresh)));
            else if (ovr > P.thresh)
                s_alive[j] = 0;
        }
        __syncthreads();
    }
    if (tid == 0) keep_cnt[b] = cnt;
}

}  // namespace udp

extern "C" int udp_coco_rescore(const float *preds, const double *box_score, int K, int N, double in_vis_thre,
                                double *score, void *stream)
{
    if (N < 0 || K < 1 || K > udp::NMS_MAXK) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!preds || !box_score || !score) return UDP_EINVAL;
    udp::coco_rescore_kernel<<<(N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(preds, box_score, K, N, (float)in_vis_thre, score);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_oks_nms(const float *preds, const double *area, const double *score, const int32_t *img_off,
                           int num_images, int max_per_image, int K, double thresh, const double *sigmas_host,
                           int use_vis_thre, double in_vis_thre, int soft, int max_dets, int32_t *keep,
                           int32_t *keep_cnt, double *oks_top, void *stream)
{
    static const double coco[17] = {.26, .25, .25, .35, .35, .79, .79, .72, .72, .62, .62, 1.07, 1.07, .87, .87, .89, .89};
    if (num_images < 0 || max_per_image < 0 || K < 1 || K > udp::NMS_MAXK || max_dets < 0) return UDP_EINVAL;
    if (!sigmas_host && K != 17) return UDP_EINVAL;  // the default sigmas are COCO's 17 (nms.py:77)
    if (num_images == 0) return 0;
    if (!preds || !area || !score || !img_off || !keep || !keep_cnt) return UDP_EINVAL;
    udp::NmsParams P;
    for (int k = 0; k < udp::NMS_MAXK; ++k) P.var[k] = 1.0;
    for (int k = 0; k < K; ++k) {
        const volatile double s = sigmas_host ? sigmas_host[k] : coco[k] / 10.0;
        const volatile double t = s * 2;
        P.var[k] = t * t;
    }
    P.thresh = thresh;
    P.vis = (float)in_vis_thre;
    P.masked = use_vis_thre != 0;
    P.soft = soft != 0;
    P.max_dets = max_dets;
    P.K = K;
    P.n_max = max_per_image;
    const size_t smem = (size_t)((max_per_image + 1) & ~1) * sizeof(double) + (size_t)max_per_image + 16;
    if (smem > 200 * 1024) return UDP_EINVAL;
    if (smem > 48 * 1024) {
        cudaError_t e = cudaFuncSetAttribute(udp::oks_nms_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) return (int)e;
    }
    udp::oks_nms_kernel<<<num_images, udp::NMS_THREADS, smem, (cudaStream_t)stream>>>(preds, area, score, img_off, P, keep,
                                                                                    keep_cnt, oks_top);
    UDP_LAUNCH_CHECK();
    return 0;
}
