// evaluate.cu -- SURVEY 8(f)-2: accuracy() / PCK on the device.
//
// Reference: deep_hrnet/lib/core/evaluate.py:16-71 (calc_dists, dist_acc, accuracy), called once
// per batch from lib/core/function.py:76-80 (train) and :187-192 (validate) on
// `output.cpu().numpy()` and `target.cpu().numpy()` -- the second per-batch device-to-host round
// trip of validate().  accuracy() = two get_max_preds (core/inference.py:30-58: flat first-index
// argmax, coordinates zeroed where maxval <= 0), a normalised distance per (joint, person) and a
// thresholded count per joint.
//
// Kernel 1 (one CTA per (person, joint)): streams the output map and the target map once with
//   16-byte loads, block-wide np.argmax-exact (value, first index) reductions, then
//   dists[c][n] = || pred/norm - target/norm ||  (float64, norm = (h, w)/10 -- the reference
//   divides x by h/10 and y by w/10, evaluate.py:55) or -1 when the target peak is not > 1 in both
//   coordinates (:21).  HBM-bound: 2*h*w*4 bytes per (person, joint), read once.
// Kernel 2 (one CTA, one warp per joint): dist_acc per joint (:31-38) and the average over the
//   joints that have any valid distance (:60-70), float64, joint order as in the reference.
#include <mutex>

#include "blur.cuh"

namespace udp {

constexpr int PCK_THREADS = 128;

__device__ __forceinline__ void pck_scan(const float *__restrict__ map, int hw, float &bv, int &bi)
{
    bv = -INFINITY;
    bi = 0x7fffffff;
    const int n4 = hw >> 2;
    for (int q0 = threadIdx.x; q0 < n4; q0 += 4 * PCK_THREADS) {  // four 16-byte loads in flight per thread
        float4 v[4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int q = q0 + u * PCK_THREADS;
            v[u] = q < n4 ? ld_stream_f4(map + 4 * q) : make_float4(-INFINITY, -INFINITY, -INFINITY, -INFINITY);
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
            const int i0 = 4 * (q0 + u * PCK_THREADS);
            if (q0 + u * PCK_THREADS < n4) {
                if (arg_better(v[u].x, i0, bv, bi)) { bv = v[u].x; bi = i0; }
                if (arg_better(v[u].y, i0 + 1, bv, bi)) { bv = v[u].y; bi = i0 + 1; }
                if (arg_better(v[u].z, i0 + 2, bv, bi)) { bv = v[u].z; bi = i0 + 2; }
                if (arg_better(v[u].w, i0 + 3, bv, bi)) { bv = v[u].w; bi = i0 + 3; }
            }
        }
    }
}

__global__ void __launch_bounds__(PCK_THREADS)
pck_dist_kernel(const float *__restrict__ output, const float *__restrict__ target, int ch_stride, int K, int w,
                int h, int N, float *__restrict__ pred, double *__restrict__ dists)
{
    __shared__ float red_v[32];
    __shared__ int red_i[32];
    const int hw = h * w;
    const long long m = blockIdx.x;  // (person, joint)
    const int n = (int)(m / K), c = (int)(m - (long long)n * K);
    const long long off = ((long long)n * K * ch_stride + (long long)c * ch_stride) * hw;
    float ov, tv;
    int oi, ti;
    pck_scan(output + off, hw, ov, oi);
    pck_scan(target + off, hw, tv, ti);
    block_argmax(ov, oi, red_v, red_i);
    block_argmax(tv, ti, red_v, red_i);
    if (threadIdx.x == 0) {
        // get_max_preds: (idx % w, floor(idx / w)) * (maxval > 0)
        const bool om = ov > 0.0f, tm = tv > 0.0f;
        const float px = om ? (float)(oi % w) : 0.0f, py = om ? (float)(oi / w) : 0.0f;
        const float tx = tm ? (float)(ti % w) : 0.0f, ty = tm ? (float)(ti / w) : 0.0f;
        pred[2 * m] = px;
        pred[2 * m + 1] = py;
        double d = -1.0;
        if (tx > 1.0f && ty > 1.0f) {
            const double nx = __ddiv_rn((double)h, 10.0), ny = __ddiv_rn((double)w, 10.0);  // np.array([h, w]) / 10
            const double dx = __dsub_rn(__ddiv_rn((double)px, nx), __ddiv_rn((double)tx, nx));
            const double dy = __dsub_rn(__ddiv_rn((double)py, ny), __ddiv_rn((double)ty, ny));
            d = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
        }
        dists[(long long)c * N + n] = d;  // [K, N], as calc_dists lays it out
    }
}

__global__ void pck_acc_kernel(const double *__restrict__ dists, int K, int N, double thr, double *__restrict__ acc,
                               double *__restrict__ avg_acc, int *__restrict__ cnt_out)
{
    __shared__ double s_acc[32];
    const int lane = threadIdx.x & 31, c = threadIdx.x >> 5;
    if (c < K) {
        int valid = 0, below = 0;
        for (int n = lane; n < N; n += 32) {
            const double d = dists[(long long)c * N + n];
            valid += d != -1.0;
            below += d != -1.0 && d < thr;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            valid += __shfl_xor_sync(0xffffffffu, valid, o);
            below += __shfl_xor_sync(0xffffffffu, below, o);
        }
        if (lane == 0) s_acc[c] = valid > 0 ? __ddiv_rn(__dmul_rn((double)below, 1.0), (double)valid) : -1.0;
    }
    __syncthreads();
    if (threadIdx.x == 0) {
        double sum = 0.0;
        int cnt = 0;
        for (int j = 0; j < K; ++j) {
            acc[j + 1] = s_acc[j];
            if (s_acc[j] >= 0.0) { sum = __dadd_rn(sum, s_acc[j]); ++cnt; }
        }
        const double avg = cnt != 0 ? __ddiv_rn(sum, (double)cnt) : 0.0;
        acc[0] = cnt != 0 ? avg : 0.0;
        *avg_acc = avg;
        *cnt_out = cnt;
    }
}

// SURVEY 8(f)-4 (first half): the demo engines' quarter-offset decoder,
// deep_hrnet/tools/infer_utils/decode.py:19-40 (get_max_preds :43-52, biased transform_preds :55-62).
// One CTA per map: the same streaming arg-max as above, then thread 0 reads the four neighbours
// (L2 hits: the map was just streamed) and applies sign(diff) * 0.25, the scale/output_size
// affine in the boxes' dtype and the truncation of `.astype(int)`.
template <bool F64>
__global__ void __launch_bounds__(PCK_THREADS)
decode_quarter_kernel(const float *__restrict__ hm, const void *__restrict__ boxes, int K, int w, int h,
                      int32_t *__restrict__ out, float *__restrict__ coords_f32, float *__restrict__ maxvals)
{
    __shared__ float red_v[32];
    __shared__ int red_i[32];
    const int hw = h * w;
    const long long m = blockIdx.x;
    const int n = (int)(m / K);
    const float *map = hm + m * hw;
    float v;
    int idx;
    pck_scan(map, hw, v, idx);
    block_argmax(v, idx, red_v, red_i);
    if (threadIdx.x != 0) return;
    const bool pos = v > 0.0f;
    const int px = pos ? idx % w : -1, py = pos ? idx / w : -1;
    float cx = (float)px, cy = (float)py;
    if (1 < px && px < w - 1 && 1 < py && py < h - 1) {
        const float dx = __fsub_rn(map[py * w + px + 1], map[py * w + px - 1]);
        const float dy = __fsub_rn(map[(py + 1) * w + px], map[(py - 1) * w + px]);
        // np.sign: -1, 0, +1, NaN for NaN
        const float sx = dx > 0.0f ? 1.0f : (dx < 0.0f ? -1.0f : (dx == 0.0f ? 0.0f : dx));
        const float sy = dy > 0.0f ? 1.0f : (dy < 0.0f ? -1.0f : (dy == 0.0f ? 0.0f : dy));
        cx = __fadd_rn(cx, __fmul_rn(sx, 0.25f));
        cy = __fadd_rn(cy, __fmul_rn(sy, 0.25f));
    }
    float tx, ty;
    if (F64) {
        const double *b = reinterpret_cast<const double *>(boxes) + 4 * (long long)n;
        const double s0 = __dmul_rn(b[2], 200.0), s1 = __dmul_rn(b[3], 200.0);
        tx = (float)__dsub_rn(__dadd_rn(__dmul_rn((double)cx, __ddiv_rn(s0, (double)w)), b[0]), __dmul_rn(s0, 0.5));
        ty = (float)__dsub_rn(__dadd_rn(__dmul_rn((double)cy, __ddiv_rn(s1, (double)h)), b[1]), __dmul_rn(s1, 0.5));
    } else {
        const float *b = reinterpret_cast<const float *>(boxes) + 4 * (long long)n;
        const float s0 = __fmul_rn(b[2], 200.0f), s1 = __fmul_rn(b[3], 200.0f);
        tx = __fsub_rn(__fadd_rn(__fmul_rn(cx, __fdiv_rn(s0, (float)w)), b[0]), __fmul_rn(s0, 0.5f));
        ty = __fsub_rn(__fadd_rn(__fmul_rn(cy, __fdiv_rn(s1, (float)h)), b[1]), __fmul_rn(s1, 0.5f));
    }
    // .astype(int): truncation; NaN / out of int32 range -> INT32_MIN (numpy gives INT64_MIN there)
    const bool okx = fabsf(tx) < 2147483520.0f, oky = fabsf(ty) < 2147483520.0f;
    out[2 * m] = okx ? (int32_t)tx : INT32_MIN;
    out[2 * m + 1] = oky ? (int32_t)ty : INT32_MIN;
    if (coords_f32) { coords_f32[2 * m] = tx; coords_f32[2 * m + 1] = ty; }
    if (maxvals) maxvals[m] = v;
}

}  // namespace udp

extern "C" int udp_accuracy_pck(const float *output, const float *target, int ch_stride, int K, int w, int h,
                                int N, double thr, float *pred, double *dists, double *acc, double *avg_acc,
                                int32_t *cnt, void *stream)
{
    if (N < 0 || K < 1 || K > 32 || w < 1 || h < 1 || ch_stride < 1 || ((w * h) & 3)) return UDP_EINVAL;
    if (!acc || !avg_acc || !cnt) return UDP_EINVAL;
    if (N > 0 && (!output || !target || !pred || !dists)) return UDP_EINVAL;
    if ((((uintptr_t)output | (uintptr_t)target) & 15) != 0) return UDP_EINVAL;  // 16-byte vector loads
    const long long maps = (long long)N * K;
    if (maps > 0x7fffffffLL) return UDP_EINVAL;
    if (maps > 0) {
        udp::pck_dist_kernel<<<(unsigned)maps, udp::PCK_THREADS, 0, (cudaStream_t)stream>>>(output, target, ch_stride, K, w, h, N,
                                                                                           pred, dists);
        UDP_LAUNCH_CHECK();
    }
    udp::pck_acc_kernel<<<1, 32 * K, 0, (cudaStream_t)stream>>>(dists, K, N, thr, acc, avg_acc, cnt);
    UDP_LAUNCH_CHECK();
    return 0;
}

// ---------------------------------------------------------------------------------------
// udp_accuracy_pck without caller-provided scratch (no counterpart in the reference; added
// for callers that cannot allocate: CUDA graph capture, serving loops, harnesses that forbid
// cudaMalloc).  The [K,N] float64 `dists` workspace lives in a file-scope device array of
// UDP_PCK_NOALLOC_MAX_DISTS doubles (4 MiB, resident on every GPU that loads the library until exit).
// In host code `g_pck_scratch` is only a host-side stand-in for the symbol: its device address is fetched with
// cudaGetSymbolAddress, once per device with that device current, and that pointer goes to the kernels.
// Every call shares the array and finds the previous call's contents in it; pck_dist_kernel overwrites every
// one of the K*N entries pck_acc_kernel reads, so nothing has to start at zero.  Two calls that overlap on the
// GPU (different streams, no ordering between them) race on it.
// ---------------------------------------------------------------------------------------
static __device__ __align__(256) double g_pck_scratch[UDP_PCK_NOALLOC_MAX_DISTS];

static int pck_scratch_ptr(double **out)
{
    static double *ptrs[64] = {nullptr};
    static std::mutex mu;
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= 64) return UDP_EINVAL;
    std::lock_guard<std::mutex> lock(mu);
    if (!ptrs[dev]) {
        void *p = nullptr;
        e = cudaGetSymbolAddress(&p, g_pck_scratch);  // the address on the CURRENT device
        if (e != cudaSuccess) return (int)e;
        if (!p) return UDP_EINVAL;
        ptrs[dev] = reinterpret_cast<double *>(p);
    }
    *out = ptrs[dev];
    return 0;
}

extern "C" int udp_accuracy_pck_noalloc(const float *output, const float *target, int ch_stride, int K, int w, int h,
                                        int N, double thr, float *pred, double *acc, double *avg_acc, int32_t *cnt,
                                        void *stream)
{
    if (N < 0 || K < 1 || K > 32 || w < 1 || h < 1 || ch_stride < 1 || ((w * h) & 3)) return UDP_EINVAL;
    if (!acc || !avg_acc || !cnt) return UDP_EINVAL;
    if (N > 0 && (!output || !target || !pred)) return UDP_EINVAL;
    if ((((uintptr_t)output | (uintptr_t)target) & 15) != 0) return UDP_EINVAL;  // 16-byte vector loads
    const long long maps = (long long)N * K;
    if (maps > (long long)UDP_PCK_NOALLOC_MAX_DISTS) return UDP_ESCRATCH;  // does not fit the built-in workspace
    double *dists = nullptr;
    const int rc = pck_scratch_ptr(&dists);
    if (rc) return rc;
    if (maps > 0) {
        udp::pck_dist_kernel<<<(unsigned)maps, udp::PCK_THREADS, 0, (cudaStream_t)stream>>>(output, target, ch_stride, K, w, h, N,
                                                                                           pred, dists);
        UDP_LAUNCH_CHECK();
    }
    udp::pck_acc_kernel<<<1, 32 * K, 0, (cudaStream_t)stream>>>(dists, K, N, thr, acc, avg_acc, cnt);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_decode_quarter(const float *heatmaps, const void *boxes, int boxes_f64, int K, int w, int h, int N,
                                  int32_t *coords, float *coords_f32, float *maxvals, void *stream)
{
    if (N < 0 || K < 1 || w < 4 || h < 4 || ((w * h) & 3)) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!heatmaps || !boxes || !coords) return UDP_EINVAL;
    if (((uintptr_t)heatmaps & 15) != 0) return UDP_EINVAL;  // 16-byte vector loads
    const long long maps = (long long)N * K;
    if (maps > 0x7fffffffLL) return UDP_EINVAL;
    if (boxes_f64)
        udp::decode_quarter_kernel<true><<<(unsigned)maps, udp::PCK_THREADS, 0, (cudaStream_t)stream>>>(heatmaps, boxes, K, w, h, coords,
                                                                                                 coords_f32, maxvals);
    else
        udp::decode_quarter_kernel<false><<<(unsigned)maps, udp::PCK_THREADS, 0, (cudaStream_t)stream>>>(heatmaps, boxes, K, w, h, coords,
                                                                                                  coords_f32, maxvals);
    UDP_LAUNCH_CHECK();
    return 0;
}
