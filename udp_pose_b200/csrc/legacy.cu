// legacy.cu -- SURVEY 8(f)-4 (second half): RSN's legacy (biased) decoder.
//
// Reference: get_results of RSN/exps/RSN18.coco/test.py:25-71 (the same function in the Res18.coco,
// 4XRSN18.coco and RSN18.mpii experiments), called from compute_on_dataset (:74-103) after the flip
// average (:88-96, RSN/lib/utils/transforms.py:59-67).  Per map the reference copies the float32
// network output into a zero float64 array with a 10-pixel border, runs cv2.GaussianBlur
// (kernel cfg.TEST.GAUSSIAN_KERNEL: 5 for COCO, 9 for MPII, RSN/dataset/attribute.py:49,92; 11 is
// the function's default), takes the arg-max, zeroes it, takes the arg-max again, moves a quarter
// pixel from the first towards the second, clamps, maps `x*4+2` through the biased affine and
// reads the score (out/255 + 0.5) at the rounded position.
//
// One CTA per map.  The (flip-averaged) float32 map is staged in shared memory inside a zero frame,
// the row pass writes float64 rows, the column pass is evaluated on the fly over the whole padded
// domain and reduced to the two best (value, first index) candidates -- the blurred map is never
// stored, and "zero the maximum, search again" becomes a comparison of the runner-up with a zero
// at the winner's index.  The blur reproduces OpenCV's CV_64F path operation by operation (probed
// against cv2, tests/test_rsn_legacy.py): row filter = k[0]*x[0] then a FUSED left-to-right chain,
// column filter = k[c]*x[c] then separately rounded + k[c+i]*(x[c+i] + x[c-i]); taps that fall on
// the zero frame are multiplied and added like any other, so signed zeros come out the same.
// FP64-bound (2n fused/unfused multiply-adds per padded element), not HBM-bound.  Maps with NaN are
// not supported (np.argmax's NaN rule is not reproduced here).
#include <math.h>

#include "common.cuh"

namespace udp {

constexpr int LG_THREADS = 256;
constexpr int LG_BORDER = 10;   // test.py:36
constexpr int LG_MAXK = 15;

struct LegacyArgs {
    const float *out, *out_flipped;
    const int32_t *flip_perm;
    const float *center, *scale;
    double *preds, *maxvals;
    double k[LG_MAXK];
    double shift;
    int K, w, h, in_w, in_h, n;
};

// (va, ia) ranks before (vb, ib): larger value, then lower index (np.argmax's first-index rule)
__device__ __forceinline__ bool lg_before(double va, int ia, double vb, int ib) { return va > vb || (va == vb && ia < ib); }

struct Top2 {
    double v1, v2;
    int i1, i2;
    __device__ __forceinline__ void init() { v1 = v2 = -INFINITY; i1 = i2 = 0x7fffffff; }
    __device__ __forceinline__ void push(double v, int i)
    {
        if (lg_before(v, i, v1, i1)) { v2 = v1; i2 = i1; v1 = v; i1 = i; }
        else if (lg_before(v, i, v2, i2)) { v2 = v; i2 = i; }
    }
};

// NK = the kernel size when known at compile time (5: COCO, 9: MPII, 11: the function's default), 0 = p.n
template <int NK>
__global__ void __launch_bounds__(LG_THREADS)
rsn_get_results_kernel(const __grid_constant__ LegacyArgs p)
{
    extern __shared__ __align__(16) unsigned char lg_smem[];
    __shared__ Top2 s_red[LG_THREADS / 32];
    const int w = p.w, h = p.h, n = NK ? NK : p.n, r = n >> 1, B = LG_BORDER;
    const int PW = w + 2 * B, PH = h + 2 * B, SW = PW + 2 * r;
    double *T0 = reinterpret_cast<double *>(lg_smem);            // [2r + h + 2r][PW] row-pass result inside 2r zero rows
    double *T = T0 + 2 * r * PW;                                 // row 0 of the map
    float *S = reinterpret_cast<float *>(T0 + (size_t)(h + 4 * r) * PW);  // [h][SW] the map inside its zero frame
    const long long m = blockIdx.x;
    const int person = (int)(m / p.K), joint = (int)(m - (long long)person * p.K);
    const float *src = p.out + m * (long long)(h * w);
    const float *fl = p.out_flipped ? p.out_flipped + ((long long)person * p.K + p.flip_perm[joint]) * (long long)(h * w) : nullptr;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int NWARP = LG_THREADS / 32;
    for (int i = threadIdx.x; i < 2 * r * PW; i += LG_THREADS) { T0[i] = 0.0; T[h * PW + i] = 0.0; }
    const bool vec = (w & 3) == 0 && (((uintptr_t)src | (uintptr_t)fl) & 15) == 0;
    if (vec) {
        // zero frame left and right of every row, then the map itself with 16-byte loads, four in flight per
        // thread (the scalar version of this loop held 44 % of the kernel's stall samples on its global loads)
        const int fwid = SW - w, lead = r + B;
        for (int i = threadIdx.x; i < h * fwid; i += LG_THREADS) {
            const int y = i / fwid, c = i - y * fwid;
            S[y * SW + (c < lead ? c : c + w)] = 0.0f;
        }
        const int w4 = w >> 2, nv = h * w4;
        for (int j0 = threadIdx.x; j0 < nv; j0 += 4 * LG_THREADS) {
            float4 a[4], f[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + u * LG_THREADS;
                if (j < nv) {
                    const int y = j / w4, x = (j - y * w4) << 2;
                    a[u] = __ldcs(reinterpret_cast<const float4 *>(src + y * w + x));
                    if (fl) f[u] = __ldcs(reinterpret_cast<const float4 *>(fl + y * w + (w - 4 - x)));
                }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const int j = j0 + u * LG_THREADS;
                if (j < nv) {
                    const int y = j / w4, x = (j - y * w4) << 2;
                    float4 v = a[u];
                    if (fl) {  // flip_back + average (test.py:92-96)
                        v.x = __fmul_rn(__fadd_rn(v.x, f[u].w), 0.5f); v.y = __fmul_rn(__fadd_rn(v.y, f[u].z), 0.5f);
                        v.z = __fmul_rn(__fadd_rn(v.z, f[u].y), 0.5f); v.w = __fmul_rn(__fadd_rn(v.w, f[u].x), 0.5f);
                    }
                    float *d = S + y * SW + lead + x;
                    d[0] = v.x; d[1] = v.y; d[2] = v.z; d[3] = v.w;
                }
            }
        }
    } else {
    // warp per row, lanes along the row: no divisions, and the row tests below are warp-uniform
    for (int y = warp; y < h; y += NWARP) {
        float *srow = S + y * SW;
        for (int q = lane; q < SW; q += 32) {
            const int x = q - r - B;
            float v = 0.0f;
            if (x >= 0 && x < w) {
                v = src[y * w + x];
                if (fl) v = __fmul_rn(__fadd_rn(v, fl[y * w + (w - 1 - x)]), 0.5f);  // flip_back + average (test.py:92-96)
            }
            srow[q] = v;
        }
    }
    }
    __syncthreads();
    for (int y = warp; y < h; y += NWARP) {                      // row pass: fused chain, left to right
        const float *srow = S + y * SW;
        for (int xp = lane; xp < PW; xp += 32) {
            double acc = __dmul_rn(p.k[0], (double)srow[xp]);
#pragma unroll
            for (int t = 1; t < n; ++t) acc = fma(p.k[t], (double)srow[xp + t], acc);
            T[y * PW + xp] = acc;
        }
    }
    __syncthreads();
    Top2 best;
    best.init();
    for (int yp = warp; yp < PH; yp += NWARP) {                  // column pass over the whole padded domain
        const int y = yp - B;
        if (y + r < 0 || y - r >= h) {
            // every tap lies in the zero frame: k*0 + k*(0 + 0) ... = +0 exactly, nothing to load
            // (a thread walks its elements in index order, so only a strictly larger value can enter its pair)
            for (int xp = lane; xp < PW; xp += 32)
                if (0.0 > best.v2) best.push(0.0, yp * PW + xp);
            continue;
        }
        for (int xp = lane; xp < PW; xp += 32) {
            const double *c = T + y * PW + xp;   // rows y - r .. y + r exist: the map sits inside 2r zero rows
            double acc = __dmul_rn(p.k[r], c[0]);
#pragma unroll
            for (int t = 1; t <= r; ++t) acc = __dadd_rn(acc, __dmul_rn(p.k[r + t], __dadd_rn(c[t * PW], c[-t * PW])));
            if (acc > best.v2) best.push(acc, yp * PW + xp);
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov1 = __shfl_xor_sync(0xffffffffu, best.v1, o), ov2 = __shfl_xor_sync(0xffffffffu, best.v2, o);
        const int oi1 = __shfl_xor_sync(0xffffffffu, best.i1, o), oi2 = __shfl_xor_sync(0xffffffffu, best.i2, o);
        best.push(ov1, oi1);
        best.push(ov2, oi2);
    }
    if ((threadIdx.x & 31) == 0) s_red[threadIdx.x >> 5] = best;
    __syncthreads();
    if (threadIdx.x != 0) return;
    for (int q = 1; q < LG_THREADS / 32; ++q) {
        best.push(s_red[q].v1, s_red[q].i1);
        best.push(s_red[q].v2, s_red[q].i2);
    }
    // first arg-max, then arg-max of the map with that element set to 0 (test.py:44-50)
    const int i1 = best.i1;
    const int i2 = lg_before(0.0, i1, best.v2, best.i2) ? i1 : best.i2;
    const int yi = i1 / PW - B, xi = i1 % PW - B;
    const int px = (i2 % PW) - (i1 % PW), py = (i2 / PW) - (i1 / PW);
    double x = (double)xi, y = (double)yi;
    const double ln = __dsqrt_rn((double)(px * px + py * py));  // int ** 0.5: numpy's sqrt fast path
    if (ln > 1e-3) {
        x = __dadd_rn(x, __ddiv_rn(__dmul_rn(p.shift, (double)px), ln));
        y = __dadd_rn(y, __ddiv_rn(__dmul_rn(p.shift, (double)py), ln));
    }
    x = fmax(0.0, fmin(x, (double)(w - 1)));
    y = fmax(0.0, fmin(y, (double)(h - 1)));
    const int rx = (int)rint(x), ry = (int)rint(y);               // round(): half to even
    const float raw = S[ry * SW + rx + r + B];
    const float score = __fadd_rn(__fdiv_rn(raw, 255.0f), 0.5f);   // score_map = out / 255 + 0.5 (float32)
    const float s0 = __fmul_rn(p.scale[2 * person], 200.0f), s1 = __fmul_rn(p.scale[2 * person + 1], 200.0f);  // scales *= 200 (float32)
    const double kx = __dadd_rn(__dmul_rn(x, 4.0), 2.0), ky = __dadd_rn(__dmul_rn(y, 4.0), 2.0);
    p.preds[2 * m] = __dsub_rn(__dadd_rn(__dmul_rn(__ddiv_rn(kx, (double)p.in_w), (double)s0), (double)p.center[2 * person]),
                               (double)__fmul_rn(s0, 0.5f));
    p.preds[2 * m + 1] = __dsub_rn(__dadd_rn(__dmul_rn(__ddiv_rn(ky, (double)p.in_h), (double)s1), (double)p.center[2 * person + 1]),
                                   (double)__fmul_rn(s1, 0.5f));
    p.maxvals[m] = (double)score;
}

// ---------------------------------------------------------------------------------------
// Second generation (kernel sizes known at compile time: 5, 9, 11).  Same arithmetic, operation by operation, as the
// kernel above; what changes is how often a value is touched.  ncu on the first kernel (profiles/r01_rsn_legacy_v2_ncu.txt):
// 24 000 warp-instructions per 64x48 map of which a tenth are the float64 blur -- five shared loads and five
// float -> double conversions per row-pass output, five loads per column-pass output, 68-element rows walked in
// 32-lane steps, a branchy double-precision top-2 merged through shuffles.  Here
//   * row pass: a thread owns (row, 17-column segment) and slides along it with the converted window in registers:
//     one load and one conversion per input value; out-of-map taps are zeros in registers (no zero frame in memory);
//   * column pass: a thread owns (column, 24-row segment) of the rows that can see the map, window in registers:
//     one load per input value; rows that only see the zero frame are exactly +0 and are represented by the two
//     lowest indices of the padded map, (0.0, 0) and (0.0, 1) -- all the arg-max rule can ever pick from them;
//   * top-2: per thread as before (one well-predicted branch per element), across threads as two plain (value,
//     index) arg-max reductions: the global best, then the best of "my second if I own the best, else my first".
// ---------------------------------------------------------------------------------------
constexpr int LG_SEGX = 17, LG_SEGY = 24;

__device__ __forceinline__ void lg_block_argmax(double &v, int &i, double *red_v, int *red_i)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const double ov = __shfl_xor_sync(0xffffffffu, v, o);
        const int oi = __shfl_xor_sync(0xffffffffu, i, o);
        if (lg_before(ov, oi, v, i)) { v = ov; i = oi; }
    }
    if (lane == 0) { red_v[warp] = v; red_i[warp] = i; }
    __syncthreads();
    v = red_v[0]; i = red_i[0];
#pragma unroll
    for (int q = 1; q < LG_THREADS / 32; ++q)
        if (lg_before(red_v[q], red_i[q], v, i)) { v = red_v[q]; i = red_i[q]; }
    __syncthreads();  // the scratch is reused
}

// three CTAs per SM for the 5-tap kernel (80 registers; 4.39 ms against 5.25 at two); the 9- and 11-tap windows spill there
template <int NK>
__global__ void __launch_bounds__(LG_THREADS, NK <= 5 ? 3 : 2)
rsn_get_results_fast_kernel(const __grid_constant__ LegacyArgs p)
{
    extern __shared__ __align__(16) unsigned char lg_smem[];
    __shared__ double s_rv[LG_THREADS / 32];
    __shared__ int s_ri[LG_THREADS / 32];
    constexpr int n = NK, r = NK >> 1, B = LG_BORDER;
    const int w = p.w, h = p.h;
    const int PW = w + 2 * B;
    double *T0 = reinterpret_cast<double *>(lg_smem);            // [2r + h + 2r][PW] row-pass result inside 2r zero rows
    double *T = T0 + 2 * r * PW;                                 // row 0 of the map
    float *S = reinterpret_cast<float *>(T0 + (size_t)(h + 4 * r) * PW);  // [h][w] the (flip-averaged) map
    const long long m = blockIdx.x;
    const int person = (int)(m / p.K), joint = (int)(m - (long long)person * p.K);
    const float *src = p.out + m * (long long)(h * w);
    const float *fl = p.out_flipped ? p.out_flipped + ((long long)person * p.K + p.flip_perm[joint]) * (long long)(h * w) : nullptr;
    const int tid = threadIdx.x;
    for (int i = tid; i < 2 * r * PW; i += LG_THREADS) { T0[i] = 0.0; T[h * PW + i] = 0.0; }
    // ---- stage the map (flip_back + average, test.py:92-96)
    if ((w & 3) == 0 && (((uintptr_t)src | (uintptr_t)fl) & 15) == 0) {
        const int w4 = w >> 2, nv = h * w4;
        int y = tid / w4, x4 = tid - y * w4;                     // one division per thread, then incremental
        const int dy = LG_THREADS / w4, dx = LG_THREADS - dy * w4;
        for (int j0 = tid; j0 < nv; j0 += 4 * LG_THREADS) {
            float4 a[4], f[4];
            int yy[4], xx[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                yy[u] = y; xx[u] = x4 << 2;
                if (j0 + u * LG_THREADS < nv) {
                    a[u] = __ldcs(reinterpret_cast<const float4 *>(src + y * w + xx[u]));
                    if (fl) f[u] = __ldcs(reinterpret_cast<const float4 *>(fl + y * w + (w - 4 - xx[u])));
                }
                y += dy; x4 += dx;
                if (x4 >= w4) { x4 -= w4; ++y; }
            }
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                if (j0 + u * LG_THREADS < nv) {
                    float4 v = a[u];
                    if (fl) {
                        v.x = __fmul_rn(__fadd_rn(v.x, f[u].w), 0.5f); v.y = __fmul_rn(__fadd_rn(v.y, f[u].z), 0.5f);
                        v.z = __fmul_rn(__fadd_rn(v.z, f[u].y), 0.5f); v.w = __fmul_rn(__fadd_rn(v.w, f[u].x), 0.5f);
                    }
                    *reinterpret_cast<float4 *>(S + yy[u] * w + xx[u]) = v;
                }
            }
        }
    } else {
        for (int y = tid >> 5; y < h; y += LG_THREADS / 32)
            for (int x = tid & 31; x < w; x += 32) {
                float v = src[y * w + x];
                if (fl) v = __fmul_rn(__fadd_rn(v, fl[y * w + (w - 1 - x)]), 0.5f);
                S[y * w + x] = v;
            }
    }
    __syncthreads();
    // ---- row pass: fused chain, left to right; thread = (row y, columns xp0 .. xp0 + 16 of the padded row)
    {
        const int nsx = (PW + LG_SEGX - 1) / LG_SEGX, items = h * nsx;
        for (int it = tid; it < items; it += LG_THREADS) {
            const int y = it / nsx, xp0 = (it - y * nsx) * LG_SEGX;
            const float *srow = S + y * w;
            double d[LG_SEGX + n - 1];                            // d[j] = padded-row element xp0 + j - r  <->  map column xp0 + j - r - B
#pragma unroll
            for (int j = 0; j < LG_SEGX + n - 1; ++j) {
                const int x = xp0 + j - r - B;
                d[j] = (x >= 0 && x < w) ? (double)srow[x] : 0.0;
            }
            double *trow = T + y * PW + xp0;
#pragma unroll
            for (int o = 0; o < LG_SEGX; ++o) {
                double acc = __dmul_rn(p.k[0], d[o]);
#pragma unroll
                for (int t = 1; t < n; ++t) acc = fma(p.k[t], d[o + t], acc);
                if (xp0 + o < PW) trow[o] = acc;
            }
        }
    }
    __syncthreads();
    // ---- column pass over the rows that can see the map (yp = B - r .. B + h + r - 1); thread = (column, 24 rows)
    Top2 best;
    best.init();
    if (tid == 0) { best.push(0.0, 0); best.push(0.0, 1); }       // the all-zero rows above (B > r: indices 0 and 1 are theirs)
    {
        const int AR = h + 2 * r, nsy = (AR + LG_SEGY - 1) / LG_SEGY, items = PW * nsy;
        for (int it = tid; it < items; it += LG_THREADS) {
            const int sy = it / PW, xp = it - sy * PW;
            const int y0 = sy * LG_SEGY - r;                      // map row of the first output of this segment
            const double *col = T + xp;                          // T row y lives at col[y * PW]; rows -2r .. h + 2r - 1 exist
            double c[LG_SEGY + 2 * r];
#pragma unroll
            for (int j = 0; j < LG_SEGY + 2 * r; ++j) {
                const int y = y0 - r + j;
                c[j] = (y < h + 2 * r) ? col[y * PW] : 0.0;       // (y >= -2r always; past the zero rows below only in a partial segment)
            }
#pragma unroll
            for (int o = 0; o < LG_SEGY; ++o) {
                double acc = __dmul_rn(p.k[r], c[o + r]);
#pragma unroll
                for (int t = 1; t <= r; ++t) acc = __dadd_rn(acc, __dmul_rn(p.k[r + t], __dadd_rn(c[o + r + t], c[o + r - t])));
                const int y = y0 + o;
                if (y < h + r && acc > best.v2) best.push(acc, (y + B) * PW + xp);
            }
        }
    }
    // ---- top-2 across the block
    double v1 = best.v1;
    int i1 = best.i1;
    lg_block_argmax(v1, i1, s_rv, s_ri);
    double v2 = best.i1 == i1 ? best.v2 : best.v1;
    int i2 = best.i1 == i1 ? best.i2 : best.i1;
    lg_block_argmax(v2, i2, s_rv, s_ri);
    if (tid != 0) return;
    // first arg-max, then arg-max of the map with that element set to 0 (test.py:44-50)
    if (lg_before(0.0, i1, v2, i2)) i2 = i1;
    const int yi = i1 / PW - B, xi = i1 % PW - B;
    const int px = (i2 % PW) - (i1 % PW), py = (i2 / PW) - (i1 / PW);
    double x = (double)xi, y = (double)yi;
    const double ln = __dsqrt_rn((double)(px * px + py * py));  // int ** 0.5: numpy's sqrt fast path
    if (ln > 1e-3) {
        x = __dadd_rn(x, __ddiv_rn(__dmul_rn(p.shift, (double)px), ln));
        y = __dadd_rn(y, __ddiv_rn(__dmul_rn(p.shift, (double)py), ln));
    }
    x = fmax(0.0, fmin(x, (double)(w - 1)));
    y = fmax(0.0, fmin(y, (double)(h - 1)));
    const int rx = (int)rint(x), ry = (int)rint(y);               // round(): half to even
    const float raw = S[ry * w + rx];
    const float score = __fadd_rn(__fdiv_rn(raw, 255.0f), 0.5f);   // score_map = out / 255 + 0.5 (float32)
    const float s0 = __fmul_rn(p.scale[2 * person], 200.0f), s1 = __fmul_rn(p.scale[2 * person + 1], 200.0f);  // scales *= 200 (float32)
    const double kx = __dadd_rn(__dmul_rn(x, 4.0), 2.0), ky = __dadd_rn(__dmul_rn(y, 4.0), 2.0);
    p.preds[2 * m] = __dsub_rn(__dadd_rn(__dmul_rn(__ddiv_rn(kx, (double)p.in_w), (double)s0), (double)p.center[2 * person]),
                               (double)__fmul_rn(s0, 0.5f));
    p.preds[2 * m + 1] = __dsub_rn(__dadd_rn(__dmul_rn(__ddiv_rn(ky, (double)p.in_h), (double)s1), (double)p.center[2 * person + 1]),
                                   (double)__fmul_rn(s1, 0.5f));
    p.maxvals[m] = (double)score;
}

}  // namespace udp

extern "C" int udp_rsn_get_results(const float *outputs, const float *outputs_flipped, const int32_t *flip_perm,
                                   const float *centers, const float *scales, int K, int w, int h, int in_w, int in_h,
                                   int kernel, const double *taps_host, double shift, double *preds, double *maxvals,
                                   int N, void *stream)
{
    if (N < 0 || K < 1 || w < 1 || h < 1 || in_w < 1 || in_h < 1) return UDP_EINVAL;
    if (kernel < 3 || kernel > udp::LG_MAXK || !(kernel & 1) || !taps_host) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!outputs || !centers || !scales || !preds || !maxvals || (outputs_flipped && !flip_perm)) return UDP_EINVAL;
    const long long maps = (long long)N * K;
    if (maps > 0x7fffffffLL) return UDP_EINVAL;
    udp::LegacyArgs a;
    a.out = outputs; a.out_flipped = outputs_flipped; a.flip_perm = flip_perm;
    a.center = centers; a.scale = scales; a.preds = preds; a.maxvals = maxvals;
    for (int i = 0; i < udp::LG_MAXK; ++i) a.k[i] = i < kernel ? taps_host[i] : 0.0;
    a.shift = shift;
    a.K = K; a.w = w; a.h = h; a.in_w = in_w; a.in_h = in_h; a.n = kernel;
    const int PW = w + 2 * udp::LG_BORDER, SW = PW + 2 * (kernel >> 1);
    const char *le = getenv("UDP_LEGACY_FIRST_GEN");  // UDP_LEGACY_FIRST_GEN=1: the first-generation kernel for every size (tests)
    const bool first_gen = (le && le[0] == '1') || !(kernel == 5 || kernel == 9 || kernel == 11);
    // the second-generation kernel keeps the map without its zero frame
    const size_t smem = (size_t)(h + 4 * (kernel >> 1)) * PW * sizeof(double) + (size_t)h * (first_gen ? SW : w) * sizeof(float);
    if (smem > 200 * 1024) return UDP_EINVAL;
    void (*kern)(udp::LegacyArgs) = first_gen ? (kernel == 5 ? udp::rsn_get_results_kernel<5> : kernel == 9 ? udp::rsn_get_results_kernel<9>
                                                : kernel == 11 ? udp::rsn_get_results_kernel<11> : udp::rsn_get_results_kernel<0>)
                                  : kernel == 5 ? udp::rsn_get_results_fast_kernel<5>
                                  : kernel == 9 ? udp::rsn_get_results_fast_kernel<9>
                                  : kernel == 11 ? udp::rsn_get_results_fast_kernel<11> : udp::rsn_get_results_kernel<0>;
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) return (int)e;
    kern<<<(unsigned)maps, udp::LG_THREADS, smem, (cudaStream_t)stream>>>(a);
    UDP_LAUNCH_CHECK();
    return 0;
}
