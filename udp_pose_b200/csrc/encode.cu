// encode.cu -- target encoders: UDP Gaussian heatmaps, UDPv1 classification/offset maps,
// RSN multi-scale full-map Gaussians.  Write-only, HBM-bound kernels.
//
// Reference (paths relative to the reference root):
//   generate_target 'gaussian'  deep_hrnet/lib/dataset/JointsDataset.py:291-348
//   generate_target 'offset'    deep_hrnet/lib/dataset/JointsDataset.py:349-381
//   joints_weight               deep_hrnet/lib/dataset/JointsDataset.py:382-383
//   RSN generate_heatmap        RSN/dataset/JointsDataset.py:276-291, loop :214-220, sigmas :19-25
//
// Shape of every kernel: a CTA owns MAPS_PER_CTA consecutive (person, joint) maps.
//   phase 1 (one warp per map): lane 0 does the float64 per-joint prologue exactly as numpy
//     does (division by feat_stride, int() truncation, window bounds), broadcasts it with
//     warp shuffles, and the 32 lanes fill the separable float64 tables in shared memory;
//   phase 2 (all threads): the CTA's output range is one contiguous span of global memory;
//     every thread produces float4s straight from registers and issues 16-byte streaming
//     stores -- zeros and window values in one pass, no memset + scatter.
// The 2-D Gaussian is evaluated as exp(-dx^2/den) * exp(-dy^2/den) in float64 and rounded
// to float32 once; against numpy's exp(-(dx^2+dy^2)/den) this differs by < 2^-50 relative,
// i.e. the float32 results are identical except at rounding ties.
#include "common.cuh"

namespace udp {

constexpr int ENC_THREADS = 256;
constexpr int ENC_MAPS = ENC_THREADS / 32;  // maps per CTA = warps per CTA
constexpr int ENC_MAX_WIN = 64;             // max Gaussian window edge (6*sigma+1)

struct GaussMap {
    double ex[ENC_MAX_WIN];
    double ey[ENC_MAX_WIN];
    int ix0, ix1, iy0, iy1;  // image range [ix0,ix1) x [iy0,iy1) that receives the window
    int gx0, gy0;            // window index of image column ix0 / row iy0
    int draw;
    int pad;
};

// ---------------------------------------------------------------------------------------
// a5: UDP Gaussian target
// ---------------------------------------------------------------------------------------
__global__ void __launch_bounds__(ENC_THREADS)
encode_gaussian_kernel(const double *__restrict__ joints, const double *__restrict__ vis,
                       const float *__restrict__ jw, int K, int W, int H, int w, int h,
                       double sigma, float *__restrict__ target, float *__restrict__ tweight,
                       long long total_maps)
{
    __shared__ GaussMap sm[ENC_MAPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long map0 = (long long)blockIdx.x * ENC_MAPS;
    const int nmaps = (int)min((long long)ENC_MAPS, total_maps - map0);

    if (warp < nmaps) {
        const long long mi = map0 + warp;
        GaussMap &g = sm[warp];
        const double tmp = __dmul_rn(sigma, 3.0);
        const double size = __dadd_rn(__dmul_rn(2.0, tmp), 1.0);
        const int nwin = (int)ceil(size);
        const double half = floor(__ddiv_rn(size, 2.0));
        const double den = __dmul_rn(2.0, __dmul_rn(sigma, sigma));
        int ul0 = 0, ul1 = 0, draw = 0;
        double x0 = 0, y0 = 0;
        if (lane == 0) {
            const double fsx = __ddiv_rn(W - 1.0, w - 1.0), fsy = __ddiv_rn(H - 1.0, h - 1.0);
            const double jx = joints[mi * 3 + 0], jy = joints[mi * 3 + 1];
            float tw = (float)vis[mi * 3 + 0];
            const double ax = __ddiv_rn(jx, fsx), ay = __ddiv_rn(jy, fsy);
            const int mu_x = (int)__dadd_rn(ax, 0.5), mu_y = (int)__dadd_rn(ay, 0.5);  // trunc toward 0
            ul0 = (int)__dsub_rn((double)mu_x, tmp);
            ul1 = (int)__dsub_rn((double)mu_y, tmp);
            const int br0 = (int)__dadd_rn(__dadd_rn((double)mu_x, tmp), 1.0);
            const int br1 = (int)__dadd_rn(__dadd_rn((double)mu_y, tmp), 1.0);
            const bool outside = ul0 >= w || ul1 >= h || br0 < 0 || br1 < 0;
            if (outside) tw = 0.0f;
            draw = (!outside && tw > 0.5f) ? 1 : 0;
            x0 = __dadd_rn(half, __dsub_rn(ax, (double)mu_x));
            y0 = __dadd_rn(half, __dsub_rn(ay, (double)mu_y));
            g.gx0 = ul0 < 0 ? -ul0 : 0;
            g.gy0 = ul1 < 0 ? -ul1 : 0;
            g.ix0 = max(ul0, 0);
            g.iy0 = max(ul1, 0);
            // numpy slicing clamps the Gaussian range to the window extent
            g.ix1 = g.ix0 + max(0, min(min(br0, w) - ul0, nwin) - g.gx0);
            g.iy1 = g.iy0 + max(0, min(min(br1, h) - ul1, nwin) - g.gy0);
            g.draw = draw;
            tweight[mi] = jw ? __fmul_rn(tw, jw[mi % K]) : tw;
        }
        draw = __shfl_sync(0xffffffffu, draw, 0);
        if (draw) {
            x0 = __shfl_sync(0xffffffffu, x0, 0);
            y0 = __shfl_sync(0xffffffffu, y0, 0);
            for (int i = lane; i < nwin; i += 32) {
                double d = __dsub_rn((double)(float)i, x0);
                g.ex[i] = exp(__ddiv_rn(-__dmul_rn(d, d), den));
                d = __dsub_rn((double)(float)i, y0);
                g.ey[i] = exp(__ddiv_rn(-__dmul_rn(d, d), den));
            }
        }
    }
    __syncthreads();

    const int w4 = w >> 2;
    const int per_map = w4 * h;  // float4s per map
    float *base = target + map0 * (long long)(h * w);
    for (int m = 0; m < nmaps; ++m) {
        const GaussMap &g = sm[m];
        float *mp = base + (size_t)m * h * w;
        for (int j = threadIdx.x; j < per_map; j += ENC_THREADS) {
            const int y = j / w4, x = (j - y * w4) << 2;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (g.draw && y >= g.iy0 && y < g.iy1 && x + 3 >= g.ix0 && x < g.ix1) {
                const double ey = g.ey[g.gy0 + y - g.iy0];
                const int gx = g.gx0 - g.ix0 + x;
                if (x >= g.ix0 && x < g.ix1) v.x = (float)__dmul_rn(g.ex[gx], ey);
                if (x + 1 >= g.ix0 && x + 1 < g.ix1) v.y = (float)__dmul_rn(g.ex[gx + 1], ey);
                if (x + 2 >= g.ix0 && x + 2 < g.ix1) v.z = (float)__dmul_rn(g.ex[gx + 2], ey);
                if (x + 3 >= g.ix0 && x + 3 < g.ix1) v.w = (float)__dmul_rn(g.ex[gx + 3], ey);
            }
            st_stream_f4(mp + (size_t)j * 4, v);
        }
    }
}

// ---------------------------------------------------------------------------------------
// a6: UDPv1 classification + offset target.  Per joint three channel maps; the offsets are
// separable ((mu_x - col)/kpd depends on the column only), so the float64 divisions happen
// once per row / column and the per-pixel test is one float64 add + compare -- exactly the
// reference's dis = xo**2 + yo**2 <= 1.
// ---------------------------------------------------------------------------------------
constexpr int OFF_MAX_DIM = 256;

struct OffsetMap {
    double xo2[OFF_MAX_DIM];
    double yo2[OFF_MAX_DIM];
    float xo[OFF_MAX_DIM];
    float yo[OFF_MAX_DIM];
    int draw;
    int xlo, xhi, ylo, yhi;  // conservative bounding box of the kept disc (half-open)
};

constexpr int OFF_THREADS = 256;
constexpr int OFF_MAPS = 4;  // joints per CTA (12 channel maps)

__global__ void __launch_bounds__(OFF_THREADS)
encode_offset_kernel(const double *__restrict__ joints, const double *__restrict__ vis,
                     const float *__restrict__ jw, int K, int W, int H, int w, int h, double kpd,
                     float *__restrict__ target, float *__restrict__ tweight, long long total_maps)
{
    __shared__ OffsetMap sm[OFF_MAPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long map0 = (long long)blockIdx.x * OFF_MAPS;
    const int nmaps = (int)min((long long)OFF_MAPS, total_maps - map0);

    for (int m = warp; m < nmaps; m += OFF_THREADS / 32) {
        const long long mi = map0 + m;
        OffsetMap &g = sm[m];
        const float tw = (float)vis[mi * 3 + 0];
        const int draw = tw > 0.5f;
        if (lane == 0) {
            g.draw = draw;
            tweight[mi] = jw ? __fmul_rn(tw, jw[mi % K]) : tw;
        }
        if (draw) {
            const double fsx = __ddiv_rn(W - 1.0, w - 1.0), fsy = __ddiv_rn(H - 1.0, h - 1.0);
            const double mu_x = __ddiv_rn(joints[mi * 3 + 0], fsx);
            const double mu_y = __ddiv_rn(joints[mi * 3 + 1], fsy);
            for (int i = lane; i < w; i += 32) {
                const double o = __ddiv_rn(__dsub_rn(mu_x, (double)i), kpd);
                g.xo2[i] = __dmul_rn(o, o);
                g.xo[i] = (float)o;
            }
            for (int i = lane; i < h; i += 32) {
                const double o = __ddiv_rn(__dsub_rn(mu_y, (double)i), kpd);
                g.yo2[i] = __dmul_rn(o, o);
                g.yo[i] = (float)o;
            }
            if (lane == 0) {  // +-1 cell of slack around |mu - i| <= kpd; exact test is per pixel
                // (clamped in double first so that huge / NaN joints cannot overflow the casts;
                //  NaN gives the full box and then fails dis <= 1 at every pixel, as numpy does)
                g.xlo = (int)floor(fmax(mu_x - kpd - 1.0, 0.0));
                g.xhi = (int)ceil(fmin(fmax(mu_x + kpd + 2.0, 0.0), (double)w));
                g.ylo = (int)floor(fmax(mu_y - kpd - 1.0, 0.0));
                g.yhi = (int)ceil(fmin(fmax(mu_y + kpd + 2.0, 0.0), (double)h));
            }
        }
    }
    __syncthreads();

    const int w4 = w >> 2;
    const int per_map = w4 * h;
    const size_t hw = (size_t)h * w;
    float *base = target + map0 * (long long)(3 * hw);
    for (int m = 0; m < nmaps; ++m) {
        const OffsetMap &g = sm[m];
        float *c0 = base + (size_t)m * 3 * hw, *c1 = c0 + hw, *c2 = c1 + hw;
        for (int j = threadIdx.x; j < per_map; j += OFF_THREADS) {
            const int y = j / w4, x = (j - y * w4) << 2;
            float4 a = make_float4(0.f, 0.f, 0.f, 0.f), b = a, c = a;
            if (g.draw && y >= g.ylo && y < g.yhi && x + 3 >= g.xlo && x < g.xhi) {
                const double y2 = g.yo2[y];
                const float yo = g.yo[y];
                double d;
                d = __dadd_rn(g.xo2[x], y2);     if (d <= 1.0) { a.x = 1.f; b.x = g.xo[x]; c.x = yo; }
                d = __dadd_rn(g.xo2[x + 1], y2); if (d <= 1.0) { a.y = 1.f; b.y = g.xo[x + 1]; c.y = yo; }
                d = __dadd_rn(g.xo2[x + 2], y2); if (d <= 1.0) { a.z = 1.f; b.z = g.xo[x + 2]; c.z = yo; }
                d = __dadd_rn(g.xo2[x + 3], y2); if (d <= 1.0) { a.w = 1.f; b.w = g.xo[x + 3]; c.w = yo; }
            }
            st_stream_f4(c0 + (size_t)j * 4, a);
            st_stream_f4(c1 + (size_t)j * 4, b);
            st_stream_f4(c2 + (size_t)j * 4, c);
        }
    }
}

// ---------------------------------------------------------------------------------------
// a7: RSN full-map Gaussians, 255*exp(-dis/(2 sigma^2)) for S sigmas.  labels [N,S,K,h,w].
// One warp prepares one (person, sigma, joint) map: separable float64 exp tables.
// ---------------------------------------------------------------------------------------
struct RsnSigmas { double den[UDP_MAX_SIGMAS]; };  // -2*sigma^2

// The separable factors are kept as float-float pairs (hi + lo = the float64 value to 2^-48), so the
// per-element product runs on the packed FP32 pipe: p = xh*yh, its exact error by FMA, the two cross
// terms, one final add -- 5 fma.rn.f32x2-class instructions per 2 elements.  The first version
// multiplied in float64 and converted each element with F2F.F32.F64, which issues on the quarter-rate
// XU pipe: ncu showed XU at 39 % and the kernel at 69 % of the store roofline.  The float-float
// product differs from RN32(RN64(x*y)) only at near-ties (relative error 2^-46 before rounding).
struct RsnMap {
    float exh[OFF_MAX_DIM], exl[OFF_MAX_DIM];  // exp(dx^2/den)
    float eyh[OFF_MAX_DIM], eyl[OFF_MAX_DIM];  // 255 * exp(dy^2/den)
    int draw;
    int pad[3];
};

constexpr int RSN_THREADS = 256;
constexpr int RSN_MAPS = 8;

__global__ void __launch_bounds__(RSN_THREADS)
encode_rsn_kernel(const double *__restrict__ joints, int jstride, const float *__restrict__ valid,
                  int K, int W, int H, int w, int h, RsnSigmas sig, int S,
                  float *__restrict__ labels, long long total_maps)
{
    __shared__ __align__(16) RsnMap sm[RSN_MAPS];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const long long map0 = (long long)blockIdx.x * RSN_MAPS;
    const int nmaps = (int)min((long long)RSN_MAPS, total_maps - map0);

    for (int m = warp; m < nmaps; m += RSN_THREADS / 32) {
        const long long mi = map0 + m;  // index into [N,S,K]
        const int k = (int)(mi % K);
        const int s = (int)((mi / K) % S);
        const long long n = mi / ((long long)K * S);
        RsnMap &g = sm[m];
        const int draw = !(valid[n * K + k] < 1.0f);
        if (lane == 0) g.draw = draw;
        if (draw) {
            const double *jp = joints + (n * K + k) * (long long)jstride;
            const double ty = __ddiv_rn(__dmul_rn(jp[1], h - 1.0), H - 1.0);
            const double tx = __ddiv_rn(__dmul_rn(jp[0], w - 1.0), W - 1.0);
            const double den = sig.den[s];
            for (int i = lane; i < w; i += 32) {
                const double o = __dsub_rn(tx, (double)i);
                const double e = exp(__ddiv_rn(__dmul_rn(o, o), den));
                const float hi = (float)e;
                g.exh[i] = hi;
                g.exl[i] = (float)__dsub_rn(e, (double)hi);
            }
            for (int i = lane; i < h; i += 32) {
                const double o = __dsub_rn(ty, (double)i);
                const double e = __dmul_rn(255.0, exp(__ddiv_rn(__dmul_rn(o, o), den)));
                const float hi = (float)e;
                g.eyh[i] = hi;
                g.eyl[i] = (float)__dsub_rn(e, (double)hi);
            }
        }
    }
    __syncthreads();

    const int w4 = w >> 2;
    const int per_map = w4 * h;
    const size_t hw = (size_t)h * w;
    float *base = labels + map0 * (long long)hw;
    for (int m = 0; m < nmaps; ++m) {
        const RsnMap &g = sm[m];
        float *mp = base + (size_t)m * hw;
        if (!g.draw) {
            for (int j = threadIdx.x; j < per_map; j += RSN_THREADS)
                st_stream_f4(mp + (size_t)j * 4, make_float4(0.f, 0.f, 0.f, 0.f));
            continue;
        }
        for (int j = threadIdx.x; j < per_map; j += RSN_THREADS) {
            const int y = j / w4, x = (j - y * w4) << 2;
            const float yh = g.eyh[y], yl = g.eyl[y];
            const f32x2 Y = pk(yh, yh), Yn = pk(-yh, -yh), YLn = pk(-yl, -yl);
            const ulonglong2 xh = *reinterpret_cast<const ulonglong2 *>(g.exh + x);
            const ulonglong2 xl = *reinterpret_cast<const ulonglong2 *>(g.exl + x);
            ulonglong2 r;
            {
                const f32x2 p = fmul2pp(xh.x, Y);
                f32x2 c = ffma2pp(xh.x, Yn, p);      // p - xh*yh, exact
                c = ffma2pp(xh.x, YLn, c);
                c = ffma2pp(xl.x, Yn, c);
                r.x = fsub2p(p, c);
            }
            {
                const f32x2 p = fmul2pp(xh.y, Y);
                f32x2 c = ffma2pp(xh.y, Yn, p);
                c = ffma2pp(xh.y, YLn, c);
                c = ffma2pp(xl.y, Yn, c);
                r.y = fsub2p(p, c);
            }
            __stcs(reinterpret_cast<float4 *>(mp + (size_t)j * 4),
                   make_float4(lo_of(r.x), hi_of(r.x), lo_of(r.y), hi_of(r.y)));
        }
    }
}

}  // namespace udp

static int check_enc(const void *joints, const void *vis, const void *target, const void *tw,
                     int K, int W, int H, int w, int h, int N)
{
    if (N < 0 || K < 1 || W < 2 || H < 2 || w < 4 || h < 2 || (w & 3)) return UDP_EINVAL;
    if (N > 0 && (!joints || !vis || !target || !tw)) return UDP_EINVAL;
    return 0;
}

extern "C" int udp_encode_gaussian(const double *joints, const double *joints_vis,
                                   const float *joints_weight, int K, int W, int H, int w, int h,
                                   double sigma, float *target, float *target_weight, int N,
                                   void *stream)
{
    int e = check_enc(joints, joints_vis, target, target_weight, K, W, H, w, h, N);
    if (e) return e;
    if (!(sigma > 0) || 6 * sigma + 1 > udp::ENC_MAX_WIN) return UDP_EINVAL;
    if (N == 0) return 0;
    long long maps = (long long)N * K;
    long long blocks = (maps + udp::ENC_MAPS - 1) / udp::ENC_MAPS;
    udp::encode_gaussian_kernel<<<(unsigned)blocks, udp::ENC_THREADS, 0, (cudaStream_t)stream>>>(
        joints, joints_vis, joints_weight, K, W, H, w, h, sigma, target, target_weight, maps);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_encode_offset(const double *joints, const double *joints_vis,
                                 const float *joints_weight, int K, int W, int H, int w, int h,
                                 double kpd, float *target, float *target_weight, int N,
                                 void *stream)
{
    int e = check_enc(joints, joints_vis, target, target_weight, K, W, H, w, h, N);
    if (e) return e;
    if (!(kpd > 0) || w > udp::OFF_MAX_DIM || h > udp::OFF_MAX_DIM) return UDP_EINVAL;
    if (N == 0) return 0;
    long long maps = (long long)N * K;
    long long blocks = (maps + udp::OFF_MAPS - 1) / udp::OFF_MAPS;
    udp::encode_offset_kernel<<<(unsigned)blocks, udp::OFF_THREADS, 0, (cudaStream_t)stream>>>(
        joints, joints_vis, joints_weight, K, W, H, w, h, kpd, target, target_weight, maps);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_encode_rsn(const double *joints, int jstride, const float *valid, int K, int W,
                              int H, int w, int h, const double *sigmas, int S, float *labels,
                              int N, void *stream)
{
    if (N < 0 || K < 1 || W < 2 || H < 2 || w < 4 || h < 2 || (w & 3) || jstride < 2) return UDP_EINVAL;
    if (S < 1 || S > UDP_MAX_SIGMAS || !sigmas || w > udp::OFF_MAX_DIM || h > udp::OFF_MAX_DIM) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!joints || !valid || !labels) return UDP_EINVAL;
    udp::RsnSigmas sg;
    for (int s = 0; s < UDP_MAX_SIGMAS; ++s) {
        double v = s < S ? sigmas[s] : 1.0;
        sg.den[s] = -2.0 * (v * v);  // RSN/dataset/JointsDataset.py:289: dis / (-2.0 * sigma ** 2)
    }
    long long maps = (long long)N * S * K;
    long long blocks = (maps + udp::RSN_MAPS - 1) / udp::RSN_MAPS;
    udp::encode_rsn_kernel<<<(unsigned)blocks, udp::RSN_THREADS, 0, (cudaStream_t)stream>>>(
        joints, jstride, valid, K, W, H, w, h, sg, S, labels, maps);
    UDP_LAUNCH_CHECK();
    return 0;
}
