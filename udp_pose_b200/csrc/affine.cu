// affine.cu -- UDP unbiased affine: warp matrices, joint forward map, fixed-point bilinear crop.
//
// Reference (paths relative to the reference root):
//   get_warpmatrix      deep_hrnet/lib/dataset/JointsDataset.py:29-49 (call :226)
//   rotate_points       deep_hrnet/lib/dataset/JointsDataset.py:51-73 (call :228)
//   cv2.warpAffine      call sites JointsDataset.py:227, RSN/dataset/JointsDataset.py:197-198
//   RSN joint map       RSN/dataset/JointsDataset.py:27-47,:203-211; RSN/lib/utils/transforms.py:38-41
#include "blur.cuh"

namespace udp {

// ---------------------------------------------------------------------------------------
// a1: one thread per person.  Mirrors numpy-2 promotion: float32 center/scale make the two
// translation entries float32 arithmetic end to end; the linear entries are float64 / (W-1).
// ---------------------------------------------------------------------------------------
__global__ void warp_matrices_kernel(const void *center, const void *scale, int cs_f64,
                                     const double *rot, int W, int H, float *M, int N)
{
    int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    double theta = rot ? rot[n] : 0.0;
    double co = 1.0, sn = 0.0;
    if (theta != 0.0) {
        double th = __dmul_rn(__ddiv_rn(theta, 180.0), kPi);
        sincos(th, &sn, &co);
    }
    CS cs = load_cs(center, scale, cs_f64, n);
    double sx = __ddiv_rn(cs.st0, W - 1.0), sy = __ddiv_rn(cs.st1, H - 1.0);
    float *m = M + 6 * (size_t)n;
    m[0] = (float)__dmul_rn(co, sx);
    m[1] = (float)__dmul_rn(sn, sy);
    m[3] = (float)__dmul_rn(-sn, sx);
    m[4] = (float)__dmul_rn(co, sy);
    if (cs_f64) {
        double si0 = __dmul_rn(cs.c0, 2.0), si1 = __dmul_rn(cs.c1, 2.0);
        double a = __dmul_rn(__dmul_rn(-0.5, cs.st0), co);
        double b = __dmul_rn(__dmul_rn(0.5, cs.st1), sn);
        m[2] = (float)__dadd_rn(__dsub_rn(a, b), __dmul_rn(0.5, si0));
        a = __dmul_rn(__dmul_rn(0.5, cs.st0), sn);
        b = __dmul_rn(__dmul_rn(0.5, cs.st1), co);
        m[5] = (float)__dadd_rn(__dsub_rn(a, b), __dmul_rn(0.5, si1));
    } else {
        float cf = (float)co, sf = (float)sn;
        float si0 = __fmul_rn(cs.c0f, 2.0f), si1 = __fmul_rn(cs.c1f, 2.0f);
        float a = __fmul_rn(__fmul_rn(-0.5f, cs.st0f), cf);
        float b = __fmul_rn(__fmul_rn(0.5f, cs.st1f), sf);
        m[2] = __fadd_rn(__fsub_rn(a, b), __fmul_rn(0.5f, si0));
        a = __fmul_rn(__fmul_rn(0.5f, cs.st0f), sf);
        b = __fmul_rn(__fmul_rn(0.5f, cs.st1f), cf);
        m[5] = __fadd_rn(__fsub_rn(a, b), __fmul_rn(0.5f, si1));
    }
}

// ---------------------------------------------------------------------------------------
// a3: one thread per (person, joint), float64.
// ---------------------------------------------------------------------------------------
__global__ void rotate_points_kernel(const double *joints, int jstride, const double *rot,
                                     const void *center, const void *scale, int cs_f64, int W,
                                     int H, int do_clip, double *out, int ostride, int K, int N)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * K) return;
    int n = i / K;
    CS cs = load_cs(center, scale, cs_f64, n);
    double theta = rot ? rot[n] : 0.0;
    double co = 1.0, sn = 0.0;
    if (theta != 0.0) sincos(__dmul_rn(__ddiv_rn(theta, 180.0), kPi), &sn, &co);
    double rsin = -sn, rcos = co;
    double scale_x = __ddiv_rn(W - 1.0, cs.st0), scale_y = __ddiv_rn(H - 1.0, cs.st1);
    double px = joints[(size_t)i * jstride + 0], py = joints[(size_t)i * jstride + 1];
    double sx = __dsub_rn(px, cs.c0), sy = __dsub_rn(py, cs.c1);
    double dx = __dadd_rn(__dmul_rn(rcos, sx), __dmul_rn(rsin, sy));
    double dy = __dadd_rn(__dmul_rn(-rsin, sx), __dmul_rn(rcos, sy));
    dx = __dadd_rn(dx, __dmul_rn(cs.st0, 0.5));
    dy = __dadd_rn(dy, __dmul_rn(cs.st1, 0.5));
    dx = __dmul_rn(dx, scale_x);
    dy = __dmul_rn(dy, scale_y);
    if (do_clip) {  // the reference clips x by shape[1]-1 and y by shape[0]-1 (:70-72)
        dx = fmin(fmax(dx, 0.0), H - 1.0);
        dy = fmin(fmax(dy, 0.0), W - 1.0);
    }
    out[(size_t)i * ostride + 0] = dx;
    out[(size_t)i * ostride + 1] = dy;
}

// ---------------------------------------------------------------------------------------
// a3': RSN forward joint map with a float32 matrix, float64 dot, visibility cull.
// One thread per (person, joint); every thread of a person recomputes the 6-entry matrix
// (cheap, avoids a second launch); joint 0's thread stores it.
// ---------------------------------------------------------------------------------------
__global__ void rsn_transform_joints_kernel(double *joints, int jstride, double *vis, int vstride,
                                            const double *rot, const float *center,
                                            const float *scale, int W, int H, float *Minv, int K,
                                            int N)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * K) return;
    int n = i / K, k = i - n * K;
    double theta = rot ? rot[n] : 0.0;
    double co = 1.0, sn = 0.0;
    if (theta != 0.0) sincos(__dmul_rn(__ddiv_rn(theta, 180.0), kPi), &sn, &co);
    float st0 = __fmul_rn(scale[2 * n], 200.0f), st1 = __fmul_rn(scale[2 * n + 1], 200.0f);
    float si0 = __fmul_rn(center[2 * n], 2.0f), si1 = __fmul_rn(center[2 * n + 1], 2.0f);
    double scx = __ddiv_rn(W - 1.0, (double)st0), scy = __ddiv_rn(H - 1.0, (double)st1);
    float cf = (float)co, sf = (float)sn;
    float m[6];
    m[0] = (float)__dmul_rn(co, scx);
    m[1] = (float)__dmul_rn(-sn, scx);
    m[3] = (float)__dmul_rn(sn, scy);
    m[4] = (float)__dmul_rn(co, scy);
    {
        float a = __fmul_rn(__fmul_rn(-0.5f, si0), cf), b = __fmul_rn(__fmul_rn(0.5f, si1), sf);
        float d = __fadd_rn(__fadd_rn(a, b), __fmul_rn(0.5f, st0));
        m[2] = (float)__dmul_rn(scx, (double)d);
        a = __fmul_rn(__fmul_rn(-0.5f, si0), sf);
        b = __fmul_rn(__fmul_rn(0.5f, si1), cf);
        d = __fadd_rn(__fsub_rn(a, b), __fmul_rn(0.5f, st1));
        m[5] = (float)__dmul_rn(scy, (double)d);
    }
    if (Minv && k == 0)
        for (int j = 0; j < 6; ++j) Minv[6 * (size_t)n + j] = m[j];
    double *v = vis + (size_t)i * vstride;
    if (!(v[0] > 0)) return;
    double *p = joints + (size_t)i * jstride;
    double x = p[0], y = p[1];
    double nx = __dadd_rn(__dadd_rn(__dmul_rn((double)m[0], x), __dmul_rn((double)m[1], y)), (double)m[2]);
    double ny = __dadd_rn(__dadd_rn(__dmul_rn((double)m[3], x), __dmul_rn((double)m[4], y)), (double)m[5]);
    p[0] = nx;
    p[1] = ny;
    if (nx < 0 || nx > W - 1 || ny < 0 || ny > H - 1) v[0] = 0;
}

// ---------------------------------------------------------------------------------------
// a4: fliplr_joints  deep_hrnet/lib/utils/transforms.py:50-64 (RSN flip_joints :70-79 is the
// same without the final product -- pass mul_vis = 0).  One thread per (person, joint, coord).
// ---------------------------------------------------------------------------------------
__global__ void fliplr_joints_kernel(const double *joints, const double *vis, const int32_t *perm,
                                     const int32_t *width, int K, int N, int mul_vis,
                                     double *joints_out, double *vis_out)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * K * 3) return;
    int c = i % 3, k = (i / 3) % K, n = i / (3 * K);
    size_t src = ((size_t)n * K + perm[k]) * 3 + c;
    double v = joints[src];
    if (c == 0) v = __dsub_rn(__dsub_rn((double)width[n], v), 1.0);
    double vv = vis[src];
    joints_out[i] = mul_vis ? __dmul_rn(v, vv) : v;
    vis_out[i] = vv;
}

// a16 standalone: transform_preds(coords, center, scale, [w,h])  inference.py:20-27
__global__ void transform_preds_kernel(const void *coords, int coords_f64, const void *center,
                                       const void *scale, int cs_f64, int w, int h, double *out,
                                       int K, int N)
{
    int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= N * K) return;
    int n = i / K;
    double cx, cy;
    if (coords_f64) { cx = ((const double *)coords)[2 * (size_t)i]; cy = ((const double *)coords)[2 * (size_t)i + 1]; }
    else { cx = ((const float *)coords)[2 * (size_t)i]; cy = ((const float *)coords)[2 * (size_t)i + 1]; }
    CS cs = load_cs(center, scale, cs_f64, n);
    double ox, oy;
    transform_pred(cx, cy, cs, cs_f64, coords_f64 != 0, w, h, ox, oy);
    if (cs_f64 && !coords_f64) {  // float32 coords * np.float64 scalars: float64 result, not re-rounded
        const double sx = __ddiv_rn(cs.st0, w - 1.0), sy = __ddiv_rn(cs.st1, h - 1.0);
        ox = __dsub_rn(__dadd_rn(__dmul_rn(cx, sx), cs.c0), __dmul_rn(cs.st0, 0.5));
        oy = __dsub_rn(__dadd_rn(__dmul_rn(cy, sy), cs.c1), __dmul_rn(cs.st1, 0.5));
    }
    out[2 * (size_t)i] = ox;
    out[2 * (size_t)i + 1] = oy;
}

}  // namespace udp

extern "C" int udp_warp_matrices(const void *center, const void *scale, int cs_f64,
                                 const double *rot_deg, int W, int H, float *M_out, int N,
                                 void *stream)
{
    if (N < 0 || W < 2 || H < 2 || (N > 0 && (!center || !scale || !M_out))) return UDP_EINVAL;
    if (N == 0) return 0;
    udp::warp_matrices_kernel<<<(N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        center, scale, cs_f64, rot_deg, W, H, M_out, N);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_rotate_points(const double *joints, int jstride, const double *rot_deg,
                                 const void *center, const void *scale, int cs_f64, int W, int H,
                                 int do_clip, double *out, int ostride, int K, int N, void *stream)
{
    if (N < 0 || K < 0 || jstride < 2 || ostride < 2 || W < 2 || H < 2) return UDP_EINVAL;
    if (N * K == 0) return 0;
    if (!joints || !out || !center || !scale) return UDP_EINVAL;
    udp::rotate_points_kernel<<<(N * K + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        joints, jstride, rot_deg, center, scale, cs_f64, W, H, do_clip, out, ostride, K, N);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_rsn_transform_joints(double *joints, int jstride, double *vis, int vstride,
                                        const double *rot_deg, const float *center,
                                        const float *scale, int W, int H, float *Minv_out, int K,
                                        int N, void *stream)
{
    if (N < 0 || K < 0 || jstride < 2 || vstride < 1 || W < 2 || H < 2) return UDP_EINVAL;
    if (N * K == 0) return 0;
    if (!joints || !vis || !center || !scale) return UDP_EINVAL;
    udp::rsn_transform_joints_kernel<<<(N * K + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        joints, jstride, vis, vstride, rot_deg, center, scale, W, H, Minv_out, K, N);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_codec_version(void) { return 100; }  // 0.1.0

extern "C" int udp_codec_sm_arch(void)
{
#ifdef UDP_SM_ARCH
    return UDP_SM_ARCH;
#else
    return 0;
#endif
}

extern "C" int udp_fliplr_joints(const double *joints, const double *joints_vis, const int32_t *flip_perm,
                                 const int32_t *width, int K, int N, int mul_vis, double *joints_out,
                                 double *vis_out, void *stream)
{
    if (N < 0 || K < 1) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!joints || !joints_vis || !flip_perm || !width || !joints_out || !vis_out) return UDP_EINVAL;
    if (joints_out == joints || vis_out == joints_vis) return UDP_EINVAL;
    int total = N * K * 3;
    udp::fliplr_joints_kernel<<<(total + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        joints, joints_vis, flip_perm, width, K, N, mul_vis, joints_out, vis_out);
    UDP_LAUNCH_CHECK();
    return 0;
}

// all_boxes rows of validate() (lib/core/function.py:211-214): (cx, cy, sx, sy, prod(s * 200), score) in float64.
// s * 200 and the product run in the records' dtype (float32 for COCO: numpy keeps float32 through `s*200` and np.prod).
namespace udp {
__global__ void pack_all_boxes_kernel(const void *center, const void *scale, int cs_f64, const double *score, double *out, int N)
{
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n >= N) return;
    double c0, c1, s0, s1, area;
    if (cs_f64) {
        const double *c = reinterpret_cast<const double *>(center) + 2 * (size_t)n, *s = reinterpret_cast<const double *>(scale) + 2 * (size_t)n;
        c0 = c[0]; c1 = c[1]; s0 = s[0]; s1 = s[1];
        area = __dmul_rn(__dmul_rn(s0, 200.0), __dmul_rn(s1, 200.0));
    } else {
        const float *c = reinterpret_cast<const float *>(center) + 2 * (size_t)n, *s = reinterpret_cast<const float *>(scale) + 2 * (size_t)n;
        c0 = c[0]; c1 = c[1]; s0 = s[0]; s1 = s[1];
        area = (double)__fmul_rn(__fmul_rn(s[0], 200.0f), __fmul_rn(s[1], 200.0f));
    }
    double *o = out + 6 * (size_t)n;
    o[0] = c0; o[1] = c1; o[2] = s0; o[3] = s1; o[4] = area; o[5] = score ? score[n] : 1.0;
}
}  // namespace udp

extern "C" int udp_pack_all_boxes(const void *center, const void *scale, int cs_f64, const double *score, double *all_boxes,
                                  int N, void *stream)
{
    if (N < 0) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!center || !scale || !all_boxes) return UDP_EINVAL;
    udp::pack_all_boxes_kernel<<<(N + 127) / 128, 128, 0, (cudaStream_t)stream>>>(center, scale, cs_f64, score, all_boxes, N);
    UDP_LAUNCH_CHECK();
    return 0;
}

extern "C" int udp_transform_preds(const void *coords, int coords_f64, const void *center,
                                   const void *scale, int cs_f64, int w, int h, double *out, int K,
                                   int N, void *stream)
{
    if (N < 0 || K < 1 || w < 2 || h < 2) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!coords || !center || !scale || !out) return UDP_EINVAL;
    udp::transform_preds_kernel<<<(N * K + 127) / 128, 128, 0, (cudaStream_t)stream>>>(
        coords, coords_f64, center, scale, cs_f64, w, h, out, K, N);
    UDP_LAUNCH_CHECK();
    return 0;
}
