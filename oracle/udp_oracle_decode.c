/*
 * udp_oracle_decode.c -- CPU restatement of flip-back, blur, argmax, DARK-style Taylor
 * refinement, offset voting and the inverse affine (TEST INFRASTRUCTURE ONLY; see the
 * header of udp_oracle.c).  Parity: pinned by tests/golden/*.npz (outputs of the
 * unmodified reference) and, for the blur, bit-exactly against the installed
 * cv2.GaussianBlur in tests/test_oracle_cpu.py.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

/* cv2.getGaussianKernel(n, 0, CV_32F): n=7 is OpenCV's fixed small table [2,7,14,18,14,7,2]/64;
 * n=15 is exp(-x^2/(2*2.6^2)) normalised in double, stored float32 (bit patterns below were
 * read from cv2 4.13.0 and are re-checked in tests/test_oracle_cpu.py). */
static const float K7[7] = {0.03125f, 0.109375f, 0.21875f, 0.28125f, 0.21875f, 0.109375f, 0.03125f};
static const uint32_t K15_BITS[15] = {
    0x3b869623u, 0x3c300526u, 0x3cc68dcbu, 0x3d412caau, 0x3da218f7u, 0x3deaa1edu, 0x3e12764eu,
    0x3e1db4b1u, 0x3e12764eu, 0x3deaa1edu, 0x3da218f7u, 0x3d412caau, 0x3cc68dcbu, 0x3c300526u,
    0x3b869623u};

void udp_o_gauss_kernel(int n, float *k)
{
    if (n == 7) memcpy(k, K7, sizeof(K7));
    else memcpy(k, K15_BITS, sizeof(K15_BITS));
}

static inline int reflect101(int i, int n)
{
    if (n == 1) return 0;
    while (i < 0 || i >= n) {
        if (i < 0) i = -i;
        else i = 2 * (n - 1) - i;
    }
    return i;
}

/* cv2.GaussianBlur(f32 [h,w], (n,n), 0), BORDER_REFLECT_101 -- call sites
 * deep_hrnet/lib/core/inference.py:76,165-167.  Row pass: sequential FMA left->right;
 * column pass: centre tap then symmetric pairs, FMA.  dst may alias src. */
void udp_o_gaussian_blur(const float *src, float *dst, int h, int w, int n)
{
    float k[15];
    int r = n / 2;
    udp_o_gauss_kernel(n, k);
    float *T = (float *)malloc(sizeof(float) * (size_t)h * w);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            float s = src[y * w + reflect101(x - r, w)] * k[0];
            for (int i = 1; i < n; ++i) s = fmaf(src[y * w + reflect101(x - r + i, w)], k[i], s);
            T[y * w + x] = s;
        }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            float s = fmaf(T[y * w + x], k[r], 0.0f);
            for (int i = 1; i <= r; ++i) {
                float p = T[reflect101(y + i, h) * w + x] + T[reflect101(y - i, h) * w + x];
                s = fmaf(p, k[r + i], s);
            }
            dst[y * w + x] = s;
        }
    free(T);
}

/* a9/a10/a11  flip_back / flip_back_offset + average
 * deep_hrnet/lib/utils/transforms.py:15-47, lib/core/function.py:171.
 * out, flipped: [C,h,w]; perm[K] = partner joint (identity where unpaired).
 * triplets = 0: C = K (gaussian); triplets = 1: C = 3K, channel (j,1) negated (:40). */
void udp_o_flip_average(const float *out, const float *flipped, const int *perm, int K, int h,
                        int w, int triplets, float *avg)
{
    int per = triplets ? 3 : 1;
    for (int j = 0; j < K; ++j)
        for (int t = 0; t < per; ++t) {
            const float *a = out + ((size_t)j * per + t) * h * w;
            const float *f = flipped + ((size_t)perm[j] * per + t) * h * w;
            float *o = avg + ((size_t)j * per + t) * h * w;
            for (int y = 0; y < h; ++y)
                for (int x = 0; x < w; ++x) {
                    float fv = f[y * w + (w - 1 - x)];
                    if (triplets && t == 1) fv = -fv;
                    o[y * w + x] = (a[y * w + x] + fv) * 0.5f;
                }
        }
}

/* a12 get_max_preds  inference.py:30-58 : first (lowest) flat index of the maximum. */
static void argmax_map(const float *m, int n, int *idx, float *maxv)
{
    int bi = 0;
    float bv = m[0];
    for (int i = 1; i < n; ++i) /* np.argmax/np.amax: the first NaN wins and propagates */
        if (m[i] > bv || (m[i] != m[i] && bv == bv)) { bv = m[i]; bi = i; }
    *idx = bi;
    *maxv = bv;
}

/* a16 transform_preds  inference.py:20-27.  f64coords: coords are float64 (post-processed)
 * else float32 (arithmetic stays float32 when center/scale are float32 too). */
static void transform_pred(double cx, double cy, const void *center, const void *scale, int cs_f64,
                           int f64coords, int w, int h, double *ox, double *oy)
{
    if (cs_f64) {
        const double *c = (const double *)center, *s = (const double *)scale;
        double s0 = s[0] * 200.0, s1 = s[1] * 200.0;
        double sx = s0 / (w - 1.0), sy = s1 / (h - 1.0);
        *ox = cx * sx + c[0] - s0 * 0.5;
        *oy = cy * sy + c[1] - s1 * 0.5;
        if (!f64coords) { /* float32 coords array: result is stored back into float32 preds */
            *ox = (double)(float)*ox;
            *oy = (double)(float)*oy;
        }
    } else {
        const float *c = (const float *)center, *s = (const float *)scale;
        float s0 = s[0] * 200.0f, s1 = s[1] * 200.0f;
        float sx = s0 / (float)(w - 1.0), sy = s1 / (float)(h - 1.0);
        if (f64coords) {
            *ox = cx * (double)sx + (double)c[0] - (double)(s0 * 0.5f);
            *oy = cy * (double)sy + (double)c[1] - (double)(s1 * 0.5f);
        } else {
            float fx = (float)cx * sx, fy = (float)cy * sy;
            fx = fx + c[0];
            fy = fy + c[1];
            *ox = (double)(fx - s0 * 0.5f);
            *oy = (double)(fy - s1 * 0.5f);
        }
    }
}

/* np.clip(x, 0.001, 50) in float32 (:81); NaN passes through. */
static inline float clip_nan(float v)
{
    if (v != v) return v;
    return v < 0.001f ? 0.001f : (v > 50.0f ? 50.0f : v);
}

/* 2x2 inverse the way np.linalg.inv does it (LAPACK gesv: LU with partial pivoting);
 * returns 0 and a zero matrix when a pivot is exactly zero (LinAlgError -> zeros,
 * inference.py:126-132). */
static int inv2x2(double a, double b, double c, double d, double *inv)
{
    /* matrix [[a,b],[c,d]] */
    double r0[2] = {a, b}, r1[2] = {c, d};
    double e0[2] = {1, 0}, e1[2] = {0, 1};
    if (fabs(c) > fabs(a)) {
        double t;
        t = r0[0]; r0[0] = r1[0]; r1[0] = t;
        t = r0[1]; r0[1] = r1[1]; r1[1] = t;
        e0[0] = 0; e0[1] = 1; e1[0] = 1; e1[1] = 0;
    }
    if (r0[0] == 0.0) { inv[0] = inv[1] = inv[2] = inv[3] = 0; return 0; }
    double l = r1[0] / r0[0];
    double u11 = r1[1] - l * r0[1];
    if (u11 == 0.0) { inv[0] = inv[1] = inv[2] = inv[3] = 0; return 0; }
    /* forward: y0 = e0, y1 = e1 - l*e0 ; back: x1 = y1/u11, x0 = (y0 - r0[1]*x1)/r0[0] */
    for (int col = 0; col < 2; ++col) {
        double y0 = e0[col], y1 = e1[col] - l * e0[col];
        double x1 = y1 / u11;
        double x0 = (y0 - r0[1] * x1) / r0[0];
        inv[0 * 2 + col] = x0;
        inv[1 * 2 + col] = x1;
    }
    return 1;
}

/* a13+a14: get_final_preds 'gaussian' for ONE person.  inference.py:149-155,176-186 and
 * post() :60-145.  hm: [K,h,w] float32 (already flip-averaged).  pre_div: 0 = none, else the
 * map is first divided by it in float32 (RSN prm test.py:185, outputs/255.0).
 * Outputs: preds [K,2] (image space), maxvals [K] (RAW maxima), preds_in [K,2]
 * (preds_in_input_space), idx [K] (raw argmax flat index), blurred (optional [K,h,w]:
 * the renormalised blurred maps post() writes back, :80). */
void udp_o_decode_gaussian(const float *hm_in, int K, int w, int h, const void *center,
                           const void *scale, int cs_f64, float pre_div, int post_process,
                           double *preds, float *maxvals, double *preds_in, int32_t *idx,
                           float *blurred)
{
    int n = h * w;
    float *m = (float *)malloc(sizeof(float) * n);
    float *b = (float *)malloc(sizeof(float) * n);
    for (int k = 0; k < K; ++k) {
        const float *src = hm_in + (size_t)k * n;
        for (int i = 0; i < n; ++i) m[i] = pre_div != 0.0f ? src[i] / pre_div : src[i];
        int bi;
        float mv;
        argmax_map(m, n, &bi, &mv);
        idx[k] = bi;
        maxvals[k] = mv;
        float px = (float)(bi % w), py = floorf((float)bi / (float)w);
        float mask = mv > 0.0f ? 1.0f : 0.0f; /* :54-57 */
        px *= mask;
        py *= mask;
        double cx = px, cy = py;
        if (post_process) {
            float maxori = mv; /* :74 np.max of the raw map */
            udp_o_gaussian_blur(m, b, h, w, 7);
            float mx = b[0], mn = b[0];
            for (int i = 1; i < n; ++i) { /* np.max / np.min propagate NaN */
                if (b[i] > mx || b[i] != b[i]) mx = b[i];
                if (b[i] < mn || b[i] != b[i]) mn = b[i];
            }
            float range = mx - mn;
            for (int i = 0; i < n; ++i) b[i] = (b[i] - mn) / range * maxori; /* :79 */
            if (blurred) memcpy(blurred + (size_t)k * n, b, sizeof(float) * n);
            int ix = (int)px, iy = (int)py; /* :100 astype(int32) */
#define TAP(yy, xx) ((double)logf(clip_nan(b[((yy) < 0 ? 0 : ((yy) > h - 1 ? h - 1 : (yy))) * w + \
                                              ((xx) < 0 ? 0 : ((xx) > w - 1 ? w - 1 : (xx)))])))
            double I = TAP(iy, ix), Ix1 = TAP(iy, ix + 1), Ix1_ = TAP(iy, ix - 1);
            double Iy1 = TAP(iy + 1, ix), Iy1_ = TAP(iy - 1, ix);
            double Ix1y1 = TAP(iy + 1, ix + 1), Ix1_y1_ = TAP(iy - 1, ix - 1);
#undef TAP
            double dx = 0.5 * (Ix1 - Ix1_), dy = 0.5 * (Iy1 - Iy1_);
            double dxx = Ix1 - 2 * I + Ix1_, dyy = Iy1 - 2 * I + Iy1_;
            double dxy = 0.5 * (Ix1y1 - Ix1 - Iy1 + I + I - Ix1_ - Iy1_ + Ix1_y1_);
            double inv[4];
            inv2x2(dxx, dxy, dxy, dyy, inv);
            cx = (double)ix - (inv[0] * dx + inv[1] * dy); /* :135-144 (coords were int32) */
            cy = (double)iy - (inv[2] * dx + inv[3] * dy);
        }
        if (post_process) {
            preds_in[k * 2 + 0] = cx / (w - 1.0) * (4 * w - 1.0); /* :178-179 */
            preds_in[k * 2 + 1] = cy / (h - 1.0) * (4 * h - 1.0);
        } else {
            preds_in[k * 2 + 0] = (double)((float)cx / (float)(w - 1.0) * (float)(4 * w - 1.0));
            preds_in[k * 2 + 1] = (double)((float)cy / (float)(h - 1.0) * (float)(4 * h - 1.0));
        }
        transform_pred(cx, cy, center, scale, cs_f64, post_process, w, h, &preds[k * 2], &preds[k * 2 + 1]);
    }
    free(m);
    free(b);
}

/* a15: get_final_preds 'offset' (UDPv1 voting) for ONE person.  inference.py:156-174.
 * out: [3K,h,w] float32 (already flip-averaged).  maxvals are maxima of the BLURRED
 * classification map; idx is its argmax. */
void udp_o_decode_offset(const float *out, int K, int w, int h, double kpd, const void *center,
                         const void *scale, int cs_f64, double *preds, float *maxvals,
                         double *preds_in, int32_t *idx)
{
    int n = h * w;
    float kf = (float)kpd;
    float *hb = (float *)malloc(sizeof(float) * n);
    float *ox = (float *)malloc(sizeof(float) * n);
    float *oy = (float *)malloc(sizeof(float) * n);
    for (int k = 0; k < K; ++k) {
        const float *c0 = out + ((size_t)k * 3 + 0) * n;
        const float *c1 = out + ((size_t)k * 3 + 1) * n;
        const float *c2 = out + ((size_t)k * 3 + 2) * n;
        for (int i = 0; i < n; ++i) { ox[i] = c1[i] * kf; oy[i] = c2[i] * kf; } /* :161-162 */
        udp_o_gaussian_blur(c0, hb, h, w, 15); /* :165 */
        udp_o_gaussian_blur(ox, ox, h, w, 7);  /* :166 */
        udp_o_gaussian_blur(oy, oy, h, w, 7);  /* :167 */
        int bi;
        float mv;
        argmax_map(hb, n, &bi, &mv);
        idx[k] = bi;
        maxvals[k] = mv;
        float px = (float)(bi % w), py = floorf((float)bi / (float)w);
        float mask = mv > 0.0f ? 1.0f : 0.0f;
        px *= mask;
        py *= mask;
        int ix = (int)px, iy = (int)py; /* :171-172 */
        px = px + ox[iy * w + ix];       /* :173-174, float32 */
        py = py + oy[iy * w + ix];
        preds_in[k * 2 + 0] = (double)(px / (float)(w - 1.0) * (float)(4 * w - 1.0));
        preds_in[k * 2 + 1] = (double)(py / (float)(h - 1.0) * (float)(4 * h - 1.0));
        transform_pred(px, py, center, scale, cs_f64, 0, w, h, &preds[k * 2], &preds[k * 2 + 1]);
    }
    free(hb);
    free(ox);
    free(oy);
}

/* ---------------------------------------------------------------------------------------
 * SURVEY 8(f)-4 (second half): the float64 Gaussian blur of RSN's legacy decoder
 * (RSN/exps/RSN18.coco/test.py:39-40: cv2.GaussianBlur on a float64 map, sigma 0).
 * The arithmetic lives in OpenCV (not vendored; container has 4.13.0).  Probed here against cv2
 * (tests/test_rsn_legacy.py repeats the probe on whatever host runs it): for CV_64F the row
 * filter is the generic one -- k[0]*x[0], then a left-to-right chain that the library's build
 * FUSES (fma) -- and the column filter is the symmetric one -- k[c]*x[c], then
 * + k[c+i]*(x[c+i] + x[c-i]) with a separately rounded multiply and add.  Border: reflect-101.
 * `k` is cv2.getGaussianKernel(n, 0, CV_64F) (tables in oracle.py, taken from cv2 itself).
 * ------------------------------------------------------------------------------------- */
void udp_o_gaussian_blur_f64(const double *src, double *dst, int h, int w, int n, const double *k)
{
    const int r = n / 2;
    double *t = (double *)malloc(sizeof(double) * (size_t)h * w);
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            double s = k[0] * src[y * w + reflect101(x - r, w)];
            for (int i = 1; i < n; ++i) s = fma(k[i], src[y * w + reflect101(x - r + i, w)], s);
            t[y * w + x] = s;
        }
    for (int y = 0; y < h; ++y)
        for (int x = 0; x < w; ++x) {
            double s = k[r] * t[y * w + x];
            for (int i = 1; i <= r; ++i) {
                const double pair = t[reflect101(y + i, h) * w + x] + t[reflect101(y - i, h) * w + x];
                const double prod = k[r + i] * pair;
                s = s + prod;
            }
            dst[y * w + x] = s;
        }
    free(t);
}
