/*
 * udp_oracle_encode.c -- CPU restatement of the target encoders (TEST INFRASTRUCTURE ONLY,
 * see the header of udp_oracle.c for the rules and the parity status: pinned by
 * tests/golden/*.npz generated from the unmodified reference).
 */
#include <math.h>
#include <stdint.h>
#include <string.h>

/* ------------------------------------------------------------------------------------ */
/* a5  JointsDataset.generate_target, 'gaussian'  deep_hrnet/lib/dataset/JointsDataset.py */
/*     :291-348 (+ joints_weight :382-383).                                               */
/* joints, joints_vis: [K,3] float64.  target: [K,h,w] float32, target_weight: [K] f32.   */
/* The window is evaluated in float64 (numpy-2: float32 array - np.float64 scalar) and    */
/* rounded to float32 on store.                                                           */
/* ------------------------------------------------------------------------------------ */
void udp_o_encode_gaussian(const double *joints, const double *joints_vis, int K,
                           int W, int H, int w, int h, double sigma,
                           const float *joints_weight /* K or NULL */,
                           float *target, float *target_weight)
{
    double fsx = (W - 1.0) / (w - 1.0), fsy = (H - 1.0) / (h - 1.0); /* :311 */
    double tmp = sigma * 3;                                          /* :307 */
    double size = 2 * tmp + 1;                                       /* :325 */
    int n = (int)ceil(size);                                         /* len(arange(0,size,1)) */
    double half = floor(size / 2);                                   /* size // 2 */
    double den = 2 * sigma * sigma;
    memset(target, 0, sizeof(float) * (size_t)K * h * w);
    for (int k = 0; k < K; ++k) {
        float tw = (float)joints_vis[k * 3 + 0]; /* :297-298 */
        double jx = joints[k * 3 + 0], jy = joints[k * 3 + 1];
        int mu_x = (int)(jx / fsx + 0.5); /* int() truncates toward zero :313-314 */
        int mu_y = (int)(jy / fsy + 0.5);
        int ul0 = (int)(mu_x - tmp), ul1 = (int)(mu_y - tmp);
        int br0 = (int)(mu_x + tmp + 1), br1 = (int)(mu_y + tmp + 1);
        if (ul0 >= w || ul1 >= h || br0 < 0 || br1 < 0) { /* :318-322 */
            target_weight[k] = 0.0f;
            continue;
        }
        double x0 = half + (jx / fsx - mu_x); /* :330-334 */
        double y0 = half + (jy / fsy - mu_y);
        int gx0 = ul0 < 0 ? -ul0 : 0, gx1 = (br0 < w ? br0 : w) - ul0; /* :339-340 */
        int gy0 = ul1 < 0 ? -ul1 : 0, gy1 = (br1 < h ? br1 : h) - ul1;
        int ix0 = ul0 > 0 ? ul0 : 0, iy0 = ul1 > 0 ? ul1 : 0; /* :342-343 */
        if (gx1 > n) gx1 = n; /* numpy slicing clamps to the array extent */
        if (gy1 > n) gy1 = n;
        if (tw > 0.5f) { /* :345-348 */
            float *t = target + (size_t)k * h * w;
            for (int gy = gy0; gy < gy1; ++gy)
                for (int gx = gx0; gx < gx1; ++gx) {
                    double dx = (double)(float)gx - x0, dy = (double)(float)gy - y0;
                    double g = exp(-(dx * dx + dy * dy) / den); /* :336 */
                    t[(iy0 + gy - gy0) * w + (ix0 + gx - gx0)] = (float)g;
                }
        }
        target_weight[k] = tw;
    }
    if (joints_weight)
        for (int k = 0; k < K; ++k) target_weight[k] *= joints_weight[k];
}

/* ------------------------------------------------------------------------------------ */
/* a6  generate_target, 'offset' (UDPv1)  JointsDataset.py:349-381                        */
/* target: [3K,h,w] float32 with channel order (j,0)=cls, (j,1)=x off, (j,2)=y off.       */
/* No window test and no weight zeroing in this mode; mu is NOT rounded.                  */
/* ------------------------------------------------------------------------------------ */
void udp_o_encode_offset(const double *joints, const double *joints_vis, int K,
                         int W, int H, int w, int h, double kpd,
                         const float *joints_weight, float *target, float *target_weight)
{
    double fsx = (W - 1.0) / (w - 1.0), fsy = (H - 1.0) / (h - 1.0); /* :365 */
    memset(target, 0, sizeof(float) * (size_t)K * 3 * h * w);
    for (int k = 0; k < K; ++k) {
        float tw = (float)joints_vis[k * 3 + 0];
        target_weight[k] = tw;
        if (!(tw > 0.5f)) continue; /* :376 */
        double mu_x = joints[k * 3 + 0] / fsx, mu_y = joints[k * 3 + 1] / fsy; /* :367-368 */
        float *c0 = target + ((size_t)k * 3 + 0) * h * w;
        float *c1 = target + ((size_t)k * 3 + 1) * h * w;
        float *c2 = target + ((size_t)k * 3 + 2) * h * w;
        for (int y = 0; y < h; ++y) {
            double yo = (mu_y - (double)y) / kpd; /* :372 */
            for (int x = 0; x < w; ++x) {
                double xo = (mu_x - (double)x) / kpd; /* :371 */
                double dis = xo * xo + yo * yo;       /* :374 */
                if (dis <= 1 && dis >= 0) {           /* :375 */
                    c0[y * w + x] = 1.0f;
                    c1[y * w + x] = (float)xo;
                    c2[y * w + x] = (float)yo;
                }
            }
        }
    }
    if (joints_weight)
        for (int k = 0; k < K; ++k) target_weight[k] *= joints_weight[k];
}

/* ------------------------------------------------------------------------------------ */
/* a7  RSN generate_heatmap x S kernel sizes  RSN/dataset/JointsDataset.py:276-291,        */
/*     loop :214-220, sigma table :19-25.  labels: [S,K,h,w] float32.                     */
/* valid: [K] (joints_vis[:,0] as float); a joint is drawn iff valid >= 1.                */
/* ------------------------------------------------------------------------------------ */
void udp_o_encode_rsn(const double *joints /* [K,jstride] */, int jstride,
                      const float *valid, int K, int W, int H, int w, int h,
                      const double *sigmas, int S, float *labels)
{
    memset(labels, 0, sizeof(float) * (size_t)S * K * h * w);
    for (int s = 0; s < S; ++s) {
        double den = -2.0 * (sigmas[s] * sigmas[s]); /* :289 */
        for (int k = 0; k < K; ++k) {
            if (valid[k] < 1) continue; /* :281 */
            double ty = joints[k * jstride + 1] * (h - 1.0) / (H - 1.0); /* :283 */
            double tx = joints[k * jstride + 0] * (w - 1.0) / (W - 1.0); /* :284 */
            float *t = labels + ((size_t)s * K + k) * h * w;
            for (int y = 0; y < h; ++y) {
                double yo = ty - (double)y;
                for (int x = 0; x < w; ++x) {
                    double xo = tx - (double)x;
                    double dis = xo * xo + yo * yo;
                    t[y * w + x] = (float)(255.0 * exp(dis / den));
                }
            }
        }
    }
}

/* ------------------------------------------------------------------------------------ */
/* a3' RSN forward joint map: get_warpmatrix_inverse + affine_transform + visibility cull */
/*     RSN/dataset/JointsDataset.py:27-47, :203-211; RSN/lib/utils/transforms.py:38-41.   */
/* center/scale float32 (RSN/dataset/COCO/coco.py builds them as float32).                */
/* joints [K,jstride] float64 in/out (x,y updated in place where vis>0), vis [K] in/out.  */
/* ------------------------------------------------------------------------------------ */
void udp_o_rsn_warp_matrix_inverse(double theta_deg, const float *c, const float *s,
                                   int W, int H, float *M)
{
    double th = theta_deg / 180.0 * 3.141592653589793;
    double co = cos(th), sn = sin(th);
    float st0 = s[0] * 200.0f, st1 = s[1] * 200.0f;
    float si0 = c[0] * 2.0f, si1 = c[1] * 2.0f;
    double scale_x = (W - 1.0) / (double)st0, scale_y = (H - 1.0) / (double)st1;
    float cf = (float)co, sf = (float)sn;
    M[0] = (float)(co * scale_x);
    M[1] = (float)(-sn * scale_x);
    M[3] = (float)(sn * scale_y);
    M[4] = (float)(co * scale_y);
    {
        float a = (-0.5f * si0) * cf;
        float b = (0.5f * si1) * sf;
        float d = a + b;
        d = d + 0.5f * st0;
        M[2] = (float)(scale_x * (double)d);
    }
    {
        float a = (-0.5f * si0) * sf;
        float b = (0.5f * si1) * cf;
        float d = a - b;
        d = d + 0.5f * st1;
        M[5] = (float)(scale_y * (double)d);
    }
}

void udp_o_rsn_transform_joints(double *joints, int jstride, double *vis, int vstride, int K,
                                const float *Minv, int W, int H)
{
    for (int k = 0; k < K; ++k) {
        if (!(vis[k * vstride] > 0)) continue;
        double x = joints[k * jstride + 0], y = joints[k * jstride + 1];
        double nx = (double)Minv[0] * x + (double)Minv[1] * y + (double)Minv[2];
        double ny = (double)Minv[3] * x + (double)Minv[4] * y + (double)Minv[5];
        joints[k * jstride + 0] = nx;
        joints[k * jstride + 1] = ny;
        if (nx < 0 || nx > W - 1 || ny < 0 || ny > H - 1) vis[k * vstride] = 0;
    }
}
