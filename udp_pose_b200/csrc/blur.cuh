// blur.cuh -- shared-memory separable Gaussian blur, bit-exact with cv2.GaussianBlur(f32,
// (n,n), 0) (BORDER_REFLECT_101) for n = 7 and n = 15, plus the flip-fused tile load and the
// first-index argmax reduction used by both decoders.
//
// Reference call sites: deep_hrnet/lib/core/inference.py:76 (7x7, post), :165 (15x15),
// :166-167 (7x7 on the offsets).  OpenCV's summation order (re-checked on the
// installed cv2 in tests/test_oracle_cpu.py):
//   row pass    s = a[x-r]*k[0];  s = fmaf(a[x-r+i], k[i], s)  i = 1..n-1   (left -> right)
//   column pass s = fmaf(T[y], k[r], 0); s = fmaf(T[y+i] + T[y-i], k[r+i], s)  i = 1..r
// CUDA's fmaf is IEEE, so following this order reproduces cv2 bit for bit.
//
// Shared-memory layout for one h x w map (w % 4 == 0):
//   A : h rows of stride SA = w + 2*PAD floats, data at [PAD, PAD+w), PAD = 4 (r=3) / 8 (r=7);
//       the pads hold the reflect-101 columns so the row pass is branch-free LDS.128;
//   T : h x w, the row-pass result; the column pass reads T with reflected row indices.
#pragma once
#include "common.cuh"

namespace udp {

template <int R> struct Taps;
template <> struct Taps<3> {
    // OpenCV's fixed table for ksize 7, sigma <= 0: [2,7,14,18,14,7,2]/64
    static __device__ __forceinline__ float k(int i)
    {
        constexpr float t[7] = {0.03125f, 0.109375f, 0.21875f, 0.28125f, 0.21875f, 0.109375f, 0.03125f};
        return t[i];
    }
};
template <> struct Taps<7> {
    // cv2.getGaussianKernel(15, 0, CV_32F): sigma 2.6, normalised in double, stored float32
    static __device__ __forceinline__ float k(int i)
    {
        constexpr float t[15] = {0x1.0d2c46p-8f, 0x1.600a4cp-7f, 0x1.8d1b96p-6f, 0x1.825954p-5f,
                                 0x1.4431eep-4f, 0x1.d543dap-4f, 0x1.24ec9cp-3f, 0x1.3b6962p-3f,
                                 0x1.24ec9cp-3f, 0x1.d543dap-4f, 0x1.4431eep-4f, 0x1.825954p-5f,
                                 0x1.8d1b96p-6f, 0x1.600a4cp-7f, 0x1.0d2c46p-8f};
        return t[i];
    }
};

template <int R> struct BlurGeom {
    static constexpr int PAD = (R <= 4) ? 4 : 8;
    static __host__ __device__ int stride_a(int w) { return w + 2 * PAD; }
    static __host__ __device__ size_t smem_floats(int w, int h) { return (size_t)h * stride_a(w) + (size_t)h * w; }
};

__device__ __forceinline__ int reflect101(int i, int n)
{
    i = i < 0 ? -i : i;
    return i >= n ? 2 * (n - 1) - i : i;
}

// np.argmax order: larger value wins, NaN beats everything, the lower flat index wins ties.
__device__ __forceinline__ bool arg_better(float v1, int i1, float v2, int i2)
{
    const bool n1 = v1 != v1, n2 = v2 != v2;
    if (n1 || n2) return n1 && (!n2 || i1 < i2);
    return v1 > v2 || (v1 == v2 && i1 < i2);
}

// Block-wide (value, index) argmax.  red_v/red_i: shared scratch of >= 32 entries.
// Result valid in every thread.
__device__ __forceinline__ void block_argmax(float &v, int &i, float *red_v, int *red_i)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = (blockDim.x + 31) >> 5;
    const unsigned lanes = warp_lanes();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ov = shfl_xor_or_self(lanes, v, o);
        const int oi = shfl_xor_or_self(lanes, i, o);
        if (arg_better(ov, oi, v, i)) { v = ov; i = oi; }
    }
    if (lane == 0) { red_v[warp] = v; red_i[warp] = i; }
    __syncthreads();
    v = red_v[0]; i = red_i[0];
    for (int q = 1; q < nwarps; ++q)
        if (arg_better(red_v[q], red_i[q], v, i)) { v = red_v[q]; i = red_i[q]; }
    __syncthreads();  // scratch may be reused
}

// Load one map into A (padded rows), fusing the flip test:
//   v = (out + sign * flipped[mirror]) * 0.5   (utils/transforms.py:15-47 + core/function.py:171)
// then the optional RSN pre-division (prm test.py:185), and track the raw first-index argmax.
// `fl` already points at the partner joint's map (or is NULL).
template <int R>
__device__ __forceinline__ void load_tile(const float *__restrict__ src, const float *__restrict__ fl,
                                          float fsign, float pre_div, int w, int h, float *A,
                                          float &best_v, int &best_i)
{
    constexpr int PAD = BlurGeom<R>::PAD;
    const int SA = w + 2 * PAD, w4 = w >> 2, per_map = w4 * h;
    best_v = -INFINITY;
    best_i = 0x7fffffff;
    bool first = true;
    for (int j = threadIdx.x; j < per_map; j += blockDim.x) {
        const int y = j / w4, x = (j - y * w4) << 2;
        float4 a = ld_stream_f4(src + (size_t)j * 4);
        if (fl) {
            const float4 f = ld_stream_f4(fl + (size_t)y * w + (w - 4 - x));
            a.x = __fmul_rn(__fadd_rn(a.x, fsign * f.w), 0.5f);
            a.y = __fmul_rn(__fadd_rn(a.y, fsign * f.z), 0.5f);
            a.z = __fmul_rn(__fadd_rn(a.z, fsign * f.y), 0.5f);
            a.w = __fmul_rn(__fadd_rn(a.w, fsign * f.x), 0.5f);
        }
        if (pre_div != 0.0f) {
            a.x = __fdiv_rn(a.x, pre_div); a.y = __fdiv_rn(a.y, pre_div);
            a.z = __fdiv_rn(a.z, pre_div); a.w = __fdiv_rn(a.w, pre_div);
        }
        *reinterpret_cast<float4 *>(A + y * SA + PAD + x) = a;
        const int base = j << 2;
        if (first) { best_v = a.x; best_i = base; first = false; }
        else if (a.x > best_v || (a.x != a.x && best_v == best_v)) { best_v = a.x; best_i = base; }
        if (a.y > best_v || (a.y != a.y && best_v == best_v)) { best_v = a.y; best_i = base + 1; }
        if (a.z > best_v || (a.z != a.z && best_v == best_v)) { best_v = a.z; best_i = base + 2; }
        if (a.w > best_v || (a.w != a.w && best_v == best_v)) { best_v = a.w; best_i = base + 3; }
    }
}

// Fill the reflect-101 pad columns of A (call after a __syncthreads following load_tile).
template <int R>
__device__ __forceinline__ void fill_pads(float *A, int w, int h)
{
    constexpr int PAD = BlurGeom<R>::PAD;
    const int SA = w + 2 * PAD;
    for (int t = threadIdx.x; t < h * 2 * R; t += blockDim.x) {
        const int y = t / (2 * R), q = t - y * 2 * R;
        float *row = A + y * SA + PAD;
        if (q < R) row[-1 - q] = row[1 + q];
        else { const int i = q - R; row[w + i] = row[w - 2 - i]; }
    }
}

// Row pass A -> T.  Each work item = 4 adjacent outputs, inputs via aligned LDS.128.
template <int R>
__device__ __forceinline__ void row_pass(const float *A, float *T, int w, int h)
{
    constexpr int PAD = BlurGeom<R>::PAD;
    constexpr int NV = (PAD * 2 + 4) / 4;  // float4 loads: 3 (r=3) / 5 (r=7)
    constexpr int OFF = PAD - R;           // first tap of output 0 inside the loaded span
    const int SA = w + 2 * PAD, w4 = w >> 2, per_map = w4 * h;
    for (int j = threadIdx.x; j < per_map; j += blockDim.x) {
        const int y = j / w4, x = (j - y * w4) << 2;
        const float4 *p = reinterpret_cast<const float4 *>(A + y * SA + x);  // = data[x - PAD]
        float v[NV * 4];
#pragma unroll
        for (int q = 0; q < NV; ++q) {
            const float4 t = p[q];
            v[4 * q] = t.x; v[4 * q + 1] = t.y; v[4 * q + 2] = t.z; v[4 * q + 3] = t.w;
        }
        float o[4];
#pragma unroll
        for (int c = 0; c < 4; ++c) {
            float s = __fmul_rn(v[OFF + c], Taps<R>::k(0));
#pragma unroll
            for (int i = 1; i < 2 * R + 1; ++i) s = fmaf(v[OFF + c + i], Taps<R>::k(i), s);
            o[c] = s;
        }
        *reinterpret_cast<float4 *>(T + y * w + x) = make_float4(o[0], o[1], o[2], o[3]);
    }
}

// Column pass over T.  Work item = 4 columns x L rows; the L + 2R input rows are held in
// registers (compile-time indexing, no sliding-window moves).  emit(y, x, float4) receives
// every finished group of 4 outputs.
template <int R, int L, typename Emit>
__device__ __forceinline__ void col_pass(const float *T, int w, int h, Emit emit)
{
    const int w4 = w >> 2, segs = (h + L - 1) / L, items = w4 * segs;
    for (int it = threadIdx.x; it < items; it += blockDim.x) {
        const int seg = it / w4, x = (it - seg * w4) << 2, y0 = seg * L;
        float4 r[L + 2 * R];
#pragma unroll
        for (int i = 0; i < L + 2 * R; ++i) {
            const int yy = reflect101(y0 + i - R, h);
            r[i] = *reinterpret_cast<const float4 *>(T + yy * w + x);
        }
#pragma unroll
        for (int o = 0; o < L; ++o) {
            if (y0 + o < h) {
                const float kc = Taps<R>::k(R);
                float4 s;
                s.x = fmaf(r[o + R].x, kc, 0.0f); s.y = fmaf(r[o + R].y, kc, 0.0f);
                s.z = fmaf(r[o + R].z, kc, 0.0f); s.w = fmaf(r[o + R].w, kc, 0.0f);
#pragma unroll
                for (int i = 1; i <= R; ++i) {
                    const float ki = Taps<R>::k(R + i);
                    s.x = fmaf(__fadd_rn(r[o + R + i].x, r[o + R - i].x), ki, s.x);
                    s.y = fmaf(__fadd_rn(r[o + R + i].y, r[o + R - i].y), ki, s.y);
                    s.z = fmaf(__fadd_rn(r[o + R + i].z, r[o + R - i].z), ki, s.z);
                    s.w = fmaf(__fadd_rn(r[o + R + i].w, r[o + R - i].w), ki, s.w);
                }
                emit(y0 + o, x, s);
            }
        }
    }
}

// UDP inverse affine of one coordinate pair (transform_preds, inference.py:20-27) honouring
// the dtypes of coords (f64 after post(), else f32) and of center/scale.
__device__ __forceinline__ void transform_pred(double cx, double cy, const CS &cs, int cs_f64,
                                               bool f64coords, int w, int h, double &ox, double &oy)
{
    if (cs_f64) {
        const double sx = __ddiv_rn(cs.st0, w - 1.0), sy = __ddiv_rn(cs.st1, h - 1.0);
        ox = __dsub_rn(__dadd_rn(__dmul_rn(cx, sx), cs.c0), __dmul_rn(cs.st0, 0.5));
        oy = __dsub_rn(__dadd_rn(__dmul_rn(cy, sy), cs.c1), __dmul_rn(cs.st1, 0.5));
        if (!f64coords) { ox = (double)(float)ox; oy = (double)(float)oy; }
    } else {
        const float sx = __fdiv_rn(cs.st0f, (float)(w - 1.0)), sy = __fdiv_rn(cs.st1f, (float)(h - 1.0));
        if (f64coords) {
            ox = __dsub_rn(__dadd_rn(__dmul_rn(cx, (double)sx), cs.c0), (double)__fmul_rn(cs.st0f, 0.5f));
            oy = __dsub_rn(__dadd_rn(__dmul_rn(cy, (double)sy), cs.c1), (double)__fmul_rn(cs.st1f, 0.5f));
        } else {
            float fx = __fadd_rn(__fmul_rn((float)cx, sx), cs.c0f);
            float fy = __fadd_rn(__fmul_rn((float)cy, sy), cs.c1f);
            ox = (double)__fsub_rn(fx, __fmul_rn(cs.st0f, 0.5f));
            oy = (double)__fsub_rn(fy, __fmul_rn(cs.st1f, 0.5f));
        }
    }
}

}  // namespace udp
