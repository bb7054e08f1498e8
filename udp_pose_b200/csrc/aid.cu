// aid.cu -- SURVEY 8(f)-5: the AID information-dropping augmentations applied to the crop.
//
// Reference: Cutout (deep_hrnet/lib/utils/transforms.py:186-223) and HideAndSeek (:144-183), applied to
// the warped uint8 crop at lib/dataset/JointsDataset.py:230-235 (train only), before ToTensor.
// Their random numbers come from numpy's global generator; the host side (reference_api.Cutout /
// HideAndSeek) draws them in the reference's order and hands the drawn PARAMETERS to this kernel, so
// the erase itself is deterministic:
//   Cutout      : per patch (cx, cy, rx, ry) float64; pixel (row y, col x) is zeroed when
//                 ((cx - x) / rx)^2 + ((cy - y) / ry)^2 <= 1   (float64, :209-213)
//   HideAndSeek : a grid size g and one bit per cell in draw order (outer loop over range(0, width, g),
//                 inner over range(0, height, g)).  The reference indexes img[x:x_end, y:y_end] with the
//                 WIDTH coordinate on the row axis (:175): reproduced, so pixel (row, col) belongs to
//                 cell (row / g, col / g) when row < width and col < height.
// One CTA per person, threads over pixels; nothing is read, erased pixels get three zero bytes.
#include "common.cuh"

namespace udp {

constexpr int AID_THREADS = 256;
constexpr int AID_MASK_WORDS = UDP_AID_MASK_WORDS;   // up to 512 cells: a 384x288 crop with the reference's smallest grid (16) has 18 * 24 = 432

__global__ void __launch_bounds__(AID_THREADS)
aid_erase_kernel(uint8_t *__restrict__ img, int H, int W, const double *__restrict__ ellipses,
                 const int32_t *__restrict__ n_ellipses, int max_ellipses, const int32_t *__restrict__ grid_size,
                 const uint32_t *__restrict__ hide_mask)
{
    __shared__ double s_e[4 * 8];
    __shared__ uint32_t s_m[AID_MASK_WORDS];
    const int n = blockIdx.x;
    const int ne = n_ellipses ? min(n_ellipses[n], min(max_ellipses, 8)) : 0;
    const int g = grid_size ? grid_size[n] : 0;
    if (ne <= 0 && g <= 0) return;
    if (threadIdx.x < 4 * ne) s_e[threadIdx.x] = ellipses[(size_t)n * max_ellipses * 4 + threadIdx.x];
    if (threadIdx.x < AID_MASK_WORDS) s_m[threadIdx.x] = g > 0 ? hide_mask[(size_t)n * AID_MASK_WORDS + threadIdx.x] : 0u;
    __syncthreads();
    const int ny = g > 0 ? (H + g - 1) / g : 0;  // cells along the inner loop (range(0, height, g))
    uint8_t *base = img + (size_t)n * H * W * 3;
    for (int y = 0; y < H; ++y) {
        for (int x = threadIdx.x; x < W; x += AID_THREADS) {
            bool kill = false;
            for (int e = 0; e < ne; ++e) {
                const double xo = __ddiv_rn(__dsub_rn(s_e[4 * e], (double)x), s_e[4 * e + 2]);
                const double yo = __ddiv_rn(__dsub_rn(s_e[4 * e + 1], (double)y), s_e[4 * e + 3]);
                kill |= __dadd_rn(__dmul_rn(xo, xo), __dmul_rn(yo, yo)) <= 1.0;
            }
            if (g > 0 && y < W && x < H) {  // the reference's transposed indexing: rows are clipped at `width`
                const int cell = (y / g) * ny + (x / g);
                kill |= (s_m[cell >> 5] >> (cell & 31)) & 1u;
            }
            if (kill) {
                uint8_t *px = base + ((size_t)y * W + x) * 3;
                px[0] = 0; px[1] = 0; px[2] = 0;
            }
        }
    }
}

}  // namespace udp

extern "C" int udp_aid_erase(uint8_t *img, int N, int H, int W, const double *ellipses, const int32_t *n_ellipses,
                             int max_ellipses, const int32_t *grid_size, const uint32_t *hide_mask, int min_grid,
                             void *stream)
{
    if (N < 0 || H < 1 || W < 1 || max_ellipses < 0 || max_ellipses > 8) return UDP_EINVAL;
    if (N == 0) return 0;
    if (!img || (n_ellipses && !ellipses) || (grid_size && !hide_mask)) return UDP_EINVAL;
    if (grid_size) {  // min_grid: the smallest non-zero grid size in use; its cells must fit the 512-bit mask
        if (min_grid < 1) return UDP_EINVAL;
        const long long cells = (long long)((W + min_grid - 1) / min_grid) * ((H + min_grid - 1) / min_grid);
        if (cells > 32 * udp::AID_MASK_WORDS) return UDP_EINVAL;
    }
    udp::aid_erase_kernel<<<N, udp::AID_THREADS, 0, (cudaStream_t)stream>>>(img, H, W, ellipses, n_ellipses, max_ellipses,
                                                                            grid_size, hide_mask);
    UDP_LAUNCH_CHECK();
    return 0;
}
