/*
 * udp_codec.h -- C ABI of the B200 (sm_100a) UDP heatmap codec.
 *
 * This is the drop-in boundary for the codec hot path of realphongha/UDP-Pose.  The
 * reference has no FFI of its own: the path sits behind plain Python functions
 * (SURVEY.md 8b).  Each entry point below names the reference function(s) it replaces,
 * paths relative to the reference root.  INTEGRATION.md shows the ctypes binding a
 * maintainer adds on the reference side.
 *
 * Conventions
 *  - every pointer is a DEVICE pointer unless the comment says HOST; row-major, dense;
 *    float map tensors (targets, network outputs, blurred) must be 16-byte aligned (UDP_EINVAL
 *    otherwise) -- any tensor that starts an allocation is;
 *  - `stream` is a cudaStream_t (NULL = default stream); calls are asynchronous;
 *  - return value: 0 on success, otherwise a cudaError_t (> 0) or UDP_EINVAL (< 0);
 *    no exceptions cross the ABI; no entry point allocates device memory per call or synchronises
 *    with the host.  Library-owned state, all of it created once per device on first use and
 *    documented at the entry points that use it: udp_warp_affine_u8 / udp_warp_affine_normalize_f32
 *    keep one side stream + event pair (the cp.async kernel for the few persons the TMA kernel
 *    cannot take runs there, forked from and joined back into `stream`; not used while `stream` is
 *    being captured) and claim slots of a static 8192-entry tensor-map pool in global memory;
 *    the TMA paths fetch cuTensorMapEncodeTiled from the driver once;
 *  - `cs_f64`: dtype of the center/scale arrays -- 1 = float64 (MPII records,
 *    lib/dataset/mpii.py:55-56), 0 = float32 (COCO records, lib/dataset/coco.py:223-227).
 *    The reference's arithmetic depends on that dtype (numpy promotion), so it is part of
 *    the contract;
 *  - N = number of person instances, K = joints, (W,H) network input, (w,h) heatmap.
 */
#ifndef UDP_CODEC_H
#define UDP_CODEC_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define UDP_EINVAL (-1)
#define UDP_ESCRATCH (-2) /* the request does not fit a library-owned, fixed-size workspace (udp_accuracy_pck_noalloc) */
#define UDP_MAX_SIGMAS 8

/* Library version (major*10000 + minor*100 + patch) and the SM architecture compiled in. */
int udp_codec_version(void);
int udp_codec_sm_arch(void);

/* a1  get_warpmatrix(r, c*2.0, image_size-1.0, s)
 *     deep_hrnet/lib/dataset/JointsDataset.py:29-49, call :226 (RSN twin RSN/dataset/JointsDataset.py:49-69).
 *     rot_deg [N] float64 or NULL (= 0).  M_out [N,2,3] float32 (dst->src, for WARP_INVERSE_MAP). */
int udp_warp_matrices(const void *center, const void *scale, int cs_f64, const double *rot_deg,
                      int W, int H, float *M_out, int N, void *stream);

/* a2  cv2.warpAffine(img u8 HWC3, M, (W,H), flags=WARP_INVERSE_MAP|INTER_LINEAR), border 0
 *     call sites JointsDataset.py:227, RSN/dataset/JointsDataset.py:197-198.
 *     src_ptrs [N] device array of device pointers (one image per person, may repeat);
 *     src_h/src_w/src_stride [N] int32 (stride in bytes).  dst [N,H,W,3] uint8.
 *     Bit-exact with OpenCV's fixed-point path (Q10 coordinates, Q5 fractions, Q15 weights).
 *     Images whose stride is a multiple of 16 bytes are fetched by the TMA unit (one tensor load
 *     per 32 x {32,16,8} output tile, hardware zero fill outside the image); any other stride /
 *     alignment works through a cp.async path.  udp_warp_affine_launches() = kernel launches
 *     one call makes for this crop size (2 when both paths are armed, else 1).  When both are armed the
 *     cp.async launch runs on a library-owned side stream forked from and joined back into `stream`
 *     (created once per device on first use; UDP_CROP_FORK=0 in the environment disables the fork) --
 *     ordering with respect to `stream` is unchanged. */
int udp_warp_affine_u8(const uint8_t *const *src_ptrs, const int32_t *src_h, const int32_t *src_w,
                       const int32_t *src_stride, const float *M, uint8_t *dst, int W, int H,
                       int N, void *stream);
int udp_warp_affine_launches(int W, int H);

/* a2 + SURVEY 8(f)-1: the crop with `transforms.ToTensor()` + `transforms.Normalize(mean, std)` fused
 *     (deep_hrnet/tools/train.py:127-143, applied at JointsDataset.py:237-238), and the mirrored
 *     network input of the flip test, `np.flip(input, 3)` (lib/core/function.py:154-155), as an
 *     optional second output.  dst, dst_flipped [N,3,H,W] float32 = ((u8 / 255) - mean[c]) / std[c],
 *     every step rounded as torch rounds it (bit-exact with torchvision on the bit-exact crop);
 *     the uint8 crop is never written.  mean, std: HOST float[3].  dst_flipped may be NULL.
 *     W and H must be multiples of 32 (every reference config), else UDP_EINVAL. */
int udp_warp_affine_normalize_f32(const uint8_t *const *src_ptrs, const int32_t *src_h,
                                  const int32_t *src_w, const int32_t *src_stride, const float *M,
                                  const float *mean, const float *std_, float *dst, float *dst_flipped,
                                  int W, int H, int N, void *stream);

/* a3  rotate_points(joints[:, 0:2], r, c, image_size, s, do_clip)  JointsDataset.py:51-73, call :228.
 *     joints [N,K,jstride] float64 (x,y at 0,1), out [N,K,ostride] float64; may alias. */
int udp_rotate_points(const double *joints, int jstride, const double *rot_deg, const void *center,
                      const void *scale, int cs_f64, int W, int H, int do_clip, double *out,
                      int ostride, int K, int N, void *stream);

/* a3' RSN: get_warpmatrix_inverse + affine_transform + visibility cull
 *     RSN/dataset/JointsDataset.py:27-47, :203-211; RSN/lib/utils/transforms.py:38-41.
 *     joints [N,K,jstride] f64 and vis [N,K,vstride] f64 are updated IN PLACE; center/scale f32.
 *     Minv_out [N,2,3] float32 (nullable). */
int udp_rsn_transform_joints(double *joints, int jstride, double *vis, int vstride,
                             const double *rot_deg, const float *center, const float *scale,
                             int W, int H, float *Minv_out, int K, int N, void *stream);

/* a4  fliplr_joints(joints, joints_vis, width, matched_parts)  deep_hrnet/lib/utils/transforms.py:50-64
 *     (mul_vis = 1: returns joints*joints_vis) and RSN flip_joints RSN/lib/utils/transforms.py:70-79
 *     (mul_vis = 0).  joints, joints_vis [N,K,3] f64; width [N] int32 (image width per person);
 *     outputs must not alias the inputs. */
int udp_fliplr_joints(const double *joints, const double *joints_vis, const int32_t *flip_perm,
                      const int32_t *width, int K, int N, int mul_vis, double *joints_out,
                      double *vis_out, void *stream);

/* a16 transform_preds(coords, center, scale, [w,h])  deep_hrnet/lib/core/inference.py:20-27
 *     (= utils/transforms.py:67-74, RSN prm test.py:29-39).  coords [N,K,2] float64 or float32
 *     (coords_f64); out [N,K,2] float64 (the reference returns np.zeros(coords.shape)). */
int udp_transform_preds(const void *coords, int coords_f64, const void *center, const void *scale,
                        int cs_f64, int w, int h, double *out, int K, int N, void *stream);

/* a5  JointsDataset.generate_target, TARGET_TYPE 'gaussian'  JointsDataset.py:291-348,382-383.
 *     joints, joints_vis [N,K,3] float64; joints_weight [K] float32 or NULL
 *     (LOSS.USE_DIFFERENT_JOINTS_WEIGHT).  target [N,K,h,w] f32, target_weight [N,K,1] f32. */
int udp_encode_gaussian(const double *joints, const double *joints_vis, const float *joints_weight,
                        int K, int W, int H, int w, int h, double sigma, float *target,
                        float *target_weight, int N, void *stream);

/* a6  generate_target, TARGET_TYPE 'offset' (UDPv1)  JointsDataset.py:349-381.
 *     target [N,3K,h,w] f32, channel order (j,cls),(j,xoff),(j,yoff). */
int udp_encode_offset(const double *joints, const double *joints_vis, const float *joints_weight,
                      int K, int W, int H, int w, int h, double kpd, float *target,
                      float *target_weight, int N, void *stream);

/* a7  RSN generate_heatmap for S kernel sizes  RSN/dataset/JointsDataset.py:276-291, :214-220.
 *     joints [N,K,jstride] f64 (crop coords), valid [N,K] f32 (drawn iff >= 1),
 *     sigmas: HOST pointer to S <= UDP_MAX_SIGMAS doubles.  labels [N,S,K,h,w] f32. */
int udp_encode_rsn(const double *joints, int jstride, const float *valid, int K, int W, int H,
                   int w, int h, const double *sigmas, int S, float *labels, int N, void *stream);

/* a9/a10/a11  flip_back / flip_back_offset (+ average)
 *     deep_hrnet/lib/utils/transforms.py:15-47 (RSN/lib/utils/transforms.py:59-67),
 *     lib/core/function.py:171.
 *     flipped [N,C,h,w]; flip_perm [K] int32 (partner joint, identity where unpaired);
 *     triplets = 1: C = 3K and the x-offset channel is negated.  out == NULL: dst = flip_back(flipped);
 *     else dst = (out + flip_back(flipped)) * 0.5.  dst must not alias flipped. */
int udp_flip_back(const float *out, const float *flipped, const int32_t *flip_perm, int K, int w,
                  int h, int triplets, float *dst, int N, void *stream);

/* a12/a13/a14/a16/a17  get_final_preds, TARGET_TYPE 'gaussian', with the flip test fused in
 *     deep_hrnet/lib/core/inference.py:30-58 (get_max_preds), :60-145 (post), :149-155,176-186,
 *     :20-27 (transform_preds); flip: utils/transforms.py:15-29 + core/function.py:171;
 *     RSN prm: RSN/exps/RSN18.coco.e1.se.36x8x132000_prm/test.py:183-196 (pre_div = 255).
 *     hm [N,K,h,w] f32; hm_flipped same shape or NULL; flip_perm [K] (ignored when NULL flipped).
 *     center/scale [N,2] or NULL (NULL: coordinates stay in heatmap space = get_max_preds / post).
 *     pre_div: 0 = none, else maps are divided by it first.  post_process = TEST.POST_PROCESS.
 *     preds, preds_in [N,K,2]: float64 when post_process else float32 (the reference's dtypes);
 *     maxvals [N,K,1] f32; argmax_idx [N,K] int32 (nullable); packed [N,K,3] f32 x,y,maxval
 *     (nullable; the all_preds layout of function.py:208-209, used for the multi-GPU gather);
 *     blurred [N,K,h,w] f32 (nullable; the maps post() writes back into its input, :80). */
int udp_decode_gaussian(const float *hm, const float *hm_flipped, const int32_t *flip_perm,
                        const void *center, const void *scale, int cs_f64, float pre_div,
                        int post_process, int K, int w, int h, void *preds, float *maxvals,
                        void *preds_in, int32_t *argmax_idx, float *packed, float *blurred, int N,
                        void *stream);

/* a10/a11/a15/a16  get_final_preds, TARGET_TYPE 'offset' (UDPv1 voting), flip test fused in
 *     inference.py:156-174,176-186; flip: utils/transforms.py:31-47 + function.py:171.
 *     out [N,3K,h,w] f32; out_flipped or NULL.  preds, preds_in [N,K,2] float32;
 *     maxvals = maxima of the 15x15-blurred classification maps; argmax_idx their argmax. */
int udp_decode_offset(const float *out, const float *out_flipped, const int32_t *flip_perm,
                      double kpd, const void *center, const void *scale, int cs_f64, int K, int w,
                      int h, float *preds, float *maxvals, float *preds_in, int32_t *argmax_idx,
                      float *packed, int N, void *stream);

/* a10/a11/a15/a16 with split sources -- the host round trip of lib/core/function.py:151-171,205 made cheaper.
 *     The voting decoder reads the K classification maps in full but only a 7x7 window of each offset map
 *     (inference.py:166-167,173-174).  cls_out / cls_flipped: [N,3K,h,w] DEVICE tensors of which only the
 *     classification channels (0::3) need to hold data; off_out / off_flipped: the same tensors' offset channels
 *     somewhere else with the same [N,3K,h,w] layout -- typically the caller's pinned HOST buffers as seen from the
 *     device (udp_host_device_pointer), so that two thirds of the network output never cross the host link.
 *     Everything else as udp_decode_offset; results are bit-identical. */
int udp_decode_offset_split(const float *cls_out, const float *cls_flipped, const float *off_out,
                            const float *off_flipped, const int32_t *flip_perm, double kpd, const void *center,
                            const void *scale, int cs_f64, int K, int w, int h, float *preds, float *maxvals,
                            float *preds_in, int32_t *argmax_idx, float *packed, int N, void *stream);

/* Host plumbing for the split call, no kernels (the .cpu().numpy() / .cuda() pairs of function.py:154-169):
 *     udp_copy2d_async: cudaMemcpy2DAsync of `height` rows of `width` bytes, kind 1 = host to device, 2 = device
 *     to host (one call moves the K classification maps of a chunk: width = h*w*4, pitches = 3*h*w*4);
 *     udp_host_device_pointer: the device-side address of a pinned, mapped HOST buffer (cudaHostGetDevicePointer). */
int udp_copy2d_async(void *dst, size_t dpitch, const void *src, size_t spitch, size_t width, size_t height,
                     int kind, void *stream);
int udp_host_device_pointer(void *host_ptr, void **device_ptr);

/* SURVEY 8(f)-2  accuracy(output, target, hm_type='gaussian', thr)  deep_hrnet/lib/core/evaluate.py:16-71,
 *     call sites lib/core/function.py:76-80 and :187-192 (offset mode passes the [:, ::3] views:
 *     ch_stride = 3).  output, target [N, K*ch_stride, h, w] float32 (16-byte aligned, h*w % 4 == 0).
 *     pred [N,K,2] float32 = get_max_preds(output) coordinates (the 4th return value);
 *     dists [K,N] float64 scratch/out (calc_dists, -1 = ignored); acc [K+1] float64 (acc[0] = average,
 *     acc[1+j] = joint j or -1), avg_acc [1] float64, cnt [1] int32 -- all DEVICE pointers.  K <= 32. */
int udp_accuracy_pck(const float *output, const float *target, int ch_stride, int K, int w, int h,
                     int N, double thr, float *pred, double *dists, double *acc, double *avg_acc,
                     int32_t *cnt, void *stream);

/* SURVEY 8(f)-2 (tail)  the all_boxes rows validate() keeps next to all_preds, lib/core/function.py:211-214:
 *     all_boxes[n] = (center x, y, scale x, y, np.prod(scale * 200), score) as float64 [N,6]; `scale * 200` and the
 *     product in the records' dtype (cs_f64 = 0: float32, as numpy does for COCO records).  score [N] float64 or NULL (1).
 *     With udp_decode_*'s `packed` (= all_preds) this is everything COCODataset.evaluate consumes. */
int udp_pack_all_boxes(const void *center, const void *scale, int cs_f64, const double *score, double *all_boxes,
                       int N, void *stream);

/* udp_accuracy_pck for callers that cannot allocate -- NO COUNTERPART IN THE REFERENCE; added
 *     (code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).
 *     Same arguments and results as udp_accuracy_pck minus `dists`: that [K,N] float64 workspace is a static
 *     device array inside the library of UDP_PCK_NOALLOC_MAX_DISTS doubles (4 MiB, held by every GPU that loads
 *     the library until exit).  N * K > UDP_PCK_NOALLOC_MAX_DISTS returns UDP_ESCRATCH (-2) and launches nothing.
 *     Calls no cudaMalloc / cudaFree / allocating routine; its only set-up is one cudaGetSymbolAddress per device.
 *     Every call shares the workspace (it keeps the previous call's contents; each call overwrites all the K*N
 *     entries it reads, nothing needs clearing): calls that overlap on the GPU -- different streams with no
 *     ordering between them -- race on it; calls on one stream, or ordered by events, are safe. */
#define UDP_PCK_NOALLOC_MAX_DISTS (512 * 1024)
int udp_accuracy_pck_noalloc(const float *output, const float *target, int ch_stride, int K, int w, int h,
                             int N, double thr, float *pred, double *acc, double *avg_acc, int32_t *cnt,
                             void *stream);

/* SURVEY 8(f)-4 (first half)  the demo engines' quarter-offset decoder:
 *     get_final_preds(heatmaps, boxes) of deep_hrnet/tools/infer_utils/decode.py:19-40, with its
 *     get_max_preds (:43-52, coordinates -1 where maxval <= 0) and the biased transform_preds
 *     (:55-62, scale*200/output_size).  heatmaps [N,K,h,w] float32; boxes [N,4] = (cx, cy, sx, sy),
 *     float32 (boxes_f64 = 0) or float64 (1): the dtype selects the affine's arithmetic, as numpy
 *     promotion does in the reference.  coords [N,K,2] int32 = the reference's `.astype(int)`
 *     (INT32_MIN where numpy yields INT64_MIN: NaN or beyond int32); coords_f32 [N,K,2] (the values
 *     before truncation) and maxvals [N,K] are optional (NULL). */
int udp_decode_quarter(const float *heatmaps, const void *boxes, int boxes_f64, int K, int w, int h, int N,
                       int32_t *coords, float *coords_f32, float *maxvals, void *stream);

/* SURVEY 8(f)-4 (second half)  RSN's legacy decoder: get_results(outputs, centers, scales, kernel, shifts)
 *     of RSN/exps/RSN18.coco/test.py:25-71 (also Res18.coco, 4XRSN18.coco, RSN18.mpii), with the flip
 *     average of compute_on_dataset (:88-96) fused when outputs_flipped is given (flip_perm [K] int32).
 *     outputs [N,K,h,w] float32 (network units, 255 * heat); centers/scales [N,2] float32 (scales NOT yet
 *     multiplied by 200 -- the reference's in-place `scales *= 200` happens inside); (in_w, in_h) =
 *     cfg.INPUT_SHAPE[1], [0]; kernel = cfg.TEST.GAUSSIAN_KERNEL (odd, 3..15) and taps_host = HOST pointer
 *     to its `kernel` float64 taps, cv2.getGaussianKernel(kernel, 0, CV_64F); shift = cfg.TEST.SHIFT_RATIOS[0].
 *     Out: preds [N,K,2] float64, maxvals [N,K,1] float64.  Maps containing NaN are not supported. */
int udp_rsn_get_results(const float *outputs, const float *outputs_flipped, const int32_t *flip_perm,
                        const float *centers, const float *scales, int K, int w, int h, int in_w, int in_h,
                        int kernel, const double *taps_host, double shift, double *preds, double *maxvals,
                        int N, void *stream);

/* SURVEY 8(f)-5  AID augmentations on the crop: Cutout (deep_hrnet/lib/utils/transforms.py:186-223) and
 *     HideAndSeek (:144-183), applied in place to the uint8 crops [N,H,W,3] (JointsDataset.py:230-235).
 *     The random draws stay on the host (numpy's global generator, in the reference's order); this entry
 *     point applies the drawn parameters:
 *     ellipses [N,max_ellipses,4] float64 (cx, cy, rx, ry; :205-206), n_ellipses [N] int32 (0 = Cutout not
 *     applied to this person), max_ellipses <= 8; either may be NULL (no Cutout).
 *     grid_size [N] int32 (0 = HideAndSeek not applied / grid 0) and hide_mask [N,UDP_AID_MASK_WORDS] uint32: bit c of
 *     the 512-bit mask = cell c hidden, cells numbered in the reference's draw order (outer range(0, width, g),
 *     inner range(0, height, g)); either may be NULL.  min_grid = smallest non-zero grid size in use
 *     (cells must fit the mask).  The reference's transposed slice img[x:x_end, y:y_end] (:175) is kept. */
#define UDP_AID_MASK_WORDS 16
int udp_aid_erase(uint8_t *img, int N, int H, int W, const double *ellipses, const int32_t *n_ellipses,
                  int max_ellipses, const int32_t *grid_size, const uint32_t *hide_mask, int min_grid,
                  void *stream);

/* SURVEY 8(f)-3  COCO rescoring: deep_hrnet/lib/dataset/coco.py:326-339.
 *     preds [N,K,3] float32 (x, y, maxval -- the rows of validate()'s all_preds, function.py:208-209),
 *     box_score [N] float64 (all_boxes[:,5]).  score[n] = mean(maxval > in_vis_thre) * box_score[n]
 *     (float32 mean in joint order, float64 product; 0 when no joint passes). */
int udp_coco_rescore(const float *preds, const double *box_score, int K, int N, double in_vis_thre,
                     double *score, void *stream);

/* SURVEY 8(f)-3  oks_nms / soft_oks_nms of deep_hrnet/lib/nms/nms.py:97-124 / :139-180 (with oks_iou
 *     :75-94 and rescore :127-136), one CTA per image.  Persons of image i are rows
 *     img_off[i] .. img_off[i+1]-1 of preds/area/score (img_off [num_images+1] int32; the caller groups
 *     them, as coco.py:317-319 does); max_per_image >= the largest group (sizes shared memory).
 *     area [N] float64 (all_boxes[:,4]), score [N] float64 (udp_coco_rescore's output).
 *     sigmas_host: HOST pointer to K float64 (already divided by 10, as the reference's `sigmas`
 *     argument is) or NULL = COCO's 17 defaults (nms.py:77).  use_vis_thre/in_vis_thre: oks_iou's
 *     optional in_vis_thre (None in COCO evaluation => use_vis_thre = 0).  soft = 0: keep while
 *     oks <= thresh; soft = 1: gaussian score decay, at most max_dets (20 at nms.py:158) per image.
 *     Out: keep [N] int32 -- for image i, keep_cnt[i] global rows in the reference's keep order,
 *     starting at keep[img_off[i]]; keep_cnt [num_images] int32 (-1 if a group exceeds
 *     max_per_image); oks_top [N] float64 or NULL -- oks of every person against its image's
 *     first pick (the pick's own entry is not written). */
int udp_oks_nms(const float *preds, const double *area, const double *score, const int32_t *img_off,
                int num_images, int max_per_image, int K, double thresh, const double *sigmas_host,
                int use_vis_thre, double in_vis_thre, int soft, int max_dets, int32_t *keep,
                int32_t *keep_cnt, double *oks_top, void *stream);

#ifdef __cplusplus
}
#endif
#endif /* UDP_CODEC_H */
