/*
 * curl_b200 -- C ABI of the B200 (sm_100a) CURL curve-layer library (libcurl_b200.so).
 *
 * The reference (danielbulhosa/CURL) is pure Python and has no FFI of its own: its boundary for this
 * path is the nn.Module / function API of colors.py, curves.py and model.py::CURLLayer.  This
 * header is the C ABI inserted *underneath* that API ("Level 2"); each entry point
 * names the reference interface it replaces.  curl_b200/ (Python) binds it with ctypes and mirrors
 * the reference module API on top; INTEGRATION.md shows the binding a reference maintainer would add.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer owned by the caller, except where noted; nothing is
 *     allocated, freed or synchronised inside the library; work is enqueued on `stream`
 *     (a cudaStream_t passed as void*; NULL = legacy default stream) of the CURRENT device.
 *   - images are planar NCHW fp32, contiguous: img[B][3][H][W].  mask is [B][1][H][W].
 *   - return 0 on success, a negative CURL_ERR_* otherwise; nothing is thrown across the ABI.
 *   - re-entrant across host threads as long as the streams (and workspaces) differ.
 *   - all calls are CUDA-graph capturable (no host synchronisation, no host-side reads).
 */
#ifndef CURL_B200_H_
#define CURL_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CURL_ABI_VERSION 10

/* error codes */
#define CURL_OK 0
#define CURL_ERR_BAD_ARG (-1)        /* null pointer, non-positive size, K < 3, unknown enum */
#define CURL_ERR_TOO_MANY_KNOTS (-2) /* total knots per image exceed CURL_MAX_KNOTS */
#define CURL_ERR_WORKSPACE (-3)      /* workspace missing or too small */
#define CURL_ERR_LAUNCH (-4)         /* cudaGetLastError() after the launch was not cudaSuccess */
#define CURL_ERR_MISALIGNED (-5)     /* a base pointer is not 4-byte aligned */
#define CURL_ERR_TOO_LARGE (-6)      /* curl_layer_bwd_noalloc_f32: shape needs more than CURL_NOALLOC_SCRATCH_BYTES */

#define CURL_MAX_KNOTS 1024

/* mask_kind: how `mask` is to be read (reference: data.py:190 gives bool, infer.py:41 gives float) */
#define CURL_MASK_NONE 0 /* mask == NULL, treated as all ones                              */
#define CURL_MASK_U8 1   /* 1 byte per pixel, zero / non-zero (torch.bool; the Python layer sends a  */
                         /* torch.uint8 mask as F32 because the reference multiplies by its VALUE)  */
#define CURL_MASK_F32 2  /* fp32 per pixel, arbitrary values, multiplied in at every stage   */

/* colour conversions (reference colors.py) */
#define CURL_RGB2LAB 0 /* colors.py:27-62   RGB2LAB.forward */
#define CURL_LAB2RGB 1 /* colors.py:88-123  LAB2RGB.forward */
#define CURL_RGB2HSV 2 /* colors.py:195-242 RGB2HSV.forward */
#define CURL_HSV2RGB 3 /* colors.py:131-177 HSV2RGB.forward */

/* curve groups (reference curves.py) */
#define CURL_ADJUST_LAB 0 /* curves.py:136-180 adjust_lab: 3 curves c->c                 */
#define CURL_ADJUST_RGB 1 /* curves.py:90-133  adjust_rgb: 3 curves c->c                 */
#define CURL_ADJUST_HSV 2 /* curves.py:41-87   adjust_hsv: 4 curves 0->0, 0->1, 1->1, 2->2 */

int curl_abi_version(void);
const char* curl_strerror(int code);

/*
 * TEST HOOK, not part of the reference-facing surface (the reference has nothing like it): caps the
 * number of blocks per image every layer / colour / curve / TriSpace kernel is launched with (0 = no cap, the
 * default) and returns the previous cap.  With a small cap a SMALL image drives each block through
 * many tiles -- shared-memory ring reuse, mbarrier parity flips, grid-stride steps -- so the parity
 * tests can compare exactly the loop structure the 4K benchmark shapes run against the CPU oracle.
 * Process-global and not thread-safe: set it, run the test, restore it.  Results do not depend on
 * the cap except for the summation order of the knot-gradient partials.
 */
int curl_test_set_grid_cap(int max_blocks_per_image);

/*
 * Fused forward of the whole layer.  Replaces CURLLayer.forward (model.py:137-176):
 *   rgb2lab -> adjust_lab -> *mask -> lab2rgb -> adjust_rgb -> *mask -> rgb2hsv -> adjust_hsv ->
 *   *mask -> hsv2rgb -> clamp(img + residual, 0, 1) * mask,   reg = reg_rgb + reg_lab + reg_hsv.
 * L, R, Hk: log-domain knots, row b at L + b*ldL (ld in elements, so column-sliced views such as
 * curves[:, :48] (model.py:197-199) need no copy); 3*K_lab, 3*K_rgb, 4*K_hsv leading columns read.
 * out[B][3][H][W], reg[B].  One kernel launch.
 */
int curl_layer_fwd_f32(const float* img, const void* mask, int mask_kind,
                       const float* L, int64_t ldL, const float* R, int64_t ldR,
                       const float* Hk, int64_t ldH,
                       float* out, float* reg,
                       int B, int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream);

/*
 * Fused backward of the layer.  Replaces torch.autograd over model.py:137-176 (SURVEY.md 8a row
 * a10).  Recomputes the forward per pixel from img (no saved activations), writes
 *   gimg[B][3][H][W]   (may be NULL: skipped, e.g. when img does not require grad),
 *   gL[B][3*K_lab], gR[B][3*K_rgb], gH[B][4*K_hsv]   dense, gradient w.r.t. the LOG-domain knots,
 * given gout[B][3][H][W] and greg[B] (may be NULL = zeros).  Knot gradients are reduced in the
 * kernel: registers -> warp shuffles -> shared memory -> one partial row per block in
 * `workspace`, one atomic ticket per block, and the last block of each image folds the partials in
 * a fixed order (bitwise deterministic).  One memset node + one kernel launch.
 */
int curl_layer_bwd_f32(const float* img, const void* mask, int mask_kind,
                       const float* L, int64_t ldL, const float* R, int64_t ldR,
                       const float* Hk, int64_t ldH,
                       const float* gout, const float* greg,
                       float* gimg, float* gL, float* gR, float* gH,
                       void* workspace, size_t workspace_bytes,
                       int B, int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream);

/*
 * Inference on 8-bit images: u8 in -> fp32 arithmetic -> u8 out, 7 bytes per pixel over HBM/PCIe
 * instead of 25.  Replaces the data-format steps the reference wraps around the layer:
 *   in : torchvision TF.to_tensor on a uint8 image (data.py:152-158, infer.py:38-41): x = u8 / 255
 *        in fp32, correctly rounded;
 *   out: (out * 255) truncated to uint8 (evaluate.py:63 `.astype('uint8')`, TF.to_pil_image's
 *        `.mul(255).byte()`, infer.py:47).
 * in_layout / out_layout: CURL_LAYOUT_CHW = planar [B][3][H][W] (what to_tensor produces),
 * CURL_LAYOUT_HWC = interleaved [B][H][W][3] (what PIL / plt.imsave hold, transpose.py).
 * Everything between the two conversions is curl_layer_fwd_f32, bit for bit.  Forward only.
 */
#define CURL_LAYOUT_CHW 0
#define CURL_LAYOUT_HWC 1
int curl_layer_fwd_u8(const uint8_t* img, int in_layout, const void* mask, int mask_kind,
                      const float* L, int64_t ldL, const float* R, int64_t ldR,
                      const float* Hk, int64_t ldH,
                      uint8_t* out, int out_layout, float* reg,
                      int B, int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream);

/* Bytes of device workspace the two *_bwd_f32 entry points need for this shape (host-only call). */
size_t curl_bwd_workspace_bytes(int B, int H, int W);

/*
 * Allocation-free variant of curl_layer_bwd_f32.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot allocate device memory (code under CUDA-graph capture, tight serving
 * loops, harnesses that forbid cudaMalloc).  Same arguments, semantics and results as curl_layer_bwd_f32 minus
 * the workspace: the scratch (tickets + per-block moment partials) is a file-scope device array of
 * CURL_NOALLOC_SCRATCH_BYTES that every GPU loading the library holds until the process exits.
 *   - shapes with curl_bwd_workspace_bytes(B, H, W) > CURL_NOALLOC_SCRATCH_BYTES return CURL_ERR_TOO_LARGE
 *     (nothing is launched); every BASELINE shape fits (4K x 64: 0.57 MB);
 *   - calls no cudaMalloc / cudaFree / allocating routine; the first call on a device looks the array's address
 *     up once (cudaGetSymbolAddress);
 *   - ALL calls share the one array: calls that overlap on the GPU (different streams) race on it.  Serialise
 *     them, or use curl_layer_bwd_f32 with a workspace per stream.  What must start at zero is cleared per call.
 */
#define CURL_NOALLOC_SCRATCH_BYTES (1 << 20)
int curl_layer_bwd_noalloc_f32(const float* img, const void* mask, int mask_kind,
                               const float* L, int64_t ldL, const float* R, int64_t ldR,
                               const float* Hk, int64_t ldH,
                               const float* gout, const float* greg,
                               float* gimg, float* gL, float* gR, float* gH,
                               int B, int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream);

/* Stand-alone colour conversion, forward.  Replaces colors.<Module>.forward(img). */
int curl_color_fwd_f32(int conversion, const float* in, float* out, int B, int H, int W, void* stream);
/* ... and its backward: gin = J(in)^T gout (recomputes the forward from `in`). */
int curl_color_bwd_f32(int conversion, const float* in, const float* gout, float* gin,
                       int B, int H, int W, void* stream);

/* Stand-alone curve group, forward.  Replaces curves.adjust_lab / adjust_rgb / adjust_hsv(img, P):
 * P[B][>= n_curves*K] log-domain knots with row stride ldP; out[B][3][H][W]; reg[B]. */
int curl_adjust_fwd_f32(int group, const float* in, const float* P, int64_t ldP,
                        float* out, float* reg, int B, int H, int W, int K, void* stream);
/* ... and its backward: gin[B][3][H][W] (nullable), gP[B][n_curves*K] dense. */
int curl_adjust_bwd_f32(int group, const float* in, const float* P, int64_t ldP,
                        const float* gout, const float* greg, float* gin, float* gP,
                        void* workspace, size_t workspace_bytes,
                        int B, int H, int W, int K, void* stream);

/* One curve, knots given directly as C[B][>=K] (NOT log-domain), scale from channel_in applied to
 * channel_out, then the whole-image clamp.  Replaces curves.apply_curve(img, C, slope_sqr_diff,
 * channel_in, channel_out) (curves.py:4-38); reg[B] receives this curve's regulariser term (the
 * caller adds it to its running slope_sqr_diff). */
int curl_apply_curve_fwd_f32(const float* in, const float* C, int64_t ldC, int channel_in, int channel_out,
                             float* out, float* reg, int B, int H, int W, int K, void* stream);
/* ... and its backward: gin[B][3][H][W] (nullable), gC[B][K] dense (gradient w.r.t. C itself). */
int curl_apply_curve_bwd_f32(const float* in, const float* C, int64_t ldC, int channel_in, int channel_out,
                             const float* gout, const float* greg, float* gin, float* gC,
                             void* workspace, size_t workspace_bytes,
                             int B, int H, int W, int K, void* stream);

/*
 * Polynomial colour regression ("TriSpace" residual): the per-pixel path of TriSpaceRegNet, which
 * is what the reference's main.py / infer.py run (model.py:411-520).
 *
 * curl_trispace_fwd_f32 replaces TriSpaceRegNet.generate_residual (model.py:486-503) and, when
 * final_clamp != 0, generate_image (model.py:505-508) as well:
 *   residual = 2(sigmoid(P_R(rgb)) - .5) + 2(lab2rgb(sigmoid(P_L(rgb2lab(rgb)))) - .5)
 *            + 2(hsv2rgb(sigmoid(P_H(rgb2hsv(rgb)))) - .5);     out = clamp(img + residual, 0, 1) | residual
 * P_* = ChannelPolyLayer(degree, num_variables = 3 + 2*spatial, num_out = 3) (model.py:206-333; the
 * Deg4MobilePolyLayer of model.py:336-408 is the same function): all monomials up to `degree` in the
 * three colour channels and, if spatial, x = column/W and y = row/H (model.py:472-484), coefficient
 * order as ChannelPolyLayer.generate_powers enumerates them.  R, L, Hc: [B][3][NC] contiguous,
 * NC = curl_trispace_num_coeffs(degree, spatial) (35 / 126 for degree 4).  One pass per pixel; the
 * reference's [B, H, W, NC] monomial tensor (504 B/px per space) is never formed.
 * Degrees 1..4 are built (4 is the reference's polynomial_order default and the only one its scripts use;
 * NC = 4/10/20/35 without and 6/21/56/126 with the spatial variables); other degrees return CURL_ERR_BAD_ARG.
 * saved: NULL, or [B][9][H][W] receiving the nine sigmoid outputs (3 spaces x 3 channels) -- the only
 * intermediate the backward needs besides the image.  Handing the same buffer to
 * curl_trispace_bwd_f32 skips re-evaluating the three polynomials there (36 B/px of HBM for about
 * 1.4x on the backward); with NULL the backward recomputes everything from the image.
 */
int curl_trispace_num_coeffs(int degree, int spatial);
int curl_trispace_fwd_f32(const float* img, const float* R, const float* L, const float* Hc,
                          int degree, int spatial, int final_clamp, float* out, float* saved,
                          int B, int H, int W, void* stream);
/*
 * Backward of the above w.r.t. the coefficients: gR, gL, gH [B][3][NC] dense, given gout[B][3][H][W].
 * Replaces torch.autograd over model.py:486-508.  The image is data on this path
 * (TriSpaceRegNet.forward feeds the input or target image, model.py:517-519), so there is no image
 * gradient.  Recomputes the forward per pixel; per image the reduction is a [3 x N].[N x NC]
 * product per space, done tile by tile in shared memory, then one partial row per block, one atomic
 * ticket per block and a fixed-order fold in double by the image's last block (deterministic).
 */
size_t curl_trispace_bwd_workspace_bytes(int B, int H, int W, int degree, int spatial);
int curl_trispace_bwd_f32(const float* img, const float* R, const float* L, const float* Hc,
                          int degree, int spatial, int final_clamp, const float* gout, const float* saved,
                          float* gR, float* gL, float* gH, void* workspace, size_t workspace_bytes,
                          int B, int H, int W, void* stream);
/*
 * Backward of the same w.r.t. the IMAGE: gimg[B][3][H][W] dense, given gout.  Replaces torch.autograd over
 * model.py:486-508 when the image handed to TriSpaceRegNet requires a gradient (the reference's own scripts feed
 * data there, model.py:517-519, and never ask; the coefficient gradients above do not depend on this call).
 * Element-wise, no workspace: the direct term of generate_image's clamp, plus per space the gradient of the
 * polynomial w.r.t. its three colour variables (x = column/W, y = row/H are constants, model.py:444-452) chained
 * through the Jacobians of RGB2LAB / RGB2HSV.  saved: as for curl_trispace_bwd_f32 (NULL: recompute).
 */
int curl_trispace_bwd_img_f32(const float* img, const float* R, const float* L, const float* Hc,
                              int degree, int spatial, int final_clamp, const float* gout, const float* saved,
                              float* gimg, int B, int H, int W, void* stream);

/*
 * Colour terms of CURLLoss (model.py:36-118): everything in that loss except MS-SSIM.
 *
 * curl_loss_terms_fwd_f32 replaces, in ONE pass over (pred, target, mask), model.py:89 (masking),
 * :91 (L1 in RGB), :95 (cosine_similarity over channels), :57-58 + :98-100 (clamped RGB2LAB of both
 * images and their L1), :66-79 + :106-108 (HSV cone of both images and their L1) and :102-103 (the
 * two Lab L planes MS-SSIM consumes).  Per image b it writes sums[b][0..4]:
 *   [0] sum |p - t|   [1] sum cos_sim(p, t)   [2] sum |lab(p) - lab(t)|   [3] sum |cone(p) - cone(t)|
 *   [4] sum mask                                     (p = pred * mask, t = target * mask)
 * from which the caller forms model.py:90-117:
 *   n = 3 * sum_b s4;  rgb = sum_b s0 / n;  lab = sum_b s2 / n;  hsv = sum_b s3 / n;
 *   cosine = 1 - (sum_b s1 + (B*H*W - sum_b s4)) / (B*H*W)      (mask is 0/1; model.py:96 broadcasts
 *   [B,H,W] + [B,1,H,W] to [B,B,H,W], whose mean is mean(cos) + mean(!mask));
 *   loss = (rgb + cosine + lab + hsv + 10 * (1 - msssim(L_pred, L_target)).mean()) / 5.
 * L_pred / L_target: [B][H][W] or NULL.  Deterministic: one partial row per block, an atomic ticket
 * per block, fixed-order fold in double by each image's last block.
 */
size_t curl_loss_terms_workspace_bytes(int B, int H, int W);
int curl_loss_terms_fwd_f32(const float* pred, const float* target, const void* mask, int mask_kind,
                            float* sums /*[B][5]*/, float* L_pred, float* L_target,
                            void* workspace, size_t workspace_bytes, int B, int H, int W, void* stream);
/*
 * Backward w.r.t. pred (the target is data).  w: DEVICE pointer to [B][4] = d loss / d sums[b][0..3]
 * (they depend on sum mask, which lives on the device; no host round trip); gL_pred: [B][H][W]
 * gradient arriving at L_pred from MS-SSIM, or NULL.  Recomputes the conversions per pixel;
 * replaces torch.autograd over model.py:89-108.
 */
int curl_loss_terms_bwd_f32(const float* pred, const float* target, const void* mask, int mask_kind,
                            const float* w, const float* gL_pred, float* gpred,
                            int B, int H, int W, void* stream);

/*
 * Opt-in: the layer with SEGMENT-CLAMPED curves -- eq. (1) of the CURL paper (arXiv 1911.13175),
 *   scale(x) = k_0 + sum_{i=0}^{K-2} (k_{i+1} - k_i) * clamp((K-1) x - i, 0, 1),
 * the piecewise-linear curve through all K knots.  This is NOT what the reference evaluates
 * (curves.py:29-32 sums the unclamped terms and leaves the last slope out, which is affine in x);
 * everything else -- curve wiring, whole-pixel clamps, regulariser, layer chain, arguments -- is as
 * in curl_layer_fwd_f32 / curl_layer_bwd_f32 above.  Scalar kernels (the segment is a per-pixel
 * lookup); knot gradients are a per-segment reduction through per-warp shared-memory histograms
 * (no atomics: same-bin lanes add in lane order, bit-reproducible), then partial rows + ticket + a
 * fixed-order fold in double.  No reference code exists for this variant: checked against the
 * oracle's restatement of the published formula only (parity unpinned).
 */
int curl_layer_seg_fwd_f32(const float* img, const void* mask, int mask_kind,
                           const float* L, int64_t ldL, const float* R, int64_t ldR,
                           const float* Hk, int64_t ldH,
                           float* out, float* reg,
                           int B, int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream);
size_t curl_layer_seg_bwd_workspace_bytes(int B, int H, int W, int K_lab, int K_rgb, int K_hsv);
int curl_layer_seg_bwd_f32(const float* img, const void* mask, int mask_kind,
                           const float* L, int64_t ldL, const float* R, int64_t ldR,
                           const float* Hk, int64_t ldH,
                           const float* gout, const float* greg,
                           float* gimg, float* gL, float* gR, float* gH,
                           void* workspace, size_t workspace_bytes,
                           int B, int H, int W, int K_lab, int K_rgb, int K_hsv, void* stream);

/*
 * Device-side TF.to_tensor for 8-bit images (data.py:152-158, infer.py:38-41): uint8 [B][3][H][W]
 * (CURL_LAYOUT_CHW) or [B][H][W][3] (CURL_LAYOUT_HWC) -> planar fp32 out[B][3][H][W] = u8 / 255,
 * correctly rounded for all 256 inputs (bit-identical to torchvision's division).  For TRAINING from
 * 8-bit data: the host uploads 3 B/px instead of 12 and the fp32 tensor the layer differentiates
 * through is produced on the device (inference does not need it: curl_layer_fwd_u8 converts in
 * registers).
 */
int curl_u8_to_unit_f32(const uint8_t* img, int in_layout, float* out, int B, int H, int W, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* CURL_B200_H_ */
