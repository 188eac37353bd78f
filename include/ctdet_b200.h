/*
 * ctdet_b200.h -- C ABI of libctdet_b200.so: the CenterNet "ctdet" heatmap hot path as
 * hand-written sm_100a CUDA kernels (no torch types, no CPU fallback).
 *
 * All pointers are DEVICE pointers on the current CUDA device unless stated otherwise; the
 * caller owns every buffer (outputs and workspaces included), the library never allocates or
 * frees device memory and keeps no state between calls.  Every entry point enqueues its work on
 * `stream` (a cudaStream_t passed as void*) and returns without synchronising.
 * Return value: 0 on success, otherwise a CTDET_E* code (see ctdet_error_string).
 *
 * Reference interfaces replaced (paths relative to the ShawnNew/Detectron2-CenterNet root):
 *   ctdet_target_gen     detectron2/data/detection_utils.py:600  gen_heatmap (+ :654 gaussian_radius,
 *                        :682 gaussian2D, :690 draw_umich_gaussian) and the torch.cat/.to(device) of
 *                        detectron2/modeling/meta_arch/centernet.py:197-201
 *   ctdet_loss_fwd_bwd   centernet.py:191-212 CenterNet.losses = :204 sigmoid_+clamp, :323-369
 *                        FocalLoss/_neg_loss, :372-381 RegL1Loss (+ :383 _transpose_and_gather_feat)
 *                        and their autograd backward
 *   ctdet_target_loss_fused  the two rows above in one pass (gen_heatmap at centernet.py:188 + losses :191-212)
 *   ctdet_head_focal_fused   centernet.py:118-123 / :152 (the hm head's final 1x1 Conv2d) + :204 + :323-369 (focal loss)
 *   ctdet_decode         centernet.py:426 ctdet_decode = :399 _nms, :408 _topk, :383 gather, boxes
 *   ctdet_nms / ctdet_gather_feat   centernet.py:399 _nms / :383 _transpose_and_gather_feat
 *   ctdet_postprocess    centernet.py:236-266 inference_single_image (slice + score threshold) and
 *                        detectron2/modeling/postprocessing.py:10-74 detector_postprocess (scale, clip, nonempty)
 *   ctdet_preprocess     centernet.py:173-184 CenterNet.preprocess_image (x/255, normalise) and
 *                        detectron2/structures/image_list.py:58-130 ImageList.from_tensors (pad + batch)
 *   ctdet_upsample_add   detectron2/modeling/backbone/dla.py:166-168 `up_i` (depth-wise ConvTranspose2d) applied at
 *                        :175 and the `+ layers[i - 1]` of :176-177 (IDAUp.forward)
 *   dcnv2_forward        detectron2/layers/csrc/deformable/deform_conv.h:262 modulated_deform_conv_forward
 *                        -> deform_conv_cuda.cu:804 modulated_deform_conv_cuda_forward
 *                        (im2col kernel deform_conv_cuda_kernel.cu:786 + addmm_ :904 + bias :924)
 *   dcnv2_backward       deform_conv.h:314 modulated_deform_conv_backward
 *                        -> deform_conv_cuda.cu:929 modulated_deform_conv_cuda_backward
 *                        (col2im_coord deform_conv_cuda_kernel.cu:952, col2im :871, im2col :786, 3 GEMMs)
 */
#ifndef CTDET_B200_H_
#define CTDET_B200_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define CTDET_ABI_VERSION 1
#define CTDET_MAX_OBJS 128   /* slot count hard-coded at detection_utils.py:611-616 */
#define CTDET_MAX_K 1024     /* largest top-K supported by ctdet_decode */

enum {
  CTDET_OK = 0,
  CTDET_EINVAL = 1,      /* bad shape / size / flag combination           */
  CTDET_ENULL = 2,       /* required pointer is NULL                       */
  CTDET_EALIGN = 3,      /* pointer not aligned as documented              */
  CTDET_EWORKSPACE = 4,  /* workspace too small                            */
  CTDET_ECUDA = 5,       /* CUDA launch / runtime error (cudaGetLastError) */
  CTDET_EUNSUPPORTED = 6 /* valid request this build does not implement    */
};

typedef void* ctdet_stream_t; /* cudaStream_t */

int ctdet_abi_version(void);
const char* ctdet_error_string(int code);
/* last CUDA error string seen by this thread inside the library ("" if none) */
const char* ctdet_last_cuda_error(void);

/* ---------------------------------------------------------------- target generation ----
 * boxes   f32 [total,4]  XYXY, input-pixel units (divided by the hard-coded 4 inside)
 * classes i64 [total]    in [0,C)
 * offsets i32 [B+1]      image b owns objects offsets[b] .. offsets[b+1]-1 (only the first 128 are used)
 * outputs: hm f32 [B,C,H,W] (every cell written exactly once, zero fill fused),
 *          reg_mask u8 [B,128], ind i64 [B,128], wh f32 [B,128,2], reg f32 [B,128,2],
 *          num_pos i32 [B] = number of cells of image b with hm == 1 (may be NULL).
 * hm must be 16-byte aligned. */
int ctdet_target_gen(const float* boxes, const int64_t* classes, const int32_t* offsets,
                     int B, int C, int H, int W,
                     float* hm, uint8_t* reg_mask, int64_t* ind, float* wh, float* reg,
                     int32_t* num_pos, ctdet_stream_t stream);

/* ---------------------------------------------------------------- fused losses ---------
 * One pass over the heatmap: p = clamp(sigmoid(hm_logits),1e-4,1-1e-4), penalty-reduced focal
 * loss, RegL1 for wh and reg, and (when d_hm != NULL) the gradients of
 *   w_hm*hm_loss + w_wh*wh_loss + w_off*off_loss   w.r.t. hm_logits / wh / reg.
 * Sections are optional: hm == NULL skips the focal term (losses[0] = 0), wh == NULL / reg == NULL
 * skip that RegL1 head; gradients are written for the sections present when their d_* pointer is
 * non-NULL (d_* must not alias an input).  With CTDET_LOSS_INPUT_PROB `hm` holds probabilities
 * (the `pred` argument of FocalLoss.forward, centernet.py:330) and d_hm is d/d(pred).
 * hm, gt_hm, d_hm: f32 [B,C,H,W];  wh, reg, d_wh, d_reg: f32 [B,2,H,W];
 * reg_mask u8 [B,M]; ind i64 [B,M]; gt_wh, gt_reg f32 [B,M,2]; alpha f32 [C] (per class,
 * applied to the positive term only, centernet.py:361).
 * Normaliser, first match wins:
 *   norm_override  f32 [2] = {num_pos, sum(reg_mask)} (e.g. all-reduced over ranks), else
 *   num_pos_img    i32 [B] as written by ctdet_target_gen (summed in-kernel), else
 *   a counting pre-pass over gt_hm (costs one extra read of gt_hm).
 * losses f32 [3] = {hm_loss*w_hm, wh_loss*w_wh, off_loss*w_off}.
 * workspace: ctdet_loss_workspace_bytes() bytes, zero-filled ONCE by the caller before first
 * use; every call leaves it zero-filled again.  Not re-entrant on two streams with the same
 * workspace. */
#define CTDET_LOSS_INPUT_PROB 1
#define CTDET_LOSS_ALPHA_UNIFORM 2 /* every class uses alpha[0] (skips the per-class lookup) */
size_t ctdet_loss_workspace_bytes(void);
int ctdet_loss_fwd_bwd(const float* hm, const float* wh, const float* reg,
                       const float* gt_hm, const uint8_t* reg_mask, const int64_t* ind,
                       const float* gt_wh, const float* gt_reg, const float* alpha,
                       const int32_t* num_pos_img, const float* norm_override,
                       int B, int C, int H, int W, int M, int flags,
                       float w_hm, float w_wh, float w_off,
                       float* losses, float* d_hm, float* d_wh, float* d_reg,
                       void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

/* ---------------------------------------------------------------- fused targets + losses -
 * gen_heatmap for the whole batch AND CenterNet.losses in one pass: the
 * Gaussian heatmap target is built per band in shared memory and consumed by the focal-loss sweep,
 * it never exists in HBM (8 B per heatmap cell instead of 4 + 12).  Objects as in ctdet_target_gen,
 * head outputs / alpha / norm_override / weights / losses / gradients as in ctdet_loss_fwd_bwd
 * (hm holds raw logits; CTDET_LOSS_ALPHA_UNIFORM is the only flag).  The slot targets
 * reg_mask u8 [B,128], ind i64 [B,128], gt_wh / gt_reg f32 [B,128,2] and num_pos i32 [B] are
 * OUTPUTS (written by a small pre-pass) and remain valid for the caller.
 * Requires W % 4 == 0 and W <= 4096.  workspace: ctdet_fused_workspace_bytes(B, C) bytes, 16-byte
 * aligned; its first 256 bytes follow the ctdet_loss_fwd_bwd workspace contract (zero-filled once
 * by the caller, left zeroed), the rest needs no initialisation. */
size_t ctdet_fused_workspace_bytes(int B, int C);
int ctdet_target_loss_fused(const float* boxes, const int64_t* classes, const int32_t* offsets,
                            const float* hm, const float* wh, const float* reg,
                            const float* alpha, const float* norm_override,
                            int B, int C, int H, int W, int flags,
                            float w_hm, float w_wh, float w_off,
                            uint8_t* reg_mask, int64_t* ind, float* gt_wh, float* gt_reg, int32_t* num_pos,
                            float* losses, float* d_hm, float* d_wh, float* d_reg,
                            void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

/* ---------------------------------------------------------------- fused hm head + focal loss --
 * SURVEY.md section 8 f2, second half: the hm head's final 1x1 convolution (centernet.py:118-123: nn.Conv2d(head_conv,
 * num_classes, kernel_size=1), bias -2.19; applied at :152) AND the focal term of CenterNet.losses (:204 sigmoid_ + clamp,
 * :323-369 FocalLoss / _neg_loss) with its backward down to the logits, in one kernel:
 *   logits[128 px, C] = feat[128 px, Cf] * weight^T + bias   per tile on the tensor cores (bf16 operands, fp32 accumulation)
 * are consumed from TMEM by the loss arithmetic of ctdet_loss_fwd_bwd; the [B,C,H,W] logits never exist in HBM.
 *   feat f32 [B,Cf,H,W] (the ReLU output of the head's 3x3 convolution)   weight f32 [C,Cf] (= the [C,Cf,1,1] kernel)
 *   bias f32 [C] or NULL   gt_hm f32 [B,C,H,W]   alpha / num_pos_img / norm_override / w_hm as in ctdet_loss_fwd_bwd
 *   loss f32 [1] = hm_loss * w_hm   d_logits f32 [B,C,H,W] or NULL: d(loss)/d(logits) -- the grad_output of the
 *   convolution's own backward (grad_input, grad_weight, grad_bias stay with the caller's Conv2d).
 * Requires Cf % 64 == 0, C % 16 == 0, C <= 128; weight 16-byte aligned (tiles of 128 pixels, the last one of an image may be ragged).
 * workspace: ctdet_loss_workspace_bytes() bytes, the ctdet_loss_fwd_bwd contract (zero on entry, left zero). */
int ctdet_head_focal_fused(const float* feat, const float* weight, const float* bias, const float* gt_hm,
                           const float* alpha, const int32_t* num_pos_img, const float* norm_override,
                           int B, int Cf, int C, int H, int W, int flags, float w_hm,
                           float* loss, float* d_logits, void* workspace, size_t workspace_bytes,
                           ctdet_stream_t stream);

/* The same with the convolution's backward inside (autograd of the nn.Conv2d at centernet.py:118-123): neither the logits nor
 * their gradient reach HBM (feat in, d_feat out).
 *   d_feat f32 [B,Cf,H,W] = d_logits x weight   d_weight f32 [C,Cf] = d_logits^T x feat   d_bias f32 [C] or NULL = sum d_logits
 *   (all OVERWRITTEN; d_logits is rounded to bf16 as the operand of the two contractions)   d_logits f32 [B,C,H,W] or NULL
 * Requires, beyond ctdet_head_focal_fused: Cf % 128 == 0, Cf <= 256 and C + (Cf/128) C + Cf <= 512 (TMEM columns: C = 80,
 * Cf = 256 uses 496). */
int ctdet_head_focal_fused_bwd(const float* feat, const float* weight, const float* bias, const float* gt_hm,
                               const float* alpha, const int32_t* num_pos_img, const float* norm_override,
                               int B, int Cf, int C, int H, int W, int flags, float w_hm,
                               float* loss, float* d_logits, float* d_feat, float* d_weight, float* d_bias,
                               void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

/* d[i] *= *g for i < n, entirely skipped on the device when *g == 1 (autograd upstream gradient,
 * no host sync).  g: f32 device scalar. */
int ctdet_scale_grad(float* d, size_t n, const float* g, ctdet_stream_t stream);

/* ---------------------------------------------------------------- decode ---------------
 * heat f32 [B,C,H,W]: probabilities in (0,1] (centernet.py:161), or raw logits with
 * CTDET_DECODE_LOGITS (ranking on logits, score = clamp(sigmoid) of the winners).
 * wh f32 [B,2,H,W] (or [B,2C,H,W] with CTDET_DECODE_CAT_SPEC_WH); reg f32 [B,2,H,W] or NULL (+0.5).
 * Outputs: boxes f32 [B,K,4] XYXY * down_ratio, scores f32 [B,K] descending, classes i32 [B,K],
 * inds i64 [B,K] spatial index y*W+x (may be NULL).
 * Tie rule: higher score first, then lower flat index c*H*W + y*W + x.  Requires K <= H*W,
 * K <= CTDET_MAX_K, C*H*W < 2^31.
 * workspace: ctdet_decode_workspace_bytes() bytes, 8-byte aligned, zero-filled ONCE by the caller
 * before first use: it starts with per-image headers and histograms that every call leaves
 * zeroed, at offsets that depend on B only, so one (large enough) workspace may serve calls of
 * different C/H/W/K with the same B; one workspace per stream. */
#define CTDET_DECODE_LOGITS 1
#define CTDET_DECODE_CAT_SPEC_WH 2
#define CTDET_DECODE_NO_NMS 4 /* rank every cell (the bare _topk of centernet.py:408) */
size_t ctdet_decode_workspace_bytes(int B, int C, int H, int W, int K);
int ctdet_decode(const float* heat, const float* wh, const float* reg,
                 int B, int C, int H, int W, int K, float down_ratio, int flags,
                 float* boxes, float* scores, int32_t* classes, int64_t* inds,
                 void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

/* ctdet_decode without a caller workspace, for callers that cannot allocate device memory (code under CUDA graph
 * capture, tight serving loops, harnesses that forbid cudaMalloc).  NO COUNTERPART IN THE REFERENCE.
 * Same arguments and results as ctdet_decode; the scratch is a file-scope
 * `static __device__` array of CTDET_DECODE_STATIC_BYTES inside the library (every GPU that loads the library holds
 * it until exit).  The call allocates nothing: no cudaMalloc / cudaFree, no allocating library routine -- one
 * cudaGetSymbolAddress per device at first use, one cudaMemsetAsync of the zero-on-entry prefix per call, then
 * ctdet_decode's launches.  Shapes whose ctdet_decode_workspace_bytes(B,C,H,W,K) exceeds CTDET_DECODE_STATIC_BYTES
 * (16 MiB: B <= 8 at 80 x 128 x 128, K = 100) return CTDET_EWORKSPACE and ctdet_last_cuda_error() names the bytes
 * needed and the limit.  ALL calls on a device share the one array (it keeps the last call's contents; what must
 * start at zero is cleared by the call): two calls that overlap on the GPU -- different streams, or graph branches --
 * RACE; serialise them, or use ctdet_decode with one workspace per stream. */
#define CTDET_DECODE_STATIC_BYTES (16u << 20)
size_t ctdet_decode_static_max_bytes(void);
int ctdet_decode_static(const float* heat, const float* wh, const float* reg,
                        int B, int C, int H, int W, int K, float down_ratio, int flags,
                        float* boxes, float* scores, int32_t* classes, int64_t* inds,
                        ctdet_stream_t stream);

/* ---------------------------------------------------------------- fused hm head + decode --
 * SURVEY.md section 8 f2, second half, decode side: the hm head's final 1x1 convolution (centernet.py:118-123, applied :152),
 * hm.sigmoid_() (:161) and ctdet_decode (:426-458) without the [B,C,H,W] logits in HBM: every map row (W <= 128 pixels) is
 *   logits[128 px, C] = feat[128 px, Cf] * weight^T + bias   on the tensor cores (bf16 operands, fp32 accumulation)
 * and the 3x3 peak test / threshold sweep of ctdet_decode runs on a three-row window of them in shared memory (a sampled
 * pass for the bound, then the full sweep; the exact fall-back for ties / plateaus / tiny maps stores the logits of the
 * images concerned in the workspace and decodes them as ctdet_decode does).  Semantics = ctdet_decode with
 * CTDET_DECODE_LOGITS on these logits: outputs as there; flags: CTDET_DECODE_CAT_SPEC_WH only.
 *   feat f32 [B,Cf,H,W]   weight f32 [C,Cf]   bias f32 [C] or NULL   wh / reg / outputs as ctdet_decode
 *   logits_out f32 [B,C,H,W] or NULL: when given, the logits are ALSO written (one extra pass; verification / callers
 *   that want both) -- ctdet_decode(logits_out, CTDET_DECODE_LOGITS) then returns the same detections bit for bit.
 * Requires W <= 128 (a map row is one tensor-core tile: 128 = a 512-pixel-wide input at stride 4 uses it fully, 104 = a
 * 416-pixel one to 81 %), Cf % 64 == 0, C % 16 == 0, C <= 128, shared memory for
 * 3 rows of logits (C = 80, Cf = 256: 213 KB); else CTDET_EUNSUPPORTED (run the convolution + ctdet_decode).
 * workspace: ctdet_head_decode_fused_workspace_bytes() bytes, 256-byte aligned, zero on first use, left zero where it must
 * be (it starts with ctdet_decode's zero-on-entry prefix and may be the same buffer for a given B). */
size_t ctdet_head_decode_fused_workspace_bytes(int B, int Cf, int C, int H, int W, int K);
int ctdet_head_decode_fused(const float* feat, const float* weight, const float* bias,
                            const float* wh, const float* reg, int B, int Cf, int C, int H, int W, int K,
                            float down_ratio, int flags, float* boxes, float* scores, int32_t* classes,
                            int64_t* inds, float* logits_out, void* workspace, size_t workspace_bytes,
                            ctdet_stream_t stream);

/* heat*keep of centernet.py:399-405, dense: out f32 [B,C,H,W]. */
int ctdet_nms(const float* heat, int B, int C, int H, int W, float* out, ctdet_stream_t stream);

/* _transpose_and_gather_feat, centernet.py:383: feat f32 [B,Ch,H*W], ind i64 [B,K] ->
 * out f32 [B,K,Ch] with out[b,k,c] = feat[b,c,ind[b,k]] (no NHWC copy). */
int ctdet_gather_feat(const float* feat, const int64_t* ind, int B, int Ch, int HW, int K,
                      float* out, ctdet_stream_t stream);

/* Batched inference_single_image + detector_postprocess: for image b keep the first
 * min(max_dets,K) detections with score > score_thresh, scale boxes by
 * (out_w/in_w, out_h/in_h), clip to the output size, drop empty boxes; survivors are compacted
 * to the front (order preserved).  sizes i32 [B,4] = {in_h, in_w, out_h, out_w}; n_keep i32 [B].
 * out_h <= 0 or out_w <= 0 for an image: CenterNet.inference alone (slice + threshold; no scale / clip / nonempty),
 * i.e. the Instances `inference` returns before the reference's forward calls detector_postprocess (:169-171).
 * In place on boxes/scores/classes [B,K,...]. */
int ctdet_postprocess(float* boxes, float* scores, int32_t* classes, const int32_t* sizes,
                      int B, int K, int max_dets, float score_thresh, int32_t* n_keep,
                      ctdet_stream_t stream);

/* CenterNet.preprocess_image (meta_arch/centernet.py:177-184) + ImageList.from_tensors
 * (structures/image_list.py:58-130), batched: out[b,c,y,x] = ((img_b[c,y,x] / 255) - mean[c]) / std[c]
 * inside image b, 0 (pad_value) outside; fp32, the reference's operation order.
 *   images   DEVICE array of B device pointers, image b = [C,h_b,w_b]
 *            contiguous, uint8 (CTDET_IMAGE_U8, what the dataset mapper yields) or f32 (CTDET_IMAGE_F32)
 *   sizes_hw i32 [B,2] = {h_b, w_b}   mean, std f32 [C]   out f32 [B,C,Hp,Wp], Wp % 4 == 0
 * Hp / Wp are the batch maxima rounded up to the backbone's size_divisibility (host side). */
#define CTDET_IMAGE_U8 0
#define CTDET_IMAGE_F32 1
int ctdet_preprocess(const void* const* images, const int32_t* sizes_hw, int dtype,
                     int B, int C, int Hp, int Wp, const float* mean, const float* std,
                     float* out, ctdet_stream_t stream);

/* IDAUp up-sampling + skip add (modeling/backbone/dla.py:166-168, :175-177):
 *   out = ConvTranspose2d(C, C, 2f, stride=f, padding=f/2, groups=C, bias=False)(x) + skip
 *   x f32 [B,C,h,w]   weight f32 [C,1,2f,2f] (the module's weight: bilinear at init, learnable)
 *   skip f32 [B,C,f*h,f*w] or NULL (plain up-sampling)   out f32 [B,C,f*h,f*w]   f in {2,4,8,16} */
int ctdet_upsample_add(const float* x, const float* weight, const float* skip, int B, int C, int h, int w,
                       int f, float* out, ctdet_stream_t stream);

/* Backward of ctdet_upsample_add (autograd of `up_i(x) + layers[i-1]`, dla.py:175-177; in the reference: the
 * grouped ConvTranspose2d's cuDNN backward-data / backward-filter kernels + the add's pass-through):
 *   grad_out f32 [B,C,f*h,f*w]   x f32 [B,C,h,w] (needed for grad_weight only)   weight f32 [C,1,2f,2f]
 *   grad_x f32 [B,C,h,w] or NULL   grad_weight f32 [C,1,2f,2f] or NULL (overwritten)   f in {2,4}
 * grad_skip is grad_out itself (no pass). */
int ctdet_upsample_add_backward(const float* grad_out, const float* x, const float* weight, int B, int C, int h,
                                int w, int f, float* grad_x, float* grad_weight, ctdet_stream_t stream);

/* ---------------------------------------------------------------- DCNv2 forward ----------
 * Modulated deformable 3x3 convolution, replaces modulated_deform_conv_forward
 * (detectron2/layers/csrc/deformable/deform_conv.h:262-312; same argument meaning, minus the
 * `ones`/`columns` scratch tensors: the im2col tile lives in shared memory and feeds tcgen05 MMAs).
 *   input  f32 [B,Cin,H,W]      weight f32 [Cout,Cin,3,3]    bias f32 [Cout] or NULL (with_bias = 0)
 *   offset f32 [B,18,Ho,Wo]     (channel 2t = dh, 2t+1 = dw of tap t)   mask f32 [B,9,Ho,Wo]
 *   output f32 [B,Cout,Ho,Wo]
 * Optional fused epilogue (DeformConvV2.actf, detectron2/layers/deform_conv.py:501-504, with the
 * BatchNorm folded for inference): y = (conv + bias) * scale[c] + shift[c], then ReLU with
 * DCN_FLAG_RELU; scale/shift NULL to skip.
 * Operands are rounded to bf16, accumulation is fp32 (tolerance: see tests/test_gpu_dcn.py).
 * DCN_FLAG_FP32_BLEND evaluates the bilinear blend in fp32 before the bf16 rounding (slower).
 * Supported: kernel 3x3, group = deformable_group = 1, square stride/pad/dilation,
 * Cin % 64 == 0, Cout % 16 == 0, 16 <= Cout <= 256; anything else returns CTDET_EUNSUPPORTED.
 * workspace: dcnv2_workspace_bytes() bytes, 16-byte aligned (NHWC bf16 copy of the input + packed
 * weights), no initialisation needed. */
#define DCN_FLAG_RELU 1
#define DCN_FLAG_FP32_BLEND 2
#define DCN_FLAG_PACKED_WEIGHT 4 /* `weight` points at dcnv2_pack_weight's output for these weights: skips the per-call pack */
size_t dcnv2_workspace_bytes(int B, int Cin, int H, int W, int Cout);
/* The bf16, swizzled UMMA image of weight f32 [Cout,Cin,3,3] (dcnv2_packed_weight_bytes() bytes, 16-byte aligned) that
 * dcnv2_forward / dcnv2_forward_raw27 otherwise rebuild on every call.  A caller keeps it per module and weight version
 * (the reference has no counterpart: its GEMM reads the fp32 weights, deform_conv_cuda.cu:904) and passes it as `weight`
 * with DCN_FLAG_PACKED_WEIGHT. */
size_t dcnv2_packed_weight_bytes(int Cin, int Cout);
int dcnv2_pack_weight(const float* weight, int Cin, int Cout, void* packed, ctdet_stream_t stream);
int dcnv2_forward(const float* input, const float* weight, const float* bias,
                  const float* offset, const float* mask, float* output,
                  int B, int Cin, int H, int W, int Cout, int kernel_h, int kernel_w,
                  int stride_h, int stride_w, int pad_h, int pad_w, int dilation_h, int dilation_w,
                  int group, int deformable_group, const float* scale, const float* shift, int flags,
                  void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

/* dcnv2_forward fed by the RAW output of DCN.conv_offset_mask (the 27-channel 3x3 convolution of `input`):
 * replaces the chunk / cat / sigmoid glue between that convolution and the contraction
 * (modeling/backbone/resnet.py:308-311: o1, o2, mask = chunk(out, 3); offset = cat(o1, o2) -- i.e. channels
 * 0..17 unchanged -- and mask = sigmoid(channels 18..26)) and, with offset_mask_bias, the convolution's bias add.
 *   offset_mask f32 [B,27,Ho,Wo]   offset_mask_bias f32 [27] or NULL (bias already added)
 * Everything else as dcnv2_forward (3x3, one group, one deformable group). */
int dcnv2_forward_raw27(const float* input, const float* weight, const float* bias,
                        const float* offset_mask, const float* offset_mask_bias, float* output,
                        int B, int Cin, int H, int W, int Cout, int stride, int pad, int dilation,
                        const float* scale, const float* shift, int flags,
                        void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

/* The same with conv_offset_mask itself inside the call: `om_weight` f32 [27,Cin,3,3] /
 * `om_bias` f32 [27] or NULL are the parameters of the DCN's `nn.Conv2d(Cin, 27, 3, stride 1, padding 1)`
 * (deform_conv.py:505-513; chunk / cat / sigmoid convention resnet.py:308-313).  x is packed to NHWC bf16 ONCE; the
 * 27-channel convolution runs on the tensor cores over that copy (the main kernel in its plain mode: zero offsets, mask 1,
 * one corner per tap, bf16 operands, fp32 accumulation) and its raw output feeds the DCN kernel's geometry stage as in
 * dcnv2_forward_raw27.  stride = padding = dilation = 1 (what DLA-34 uses).
 *   offset_mask_out f32 [B,27,H,W] or NULL: where the raw convolution output (without bias) goes; NULL = in the workspace.
 * workspace: dcnv2_fused_offsets_workspace_bytes() bytes, 16-byte aligned.  Other arguments / flags as dcnv2_forward. */
size_t dcnv2_fused_offsets_workspace_bytes(int B, int Cin, int H, int W, int Cout);
int dcnv2_forward_fused_offsets(const float* input, const float* weight, const float* bias,
                                const float* om_weight, const float* om_bias, float* output,
                                float* offset_mask_out, int B, int Cin, int H, int W, int Cout,
                                const float* scale, const float* shift, int flags,
                                void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

/* An ordinary 3x3 convolution, stride 1, padding 1, + bias + optional ReLU (DCN_FLAG_RELU) on the tensor cores (bf16
 * operands, fp32 accumulation): the first layer of the ctdet heads, nn.Conv2d(64, head_conv, 3, padding=1) + ReLU
 * (centernet.py:112-121), as the same kernel in its plain mode.  input f32 [B,Cin,H,W], weight f32 [Cout,Cin,3,3], bias
 * f32 [Cout] or NULL, output f32 [B,Cout,H,W].  Requires Cin % 64 == 0.  workspace: ctdet_conv3x3_workspace_bytes(). */
size_t ctdet_conv3x3_workspace_bytes(int B, int Cin, int H, int W, int Cout);
int ctdet_conv3x3(const float* input, const float* weight, const float* bias, float* output,
                  int B, int Cin, int H, int W, int Cout, int flags,
                  void* workspace, size_t workspace_bytes, ctdet_stream_t stream);
/* `n` such convolutions of ONE input (the heads' hm / wh / reg branches): the input is packed to NHWC bf16 once.
 * weights / biases / outputs: host arrays of n device pointers (biases or its entries may be NULL), couts: host array of n
 * channel counts; workspace: ctdet_conv3x3_workspace_bytes() with the largest Cout. */
int ctdet_conv3x3_multi(const float* input, int n, const float* const* weights, const float* const* biases,
                        float* const* outputs, const int* couts, int B, int Cin, int H, int W, int flags,
                        void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

/* ---------------------------------------------------------------- DCNv2 backward ---------
 * Replaces modulated_deform_conv_backward (deform_conv.h:314-338; same argument meaning, minus the
 * `ones`/`columns` scratch tensors and the unused `bias`).  All gradient outputs are OVERWRITTEN
 * (the reference accumulates into caller-zeroed tensors, deform_conv.py:230-234 -- same result).
 *   grad_output f32 [B,Cout,Ho,Wo]
 *   grad_input  f32 [B,Cin,H,W]   grad_offset f32 [B,18,Ho,Wo]   grad_mask f32 [B,9,Ho,Wo]
 *       -- produced together: pass all three or none (NULL)
 *   grad_weight f32 [Cout,Cin,3,3] or NULL     grad_bias f32 [Cout] or NULL
 * Operands (input, weight, grad_output and the column gradient) are rounded to bf16, accumulation
 * is fp32 (tolerance: tests/test_gpu_dcn_bwd.py).  grad_input / grad_weight are accumulated with
 * fp32 atomics, so their low bits vary from run to run -- exactly like the reference's col2im
 * (deform_conv_cuda_kernel.cu:944).
 * Supported: as dcnv2_forward, and Cout % 64 == 0.  flags must be 0.
 * workspace: dcnv2_backward_workspace_bytes() bytes, 16-byte aligned, no initialisation needed. */
size_t dcnv2_backward_workspace_bytes(int B, int Cin, int H, int W, int Cout, int Ho, int Wo);
int dcnv2_backward(const float* input, const float* weight, const float* offset, const float* mask,
                   const float* grad_output, float* grad_input, float* grad_weight, float* grad_bias,
                   float* grad_offset, float* grad_mask,
                   int B, int Cin, int H, int W, int Cout, int kernel_h, int kernel_w,
                   int stride_h, int stride_w, int pad_h, int pad_w, int dilation_h, int dilation_w,
                   int group, int deformable_group, int flags,
                   void* workspace, size_t workspace_bytes, ctdet_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* CTDET_B200_H_ */
