/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference's modulated deformable
 * convolution forward and backward.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
 * load this; the product (libctdet_b200.so) never does.
 *
 * Restates, in plain C and fp32 like the reference:
 *   detectron2/layers/csrc/deformable/deform_conv_cuda_kernel.cu:786-868
 *       modulated_deformable_im2col_gpu_kernel (tap order, offset channel 2t = dh, 2t+1 = dw,
 *       sampling allowed iff h_im > -1 && w_im > -1 && h_im < H && w_im < W, :852)
 *   deform_conv_cuda_kernel.cu:666-699  dmcn_im2col_bilinear (each corner zero when outside)
 *   deform_conv_cuda.cu:904-909 (columns x weight GEMM per image) and :924-926 (bias add).
 * The reference has NO CPU implementation of this op (deform_conv.h:311), so parity is pinned
 * against torchvision.ops.deform_conv2d(mask=...) in tests/test_dcn_oracle.py instead of against
 * the reference's own code; groups and deformable_groups are fixed to 1 (what DLA-34 uses,
 * detectron2/layers/deform_conv.py:505-513).
 */
#include <math.h>
#include <stdlib.h>
#include <string.h>

static float bilinear(const float* im, int H, int W, float h, float w) {
  const int h_low = (int)floorf(h), w_low = (int)floorf(w);
  const int h_high = h_low + 1, w_high = w_low + 1;
  const float lh = h - (float)h_low, lw = w - (float)w_low;
  const float hh = 1.f - lh, hw = 1.f - lw;
  const float v1 = (h_low >= 0 && w_low >= 0) ? im[h_low * W + w_low] : 0.f;
  const float v2 = (h_low >= 0 && w_high <= W - 1) ? im[h_low * W + w_high] : 0.f;
  const float v3 = (h_high <= H - 1 && w_low >= 0) ? im[h_high * W + w_low] : 0.f;
  const float v4 = (h_high <= H - 1 && w_high <= W - 1) ? im[h_high * W + w_high] : 0.f;
  return hh * hw * v1 + hh * lw * v2 + lh * hw * v3 + lh * lw * v4;
}

/* x [B,Cin,H,W], offset [B,18,Ho,Wo], mask [B,9,Ho,Wo], weight [Cout,Cin,3,3], bias [Cout] or NULL,
 * out [B,Cout,Ho,Wo].  Returns 0, or 1 when the scratch allocation fails. */
int dcn_oracle_forward(const float* x, const float* offset, const float* mask, const float* weight,
                       const float* bias, int B, int Cin, int H, int W, int Cout, int stride, int pad,
                       int dil, float* out) {
  const int Ho = (H + 2 * pad - (dil * 2 + 1)) / stride + 1;
  const int Wo = (W + 2 * pad - (dil * 2 + 1)) / stride + 1;
  const int P = Ho * Wo, Kdim = Cin * 9;
  int rc = 0;

  for (int b = 0; b < B; ++b) {
    float* col = (float*)malloc((size_t)Kdim * P * sizeof(float));      /* columns [Cin*9, Ho*Wo] */
    if (!col) { rc = 1; continue; }
    for (int c = 0; c < Cin; ++c) {
      const float* im = x + ((size_t)b * Cin + c) * H * W;
      for (int t = 0; t < 9; ++t) {
        const int i = t / 3, j = t % 3;
        const float* oh = offset + ((size_t)b * 18 + 2 * t) * P;
        const float* ow = offset + ((size_t)b * 18 + 2 * t + 1) * P;
        const float* mk = mask + ((size_t)b * 9 + t) * P;
        float* dst = col + (size_t)(c * 9 + t) * P;
        for (int ho = 0; ho < Ho; ++ho)
          for (int wo = 0; wo < Wo; ++wo) {
            const int p = ho * Wo + wo;
            const float h_im = (float)(ho * stride - pad + i * dil) + oh[p];
            const float w_im = (float)(wo * stride - pad + j * dil) + ow[p];
            float v = 0.f;
            if (h_im > -1.f && w_im > -1.f && h_im < (float)H && w_im < (float)W) v = bilinear(im, H, W, h_im, w_im);
            dst[p] = v * mk[p];
          }
      }
    }
    for (int co = 0; co < Cout; ++co) {
      float* o = out + ((size_t)b * Cout + co) * P;
      const float bv = bias ? bias[co] : 0.f;
      for (int p = 0; p < P; ++p) o[p] = 0.f;
      const float* wrow = weight + (size_t)co * Kdim;
      for (int k = 0; k < Kdim; ++k) {
        const float wv = wrow[k];
        const float* cr = col + (size_t)k * P;
        for (int p = 0; p < P; ++p) o[p] += wv * cr[p];
      }
      for (int p = 0; p < P; ++p) o[p] += bv;
    }
    free(col);
  }
  return rc;
}

/* ---------------------------------------------------------------------------------------------
 * Backward.  Restates modulated_deform_conv_cuda_backward (deform_conv_cuda.cu:929-1129) per image:
 *   columns = weight^T x grad_output                                   (:1003-1008)
 *   modulated_deformable_col2im_coord_gpu_kernel  -> grad_offset, grad_mask
 *       (deform_conv_cuda_kernel.cu:952-1060; dmcn_get_coordinate_weight :727-784)
 *   modulated_deformable_col2im_gpu_kernel        -> grad_input        (:871-949; the 5x5 window test
 *       |h_im - y| < 1 && |w_im - x| < 1 selects exactly the in-image bilinear corners, weight
 *       dmcn_get_gradient_weight :700-725)
 *   im2col again, grad_weight += grad_output x columns^T, grad_bias += grad_output x ones (:1055-1100)
 * All gradient outputs are overwritten.  fp32 storage, double accumulation for the long sums.
 */
int dcn_oracle_backward(const float* x, const float* offset, const float* mask, const float* weight,
                        const float* gout, int B, int Cin, int H, int W, int Cout, int stride, int pad,
                        int dil, float* gx, float* goffset, float* gmask, float* gweight, float* gbias) {
  const int Ho = (H + 2 * pad - (dil * 2 + 1)) / stride + 1;
  const int Wo = (W + 2 * pad - (dil * 2 + 1)) / stride + 1;
  const int P = Ho * Wo, Kdim = Cin * 9;
  double* gw64 = (double*)calloc((size_t)Cout * Kdim, sizeof(double));
  double* gb64 = (double*)calloc((size_t)Cout, sizeof(double));
  float* gcol = (float*)malloc((size_t)Kdim * P * sizeof(float));
  float* col = (float*)malloc((size_t)Kdim * P * sizeof(float));
  if (!gw64 || !gb64 || !gcol || !col) { free(gw64); free(gb64); free(gcol); free(col); return 1; }
  memset(gx, 0, (size_t)B * Cin * H * W * sizeof(float));
  memset(goffset, 0, (size_t)B * 18 * P * sizeof(float));
  memset(gmask, 0, (size_t)B * 9 * P * sizeof(float));

  for (int b = 0; b < B; ++b) {
    const float* go = gout + (size_t)b * Cout * P;
    /* columns[(c*9+t), p] = sum_co weight[co, c, t] * grad_output[co, p] */
    for (int k = 0; k < Kdim; ++k) {
      float* gc = gcol + (size_t)k * P;
      for (int p = 0; p < P; ++p) gc[p] = 0.f;
      for (int co = 0; co < Cout; ++co) {
        const float wv = weight[(size_t)co * Kdim + k];
        const float* g = go + (size_t)co * P;
        for (int p = 0; p < P; ++p) gc[p] += wv * g[p];
      }
    }
    for (int c = 0; c < Cin; ++c) {
      const float* im = x + ((size_t)b * Cin + c) * H * W;
      float* gim = gx + ((size_t)b * Cin + c) * H * W;
      for (int t = 0; t < 9; ++t) {
        const int i = t / 3, j = t % 3;
        const float* oh = offset + ((size_t)b * 18 + 2 * t) * P;
        const float* ow = offset + ((size_t)b * 18 + 2 * t + 1) * P;
        const float* mk = mask + ((size_t)b * 9 + t) * P;
        float* goh = goffset + ((size_t)b * 18 + 2 * t) * P;
        float* gow = goffset + ((size_t)b * 18 + 2 * t + 1) * P;
        float* gmk = gmask + ((size_t)b * 9 + t) * P;
        const float* gc = gcol + (size_t)(c * 9 + t) * P;
        float* cl = col + (size_t)(c * 9 + t) * P;
        for (int ho = 0; ho < Ho; ++ho)
          for (int wo = 0; wo < Wo; ++wo) {
            const int p = ho * Wo + wo;
            const float h_im = (float)(ho * stride - pad + i * dil) + oh[p];
            const float w_im = (float)(wo * stride - pad + j * dil) + ow[p];
            cl[p] = 0.f;
            if (!(h_im > -1.f && w_im > -1.f && h_im < (float)H && w_im < (float)W)) continue;
            const int h_low = (int)floorf(h_im), w_low = (int)floorf(w_im);
            const int h_high = h_low + 1, w_high = w_low + 1;
            const float lh = h_im - (float)h_low, lw = w_im - (float)w_low, hh = 1.f - lh, hw = 1.f - lw;
            const int ok1 = h_low >= 0 && w_low >= 0, ok2 = h_low >= 0 && w_high <= W - 1;
            const int ok3 = h_high <= H - 1 && w_low >= 0, ok4 = h_high <= H - 1 && w_high <= W - 1;
            const float v1 = ok1 ? im[h_low * W + w_low] : 0.f, v2 = ok2 ? im[h_low * W + w_high] : 0.f;
            const float v3 = ok3 ? im[h_high * W + w_low] : 0.f, v4 = ok4 ? im[h_high * W + w_high] : 0.f;
            const float val = hh * hw * v1 + hh * lw * v2 + lh * hw * v3 + lh * lw * v4;
            const float g = gc[p], m = mk[p];
            cl[p] = val * m;                                            /* im2col for grad_weight */
            gmk[p] += g * val;                                          /* :1019-1030 */
            goh[p] += (-hw * v1 - lw * v2 + hw * v3 + lw * v4) * g * m; /* bp_dir 0, :744-757 */
            gow[p] += (-hh * v1 + hh * v2 - lh * v3 + lh * v4) * g * m; /* bp_dir 1, :758-771 */
            const float gm = g * m;                                     /* cur_top_grad, :925 */
            if (ok1) gim[h_low * W + w_low] += hh * hw * gm;
            if (ok2) gim[h_low * W + w_high] += hh * lw * gm;
            if (ok3) gim[h_high * W + w_low] += lh * hw * gm;
            if (ok4) gim[h_high * W + w_high] += lh * lw * gm;
          }
      }
    }
    for (int co = 0; co < Cout; ++co) {
      const float* g = go + (size_t)co * P;
      double sb = 0.0;
      for (int p = 0; p < P; ++p) sb += g[p];
      gb64[co] += sb;
      for (int k = 0; k < Kdim; ++k) {
        const float* cr = col + (size_t)k * P;
        double s = 0.0;
        for (int p = 0; p < P; ++p) s += (double)g[p] * cr[p];
        gw64[(size_t)co * Kdim + k] += s;
      }
    }
  }
  for (size_t i = 0; i < (size_t)Cout * Kdim; ++i) gweight[i] = (float)gw64[i];
  for (int i = 0; i < Cout; ++i) gbias[i] = (float)gb64[i];
  free(gw64); free(gb64); free(gcol); free(col);
  return 0;
}
