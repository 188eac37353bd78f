/* TEST INFRASTRUCTURE ONLY -- CPU restatement ("port") of the reference's modulated deformable
 * convolution forward.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg may
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
