// Fused ctdet losses, forward AND backward, in one pass over the heatmap.
//
// Replaces CenterNet.losses (detectron2/modeling/meta_arch/centernet.py:191-212): the in-place
// sigmoid + clamp (:204), _neg_loss (:333-369, ~14 full-tensor ATen passes), two RegL1Loss calls
// (:372-381) with their full-map NHWC permute (:383-387), and the autograd backward of all of it.
//
// Algorithmic traffic: read logits (4) + read gt (4) + write d_logits (4) = 12 B per heatmap cell
// (8 B/cell forward-only), plus zero-fill + scatter of the dense wh/reg gradients.
//
// One launch:  CTAs [0, 2B)   -> RegL1 for (image, head): zero-fill d[b,0:2,:,:], gather at ind,
//                                 |.| sum, scatter-add the gradient (duplicates accumulate)
//              CTAs [2B, grid) -> grid-stride float4 streaming sweep of the heatmap
//              last CTA to finish -> the three loss scalars; resets the workspace to zero.
#include "common.cuh"

namespace ctdet {

constexpr int kLossThreads = 256;
constexpr int kUnroll = 4;         // float4 chunks in flight per thread (x and gt: 8 x LDG.128)

struct LossWs {                    // must be zero on entry; left zero on exit
  double pos_sum, neg_sum, wh_sum, off_sum;
  unsigned long long npos_count;   // counted positives (pre-pass), when no normaliser is given
  unsigned int done;
  float nmask;                     // sum(reg_mask), published by RegL1 CTA 0 for the finaliser
};

__global__ void __launch_bounds__(kLossThreads)
count_pos_kernel(const float* __restrict__ gt, size_t n4, LossWs* ws) {
  int cnt = 0;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n4;
       i += (size_t)gridDim.x * blockDim.x) {
    const float4 g = ldg_stream4(gt + 4 * i);
    cnt += (g.x == 1.f) + (g.y == 1.f) + (g.z == 1.f) + (g.w == 1.f);
  }
  cnt = warp_sum(cnt);
  if ((threadIdx.x & 31) == 0 && cnt) atomicAdd(&ws->npos_count, (unsigned long long)cnt);
}

// One heatmap cell.  Follows the reference's fp32 evaluation order (:350-357): pred =
// clamp(sigmoid(x)); log(pred)*(1-pred)^2 for positives, log(1-pred)*pred^2*(1-gt)^4 for
// negatives.  Returns d(term)/dx (zero where the clamp is active: torch.clamp passes gradient only
// for 1e-4 <= p <= 1-1e-4) and accumulates the un-normalised loss terms in pos/neg.
// Instruction budget: the sweep has ~60 issue slots per cell at the HBM roofline, so the sigmoid
// uses the MUFU paths (ex2.approx / rcp.approx: <= 1e-6 relative on p over the unclamped range
// |x| <= 9.3, measured against torch in the tests); log(1-p) stays the accurate logf because its
// value can be as small as 1e-4 and needs relative, not absolute, accuracy.
__device__ __forceinline__ float rcp_approx(float a) {
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(a));
  return r;
}

template <bool kProb>
__device__ __forceinline__ float cell(float x, float g, float alpha, float& pos, float& neg) {
  float pc, dsig;
  if (kProb) {                                       // x already is pred (FocalLoss drop-in, :323)
    pc = x; dsig = 1.f;
  } else {
    const float p = rcp_approx(1.f + __expf(-x));    // sigmoid
    pc = fminf(fmaxf(p, 1e-4f), 1.f - 1e-4f);
    dsig = (pc == p) ? p * (1.f - p) : 0.f;          // d clamp(sigmoid(x)) / dx
  }
  const float om = 1.f - pc;
  // Branch-free over the (rare) positives: with A = the factor that is squared and L = the log's
  // argument, term = coef * log(L) * A^2 and d(term)/dp = sgn * coef * (2 A log(L) - A^2 / L):
  //   negatives: A = p,   L = 1-p, coef = (1-g)^4, sgn = +1      (:355)
  //   positives: A = 1-p, L = p,   coef = alpha,   sgn = -1      (:354, :361)
  const bool is_pos = (g == 1.f);
  const float w = (1.f - g) * (1.f - g);
  const float A = is_pos ? om : pc;
  const float Lg = is_pos ? pc : om;
  const float coef = is_pos ? alpha : w * w;
  const float l = logf(Lg);                           // accurate: can be as small as 1e-4
  const float aa = A * A;
  const float term = coef * (l * aa);
  if (is_pos) pos += term; else neg += term;
  float dterm = coef * ((A + A) * l - aa * rcp_approx(Lg));
  dterm = is_pos ? -dterm : dterm;
  return dterm * dsig;
}

template <bool kProb>
__global__ void __launch_bounds__(kLossThreads)
loss_kernel(const float* __restrict__ hm, const float* __restrict__ wh, const float* __restrict__ reg,
            const float* __restrict__ gt, const uint8_t* __restrict__ reg_mask,
            const int64_t* __restrict__ ind, const float* __restrict__ gt_wh,
            const float* __restrict__ gt_reg, const float* __restrict__ alpha,
            const int32_t* __restrict__ num_pos_img, const float* __restrict__ norm_override,
            int B, int C, int HW, int M, int n_reg_ctas, FastDiv div_hw4, FastDiv div_c,
            float w_hm, float w_wh, float w_off, float* __restrict__ losses, float* __restrict__ d_hm, float* __restrict__ d_wh,
            float* __restrict__ d_reg, LossWs* ws) {
  __shared__ float s_norm[2];
  __shared__ double s_red[2][kLossThreads / 32];
  __shared__ bool s_last;
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const bool is_reg = (int)blockIdx.x < n_reg_ctas;

  // ---- normalisers: num_pos (focal) and sum(reg_mask) (RegL1), identical in every CTA ------
  {
    float npos = 0.f, nmask = 0.f;
    if (norm_override != nullptr) {
      if (tid == 0) { npos = norm_override[0]; nmask = norm_override[1]; }
    } else {
      if (warp == 0) {
        if (num_pos_img != nullptr) {
          int a = 0;
          for (int i = lane; i < B; i += 32) a += num_pos_img[i];
          npos = (float)warp_sum(a);
        } else {
          npos = (float)ws->npos_count;     // written by count_pos_kernel (stream order)
        }
      }
      if (is_reg && warp == 1) {
        int a = 0;
        for (int i = lane; i < B * M; i += 32) a += reg_mask[i];
        nmask = (float)warp_sum(a);
        if (lane == 0) s_norm[1] = nmask;
      }
    }
    if (tid == 0) { s_norm[0] = npos; if (norm_override != nullptr) s_norm[1] = nmask; }
    __syncthreads();
  }

  double acc0 = 0.0, acc1 = 0.0;   // focal: (pos, neg)   RegL1: (abs-sum, unused)

  if (is_reg) {
    // ---- RegL1 (:376-381) for image b, head h (0 = wh, 1 = reg/offset) ----------------------
    const int b = blockIdx.x >> 1, h = blockIdx.x & 1;
    const float* out = (h == 0 ? wh : reg);
    const float* tgt = (h == 0 ? gt_wh : gt_reg) + (size_t)b * M * 2;
    float* dout = (h == 0 ? d_wh : d_reg);
    const int Mh = (out != nullptr) ? M : 0;                  // head absent: contributes nothing
    if (out != nullptr) out += (size_t)b * 2 * HW; else dout = nullptr;
    const float denom = 2.f * s_norm[1] + 1e-4f;             // mask expanded to 2 channels (:378-380)
    const float gscale = (h == 0 ? w_wh : w_off) / denom;
    if (dout != nullptr) {
      dout += (size_t)b * 2 * HW;
      for (int i = tid; i < 2 * HW; i += kLossThreads) dout[i] = 0.f;
      __syncthreads();
    }
    float a = 0.f;
    for (int k = tid; k < Mh; k += kLossThreads) {
      const float m = (float)reg_mask[(size_t)b * M + k];
      const int64_t i = ind[(size_t)b * M + k];
      if (i < 0 || i >= HW) continue;                         // the reference's gather would raise
#pragma unroll
      for (int ch = 0; ch < 2; ++ch) {
        const float diff = out[(size_t)ch * HW + i] * m - tgt[2 * k + ch] * m;
        a += fabsf(diff);
        if (dout != nullptr && m != 0.f && diff != 0.f)       // sign(0) = 0
          atomicAdd(dout + (size_t)ch * HW + i, (diff > 0.f ? gscale : -gscale) * m);
      }
    }
    acc0 = (double)a;
  } else {
    // ---- focal loss sweep ---------------------------------------------------------------------
    const float npos = s_norm[0];
    const float gscale = -(npos == 0.f ? w_hm : w_hm / npos);  // :365-368
    const size_t n4 = (size_t)B * C * HW / 4;      // < 2^31 (checked on the host)
    const size_t stride = (size_t)(gridDim.x - n_reg_ctas) * kLossThreads;
    size_t i = (size_t)(blockIdx.x - n_reg_ctas) * kLossThreads + tid;
    for (; i < n4; i += stride * kUnroll) {
      float4 x[kUnroll], g[kUnroll];
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t j = i + u * stride;
        if (j < n4) { x[u] = ldg_stream4(hm + 4 * j); g[u] = ldg_stream4(gt + 4 * j); }
      }
      float pos = 0.f, neg = 0.f;
#pragma unroll
      for (int u = 0; u < kUnroll; ++u) {
        const size_t j = i + u * stride;
        if (j < n4) {
          const uint32_t plane = div_hw4.div((uint32_t)j);           // b*C + c
          const float al = __ldg(alpha + (plane - div_c.div(plane) * (uint32_t)C));
          float4 d;
          d.x = gscale * cell<kProb>(x[u].x, g[u].x, al, pos, neg);
          d.y = gscale * cell<kProb>(x[u].y, g[u].y, al, pos, neg);
          d.z = gscale * cell<kProb>(x[u].z, g[u].z, al, pos, neg);
          d.w = gscale * cell<kProb>(x[u].w, g[u].w, al, pos, neg);
          if (d_hm != nullptr) stg_stream4(d_hm + 4 * j, d);
        }
      }
      acc0 += (double)pos;
      acc1 += (double)neg;
    }
  }

  // ---- CTA reduction -> fp64 atomics ----------------------------------------------------------
  acc0 = warp_sum(acc0);
  acc1 = warp_sum(acc1);
  if (lane == 0) { s_red[0][warp] = acc0; s_red[1][warp] = acc1; }
  __syncthreads();
  if (tid == 0) {
    double a0 = 0.0, a1 = 0.0;
    for (int w = 0; w < kLossThreads / 32; ++w) { a0 += s_red[0][w]; a1 += s_red[1][w]; }
    if (is_reg) {
      atomicAdd((blockIdx.x & 1) ? &ws->off_sum : &ws->wh_sum, a0);
      if (blockIdx.x == 0) ws->nmask = s_norm[1];
    } else {
      if (a0 != 0.0) atomicAdd(&ws->pos_sum, a0);
      atomicAdd(&ws->neg_sum, a1);
    }
    __threadfence();
    s_last = (atomicAdd(&ws->done, 1u) == gridDim.x - 1);
  }
  __syncthreads();
  if (s_last && tid == 0) {
    __threadfence();
    const double pos = *((volatile double*)&ws->pos_sum), neg = *((volatile double*)&ws->neg_sum);
    const double whs = *((volatile double*)&ws->wh_sum), offs = *((volatile double*)&ws->off_sum);
    const float npos = s_norm[0];
    const float nmask = *((volatile float*)&ws->nmask);
    const float denom = 2.f * nmask + 1e-4f;
    const float hm_loss = (npos == 0.f) ? (float)(-neg) : (float)(-(pos + neg) / (double)npos);
    losses[0] = hm_loss * w_hm;
    losses[1] = ((float)whs / denom) * w_wh;
    losses[2] = ((float)offs / denom) * w_off;
    ws->pos_sum = 0.0; ws->neg_sum = 0.0; ws->wh_sum = 0.0; ws->off_sum = 0.0;
    ws->npos_count = 0ull; ws->done = 0u; ws->nmask = 0.f;
  }
}

}  // namespace ctdet

extern "C" size_t ctdet_loss_workspace_bytes(void) { return sizeof(ctdet::LossWs); }

// In-place d *= *g, skipped entirely (early exit) when *g == 1: lets autograd's backward honour an
// arbitrary upstream gradient without a host sync and without a pass in the common case.
__global__ void scale_grad_kernel(float* __restrict__ d, size_t n, const float* __restrict__ g) {
  const float s = *g;
  if (s == 1.f) return;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
    d[i] *= s;
}

extern "C" int ctdet_scale_grad(float* d, size_t n, const float* g, ctdet_stream_t stream) {
  using namespace ctdet;
  if (n == 0) return CTDET_OK;
  if (!d || !g) return CTDET_ENULL;
  size_t blocks = (n + 1023) / 1024;
  if (blocks > (size_t)kNumSMs * 8) blocks = (size_t)kNumSMs * 8;
  scale_grad_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(d, n, g);
  return check_launch();
}

extern "C" int ctdet_loss_fwd_bwd(const float* hm, const float* wh, const float* reg,
                                  const float* gt_hm, const uint8_t* reg_mask, const int64_t* ind,
                                  const float* gt_wh, const float* gt_reg, const float* alpha,
                                  const int32_t* num_pos_img, const float* norm_override,
                                  int B, int C, int H, int W, int M, int flags,
                                  float w_hm, float w_wh, float w_off,
                                  float* losses, float* d_hm, float* d_wh, float* d_reg,
                                  void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || M <= 0) return CTDET_EINVAL;
  if (flags & ~CTDET_LOSS_INPUT_PROB) return CTDET_EINVAL;
  if (!losses || !workspace) return CTDET_ENULL;
  if (hm && (!gt_hm || !alpha)) return CTDET_ENULL;
  if ((wh || reg) && (!reg_mask || !ind)) return CTDET_ENULL;
  if ((wh && !gt_wh) || (reg && !gt_reg)) return CTDET_ENULL;
  if (!hm && !wh && !reg) return CTDET_EINVAL;
  if (workspace_bytes < sizeof(LossWs)) return CTDET_EWORKSPACE;
  const int64_t HW = (int64_t)H * W;
  if (HW % 4 != 0) return CTDET_EUNSUPPORTED;    // float4 chunks must not straddle class planes
  if (HW >= ((int64_t)1 << 30) || (int64_t)B * M >= ((int64_t)1 << 30)) return CTDET_EINVAL;
  if ((hm && !aligned16(hm)) || (gt_hm && !aligned16(gt_hm)) || (d_hm && !aligned16(d_hm)) ||
      (reinterpret_cast<uintptr_t>(workspace) & 7u))
    return CTDET_EALIGN;
  cudaStream_t st = (cudaStream_t)stream;
  LossWs* ws = reinterpret_cast<LossWs*>(workspace);
  const size_t n4 = hm ? (size_t)B * C * HW / 4 : 0;
  if (n4 >= ((size_t)1 << 31)) return CTDET_EUNSUPPORTED;
  if (hm && num_pos_img == nullptr && norm_override == nullptr) {
    count_pos_kernel<<<kNumSMs * 8, kLossThreads, 0, st>>>(gt_hm, n4, ws);
    int rc = check_launch();
    if (rc) return rc;
  }
  // persistent sweep: as many CTAs as are co-resident (148 SMs x occupancy); each thread keeps
  // 2 x kUnroll 16-byte loads in flight
  static int occ = 0;
  if (occ == 0) {
    int o = 0;
    if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, loss_kernel<false>, kLossThreads, 0) != cudaSuccess || o < 1) o = 3;
    occ = o;
  }
  size_t want = (n4 + (size_t)kLossThreads * kUnroll - 1) / ((size_t)kLossThreads * kUnroll);
  size_t cap = (size_t)kNumSMs * occ;
  int sweep = (int)(want < cap ? want : cap);
  const int n_reg = (wh || reg) ? 2 * B : 0;
  if (n_reg + sweep < 1) return CTDET_EINVAL;
  auto kern = (flags & CTDET_LOSS_INPUT_PROB) ? loss_kernel<true> : loss_kernel<false>;
  kern<<<n_reg + sweep, kLossThreads, 0, st>>>(
      hm, wh, reg, gt_hm, reg_mask, ind, gt_wh, gt_reg, alpha, num_pos_img, norm_override,
      B, C, (int)HW, M, n_reg, FastDiv((uint32_t)(HW / 4)), FastDiv((uint32_t)C), w_hm, w_wh, w_off,
      losses, d_hm, d_wh, d_reg, ws);
  return check_launch();
}
