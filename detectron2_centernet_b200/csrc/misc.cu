// Small operators of the reference surface that sit either side of the hot kernels, plus the
// library's version / error plumbing.
#include "common.cuh"

namespace ctdet {

static thread_local char g_last_cuda_error[256] = "";

void set_last_cuda_error(cudaError_t e) {
  const char* s = cudaGetErrorString(e);
  int i = 0;
  for (; s && s[i] && i < 255; ++i) g_last_cuda_error[i] = s[i];
  g_last_cuda_error[i] = 0;
}

// Free-text detail for an error code that has no CUDA error behind it (read back through ctdet_last_cuda_error).
void set_last_error_text(const char* s) {
  int i = 0;
  for (; s && s[i] && i < 255; ++i) g_last_cuda_error[i] = s[i];
  g_last_cuda_error[i] = 0;
}

// heat*keep of _nms (centernet.py:399-405), dense output.  One thread per cell; the 3x3
// neighbourhood comes through L1/L2 (this is the API-compatible standalone op, not the hot path:
// ctdet_decode fuses the peak test and never materialises this tensor).
__global__ void nms_kernel(const float* __restrict__ heat, int H, int W, size_t total,
                           float* __restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int x = (int)(i % W);
  const int y = (int)((i / W) % H);
  const float v = heat[i];
  float m = v;
#pragma unroll
  for (int dy = -1; dy <= 1; ++dy) {
    const int yy = y + dy;
    if (yy < 0 || yy >= H) continue;
#pragma unroll
    for (int dx = -1; dx <= 1; ++dx) {
      const int xx = x + dx;
      if (xx < 0 || xx >= W) continue;
      m = fmaxf(m, __ldg(heat + i + (ptrdiff_t)dy * W + dx));
    }
  }
  out[i] = (m == v) ? v : v * 0.f;      // heat * keep.float()
}

// out[b,k,c] = feat[b,c,ind[b,k]]  (_transpose_and_gather_feat, centernet.py:383-397)
__global__ void gather_feat_kernel(const float* __restrict__ feat, const int64_t* __restrict__ ind,
                                   int Ch, int HW, int K, size_t total, float* __restrict__ out) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= total) return;
  const int c = (int)(i % Ch);
  const size_t bk = i / Ch;
  const size_t b = bk / K;
  const int64_t sp = ind[bk];
  out[i] = (sp >= 0 && sp < HW) ? feat[(b * Ch + c) * (size_t)HW + sp] : 0.f;
}

// inference_single_image (centernet.py:251-261) + detector_postprocess
// (modeling/postprocessing.py:10-74), one warp-sized CTA group per image; order preserving.
__global__ void postprocess_kernel(float* __restrict__ boxes, float* __restrict__ scores,
                                   int32_t* __restrict__ classes, const int32_t* __restrict__ sizes,
                                   int K, int max_dets, float thresh, int32_t* __restrict__ n_keep) {
  extern __shared__ float s_box[];                  // [K][4] scaled + clipped, then flags
  __shared__ int s_pos[CTDET_MAX_K + 1];
  const int b = blockIdx.x, tid = threadIdx.x;
  const int in_h = sizes[4 * b], in_w = sizes[4 * b + 1];
  const int out_h = sizes[4 * b + 2], out_w = sizes[4 * b + 3];
  const float sx = (float)((double)out_w / (double)max(in_w, 1));   // postprocessing.py:39-42: python-float
  const float sy = (float)((double)out_h / (double)max(in_h, 1));   // quotient, applied to an fp32 tensor
  const int n = min(max_dets, K);                   // centernet.py:252
  float* bx = boxes + (size_t)b * K * 4;
  float* sc = scores + (size_t)b * K;
  int32_t* cl = classes + (size_t)b * K;
  // out_h <= 0: CenterNet.inference alone (centernet.py:214-266) -- slice + threshold, the boxes stay as decoded;
  // detector_postprocess (scale, clip, nonempty) is what the reference's forward applies afterwards (:169-171)
  const bool resize = out_h > 0 && out_w > 0;
  for (int k = tid; k < K; k += blockDim.x) {
    float x1 = bx[4 * k], y1 = bx[4 * k + 1], x2 = bx[4 * k + 2], y2 = bx[4 * k + 3];
    bool keep = (k < n) && (sc[k] > thresh);                      // strict > (centernet.py:258)
    if (resize) {
      x1 *= sx; y1 *= sy; x2 *= sx; y2 *= sy;                     // Boxes.scale (boxes.py:271)
      x1 = fminf(fmaxf(x1, 0.f), (float)out_w); x2 = fminf(fmaxf(x2, 0.f), (float)out_w);   // Boxes.clip (:184)
      y1 = fminf(fmaxf(y1, 0.f), (float)out_h); y2 = fminf(fmaxf(y2, 0.f), (float)out_h);
      keep = keep && ((x2 - x1) > 0.f) && ((y2 - y1) > 0.f);      // Boxes.nonempty (:199)
    }
    s_box[4 * k] = x1; s_box[4 * k + 1] = y1; s_box[4 * k + 2] = x2; s_box[4 * k + 3] = y2;
    s_pos[k] = keep ? 1 : 0;
  }
  __syncthreads();
  if (tid == 0) {                                   // K <= 1024: a serial scan is negligible
    int run = 0;
    for (int k = 0; k < K; ++k) { const int f = s_pos[k]; s_pos[k] = f ? run : -1; run += f; }
    s_pos[K] = run;
    n_keep[b] = run;
  }
  __syncthreads();
  // read everything that moves before anything is overwritten
  for (int base = 0; base < K; base += blockDim.x) {
    const int k = base + tid;
    float s = 0.f; int c = 0; int dst = -1;
    if (k < K) { dst = s_pos[k]; s = sc[k]; c = cl[k]; }
    __syncthreads();
    if (dst >= 0) {
      sc[dst] = s; cl[dst] = c;
      bx[4 * dst] = s_box[4 * k]; bx[4 * dst + 1] = s_box[4 * k + 1];
      bx[4 * dst + 2] = s_box[4 * k + 2]; bx[4 * dst + 3] = s_box[4 * k + 3];
    }
    __syncthreads();
  }
}

// CenterNet.preprocess_image (meta_arch/centernet.py:177-184) + ImageList.from_tensors
// (structures/image_list.py:58-130) for the whole batch in one pass: x / 255, (x - mean) / std per
// channel in the reference's fp32 operation order (IEEE division, no FMA contraction), written into
// the zero-padded batch tensor.  HBM-bound: 1 or 4 B read + 4 B written per pixel and channel; the
// padding is written by the same launch (no separate fill).  One thread = 4 consecutive x of one
// (image, channel) plane: grid = (plane chunks, C, B).
template <typename T>
__global__ void preprocess_kernel(const T* const* __restrict__ images, const int32_t* __restrict__ sizes_hw,
                                  int C, int Hp, int Wp4, FastDiv div_wp4, const float* __restrict__ mean,
                                  const float* __restrict__ stdv, float* __restrict__ out) {
  const int c = blockIdx.y, b = blockIdx.z;
  // uint8 input: a pixel can take 256 values, so the CTA (256 threads, one channel) tabulates the
  // whole expression once with the IEEE operations and the per-pixel work is one shared-memory lookup
  __shared__ float lut[256];
  if (sizeof(T) == 1) {
    lut[threadIdx.x] = __fdiv_rn(__fsub_rn(__fdiv_rn((float)threadIdx.x, 255.f), __ldg(mean + c)), __ldg(stdv + c));
    __syncthreads();
  }
  const int idx = blockIdx.x * blockDim.x + threadIdx.x;                // float4 index inside the padded plane
  if (idx >= Hp * Wp4) return;
  const int y = (int)div_wp4.div((uint32_t)idx), x4 = idx - y * Wp4;
  const int h = __ldg(sizes_hw + 2 * b), w = __ldg(sizes_hw + 2 * b + 1);
  float v[4] = {0.f, 0.f, 0.f, 0.f};                         // pad_value of from_tensors
  const int x = x4 * 4;
  if (y < h && x < w) {
    const T* src = images[b] + ((size_t)c * h + y) * w + x;
    const float m = __ldg(mean + c), sd = __ldg(stdv + c);
    float r[4] = {0.f, 0.f, 0.f, 0.f};
    if (x + 4 <= w && (reinterpret_cast<uintptr_t>(src) & (4 * sizeof(T) - 1)) == 0) {        // one aligned 4-pixel load
      if (sizeof(T) == 1) {
        const uchar4 q = *reinterpret_cast<const uchar4*>(src);
        v[0] = lut[q.x]; v[1] = lut[q.y]; v[2] = lut[q.z]; v[3] = lut[q.w];
      } else {
        const float4 q = *reinterpret_cast<const float4*>(src);
        r[0] = q.x; r[1] = q.y; r[2] = q.z; r[3] = q.w;
      }
      if (sizeof(T) != 1) {
#pragma unroll
        for (int k = 0; k < 4; ++k) v[k] = __fdiv_rn(__fsub_rn(__fdiv_rn(r[k], 255.f), m), sd);
      }
    } else {
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (x + k < w) v[k] = sizeof(T) == 1 ? lut[(int)src[k]] : __fdiv_rn(__fsub_rn(__fdiv_rn((float)src[k], 255.f), m), sd);
    }
  }
  *reinterpret_cast<float4*>(out + (((size_t)b * C + c) * Hp + y) * (size_t)(Wp4 * 4) + x) = make_float4(v[0], v[1], v[2], v[3]);
}


// ------------------------------------------------------------------------------------------------
// IDAUp up-sampling + skip add: out = ConvTranspose2d(o, o, 2f, stride=f, padding=f/2, groups=o,
// bias=False)(x) + skip   (dla.py:166-168 builds `up_i`, :175-177 applies it and adds the previous level).
// Depth-wise, so every output cell is 2 x 2 input cells of its own channel times 4 of the channel's (2f)^2
// weights: out[Y,X] = sum over ky = (Y+p) mod f (+f), iy = (Y+p-ky)/f and the same in x.  HBM-bound: x is read
// once (neighbouring outputs share it through L1), skip read once, out written once -- the reference runs a
// grouped cuDNN dgrad kernel (5.7 ms per DLA-34 step at batch 32) plus a separate add.
// One thread = 4 consecutive outputs of a row (float4 skip load / out store); a CTA stays inside one (b, c) plane
// and keeps the plane's weights in shared memory.
// LF > 0: f = 2^LF known at compile time (the neck's f = 2 and 4).  With 4 outputs per thread X0 = 4q is a multiple
// of f, so the tap kx0 and the input column of output j depend on j only: the weight offsets and the column
// selects below become constants (the capture of the runtime-f form showed the kernel issue-bound, 67
// instructions per thread, most of them those selects and the row division).  LF == 0: runtime f (8, 16, odd widths).
template <int VEC, int LF>
__global__ void __launch_bounds__(256)
upsample_add_kernel(const float* __restrict__ x, const float* __restrict__ wt, const float* __restrict__ skip,
                    int C, int h, int w, int lf_rt, FastDiv div_wq, float* __restrict__ out) {
  __shared__ float s_w[32 * 32];                   // (2f)^2 <= 1024 (f <= 16)
  const int lf = LF > 0 ? LF : lf_rt;
  const int f = 1 << lf, k = 2 * f, p = f >> 1;
  const int c = blockIdx.y, b = blockIdx.z;
  const int Ho = h << lf, Wo = w << lf, Wq = (Wo + VEC - 1) / VEC;
  for (int i = threadIdx.x; i < k * k; i += blockDim.x) s_w[i] = wt[(size_t)c * k * k + i];
  __syncthreads();
  const size_t plane = (size_t)b * C + c;
  const float* xp = x + plane * h * w;
  const float* sp = skip ? skip + plane * Ho * Wo : nullptr;
  float* op = out + plane * Ho * Wo;
  for (int item = blockIdx.x * blockDim.x + threadIdx.x; item < Ho * Wq; item += gridDim.x * blockDim.x) {
    const int Y = (int)div_wq.div((uint32_t)item), q = item - Y * Wq, X0 = q * VEC;
    const int ty = Y + p, ky0 = ty & (f - 1), iy0 = ty >> lf, iy1 = iy0 - 1;
    const float* w0 = s_w + ky0 * k;               // weights of input row iy0
    const float* w1 = s_w + (ky0 + f) * k;         //            of input row iy1
    const bool r0 = iy0 < h, r1 = iy1 >= 0;
    // the VEC outputs read input columns c_lo .. c_lo+3 (f >= 2): load them once for both rows
    const int c_lo = ((X0 + p) >> lf) - 1;
    float a0[4], a1[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int ix = c_lo + j;
      const bool ok = ix >= 0 && ix < w;
      a0[j] = (ok && r0) ? __ldg(xp + (size_t)iy0 * w + ix) : 0.f;
      a1[j] = (ok && r1) ? __ldg(xp + (size_t)iy1 * w + ix) : 0.f;
    }
    float res[VEC];
#pragma unroll
    for (int j = 0; j < VEC; ++j) {
      int kx0, d;                                                              // d in {1, 2, 3}
      if (LF > 0 && VEC == 4 && (1 << LF) <= 4) {          // X0 % f == 0: both depend on j only (compile-time)
        kx0 = (p + j) & (f - 1); d = 1 + ((p + j) >> lf);
      } else {
        const int tx = X0 + j + p;
        kx0 = tx & (f - 1); d = (tx >> lf) - c_lo;
      }
      const float i00 = d == 1 ? a0[1] : (d == 2 ? a0[2] : a0[3]);               // (iy0, ix0)
      const float i01 = d == 1 ? a0[0] : (d == 2 ? a0[1] : a0[2]);               // (iy0, ix0 - 1)
      const float i10 = d == 1 ? a1[1] : (d == 2 ? a1[2] : a1[3]);
      const float i11 = d == 1 ? a1[0] : (d == 2 ? a1[1] : a1[2]);
      res[j] = w0[kx0] * i00 + w0[kx0 + f] * i01 + w1[kx0] * i10 + w1[kx0 + f] * i11;
    }
    const size_t o = (size_t)Y * Wo + X0;
    if (VEC == 4) {
      float4 r = make_float4(res[0], res[1], res[2], res[3]);
      if (sp) { const float4 s4 = ldg_stream4(sp + o); r.x += s4.x; r.y += s4.y; r.z += s4.z; r.w += s4.w; }
      *reinterpret_cast<float4*>(op + o) = r;
    } else {
      op[o] = sp ? res[0] + __ldg(sp + o) : res[0];
    }
  }
}

}  // namespace ctdet

extern "C" int ctdet_abi_version(void) { return CTDET_ABI_VERSION; }

extern "C" const char* ctdet_error_string(int code) {
  switch (code) {
    case CTDET_OK: return "ok";
    case CTDET_EINVAL: return "invalid shape, size or flag";
    case CTDET_ENULL: return "required pointer is NULL";
    case CTDET_EALIGN: return "pointer is not aligned as required";
    case CTDET_EWORKSPACE: return "workspace too small";
    case CTDET_ECUDA: return "CUDA error";
    case CTDET_EUNSUPPORTED: return "unsupported configuration";
    default: return "unknown error code";
  }
}

extern "C" const char* ctdet_last_cuda_error(void) { return ctdet::g_last_cuda_error; }

extern "C" int ctdet_nms(const float* heat, int B, int C, int H, int W, float* out,
                         ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || C <= 0 || H <= 0 || W <= 0) return CTDET_EINVAL;
  if (B == 0) return CTDET_OK;
  if (!heat || !out) return CTDET_ENULL;
  const size_t total = (size_t)B * C * H * W;
  const size_t blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffull) return CTDET_EINVAL;
  nms_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(heat, H, W, total, out);
  return check_launch();
}

extern "C" int ctdet_gather_feat(const float* feat, const int64_t* ind, int B, int Ch, int HW, int K,
                                 float* out, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || Ch <= 0 || HW <= 0 || K < 0) return CTDET_EINVAL;
  const size_t total = (size_t)B * K * Ch;
  if (total == 0) return CTDET_OK;
  if (!feat || !ind || !out) return CTDET_ENULL;
  const size_t blocks = (total + 255) / 256;
  if (blocks > 0x7fffffffull) return CTDET_EINVAL;
  gather_feat_kernel<<<(unsigned)blocks, 256, 0, (cudaStream_t)stream>>>(feat, ind, Ch, HW, K, total, out);
  return check_launch();
}

extern "C" int ctdet_postprocess(float* boxes, float* scores, int32_t* classes, const int32_t* sizes,
                                 int B, int K, int max_dets, float score_thresh, int32_t* n_keep,
                                 ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || K <= 0 || K > CTDET_MAX_K || max_dets < 0) return CTDET_EINVAL;
  if (B == 0) return CTDET_OK;
  if (!boxes || !scores || !classes || !sizes || !n_keep) return CTDET_ENULL;
  postprocess_kernel<<<B, 128, (size_t)K * 4 * sizeof(float), (cudaStream_t)stream>>>(
      boxes, scores, classes, sizes, K, max_dets, score_thresh, n_keep);
  return check_launch();
}

extern "C" int ctdet_preprocess(const void* const* images, const int32_t* sizes_hw, int dtype,
                                int B, int C, int Hp, int Wp, const float* mean, const float* stdv,
                                float* out, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || C <= 0 || Hp <= 0 || Wp <= 0 || (dtype != CTDET_IMAGE_U8 && dtype != CTDET_IMAGE_F32)) return CTDET_EINVAL;
  if (B == 0) return CTDET_OK;
  if (!images || !sizes_hw || !mean || !stdv || !out) return CTDET_ENULL;
  if (Wp % 4 != 0) return CTDET_EUNSUPPORTED;              // size_divisibility is 16 / 32 in every config
  if (!aligned16(out)) return CTDET_EALIGN;
  if (B > 65535 || C > 65535 || (long long)Hp * (Wp / 4) >= (1ll << 31)) return CTDET_EINVAL;
  const int Wp4 = Wp / 4;
  const dim3 grid((unsigned)((Hp * Wp4 + 255) / 256), (unsigned)C, (unsigned)B);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == CTDET_IMAGE_U8)
    preprocess_kernel<uint8_t><<<grid, 256, 0, st>>>(reinterpret_cast<const uint8_t* const*>(images), sizes_hw, C, Hp, Wp4,
                                                    FastDiv((uint32_t)Wp4), mean, stdv, out);
  else
    preprocess_kernel<float><<<grid, 256, 0, st>>>(reinterpret_cast<const float* const*>(images), sizes_hw, C, Hp, Wp4,
                                                  FastDiv((uint32_t)Wp4), mean, stdv, out);
  return check_launch();
}

// Backward of the IDAUp up-sampling + skip add (out = ConvTranspose2d_depthwise(x, wt) + skip, dla.py:166-177):
//   grad_skip = grad_out (the caller aliases it: no pass)
//   grad_x[c,i,j]  = sum_{u,v} grad_out[c, f*i - f/2 + u, f*j - f/2 + v] * wt[c,u,v]        (u, v < 2f)
//   grad_wt[c,u,v] = sum_{b,i,j} x[b,c,i,j] * grad_out[b,c, f*i - f/2 + u, f*j - f/2 + v]
// One pass over grad_out: a thread owns (input cell, kernel row u) -- the 2f lanes of a cell sit next to each other,
// so grad_x is a shuffle reduction over them -- and keeps the 2f products of its row as running grad_wt sums over
// every cell it visits; per CTA they meet in shared memory and leave as ONE atomic per (u, v).  grid = (chunks of
// the plane, C, B).  HBM-bound: grad_out is read once from DRAM (each element 4x through L1), x once, grad_x written.
template <int F>
__global__ void __launch_bounds__(256)
upsample_add_bwd_kernel(const float* __restrict__ gout, const float* __restrict__ x, const float* __restrict__ wt,
                        int C, int h, int w, int cells_per_cta, float* __restrict__ gx, float* __restrict__ gw) {
  constexpr int K = 2 * F;                       // kernel side
  constexpr int CELLS = 256 / K;                 // input cells in flight per CTA
  __shared__ float s_w[K * K], s_gw[K * K];
  const int c = blockIdx.y, b = blockIdx.z, tid = threadIdx.x;
  const int u = tid % K, slot = tid / K;
  for (int i = tid; i < K * K; i += 256) { s_w[i] = __ldg(wt + (size_t)c * K * K + i); s_gw[i] = 0.f; }
  __syncthreads();
  float wrow[K], acc[K];
#pragma unroll
  for (int v = 0; v < K; ++v) { wrow[v] = s_w[u * K + v]; acc[v] = 0.f; }
  const int Ho = h * F, Wo = w * F, n_cells = h * w;
  const float* go = gout + ((size_t)b * C + c) * Ho * Wo;
  const float* xp = x ? x + ((size_t)b * C + c) * n_cells : nullptr;
  float* gxp = gx ? gx + ((size_t)b * C + c) * n_cells : nullptr;
  const int cell0 = blockIdx.x * cells_per_cta, cell1 = min(n_cells, cell0 + cells_per_cta);
  for (int base = cell0; base < cell1; base += CELLS) {
    const int cell = base + slot;
    const bool ok = cell < cell1;                // (the K lanes of a cell agree)
    const int i = ok ? cell / w : 0, j = ok ? cell - i * w : 0;
    const int Y = F * i - F / 2 + u, X0 = F * j - F / 2;
    const bool row_ok = ok && Y >= 0 && Y < Ho;
    const float xv = (ok && xp) ? __ldg(xp + cell) : 0.f;
    float part = 0.f;
#pragma unroll
    for (int v = 0; v < K; ++v) {
      const int X = X0 + v;
      const float g = (row_ok && X >= 0 && X < Wo) ? __ldg(go + (size_t)Y * Wo + X) : 0.f;
      part = fmaf(g, wrow[v], part);
      acc[v] = fmaf(g, xv, acc[v]);
    }
#pragma unroll
    for (int o = K / 2; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);    // over the cell's K lanes
    if (ok && u == 0 && gxp) gxp[cell] = part;
  }
  if (gw) {
#pragma unroll
    for (int v = 0; v < K; ++v) atomicAdd(&s_gw[u * K + v], acc[v]);
    __syncthreads();
    for (int i = tid; i < K * K; i += 256) atomicAdd(gw + (size_t)c * K * K + i, s_gw[i]);
  }
}

extern "C" int ctdet_upsample_add(const float* x, const float* weight, const float* skip, int B, int C, int h, int w,
                                  int f, float* out, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || C <= 0 || h <= 0 || w <= 0 || f < 2) return CTDET_EINVAL;
  if (f > 16 || (f & (f - 1))) return CTDET_EUNSUPPORTED;  // IDAUp uses 2, 4, 8
  if (B == 0) return CTDET_OK;
  if (!x || !weight || !out) return CTDET_ENULL;
  if (B > 65535 || C > 65535 || (long long)h * f * w * f >= (1ll << 31)) return CTDET_EINVAL;
  int lf = 0; while ((1 << lf) < f) ++lf;
  const int Ho = h * f, Wo = w * f;
  const bool vec = (Wo % 4) == 0 && aligned16(out) && (!skip || aligned16(skip));
  const int items = Ho * (vec ? Wo / 4 : Wo);
  int bx = (items + 256 * 4 - 1) / (256 * 4);              // ~4 items per thread
  if (bx < 1) bx = 1;
  const dim3 grid((unsigned)bx, (unsigned)C, (unsigned)B);
  cudaStream_t st = (cudaStream_t)stream;
  const FastDiv dq((uint32_t)(vec ? Wo / 4 : Wo));
  if (vec && f == 2) upsample_add_kernel<4, 1><<<grid, 256, 0, st>>>(x, weight, skip, C, h, w, lf, dq, out);
  else if (vec && f == 4) upsample_add_kernel<4, 2><<<grid, 256, 0, st>>>(x, weight, skip, C, h, w, lf, dq, out);
  else if (vec) upsample_add_kernel<4, 0><<<grid, 256, 0, st>>>(x, weight, skip, C, h, w, lf, dq, out);
  else upsample_add_kernel<1, 0><<<grid, 256, 0, st>>>(x, weight, skip, C, h, w, lf, dq, out);
  return check_launch();
}

extern "C" int ctdet_upsample_add_backward(const float* grad_out, const float* x, const float* weight, int B, int C, int h,
                                           int w, int f, float* grad_x, float* grad_weight, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || C <= 0 || h <= 0 || w <= 0) return CTDET_EINVAL;
  if (f != 2 && f != 4) return CTDET_EUNSUPPORTED;        // IDAUp in DLA-34 uses 2 and 4 (8 only with a deeper base)
  if (!grad_out || !weight || (!grad_x && !grad_weight) || (grad_weight && !x)) return CTDET_ENULL;
  if (B > 65535 || C > 65535 || (long long)h * f * w * f >= (1ll << 31)) return CTDET_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (grad_weight) {                                       // accumulated with atomics: zero first (same stream)
    cudaError_t e = cudaMemsetAsync(grad_weight, 0, (size_t)C * 4 * f * f * sizeof(float), st);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  }
  if (B == 0) return CTDET_OK;
  const int n_cells = h * w;
  int per = 2048;                                          // input cells per CTA: >= 32 visits of every thread
  if (per > n_cells) per = n_cells;
  const dim3 grid((unsigned)((n_cells + per - 1) / per), (unsigned)C, (unsigned)B);
  if (f == 2) upsample_add_bwd_kernel<2><<<grid, 256, 0, st>>>(grad_out, x, weight, C, h, w, per, grad_x, grad_weight);
  else upsample_add_bwd_kernel<4><<<grid, 256, 0, st>>>(grad_out, x, weight, C, h, w, per, grad_x, grad_weight);
  return check_launch();
}
