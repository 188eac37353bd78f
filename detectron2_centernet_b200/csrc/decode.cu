// ctdet decode: 3x3 peak NMS + exact top-K over C*H*W + wh/reg gather + box assembly, batched.
//
// Replaces ctdet_decode (detectron2/modeling/meta_arch/centernet.py:426-458): _nms (:399-405,
// max_pool2d + == + mul: three full passes), _topk (:408-424, two torch.topk + three gathers), two
// full-map NHWC permutes (:435,:442) and the per-image Python loop of CenterNet.inference (:224).
//
// Kernel 1 (HBM-bound, reads every heatmap cell exactly once = 4 B/cell):
//   one CTA per (image, class plane, row band).  Warps walk down the rows of a 128-column strip
//   with float4 loads kept in registers (3-row sliding window, horizontal neighbours by warp
//   shuffle), so the 3x3 peak test needs no shared-memory staging.  Peaks are packed into 64-bit
//   keys  (ordered(score) << 32) | ~flat_index  -- one integer compare implements the tie rule
//   "higher score, then lower flat index" -- and appended to a shared-memory candidate buffer.
//   When the buffer fills, an exact MSB-first radix select prunes it to the K best and the K-th
//   key becomes the admission threshold (streaming top-K).  Each CTA emits its K best keys.
// Kernel 2 (one CTA per image): exact radix select of the K best among the emitted keys
//   (equivalent to the reference's per-class top-K followed by a top-K over C*K), bitonic sort,
//   gather of wh/reg, boxes, outputs.
#include "common.cuh"

namespace ctdet {

constexpr int kDecThreads = 128;                 // kernel 1: 4 warps
constexpr int kDecWarps = kDecThreads / 32;
constexpr int kRowsPerBatch = 4;                 // rows a warp loads before a CTA-wide check
constexpr int kBatchMax = kDecWarps * kRowsPerBatch * 128;   // most keys one batch can append
constexpr int kFinThreads = 1024;                // kernel 2

typedef unsigned long long u64;

struct SelectScratch {
  unsigned int hist[256];
  u64 prefix;       // selected high bits so far
  int need;         // how many still to take from the current bucket
  int bucket_count;
  int digit;
};

// Exact "k-th largest" of n distinct 64-bit keys: returns T such that exactly min(n,k) keys are
// >= T.  MSB-first radix select, 8 bits per pass; stops as soon as the bucket holding the k-th
// key must be taken whole.  `keyat(i)` yields key i.  All threads of the CTA must call it.
template <int kThreads, typename KeyAt>
__device__ u64 radix_select_kth(KeyAt keyat, int n, int k, SelectScratch* sc) {
  const int tid = threadIdx.x;
  if (n <= k) return 0ull;
  u64 prefix = 0ull, mask = 0ull;
  int need = k;
  for (int shift = 56; shift >= 0; shift -= 8) {
    for (int i = tid; i < 256; i += kThreads) sc->hist[i] = 0u;
    __syncthreads();
    for (int i = tid; i < n; i += kThreads) {
      const u64 key = keyat(i);
      if ((key & mask) == prefix) atomicAdd(&sc->hist[(unsigned)(key >> shift) & 255u], 1u);
    }
    __syncthreads();
    if (tid < 32) {
      // lane l owns digits 255-8l .. 248-8l (descending); find where the running count reaches need
      unsigned int loc[8];
      unsigned int sum = 0;
#pragma unroll
      for (int j = 0; j < 8; ++j) { loc[j] = sc->hist[255 - (8 * tid + j)]; sum += loc[j]; }
      unsigned int incl = sum;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
        if (tid >= o) incl += t;
      }
      const unsigned int excl = incl - sum;
      if (excl < (unsigned)need && incl >= (unsigned)need) {
        unsigned int run = excl;
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          if (run < (unsigned)need && run + loc[j] >= (unsigned)need) {
            sc->digit = 255 - (8 * tid + j);
            sc->need = need - (int)run;
            sc->bucket_count = (int)loc[j];
          }
          run += loc[j];
        }
      }
    }
    __syncthreads();
    prefix |= (u64)sc->digit << shift;
    mask |= 0xffull << shift;
    need = sc->need;
    const int bc = sc->bucket_count;
    __syncthreads();
    if (bc == need) break;     // the whole bucket is taken (always true once bc == 1)
  }
  return prefix;               // low (unselected) bits are zero
}

struct DecodeShared {
  SelectScratch sel;
  int count;        // keys in buf
  int kept;
  u64 thresh;       // admit keys > thresh
  u64 kept_min;
};

// Prunes buf[0..count) to its K largest keys (compacted to the front) and raises the threshold.
__device__ void prune_topk(u64* buf, u64* tmp, DecodeShared* sh, int K) {
  const int tid = threadIdx.x;
  const int n = sh->count;
  __syncthreads();
  if (n <= K) return;
  const u64 T = radix_select_kth<kDecThreads>([&](int i) { return buf[i]; }, n, K, &sh->sel);
  if (tid == 0) { sh->kept = 0; sh->kept_min = ~0ull; }
  __syncthreads();
  for (int i = tid; i < n; i += kDecThreads) {
    const u64 key = buf[i];
    if (key >= T) {
      tmp[atomicAdd(&sh->kept, 1)] = key;
      atomicMin(&sh->kept_min, key);
    }
  }
  __syncthreads();
  for (int i = tid; i < K; i += kDecThreads) buf[i] = tmp[i];
  if (tid == 0) { sh->count = K; sh->thresh = sh->kept_min; }
  __syncthreads();
}

// heat [B,C,H,W]; out_keys [B, tiles_per_image, K] (0 = empty slot)
__global__ void __launch_bounds__(kDecThreads)
decode_peaks_kernel(const float* __restrict__ heat, int C, int H, int W, int K, int band_rows,
                    int bands, int flags, int cap, u64* __restrict__ out_keys) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  u64* buf = reinterpret_cast<u64*>(s_raw);             // [cap]
  u64* tmp = buf + cap;                                 // [K]
  __shared__ DecodeShared sh;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int band = blockIdx.x % bands;
  const int plane = blockIdx.x / bands;                 // b*C + c
  const int c = plane % C;
  const bool logits = flags & CTDET_DECODE_LOGITS;
  const bool no_nms = flags & CTDET_DECODE_NO_NMS;
  const int y_lo = band * band_rows, y_hi = min(H, y_lo + band_rows);
  const float* p = heat + (size_t)plane * H * W;
  const float NEG = -__int_as_float(0x7f800000);        // -inf: max_pool2d padding (:401)

  if (tid == 0) { sh.count = 0; sh.thresh = 0ull; }
  __syncthreads();

  // work items of this CTA: (128-column strip, sub-band of rows); a warp takes items round-robin
  const int strips = (W + 127) / 128;
  int subs = kDecWarps / strips; if (subs < 1) subs = 1;
  const int sub_rows = (y_hi - y_lo + subs - 1) / subs;
  const int items = strips * subs;
  const int rounds = (items + kDecWarps - 1) / kDecWarps;
  const int batches = (sub_rows + kRowsPerBatch - 1) / kRowsPerBatch;
  const int limit = cap - kBatchMax;
  const bool vec = (W % 4) == 0;

  for (int rnd = 0; rnd < rounds; ++rnd) {
    const int item = rnd * kDecWarps + warp;
    const bool active = item < items;
    const int strip = active ? item % strips : 0;
    const int sub = active ? item / strips : 0;
    const int r0 = y_lo + sub * sub_rows;
    const int r1 = active ? min(y_hi, r0 + sub_rows) : r0;      // rows [r0, r1)
    const int x0 = strip * 128 + lane * 4;                      // this lane's 4 columns

    // loads row y (4 values + left/right neighbours); out-of-map cells read as -inf
    auto load_row = [&](int y, float (&v)[4], float& lft, float& rgt) {
      v[0] = v[1] = v[2] = v[3] = NEG; lft = NEG; rgt = NEG;
      if (y < 0 || y >= H) return;
      const float* row = p + (size_t)y * W;
      if (vec) {
        if (x0 < W) { const float4 t = ldg_stream4(row + x0); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
      } else {
#pragma unroll
        for (int j = 0; j < 4; ++j) if (x0 + j < W) v[j] = __ldg(row + x0 + j);
      }
      // neighbours across the strip edge come from memory, inside the strip from the next lane
      if (lane == 0 && x0 > 0 && x0 - 1 < W) lft = __ldg(row + x0 - 1);
      if (lane == 31 && x0 + 4 < W) rgt = __ldg(row + x0 + 4);
    };
    // horizontal 3-max of a row
    auto hmax3 = [&](const float (&v)[4], float lft, float rgt, float (&m)[4]) {
      const float l = __shfl_up_sync(0xffffffffu, v[3], 1);
      const float r = __shfl_down_sync(0xffffffffu, v[0], 1);
      const float L = lane == 0 ? lft : l;
      const float R = lane == 31 ? rgt : r;
      m[0] = fmaxf(fmaxf(L, v[0]), v[1]);
      m[1] = fmaxf(fmaxf(v[0], v[1]), v[2]);
      m[2] = fmaxf(fmaxf(v[1], v[2]), v[3]);
      m[3] = fmaxf(fmaxf(v[2], v[3]), R);
    };

    float cur[4], hm_prev[4], hm_cur[4];                 // sliding window state
    {
      float t[4], l, r;
      load_row(active ? r0 - 1 : -1, t, l, r);  hmax3(t, l, r, hm_prev);
      load_row(active ? r0 : -1, cur, l, r);    hmax3(cur, l, r, hm_cur);
    }

    for (int bt = 0; bt < batches; ++bt) {
      const int yb = r0 + bt * kRowsPerBatch;
      float nxt[kRowsPerBatch][4], nl[kRowsPerBatch], nr[kRowsPerBatch];
#pragma unroll
      for (int q = 0; q < kRowsPerBatch; ++q) {
        const int y = yb + q;                            // row being tested; needs row y+1
        load_row((active && y < r1) ? y + 1 : -1, nxt[q], nl[q], nr[q]);
      }
      bool crossed = false;
#pragma unroll
      for (int q = 0; q < kRowsPerBatch; ++q) {
        const int y = yb + q;
        float hm_next[4];
        hmax3(nxt[q], nl[q], nr[q], hm_next);
        if (active && y < r1) {
          const u64 th = sh.thresh;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float v = cur[j];
            const float nb = fmaxf(fmaxf(hm_prev[j], hm_cur[j]), hm_next[j]);
            bool cand = (x0 + j < W) && (no_nms || v >= nb) && (logits || v > 0.f);
            u64 key = 0ull;
            if (cand) {
              const uint32_t flat = (uint32_t)((c * H + y) * W + x0 + j);
              key = ((u64)float_to_ordered(v) << 32) | (u64)(0xffffffffu - flat);
              cand = key > th;
            }
            const unsigned m = __ballot_sync(0xffffffffu, cand);
            if (m) {
              int base = 0;
              if (lane == 0) {
                base = atomicAdd(&sh.count, __popc(m));
                crossed |= (base + __popc(m) > limit);
              }
              base = __shfl_sync(0xffffffffu, base, 0);
              if (cand) buf[base + __popc(m & ((1u << lane) - 1u))] = key;
            }
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) { hm_prev[j] = hm_cur[j]; hm_cur[j] = hm_next[j]; cur[j] = nxt[q][j]; }
      }
      // CTA-uniform decision: did any warp push the buffer past the safe fill level?
      if (__syncthreads_or(crossed)) prune_topk(buf, tmp, &sh, K);
    }
  }

  // final: exact K best of this CTA, written as a fixed-size record (zero padded)
  prune_topk(buf, tmp, &sh, K);
  __syncthreads();
  const int n = min(sh.count, K);
  u64* dst = out_keys + (size_t)blockIdx.x * K;
  for (int i = tid; i < K; i += kDecThreads) dst[i] = i < n ? buf[i] : 0ull;
}

// ------------------------------------------------------------------------------------------------
// Kernel 2: per image, exact top-K of the emitted keys, sort, gather, boxes.
__global__ void __launch_bounds__(kFinThreads)
decode_finalize_kernel(const u64* __restrict__ keys, int n_keys, const float* __restrict__ wh,
                       const float* __restrict__ reg, int C, int H, int W, int K, float down_ratio,
                       int flags, float* __restrict__ boxes, float* __restrict__ scores,
                       int32_t* __restrict__ classes, int64_t* __restrict__ inds) {
  __shared__ SelectScratch sc;
  __shared__ u64 s_sel[CTDET_MAX_K];
  __shared__ int s_n, s_valid;
  const int tid = threadIdx.x;
  const int b = blockIdx.x;
  const u64* kb = keys + (size_t)b * n_keys;
  const int HW = H * W;

  // count non-empty keys (empty slots are 0 and sort last; they are never >= a non-zero T)
  if (tid == 0) { s_n = 0; s_valid = 0; }
  __syncthreads();
  int local = 0;
  for (int i = tid; i < n_keys; i += kFinThreads) local += (kb[i] != 0ull);
  local = warp_sum(local);
  if ((tid & 31) == 0 && local) atomicAdd(&s_valid, local);
  __syncthreads();
  const int n_valid = s_valid;
  u64 T = radix_select_kth<kFinThreads>([&](int i) { return kb[i]; }, n_keys, K, &sc);
  if (n_valid <= K) T = 1ull;                             // take every non-empty key
  for (int i = tid; i < n_keys; i += kFinThreads) {
    const u64 key = kb[i];
    if (key >= T) s_sel[atomicAdd(&s_n, 1)] = key;
  }
  __syncthreads();
  int n = s_n;                                            // == min(n_valid, K)
  // fewer than K peaks: the reference's heat*keep zeros fill the rest (:405); by the tie rule
  // these are the lowest flat indices that are not selected peaks.
  if (n < K && tid == 0) {
    uint32_t flat = 0;
    for (int m = n; m < K; ++flat) {
      bool taken = false;
      for (int j = 0; j < n && !taken; ++j) taken = (0xffffffffu - (uint32_t)s_sel[j]) == flat;
      if (!taken) s_sel[m++] = (u64)(0xffffffffu - flat);  // ordered-score part 0 => "score 0"
    }
  }
  __syncthreads();
  // bitonic sort, descending, K padded to a power of two with zeros
  int P = 1; while (P < K) P <<= 1;
  for (int i = K + tid; i < P; i += kFinThreads) s_sel[i] = 0ull;
  __syncthreads();
  for (int size = 2; size <= P; size <<= 1) {
    for (int stride = size >> 1; stride > 0; stride >>= 1) {
      for (int i = tid; i < P / 2; i += kFinThreads) {
        const int lo = (i / stride) * 2 * stride + (i % stride);
        const int hi = lo + stride;
        const bool desc = ((lo & size) == 0);
        const u64 a = s_sel[lo], bb = s_sel[hi];
        if (desc ? (a < bb) : (a > bb)) { s_sel[lo] = bb; s_sel[hi] = a; }
      }
      __syncthreads();
    }
  }
  for (int k = tid; k < K; k += kFinThreads) {
    const u64 key = s_sel[k];
    const uint32_t ord = (uint32_t)(key >> 32);
    const uint32_t flat = 0xffffffffu - (uint32_t)key;
    const int cls = (int)(flat / (uint32_t)HW);
    const int sp = (int)(flat - (uint32_t)cls * HW);
    float score = 0.f;
    if (ord != 0u) {
      score = ordered_to_float(ord);
      if (flags & CTDET_DECODE_LOGITS) {
        score = 1.f / (1.f + expf(-score));
        score = fminf(fmaxf(score, 1e-4f), 1.f - 1e-4f);    // centernet.py:161
      }
    }
    float xs = (float)(sp % W), ys = (float)(sp / W);        // :414-415
    if (reg != nullptr) {
      xs = __fadd_rn(xs, reg[((size_t)b * 2 + 0) * HW + sp]);  // :436-439
      ys = __fadd_rn(ys, reg[((size_t)b * 2 + 1) * HW + sp]);
    } else {
      xs += 0.5f; ys += 0.5f;                                  // :440-441
    }
    float w_, h_;
    if (flags & CTDET_DECODE_CAT_SPEC_WH) {                    // :443-446
      w_ = wh[((size_t)b * 2 * C + 2 * cls) * HW + sp];
      h_ = wh[((size_t)b * 2 * C + 2 * cls + 1) * HW + sp];
    } else {
      w_ = wh[((size_t)b * 2 + 0) * HW + sp];
      h_ = wh[((size_t)b * 2 + 1) * HW + sp];
    }
    const float hw_ = w_ * 0.5f, hh_ = h_ * 0.5f;             // "/ 2" is exact
    float* bx = boxes + ((size_t)b * K + k) * 4;               // :452-455
    bx[0] = __fmul_rn(__fsub_rn(xs, hw_), down_ratio);
    bx[1] = __fmul_rn(__fsub_rn(ys, hh_), down_ratio);
    bx[2] = __fmul_rn(__fadd_rn(xs, hw_), down_ratio);
    bx[3] = __fmul_rn(__fadd_rn(ys, hh_), down_ratio);
    scores[(size_t)b * K + k] = score;
    classes[(size_t)b * K + k] = cls;                          // int32 (:418)
    if (inds != nullptr) inds[(size_t)b * K + k] = sp;
  }
}

struct DecodePlan {
  int band_rows, bands, tiles_per_image, cap;
  size_t smem;
};

static DecodePlan make_plan(int B, int C, int H, int W, int K) {
  DecodePlan p;
  // enough CTAs to fill the machine (>= 4 per SM) without shrinking bands below 8 rows
  int bands = 1;
  while ((long long)B * C * bands < 4LL * kNumSMs && H / (bands * 2) >= 8) bands *= 2;
  // a band of more than 128 rows x 512 columns is split further to bound the work per CTA
  while ((long long)(H / bands) * W > 128 * 512 && H / (bands * 2) >= 8) bands *= 2;
  p.band_rows = (H + bands - 1) / bands;
  p.bands = (H + p.band_rows - 1) / p.band_rows;
  p.tiles_per_image = C * p.bands;
  p.cap = K + 2 * kBatchMax;
  p.smem = (size_t)(p.cap + K) * sizeof(u64);
  return p;
}

}  // namespace ctdet

extern "C" size_t ctdet_decode_workspace_bytes(int B, int C, int H, int W, int K) {
  using namespace ctdet;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || K <= 0) return 0;
  const DecodePlan p = make_plan(B, C, H, W, K);
  return (size_t)B * p.tiles_per_image * K * sizeof(u64);
}

extern "C" int ctdet_decode(const float* heat, const float* wh, const float* reg,
                            int B, int C, int H, int W, int K, float down_ratio, int flags,
                            float* boxes, float* scores, int32_t* classes, int64_t* inds,
                            void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || C <= 0 || H <= 0 || W <= 0 || K <= 0) return CTDET_EINVAL;
  if (K > CTDET_MAX_K || (int64_t)K > (int64_t)H * W) return CTDET_EINVAL;   // torch.topk at :411 raises
  if ((int64_t)C * H * W >= ((int64_t)1 << 31)) return CTDET_EINVAL;
  if (flags & ~(CTDET_DECODE_LOGITS | CTDET_DECODE_CAT_SPEC_WH | CTDET_DECODE_NO_NMS)) return CTDET_EINVAL;
  if (B == 0) return CTDET_OK;
  if (!heat || !wh || !boxes || !scores || !classes || !workspace) return CTDET_ENULL;
  if ((W % 4) == 0 && !aligned16(heat)) return CTDET_EALIGN;
  if (reinterpret_cast<uintptr_t>(workspace) & 7u) return CTDET_EALIGN;
  const DecodePlan p = make_plan(B, C, H, W, K);
  if (workspace_bytes < (size_t)B * p.tiles_per_image * K * sizeof(u64)) return CTDET_EWORKSPACE;
  if ((long long)B * p.tiles_per_image > 0x7fffffffLL) return CTDET_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (p.smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(decode_peaks_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  }
  u64* keys = reinterpret_cast<u64*>(workspace);
  decode_peaks_kernel<<<B * p.tiles_per_image, kDecThreads, p.smem, st>>>(
      heat, C, H, W, K, p.band_rows, p.bands, flags, p.cap, keys);
  int rc = check_launch();
  if (rc) return rc;
  decode_finalize_kernel<<<B, kFinThreads, 0, st>>>(keys, p.tiles_per_image * K, wh, reg, C, H, W, K,
                                                    down_ratio, flags, boxes, scores, classes, inds);
  return check_launch();
}
