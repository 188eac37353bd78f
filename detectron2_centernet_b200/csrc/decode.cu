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
//   Every admitted key also bumps a shared-memory histogram over the top bits of its score; the
//   histogram yields, in one scan, the score bin that holds the CTA's K-th best peak.  Keys below
//   that bin are dropped and its lower edge becomes the admission threshold (streaming top-K: one
//   float compare rejects almost every later cell).  CTAs of one image share the best bound found
//   so far through a global atomicMax, and append their survivors (>= K of them, or everything
//   above the shared bound) to a per-image list.  Exactness never depends on the thresholds: they
//   are always lower bounds of the image's K-th best peak.  An exact MSB-first radix select is the
//   fallback when one bin holds too many ties.
// Kernel 2 (one CTA per image): histogram threshold + bitonic sort of the survivors = the exact K
//   best (equivalent to the reference's per-class top-K followed by a top-K over C*K), gather of
//   wh/reg, boxes, outputs; resets the per-image header.
// Workspace: per-image header {count, bound} (zero on entry, left zero) + the key lists.
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

constexpr int kHistBits = 11;                    // bin = ordered(score) >> 21: sign + exponent + 2 mantissa bits
constexpr int kHistBins = 1 << kHistBits;
constexpr int kHistShift = 32 - kHistBits;
constexpr int kKeepSlack = 512;                  // a CTA keeps / emits at most K + kKeepSlack keys

// Finds the highest bin b such that the number of histogram entries in bins >= b reaches `need`.
// Returns (via smem scratch) b and that count; if the whole histogram holds fewer than `need`
// entries, b = 0 and the count is the total.  All kThreads threads must call it.
struct BinScratch { unsigned int warp_tot[32]; int bin; unsigned int cnt_ge; };

template <int kThreads>
__device__ void find_threshold_bin(const unsigned int* hist, int need, BinScratch* bs) {
  constexpr int PER = kHistBins / kThreads;      // descending slots owned by one thread
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned int sum = 0;
#pragma unroll 8
  for (int j = 0; j < PER; ++j) sum += hist[kHistBins - 1 - (tid * PER + j)];
  unsigned int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) bs->warp_tot[warp] = incl;
  if (tid == 0) { bs->bin = 0; bs->cnt_ge = 0xffffffffu; }
  __syncthreads();
  unsigned int before = 0, total = 0;
  for (int w = 0; w < kThreads / 32; ++w) {
    const unsigned int t = bs->warp_tot[w];
    if (w < warp) before += t;
    total += t;
  }
  incl += before;
  const unsigned int excl = incl - sum;
  if (excl < (unsigned)need && incl >= (unsigned)need) {
    unsigned int run = excl;
    for (int j = 0; j < PER; ++j) {
      const int b = kHistBins - 1 - (tid * PER + j);
      run += hist[b];
      if (run >= (unsigned)need) { bs->bin = b; bs->cnt_ge = run; break; }
    }
  }
  __syncthreads();
  if (total < (unsigned)need && tid == 0) { bs->bin = 0; bs->cnt_ge = total; }
  __syncthreads();
}

struct DecodeShared {
  SelectScratch sel;
  BinScratch bs;
  int count;        // keys in buf
  int kept;
  u64 thresh;       // admit keys > thresh
  u64 kept_min;
};

// Exact prune (slow path: many equal / near-equal scores): buf[0..count) -> its K largest keys.
__device__ void prune_exact(u64* buf, u64* tmp, DecodeShared* sh, unsigned int* hist, int K) {
  const int tid = threadIdx.x;
  const int n = sh->count;
  __syncthreads();
  if (n <= K) return;
  const u64 T = radix_select_kth<kDecThreads>([&](int i) { return buf[i]; }, n, K, &sh->sel);
  if (tid == 0) { sh->kept = 0; sh->kept_min = ~0ull; }
  for (int i = tid; i < kHistBins; i += kDecThreads) hist[i] = 0u;
  __syncthreads();
  for (int i = tid; i < n; i += kDecThreads) {
    const u64 key = buf[i];
    if (key >= T) {
      tmp[atomicAdd(&sh->kept, 1)] = key;
      atomicMin(&sh->kept_min, key);
      atomicAdd(&hist[(unsigned)(key >> 32) >> kHistShift], 1u);
    }
  }
  __syncthreads();
  for (int i = tid; i < K; i += kDecThreads) buf[i] = tmp[i];
  if (tid == 0) { sh->count = K; sh->thresh = sh->kept_min - 1ull; }
  __syncthreads();
}

// Histogram prune (fast path): keeps every key whose score bin is >= the bin that holds the K-th
// best (so K <= kept), raises the admission threshold to that bin's lower edge and forgets the
// bins below.  Falls back to the exact prune when more than K + kKeepSlack keys would survive.
// `floor_ord` is an externally known lower bound (the image-wide threshold published by other CTAs).
__device__ void prune(u64* buf, u64* tmp, DecodeShared* sh, unsigned int* hist, int K,
                      uint32_t floor_ord) {
  const int tid = threadIdx.x;
  __syncthreads();
  const int n = sh->count;
  find_threshold_bin<kDecThreads>(hist, K, &sh->bs);
  int bin = sh->bs.bin;
  const int floor_bin = (int)(floor_ord >> kHistShift);
  if (bin < floor_bin) bin = floor_bin;               // the image already has K better peaks
  const u64 edge = ((u64)bin << kHistShift) << 32;    // smallest key of that bin
  if (tid == 0) sh->kept = 0;
  __syncthreads();
  for (int i = tid; i < n; i += kDecThreads) {
    const u64 key = buf[i];
    if (key >= edge) {
      const int pos = atomicAdd(&sh->kept, 1);
      if (pos < K + kKeepSlack) tmp[pos] = key;
    }
  }
  for (int i = tid; i < bin; i += kDecThreads) hist[i] = 0u;
  __syncthreads();
  const int kept = sh->kept;
  if (kept > K + kKeepSlack) {                          // too many ties in the threshold bin
    prune_exact(buf, tmp, sh, hist, K);
    return;
  }
  for (int i = tid; i < kept; i += kDecThreads) buf[i] = tmp[i];
  if (tid == 0) { sh->count = kept; if (edge > sh->thresh + 1ull) sh->thresh = edge - 1ull; }
  __syncthreads();
}

struct DecodeHeader {            // per image, zero on entry, reset by the finalize kernel
  unsigned int count;            // keys emitted so far
  unsigned int thr_ord;          // a valid lower bound (ordered score) of the image's K-th best peak
};

// heat [B,C,H,W]; per image: hdr[b], keys[b][0 .. tiles*(K+kKeepSlack))
__global__ void __launch_bounds__(kDecThreads)
decode_peaks_kernel(const float* __restrict__ heat, int C, int H, int W, int K, int band_rows,
                    int bands, int flags, int cap, DecodeHeader* __restrict__ hdr,
                    u64* __restrict__ out_keys, size_t keys_per_image) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  u64* buf = reinterpret_cast<u64*>(s_raw);             // [cap]
  u64* tmp = buf + cap;                                 // [K + kKeepSlack]
  unsigned int* hist = reinterpret_cast<unsigned int*>(tmp + K + kKeepSlack);   // [kHistBins]
  __shared__ DecodeShared sh;
  __shared__ unsigned int s_base;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int band = blockIdx.x % bands;
  const int plane = blockIdx.x / bands;                 // b*C + c
  const int c = plane % C, b = plane / C;
  const bool logits = flags & CTDET_DECODE_LOGITS;
  const bool no_nms = flags & CTDET_DECODE_NO_NMS;
  const int y_lo = band * band_rows, y_hi = min(H, y_lo + band_rows);
  const float* p = heat + (size_t)plane * H * W;
  const float NEG = -__int_as_float(0x7f800000);        // -inf: max_pool2d padding (:401)

  for (int i = tid; i < kHistBins; i += kDecThreads) hist[i] = 0u;
  if (tid == 0) {
    sh.count = 0;
    const unsigned int g = *((volatile unsigned int*)&hdr[b].thr_ord);
    sh.thresh = g ? (((u64)g << 32) - 1ull) : 0ull;
  }
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
      const u64 th = sh.thresh;
      // threshold in the float domain: one FSETP per cell rejects almost everything early
      const uint32_t th_ord = (uint32_t)(th >> 32);
      const float th_f = th_ord ? ordered_to_float(th_ord) : NEG;
#pragma unroll
      for (int q = 0; q < kRowsPerBatch; ++q) {
        const int y = yb + q;
        float hm_next[4];
        hmax3(nxt[q], nl[q], nr[q], hm_next);
        if (active && y < r1) {
          bool cand[4];
          bool any = false;
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float v = cur[j];
            const float nb = fmaxf(fmaxf(hm_prev[j], hm_cur[j]), hm_next[j]);
            cand[j] = (no_nms || v >= nb) && (v >= th_f) && (logits || v > 0.f) && (x0 + j < W);
            any |= cand[j];
          }
          if (__any_sync(0xffffffffu, any)) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              u64 key = 0ull;
              bool cj = cand[j];
              if (cj) {
                const uint32_t flat = (uint32_t)((c * H + y) * W + x0 + j);
                const uint32_t ord = float_to_ordered(cur[j]);
                key = ((u64)ord << 32) | (u64)(0xffffffffu - flat);
                cj = key > th;
                if (cj) atomicAdd(&hist[ord >> kHistShift], 1u);
              }
              const unsigned m = __ballot_sync(0xffffffffu, cj);
              if (m) {
                int base = 0;
                if (lane == 0) {
                  base = atomicAdd(&sh.count, __popc(m));
                  crossed |= (base + __popc(m) > limit);
                }
                base = __shfl_sync(0xffffffffu, base, 0);
                if (cj) buf[base + __popc(m & ((1u << lane) - 1u))] = key;
              }
            }
          }
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) { hm_prev[j] = hm_cur[j]; hm_cur[j] = hm_next[j]; cur[j] = nxt[q][j]; }
      }
      // CTA-uniform decision: did any warp push the buffer past the safe fill level?
      if (__syncthreads_or(crossed))
        prune(buf, tmp, &sh, hist, K, *((volatile unsigned int*)&hdr[b].thr_ord));
    }
  }

  // final: keep the K best (plus ties of the threshold bin), publish the bound, emit
  const unsigned int g_thr = *((volatile unsigned int*)&hdr[b].thr_ord);
  prune(buf, tmp, &sh, hist, K, g_thr);
  const int n = sh.count;
  if (tid == 0) {
    s_base = n ? atomicAdd(&hdr[b].count, (unsigned)n) : 0u;
    // this CTA alone holds >= K peaks at or above thresh+1: a valid bound for the whole image
    if (n >= K) atomicMax(&hdr[b].thr_ord, (unsigned int)((sh.thresh + 1ull) >> 32));
  }
  __syncthreads();
  u64* dst = out_keys + (size_t)b * keys_per_image + s_base;
  for (int i = tid; i < n; i += kDecThreads) dst[i] = buf[i];
}

// ------------------------------------------------------------------------------------------------
// Kernel 2: per image, exact top-K of the emitted keys, sort, gather, boxes.
constexpr int kSelCap = 2048;

__global__ void __launch_bounds__(kFinThreads)
decode_finalize_kernel(DecodeHeader* __restrict__ hdr, const u64* __restrict__ keys,
                       size_t keys_per_image, const float* __restrict__ wh,
                       const float* __restrict__ reg, int C, int H, int W, int K, float down_ratio,
                       int flags, float* __restrict__ boxes, float* __restrict__ scores,
                       int32_t* __restrict__ classes, int64_t* __restrict__ inds) {
  __shared__ SelectScratch sc;
  __shared__ BinScratch bs;
  __shared__ unsigned int hist[kHistBins];
  __shared__ u64 s_sel[kSelCap];
  __shared__ int s_n;
  const int tid = threadIdx.x;
  const int b = blockIdx.x;
  const u64* kb = keys + (size_t)b * keys_per_image;
  const int n_keys = (int)hdr[b].count;
  const int HW = H * W;

  for (int i = tid; i < kHistBins; i += kFinThreads) hist[i] = 0u;
  if (tid == 0) s_n = 0;
  __syncthreads();
  for (int i = tid; i < n_keys; i += kFinThreads)
    atomicAdd(&hist[(unsigned)(kb[i] >> 32) >> kHistShift], 1u);
  __syncthreads();
  find_threshold_bin<kFinThreads>(hist, K, &bs);
  u64 T = ((u64)bs.bin << kHistShift) << 32;
  if (bs.cnt_ge > (unsigned)kSelCap) {
    // too many keys share the threshold bin: exact radix select instead
    T = radix_select_kth<kFinThreads>([&](int i) { return kb[i]; }, n_keys, K, &sc);
  }
  for (int i = tid; i < n_keys; i += kFinThreads) {
    const u64 key = kb[i];
    if (key >= T) s_sel[atomicAdd(&s_n, 1)] = key;
  }
  __syncthreads();
  if (tid == 0) { hdr[b].count = 0u; hdr[b].thr_ord = 0u; }      // leave the header zeroed
  int n = s_n;                                            // >= min(n_keys, K), <= kSelCap
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
  if (n < K) n = K;
  // bitonic sort, descending, padded to a power of two with zeros; the first K are the answer
  int P = 1; while (P < n) P <<= 1;
  for (int i = n + tid; i < P; i += kFinThreads) s_sel[i] = 0ull;
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
  size_t smem, keys_per_image, header_bytes, total_bytes;
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
  const int keep_max = K + kKeepSlack;
  p.cap = keep_max + 256 + kBatchMax;          // prune when count > cap - kBatchMax
  p.smem = (size_t)(p.cap + keep_max) * sizeof(u64) + (size_t)kHistBins * sizeof(unsigned int);
  p.keys_per_image = (size_t)p.tiles_per_image * keep_max;
  p.header_bytes = (((size_t)B * sizeof(DecodeHeader)) + 255) / 256 * 256;
  p.total_bytes = p.header_bytes + (size_t)B * p.keys_per_image * sizeof(u64);
  return p;
}

}  // namespace ctdet

extern "C" size_t ctdet_decode_workspace_bytes(int B, int C, int H, int W, int K) {
  using namespace ctdet;
  if (B <= 0 || C <= 0 || H <= 0 || W <= 0 || K <= 0) return 0;
  return make_plan(B, C, H, W, K).total_bytes;
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
  if (workspace_bytes < p.total_bytes) return CTDET_EWORKSPACE;
  if ((long long)B * p.tiles_per_image > 0x7fffffffLL) return CTDET_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  if (p.smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(decode_peaks_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  }
  DecodeHeader* hdr = reinterpret_cast<DecodeHeader*>(workspace);
  u64* keys = reinterpret_cast<u64*>(reinterpret_cast<unsigned char*>(workspace) + p.header_bytes);
  decode_peaks_kernel<<<B * p.tiles_per_image, kDecThreads, p.smem, st>>>(
      heat, C, H, W, K, p.band_rows, p.bands, flags, p.cap, hdr, keys, p.keys_per_image);
  int rc = check_launch();
  if (rc) return rc;
  decode_finalize_kernel<<<B, kFinThreads, 0, st>>>(hdr, keys, p.keys_per_image, wh, reg, C, H, W, K,
                                                    down_ratio, flags, boxes, scores, classes, inds);
  return check_launch();
}
