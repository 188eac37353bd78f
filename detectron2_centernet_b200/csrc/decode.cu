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
//   Every admitted key also bumps (fire-and-forget RED) a per-IMAGE histogram in global memory over
//   the top bits of its score.  All CTAs of an image feed it concurrently, so after each CTA's first
//   few rows it already holds thousands of verified peaks; the bin where its suffix count reaches
//   K is a valid lower bound of the image's K-th best peak.  CTAs re-derive that bound (after their
//   first batch and whenever their buffer fills), publish it with atomicMax and use it as the
//   admission threshold (streaming top-K: one float compare rejects almost every later cell), so
//   only a few thousand keys per image ever reach the buffers.  Survivors are appended to a
//   per-image list.  Exactness never depends on the thresholds: they are always lower bounds.  An
//   exact MSB-first radix select is the fallback when one bin holds too many near-ties.
// Kernel 2 (one CTA per image): final bound from the histogram + bitonic sort of the survivors = the
//   exact K best (equivalent to the reference's per-class top-K followed by a top-K over C*K),
//   gather of wh/reg, boxes, outputs; resets the per-image header and histogram.
// Workspace: per-image header {count, bound} + histogram (zero on entry, left zero) + key lists.
#include <type_traits>
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
constexpr int kKeepSlack = 256;                  // a CTA keeps / emits at most K + kKeepSlack keys

// Finds the highest bin b such that the number of histogram entries in bins >= b reaches `need`
// (bin 0 / the total if the histogram holds fewer).  The histogram lives in global memory and may
// grow while it is read: counts only increase, so the reported bin is still a valid lower bound.
// All kThreads threads must call it; results land in *bs.
struct BinScratch { unsigned int warp_tot[32]; int bin; unsigned int cnt_ge; };

template <int kThreads>
__device__ void find_threshold_bin(const unsigned int* hist, int need, BinScratch* bs) {
  constexpr int PER = kHistBins / kThreads;      // descending slots owned by one thread
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned int loc[PER];
  unsigned int sum = 0;
#pragma unroll
  for (int j = 0; j < PER; ++j) { loc[j] = __ldcg(hist + (kHistBins - 1 - (tid * PER + j))); sum += loc[j]; }
  unsigned int incl = sum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) bs->warp_tot[warp] = incl;
  if (tid == 0) { bs->bin = 0; bs->cnt_ge = 0u; }
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
    int b = 0; unsigned int cnt = 0; bool found = false;
#pragma unroll
    for (int j = 0; j < PER; ++j) {
      run += loc[j];
      if (!found && run >= (unsigned)need) { found = true; b = kHistBins - 1 - (tid * PER + j); cnt = run; }
    }
    bs->bin = b; bs->cnt_ge = cnt;
  } else if (tid == 0 && total < (unsigned)need) {
    bs->cnt_ge = total;
  }
  __syncthreads();
}

struct DecodeHeader {            // per image, zero on entry, reset by the finalize kernel
  unsigned int count;            // keys emitted so far
  unsigned int thr_ord;          // a valid lower bound (ordered score) of the image's K-th best peak
};

struct DecodeShared {
  SelectScratch sel;
  BinScratch bs;
  int count;        // keys in buf
  int kept;
  u64 thresh;       // admit keys > thresh
  u64 kept_min;
};

// Re-derives the image-wide bound from the global histogram, publishes it and raises the local
// admission threshold.  All threads.
__device__ void refresh_bound(const unsigned int* ghist, DecodeHeader* hdr, DecodeShared* sh, int K) {
  find_threshold_bin<kDecThreads>(ghist, K, &sh->bs);
  if (threadIdx.x == 0) {
    unsigned int ord = ((unsigned int)sh->bs.bin) << kHistShift;      // lower edge of that bin
    if (sh->bs.cnt_ge < (unsigned)K) ord = 0u;
    if (ord) atomicMax(&hdr->thr_ord, ord);
    const unsigned int g = max(ord, *((volatile unsigned int*)&hdr->thr_ord));
    if (g) { const u64 t = ((u64)g << 32) - 1ull; if (t > sh->thresh) sh->thresh = t; }
  }
  __syncthreads();
}

// Shrinks buf[0..count): first by the freshest image-wide bound; if more than K + kKeepSlack keys
// survive (a bin full of near-ties, or no bound yet), by an exact radix select of the K best.
__device__ void prune(u64* buf, u64* tmp, DecodeShared* sh, const unsigned int* ghist,
                      DecodeHeader* hdr, int K) {
  const int tid = threadIdx.x;
  __syncthreads();
  refresh_bound(ghist, hdr, sh, K);
  const int n = sh->count;
  const u64 th = sh->thresh;
  const int keep_max = K + kKeepSlack;
  if (tid == 0) sh->kept = 0;
  __syncthreads();
  for (int i = tid; i < n; i += kDecThreads) {
    const u64 key = buf[i];
    if (key > th) {
      const int pos = atomicAdd(&sh->kept, 1);
      if (pos < keep_max) tmp[pos] = key;
    }
  }
  __syncthreads();
  const int kept = sh->kept;
  if (kept <= keep_max) {
    for (int i = tid; i < kept; i += kDecThreads) buf[i] = tmp[i];
    __syncthreads();
    if (tid == 0) sh->count = kept;
    __syncthreads();
    return;
  }
  // slow path: exact K best of buf[0..n)
  const u64 T = radix_select_kth<kDecThreads>([&](int i) { return buf[i]; }, n, K, &sh->sel);
  if (tid == 0) { sh->kept = 0; sh->kept_min = ~0ull; }
  __syncthreads();
  for (int i = tid; i < n; i += kDecThreads) {
    const u64 key = buf[i];
    if (key >= T) { tmp[atomicAdd(&sh->kept, 1)] = key; atomicMin(&sh->kept_min, key); }
  }
  __syncthreads();
  for (int i = tid; i < K; i += kDecThreads) buf[i] = tmp[i];
  if (tid == 0) {
    sh->count = K;
    if (sh->kept_min - 1ull > sh->thresh) sh->thresh = sh->kept_min - 1ull;
    // this CTA alone holds K peaks >= kept_min: publish the (bin-rounded) bound
    atomicMax(&hdr->thr_ord, (unsigned int)(sh->kept_min >> 32) >> kHistShift << kHistShift);
  }
  __syncthreads();
}

// heat [B,C,H,W]; per image b: hdr[b], ghist[b][kHistBins], keys[b][0 .. tiles*(K+kKeepSlack))
template <bool kVec>
__global__ void __launch_bounds__(kDecThreads, 7)
decode_peaks_kernel(const float* __restrict__ heat, int B, int C, int H, int W, int K, int band_rows,
                    int bands, int flags, int cap, DecodeHeader* __restrict__ hdrs,
                    unsigned int* __restrict__ ghists, u64* __restrict__ out_keys,
                    size_t keys_per_image) {
  extern __shared__ __align__(16) unsigned char s_raw[];
  u64* buf = reinterpret_cast<u64*>(s_raw);             // [cap]
  u64* tmp = buf + cap;                                 // [K + kKeepSlack]
  __shared__ DecodeShared sh;
  __shared__ unsigned int s_base;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  // CTA order: band fastest, then IMAGE, then class.  CTAs that run concurrently therefore belong
  // to many images, and most CTAs of an image start after its first CTAs have published a bound.
  const int band = blockIdx.x % bands;
  const int b = (blockIdx.x / bands) % B;
  const int c = blockIdx.x / (bands * B);
  const int plane = b * C + c;
  const bool logits = flags & CTDET_DECODE_LOGITS;
  const bool no_nms = flags & CTDET_DECODE_NO_NMS;
  const int y_lo = band * band_rows, y_hi = min(H, y_lo + band_rows);
  const float* p = heat + (size_t)plane * H * W;
  DecodeHeader* hdr = hdrs + b;
  unsigned int* ghist = ghists + (size_t)b * kHistBins;
  const float NEG = -__int_as_float(0x7f800000);        // -inf: max_pool2d padding (:401)
  const float POS = __int_as_float(0x7f800000);
  // probabilities: only cells > 0 are candidates (heat*keep zeros rank last) -> smallest denormal
  const float floor_f = logits ? NEG : __int_as_float(1);

  if (tid == 0) {
    sh.count = 0;
    const unsigned int g = *((volatile unsigned int*)&hdr->thr_ord);
    sh.thresh = g ? (((u64)g << 32) - 1ull) : 0ull;
  }
  __syncthreads();

  // work items of this CTA: (128-column strip, sub-band of rows); a warp takes items round-robin
  const int strips = (W + 127) / 128;
  int subs = kDecWarps / strips; if (subs < 1) subs = 1;
  const int sub_rows = (y_hi - y_lo + subs - 1) / subs;
  const int items = strips * subs;
  const int rounds = (items + kDecWarps - 1) / kDecWarps;
  // the first batch of an item is a single row: its peaks reach the image histogram early and
  // give a first bound after 1/32 of the work instead of 1/8
  const int batches = 1 + (sub_rows - 1 + kRowsPerBatch - 1) / kRowsPerBatch;
  const int limit = cap - kBatchMax;

  for (int rnd = 0; rnd < rounds; ++rnd) {
    const int item = rnd * kDecWarps + warp;
    const bool active = item < items;
    const int strip = active ? item % strips : 0;
    const int sub = active ? item / strips : 0;
    const int r0 = y_lo + sub * sub_rows;
    const int r1 = active ? min(y_hi, r0 + sub_rows) : r0;      // rows [r0, r1) are tested
    const int last_row = active ? min(r1, H - 1) : -1;          // last row that may be loaded
    const int x0 = strip * 128 + lane * 4;                      // this lane's 4 columns
    const bool edge_l = (strips > 1) && (lane == 0) && (x0 > 0) && (x0 - 1 < W);
    const bool edge_r = (strips > 1) && (lane == 31) && (x0 + 4 < W);
    const bool lane_in = x0 < W;
    const float* colp = p + x0;

    // loads row y if 0 <= y <= last_row (warp-uniform), else the row reads as -inf
    auto load_row = [&](int y, float (&v)[4], float& lft, float& rgt) {
      v[0] = v[1] = v[2] = v[3] = NEG; lft = NEG; rgt = NEG;
      if (y >= 0 && y <= last_row) {
        const float* row = colp + (size_t)y * W;
        if (kVec) {
          if (lane_in) { const float4 t = ldg_stream4(row); v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w; }
        } else {
#pragma unroll
          for (int j = 0; j < 4; ++j) if (x0 + j < W) v[j] = __ldg(row + j);
        }
        if (edge_l) lft = __ldg(row - 1);                  // neighbour lives in another strip
        if (edge_r) rgt = __ldg(row + 4);
      }
    };
    // horizontal 3-max of a row; inside the strip the neighbours come from the adjacent lanes
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

    float cur[4], hm_prev[4], hm_cur[4];                 // sliding 3-row window
    {
      float t[4], l, r;
      load_row(r0 - 1, t, l, r);   hmax3(t, l, r, hm_prev);
      load_row(r0, cur, l, r);     hmax3(cur, l, r, hm_cur);
    }

    // tests rows yb .. yb+NR-1 (NR loads in flight per lane before the first use)
    auto run_batch = [&](auto nr_tag, int yb) {
      constexpr int NR = decltype(nr_tag)::value;
      float nxt[NR][4], nl[NR], nr[NR];
#pragma unroll
      for (int q = 0; q < NR; ++q) load_row(yb + q + 1, nxt[q], nl[q], nr[q]);

      bool crossed = false;
      const u64 th = sh.thresh;
      const uint32_t th_ord = (uint32_t)(th >> 32);
      // admission test in the float domain (one FSETP per cell); cells outside the map never pass
      const float thv = fmaxf(th_ord ? ordered_to_float(th_ord) : NEG, floor_f);
      float th_f[4];
#pragma unroll
      for (int j = 0; j < 4; ++j) th_f[j] = (x0 + j < W) ? thv : POS;
      uint32_t flat_row = (uint32_t)((c * H + yb) * W + x0);
#pragma unroll
      for (int q = 0; q < NR; ++q) {
        float hm_next[4];
        hmax3(nxt[q], nl[q], nr[q], hm_next);
        if (yb + q < r1) {                                 // warp-uniform
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const float v = cur[j];
            const float nb = fmaxf(fmaxf(hm_prev[j], hm_cur[j]), hm_next[j]);
            if ((no_nms || v >= nb) && v >= th_f[j]) {     // rare once a bound exists
              const uint32_t ord = float_to_ordered(v);
              const u64 key = ((u64)ord << 32) | (u64)(0xffffffffu - (flat_row + j));
              if (key > th) {
                const int pos = atomicAdd(&sh.count, 1);
                buf[pos] = key;
                atomicAdd(ghist + (ord >> kHistShift), 1u);   // image-wide score histogram (RED)
                crossed |= (pos >= limit);
              }
            }
          }
        }
        flat_row += (uint32_t)W;
#pragma unroll
        for (int j = 0; j < 4; ++j) { hm_prev[j] = hm_cur[j]; hm_cur[j] = hm_next[j]; cur[j] = nxt[q][j]; }
      }
      return crossed;
    };

    for (int bt = 0; bt < batches; ++bt) {
      bool crossed;
      if (bt == 0) crossed = run_batch(std::integral_constant<int, 1>{}, r0);
      else crossed = run_batch(std::integral_constant<int, kRowsPerBatch>{}, r0 + 1 + (bt - 1) * kRowsPerBatch);
      // pick up a better image-wide bound published by other CTAs (visible after the barrier)
      if (tid == 0) {
        const unsigned int g = *((volatile unsigned int*)&hdr->thr_ord);
        if (g) { const u64 t = ((u64)g << 32) - 1ull; if (t > sh.thresh) sh.thresh = t; }
      }
      // CTA-uniform decision: did any thread push the buffer past the safe fill level?
      if (__syncthreads_or(crossed)) prune(buf, tmp, &sh, ghist, hdr, K);
      else if (bt <= 1 && rnd == 0) refresh_bound(ghist, hdr, &sh, K);
    }
  }

  // final: drop everything below the freshest bound, append the survivors to the image's list
  __threadfence();
  prune(buf, tmp, &sh, ghist, hdr, K);
  const int n = sh.count;
  if (tid == 0) s_base = n ? atomicAdd(&hdr->count, (unsigned)n) : 0u;
  __syncthreads();
  u64* dst = out_keys + (size_t)b * keys_per_image + s_base;
  for (int i = tid; i < n; i += kDecThreads) dst[i] = buf[i];
}

// ------------------------------------------------------------------------------------------------
// Kernel 2: per image, exact top-K of the emitted keys, sort, gather, boxes.
constexpr int kSelCap = 2048;

__global__ void __launch_bounds__(kFinThreads)
decode_finalize_kernel(DecodeHeader* __restrict__ hdrs, unsigned int* __restrict__ ghists,
                       const u64* __restrict__ keys, size_t keys_per_image,
                       const float* __restrict__ wh, const float* __restrict__ reg, int C, int H,
                       int W, int K, float down_ratio, int flags, float* __restrict__ boxes,
                       float* __restrict__ scores, int32_t* __restrict__ classes,
                       int64_t* __restrict__ inds) {
  __shared__ SelectScratch sc;
  __shared__ BinScratch bs;
  __shared__ u64 s_sel[kSelCap];
  __shared__ int s_n;
  const int tid = threadIdx.x;
  const int b = blockIdx.x;
  const u64* kb = keys + (size_t)b * keys_per_image;
  DecodeHeader* hdr = hdrs + b;
  unsigned int* ghist = ghists + (size_t)b * kHistBins;
  const int n_keys = (int)hdr->count;
  const int HW = H * W;

  if (tid == 0) s_n = 0;
  // every admitted peak of the image was counted once in ghist: its K-th best bin is exact now
  find_threshold_bin<kFinThreads>(ghist, K, &bs);
  u64 T = ((u64)bs.bin << kHistShift) << 32;
  if (bs.cnt_ge < (unsigned)K) T = 0ull;                  // fewer than K peaks in the whole image
  if (bs.cnt_ge > (unsigned)kSelCap) {
    // too many keys share the threshold bin: exact radix select over the list instead
    T = radix_select_kth<kFinThreads>([&](int i) { return kb[i]; }, n_keys, K, &sc);
  }
  __syncthreads();
  for (int i = tid; i < n_keys; i += kFinThreads) {
    const u64 key = kb[i];
    if (key >= T) { const int pos = atomicAdd(&s_n, 1); if (pos < kSelCap) s_sel[pos] = key; }
  }
  for (int i = tid; i < kHistBins; i += kFinThreads) ghist[i] = 0u;     // leave the workspace zeroed
  __syncthreads();
  if (tid == 0) { hdr->count = 0u; hdr->thr_ord = 0u; }
  int n = min(s_n, kSelCap);                              // >= min(#peaks, K)
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
  size_t smem, keys_per_image, header_bytes, hist_bytes, total_bytes;
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
  p.cap = keep_max + 128 + kBatchMax;          // prune when count > cap - kBatchMax
  p.smem = (size_t)(p.cap + keep_max) * sizeof(u64);
  p.keys_per_image = (size_t)p.tiles_per_image * keep_max;
  p.header_bytes = (((size_t)B * sizeof(DecodeHeader)) + 255) / 256 * 256;
  p.hist_bytes = (size_t)B * kHistBins * sizeof(unsigned int);
  p.total_bytes = p.header_bytes + p.hist_bytes + (size_t)B * p.keys_per_image * sizeof(u64);
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
  const bool vec = (W % 4) == 0;
  auto kern = vec ? decode_peaks_kernel<true> : decode_peaks_kernel<false>;
  if (p.smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  }
  unsigned char* base = reinterpret_cast<unsigned char*>(workspace);
  DecodeHeader* hdr = reinterpret_cast<DecodeHeader*>(base);
  unsigned int* ghist = reinterpret_cast<unsigned int*>(base + p.header_bytes);
  u64* keys = reinterpret_cast<u64*>(base + p.header_bytes + p.hist_bytes);
  kern<<<B * p.tiles_per_image, kDecThreads, p.smem, st>>>(
      heat, B, C, H, W, K, p.band_rows, p.bands, flags, p.cap, hdr, ghist, keys, p.keys_per_image);
  int rc = check_launch();
  if (rc) return rc;
  decode_finalize_kernel<<<B, kFinThreads, 0, st>>>(hdr, ghist, keys, p.keys_per_image, wh, reg, C, H, W,
                                                    K, down_ratio, flags, boxes, scores, classes, inds);
  return check_launch();
}
