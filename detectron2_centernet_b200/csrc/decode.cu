// ctdet decode: 3x3 peak NMS + exact top-K over C*H*W + wh/reg gather + box assembly, batched.
//
// Replaces ctdet_decode (detectron2/modeling/meta_arch/centernet.py:426-458): _nms (:399-405,
// max_pool2d + == + mul: three full passes), _topk (:408-424, two torch.topk + three gathers), two
// full-map NHWC permutes (:435,:442) and the per-image Python loop of CenterNet.inference (:224).
//
// Two paths produce the per-image candidate keys (ordered(score) << 32) | ~flat_index -- one integer compare
// implements the tie rule "higher score, then lower flat index" -- both HBM-bound at 4 B/cell:
//
// STREAMING (default when W % 4 == 0): decode_seed_kernel histograms a ~1/16 sample of every image's peaks;
//   decode_stream_kernel -- a lean TMA-staged sweep, one CTA per ~16 KB band of full-width rows moved into
//   shared memory by ONE 1-D bulk async copy -- derives a valid lower bound of the image's K-th best peak from
//   that histogram while its tile is in flight, rejects ~99.7 % of the cells with one compare and peak-tests
//   the rest in shared memory; the keys go to the TILE's own segment (no atomics).  decode_tail_kernel (one
//   CTA per image) gathers the segments and selects.  See the comments above those kernels.
// STAGED (any shape; also the exact fall-back for images the streaming path hands over: no usable sample
//   bound, or more candidates at the bound than a tile segment holds -- ties, plateaus; the fall-back runs
//   INSIDE decode_tail_kernel, shared by a few helper CTAs per image): decode_peaks_tile,
//   one CTA per (image, class plane, row band).  The band (+ halo rows) is staged into shared
//   memory by ONE bulk async copy (TMA 1-D, mbarrier completion); several CTAs are resident per SM
//   so the copies of some overlap the peak tests of others.  Warps walk down the rows of a
//   128-column strip (float4 per lane per row from shared memory, 3-row sliding window in
//   registers, horizontal neighbours by warp shuffle) and record the candidate cells -- 3x3 peaks
//   at or above the current bound -- as a bitmap (4 ballots per row).  No barriers, no atomics in
//   this sweep.  The few candidates are then collected from the bitmap into 64-bit keys
//   and each bumps (fire-and-forget RED) a per-IMAGE histogram in
//   global memory over the top bits of its score.  The bin where that histogram's suffix count
//   reaches K is a valid lower bound of the image's K-th best peak: the first CTAs of an image
//   derive and publish it (atomicMax), later CTAs start from it, so one float compare rejects
//   almost every cell and only a few thousand keys per image survive.  Survivors are appended to a
//   per-image list.  Exactness never depends on the bounds: they are always lower bounds; an exact
//   MSB-first radix select is the fallback when a tile holds too many near-ties.
// TAIL (one CTA per image): threshold from a histogram of the survivors + rank sort = the
//   exact K best (equivalent to the reference's per-class top-K followed by a top-K over C*K),
//   gather of wh/reg, boxes, outputs; resets the per-image header and histograms.
// With CTDET_DECODE_LOGITS the 3x3 peak test is defined on the fp32 probabilities clamp(sigmoid(x)) -- the values
//   the reference compares (centernet.py:161, :401-404) -- so cells that share a probability with their largest
//   neighbour (clamp saturation, rounding near 1) are all kept, like the reference's plateaus; the ranking stays
//   on the logits (same order; among equal probabilities: higher logit, then lower flat index).
// Workspace: per-image headers + histograms (zero on entry, left zero) + tile counts / key segments + staged lists.
#include <algorithm>
#include <cstdlib>
#include "dcn_common.cuh"      // common.cuh + the tcgen05 / TMEM / mbarrier wrappers (head_decode.cuh at the end of this file)

namespace ctdet {

constexpr int kDecThreads = 128;                 // kernel 1: 4 warps
constexpr int kDecWarps = kDecThreads / 32;
constexpr int kFinThreads = 512;                 // kernel 2

typedef unsigned long long u64;

// Programmatic dependent launch (PDL): a kernel launched with the programmatic-serialization attribute may
// start while its predecessor in the stream drains; everything that depends on the predecessor's results
// comes after pdl_wait() (= the predecessor grid has completed and its writes are visible).  Both are
// no-ops in a kernel launched without the attribute.
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

// Whole-decode timeline (only with -DCTDET_DECODE_TIMING; scripts/decode_timeline.py): first / last %globaltimer stamp per
// event over all CTAs -- 0/1 seed first start / last end, 2/3 first / last bound published, 4/5 sweep first start / last
// start, 6 first sweep CTA past its bound, 7 sweep last end, 8 tail first start, 9/10 first / last tail CTA past
// griddepcontrol.wait, 11 tail last end.
#ifdef CTDET_DECODE_TIMING
__device__ unsigned long long g_dec_ts[16];
__device__ __forceinline__ unsigned long long gt_ns() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
#define DEC_TS_MIN(i) do { if (threadIdx.x == 0) atomicMin(&g_dec_ts[i], gt_ns()); } while (0)
#define DEC_TS_MAX(i) do { if (threadIdx.x == 0) atomicMax(&g_dec_ts[i], gt_ns()); } while (0)
#else
#define DEC_TS_MIN(i) do {} while (0)
#define DEC_TS_MAX(i) do {} while (0)
#endif

template <typename... KArgs, typename... Args>
static cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st, Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = smem; cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at; cfg.numAttrs = 1;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

struct SelectScratch {
  unsigned int hist[256];
  int need;         // how many still to take from the current bucket
  int bucket_count;
  int digit;
};

// Exact "k-th largest" of n distinct 64-bit keys: returns T such that exactly min(n,k) keys are
// >= T.  MSB-first radix select, 8 bits per pass; stops as soon as the bucket holding the k-th
// key must be taken whole.  visit(f) must call f(key) for every key owned by the calling thread
// (the union over the CTA's threads = the n keys).  All threads of the CTA must call it.
template <int kThreads, typename Visit>
__device__ u64 radix_select_kth(Visit visit, int n, int k, SelectScratch* sc) {
  const int tid = threadIdx.x;
  if (n <= k) return 0ull;
  u64 prefix = 0ull, mask = 0ull;
  int need = k;
  for (int shift = 56; shift >= 0; shift -= 8) {
    for (int i = tid; i < 256; i += kThreads) sc->hist[i] = 0u;
    __syncthreads();
    visit([&](u64 key) {
      if ((key & mask) == prefix) atomicAdd(&sc->hist[(unsigned)(key >> shift) & 255u], 1u);
    });
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
constexpr int kKeepSlack = 384;                  // a CTA collects / emits at most K + kKeepSlack keys

// Finds the highest bin b such that the number of histogram entries in bins >= b reaches `need`
// (bin 0 / the total if the histogram holds fewer).  The histogram lives in global memory and may
// grow while it is read: counts only increase, so the reported bin is still a valid lower bound.
// All kThreads threads must call it; results land in *bs.
struct BinScratch { unsigned int warp_tot[32]; int bin; unsigned int cnt_ge; };

template <int kThreads, bool kSharedHist = false>
__device__ void find_threshold_bin(const unsigned int* hist, int need, BinScratch* bs) {
  constexpr int PER = kHistBins / kThreads;      // descending slots owned by one thread
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  unsigned int loc[PER];
  unsigned int sum = 0;
#pragma unroll
  for (int j = 0; j < PER; ++j) {
    const unsigned int* q = hist + (kHistBins - 1 - (tid * PER + j));
    loc[j] = kSharedHist ? *q : __ldcg(q);
    sum += loc[j];
  }
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

// clamp(sigmoid(x), 1e-4, 1 - 1e-4) exactly as the finalize step scores a winner (centernet.py:161).  With
// CTDET_DECODE_LOGITS the 3x3 peak test is defined on THESE fp32 values, like the reference's hmax == heat
// (:401-404): distinct logits can share one probability (clamp saturation, rounding near 1), and the reference
// then keeps every cell of the plateau.  sigmoid is monotone, so "x_c >= max neighbour logit" already implies
// the peak; only a centre BELOW its largest neighbour needs the two sigmoids.
__device__ __forceinline__ float prob_of_logit(float x) {
  const float p = 1.f / (1.f + expf(-x));
  return fminf(fmaxf(p, 1e-4f), 1.f - 1e-4f);
}
__device__ __noinline__ bool same_prob(float a, float b) { return prob_of_logit(a) == prob_of_logit(b); }   // rare: out of line
__device__ __forceinline__ bool is_peak(float v, float nb, bool logits) {
  return v >= nb || (logits && same_prob(v, nb));
}

struct DecodeHeader {            // per image, zero on entry, reset by the finalize kernel
  unsigned int count;            // keys emitted so far by decode_peaks_kernel (the staged path)
  unsigned int thr_ord;          // a valid lower bound (ordered score) of the image's K-th best peak
  unsigned int s_count;          // keys emitted by decode_stream_kernel
  unsigned int staged;           // 1: the image is decoded by decode_peaks_kernel (no seed bound, or the
                                 //    streaming list overflowed); finalize then reads that path's list
  unsigned int ticket;           // decode_tail_kernel: helper CTAs of a handed-over image that finished their tiles
  unsigned int seed_ticket;      // decode_seed_kernel: CTAs of the image that have flushed their sample histogram
  unsigned int pad[2];
};

struct DecodeShared {
  SelectScratch sel;
  BinScratch bs;
  int ncand;        // candidate bits set in the bitmap
  int count;        // keys in buf
  u64 thresh;       // candidates must be > thresh
};

// Re-derives the image-wide bound from the global histogram, publishes it and raises the local
// threshold.  All threads.
template <int NT>
__device__ void refresh_bound(const unsigned int* ghist, DecodeHeader* hdr, DecodeShared* sh, int K) {
  find_threshold_bin<NT>(ghist, K, &sh->bs);
  if (threadIdx.x == 0) {
    unsigned int ord = ((unsigned int)sh->bs.bin) << kHistShift;      // lower edge of that bin
    if (sh->bs.cnt_ge < (unsigned)K) ord = 0u;
    if (ord) atomicMax(&hdr->thr_ord, ord);
    const unsigned int g = max(ord, *((volatile unsigned int*)&hdr->thr_ord));
    if (g) { const u64 t = ((u64)g << 32) - 1ull; if (t > sh->thresh) sh->thresh = t; }
  }
  __syncthreads();
}

// ---- mbarrier / bulk-copy helpers (TMA 1-D bulk load: "UBLKCP" in SASS) ---------------------------
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;"
               :: "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      :: "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc), "r"(bytes),
                  "r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}

// heat [B,C,H,W]; per image b: hdr[b], ghist[b][kHistBins], keys[b][0 .. tiles*(K+kKeepSlack))
// grid = (bands, B, C): band fastest, then IMAGE, then class.  CTAs that run concurrently therefore
// belong to many images, and most CTAs of an image start after its first CTAs published a bound.
template <int NT>
__device__ __forceinline__ void
decode_peaks_tile(const float* __restrict__ heat, int C, int H, int W, int Wp, int K,
                  int band_rows, int sub_rows, FastDiv div_strips, int flags,
                  DecodeHeader* __restrict__ hdrs, unsigned int* __restrict__ ghists,
                  u64* __restrict__ out_keys, size_t keys_per_image,
                  const int band, const int b, const int c, unsigned char* s_raw, const bool reinit) {
  const int strips = (int)div_strips.d;
  const int cap = (K + kKeepSlack + 1) & ~1;                            // even: keeps the bitmap 16-byte aligned
  float* tile = reinterpret_cast<float*>(s_raw);                        // [(band_rows+2)][Wp]
  u64* buf = reinterpret_cast<u64*>(s_raw + (size_t)(band_rows + 2) * Wp * sizeof(float));   // [cap]
  uint32_t* bitmap = reinterpret_cast<uint32_t*>(buf + cap);            // [band_rows][strips][4]
  __shared__ DecodeShared sh;
  __shared__ unsigned int s_base;
  __shared__ __align__(8) uint64_t s_bar;

  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int plane = b * C + c;
  const bool logits = flags & CTDET_DECODE_LOGITS;
  const bool no_nms = flags & CTDET_DECODE_NO_NMS;
  const int y_lo = band * band_rows, y_hi = min(H, y_lo + band_rows);
  const int nrows = y_hi - y_lo;                                        // tile row t <-> map row y_lo-1+t
  const float* p = heat + (size_t)plane * H * W;
  DecodeHeader* hdr = hdrs + b;
  unsigned int* ghist = ghists + (size_t)b * kHistBins;
  const float NEG = -__int_as_float(0x7f800000);                        // max_pool2d padding (:401)
  const float POS = __int_as_float(0x7f800000);
  // candidates: probabilities must be > 0 (heat*keep zeros rank last, :405); logits must be finite
  const float floor_f = (logits || no_nms) ? -3.402823466e38f : __int_as_float(1);

  // ---- stage the band (+ halo rows) -------------------------------------------------------------
  const int g_first = max(y_lo - 1, 0), g_last = min(y_hi, H - 1);      // map rows that exist
  const bool bulk = (Wp == W);
  const int n_words = nrows * strips * 4;
  if (tid == 0) {
    if (reinit) asm volatile("mbarrier.inval.shared::cta.b64 [%0];" :: "r"((uint32_t)__cvta_generic_to_shared(&s_bar)) : "memory");
    mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    if (bulk) {
      const unsigned bytes = (unsigned)((g_last - g_first + 1) * W * sizeof(float));
      mbar_expect_tx(&s_bar, bytes);
      bulk_g2s(tile + (size_t)(g_first - (y_lo - 1)) * Wp, p + (size_t)g_first * W, bytes, &s_bar);
    }
    sh.ncand = 0; sh.count = 0;
    const unsigned int g = *((volatile unsigned int*)&hdr->thr_ord);   // overlaps the copy
    sh.thresh = g ? (((u64)g << 32) - 1ull) : 0ull;
  }
  if (!bulk) {                                          // W % 4 != 0: padded rows, plain copies
    for (int i = tid; i < (g_last - g_first + 1) * Wp; i += NT) {
      const int r = i / Wp, x = i - r * Wp;
      tile[(size_t)(g_first - (y_lo - 1) + r) * Wp + x] = x < W ? __ldg(p + (size_t)(g_first + r) * W + x) : NEG;
    }
  }
  if (y_lo == 0) for (int x = tid; x < Wp; x += NT) tile[x] = NEG;
  if (y_hi == H) for (int x = tid; x < Wp; x += NT) tile[(size_t)(nrows + 1) * Wp + x] = NEG;
  for (int w = tid; w < n_words; w += NT) bitmap[w] = 0u;
  __syncthreads();
  const u64 th0 = sh.thresh;
  const bool had_bound = th0 != 0ull;
  if (bulk) mbar_wait(&s_bar, 0);

  // ---- sweep: candidate bitmap ------------------------------------------------------------------
  // work items: (128-column strip, sub-band of rows); a warp takes items round-robin.  Per row a
  // lane reads its 4 cells and the warp votes whether ANY of them reaches the bound; with a bound
  // in place that is false for ~97% of the rows and the row costs ~8 instructions.  Only the other
  // rows run the 3x3 test (rows t-1, t, t+1 are all in shared memory).
  {
    const int items = strips * ((nrows + sub_rows - 1) / sub_rows);
    const uint32_t th_ord = (uint32_t)(th0 >> 32);
    const float thv = fmaxf(th_ord ? ordered_to_float(th_ord) : NEG, floor_f);
    for (int item = warp; item < items; item += (NT / 32)) {
      const int sub = (int)div_strips.div((uint32_t)item), strip = item - sub * strips;
      const int t0 = 1 + sub * sub_rows;                        // tile rows [t0, t1) are tested
      const int t1 = min(nrows + 1, t0 + sub_rows);
      const int x0 = strip * 128 + lane * 4;                    // this lane's 4 columns
      const bool lane_in = x0 < Wp;
      const bool has_l = (lane == 0) && (x0 > 0);               // neighbour cell in another strip
      const bool has_r = (lane == 31) && (x0 + 4 < Wp);
      const float* colp = tile + (lane_in ? x0 : 0);
      const float th_lane = lane_in ? thv : POS;

      // the 3x3 test of tile row t for this strip; warp-uniform call
      auto full_test = [&](int t) {
        float hm[3][4], v1[4];
#pragma unroll
        for (int r = 0; r < 3; ++r) {
          const float* row = colp + (size_t)(t - 1 + r) * Wp;
          const float4 q = *reinterpret_cast<const float4*>(row);
          const float v[4] = {lane_in ? q.x : NEG, lane_in ? q.y : NEG, lane_in ? q.z : NEG, lane_in ? q.w : NEG};
          const float lft = has_l ? row[-1] : NEG;
          const float rgt = has_r ? row[4] : NEG;
          const float l = __shfl_up_sync(0xffffffffu, v[3], 1);
          const float rr = __shfl_down_sync(0xffffffffu, v[0], 1);
          const float L = lane == 0 ? lft : l;
          const float R = lane == 31 ? rgt : rr;
          hm[r][0] = fmaxf(fmaxf(L, v[0]), v[1]);
          hm[r][1] = fmaxf(fmaxf(v[0], v[1]), v[2]);
          hm[r][2] = fmaxf(fmaxf(v[1], v[2]), v[3]);
          hm[r][3] = fmaxf(fmaxf(v[2], v[3]), R);
          if (r == 1) { v1[0] = v[0]; v1[1] = v[1]; v1[2] = v[2]; v1[3] = v[3]; }
        }
        uint32_t m[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const float nb = fmaxf(fmaxf(hm[0][j], hm[1][j]), hm[2][j]);
          const float thj = (x0 + j < W) ? thv : POS;          // pad cells never pass
          m[j] = __ballot_sync(0xffffffffu, v1[j] >= thj && (no_nms || is_peak(v1[j], nb, logits)));
        }
        if (lane == 0 && (m[0] | m[1] | m[2] | m[3]))
          *reinterpret_cast<uint4*>(bitmap + ((size_t)(t - 1) * strips + strip) * 4) = make_uint4(m[0], m[1], m[2], m[3]);
      };

      const float* rowp = colp + (size_t)t0 * Wp;
#pragma unroll 4
      for (int t = t0; t < t1; ++t, rowp += Wp) {
        const float4 q = *reinterpret_cast<const float4*>(rowp);
        const float mx = fmaxf(fmaxf(q.x, q.y), fmaxf(q.z, q.w));
        if (__any_sync(0xffffffffu, mx >= th_lane)) full_test(t);
      }
    }
  }
  __syncthreads();

  // ---- candidates -> keys -----------------------------------------------------------------------
  // calls f(key, word index, bit) for every candidate bit owned by this thread
  auto for_each_candidate = [&](auto f) {
    for (int w = tid; w < n_words; w += NT) {
      uint32_t bits = bitmap[w];
      if (!bits) continue;
      const int j = w & 3, rs = w >> 2;
      const int t = (int)div_strips.div((uint32_t)rs), strip = rs - t * strips;   // tile row t+1
      while (bits) {
        const int ln = __ffs(bits) - 1;
        bits &= bits - 1;
        const int x = strip * 128 + ln * 4 + j;
        const float v = tile[(size_t)(t + 1) * Wp + x];
        const uint32_t flat = (uint32_t)((c * H + (y_lo + t)) * W + x);
        f(((u64)float_to_ordered(v) << 32) | (u64)(0xffffffffu - flat), w, ln);
      }
    }
  };
  // Fast path (the common case once the image has a bound): append the candidates above the
  // threshold straight into the key buffer -- no counting pass, one barrier.
  {
    const u64 th = sh.thresh;
    for_each_candidate([&](u64 key, int, int) {
      if (key > th) { const int pos = atomicAdd(&sh.count, 1); if (pos < cap) buf[pos] = key; }
    });
  }
  __syncthreads();
  int n = sh.count;
  if (n == 0) return;                                    // CTA-uniform: nothing in this tile
  if (n > cap) {
    // Slow path (weak or no bound yet, plateaus, ties).  1: count every candidate into the image
    // histogram and re-derive the bound; 2: drop the candidates below it; 3: if still too many,
    // keep exactly the tile's K best (exact radix select); 4: collect again.  CTA-uniform branch.
    for_each_candidate([&](u64 key, int, int) { atomicAdd(ghist + ((uint32_t)(key >> 32) >> kHistShift), 1u); });
    __threadfence();
    __syncthreads();
    refresh_bound<NT>(ghist, hdr, &sh, K);
    if (tid == 0) { sh.ncand = 0; sh.count = 0; }
    __syncthreads();
    {
      const u64 th = sh.thresh;
      int kept = 0;
      for_each_candidate([&](u64 key, int w, int ln) {
        if (key > th) ++kept; else atomicAnd(&bitmap[w], ~(1u << ln));
      });
      kept = warp_sum(kept);
      if (lane == 0 && kept) atomicAdd(&sh.ncand, kept);
    }
    __syncthreads();
    if (sh.ncand > cap) {
      const int nc = sh.ncand;
      const u64 T = radix_select_kth<NT>(
          [&](auto f) { for_each_candidate([&](u64 key, int, int) { f(key); }); }, nc, K, &sh.sel);
      for_each_candidate([&](u64 key, int w, int ln) { if (key < T) atomicAnd(&bitmap[w], ~(1u << ln)); });
      if (tid == 0 && T - 1ull > sh.thresh) sh.thresh = T - 1ull;
      __syncthreads();
    }
    {                                                    // at most cap keys by construction
      const u64 th = sh.thresh;
      for_each_candidate([&](u64 key, int, int) {
        if (key > th) { const int pos = atomicAdd(&sh.count, 1); if (pos < cap) buf[pos] = key; }
      });
    }
    __syncthreads();
    n = min(sh.count, cap);
  } else {
    // every collected key is counted exactly once in the image histogram
    for (int i = tid; i < n; i += NT) atomicAdd(ghist + ((uint32_t)(buf[i] >> 32) >> kHistShift), 1u);
  }

  // ---- publish / pick up the image bound, emit --------------------------------------------------
  // CTAs that started without a bound (the first of their image) or still hold many keys derive a
  // new bound from the histogram; the others just emit.
  u64 th = sh.thresh;
  if (!had_bound || n > K / 2 + 32) {
    __threadfence();
    refresh_bound<NT>(ghist, hdr, &sh, K);
    th = sh.thresh;
    // the bound may have risen: count, reserve and compact the keys that still pass
    int mine = 0;
    for (int i = tid; i < n; i += NT) mine += (buf[i] > th);
    mine = warp_sum(mine);
    if (tid == 0) sh.count = 0;
    __syncthreads();
    if (lane == 0 && mine) atomicAdd(&sh.count, mine);
    __syncthreads();
    if (tid == 0) { const int tot = sh.count; s_base = tot ? atomicAdd(&hdr->count, (unsigned)tot) : 0u; }
    __syncthreads();
    if (tid == 0) sh.count = 0;
    __syncthreads();
    u64* dst = out_keys + (size_t)b * keys_per_image + s_base;
    for (int i = tid; i < n; i += NT) {
      const u64 key = buf[i];
      if (key > th) dst[atomicAdd(&sh.count, 1)] = key;
    }
  } else {
    // threshold unchanged since the collect: all n keys pass
    if (tid == 0) s_base = atomicAdd(&hdr->count, (unsigned)n);
    __syncthreads();
    u64* dst = out_keys + (size_t)b * keys_per_image + s_base;
    for (int i = tid; i < n; i += NT) dst[i] = buf[i];
  }
}

// grid (bands, B, C), one tile per CTA -- band fastest, then IMAGE, then class, so CTAs that run concurrently
// belong to many images and most CTAs of an image start after its first CTAs published a bound.  (The fall-back
// behind the sweep runs the same tile function inside decode_tail_kernel.)
__global__ void __launch_bounds__(kDecThreads)
decode_peaks_kernel(const float* __restrict__ heat, int C, int H, int W, int Wp, int K,
                    int band_rows, int sub_rows, FastDiv div_strips, int flags,
                    DecodeHeader* __restrict__ hdrs, unsigned int* __restrict__ ghists,
                    u64* __restrict__ out_keys, size_t keys_per_image) {
  extern __shared__ __align__(128) unsigned char s_raw[];
  pdl_launch_dependents();
  decode_peaks_tile<kDecThreads>(heat, C, H, W, Wp, K, band_rows, sub_rows, div_strips, flags, hdrs, ghists, out_keys,
                                 keys_per_image, blockIdx.x, blockIdx.y, blockIdx.z, s_raw, false);
}

// ------------------------------------------------------------------------------------------------
// Streaming path (the default when C*H*W and W are multiples of 4): seed bound from a sample, then a
// pure threshold sweep over the logits.
//
// decode_seed_kernel: 3x3-peak-tests a SAMPLE of every image -- groups of kSeedRows rows every
// kSeedPeriod rows of every class plane, interior columns only, ~1/8 of the cells for ~3/16 of the
// bytes -- and histograms the sampled peaks per image over the same 11-bit bins as above.  Every
// counted cell is a real peak of the image, so the bin where the sample's suffix count reaches K is
// a valid lower bound of the image's K-th best peak (typically near its 8K-th best).  The last CTA of
// an image (ticket) publishes it in hdr->thr_ord and leaves the sample histogram zeroed; with fewer
// than K sampled peaks the image is handed to the staged path (hdr->staged = 1).
constexpr int kSeedThreads = 256;
constexpr int kSeedWarps = kSeedThreads / 32;
constexpr int kSeedPeriod = 64, kSeedRows = 4;
constexpr int kSeedHist = kHistBins + kHistBins / 32;   // per image: the 2048 fine bins, then 64 coarse bins (32 fine each)

// The bound of one image from its sample histogram (a whole warp, warp-uniform result): the highest bin whose
// suffix count reaches K -- two dependent 256/128-byte reads: the 64 coarse bins (32 fine bins each), then the 32
// fine bins of the coarse bin that holds the K-th entry; lane 0 owns the highest bins.  0 = fewer than K samples.
__device__ unsigned int seed_bound_of_image(const unsigned int* ph, const unsigned int* pv, int n_vals, int K,
                                            unsigned int* sub /* 32 shared words of this warp */) {
  const int lane = threadIdx.x & 31;
    const uint2 c2 = __ldcg(reinterpret_cast<const uint2*>(ph + kHistBins) + (31 - lane));  // coarse 62-2l, 63-2l
    const unsigned int sum = c2.x + c2.y;
    unsigned int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const unsigned int t = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += t;
    }
    const unsigned int excl = incl - sum;
    const unsigned int hit = __ballot_sync(0xffffffffu, excl < (unsigned)K && incl >= (unsigned)K);
    unsigned int ord = 0u;
    if (hit) {
      const int src = __ffs(hit) - 1;
      const bool upper = excl + c2.y >= (unsigned)K;          // meaningful on lane src only
      const int coarse = __shfl_sync(0xffffffffu, upper ? 63 - 2 * lane : 62 - 2 * lane, src);
      const unsigned int before = __shfl_sync(0xffffffffu, upper ? excl : excl + c2.y, src);
      const unsigned int f = __ldcg(ph + coarse * 32 + (31 - lane));                         // descending by lane
      unsigned int fi = f;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) {
        const unsigned int t = __shfl_up_sync(0xffffffffu, fi, o);
        if (lane >= o) fi += t;
      }
      const unsigned int hit2 = __ballot_sync(0xffffffffu, before + fi - f < (unsigned)K && before + fi >= (unsigned)K);
      if (hit2) {
        const int l2 = __ffs(hit2) - 1;
        ord = (unsigned int)(coarse * 32 + 31 - l2) << kHistShift;
        // Concentrated scores (e.g. a random-init head: every logit within +-0.3 of the -2.19 bias) put most of the
        // sample into this one quarter-octave bin, and its lower edge would pass the whole image.  Then refine
        // on the sample itself, 5 more bits per round: a per-warp 32-bin histogram of the samples that share the
        // prefix found so far, same suffix-count rule.  Still a valid lower bound: >= K samples are >= it.
        unsigned int in_bin = __shfl_sync(0xffffffffu, f, l2);
        unsigned int need = (unsigned)K - __shfl_sync(0xffffffffu, before + fi - f, l2);
        unsigned int prefix = ord >> kHistShift;
        int shift = kHistShift;
        
        
        // only when the bin really is crowded (a normal image has n_vals / 30 or so samples in it, and every warp
        // of every CTA runs this code): then keep going until the sub-bin is small
        const unsigned int crowded = max(2u * (unsigned)K, (unsigned)n_vals / 8u);
        if (in_bin > crowded)
        while (in_bin > 24u && shift >= 5) {
          sub[lane] = 0u;
          __syncwarp();
          for (int i = lane; i < n_vals; i += 32) {
            const unsigned int v = __ldcg(pv + i);
            if ((v >> shift) == prefix) atomicAdd(&sub[(v >> (shift - 5)) & 31u], 1u);
          }
          __syncwarp();
          const unsigned int cnt = sub[31 - lane];                // descending by lane
          unsigned int ci = cnt;
#pragma unroll
          for (int o = 1; o < 32; o <<= 1) {
            const unsigned int t = __shfl_up_sync(0xffffffffu, ci, o);
            if (lane >= o) ci += t;
          }
          const unsigned int h3 = __ballot_sync(0xffffffffu, ci - cnt < need && ci >= need);
          __syncwarp();
          if (!h3) break;                                          // (cannot happen: the bin holds >= need samples)
          const int l3 = __ffs(h3) - 1;
          prefix = (prefix << 5) | (unsigned)(31 - l3);
          shift -= 5;
          need -= __shfl_sync(0xffffffffu, ci - cnt, l3);
          in_bin = __shfl_sync(0xffffffffu, cnt, l3);
        }
        ord = prefix << shift;
      }
    }
    return ord;
}

#ifndef CTDET_SEED_MINB
#define CTDET_SEED_MINB 1
#endif
#ifndef CTDET_SEED_ITEMS_PER_WARP
#define CTDET_SEED_ITEMS_PER_WARP 1
#endif
__global__ void __launch_bounds__(kSeedThreads, CTDET_SEED_MINB)
decode_seed_kernel(const float* __restrict__ heat, int C, int H, int W, int K, int flags,
                   int groups, FastDiv div_strips, FastDiv div_groups, DecodeHeader* __restrict__ hdrs,
                   unsigned int* __restrict__ phists, unsigned int* __restrict__ pvals_all, int n_vals) {
  __shared__ unsigned int s_sub[32];
  __shared__ unsigned int s_last;
  pdl_launch_dependents();             // the sweep may be scheduled as soon as every CTA of this grid runs
  DEC_TS_MIN(0);
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  const int b = blockIdx.y;
  const int strips = (int)div_strips.d;
  const bool logits = flags & CTDET_DECODE_LOGITS;
  const bool no_nms = flags & CTDET_DECODE_NO_NMS;
  const float NEG = -__int_as_float(0x7f800000);
  const float floor_f = (logits || no_nms) ? -3.402823466e38f : __int_as_float(1);
  unsigned int* phist = phists + (size_t)b * kSeedHist;
  unsigned int* pvals = pvals_all + (size_t)b * n_vals;  // [items][kSeedRows] ordered scores, 0 = no peak

  const int items = C * groups * strips;
  for (int item = blockIdx.x * kSeedWarps + warp; item < items; item += gridDim.x * kSeedWarps) {
    const int cg = (int)div_strips.div((uint32_t)item), strip = item - cg * strips;
    const int c = (int)div_groups.div((uint32_t)cg), g = cg - c * groups;
    const int r0 = 1 + g * kSeedPeriod;                       // tested rows r0 .. r0+kSeedRows-1, all interior
    const int x0 = strip * 128 + lane * 4;
    const bool in = x0 < W;                                   // W % 4 == 0: the lane's 4 cells exist
    const float* base = heat + (((size_t)b * C + c) * H + (r0 - 1)) * W + (in ? x0 : 0);
    float4 q[kSeedRows + 2];
#pragma unroll
    for (int r = 0; r < kSeedRows + 2; ++r)
      q[r] = in ? ldg_stream4(base + (size_t)r * W) : make_float4(NEG, NEG, NEG, NEG);
    float hm[kSeedRows + 2][4];
#pragma unroll
    for (int r = 0; r < kSeedRows + 2; ++r) {
      const float l = __shfl_up_sync(0xffffffffu, q[r].w, 1);
      const float rr = __shfl_down_sync(0xffffffffu, q[r].x, 1);
      hm[r][0] = fmaxf(fmaxf(l, q[r].x), q[r].y);             // lane 0 / lane 31 edge cells are never tested
      hm[r][1] = fmaxf(fmaxf(q[r].x, q[r].y), q[r].z);
      hm[r][2] = fmaxf(fmaxf(q[r].y, q[r].z), q[r].w);
      hm[r][3] = fmaxf(fmaxf(q[r].z, q[r].w), rr);
    }
    // one histogram entry per (row, 128-column segment): its best sampled peak.  Distinct entries are
    // distinct peaks, and near the top of the distribution almost every high peak is its segment's best.
    float best[kSeedRows];
#pragma unroll
    for (int t = 1; t <= kSeedRows; ++t) {
      const float v[4] = {q[t].x, q[t].y, q[t].z, q[t].w};
      float bt = NEG;
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int x = x0 + j;
        const bool testable = in && x >= 1 && x <= W - 2 && !(lane == 0 && j == 0) && !(lane == 31 && j == 3);
        const float nb = fmaxf(fmaxf(hm[t - 1][j], hm[t][j]), hm[t + 1][j]);
        if (testable && (no_nms || v[j] >= nb) && v[j] >= floor_f) bt = fmaxf(bt, v[j]);
      }
      best[t - 1] = bt;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int t = 0; t < kSeedRows; ++t) best[t] = fmaxf(best[t], __shfl_xor_sync(0xffffffffu, best[t], o));
    }
    float mine = NEG;
#pragma unroll
    for (int t = 0; t < kSeedRows; ++t) if (lane == t) mine = best[t];
    if (lane < kSeedRows) {
      const bool has = mine >= floor_f;
      const uint32_t ord = has ? float_to_ordered(mine) : 0u;
      pvals[item * kSeedRows + lane] = ord;                  // the sample itself, for the rare refinement below 11 bits
      if (has) {
        atomicAdd(phist + (ord >> kHistShift), 1u);
        atomicAdd(phist + kHistBins + (ord >> (kHistShift + 5)), 1u);
      }
    }
  }
  // ---- the last CTA of the image (ticket) derives the bound ONCE and publishes it.  The sweep's CTAs then need a
  // single header read, requested before their tile copy even starts: with 30 MB of bulk copies in flight a small
  // dependent load queues behind them for ~2 us, and three of them in a row (header, coarse bins, fine bins) per
  // sweep CTA were the longest thing in its life (clock64 stamps: 5 us of ~7) ------------------------------------
  DecodeHeader* hdr = hdrs + b;
  __threadfence();
  __syncthreads();
  if (tid == 0) {
    __threadfence();
    s_last = (atomicAdd(&hdr->seed_ticket, 1u) == gridDim.x - 1) ? 1u : 0u;
    __threadfence();
  }
  __syncthreads();
  DEC_TS_MAX(1);
  if (!s_last || warp != 0) return;
  const unsigned int ord = seed_bound_of_image(phist, pvals, n_vals, K, s_sub);
  if (lane == 0) {
    if (ord == 0u) *((volatile unsigned int*)&hdr->staged) = 1u;      // no usable sample bound: the staged path owns the image
    else *((volatile unsigned int*)&hdr->thr_ord) = ord;              // (also the head start of the staged path after an overflow)
  }
  DEC_TS_MIN(2); DEC_TS_MAX(3); DEC_TS_MAX(1);
}

// decode_stream_kernel: the sweep proper -- a lean TMA-staged threshold sweep.  One CTA (4 warps) per tile = a
// band of `rows` full-width rows of one class plane (~32 KB) plus its two halo rows: CONTIGUOUS in memory, so one
// elected thread moves the whole tile into shared memory with ONE 1-D bulk async copy (mbarrier completion; issued
// before griddepcontrol.wait -- the logits do not depend on the seed kernel).  Shared memory, not registers, is
// the in-flight buffer: 6 CTAs x 34 KB = 200 KB per SM on the way at any time, with ~40 registers per thread.
// While the tile is in flight warp 0 derives the image's bound from the sample histogram (two dependent L2 reads
// that cost nothing behind the copy).  After the wait each warp walks its rows in a small ROLLED loop: one
// conflict-free 16-byte shared-memory load per lane and row, "max of the lane's 4 cells >= bound", a vote; the ~1/4
// of the rows where some lane passes run the 3x3 test on the neighbouring rows in shared memory.  Passing cells are
// recorded in a per-lane 64-bit map and written to the TILE's own segment of the key list (count + keys): no list
// reservation, no atomic at all.  One CTA barrier, no bitmap, no histogram, nothing read twice from DRAM (a band's
// halo rows are its neighbours' own rows, in flight at the same time: L2 hits).
// Forms of this kernel that were built, parity-green, and rejected (B = 32, 128x128x80, ncu durations):
//   flat float4 stream, hits re-tested with neighbours through L2 (round 1)                          40.8 us
//   18 rows in registers per warp, fully unrolled: 8 K instructions, instruction-cache-bound
//     (13 "no instruction" stall cycles per issue)                                                   180 us
//   the same with the row test out of line                                                           70 us
//   rolled ring of 8 register rows per lane (2.5 KB in flight per warp, 24 warps / SM: latency)      66 us
//   this kernel with ONE key list per image and a warp-aggregated reservation: ~1200 returning
//     same-address atomics per image, ~75 ns each                                                    87 us
//   per-tile segments, 16 KB tiles, generic strips: issue-bound (755 instructions / warp, 77 %)      38 us
// A segment overflow (ties / plateaus at the bound) hands the image to the staged path, exact for any input.
constexpr int kStrThreads = 128;
constexpr int kStrWarps = kStrThreads / 32;
constexpr int kTileCap = 64;                     // keys per tile segment of the fused head decode (head_decode.cuh)
#ifndef CTDET_DECODE_SW_TILE_CAP
#define CTDET_DECODE_SW_TILE_CAP 16
#endif
constexpr int kSwTileCap = CTDET_DECODE_SW_TILE_CAP;   // keys per tile segment of decode_stream_kernel: ~5 used at K = 100, the rest spills.
// 64 -> 16 (measured 64 / 32 / 16): the tail CTA of an image reads every segment, zeros included -- 327 KB per 512^2 image
// at 64, 1.3 MB per 1024^2 image; B = 32: 53.3 -> 51.2 us, 1024^2 B = 8: 63.5 -> 55.4 us, class-concentrated scores unchanged
constexpr int kSpillCap = 8192;                  // keys per image beyond their tiles' segments (class- / object-concentrated scores)

template <bool kOneStrip>                        // W <= 128: a row is one warp-wide float4 load
__global__ void __launch_bounds__(kStrThreads)
decode_stream_kernel(const float* __restrict__ heat, int C, int H, int W, int band_rows, int bands, int flags, int K,
                     FastDiv div_bands, FastDiv div_strips, DecodeHeader* __restrict__ hdrs,
                     const unsigned int* __restrict__ phists, const unsigned int* __restrict__ pvals_all, int n_vals,
                     u64* __restrict__ skeys_all, u64* __restrict__ spill_all, size_t keys_per_image, int spill_cap) {
  extern __shared__ __align__(128) unsigned char s_raw[];
  __shared__ __align__(8) uint64_t s_bar;
  __shared__ int s_wtot[kStrWarps];
  __shared__ unsigned int s_spill;
  __shared__ unsigned int s_sub[32];
  __shared__ unsigned int s_thr, s_skip;
  float* tile = reinterpret_cast<float*>(s_raw);              // [(band_rows + 2)][W]; tile row t <-> map row y_lo - 1 + t
  const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
  DEC_TS_MIN(4); DEC_TS_MAX(5);
  const int b = blockIdx.y;
  const int c = (int)div_bands.div(blockIdx.x), band = (int)blockIdx.x - c * bands;
  const int y_lo = band * band_rows, y_hi = min(H, y_lo + band_rows);
  const int nrows = y_hi - y_lo;
  const float* plane = heat + ((size_t)b * C + c) * H * W;
  DecodeHeader* hdr = hdrs + b;
  const float NEG = -__int_as_float(0x7f800000);

  const int g_first = max(y_lo - 1, 0), g_last = min(y_hi, H - 1);      // map rows that exist
  if (tid == 0) {
    mbar_init(&s_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    const unsigned bytes = (unsigned)((g_last - g_first + 1) * W * sizeof(float));
    mbar_expect_tx(&s_bar, bytes);
    // The logits are read once: the tile copy carries an L2 evict-first policy, so the sweep does not push out what the
    // neighbouring kernels of a step keep in L2 (measured in the bench step, scripts/step_overlap.py: one stream 160.9 ->
    // 154.7 us, two-branch graph 145.7 -> 144.1 us; evict-last: no change).  -DCTDET_SWEEP_L2_HINT=0 = no hint, 1 = evict-last.
#ifndef CTDET_SWEEP_L2_HINT
#define CTDET_SWEEP_L2_HINT 2
#endif
#if CTDET_SWEEP_L2_HINT != 0
    {
      uint64_t pol;
#if CTDET_SWEEP_L2_HINT == 1
      asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
#else
      asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
#endif
      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;"
                   :: "r"((uint32_t)__cvta_generic_to_shared(tile + (size_t)(g_first - (y_lo - 1)) * W)),
                      "l"(plane + (size_t)g_first * W), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(&s_bar)), "l"(pol) : "memory");
    }
#else
    bulk_g2s(tile + (size_t)(g_first - (y_lo - 1)) * W, plane + (size_t)g_first * W, bytes, &s_bar);
#endif
  }
  if (y_lo == 0) for (int x = tid; x < W; x += kStrThreads) tile[x] = NEG;                      // max_pool2d padding (:401)
  if (y_hi == H) for (int x = tid; x < W; x += kStrThreads) tile[(size_t)(nrows + 1) * W + x] = NEG;
  // NO griddepcontrol.wait: this grid is launched as soon as every seed CTA has started (programmatic dependent launch)
  // and each of its CTAs waits only for the bound of ITS image, which the last seed CTA of that image publishes
  // (thr_ord != 0, or staged = 1 when the sample gives no bound; both are zero on entry).  Image 0 is swept while images
  // 1.. are still sampled, and the seed's ticket + derivation latency hides behind the first tile copies.  Every
  // seed CTA is resident by then, so the wait cannot deadlock; it is bounded anyway (~4 ms, then the image is handed
  // to the staged path, which needs no sample).
  if (tid == 32) {                     // (not the thread that issues the copy)
    unsigned int ord = 0u, st_ = 0u, polls = 0u;
    for (;;) {
      asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(ord) : "l"(&hdr->thr_ord) : "memory");
      st_ = *((volatile unsigned int*)&hdr->staged);          // no sample bound, or an overflow elsewhere
      if (ord != 0u || st_ != 0u) break;
      if (++polls > 65536u) { hdr->staged = 1u; st_ = 1u; break; }
      __nanosleep(64);
    }
    s_thr = ord;
    s_skip = (ord == 0u || st_ != 0u) ? 1u : 0u;
  }
  __syncthreads();                                            // bound, barrier init, padding rows
  DEC_TS_MIN(6);
  const size_t tile_id = (size_t)b * gridDim.x + blockIdx.x;
  const unsigned int thr_ord = s_thr;
  if (s_skip) {                                               // CTA-uniform (the tail ignores this image's segments)
    mbar_wait(&s_bar, 0);                                     // (the copy must land before the CTA's memory is released)
    pdl_launch_dependents();
    return;
  }
  const bool logits = flags & CTDET_DECODE_LOGITS;
  const bool no_nms = flags & CTDET_DECODE_NO_NMS;
  const float thv = fmaxf(ordered_to_float(thr_ord), (logits || no_nms) ? -3.402823466e38f : __int_as_float(1));
  const int strips = kOneStrip ? 1 : (int)div_strips.d;
  const int pairs = nrows * strips;                           // (row, 128-column strip) work items of the tile
  // a warp takes a contiguous run of at most 16 pairs: the per-lane hit map is 64 bits
  const int per_warp = (pairs + kStrWarps - 1) / kStrWarps;
  const int p0 = warp * per_warp, pn = min(per_warp, pairs - p0);
  const bool lane_in = kOneStrip ? (lane * 4 < W) : true;
  const float thl = lane_in ? thv : __int_as_float(0x7f800000);         // lanes beyond the map never pass
  mbar_wait(&s_bar, 0);

  unsigned long long hits = 0ull;
  const float* rowp1 = tile + (size_t)(p0 + 1) * W + (lane_in ? lane * 4 : 0);      // kOneStrip: pair = row
  for (int i = 0; i < pn; ++i) {
    const float* rowp;
    int x0;
    bool in;
    if (kOneStrip) {
      rowp = rowp1 + (size_t)i * W; x0 = lane * 4; in = lane_in;
    } else {
      const int pr = p0 + i;
      const int t = (int)div_strips.div((uint32_t)pr), strip = pr - t * strips;    // tile row t + 1
      x0 = strip * 128 + lane * 4;
      in = x0 < W;
      rowp = tile + (size_t)(t + 1) * W + (in ? x0 : 0);
    }
    const float4 cur = *reinterpret_cast<const float4*>(rowp);
    const float m = fmaxf(fmaxf(cur.x, cur.y), fmaxf(cur.z, cur.w));
    const bool mine = m >= (kOneStrip ? thl : (in ? thv : __int_as_float(0x7f800000)));
    {
      if (mine) {                                             // ~1/4 of the rows, 1-2 lanes of the warp
        // per passing cell (usually one of the four): its 8 neighbours straight from shared memory
        const float v[4] = {cur.x, cur.y, cur.z, cur.w};
        unsigned bits = 0u;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          if (v[j] >= thv && float_to_ordered(v[j]) >= thr_ord) {                 // (-0.0 against a +0.0 bound)
            bool pk = no_nms;
            if (!pk) {
              const float* q = rowp + j;
              float nbm = fmaxf(q[-W], q[W]);
              if (j > 0 || x0 > 0) nbm = fmaxf(nbm, fmaxf(fmaxf(q[-W - 1], q[-1]), q[W - 1]));
              if (j < 3 || x0 + 4 < W) nbm = fmaxf(nbm, fmaxf(fmaxf(q[-W + 1], q[1]), q[W + 1]));
              pk = v[j] >= nbm || (logits && same_prob(v[j], nbm));
            }
            bits |= (unsigned)pk << j;
          }
        }
        hits |= (unsigned long long)bits << (4 * i);
      }
    }
  }
  // ---- emit into the TILE's own segment of the key list; the tail kernel gathers the segments ------------------
  const int mine_n = __popcll(hits);
  int incl = mine_n;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int t = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += t;
  }
  if (lane == 31) s_wtot[warp] = incl;
  __syncthreads();
  int base = 0, total = 0;
#pragma unroll
  for (int w = 0; w < kStrWarps; ++w) { const int t = s_wtot[w]; if (w < warp) base += t; total += t; }
  u64* spill = nullptr;
  if (total > kSwTileCap) {
    // More keys than the tile's segment holds (scores concentrated in a few classes or around a few objects; ties and
    // plateaus at the bound): the first kSwTileCap keys go to the segment, the others to the image's spill list -- ONE
    // returning atomic per overflowing tile.  Only when that list is full too is the image handed to the staged path.
    if (tid == 0) s_spill = atomicAdd(&hdr->s_count, (unsigned int)(total - kSwTileCap));
    __syncthreads();                                          // (CTA-uniform branch)
    if (s_spill + (unsigned int)(total - kSwTileCap) > (unsigned int)spill_cap) {
      if (tid == 0) hdr->staged = 1u;
      pdl_launch_dependents();
      return;
    }
    spill = spill_all + (size_t)b * keys_per_image + s_spill;
  }
  pdl_launch_dependents();
  // the segment is always written in full -- keys, then zeros -- so the tail reads it as one dense, coalesced vector
  // without knowing the counts (no count array, no dependent second trip)
  for (int i = total + tid; i < kSwTileCap; i += kStrThreads) skeys_all[tile_id * kSwTileCap + i] = 0ull;
  int pos = base + (incl - mine_n);                           // this lane's first slot in CTA order
  u64* seg = skeys_all + tile_id * kSwTileCap;
  while (hits) {
    const int bit = __ffsll((long long)hits) - 1;
    hits &= hits - 1ull;
    const int pr = p0 + (bit >> 2);
    const int t = kOneStrip ? pr : (int)div_strips.div((uint32_t)pr), strip = kOneStrip ? 0 : pr - t * strips;
    const int x = strip * 128 + lane * 4 + (bit & 3);
    const float v = tile[(size_t)(t + 1) * W + x];
    const uint32_t flat = (uint32_t)((c * H + (y_lo + t)) * W + x);
    const u64 key = ((u64)float_to_ordered(v) << 32) | (u64)(0xffffffffu - flat);
    if (pos < kSwTileCap) seg[pos] = key; else spill[pos - kSwTileCap] = key;
    ++pos;
  }
  DEC_TS_MAX(7);
}

// ------------------------------------------------------------------------------------------------
// Tail kernel: per image, exact top-K of the emitted keys, sort, gather, boxes -- and, for the images the sweep
// handed over, the staged fall-back itself (no separate, normally empty launch).
constexpr int kSelCap = 2048;
#ifdef CTDET_DECODE_TIMING
__device__ unsigned long long g_tail_ts[64][8];
__device__ __forceinline__ unsigned long long gtime() { return (unsigned long long)clock64(); }   // SM cycles (1.965 GHz)
#define TAIL_TS(i) do { if (threadIdx.x == 0 && blockIdx.x < 64) g_tail_ts[blockIdx.x][i] = gtime(); } while (0)
#else
#define TAIL_TS(i) do {} while (0)
#endif
constexpr int kTailHelpers = 4;                  // CTAs per image that share the tiles of a handed-over image
constexpr int kTailSmemKeys = 6144;              // keys of an image held in shared memory by its tail CTA (48 KB)

struct FinShared {
  SelectScratch sc;
  BinScratch bs;
  __align__(16) u64 sel[kSelCap];
  unsigned int h2[kHistBins];
  int rank[kFinThreads];
  int n, above;
  unsigned int last, maxord;
};

// Output stage shared by both list sources: fs->sel holds fs->n >= min(#candidates, K) keys that contain the K best.
// One detection from its key: score, gathered wh / reg, box in the reference's fp32 operation order.
struct DetGather { float rx, ry, w, h; };
__device__ __forceinline__ DetGather det_gather(u64 key, int b, const float* __restrict__ wh, const float* __restrict__ reg,
                                                int C, int HW, int flags) {
  const uint32_t flat = 0xffffffffu - (uint32_t)key;
  const int cls = (int)(flat / (uint32_t)HW);
  const int sp = (int)(flat - (uint32_t)cls * HW);
  DetGather g;
  g.rx = g.ry = 0.5f;                                              // :440-441
  if (reg != nullptr) {
    g.rx = __ldg(reg + ((size_t)b * 2 + 0) * HW + sp);             // :436-439
    g.ry = __ldg(reg + ((size_t)b * 2 + 1) * HW + sp);
  }
  if (flags & CTDET_DECODE_CAT_SPEC_WH) {                          // :443-446
    g.w = __ldg(wh + ((size_t)b * 2 * C + 2 * cls) * HW + sp);
    g.h = __ldg(wh + ((size_t)b * 2 * C + 2 * cls + 1) * HW + sp);
  } else {
    g.w = __ldg(wh + ((size_t)b * 2 + 0) * HW + sp);
    g.h = __ldg(wh + ((size_t)b * 2 + 1) * HW + sp);
  }
  return g;
}
__device__ __forceinline__ void det_store(u64 key, const DetGather& g, int b, int k, int K, int HW, int W, float down_ratio,
                                          int flags, float* __restrict__ boxes, float* __restrict__ scores,
                                          int32_t* __restrict__ classes, int64_t* __restrict__ inds) {
  const uint32_t ord = (uint32_t)(key >> 32);
  const uint32_t flat = 0xffffffffu - (uint32_t)key;
  const int cls = (int)(flat / (uint32_t)HW);
  const int sp = (int)(flat - (uint32_t)cls * HW);
  float score = 0.f;
  if (ord != 0u) {
    score = ordered_to_float(ord);
    if (flags & CTDET_DECODE_LOGITS) score = prob_of_logit(score);   // centernet.py:161
  }
  float xs = (float)(sp % W), ys = (float)(sp / W);                // :414-415
  if (flags & 0x40000000) { xs += 0.5f; ys += 0.5f; }              // (reg == nullptr, :440-441)
  else { xs = __fadd_rn(xs, g.rx); ys = __fadd_rn(ys, g.ry); }     // :436-439
  const float hw_ = g.w * 0.5f, hh_ = g.h * 0.5f;                  // "/ 2" is exact
  float* bx = boxes + ((size_t)b * K + k) * 4;                     // :452-455
  bx[0] = __fmul_rn(__fsub_rn(xs, hw_), down_ratio);
  bx[1] = __fmul_rn(__fsub_rn(ys, hh_), down_ratio);
  bx[2] = __fmul_rn(__fadd_rn(xs, hw_), down_ratio);
  bx[3] = __fmul_rn(__fadd_rn(ys, hh_), down_ratio);
  scores[(size_t)b * K + k] = score;
  classes[(size_t)b * K + k] = cls;                                // int32 (:418)
  if (inds != nullptr) inds[(size_t)b * K + k] = sp;
}

// Output stage shared by both list sources: fs->sel holds fs->n >= min(#candidates, K) keys that contain the K best.
__device__ void finalize_emit(FinShared* fs, int b, const float* __restrict__ wh, const float* __restrict__ reg,
                              int C, int H, int W, int K, float down_ratio, int flags, float* __restrict__ boxes,
                              float* __restrict__ scores, int32_t* __restrict__ classes, int64_t* __restrict__ inds) {
  const int tid = threadIdx.x;
  const int HW = H * W;
  const int sflags = flags | (reg == nullptr ? 0x40000000 : 0);
  u64* s_sel = fs->sel;
  int n = min(fs->n, kSelCap);                            // >= min(#peaks, K)
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
  if (n <= kFinThreads) {
    // rank sort: keys are distinct, so rank = number of larger keys.  All 512 threads work: with
    // P2 = next power of two >= n, thread t owns key t % P2 and compares it against slice t / P2
    // of the list (two keys per 16-byte shared-memory broadcast load); partial ranks are combined
    // with shared-memory atomics.  No barriers inside the compare loop (a bitonic network of this
    // size costs ~36 __syncthreads).  The owner of a key (slice 0) requests its wh / reg values BEFORE the compare
    // loop -- their L2 round trip hides behind the ranking -- and writes the detection at its rank itself.
    int* s_rank = fs->rank;
    int P2 = 32; while (P2 < n) P2 <<= 1;
    const int parts = kFinThreads / P2, idx = tid & (P2 - 1), part = tid / P2;
    s_rank[tid] = 0;
    if (n & 1) { if (tid == 0) s_sel[n] = 0ull; }          // pad to an even count for the paired loads
    __syncthreads();
    const int npair = (n + 1) >> 1;
    const int j_lo = (int)(((long long)part * npair) / parts), j_hi = (int)(((long long)(part + 1) * npair) / parts);
    u64 mine = 0ull;
    DetGather g = {0.f, 0.f, 0.f, 0.f};
    if (idx < n) {
      mine = s_sel[idx];
      if (part == 0) g = det_gather(mine, b, wh, reg, C, HW, flags);
      int rank = 0;
      const ulonglong2* pairs = reinterpret_cast<const ulonglong2*>(s_sel);
#pragma unroll 4
      for (int j2 = j_lo; j2 < j_hi; ++j2) {
        const ulonglong2 kk = pairs[j2];
        rank += (kk.x > mine) + (kk.y > mine);
      }
      if (rank) atomicAdd(&s_rank[idx], rank);
    }
    __syncthreads();
    if (part == 0 && idx < n) {
      const int rank = s_rank[idx];
      if (rank < K) det_store(mine, g, b, rank, K, HW, W, down_ratio, sflags, boxes, scores, classes, inds);
    }
    return;
  }
  // bitonic sort, descending, padded to a power of two with zeros; the first K are the answer
  {
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
  }
  for (int k = tid; k < K; k += kFinThreads) {
    const u64 key = s_sel[k];
    const DetGather g = det_gather(key, b, wh, reg, C, HW, flags);
    det_store(key, g, b, k, K, HW, W, down_ratio, sflags, boxes, scores, classes, inds);
  }
}

// Two-level threshold over a key list `kb[0..n_keys)`: level 1 = 2048 bins over the top 11 bits of the ordered
// score (hist1; every key counted once), level 2 = the next 11 bits inside the threshold bin, histogrammed in
// shared memory.  Leaves in fs->sel the keys >= the refined threshold: K plus a handful; the exact radix select
// is the fall-back when ties keep the count above kSelCap.  scan(f) calls f(key) for every key (0 = none).
template <bool kSharedHist, typename Scan>
__device__ void finalize_select(FinShared* fs, const unsigned int* hist1, Scan scan, const u64* kb, int n_keys, int K) {
  const int tid = threadIdx.x;
  find_threshold_bin<kFinThreads, kSharedHist>(hist1, K, &fs->bs);
  u64 T = ((u64)fs->bs.bin << kHistShift) << 32;
  if (fs->bs.cnt_ge < (unsigned)K) T = 0ull;              // fewer than K candidates in the whole image
  if (T != 0ull && fs->bs.cnt_ge > (unsigned)kFinThreads) {
    // The 11-bit bins are coarse near the top (sign + exponent + 2 mantissa bits: all of [0.875, 1) is one
    // bin, ~3800 peaks of a 512^2 image).  Refine once on the NEXT 11 bits of the keys inside the bin.
    unsigned int* s_h2 = fs->h2;
    const uint32_t tbin = (uint32_t)fs->bs.bin;
    for (int i = tid; i < kHistBins; i += kFinThreads) s_h2[i] = 0u;
    if (tid == 0) fs->above = 0;
    __syncthreads();
    int above = 0;
    scan([&](u64 key) {
      const uint32_t ord = (uint32_t)(key >> 32), ob = ord >> kHistShift;
      if (key == 0ull || ob < tbin) return;
      if (ob > tbin) ++above;
      else atomicAdd(&s_h2[(ord >> (kHistShift - kHistBits)) & (kHistBins - 1)], 1u);
    });
    above = warp_sum(above);
    if ((tid & 31) == 0 && above) atomicAdd(&fs->above, above);
    __syncthreads();
    const int hi = fs->above;                             // < K: that is what made tbin the threshold bin
    find_threshold_bin<kFinThreads, true>(s_h2, K - hi, &fs->bs);
    if ((unsigned)hi + fs->bs.cnt_ge <= (unsigned)kSelCap) {
      T = (u64)((tbin << kHistShift) | ((uint32_t)fs->bs.bin << (kHistShift - kHistBits))) << 32;
    } else {
      T = radix_select_kth<kFinThreads>(
          [&](auto f) { for (int i = tid; i < n_keys; i += kFinThreads) f(kb[i]); }, n_keys, K, &fs->sc);
    }
  }
  __syncthreads();
  scan([&](u64 key) {
    if (key != 0ull && key >= T) { const int pos = atomicAdd(&fs->n, 1); if (pos < kSelCap) fs->sel[pos] = key; }
  });
}

// (out of line on purpose: the tail kernel runs once, cold -- its time is instruction fetch, so the rarely taken
// fall-back must not sit in the middle of the hot path)
__device__ __noinline__ void
tail_fallback_tile(const float* __restrict__ heat, int C, int H, int W, int Wp, int K, int band_rows, int sub_rows,
                   FastDiv div_strips, int flags, DecodeHeader* __restrict__ hdrs, unsigned int* __restrict__ ghists,
                   u64* __restrict__ out_keys, size_t keys_per_image, int band, int b, int c, unsigned char* s_raw, bool reinit) {
  decode_peaks_tile<kFinThreads>(heat, C, H, W, Wp, K, band_rows, sub_rows, div_strips, flags, hdrs, ghists, out_keys,
                                 keys_per_image, band, b, c, s_raw, reinit);
}

// grid = (B + (helpers - 1) * B): CTAs [0, B) are the images' own tail CTAs (scheduled first), the others
// are helpers that return at once unless their image was handed over.
//   mode 0 (behind decode_stream_kernel): image not handed over -> CTA b copies the image's key list (<= 16 K
//     keys, ~3 K typically) into shared memory ONCE, histograms it there and selects -- no global histogram, no
//     second trip to L2.  Image handed over (hdr->staged) -> its `helpers` CTAs run the staged tiles of the image
//     (decode_peaks_tile, exact for any input), and the last of them to finish (ticket) finalizes from that list.
//   mode 1 (behind decode_peaks_kernel, shapes the sweep does not take): finalize from the staged list only.
__global__ void __launch_bounds__(kFinThreads, 1)
decode_tail_kernel(const float* __restrict__ heat, int B, int mode, int helpers,
                   DecodeHeader* __restrict__ hdrs, unsigned int* __restrict__ ghists,
                   u64* __restrict__ keys, size_t keys_per_image,
                   const u64* __restrict__ skeys, int sw_tiles, int tile_cap, int spill_cap,
                   unsigned int* __restrict__ phists,
                   int Wp, int band_rows, int sub_rows, FastDiv div_strips, int bands,
                   const float* __restrict__ wh, const float* __restrict__ reg, int C, int H,
                   int W, int K, float down_ratio, int flags, float* __restrict__ boxes,
                   float* __restrict__ scores, int32_t* __restrict__ classes,
                   int64_t* __restrict__ inds) {
  extern __shared__ __align__(128) unsigned char s_dyn[];   // staged tile | the image's key list (mode 0)
  __shared__ FinShared fs;
  const int tid = threadIdx.x;
  const int b = blockIdx.x < B ? blockIdx.x : (blockIdx.x - B) % B;
  const int helper = blockIdx.x < B ? 0 : 1 + (blockIdx.x - B) / B;
  DecodeHeader* hdr = hdrs + b;
  TAIL_TS(0);
  DEC_TS_MIN(8);
  pdl_wait();
  TAIL_TS(1);
  DEC_TS_MIN(9); DEC_TS_MAX(10);
  // The image's key segments (every tile's segment is fully written: keys, then zeros) are ONE dense vector: the first
  // batch of 16-byte loads is requested together with the header -- independent loads, one L2 trip.
#ifndef CTDET_TAIL_BATCH
#define CTDET_TAIL_BATCH 5
#endif
  constexpr int kBatch = CTDET_TAIL_BATCH;                   // uint4 (2 keys) per thread and batch (40 KB per CTA and trip); measured 20 / 10 / 5 / 4 / 2:
  // B = 32 52.2 / 49.9 / 49.3 / 49.7 / 50.0 us -- the tail runs its code ONCE per CTA and waits for instruction fetch, so the
  // shorter unrolled body wins over fewer trips
  const int n_vec = (mode == 0 && helper == 0) ? sw_tiles * (tile_cap / 2) : 0;
  const uint4* segv = reinterpret_cast<const uint4*>(skeys + (mode == 0 ? (size_t)b * sw_tiles * tile_cap : 0));
  uint4 kq[kBatch];
#pragma unroll
  for (int u = 0; u < kBatch; ++u) {
    const int i = u * kFinThreads + tid;
    kq[u] = make_uint4(0u, 0u, 0u, 0u);
    if (i < n_vec) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(kq[u].x), "=r"(kq[u].y), "=r"(kq[u].z), "=r"(kq[u].w) : "l"(segv + i));
  }
  // which path produced this image's list (CTA-uniform; the header is reset only at the very end)
  const bool staged = (mode == 1) || hdr->staged != 0u;
  if (!staged && helper != 0) return;
  const unsigned int thr_lo = *((volatile unsigned int*)&hdr->thr_ord);   // the sweep's bound (every streamed key is >= it)
  if (tid == 0) { fs.n = 0; fs.last = 1u; }
  if (staged && mode == 0) {
    // ---- exact fall-back: this image's tiles, shared by its `helpers` CTAs -----------------------------
    bool reinit = false;
    const int tiles = bands * C;
    for (int t = helper; t < tiles; t += helpers) {
      const int band = t % bands, c = t / bands;
      __syncthreads();                                       // the previous tile's shared state is dead
      tail_fallback_tile(heat, C, H, W, Wp, K, band_rows, sub_rows, div_strips, flags, hdrs, ghists, keys,
                         keys_per_image, band, b, c, s_dyn, reinit);
      reinit = true;
    }
    __threadfence();                                         // every thread's keys / histogram counts, then the ticket
    __syncthreads();
    if (tid == 0) {
      __threadfence();
      fs.last = (atomicAdd(&hdr->ticket, 1u) == (unsigned)(helpers - 1)) ? 1u : 0u;
      __threadfence();
    }
  }
  __syncthreads();
  if (!fs.last) return;                                      // CTA-uniform: another CTA of the image finalizes

  if (!staged) {
    // ---- streaming path: the image's keys sit in per-tile segments, read as one dense vector (zeros = empty slots)
    // and compacted into shared memory: per batch each warp reserves its slots with ONE shared-memory atomic.
    u64* s_keys = reinterpret_cast<u64*>(s_dyn);
    unsigned int* s_h1 = reinterpret_cast<unsigned int*>(s_keys + kTailSmemKeys);
    const int lane = tid & 31;
    for (int i = tid; i < kHistBins; i += kFinThreads) s_h1[i] = 0u;
    if (tid == 0) { fs.above = 0; fs.maxord = 0u; }
    __syncthreads();
    unsigned int mx = 0u;
    for (int v0 = 0; v0 < n_vec; v0 += kBatch * kFinThreads) {
      if (v0 > 0) {                                           // (maps with more than 320 tiles per image: further batches)
#pragma unroll
        for (int u = 0; u < kBatch; ++u) {
          const int i = v0 + u * kFinThreads + tid;
          kq[u] = make_uint4(0u, 0u, 0u, 0u);
          if (i < n_vec) asm volatile("ld.global.cg.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(kq[u].x), "=r"(kq[u].y), "=r"(kq[u].z), "=r"(kq[u].w) : "l"(segv + i));
        }
      }
      int cnt = 0;
#pragma unroll
      for (int u = 0; u < kBatch; ++u) cnt += ((kq[u].x | kq[u].y) != 0u) + ((kq[u].z | kq[u].w) != 0u);
      if (v0 == 0 && cnt >= 0) TAIL_TS(2);
      int incl = cnt;
#pragma unroll
      for (int o = 1; o < 32; o <<= 1) { const int t = __shfl_up_sync(0xffffffffu, incl, o); if (lane >= o) incl += t; }
      int wbase = 0;
      if (lane == 31 && incl) wbase = atomicAdd(&fs.above, incl);
      int slot = __shfl_sync(0xffffffffu, wbase, 31) + incl - cnt;
#pragma unroll
      for (int u = 0; u < kBatch; ++u) {
        if ((kq[u].x | kq[u].y) != 0u) { if (slot < kTailSmemKeys) s_keys[slot] = ((u64)kq[u].y << 32) | kq[u].x; ++slot; mx = max(mx, kq[u].y); }
        if ((kq[u].z | kq[u].w) != 0u) { if (slot < kTailSmemKeys) s_keys[slot] = ((u64)kq[u].w << 32) | kq[u].z; ++slot; mx = max(mx, kq[u].w); }
      }
    }
    // the image's spill list (keys of tiles whose segment was full; normally empty): it lives at the start of the image's
    // staged-list region, which only a handed-over image would use
    const u64* spill = keys + (size_t)b * keys_per_image;
    const int n_spill = (int)min(*((volatile unsigned int*)&hdr->s_count), (unsigned int)spill_cap);
    for (int i = tid; i < n_spill; i += kFinThreads) {
      const u64 key = __ldcg(spill + i);
      const int slot = atomicAdd(&fs.above, 1);
      if (slot < kTailSmemKeys) s_keys[slot] = key;
      mx = max(mx, (unsigned int)(key >> 32));
    }
    mx = __reduce_max_sync(0xffffffffu, mx);
    if (lane == 0 && mx) atomicMax(&fs.maxord, mx);
    __syncthreads();
    TAIL_TS(3);
    const int n_keys = fs.above;                             // the slot counter = number of keys
    const bool fits = n_keys <= kTailSmemKeys;               // CTA-uniform
    const u64* segs = skeys + (size_t)b * sw_tiles * tile_cap;
    // (only when the list does not fit shared memory: f(key) for every key, straight from the segments and the spill list)
    auto scan_tiles = [&](auto f) {
      for (int i = tid; i < sw_tiles * tile_cap; i += kFinThreads) { const u64 key = __ldcg(segs + i); if (key) f(key, 0); }
      for (int i = tid; i < n_spill; i += kFinThreads) f(__ldcg(spill + i), 0);
    };
    __syncthreads();
    if (tid == 0) fs.above = 0;
    if (n_keys <= kFinThreads) {                             // small list: the rank sort takes it whole
      for (int i = tid; i < n_keys; i += kFinThreads) fs.sel[i] = s_keys[i];
      if (tid == 0) fs.n = n_keys;
    } else {
      // Every key is >= the sweep's bound, and the K-th best sits a little above it: histogram the keys over
      // 2048 bins of (ordered score - bound) >> sh, sh chosen from the largest score so that the bins span
      // exactly [bound, max] -- the keys spread over the whole histogram instead of piling into the 2-3
      // quarter-octave bins of the plain top bits (shared-memory atomics on one address serialise).
      const unsigned int lo = thr_lo;
      const unsigned int range = fs.maxord - lo;
      int sh = 0;
      while ((range >> sh) >= (unsigned)kHistBins) ++sh;
      auto scan = [&](auto f) {
        if (fits) { for (int i = tid; i < n_keys; i += kFinThreads) f(s_keys[i]); }
        else scan_tiles([&](u64 key, int) { f(key); });
      };
      scan([&](u64 key) { atomicAdd(&s_h1[((unsigned int)(key >> 32) - lo) >> sh], 1u); });
      __syncthreads();
      find_threshold_bin<kFinThreads, true>(s_h1, K, &fs.bs);
      u64 T = 0ull;
      if (fs.bs.cnt_ge >= (unsigned)K) T = (u64)(lo + ((unsigned int)fs.bs.bin << sh)) << 32;
      if (fs.bs.cnt_ge > (unsigned)kSelCap)                  // ties: too many keys share the threshold bin
        T = radix_select_kth<kFinThreads>([&](auto f) { scan([&](u64 key) { f(key); }); }, n_keys, K, &fs.sc);
      __syncthreads();
      scan([&](u64 key) {
        if (key >= T) { const int pos = atomicAdd(&fs.n, 1); if (pos < kSelCap) fs.sel[pos] = key; }
      });
    }
  } else {
    // ---- staged list (global memory; its per-image histogram was filled by the tiles) ---------------------
    const u64* kb = keys + (size_t)b * keys_per_image;
    unsigned int* ghist = ghists + (size_t)b * kHistBins;
    const int n_keys = (int)*((volatile unsigned int*)&hdr->count);
    // The list is read by ONE CTA: every pass keeps 8 loads per thread in flight (a plain loop exposes
    // one L2 round trip per 512 keys).
    auto scan = [&](auto f) {
      for (int i0 = 0; i0 < n_keys; i0 += 8 * kFinThreads) {
        u64 kk[8];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int i = i0 + u * kFinThreads + tid;
          kk[u] = i < n_keys ? __ldcg(kb + i) : 0ull;       // real keys are never 0 (low word = ~flat index)
        }
#pragma unroll
        for (int u = 0; u < 8; ++u) f(kk[u]);
      }
    };
    finalize_select<false>(&fs, ghist, scan, kb, n_keys, K);
    for (int i = tid; i < kHistBins; i += kFinThreads) ghist[i] = 0u;     // leave the workspace zeroed
  }
  TAIL_TS(4);
  if (phists != nullptr)
    for (int i = tid; i < kSeedHist; i += kFinThreads) phists[(size_t)b * kSeedHist + i] = 0u;   // sample histogram
  __syncthreads();
  TAIL_TS(5);
  if (tid == 0) { hdr->count = 0u; hdr->thr_ord = 0u; hdr->s_count = 0u; hdr->staged = 0u; hdr->ticket = 0u; hdr->seed_ticket = 0u; }
  finalize_emit(&fs, b, wh, reg, C, H, W, K, down_ratio, flags, boxes, scores, classes, inds);
  __syncthreads();
  TAIL_TS(6);
  DEC_TS_MAX(11);
}

struct DecodePlan {
  int band_rows, bands, tiles_per_image, Wp, strips, sub_rows;
  size_t smem, keys_per_image, header_bytes, hist_bytes, total_bytes;
  // streaming path
  bool stream;
  int seed_groups, seed_ctas, seed_vals;
  int sw_rows, sw_bands; size_t sw_smem;          // sweep: rows per tile, tiles per class plane, shared memory per CTA
  int tail_sub_rows; size_t tail_smem;            // tail kernel (512 threads)
  size_t skeys_bytes, scnt_bytes, phist_bytes, pvals_bytes;
};

static DecodePlan make_plan(int B, int C, int H, int W, int K) {
  DecodePlan p;
  p.Wp = (W + 3) / 4 * 4;                       // shared-memory row stride (floats)
  // band = as many rows as fit ~16 KB of shared memory, at least 8; enough CTAs to fill the machine
  static int band_kb = 0;                       // tuning knob: CTDET_DECODE_BAND_KB (default 16: measured best of 8/16/32/64 on B200)
  if (band_kb == 0) { const char* e = getenv("CTDET_DECODE_BAND_KB"); band_kb = e ? atoi(e) : 16; if (band_kb < 1) band_kb = 16; }
  int rows = (band_kb * 1024) / (p.Wp * 4);
  if (rows < 8) rows = 8;
  if (rows > H) rows = H;
  while ((long long)B * C * ((H + rows - 1) / rows) < 4LL * kNumSMs && rows > 8) rows = (rows + 1) / 2;
  p.band_rows = rows;
  p.bands = (H + rows - 1) / rows;
  p.tiles_per_image = C * p.bands;
  const int cap = (K + kKeepSlack + 1) & ~1;
  const int strips = (p.Wp + 127) / 128;
  p.strips = strips;
  int subs = kDecWarps / strips; if (subs < 1) subs = 1;
  p.sub_rows = (rows + subs - 1) / subs;           // rows of a band walked by one warp
  p.smem = (size_t)(rows + 2) * p.Wp * sizeof(float) + (size_t)cap * sizeof(u64) +
           (size_t)rows * strips * 4 * sizeof(uint32_t);
  p.keys_per_image = (size_t)p.tiles_per_image * cap;
  p.header_bytes = (((size_t)B * sizeof(DecodeHeader)) + 255) / 256 * 256;
  p.hist_bytes = (size_t)B * kHistBins * sizeof(unsigned int);
  // streaming path: flat float4 view of an image + interior sample rows; CTDET_DECODE_STREAM=0 disables it
  static int stream_on = -1;
  if (stream_on < 0) { const char* e = getenv("CTDET_DECODE_STREAM"); stream_on = e ? (atoi(e) != 0) : 1; }
  p.seed_groups = (H - 2 - kSeedRows) >= 0 ? (H - 2 - kSeedRows) / kSeedPeriod + 1 : 0;   // r0 + kSeedRows <= H - 1
  p.stream = stream_on && (W % 4) == 0 && W >= 8 && p.seed_groups > 0 && strips <= 64;   // (+ the tile count, below)
  {
    static int sw_kb_env = -1;                                  // tuning knob: CTDET_DECODE_TILE_KB (read once; 0 = automatic)
    if (sw_kb_env < 0) { const char* e = getenv("CTDET_DECODE_TILE_KB"); sw_kb_env = e ? atoi(e) : 0; if (sw_kb_env < 0) sw_kb_env = 0; }
    // ~16 KB of full-width rows per tile (measured best of 8 / 16 / 32 / 64 at W = 128: 66 / 58 / 59 / 59 us); rows of
    // 1 KB and more get 32 KB (W = 256, B = 8: 82 / 70 / 63 / 63 us) -- a tile should hold a few dozen rows per halo pair
    const int sw_kb = sw_kb_env ? sw_kb_env : (W >= 256 ? 32 : 16);
    int sr = (sw_kb * 1024) / (W * 4);
    if (sr * strips > 64) sr = 64 / strips;                     // <= 16 (row, strip) pairs per warp
    if (sr < 1) sr = 1;
    if (sr > H) sr = H;
    p.sw_rows = sr;
    p.sw_bands = (H + sr - 1) / sr;
    p.sw_smem = (size_t)(sr + 2) * W * sizeof(float);
    if ((long long)C * p.sw_bands * kSwTileCap >= (1ll << 28)) p.stream = false;
  }
  {
    int tsubs = (kFinThreads / 32) / strips; if (tsubs < 1) tsubs = 1;
    p.tail_sub_rows = (rows + tsubs - 1) / tsubs;
    const size_t list = (size_t)kTailSmemKeys * sizeof(u64) + (size_t)kHistBins * sizeof(unsigned int);
    p.tail_smem = p.stream ? (p.smem > list ? p.smem : list) : 16;
  }
  {
    const int items = C * p.seed_groups * strips;               // one warp-item = 6 row segments of 512 B
    // one item per warp while the seed grid is about one wave (4 CTAs of 16 K registers per SM; B = 32: 640 CTAs); two per
    // warp beyond that -- the sweep's CTAs only find room as seed CTAs leave (B = 64: 79.8 -> 77.9 us; B = 32: 52.2 -> 53.3 us
    // if applied there, the second item of a warp delays its image's bound)
    int ipw = CTDET_SEED_ITEMS_PER_WARP;
    if ((long long)B * ((items + kSeedWarps - 1) / kSeedWarps) >= 1024) ipw *= 2;
    p.seed_ctas = (items + kSeedWarps * ipw - 1) / (kSeedWarps * ipw);
    p.seed_vals = items * kSeedRows;
    if (p.seed_ctas < 1) p.seed_ctas = 1;
  }
  p.scnt_bytes = 0;                                             // (tile counts of an earlier layout)
  p.skeys_bytes = p.stream ? (size_t)B * C * p.sw_bands * kSwTileCap * sizeof(u64) : 0;
  // sample histograms (zero on entry, left zero: a fixed stride per image, whatever the map shape, because calls
  // of different shapes share one workspace), then the samples themselves (rewritten by every call)
  p.phist_bytes = (((size_t)B * kSeedHist * sizeof(unsigned int)) + 255) / 256 * 256;
  p.pvals_bytes = (((size_t)B * p.seed_vals * sizeof(unsigned int)) + 255) / 256 * 256;
  // layout: headers | staged histograms | streaming histograms | sample histograms | samples | streaming keys | staged keys
  // The zero-on-entry regions (headers, the three histogram blocks) come first, at offsets that depend on B only:
  // calls of different map shapes -- with or without the streaming path -- share one workspace per B, and the keys
  // of one layout must never land in what another layout expects to be zero.
  if (!p.stream) { p.pvals_bytes = 0; p.skeys_bytes = 0; p.scnt_bytes = 0; }
  p.total_bytes = p.header_bytes + 2 * p.hist_bytes + p.phist_bytes + p.pvals_bytes + p.scnt_bytes + p.skeys_bytes +
                  (size_t)B * p.keys_per_image * sizeof(u64);
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
  if ((W % 4) == 0 && !aligned16(heat)) return CTDET_EALIGN;     // bulk copies need 16-byte alignment
  if (reinterpret_cast<uintptr_t>(workspace) & 7u) return CTDET_EALIGN;
  const DecodePlan p = make_plan(B, C, H, W, K);
  if (workspace_bytes < p.total_bytes) return CTDET_EWORKSPACE;
  if ((long long)B * p.tiles_per_image > 0x7fffffffLL) return CTDET_EINVAL;
  if (p.smem > 200 * 1024) return CTDET_EUNSUPPORTED;          // W > ~5000 columns
  cudaStream_t st = (cudaStream_t)stream;
  {
    // (the attribute is per function and sticky; setting it costs ~1 us of host time, not a launch)
    cudaError_t e = cudaFuncSetAttribute(decode_peaks_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.smem);
    if (e == cudaSuccess)
      e = cudaFuncSetAttribute(decode_tail_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.tail_smem);

    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  }
  unsigned char* base = reinterpret_cast<unsigned char*>(workspace);
  DecodeHeader* hdr = reinterpret_cast<DecodeHeader*>(base);
  unsigned int* ghist = reinterpret_cast<unsigned int*>(base + p.header_bytes);
  unsigned char* q = base + p.header_bytes + 2 * p.hist_bytes;    // (the second histogram block is unused since r2)
  unsigned int* phist = reinterpret_cast<unsigned int*>(q); q += p.phist_bytes;
  unsigned int* pvals = reinterpret_cast<unsigned int*>(q); q += p.pvals_bytes;
  q += p.scnt_bytes;                                            // (the tile-count array of an earlier layout: unused)
  u64* skeys = reinterpret_cast<u64*>(q); q += p.skeys_bytes;
  u64* keys = reinterpret_cast<u64*>(q);
  if (!p.stream) { phist = nullptr; pvals = nullptr; skeys = nullptr; }
  if (B > 65535 || C > 65535) return CTDET_EINVAL;
  int rc;
  const FastDiv div_strips((uint32_t)p.strips);
  if (p.stream) {
    // seed (sample bound) -> sweep -> tail, chained with programmatic dependent launch
    decode_seed_kernel<<<dim3(p.seed_ctas, B), kSeedThreads, 0, st>>>(
        heat, C, H, W, K, flags, p.seed_groups, div_strips, FastDiv((uint32_t)p.seed_groups), hdr, phist, pvals, p.seed_vals);
    if ((rc = check_launch())) return rc;
    auto sweep = p.strips == 1 ? decode_stream_kernel<true> : decode_stream_kernel<false>;
    cudaError_t e = cudaFuncSetAttribute(sweep, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)p.sw_smem);
    if (e == cudaSuccess)
      e = launch_pdl(sweep, dim3(C * p.sw_bands, B), dim3(kStrThreads), p.sw_smem, st,
                     heat, C, H, W, p.sw_rows, p.sw_bands, flags, K, FastDiv((uint32_t)p.sw_bands), div_strips, hdr,
                     (const unsigned int*)phist, (const unsigned int*)pvals, p.seed_vals, skeys, keys, p.keys_per_image,
                     (int)std::min<size_t>(kSpillCap, p.keys_per_image));
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
    if ((rc = check_launch())) return rc;
    e = launch_pdl(decode_tail_kernel, dim3(B * kTailHelpers), dim3(kFinThreads), p.tail_smem, st,
                   heat, B, 0, kTailHelpers, hdr, ghist, keys, p.keys_per_image, (const u64*)skeys, C * p.sw_bands, kSwTileCap,
                   (int)std::min<size_t>(kSpillCap, p.keys_per_image), phist,
                   p.Wp, p.band_rows, p.tail_sub_rows, div_strips, p.bands,
                   wh, reg, C, H, W, K, down_ratio, flags, boxes, scores, classes, inds);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
    return check_launch();
  }
  decode_peaks_kernel<<<dim3(p.bands, B, C), kDecThreads, p.smem, st>>>(
      heat, C, H, W, p.Wp, K, p.band_rows, p.sub_rows, div_strips, flags, hdr, ghist, keys, p.keys_per_image);
  if ((rc = check_launch())) return rc;
  {
    cudaError_t e = launch_pdl(decode_tail_kernel, dim3(B), dim3(kFinThreads), p.tail_smem, st,
                               heat, B, 1, 1, hdr, ghist, keys, p.keys_per_image, (const u64*)nullptr, 0, kSwTileCap, 0, (unsigned int*)nullptr,
                               p.Wp, p.band_rows, p.tail_sub_rows, div_strips, p.bands,
                               wh, reg, C, H, W, K, down_ratio, flags, boxes, scores, classes, inds);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  }
  return check_launch();
}

#ifdef CTDET_DECODE_TIMING
extern "C" int ctdet_debug_dec_ts(unsigned long long* out /* [16] host */, int reset) {
  if (reset) {
    unsigned long long z[16];
    for (int i = 0; i < 16; ++i) z[i] = (i == 0 || i == 2 || i == 4 || i == 6 || i == 8 || i == 9) ? ~0ull : 0ull;   // min slots
    return cudaMemcpyToSymbol(ctdet::g_dec_ts, z, sizeof(z)) == cudaSuccess ? 0 : 1;
  }
  return cudaMemcpyFromSymbol(out, ctdet::g_dec_ts, sizeof(unsigned long long) * 16) == cudaSuccess ? 0 : 1;
}
extern "C" int ctdet_debug_tail_times(unsigned long long* out /* [64][8] host */) {
  return cudaMemcpyFromSymbol(out, ctdet::g_tail_ts, sizeof(unsigned long long) * 64 * 8) == cudaSuccess ? 0 : 1;
}
#endif

// ---------------------------------------------------------------------------------------------------------------
// ctdet_decode_static: ctdet_decode for callers that cannot allocate device memory (code under CUDA graph
// capture, tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The decode workspace lives in a file-scope device array of
// CTDET_DECODE_STATIC_BYTES that every GPU loading the library holds until exit.  Nothing on this path calls
// cudaMalloc / cudaFree or an allocating library routine: one cudaGetSymbolAddress per device (set-up), one
// cudaMemsetAsync per call, then the launches of ctdet_decode.
//
// In HOST code `g_decode_static_ws` is only a stand-in the runtime maps to the device symbol -- not a device
// address.  The address the kernels are given comes from cudaGetSymbolAddress, resolved once per device with that
// device current (each device holds its own copy of the array at its own address).
#include <cstdio>
namespace ctdet {

static __device__ __align__(256) unsigned char g_decode_static_ws[CTDET_DECODE_STATIC_BYTES];

constexpr int kStaticMaxDevices = 64;
static void* g_decode_static_addr[kStaticMaxDevices];      // device address of the array on device i; nullptr = not resolved yet
static int g_decode_static_lock = 0;                       // spin lock around the table (set-up only)

// -> CTDET_OK and *out = device address of the scratch array on the CURRENT device
static int decode_static_address(void** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  if (dev < 0 || dev >= kStaticMaxDevices) {
    set_last_error_text("ctdet_decode_static: device ordinal beyond the 64 this library keeps a scratch address for");
    return CTDET_EUNSUPPORTED;
  }
  while (__atomic_exchange_n(&g_decode_static_lock, 1, __ATOMIC_ACQUIRE)) {}
  void* p = g_decode_static_addr[dev];
  int rc = CTDET_OK;
  if (p == nullptr) {
    e = cudaGetSymbolAddress(&p, g_decode_static_ws);
    if (e != cudaSuccess || p == nullptr) {
      if (e != cudaSuccess) set_last_cuda_error(e);
      else set_last_error_text("ctdet_decode_static: cudaGetSymbolAddress returned a null address");
      rc = CTDET_ECUDA;
      p = nullptr;
    } else {
      g_decode_static_addr[dev] = p;
    }
  }
  __atomic_store_n(&g_decode_static_lock, 0, __ATOMIC_RELEASE);
  *out = p;
  return rc;
}

}  // namespace ctdet

extern "C" size_t ctdet_decode_static_max_bytes(void) { return (size_t)CTDET_DECODE_STATIC_BYTES; }

extern "C" int ctdet_decode_static(const float* heat, const float* wh, const float* reg,
                                   int B, int C, int H, int W, int K, float down_ratio, int flags,
                                   float* boxes, float* scores, int32_t* classes, int64_t* inds,
                                   ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || C <= 0 || H <= 0 || W <= 0 || K <= 0) return CTDET_EINVAL;
  if (K > CTDET_MAX_K || (int64_t)K > (int64_t)H * W) return CTDET_EINVAL;
  if ((int64_t)C * H * W >= ((int64_t)1 << 31)) return CTDET_EINVAL;
  if (B == 0) return CTDET_OK;
  const DecodePlan p = make_plan(B, C, H, W, K);
  if (p.total_bytes > (size_t)CTDET_DECODE_STATIC_BYTES) {
    char msg[256];
    snprintf(msg, sizeof(msg),
             "ctdet_decode_static: B=%d C=%d H=%d W=%d K=%d needs %zu bytes of scratch, the static array holds %zu "
             "(CTDET_DECODE_STATIC_BYTES); use ctdet_decode with a caller-owned workspace",
             B, C, H, W, K, p.total_bytes, (size_t)CTDET_DECODE_STATIC_BYTES);
    set_last_error_text(msg);
    return CTDET_EWORKSPACE;
  }
  void* ws = nullptr;
  int rc = decode_static_address(&ws);
  if (rc != CTDET_OK) return rc;
  // The array is shared by every call and keeps the last call's contents: a call with another B leaves key segments
  // where this B's headers / histograms (zero on entry) sit, so that prefix is cleared on the stream every time.
  const size_t zero_bytes = p.header_bytes + 2 * p.hist_bytes + p.phist_bytes;
  cudaError_t e = cudaMemsetAsync(ws, 0, zero_bytes, (cudaStream_t)stream);
  if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  return ctdet_decode(heat, wh, reg, B, C, H, W, K, down_ratio, flags, boxes, scores, classes, inds,
                      ws, (size_t)CTDET_DECODE_STATIC_BYTES, stream);
}

#include "head_decode.cuh"
