// dgrad, second form: the grad_input scatter as a SECOND tensor-core GEMM (included by dcn_bwd.cu inside
// namespace ctdet).  Replaces the col2im kernel of the reference (deform_conv_cuda_kernel.cu:871-949: one scalar
// atomicAdd per (pixel, tap, channel, corner)) and the vector REDs of dcn_dgrad_kernel, whose way out of the SM
// (L1 -> crossbar request port) bounded that kernel.
//
// For one 8 x 16 pixel tile and one 64-channel block the gradient of every input cell the tile can reach,
//     patch[q, ci] = sum_t sum_p S_t[q, p] * G_t[p, ci],      S_t[q, p] = mask[p, t] * (bilinear weight of cell q)
// is a GEMM with K = 9 taps x 128 pixels.  S_t has four non-zeros per column (the four corners of the tap), distinct
// pixels are distinct columns, so it is WRITTEN WITH PLAIN STORES into a zeroed shared-memory tile (and un-written
// after use); G_t is the bf16 column-gradient tile that already sits in shared memory for the consumers, read here
// as the MN-major B operand.  The patch -- 16 x 24 input cells around the tile = 3 M blocks of 128 -- accumulates
// over the 9 taps in TMEM and leaves the SM ONCE per (tile, channel block): 98 KB of REDs instead of 1.18 MB.
// Taps whose top-left corner falls outside the patch (offsets beyond +-3 pixels) keep the old path: the consumer
// warps RED them directly (flag = sign bit of the geometry's lh field).
//
// TMEM (512 columns): [0,128) two G accumulators, [128,512) two patch buffers of 3 x 64 columns.
// Requires stride = dilation = 1 and Cout <= 128 (shared-memory budget); the host falls back to dcn_dgrad_kernel.
#pragma once

constexpr int kPatchH = 16, kPatchW = 24, kPatchR = 3;     // cells; the tile's 10 x 18 tap bases + 3 either side
constexpr int kPatchMB = kPatchH * kPatchW / kTileM;       // 3 M blocks
constexpr int kSChunkBytes = kTileM * kTileM * 2;          // S[128 q][128 p] bf16 = two K halves of [128][128 B]
constexpr int kDg2MaxS = 3;
static_assert(kPatchH * kPatchW == kPatchMB * kTileM, "patch must be whole M blocks");

struct Dgrad2Barriers {
  uint64_t geo_full[2], geo_empty[2];       // raw geometry + patch index of a tile   (geometry -> consumers, builders)
  uint64_t gout_full[2], gout_empty[2];     // grad_output tile                       (geometry -> MMA)
  uint64_t wt_full[kDgMaxRing], wt_empty[kDgMaxRing];   // transposed weight tiles    (bulk copies -> MMA)
  uint64_t acc_full[2], acc_empty[2];       // TMEM column-gradient accumulators      (MMA -> transfer)
  uint64_t g_full[kDgMaxRing], g_empty[kDgMaxRing];     // bf16 column-gradient tile  (transfer -> consumers + MMA)
  uint64_t s_full[kDg2MaxS], s_empty[kDg2MaxS];         // scatter-matrix chunk       (builders -> MMA)
  uint64_t patch_full[2], patch_empty[2];   // TMEM patch accumulators                (MMA -> transfer)
};

// Sampling geometry of one (pixel, tap): {pixel << 9 | cell, lh | flag, lw, mask}.  `pixel` = index of the top-left corner
// in the bordered NHWC image ((h_low + 1) * (W + 2) + w_low + 1 < 2^23, checked by the host), `cell` = its cell in the
// tile's 16 x 24 patch.  A tap the reference does not sample (deform_conv_cuda_kernel.cu:852) has pixel 0 (the zero
// corner of the border) and mask 0.  Bit 31 of .y (lh >= 0) is set when all four corners lie inside the patch: the GEMM
// scatters this tap; otherwise `cell` is 0 and the consumers RED it.
__device__ __forceinline__ uint4 tap_geometry_patch(const DcnParams& prm, float off_h, float off_w, float m,
                                                    int ho, int wo, bool valid, int t, int py0, int px0) {
  uint4 o = make_uint4(0u, 0u, 0u, 0u);
  if (valid) {
    const int i = t / 3, j = t - 3 * i;
    const float h_im = (float)(ho - prm.pad + i) + off_h;
    const float w_im = (float)(wo - prm.pad + j) + off_w;
    if (h_im > -1.f && w_im > -1.f && h_im < (float)prm.H && w_im < (float)prm.W) {
      const int h_low = (int)floorf(h_im), w_low = (int)floorf(w_im);
      o.x = (uint32_t)((h_low + 1) * prm.xw + w_low + 1) << 9;
      o.y = __float_as_uint(h_im - (float)h_low);
      o.z = __float_as_uint(w_im - (float)w_low);
      o.w = __float_as_uint(m);
      const int qy = h_low - py0, qx = w_low - px0;
      if (qy >= 0 && qy <= kPatchH - 2 && qx >= 0 && qx <= kPatchW - 2) {
        o.x |= (uint32_t)(qy * kPatchW + qx);
        o.y |= 0x80000000u;
      }
    }
  }
  return o;
}
// element offset of the top-left corner in the bordered image
__device__ __forceinline__ uint32_t geo_corner_offset(uint32_t x, int Cin) { return (x >> 9) * (uint32_t)Cin; }

// byte offset of S[qrow][p] inside a chunk: K-major, 128-byte swizzle, K halves of 64 pixels 16 KB apart
__device__ __forceinline__ uint32_t s_chunk_offset(uint32_t qrow, uint32_t p) {
  return ((p >> 6) << 14) + (qrow << 7) + (((((p & 63u) >> 3) ^ (qrow & 7u))) << 4) + ((p & 7u) << 1);
}

// bf16 x bf16 + fp32 -> fp32 on the FMA pipe (SASS: FHFMA.BF16, operand halves selected for free)
__device__ __forceinline__ float fma_bf16(uint32_t a, uint32_t b, float c, bool hi) {
  float d;
  const unsigned short ah = (unsigned short)(hi ? (a >> 16) : (a & 0xffffu)), bh = (unsigned short)(hi ? (b >> 16) : (b & 0xffffu));
  asm("fma.rn.f32.bf16 %0, %1, %2, %3;" : "=f"(d) : "h"(ah), "h"(bh), "f"(c));
  return d;
}
__device__ __forceinline__ float dot8_bf(uint4 g, uint4 v) {     // two chains of four
  float a = fma_bf16(g.x, v.x, 0.f, false), b = fma_bf16(g.x, v.x, 0.f, true);
  a = fma_bf16(g.y, v.y, a, false); b = fma_bf16(g.y, v.y, b, true);
  a = fma_bf16(g.z, v.z, a, false); b = fma_bf16(g.z, v.z, b, true);
  a = fma_bf16(g.w, v.w, a, false); b = fma_bf16(g.w, v.w, b, true);
  return a + b;
}
// grad_input += G[8 channels] * c for one corner (the rare out-of-patch path): 32 B = one sector
__device__ __noinline__ void dgrad_scatter8(float* q, uint4 g, float c) {
  if (c != 0.f) {
    red_add_v4(q, bf_lo(g.x) * c, bf_hi(g.x) * c, bf_lo(g.y) * c, bf_hi(g.y) * c);
    red_add_v4(q + 4, bf_lo(g.z) * c, bf_hi(g.z) * c, bf_lo(g.w) * c, bf_hi(g.w) * c);
  }
}

#ifdef CTDET_DGRAD_TIMING
__device__ unsigned long long g_dg2_ts[16][64];
#define DG2_TS(role, n) do { if (blockIdx.x == 0 && (n) < 64) g_dg2_ts[role][n] = clock64(); } while (0)
#else
#define DG2_TS(role, n) do {} while (0)
#endif

// Persistent, warp-specialised; one CTA per SM over 8 x 16 pixel tiles; K block n = (tile, channel block, tap):
//   warps [0, NP)          consumers: grad_mask / grad_offset from G and the four gathered corners of x (as
//                          dcn_dgrad_kernel); one non-zero of S per lane and K block (written / un-written with plain
//                          stores); grad_input REDs only for taps the patch does not cover
//   warps [NP, NP+4)       transfer: TMEM G accumulator -> bf16 G tile in shared memory; every 9 K blocks the
//                          finished patch TMEM -> fp32 REDs into the bordered NHWC grad_input workspace
//   warps [NP+4, NP+8)     grad_output tile + sampling geometry + patch cell of the tiles ahead (thread = pixel)
//   warp  NP+8             one lane issues GEMM 1 (G = gout * Wt) and streams the transposed weight tiles (bulk copies)
//   warp  NP+9             one lane issues GEMM 2 (patch += S * G), gated by the G tile and the S chunks
// (a separate pair of builder warps for S was the bottleneck of the first form: ~1100 cycles per chunk)
template <int NP>
__global__ void __launch_bounds__((NP + 10) * 32, 1)
dcn_dgrad2_kernel(const DcnParams prm, const DcnBwdParams bw) {
  static_assert(NP % 4 == 0 && 32 % NP == 0, "consumer warps must tile the 32 row groups");
  extern __shared__ __align__(1024) uint8_t dcn_smem[];
  __shared__ Dgrad2Barriers bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Cin = prm.Cin, Cout = prm.Cout;
  const int P = prm.Ho * prm.Wo;
  const int NG = bw.g_stages, NW = bw.w_stages, GB = bw.gout_bufs, NSB = bw.s_bufs;
  const uint32_t gout_bytes = (uint32_t)(Cout / 64) * kTileM * 128u, wt_bytes = (uint32_t)Cout * 128u;
  const uint32_t g_bytes = (uint32_t)kTileM * 128u;
  uint8_t* smem = dcn_smem + ((1024u - (smem_u32(dcn_smem) & 1023u)) & 1023u);
  uint8_t* gout = smem;                                               // [GB][Cout/64][128][128 B]
  uint8_t* wtr = gout + (size_t)GB * gout_bytes;                      // [NW][Cout/64][64][128 B]
  uint8_t* gring = wtr + (size_t)NW * wt_bytes;                       // [NG][128][128 B]
  uint8_t* sring = gring + (size_t)NG * g_bytes;                      // [NSB][2][128][128 B]
  uint4* geo = reinterpret_cast<uint4*>(sring + (size_t)NSB * kSChunkBytes);   // [2][9][128] {pixel << 9 | cell, lh | flag, lw, mask}
  float* acc = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(geo) + kDgGeoBytes);  // [2][27][128] per-tile sums
  const int cpb = Cin / kBK, KB = 9 * cpb;
  const int tile_first = (int)(((long long)blockIdx.x * prm.num_tiles) / gridDim.x);
  const int my_tiles = (int)(((long long)(blockIdx.x + 1) * prm.num_tiles) / gridDim.x) - tile_first;
  const int total = my_tiles * KB;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      dcn_mbar_init(&bars.geo_full[i], 4);    dcn_mbar_init(&bars.geo_empty[i], NP);
      dcn_mbar_init(&bars.gout_full[i], 4);   dcn_mbar_init(&bars.gout_empty[i], 1);
      dcn_mbar_init(&bars.acc_full[i], 1);    dcn_mbar_init(&bars.acc_empty[i], 4);
      dcn_mbar_init(&bars.patch_full[i], 1);  dcn_mbar_init(&bars.patch_empty[i], 4);
    }
    for (int i = 0; i < kDgMaxRing; ++i) {
      dcn_mbar_init(&bars.wt_full[i], 1);     dcn_mbar_init(&bars.wt_empty[i], 1);
      dcn_mbar_init(&bars.g_full[i], 4);      dcn_mbar_init(&bars.g_empty[i], NP + 1);
    }
    for (int i = 0; i < kDg2MaxS; ++i) { dcn_mbar_init(&bars.s_full[i], NP); dcn_mbar_init(&bars.s_empty[i], 1); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == NP + 8) tmem_alloc(&s_tmem, 512);
  if (warp >= NP + 4 && warp < NP + 8) {       // the S ring starts (and is always left) all zero
    const int bt = tid - (NP + 4) * 32;        // 0..127
    uint4* z = reinterpret_cast<uint4*>(sring);
    const int n16 = NSB * kSChunkBytes / 16;
    for (int i = bt; i < n16; i += 128) z[i] = make_uint4(0u, 0u, 0u, 0u);
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  if (warp < NP) {
    // ===================================== consumers ==============================================
    // lane = (row group `sub`, 16-byte channel chunk `chunk`): 8 consecutive channels per lane, so one LDG.128
    // per corner and one conflict-free LDS.128 of the G row; the dot products run on the mixed-precision FMA
    // (bf16 x bf16 + fp32 -> fp32, FHFMA.BF16): no unpacking of the operands.
    if (NP == 16) asm volatile("setmaxnreg.inc.sync.aligned.u32 80;");
    constexpr int ITERS = 32 / NP;
    const int sub = lane >> 3, chunk = lane & 7;
    const uint32_t r7 = (uint32_t)(((warp & 1) << 2) + sub);                       // row & 7 is fixed per thread
    const uint32_t row0 = (uint32_t)(warp << 2) + sub;
    const uint32_t g_off = row0 * 128u + (((uint32_t)chunk ^ r7) << 4);           // this lane's 16 B of G row `row0`
    const uint32_t ex = (uint32_t)Cin, ey = (uint32_t)prm.xw * (uint32_t)Cin;      // bordered rows
    const bool b0 = (chunk & 1) != 0, b1 = (chunk & 2) != 0;
    uint4 va[ITERS], vb[ITERS], vc[ITERS], vd[ITERS];
    auto gather = [&](int i, const __nv_bfloat16* xc, uint32_t e0) {
      const __nv_bfloat16* p = xc + e0;
      va[i] = ldg_gather16(p);
      vb[i] = ldg_gather16(p + ex);
      vc[i] = ldg_gather16(p + ey);
      vd[i] = ldg_gather16(p + ey + ex);
    };
    int l_it = 0, l_cb = 0, l_t = 0;             // load cursor: the K block whose gathers are in flight
    const __nv_bfloat16* l_x = prm.xh + chunk * 8;
    if (total > 0) {
      l_x = prm.xh + (size_t)(tile_first / prm.tiles_per_image) * prm.ximg + chunk * 8;
      dcn_mbar_wait(&bars.geo_full[0], 0u);
#pragma unroll
      for (int i = 0; i < ITERS; ++i) gather(i, l_x, geo_corner_offset(geo[row0 + i * NP * 4].x, Cin));
    }
    // S (the scatter matrix of GEMM 2): lane (chunk & 3) of a row's 8 lanes owns corner (chunk & 3) of pixel row
    // row0 (chunk < 4) or row0 + 64 (chunk >= 4) -- one non-zero of S per lane and K block, written with a plain store.
    // An entry = (M block << 16) | byte offset inside the chunk, ~0 = none.  Chunk c = 3 n + mb lives in buffer c & 1,
    // whose previous content was chunk c - 2 = (n, 0) for mb = 2, (n - 1, 2) for mb = 1, (n - 1, 1) for mb = 0: it is
    // un-written from this lane's entry of this / the previous K block.
    const uint32_t s_p = row0 + (uint32_t)(chunk >> 2) * 64u;                      // column of S = pixel of the tile
    const uint32_t s_k = (uint32_t)chunk & 3u;                                     // corner
    uint32_t s_cur = 0xffffffffu, s_prv = 0xffffffffu, s_c = 0u;
    unsigned short s_w = 0;
    auto s_chunk = [&](uint32_t mb) {
      const uint32_t sbuf = s_c & 1u;
      dcn_mbar_wait(&bars.s_empty[sbuf], ((s_c >> 1) & 1u) ^ 1u);
      uint8_t* chunk_base = sring + (size_t)sbuf * kSChunkBytes;
      const uint32_t old = mb == 2u ? s_cur : s_prv, old_mb = mb == 2u ? 0u : (mb == 1u ? 2u : 1u);
      if ((old >> 16) == old_mb) *reinterpret_cast<unsigned short*>(chunk_base + (old & 0xffffu)) = 0;
      if ((s_cur >> 16) == mb) *reinterpret_cast<unsigned short*>(chunk_base + (s_cur & 0xffffu)) = s_w;
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.s_full[sbuf]);
      ++s_c;
    };
    int gs = 0; unsigned gph = 0;
    for (int n = 0; n < total; ++n) {
      const int c_it = l_it, c_cb = l_cb, c_t = l_t;
      const bool have_next = n + 1 < total;
      if (++l_t == 9) {
        l_t = 0;
        if (++l_cb == cpb) {
          l_cb = 0; ++l_it;
          if (have_next) {
            const int tile = tile_first + l_it;
            l_x = prm.xh + (size_t)(tile / prm.tiles_per_image) * prm.ximg + chunk * 8;
            dcn_mbar_wait(&bars.geo_full[l_it & 1], (unsigned)((l_it >> 1) & 1));
          }
        }
      }
      const __nv_bfloat16* l_xc = l_x + l_cb * kBK;
      const uint4* geo_n = geo + ((l_it & 1) * 9 + l_t) * kTileM + row0;
      const uint4* geo_c = geo + ((c_it & 1) * 9 + c_t) * kTileM + row0;
      {                                          // this lane's non-zero of S for this K block (geometry only: before G)
        const uint4 g4 = geo_c[(chunk >> 2) * NP * 4];
        const float lh = __uint_as_float(g4.y & 0x7fffffffu), lw = __uint_as_float(g4.z), m = __uint_as_float(g4.w);
        const float wy = (s_k & 2u) ? lh : 1.f - lh, wx = (s_k & 1u) ? lw : 1.f - lw;
        s_w = __bfloat16_as_ushort(__float2bfloat16_rn(m * (wy * wx)));
        const uint32_t qc = (g4.x & 511u) + (s_k & 1u) + (s_k >> 1) * kPatchW;
        s_prv = s_cur;
        s_cur = (g4.y >> 31) ? (((qc >> 7) << 16) | s_chunk_offset(qc & 127u, s_p)) : 0xffffffffu;
      }
      s_chunk(0u);
      dcn_mbar_wait(&bars.g_full[gs], gph);
      if (tid == 0) DG2_TS(13, n);
      {
      const uint8_t* g_lane = gring + (size_t)gs * g_bytes + g_off;
      // both rows of the thread in lock step (16 independent FMA chains, the shuffles of the two reductions
      // overlap): the loop is latency-bound with 4 consumer warps per scheduler
      uint4 q4[ITERS], gq[ITERS];
      float d[ITERS][4];
#pragma unroll
      for (int i = 0; i < ITERS; ++i) {
        q4[i] = geo_c[i * NP * 4];
        gq[i] = *reinterpret_cast<const uint4*>(g_lane + (uint32_t)(i * NP * 4) * 128u);   // G channels [8c, 8c+8)
      }
#pragma unroll
      for (int i = 0; i < ITERS; ++i) {
        d[i][0] = dot8_bf(gq[i], va[i]); d[i][1] = dot8_bf(gq[i], vb[i]);
        d[i][2] = dot8_bf(gq[i], vc[i]); d[i][3] = dot8_bf(gq[i], vd[i]);
      }
      if (have_next) {
#pragma unroll
        for (int i = 0; i < ITERS; ++i) gather(i, l_xc, geo_corner_offset(geo_n[i * NP * 4].x, Cin));   // refill: K block n+1
      }
      s_chunk(1u);
      float red[ITERS], msk[ITERS];
#pragma unroll
      for (int i = 0; i < ITERS; ++i) {
        const float lh = __uint_as_float(q4[i].y & 0x7fffffffu), lw = __uint_as_float(q4[i].z), m = __uint_as_float(q4[i].w);
        const float hh = 1.f - lh, hw = 1.f - lw;
        const float w1 = hh * hw, w2 = hh * lw, w3 = lh * hw, w4 = lh * lw;
        msk[i] = m;
        if ((q4[i].y >> 31) == 0u && m != 0.f) {  // tap outside the patch (rare): scatter it here (deform_conv_cuda_kernel.cu:944)
          float* gp = bw.gx + (size_t)((tile_first + c_it) / prm.tiles_per_image) * prm.ximg + c_cb * kBK + chunk * 8 +
                      geo_corner_offset(q4[i].x, Cin);
          dgrad_scatter8(gp, gq[i], m * w1); dgrad_scatter8(gp + ex, gq[i], m * w2);
          dgrad_scatter8(gp + ey, gq[i], m * w3); dgrad_scatter8(gp + ey + ex, gq[i], m * w4);
        }
        // this lane's share of the three per-(pixel, tap) sums
        const float vm = w1 * d[i][0] + w2 * d[i][1] + w3 * d[i][2] + w4 * d[i][3];   // :1019-1030
        const float vh = hw * (d[i][2] - d[i][0]) + lw * (d[i][3] - d[i][1]);         // dmcn_get_coordinate_weight, bp_dir 0 (x mask below)
        const float vw = hh * (d[i][1] - d[i][0]) + lh * (d[i][3] - d[i][2]);         // bp_dir 1
        // 8-lane reduction of the three sums in 4 shuffles: after the first exchange the even lanes carry
        // {vm, vh} and the odd lanes vw, after the second the lanes = 0 (mod 4) carry vm, = 2 (mod 4) vh
        const float r1 = __shfl_xor_sync(0xffffffffu, b0 ? vm : vw, 1);
        const float r2 = __shfl_xor_sync(0xffffffffu, vh, 1);
        float k0 = (b0 ? vw : vm) + r1;                                // even: vm, odd: vw
        const float k1 = vh + r2;                                     // even lanes: vh
        const float r3 = __shfl_xor_sync(0xffffffffu, (b0 || b1) ? k0 : k1, 2);
        k0 = ((!b0 && b1) ? k1 : k0) + r3;                             // 0 mod 4: vm, 2 mod 4: vh, odd: vw
        red[i] = k0 + __shfl_xor_sync(0xffffffffu, k0, 4);
      }
      if (chunk < 3) {                           // lane 0: grad_mask, lane 2: grad_offset h, lane 1: grad_offset w
        // into the tile's table of sums over channel blocks; the geometry warps write it out once the tile is done
        const int which = chunk == 0 ? 0 : (chunk == 2 ? 1 : 2);      // slot: {vm, vh, vw}
        float* a = acc + ((c_it & 1) * 27 + 3 * c_t + which) * kTileM + row0;
#pragma unroll
        for (int i = 0; i < ITERS; ++i) {
          float v = which ? red[i] * msk[i] : red[i];
          if (c_cb > 0) v += a[i * NP * 4];
          a[i * NP * 4] = v;
        }
      }
      }
      s_chunk(2u);
      if (tid == 0) DG2_TS(14, n);
      if (lane == 0) {
        dcn_mbar_arrive(&bars.g_empty[gs]);
        if (l_it != c_it) dcn_mbar_arrive(&bars.geo_empty[c_it & 1]);
      }
      if (++gs == NG) { gs = 0; gph ^= 1u; }
    }
  } else if (warp < NP + 4) {
    // ===================================== transfer: TMEM -> bf16 G tile; patch flush =============
    const int quarter = warp - NP;               // = warp % 4: the TMEM lane quarter this warp may read
    const int r = (quarter << 5) + lane;
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    // group = (tile, channel block): its patch sits in TMEM buffer group & 1 once the MMA warp has committed tap 8
    auto flush = [&](int group) {
      const int pb = group & 1;
      const int it = group / cpb, cb = group - it * cpb;
      int tile = tile_first + it;
      const int img = tile / prm.tiles_per_image;
      tile -= img * prm.tiles_per_image;
      const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
      const int by0 = tile_y * kTileH - prm.pad - kPatchR + 1, bx0 = tile_x * kTileW - prm.pad - kPatchR + 1;   // bordered coordinates
      float* base = bw.gx + (size_t)img * prm.ximg + cb * kBK;
      dcn_mbar_wait(&bars.patch_full[pb], (unsigned)((group >> 1) & 1));
      tc_fence_after();
#pragma unroll 1
      for (int mb = 0; mb < kPatchMB; ++mb) {
        const int q = mb * kTileM + r;
        const int qy = q / kPatchW, qx = q - qy * kPatchW;
        const int by = by0 + qy, bx = bx0 + qx;
        const bool ok = by >= 0 && by <= prm.H + 1 && bx >= 0 && bx <= prm.W + 1;
        float* dst = base + ((size_t)(ok ? by : 0) * prm.xw + (ok ? bx : 0)) * Cin;
#pragma unroll
        for (int c0 = 0; c0 < kBK; c0 += 16) {
          uint32_t v[16];
          tmem_ld16(tmem + lane_base + (uint32_t)(128 + pb * (kPatchMB * kBK) + mb * kBK + c0), v);
          uint32_t any = 0u;
#pragma unroll
          for (int k = 0; k < 16; ++k) any |= v[k] << 1;            // +-0 -> 0
          if (ok && any != 0u) {
#pragma unroll
            for (int k = 0; k < 16; k += 4)
              red_add_v4(dst + c0 + k, __uint_as_float(v[k]), __uint_as_float(v[k + 1]), __uint_as_float(v[k + 2]),
                         __uint_as_float(v[k + 3]));
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.patch_empty[pb]);
    };
    int gs = 0; unsigned gph = 0;
    for (int n = 0; n < total; ++n) {
      if (n >= 9 && (n % 9) == 5) flush(n / 9 - 1);                   // (GEMM 2 of that group has normally retired by now)
      const int ab = n & 1;
      dcn_mbar_wait(&bars.acc_full[ab], (unsigned)((n >> 1) & 1));
      if (quarter == 0 && lane == 0) DG2_TS(1, n);
      tc_fence_after();
      uint32_t pk[32];
#pragma unroll
      for (int c0 = 0; c0 < kBK; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tmem + lane_base + (uint32_t)(ab * kBK + c0), v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
          pk[c0 / 2 + k] = *reinterpret_cast<const uint32_t*>(&h);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.acc_empty[ab]);            // TMEM buffer may be overwritten
      if (quarter == 0 && lane == 0) DG2_TS(15, n);
      dcn_mbar_wait(&bars.g_empty[gs], gph ^ 1u);
      if (quarter == 0 && lane == 0) DG2_TS(2, n);
      uint8_t* dst = gring + (size_t)gs * g_bytes + r * 128;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(dst + ((c ^ (r & 7)) << 4)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the tile is also a tensor-core operand now
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.g_full[gs]);
      if (quarter == 0 && lane == 0) DG2_TS(3, n);
      if (++gs == NG) { gs = 0; gph ^= 1u; }
    }
    if (total > 0) flush(total / 9 - 1);
  } else if (warp < NP + 8) {
    // ===================================== grad_output tile + geometry ============================
    if (NP == 16) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    const int r = ((warp - NP - 4) << 5) + lane;
    // grad_offset / grad_mask of a finished tile: acc[tile & 1] -> global (the consumers' arrivals on geo_empty
    // order their shared-memory writes before this)
    auto drain = [&](int itd) {
      int tile = tile_first + itd;
      const int b = tile / prm.tiles_per_image;
      tile -= b * prm.tiles_per_image;
      const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
      const int ho = tile_y * kTileH + (r >> 4), wo = tile_x * kTileW + (r & 15);
      if (ho < prm.Ho && wo < prm.Wo) {
        const float* a = acc + (itd & 1) * 27 * kTileM + r;
        float* gm = bw.gmask + (size_t)b * 9 * P + ho * prm.Wo + wo;
        float* go = bw.goff + (size_t)b * 18 * P + ho * prm.Wo + wo;
#pragma unroll
        for (int t = 0; t < 9; ++t) {
          gm[(size_t)t * P] = a[(3 * t) * kTileM];
          go[(size_t)(2 * t) * P] = a[(3 * t + 1) * kTileM];
          go[(size_t)(2 * t + 1) * P] = a[(3 * t + 2) * kTileM];
        }
      }
    };
    for (int it = 0; it < my_tiles; ++it) {
      // geometry first: its buffer frees up a whole tile before the grad_output buffer does (one buffer at Cout >= 64...)
      int tile = tile_first + it;
      const int b = tile / prm.tiles_per_image;
      const int tl = tile - b * prm.tiles_per_image;
      const int tile_y = tl / prm.tiles_x, tile_x = tl - tile_y * prm.tiles_x;
      const int ho = tile_y * kTileH + (r >> 4), wo = tile_x * kTileW + (r & 15);
      const int py0 = tile_y * kTileH - prm.pad - kPatchR, px0 = tile_x * kTileW - prm.pad - kPatchR;
      const bool valid = ho < prm.Ho && wo < prm.Wo;
      const int p = valid ? ho * prm.Wo + wo : 0;
      const float* offb = prm.offset + (size_t)b * 18 * P + p;
      const float* mskb = prm.mask + (size_t)b * 9 * P + p;
      const int buf = it & 1;
      dcn_mbar_wait_relaxed(&bars.geo_empty[buf], (unsigned)(((it >> 1) & 1) ^ 1));   // tile it - 2 is done
      if (it >= 2) drain(it - 2);
      float ov[18], mv[9];
#pragma unroll
      for (int k = 0; k < 18; ++k) ov[k] = ldg_stream1(offb + (size_t)k * P);
#pragma unroll
      for (int k = 0; k < 9; ++k) mv[k] = ldg_stream1(mskb + (size_t)k * P);
#pragma unroll
      for (int t = 0; t < 9; ++t)
        geo[(buf * 9 + t) * kTileM + r] = tap_geometry_patch(prm, ov[2 * t], ov[2 * t + 1], mv[t], ho, wo, valid, t, py0, px0);
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.geo_full[buf]);

      const int gb = it % GB;
      dcn_mbar_wait_relaxed(&bars.gout_empty[gb], (unsigned)(((it / GB) & 1) ^ 1));
      dcn_load_gout_tile(prm, bw.gh, tile, r, gout + (size_t)gb * gout_bytes);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.gout_full[gb]);
    }
    for (int it = my_tiles; it < my_tiles + 2; ++it) {            // the last two tiles
      if (it < 2) continue;
      dcn_mbar_wait_relaxed(&bars.geo_empty[it & 1], (unsigned)(((it >> 1) & 1) ^ 1));
      drain(it - 2);
    }
  } else if (warp == NP + 8) {
    // ===================================== GEMM 1 issuer + weight loader ===========================
    // One lane: G = gout[128 px, Cout] * Wt_kb[Cout, 64 ci] per K block, and the 1-D bulk copies of the
    // transposed weight tiles (the slot of K block n - 1 is refilled with K block n - 1 + NW right after GEMM 1
    // of K block n has been issued: its predecessor has normally retired by then).
    if (lane == 0) {
      const uint32_t idesc1 = make_idesc_bf16(kTileM, kBK);                       // A, B K-major
      auto load_w = [&](int slot, int kblock) {
        dcn_mbar_arrive_expect_tx(&bars.wt_full[slot], wt_bytes);
        dcn_bulk_g2s(wtr + (size_t)slot * wt_bytes, bw.wt + (size_t)(kblock % KB) * Cout * kBK, wt_bytes, &bars.wt_full[slot]);
      };
      for (int k = 0; k < NW && k < total; ++k) load_w(k, k);
      int ws = 0; unsigned wph = 0;
      int ps = 0; unsigned pph = 0;                // slot / phase of the K block issued last
      uint64_t a0 = 0ull;
      int kb1 = 0, it1 = 0;
      for (int n = 0; n < total; ++n) {
        const int gb = it1 % GB;
        if (kb1 == 0) {
          dcn_mbar_wait_relaxed(&bars.gout_full[gb], (unsigned)((it1 / GB) & 1), 0u);
          a0 = make_sw128_desc(smem_u32(gout + (size_t)gb * gout_bytes));
        }
        const int ab = n & 1;
        dcn_mbar_wait_relaxed(&bars.acc_empty[ab], (unsigned)(((n >> 1) & 1) ^ 1), 0u);
        dcn_mbar_wait_relaxed(&bars.wt_full[ws], wph, 0u);
        DG2_TS(0, n);
        tc_fence_after();
        const uint64_t b0 = make_sw128_desc(smem_u32(wtr + (size_t)ws * wt_bytes));
        const uint32_t d_tmem = tmem + (uint32_t)(ab * kBK);
        for (int j = 0; j < Cout / 64; ++j) {
          const uint64_t adesc = a0 + (uint64_t)(j * (kTileM * 128 / 16)), bdesc = b0 + (uint64_t)(j * (64 * 128 / 16));
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)
            umma_bf16(d_tmem, adesc + 2 * k4, bdesc + 2 * k4, idesc1, (uint32_t)((j | k4) != 0));
        }
        umma_commit(&bars.wt_empty[ws]);
        umma_commit(&bars.acc_full[ab]);
        if (++kb1 == KB) { kb1 = 0; ++it1; umma_commit(&bars.gout_empty[gb]); }
        if (n > 0 && n - 1 + NW < total) {
          dcn_mbar_wait_relaxed(&bars.wt_empty[ps], pph, 0u);
          load_w(ps, n - 1 + NW);
        }
        ps = ws; pph = wph;
        if (++ws == NW) { ws = 0; wph ^= 1u; }
      }
    }
    __syncwarp();
  } else {
    // ===================================== GEMM 2 issuer ===========================================
    // One lane: patch[mb] += S_chunk[128 q, 128 px] * G[128 px, 64 ci] for the three M blocks of every K block
    if (lane == 0) {
      const uint32_t idesc2 = make_idesc_bf16(kTileM, kBK) | (1u << 16);          // B (= the G tile) MN-major
      int t2 = 0, grp2 = 0;
      int gs2 = 0; unsigned gph2 = 0;
      int sb = 0; unsigned sph = 0;
      for (int n = 0; n < total; ++n) {
        const int pb = grp2 & 1;
        if (t2 == 0) dcn_mbar_wait_relaxed(&bars.patch_empty[pb], (unsigned)(((grp2 >> 1) & 1) ^ 1), 0u);
        dcn_mbar_wait_relaxed(&bars.g_full[gs2], gph2, 0u);
        DG2_TS(4, n);
        const uint64_t b0 = make_sw128_desc_mn(smem_u32(gring + (size_t)gs2 * g_bytes), 1024u);
        const uint32_t d0 = tmem + (uint32_t)(128 + pb * (kPatchMB * kBK));
#pragma unroll
        for (int mb = 0; mb < kPatchMB; ++mb) {
          dcn_mbar_wait_relaxed(&bars.s_full[sb], sph, 0u);
          DG2_TS(5 + mb, n);
          tc_fence_after();
          const uint64_t a0 = make_sw128_desc(smem_u32(sring + (size_t)sb * kSChunkBytes));
#pragma unroll
          for (int k8 = 0; k8 < kTileM / 16; ++k8)         // K = 16 pixels: 32 B of an S row, 2 swizzle atoms (2 KB) of G
            umma_bf16(d0 + (uint32_t)(mb * kBK), a0 + (uint64_t)((k8 >> 2) * 1024 + 2 * (k8 & 3)), b0 + (uint64_t)(k8 * 128),
                      idesc2, (uint32_t)((t2 | k8) != 0));
          umma_commit(&bars.s_empty[sb]);
          sb ^= 1; if (sb == 0) sph ^= 1u;
        }
        DG2_TS(8, n);
        umma_commit(&bars.g_empty[gs2]);
        if (++gs2 == NG) { gs2 = 0; gph2 ^= 1u; }
        if (++t2 == 9) { t2 = 0; umma_commit(&bars.patch_full[pb]); ++grp2; }
      }
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == NP + 8) tmem_dealloc(tmem, 512);
}
