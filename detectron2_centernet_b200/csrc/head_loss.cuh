// SURVEY section 8 f2, second half: the hm head's final 1x1 convolution fused with the focal loss (included by loss.cu
// inside namespace ctdet).
//
// Reference chain replaced: `self.hm[-1]` = nn.Conv2d(head_conv, num_classes, kernel_size=1) with bias -2.19
// (detectron2/modeling/meta_arch/centernet.py:118-123) applied at :152, then CenterNet.losses' sigmoid_ + clamp (:204) and
// FocalLoss / _neg_loss (:323-369) and their autograd backward down to the logits.  Here
//     logits[128 px, C] = feat[128 px, Cf] * W^T[Cf, C] + bias
// is formed per 128-pixel tile on the tensor cores (bf16 operands, fp32 accumulation in TMEM) and consumed from TMEM by
// the cell arithmetic of loss_kernel: the [B, C, H, W] logits tensor never exists in HBM (the convolution's 4 B/cell
// write and the loss's 4 B/cell read disappear).  Outputs: the focal loss and d(loss)/d(logits).
//
// Persistent, warp-specialised, one CTA per SM, tiles = 128 consecutive pixels of one image:
//   warps [0, 8)    producers: feat is NCHW fp32, the UMMA A operand wants [pixel][64 channels] bf16 rows -- warp w
//                   owns channels 8w..8w+7 of the K block, lane l pixels l, l+32, l+64, l+96 (every load instruction is
//                   one coalesced 128-byte line), 8 channels of a pixel leave as one 16-byte store into the
//                   128-byte-swizzled row (conflict-free: 8 consecutive lanes = 8 distinct chunks)
//   warps [8, 16)   epilogue: warp & 3 = TMEM lane quarter (thread = pixel), (warp - 8) >> 2 = half of the classes;
//                   tcgen05.ld, + bias, gt loads (coalesced over pixels), cell2 (the arithmetic of loss_kernel),
//                   d_logits stores
//   warp 16         one lane issues tcgen05.mma (M = 128 pixels, N = C, K = 16 x 4 per K block) into one of two TMEM
//                   accumulators; W^T (bf16, swizzled) is built in shared memory once per CTA
#pragma once

constexpr int kHfProducerWarps = 8, kHfEpiWarps = 8;
constexpr int kHfThreads = (kHfProducerWarps + kHfEpiWarps + 1) * 32;
constexpr int kHfMaxStages = 8;

struct HeadFusedBarriers {
  uint64_t full[kHfMaxStages], empty[kHfMaxStages];   // A ring (producers -> MMA)
  uint64_t acc_full[2], acc_empty[2];                 // TMEM accumulators (MMA -> epilogue)
};

__device__ __forceinline__ void tmem_ld8(uint32_t taddr, uint32_t (&v)[8]) {      // 32 lanes x 8 columns
  asm volatile("tcgen05.ld.sync.aligned.32x32b.x8.b32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
               : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7])
               : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

__global__ void __launch_bounds__(kHfThreads, 1)
head_focal_fused_kernel(const float* __restrict__ feat, const float* __restrict__ weight, const float* __restrict__ bias,
                        const float* __restrict__ gt, const float* __restrict__ alpha, const int32_t* __restrict__ num_pos_img,
                        const float* __restrict__ norm_override, int B, int Cf, int C, int P, int flags, int stages,
                        float w_hm, float* __restrict__ losses, float* __restrict__ d_logits, LossWs* ws) {
  extern __shared__ __align__(1024) uint8_t hf_smem[];
  __shared__ HeadFusedBarriers bars;
  __shared__ uint32_t s_tmem;
  __shared__ float s_gscale;
  __shared__ bool s_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int KB = Cf / kBK;                                  // K blocks of 64 channels
  const int NS = stages;
  uint8_t* smem = hf_smem + ((1024u - (smem_u32(hf_smem) & 1023u)) & 1023u);
  uint8_t* ring = smem;                                     // [NS][128 px][128 B]
  uint8_t* wsm = ring + (size_t)NS * (kTileM * 128);        // [KB][C rows][128 B]  W^T, K-major over the channels
  float* s_bias = reinterpret_cast<float*>(wsm + (size_t)KB * C * 128);   // [C]
  float* s_alpha = s_bias + C;                              // [C]
  const int tiles_per_image = (P + kTileM - 1) / kTileM;      // the last tile of an image may be ragged
  const int num_tiles = B * tiles_per_image;
  const int tile_first = (int)(((long long)blockIdx.x * num_tiles) / gridDim.x);
  const int my_tiles = (int)(((long long)(blockIdx.x + 1) * num_tiles) / gridDim.x) - tile_first;

  if (tid == 0) {
    for (int s = 0; s < kHfMaxStages; ++s) { dcn_mbar_init(&bars.full[s], kHfProducerWarps); dcn_mbar_init(&bars.empty[s], 1); }
    for (int i = 0; i < 2; ++i) { dcn_mbar_init(&bars.acc_full[i], 1); dcn_mbar_init(&bars.acc_empty[i], kHfEpiWarps); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    // normaliser (:360-368): num_pos over the batch, or the caller's (all-reduced) value
    float npos = 0.f;
    if (norm_override != nullptr) npos = norm_override[0];
    else if (num_pos_img != nullptr) { int a = 0; for (int i = 0; i < B; ++i) a += num_pos_img[i]; npos = (float)a; }
    else npos = (float)ws->npos_count;                       // written by count_pos_kernel (stream order)
    s_gscale = -(npos == 0.f ? w_hm : w_hm / npos);
  }
  if (warp == kHfProducerWarps + kHfEpiWarps) tmem_alloc(&s_tmem, 256);
  // W^T -> bf16, swizzled: slab kb, row n (class), 16-byte chunk j (8 channels) at ((j ^ (n & 7)) << 4)
  for (int i = tid; i < C * (Cf / 8); i += kHfThreads) {
    const int n = i / (Cf / 8), j8 = i - n * (Cf / 8);      // class, group of 8 channels
    const float4 a = __ldg(reinterpret_cast<const float4*>(weight + (size_t)n * Cf + j8 * 8));
    const float4 b = __ldg(reinterpret_cast<const float4*>(weight + (size_t)n * Cf + j8 * 8 + 4));
    const __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w);
    const __nv_bfloat162 p2 = __floats2bfloat162_rn(b.x, b.y), p3 = __floats2bfloat162_rn(b.z, b.w);
    const int kb = j8 >> 3, j = j8 & 7;
    *reinterpret_cast<uint4*>(wsm + (size_t)kb * C * 128 + n * 128 + ((j ^ (n & 7)) << 4)) =
        make_uint4(*reinterpret_cast<const uint32_t*>(&p0), *reinterpret_cast<const uint32_t*>(&p1),
                   *reinterpret_cast<const uint32_t*>(&p2), *reinterpret_cast<const uint32_t*>(&p3));
  }
  for (int i = tid; i < C; i += kHfThreads) {
    s_bias[i] = bias ? __ldg(bias + i) : 0.f;
    s_alpha[i] = (flags & CTDET_LOSS_ALPHA_UNIFORM) ? __ldg(alpha) : __ldg(alpha + i);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  double acc = 0.0;                                          // this thread's share of the weighted focal loss
  if (warp < kHfProducerWarps) {
    // ===================================== producers ==============================================
    // the loads of K block n + 1 are in flight while K block n is converted and stored (64 x 4 B per thread in flight)
    int s = 0; unsigned ph = 0;
    const uint32_t swz = (uint32_t)((warp ^ (lane & 7)) << 4);              // rows l + 32 j: row & 7 = lane & 7
    const int total_kb = my_tiles * KB;
    float v[2][8][4];
    auto load_kb = [&](int n, float (&dst)[8][4]) {
      const int it = n / KB, kb = n - it * KB;
      const int tile = tile_first + it;
      const int b = tile / tiles_per_image, p0 = (tile - b * tiles_per_image) * kTileM;
      const float* src = feat + ((size_t)b * Cf + kb * kBK + warp * 8) * P + p0 + lane;
      if (p0 + kTileM <= P) {
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dst[i][j] = ldg_stream1(src + (size_t)i * P + 32 * j);
      } else {                                   // ragged tile: rows past the image are zero
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) dst[i][j] = (p0 + lane + 32 * j < P) ? ldg_stream1(src + (size_t)i * P + 32 * j) : 0.f;
      }
    };
    if (total_kb > 0) load_kb(0, v[0]);
#pragma unroll 1
    for (int n2 = 0; n2 < total_kb; n2 += 2) {
#pragma unroll
      for (int u = 0; u < 2; ++u) {
        const int n = n2 + u;
        if (n >= total_kb) break;
        if (n + 1 < total_kb) load_kb(n + 1, v[u ^ 1]);
        dcn_mbar_wait(&bars.empty[s], ph ^ 1u);
        uint8_t* a_tile = ring + (size_t)s * (kTileM * 128);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const __nv_bfloat162 q0 = __floats2bfloat162_rn(v[u][0][j], v[u][1][j]), q1 = __floats2bfloat162_rn(v[u][2][j], v[u][3][j]);
          const __nv_bfloat162 q2 = __floats2bfloat162_rn(v[u][4][j], v[u][5][j]), q3 = __floats2bfloat162_rn(v[u][6][j], v[u][7][j]);
          *reinterpret_cast<uint4*>(a_tile + (uint32_t)(lane + 32 * j) * 128u + swz) =
              make_uint4(*reinterpret_cast<const uint32_t*>(&q0), *reinterpret_cast<const uint32_t*>(&q1),
                         *reinterpret_cast<const uint32_t*>(&q2), *reinterpret_cast<const uint32_t*>(&q3));
        }
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) dcn_mbar_arrive(&bars.full[s]);
        if (++s == NS) { s = 0; ph ^= 1u; }
      }
    }
  } else if (warp < kHfProducerWarps + kHfEpiWarps) {
    // ===================================== epilogue ================================================
    const int quarter = warp & 3, half = (warp - kHfProducerWarps) >> 2;
    const int ch = C / 2, c_lo = half * ch;                  // this warp's classes
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const float gscale_all = s_gscale;
    pk2 tot = pack2(0.f, 0.f);
    float tot_pos = 0.f;
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = tile_first + it;
      const int b = tile / tiles_per_image, p = (tile - b * tiles_per_image) * kTileM + quarter * 32 + lane;
      const int ab = it & 1;
      const bool valid = p < P;                  // ragged tile: the pixel contributes nothing (coefficient 0) and is not stored
      const float gscale = valid ? gscale_all : 0.f;
      const float* g_src = gt + ((size_t)b * C + c_lo) * P + (valid ? p : 0);
      float* d_dst = (d_logits && valid) ? d_logits + ((size_t)b * C + c_lo) * P + p : nullptr;
      // gt does not depend on the logits: the first group's loads are issued before the accumulator is waited for,
      // group g + 1's while group g is computed
      float g[2][8];
#pragma unroll
      for (int k = 0; k < 8; ++k) g[0][k] = ldg_stream1(g_src + (size_t)k * P);
      dcn_mbar_wait(&bars.acc_full[ab], (unsigned)((it >> 1) & 1));
      tc_fence_after();
#pragma unroll
      for (int g8 = 0; g8 < 8; ++g8) {
        const int c0 = g8 * 8;
        if (c0 < ch) {
          if (c0 + 8 < ch) {
#pragma unroll
            for (int k = 0; k < 8; ++k) g[(g8 + 1) & 1][k] = ldg_stream1(g_src + (size_t)(c0 + 8 + k) * P);
          }
          uint32_t v[8];
          tmem_ld8(tmem + lane_base + (uint32_t)(ab * 128 + c_lo + c0), v);
#pragma unroll
          for (int k = 0; k < 8; k += 2) {
            const int c = c_lo + c0 + k;
            const pk2 d = cell2<false>(__uint_as_float(v[k]) + s_bias[c], __uint_as_float(v[k + 1]) + s_bias[c + 1],
                                       g[g8 & 1][k], g[g8 & 1][k + 1], s_alpha[c], gscale, tot, tot_pos, s_alpha[c + 1]);
            if (d_dst != nullptr) {
              float d0, d1;
              unpack2(d, d0, d1);
              d_dst[(size_t)(c0 + k) * P] = d0;
              d_dst[(size_t)(c0 + k + 1) * P] = d1;
            }
          }
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.acc_empty[ab]);
    }
    float t0, t1;
    unpack2(tot, t0, t1);
    acc = (double)t0 + (double)t1 + (double)tot_pos;
  } else {
    // ===================================== MMA issuer ==============================================
    if (elect_one()) {
      const uint32_t idesc = make_idesc_bf16(kTileM, C);
      const uint32_t w_addr = smem_u32(wsm);
      int s = 0; unsigned ph = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int ab = it & 1;
        dcn_mbar_wait_relaxed(&bars.acc_empty[ab], (unsigned)(((it >> 1) & 1) ^ 1), 0u);
        tc_fence_after();
        for (int kb = 0; kb < KB; ++kb) {
          dcn_mbar_wait_relaxed(&bars.full[s], ph, 0u);
          tc_fence_after();
          const uint64_t adesc = make_sw128_desc(smem_u32(ring + (size_t)s * (kTileM * 128)));
          const uint64_t bdesc = make_sw128_desc(w_addr + (uint32_t)kb * (uint32_t)C * 128u);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4)
            umma_bf16(tmem + (uint32_t)(ab * 128), adesc + 2 * k4, bdesc + 2 * k4, idesc, (uint32_t)((kb | k4) != 0));
          umma_commit(&bars.empty[s]);
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        umma_commit(&bars.acc_full[ab]);
      }
    }
    __syncwarp();
  }
  // ---- loss: warp sums -> fp64 atomic; the last CTA writes the scalar and re-zeroes the workspace -----
  acc = warp_sum(acc);
  if (lane == 0 && acc != 0.0) atomicAdd(&ws->pos_sum, acc);
  tc_fence_before();
  __syncthreads();
  if (warp == kHfProducerWarps + kHfEpiWarps) tmem_dealloc(tmem, 256);
  if (tid == 0) {
    __threadfence();
    s_last = (atomicAdd(&ws->done, 1u) == gridDim.x - 1);
    if (s_last) {
      __threadfence();
      losses[0] = (float)*((volatile double*)&ws->pos_sum);
      ws->pos_sum = 0.0; ws->neg_sum = 0.0; ws->wh_sum = 0.0; ws->off_sum = 0.0;
      ws->npos_count = 0ull; ws->done = 0u; ws->nmask = 0.f;
    }
  }
}

// ================================================================================================================
// The same with the convolution's backward inside: d_feat = d_logits * W, d_weight = d_logits^T * feat, d_bias = sum
// d_logits -- neither the logits nor their gradient reach HBM (feat in, d_feat out: 8 * Cf + 4 * C bytes per pixel).
//   GEMM 1  logits[128 px, C]     = feat[128 px, Cf] * W^T                 (as above; one TMEM buffer)
//   epilogue 1: loss + d_logits -> bf16 tile D[128 px][C] in shared memory (two 64-class atoms, K-major rows);
//               d_bias by a reduce-scatter butterfly over the warp (9 shuffles per 8 classes)
//   GEMM 2  d_feat[128 px, Cf]    = D[128 px, C] * W[C, Cf]                (B = the W image read MN-major: K = classes)
//   GEMM 3  dW^T[Cf, C]          += feat^T[Cf, 128 px] * D[128 px, C]      (A = two adjacent ring stages read MN-major,
//                                                                           B = D read MN-major; persistent in TMEM)
//   epilogue 2: d_feat TMEM -> NCHW fp32 (lanes = pixels: coalesced)
// TMEM (512 columns): logits [0, C) | dW^T tiles [C, C + (Cf/128) C) | d_feat [512 - Cf, 512): C = 80, Cf = 256 uses 496.
struct HeadFusedBwdBarriers {
  uint64_t full[kHfMaxStages], empty[kHfMaxStages];   // feat ring (producers -> GEMM 1 / GEMM 3)
  uint64_t lg_full, lg_empty;                         // logits in TMEM          (MMA -> epilogue 1)
  uint64_t d_full, d_empty;                           // d_logits tile in smem   (epilogue 1 -> GEMM 2 / 3)
  uint64_t df_full, df_empty;                         // d_feat in TMEM          (MMA -> epilogue 2)
  uint64_t dw_full;                                   // dW^T complete           (MMA -> final flush)
};

__global__ void __launch_bounds__(kHfThreads, 1)
head_focal_fused_bwd_kernel(const float* __restrict__ feat, const float* __restrict__ weight, const float* __restrict__ bias,
                            const float* __restrict__ gt, const float* __restrict__ alpha, const int32_t* __restrict__ num_pos_img,
                            const float* __restrict__ norm_override, int B, int Cf, int C, int P, int flags, int stages,
                            float w_hm, float* __restrict__ losses, float* __restrict__ d_logits, float* __restrict__ d_feat,
                            float* __restrict__ d_weight, float* __restrict__ d_bias, LossWs* ws) {
  extern __shared__ __align__(1024) uint8_t hf_smem[];
  __shared__ HeadFusedBwdBarriers bars;
  __shared__ uint32_t s_tmem;
  __shared__ float s_gscale;
  __shared__ bool s_last;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int KB = Cf / kBK, MT = Cf / kTileM;
  const int NS = stages;
  uint8_t* smem = hf_smem + ((1024u - (smem_u32(hf_smem) & 1023u)) & 1023u);
  uint8_t* ring = smem;                                     // [NS][128 px][128 B]
  uint8_t* dtile = ring + (size_t)NS * (kTileM * 128);      // [2 atoms of 64 classes][128 px][128 B]
  uint8_t* wsm = dtile + 2 * (kTileM * 128);                // [KB][C rows][128 B]
  float* s_bias = reinterpret_cast<float*>(wsm + (size_t)KB * C * 128);
  float* s_alpha = s_bias + C;
  const int tiles_per_image = (P + kTileM - 1) / kTileM;      // the last tile of an image may be ragged
  const int num_tiles = B * tiles_per_image;
  const int tile_first = (int)(((long long)blockIdx.x * num_tiles) / gridDim.x);
  const int my_tiles = (int)(((long long)(blockIdx.x + 1) * num_tiles) / gridDim.x) - tile_first;
  const uint32_t col_dw = (uint32_t)C, col_df = (uint32_t)(512 - Cf);

  if (tid == 0) {
    for (int s = 0; s < kHfMaxStages; ++s) { dcn_mbar_init(&bars.full[s], kHfProducerWarps); dcn_mbar_init(&bars.empty[s], 1); }
    dcn_mbar_init(&bars.lg_full, 1);   dcn_mbar_init(&bars.lg_empty, kHfEpiWarps);
    dcn_mbar_init(&bars.d_full, kHfEpiWarps);  dcn_mbar_init(&bars.d_empty, 1);
    dcn_mbar_init(&bars.df_full, 1);   dcn_mbar_init(&bars.df_empty, kHfProducerWarps);
    dcn_mbar_init(&bars.dw_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    float npos = 0.f;
    if (norm_override != nullptr) npos = norm_override[0];
    else if (num_pos_img != nullptr) { int a = 0; for (int i = 0; i < B; ++i) a += num_pos_img[i]; npos = (float)a; }
    else npos = (float)ws->npos_count;
    s_gscale = -(npos == 0.f ? w_hm : w_hm / npos);
  }
  if (warp == kHfProducerWarps + kHfEpiWarps) tmem_alloc(&s_tmem, 512);
  for (int i = tid; i < C * (Cf / 8); i += kHfThreads) {      // W -> bf16, swizzled (see head_focal_fused_kernel)
    const int n = i / (Cf / 8), j8 = i - n * (Cf / 8);
    const float4 a = __ldg(reinterpret_cast<const float4*>(weight + (size_t)n * Cf + j8 * 8));
    const float4 b = __ldg(reinterpret_cast<const float4*>(weight + (size_t)n * Cf + j8 * 8 + 4));
    const __nv_bfloat162 p0 = __floats2bfloat162_rn(a.x, a.y), p1 = __floats2bfloat162_rn(a.z, a.w);
    const __nv_bfloat162 p2 = __floats2bfloat162_rn(b.x, b.y), p3 = __floats2bfloat162_rn(b.z, b.w);
    const int kb = j8 >> 3, j = j8 & 7;
    *reinterpret_cast<uint4*>(wsm + (size_t)kb * C * 128 + n * 128 + ((j ^ (n & 7)) << 4)) =
        make_uint4(*reinterpret_cast<const uint32_t*>(&p0), *reinterpret_cast<const uint32_t*>(&p1),
                   *reinterpret_cast<const uint32_t*>(&p2), *reinterpret_cast<const uint32_t*>(&p3));
  }
  for (int i = tid; i < 2 * kTileM * 128 / 16; i += kHfThreads) reinterpret_cast<uint4*>(dtile)[i] = make_uint4(0u, 0u, 0u, 0u);
  for (int i = tid; i < C; i += kHfThreads) {
    s_bias[i] = bias ? __ldg(bias + i) : 0.f;
    s_alpha[i] = (flags & CTDET_LOSS_ALPHA_UNIFORM) ? __ldg(alpha) : __ldg(alpha + i);
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  double acc = 0.0;
  if (warp < kHfProducerWarps) {
    // ===================================== producers + epilogue 2 ==================================
    // Fill the ring with the next tile (the loads of K block kb + 1 in flight while kb is converted and stored), then
    // drain d_feat of the PREVIOUS tile from TMEM (any warp reads the TMEM lane quarter warp & 3): by the time GEMM 2 of
    // that tile has run, its GEMM 3 is about to free the ring slots this warp would otherwise be waiting for.
    int s = 0; unsigned ph = 0;
    const uint32_t swz = (uint32_t)((warp ^ (lane & 7)) << 4);              // rows l + 32 j: row & 7 = lane & 7
    const int quarter = warp & 3, fh = Cf / 2, f_lo = (warp >> 2) * fh;     // epilogue 2: this warp's pixels and channels
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    auto drain = [&](int itd) {
      const int tile = tile_first + itd;
      const int b = tile / tiles_per_image, p = (tile - b * tiles_per_image) * kTileM + quarter * 32 + lane;
      const bool valid = p < P;
      float* fq = d_feat + ((size_t)b * Cf + f_lo) * P + (valid ? p : 0);
      dcn_mbar_wait(&bars.df_full, (unsigned)(itd & 1));
      tc_fence_after();
#pragma unroll 1
      for (int c0 = 0; c0 < fh; c0 += 16) {
        uint32_t w[16];
        tmem_ld16(tmem + lane_base + col_df + (uint32_t)(f_lo + c0), w);
#pragma unroll
        for (int k = 0; k < 16; ++k) { if (valid) *fq = __uint_as_float(w[k]); fq += P; }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.df_empty);
    };
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = tile_first + it;
      const int b = tile / tiles_per_image, p0 = (tile - b * tiles_per_image) * kTileM;
      const float* src = feat + ((size_t)b * Cf + warp * 8) * P + p0 + lane;
      float v[2][8][4];
      auto load_kb = [&](int kb, float (&dst)[8][4]) {
        const float* q = src + (size_t)kb * kBK * P;
        if (p0 + kTileM <= P) {
#pragma unroll
          for (int i = 0; i < 8; ++i) {
#pragma unroll
            for (int j = 0; j < 4; ++j) dst[i][j] = ldg_stream1(q + 32 * j);
            q += P;
          }
        } else {                                 // ragged tile: rows past the image are zero
#pragma unroll
          for (int i = 0; i < 8; ++i) {
#pragma unroll
            for (int j = 0; j < 4; ++j) dst[i][j] = (p0 + lane + 32 * j < P) ? ldg_stream1(q + 32 * j) : 0.f;
            q += P;
          }
        }
      };
      load_kb(0, v[0]);
#pragma unroll 1
      for (int kb2 = 0; kb2 < KB; kb2 += 2) {
#pragma unroll
        for (int u = 0; u < 2; ++u) {
          const int kb = kb2 + u;
          if (kb >= KB) break;
          if (kb + 1 < KB) load_kb(kb + 1, v[u ^ 1]);
          dcn_mbar_wait(&bars.empty[s], ph ^ 1u);
          uint8_t* a_tile = ring + (size_t)s * (kTileM * 128);
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            const __nv_bfloat162 q0 = __floats2bfloat162_rn(v[u][0][j], v[u][1][j]), q1 = __floats2bfloat162_rn(v[u][2][j], v[u][3][j]);
            const __nv_bfloat162 q2 = __floats2bfloat162_rn(v[u][4][j], v[u][5][j]), q3 = __floats2bfloat162_rn(v[u][6][j], v[u][7][j]);
            *reinterpret_cast<uint4*>(a_tile + (uint32_t)(lane + 32 * j) * 128u + swz) =
                make_uint4(*reinterpret_cast<const uint32_t*>(&q0), *reinterpret_cast<const uint32_t*>(&q1),
                           *reinterpret_cast<const uint32_t*>(&q2), *reinterpret_cast<const uint32_t*>(&q3));
          }
          asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
          __syncwarp();
          if (lane == 0) dcn_mbar_arrive(&bars.full[s]);
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
      }
      if (it >= 1) drain(it - 1);
    }
    if (my_tiles > 0) drain(my_tiles - 1);
  } else if (warp < kHfProducerWarps + kHfEpiWarps) {
    // ===================================== epilogues ===============================================
    const int quarter = warp & 3, half = (warp - kHfProducerWarps) >> 2;
    const int ch = C / 2, c_lo = half * ch;
    const int r = quarter * 32 + lane;                        // row of the tile = TMEM lane
    const uint32_t lane_base = (uint32_t)(quarter * 32) << 16;
    const float gscale_all = s_gscale;
    const bool b4 = (lane & 4) != 0, b2 = (lane & 2) != 0, b1 = (lane & 1) != 0;
    pk2 tot = pack2(0.f, 0.f);
    float tot_pos = 0.f;
    float accb[8];                                            // d_bias: lane (l & 7) of group g owns class c_lo + 8 g + (l & 7)
#pragma unroll
    for (int g8 = 0; g8 < 8; ++g8) accb[g8] = 0.f;
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = tile_first + it;
      const int b = tile / tiles_per_image, p = (tile - b * tiles_per_image) * kTileM + r;
      const unsigned par = (unsigned)(it & 1);
      const bool valid = p < P;                  // ragged tile: coefficient 0 (no loss, zero row of D), nothing stored
      const float gscale = valid ? gscale_all : 0.f;
      const float* g_src = gt + ((size_t)b * C + c_lo) * P + (valid ? p : 0);
      float* d_dst = (d_logits && valid) ? d_logits + ((size_t)b * C + c_lo) * P + p : nullptr;
      float g[2][8];                                          // gt: group g + 1 in flight while group g is computed
#pragma unroll
      for (int k = 0; k < 8; ++k) g[0][k] = ldg_stream1(g_src + (size_t)k * P);
      dcn_mbar_wait(&bars.lg_full, par);
      tc_fence_after();
      dcn_mbar_wait(&bars.d_empty, par ^ 1u);                 // GEMM 2 / 3 of the previous tile are done with D
#pragma unroll
      for (int g8 = 0; g8 < 8; ++g8) {
        const int c0 = g8 * 8;
        if (c0 < ch) {
          if (c0 + 8 < ch) {
#pragma unroll
            for (int k = 0; k < 8; ++k) g[(g8 + 1) & 1][k] = ldg_stream1(g_src + (size_t)(c0 + 8 + k) * P);
          }
          uint32_t v[8];
          tmem_ld8(tmem + lane_base + (uint32_t)(c_lo + c0), v);
          float dd[8];
#pragma unroll
          for (int k = 0; k < 8; k += 2) {
            const int c = c_lo + c0 + k;
            const pk2 d = cell2<false>(__uint_as_float(v[k]) + s_bias[c], __uint_as_float(v[k + 1]) + s_bias[c + 1],
                                       g[g8 & 1][k], g[g8 & 1][k + 1], s_alpha[c], gscale, tot, tot_pos, s_alpha[c + 1]);
            unpack2(d, dd[k], dd[k + 1]);
            if (d_dst != nullptr) { d_dst[(size_t)(c0 + k) * P] = dd[k]; d_dst[(size_t)(c0 + k + 1) * P] = dd[k + 1]; }
          }
          {                                                   // 8 classes of this pixel -> one 16-byte chunk of D
            const __nv_bfloat162 q0 = __floats2bfloat162_rn(dd[0], dd[1]), q1 = __floats2bfloat162_rn(dd[2], dd[3]);
            const __nv_bfloat162 q2 = __floats2bfloat162_rn(dd[4], dd[5]), q3 = __floats2bfloat162_rn(dd[6], dd[7]);
            const int j = (c_lo + c0) >> 3;                   // chunk of 8 classes: atom j >> 3, chunk j & 7
            *reinterpret_cast<uint4*>(dtile + (size_t)(j >> 3) * (kTileM * 128) + (uint32_t)r * 128u + (uint32_t)(((j & 7) ^ (r & 7)) << 4)) =
                make_uint4(*reinterpret_cast<const uint32_t*>(&q0), *reinterpret_cast<const uint32_t*>(&q1),
                           *reinterpret_cast<const uint32_t*>(&q2), *reinterpret_cast<const uint32_t*>(&q3));
          }
          {                                                   // d_bias: reduce-scatter over the warp
            float r4[4], r2[2];
#pragma unroll
            for (int i = 0; i < 4; ++i) r4[i] = (b4 ? dd[i + 4] : dd[i]) + __shfl_xor_sync(0xffffffffu, b4 ? dd[i] : dd[i + 4], 4);
#pragma unroll
            for (int i = 0; i < 2; ++i) r2[i] = (b2 ? r4[i + 2] : r4[i]) + __shfl_xor_sync(0xffffffffu, b2 ? r4[i] : r4[i + 2], 2);
            float r1 = (b1 ? r2[1] : r2[0]) + __shfl_xor_sync(0xffffffffu, b1 ? r2[0] : r2[1], 1);
            r1 += __shfl_xor_sync(0xffffffffu, r1, 8);
            r1 += __shfl_xor_sync(0xffffffffu, r1, 16);
            accb[g8] += r1;
          }
        }
      }
      tc_fence_before();
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) { dcn_mbar_arrive(&bars.lg_empty); dcn_mbar_arrive(&bars.d_full); }
    }
    float t0, t1;
    unpack2(tot, t0, t1);
    acc = (double)t0 + (double)t1 + (double)tot_pos;
    // ---- d_bias and dW^T (persistent TMEM accumulators) -> fp32 atomics ----
#pragma unroll
    for (int g8 = 0; g8 < 8; ++g8)
      if (g8 * 8 < ch && lane < 8 && d_bias != nullptr) atomicAdd(d_bias + c_lo + g8 * 8 + lane, accb[g8]);
    if (my_tiles > 0) {
      dcn_mbar_wait(&bars.dw_full, 0u);
      tc_fence_after();
      for (int mt = 0; mt < MT; ++mt) {
        float* w_dst = d_weight + (size_t)c_lo * Cf + mt * kTileM + r;        // [class][channel]: lanes = channels
        for (int c0 = 0; c0 < ch; c0 += 8) {
          uint32_t v[8];
          tmem_ld8(tmem + lane_base + col_dw + (uint32_t)(mt * C + c_lo + c0), v);
#pragma unroll
          for (int k = 0; k < 8; ++k) atomicAdd(w_dst + (size_t)(c0 + k) * Cf, __uint_as_float(v[k]));
        }
      }
      tc_fence_before();
    }
  } else {
    // ===================================== MMA issuer ==============================================
    if (elect_one()) {
      const uint32_t idesc1 = make_idesc_bf16(kTileM, C);                          // logits: A, B K-major
      const uint32_t idesc2 = make_idesc_bf16(kTileM, Cf) | (1u << 16);            // d_feat: B = W read MN-major
      const uint32_t idesc3 = make_idesc_bf16(kTileM, C) | (1u << 15) | (1u << 16);   // dW^T: A = feat^T, B = D, both MN-major
      const uint32_t w_addr = smem_u32(wsm), d_addr = smem_u32(dtile);
      int s = 0; unsigned ph = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const unsigned par = (unsigned)(it & 1);
        const int s0 = s;
        dcn_mbar_wait_relaxed(&bars.lg_empty, par ^ 1u, 0u);
        tc_fence_after();
        for (int kb = 0; kb < KB; ++kb) {                     // GEMM 1
          dcn_mbar_wait_relaxed(&bars.full[s], ph, 0u);
          tc_fence_after();
          const uint64_t adesc = make_sw128_desc(smem_u32(ring + (size_t)s * (kTileM * 128)));
          const uint64_t bdesc = make_sw128_desc(w_addr + (uint32_t)kb * (uint32_t)C * 128u);
#pragma unroll
          for (int k4 = 0; k4 < 4; ++k4) umma_bf16(tmem, adesc + 2 * k4, bdesc + 2 * k4, idesc1, (uint32_t)((kb | k4) != 0));
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        umma_commit(&bars.lg_full);
        dcn_mbar_wait_relaxed(&bars.d_full, par, 0u);
        dcn_mbar_wait_relaxed(&bars.df_empty, par ^ 1u, 0u);
        tc_fence_after();
        for (int st = 0; st < C / 16; ++st) {                 // GEMM 2: K = 16 classes per step
          const uint64_t adesc = make_sw128_desc(d_addr + (uint32_t)(st >> 2) * (kTileM * 128)) + 2 * (st & 3);
          const uint64_t bdesc = make_sw128_desc_mn(w_addr + (uint32_t)st * 2048u, (uint32_t)C * 128u);
          umma_bf16(tmem + col_df, adesc, bdesc, idesc2, (uint32_t)(st != 0));
        }
        umma_commit(&bars.df_full);
        for (int mt = 0; mt < MT; ++mt) {                     // GEMM 3: K = 16 pixels per step
          const uint32_t a_addr = smem_u32(ring + (size_t)(s0 + 2 * mt) * (kTileM * 128));
#pragma unroll
          for (int k8 = 0; k8 < kTileM / 16; ++k8) {
            const uint64_t adesc = make_sw128_desc_mn(a_addr + k8 * 2048u, (uint32_t)(kTileM * 128));
            const uint64_t bdesc = make_sw128_desc_mn(d_addr + k8 * 2048u, (uint32_t)(kTileM * 128));
            umma_bf16(tmem + col_dw + (uint32_t)(mt * C), adesc, bdesc, idesc3, (uint32_t)((it | k8) != 0));
          }
        }
        for (int kb = 0; kb < KB; ++kb) umma_commit(&bars.empty[s0 + kb]);
        umma_commit(&bars.d_empty);
      }
      umma_commit(&bars.dw_full);
    }
    __syncwarp();
  }
  acc = warp_sum(acc);
  if (lane == 0 && acc != 0.0) atomicAdd(&ws->pos_sum, acc);
  tc_fence_before();
  __syncthreads();
  if (warp == kHfProducerWarps + kHfEpiWarps) tmem_dealloc(tmem, 512);
  if (tid == 0) {
    __threadfence();
    s_last = (atomicAdd(&ws->done, 1u) == gridDim.x - 1);
    if (s_last) {
      __threadfence();
      losses[0] = (float)*((volatile double*)&ws->pos_sum);
      ws->pos_sum = 0.0; ws->neg_sum = 0.0; ws->wh_sum = 0.0; ws->off_sum = 0.0;
      ws->npos_count = 0ull; ws->done = 0u; ws->nmask = 0.f;
    }
  }
}
