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
  const int tiles_per_image = P / kTileM;
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
    int s = 0; unsigned ph = 0;
    const uint32_t swz = (uint32_t)((warp ^ (lane & 7)) << 4);              // rows l + 32 j: row & 7 = lane & 7
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = tile_first + it;
      const int b = tile / tiles_per_image, p0 = (tile - b * tiles_per_image) * kTileM;
      const float* src = feat + ((size_t)b * Cf + warp * 8) * P + p0 + lane;
      for (int kb = 0; kb < KB; ++kb) {
        float v[8][4];
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) v[i][j] = ldg_stream1(src + ((size_t)kb * kBK + i) * P + 32 * j);
        dcn_mbar_wait(&bars.empty[s], ph ^ 1u);
        uint8_t* a_tile = ring + (size_t)s * (kTileM * 128);
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const __nv_bfloat162 q0 = __floats2bfloat162_rn(v[0][j], v[1][j]), q1 = __floats2bfloat162_rn(v[2][j], v[3][j]);
          const __nv_bfloat162 q2 = __floats2bfloat162_rn(v[4][j], v[5][j]), q3 = __floats2bfloat162_rn(v[6][j], v[7][j]);
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
    const float gscale = s_gscale;
    pk2 tot = pack2(0.f, 0.f);
    float tot_pos = 0.f;
    for (int it = 0; it < my_tiles; ++it) {
      const int tile = tile_first + it;
      const int b = tile / tiles_per_image, p = (tile - b * tiles_per_image) * kTileM + quarter * 32 + lane;
      const int ab = it & 1;
      const float* g_src = gt + ((size_t)b * C + c_lo) * P + p;
      float* d_dst = d_logits ? d_logits + ((size_t)b * C + c_lo) * P + p : nullptr;
      dcn_mbar_wait(&bars.acc_full[ab], (unsigned)((it >> 1) & 1));
      tc_fence_after();
      for (int c0 = 0; c0 < ch; c0 += 8) {
        float g[8];
#pragma unroll
        for (int k = 0; k < 8; ++k) g[k] = ldg_stream1(g_src + (size_t)(c0 + k) * P);
        uint32_t v[8];
        tmem_ld8(tmem + lane_base + (uint32_t)(ab * 128 + c_lo + c0), v);
#pragma unroll
        for (int k = 0; k < 8; k += 2) {
          const int c = c_lo + c0 + k;
          const pk2 d = cell2<false>(__uint_as_float(v[k]) + s_bias[c], __uint_as_float(v[k + 1]) + s_bias[c + 1], g[k], g[k + 1],
                                     s_alpha[c], gscale, tot, tot_pos, s_alpha[c + 1]);
          if (d_dst != nullptr) {
            float d0, d1;
            unpack2(d, d0, d1);
            d_dst[(size_t)(c0 + k) * P] = d0;
            d_dst[(size_t)(c0 + k + 1) * P] = d1;
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
    if (lane == 0) {
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
