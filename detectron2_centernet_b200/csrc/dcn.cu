// DCNv2 (modulated deformable 3x3 convolution) forward as ONE implicit GEMM on the 5th-generation
// tensor cores (tcgen05.mma, accumulators in TMEM) -- the only dense contraction of the ctdet path.
//
// Replaces modulated_deform_conv_cuda_forward (detectron2/layers/csrc/deformable/deform_conv_cuda.cu:
// 804-927): per IMAGE an im2col kernel (deform_conv_cuda_kernel.cu:786-868) writes fp32
// columns[Cin*9, H*W] to HBM, an fp32 SGEMM (at::addmm_, :904-909) reads them back, and a separate
// elementwise pass adds the bias (:924-926).  Here the bilinear-sampled, mask-modulated im2col tile
// only ever exists in shared memory:
//
//   GEMM view per 128-pixel tile:  D[128 pixels, Cout] = A[128 pixels, K = 9*Cin] * Wk[Cout, K]^T
//   K is ordered tap-major (k = tap*Cin + ci) so a 64-wide K block shares one sampling position.
//
//   * pre-pass 1: x NCHW fp32 -> NHWC bf16 (a sampled corner becomes one contiguous channel vector:
//     16-byte loads, 8 channels each; x is read ~36 times by the sampler, once by this pass)
//   * pre-pass 2: weight [Cout,Cin,3,3] fp32 -> Wk [Cout, 9*Cin] bf16, tap-major
//   * main kernel, one CTA (4 warps) per tile, 2-stage shared-memory pipeline:
//       all 128 threads (thread = pixel) build the A stage: 4 corners x Cin/8 16-byte gathers,
//       bilinear * mask blend in packed bf16x2 FMAs (or fp32 with DCN_FLAG_FP32_BLEND), 16-byte
//       stores into the 128B-swizzled K-major layout the UMMA descriptor expects; the Wk block is
//       copied beside it; fence.proxy.async + barrier; ONE elected thread issues 4 tcgen05.mma
//       (M=128, N=Cout, K=16, bf16 x bf16 -> fp32 in TMEM) and a tcgen05.commit onto the stage's
//       mbarrier, so the tensor pipe works on stage s while the SIMT sampler fills stage s^1.
//     epilogue: tcgen05.ld (32 lanes x 16 columns per warp) -> bias (+ optional per-channel affine
//     and ReLU: DeformConvV2.actf = BN + ReLU, deform_conv.py:501-504, folded for inference) ->
//     coalesced NCHW fp32 stores.
// Border rule identical to the reference: a tap samples iff h_im > -1 && w_im > -1 && h_im < H &&
// w_im < W (:852); each bilinear corner outside the image contributes 0 (:683-693).
#include "dcn_common.cuh"

namespace ctdet {


// ------------------------------------------------------------------------------------------------
// pre-pass 1: NCHW fp32 -> NHWC bf16.  One CTA transposes a 64-channel x 64-pixel block through
// shared memory: float4 loads along the pixel axis, 16-byte (8-channel) stores along the channel
// axis (C % 64 == 0 on this path).  Optionally accumulates per-channel sums over (batch, pixels).
__global__ void __launch_bounds__(256)
dcn_pack_input_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ xh, int C, int P, int vec,
                      float* __restrict__ chan_sum, int chunks, int border_w) {
  // [pixel][32 words of 2 channels], word w of pixel row r stored at w ^ s(r), s(r) = ((q & 7) << 2) | ((q >> 3) << 1)
  // with q = r >> 2: the transposing stores of a warp (16 pixel groups q x 2 channel pairs) hit 32 distinct banks, and a
  // 16-byte read of 8 channels stays one aligned vector (its two halves swapped when q >= 8)
  __shared__ uint32_t t[64][32];
  const int b = blockIdx.z, c0 = blockIdx.y * 64;
  const float* src = x + (size_t)b * C * P;
  // border_w = W > 0: the image is written inside a one-pixel border (rows of W + 2 pixels, pixel (y, x) at
  // (y + 1, x + 1)); the border itself is zeroed by the caller
  const size_t img_pixels = border_w ? (size_t)(P / border_w + 2) * (border_w + 2) : (size_t)P;
  __nv_bfloat16* dst = xh + (size_t)b * img_pixels * C;
  const int tid = threadIdx.x;
  // with channel sums a CTA walks `chunks` 64-pixel blocks and keeps the sums in registers: one
  // atomic per channel per CTA instead of one per 64 pixels (8192 same-address atomics per channel
  // on a 64-channel 128^2 x 32 tensor cost 4x the pack itself)
  float csum[4] = {0.f, 0.f, 0.f, 0.f};                   // [k][channel of the pair]
  for (int it = 0; it < chunks; ++it) {
    const int p0 = (blockIdx.x * chunks + it) * 64;
    if (p0 >= P) break;
    if (it) __syncthreads();                              // the previous block's stores are done with t
    float4 v[2][2];
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int idx = tid + k * 256;                      // 32 channel pairs x 16 float4
      const int cp = idx >> 4, q = idx & 15;
      const int p = p0 + q * 4;
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        const int c = c0 + 2 * cp + h;
        float4 u = make_float4(0.f, 0.f, 0.f, 0.f);
        if (c < C && p < P) {
          const float* sp = src + (size_t)c * P + p;
          if (vec) {
            u = ldg_stream4(sp);
          } else {                                        // P % 4 != 0 or unaligned base: scalar loads
            u.x = __ldg(sp);
            if (p + 1 < P) u.y = __ldg(sp + 1);
            if (p + 2 < P) u.z = __ldg(sp + 2);
            if (p + 3 < P) u.w = __ldg(sp + 3);
          }
        }
        v[k][h] = u;
      }
    }
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int idx = tid + k * 256;
      const int cp = idx >> 4, q = idx & 15;
      const int col = cp ^ (((q & 7) << 2) | ((q >> 3) << 1));
      csum[2 * k] += (v[k][0].x + v[k][0].y) + (v[k][0].z + v[k][0].w);
      csum[2 * k + 1] += (v[k][1].x + v[k][1].y) + (v[k][1].z + v[k][1].w);
      __nv_bfloat162 w;
      w = __floats2bfloat162_rn(v[k][0].x, v[k][1].x); t[q * 4 + 0][col] = *reinterpret_cast<const uint32_t*>(&w);
      w = __floats2bfloat162_rn(v[k][0].y, v[k][1].y); t[q * 4 + 1][col] = *reinterpret_cast<const uint32_t*>(&w);
      w = __floats2bfloat162_rn(v[k][0].z, v[k][1].z); t[q * 4 + 2][col] = *reinterpret_cast<const uint32_t*>(&w);
      w = __floats2bfloat162_rn(v[k][0].w, v[k][1].w); t[q * 4 + 3][col] = *reinterpret_cast<const uint32_t*>(&w);
    }
    __syncthreads();
#pragma unroll
    for (int k = 0; k < 2; ++k) {
      const int idx = tid + k * 256;                      // 64 pixels x 8 chunks of 8 channels
      const int pp = idx >> 3, ch = idx & 7;
      if (p0 + pp < P && c0 + ch * 8 < C) {
        int q = p0 + pp;
        if (border_w) { const int y = q / border_w; q += (y + 1) * 2 + border_w + 1; }   // (y+1)*(W+2) + x+1
        const int qq = pp >> 2;
        uint4 u = *reinterpret_cast<const uint4*>(&t[pp][(ch ^ (qq & 7)) << 2]);
        if (qq >> 3) u = make_uint4(u.z, u.w, u.x, u.y);
        *reinterpret_cast<uint4*>(dst + (size_t)q * C + c0 + ch * 8) = u;
      }
    }
  }
  if (chan_sum) {                                         // per-channel sums (grad_bias of the backward):
#pragma unroll
    for (int k = 0; k < 4; ++k) {                         // 16 consecutive lanes share one channel pair
      float sum = csum[k];
#pragma unroll
      for (int o = 8; o > 0; o >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
      const int c = 2 * ((tid + (k >> 1) * 256) >> 4) + (k & 1);
      if ((tid & 15) == 0 && c0 + c < C) atomicAdd(chan_sum + c0 + c, sum);
    }
  }
}

// pre-pass 2: weight [Cout][Cin][9] fp32 -> bf16 B-operand tiles, one per K block kb = cb*9 + t
// (cb = 64-channel block, t = tap): tile kb is the [Cout][64] K-major slab in EXACTLY the 128-byte
// swizzled shared-memory image the UMMA descriptor expects (row n at n*128 B, 16-byte chunk ch at
// (ch ^ (n & 7)) << 4), so the main kernel moves it with one 1-D bulk async copy.
__global__ void dcn_pack_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wk,
                                       int Cout, int Cin, int cout_real) {    // rows >= cout_real are zero (N padding)
  const size_t n_el = (size_t)Cout * Cin * 9;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_el; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i & 63);                         // channel inside the block
    const int n = (int)((i >> 6) % Cout);
    const int kb = (int)(i / ((size_t)Cout * 64));
    const int cb = kb / 9, t = kb - 9 * cb;
    const int ci = cb * 64 + c;
    const int ch = c >> 3, e = c & 7;
    wk[(size_t)kb * Cout * 64 + (size_t)n * 64 + ((ch ^ (n & 7)) << 3) + e] =
        __float2bfloat16_rn(n < cout_real ? w[((size_t)n * Cin + ci) * 9 + t] : 0.f);
  }
}

// ------------------------------------------------------------------------------------------------
// Persistent, warp-specialised main kernel: one CTA per SM loops over 8 x 16 pixel tiles (128 = UMMA
// M = TMEM lanes).  2-D tiles keep the sampled neighbourhood of the 9 taps (~14 x 22 pixels x 64
// channels bf16 = 39 KB per channel block) resident in L1; K runs channel-block-major so that
// working set does not grow with Cin.
//   warps [0, NP)        producers: gather 4 corners (8 lanes x 16 B per pixel: one warp-wide load
//                        touches 4 full 128-byte lines), blend, store A rows (swizzled), arrive on
//                        full[s].  The gathers of K block n+1 are issued while block n is blended
//                        (register double-buffering across K blocks AND tiles): 16 loads per thread
//                        stay in flight, the L1/L2 latency is never exposed
//   warps [NP, NP+4)     epilogue: TMEM -> registers -> affine/ReLU -> NCHW fp32 (lane quarter = warp % 4)
//   warps [NP+4, NP+8)   sampling geometry of the tiles ahead (thread = pixel, all 9 taps)
//   warp  NP+8           one lane issues tcgen05.mma / tcgen05.commit
//   warp  NP+9           one lane streams the pre-swizzled B (weight) tiles with 1-D bulk copies
// TMEM holds two accumulator buffers, so the epilogue of tile i overlaps the MMAs of tile i+1.
#ifdef CTDET_PLAIN_TIMING
__device__ unsigned long long g_pl_ts[16];
#endif
template <bool kFp32Blend, int NP, bool kPlain = false>
__global__ void __launch_bounds__((NP + 10) * 32, 1)
dcn_forward_kernel(const DcnParams prm) {
  static_assert(NP % 4 == 0 && 32 % NP == 0, "producer warps must tile the 32 row groups and keep warp%4 aligned");
  extern __shared__ __align__(1024) uint8_t dcn_smem[];
  __shared__ DcnBarriers bars;
  __shared__ uint32_t s_tmem;
  __shared__ __align__(16) float s_sc[256], s_sh[256];         // epilogue: y = acc * sc[c] + sh[c]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Cin = prm.Cin, Cout = prm.Cout, NS = prm.stages;
  const int P = prm.Ho * prm.Wo, Pin = prm.H * prm.W;
  const uint32_t a_bytes = (uint32_t)kTileM * kBK * 2u, b_bytes = (uint32_t)Cout * kBK * 2u;
  const uint32_t stage_bytes = a_bytes + b_bytes;
  // dynamic shared memory is only guaranteed 16-byte aligned: round up to the 1024-byte swizzle atom
  // (pointer arithmetic on the array itself, so the compiler keeps the shared address space)
  uint8_t* smem = dcn_smem + ((1024u - (smem_u32(dcn_smem) & 1023u)) & 1023u);
  uint4* geo_w = reinterpret_cast<uint4*>(smem + (size_t)NS * stage_bytes);          // [2][9][128]
  uint32_t* geo_g = reinterpret_cast<uint32_t*>(geo_w + 2 * 9 * kTileM);             // [2][9][128]
  const int cpb = Cin / kBK;                     // channel blocks
  const int KB = 9 * cpb;
  // each CTA owns a contiguous run of tiles (row-major inside an image): consecutive tiles share
  // their halo columns in L1
  const int tile_first = (int)(((long long)blockIdx.x * prm.num_tiles) / gridDim.x);
  const int my_tiles = (int)(((long long)(blockIdx.x + 1) * prm.num_tiles) / gridDim.x) - tile_first;
  const int tile_step = 1, tile_base = tile_first;

  if (tid == 0) {
    for (int s = 0; s < NS; ++s) { dcn_mbar_init(&bars.full[s], kPlain ? NP / NS + 1 : NP + 1); dcn_mbar_init(&bars.empty[s], 1); }
    for (int i = 0; i < 2; ++i) {
      dcn_mbar_init(&bars.geo_full[i], 4);  dcn_mbar_init(&bars.geo_empty[i], NP);
      dcn_mbar_init(&bars.acc_full[i], 1);  dcn_mbar_init(&bars.acc_empty[i], kPlain ? 8 : 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  const int cout_store = prm.cout_store ? prm.cout_store : Cout;
  for (int c = tid; c < Cout; c += blockDim.x) {
    const float sc = (prm.scale && c < cout_store) ? __ldg(prm.scale + c) : 1.f;
    const float bi = (prm.bias && c < cout_store) ? __ldg(prm.bias + c) : 0.f;
    s_sc[c] = sc;
    s_sh[c] = (prm.scale && c < cout_store) ? bi * sc + __ldg(prm.shift + c) : bi;
  }
  if (warp == NP + 8) tmem_alloc(&s_tmem, prm.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const uint32_t acc_stride = prm.tmem_cols >> 1;              // columns between the two accumulators

  if (warp < NP) {
    // ===================================== producers ==============================================
    if (kPlain) dcn_plain_producer_loop<NP>(prm, smem, stage_bytes, NS, bars, warp, lane, tile_first, my_tiles, KB);
    else dcn_producer_loop<kFp32Blend, NP>(prm, smem, stage_bytes, NS, geo_w, geo_g, bars, warp, lane,
                                           tile_first, my_tiles, 0, KB);
  } else if (warp < NP + 4 || (kPlain && warp < NP + 8)) {
    // ===================================== epilogue ===============================================
    // (plain mode: the four geometry warps have no table to build and take the second half of the channels -- with
    // N = 256 the stores of one warp per TMEM lane quarter were what the kernel waited for)
    const int quarter = (warp - NP) & 3;                                  // == warp % 4: TMEM lane quarter
    const int ehalf = (warp - NP) >> 2;
    const int c_split = kPlain ? ((Cout / 16 + 1) / 2) * 16 : Cout;
    const int c_begin = ehalf ? c_split : 0, c_end = ehalf ? Cout : c_split;
    const int r = (quarter << 5) + lane;                                  // tile row = TMEM lane
    const int relu = prm.relu;
    for (int j = 0; j < my_tiles; ++j) {
      const int ab = j & 1;
      int tile = tile_base + j * tile_step;
      const int b = tile / prm.tiles_per_image;
      tile -= b * prm.tiles_per_image;
      const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
      const int eho = tile_y * kTileH + (r >> 4), ewo = tile_x * kTileW + (r & 15);
      const bool evalid = eho < prm.Ho && ewo < prm.Wo;
      float* yp = prm.y + (size_t)b * (prm.y_bstride ? (size_t)prm.y_bstride : (size_t)cout_store * P) + (evalid ? eho * prm.Wo + ewo : 0) +
                  (size_t)c_begin * P;
      dcn_mbar_wait_relaxed(&bars.acc_full[ab], (unsigned)((j >> 1) & 1));
      tc_fence_after();
      const uint32_t tbase = tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)ab * acc_stride;
      for (int c0 = c_begin; c0 < c_end; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tbase + (uint32_t)c0, v);
        float sc[16], sh[16];
#pragma unroll
        for (int k = 0; k < 4; ++k) {
          *reinterpret_cast<float4*>(&sc[4 * k]) = *reinterpret_cast<const float4*>(&s_sc[c0 + 4 * k]);
          *reinterpret_cast<float4*>(&sh[4 * k]) = *reinterpret_cast<const float4*>(&s_sh[c0 + 4 * k]);
        }
#pragma unroll
        for (int k = 0; k < 16; ++k) {
          float q = fmaf(__uint_as_float(v[k]), sc[k], sh[k]);
          if (relu) q = fmaxf(q, 0.f);
          if (evalid && c0 + k < cout_store) *yp = q;
          yp += P;
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.acc_empty[ab]);
    }
  } else if (warp < NP + 8) {
    // ===================================== geometry ===============================================
    const int r = ((warp - NP - 4) << 5) + lane;                          // tile row = pixel of the tile
    for (int it = 0; it < (kPlain ? 0 : my_tiles); ++it) {                // (plain mode: integer positions, no table)
      dcn_geometry_tile<kFp32Blend>(prm, tile_first + it, it, r, geo_w, geo_g, bars);
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.geo_full[it & 1]);
    }
  } else if (warp == NP + 8) {
    // ===================================== MMA issuer =============================================
    if (kPlain ? elect_one() : lane == 0) {
      const uint32_t idesc = make_idesc_bf16(kTileM, Cout);
      int s = 0; unsigned ph = 0;
#ifdef CTDET_PLAIN_TIMING
      long long t_acc = 0, t_full = 0, t_issue = 0, t_all0 = clock64(), nblk = 0;
#endif
      for (int it = 0; it < my_tiles; ++it) {
        const int ab = it & 1;
#ifdef CTDET_PLAIN_TIMING
        long long c0 = clock64();
#endif
        dcn_mbar_wait(&bars.acc_empty[ab], (unsigned)(((it >> 1) & 1) ^ 1));   // epilogue drained this buffer
#ifdef CTDET_PLAIN_TIMING
        t_acc += clock64() - c0;
#endif
        tc_fence_after();
        const uint32_t d_tmem = tmem + (uint32_t)ab * acc_stride;
        for (int kb = 0; kb < KB; ++kb) {
#ifdef CTDET_PLAIN_TIMING
          long long c1 = clock64();
#endif
          dcn_mbar_wait_relaxed(&bars.full[s], ph, (unsigned)prm.mma_sleep);
#ifdef CTDET_PLAIN_TIMING
          long long c2 = clock64(); t_full += c2 - c1; ++nblk;
#endif
          tc_fence_after();
          const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
          const uint64_t adesc = make_sw128_desc(a_addr);
          const uint64_t bdesc = make_sw128_desc(a_addr + a_bytes);
#pragma unroll
          for (int k4 = 0; k4 < kBK / 16; ++k4)      // +32 bytes (16 bf16) along K inside the swizzle row
            umma_bf16(d_tmem, adesc + 2 * k4, bdesc + 2 * k4, idesc, (uint32_t)((kb | k4) != 0));
          umma_commit(&bars.empty[s]);               // arrives when the MMAs above have read the stage
#ifdef CTDET_PLAIN_TIMING
          t_issue += clock64() - c2;
#endif
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        umma_commit(&bars.acc_full[ab]);             // accumulator of tile it complete
      }
#ifdef CTDET_PLAIN_TIMING
      if (kPlain && blockIdx.x == 0) {
        g_pl_ts[0] = (unsigned long long)t_acc; g_pl_ts[1] = (unsigned long long)t_full; g_pl_ts[2] = (unsigned long long)t_issue;
        g_pl_ts[3] = (unsigned long long)(clock64() - t_all0); g_pl_ts[4] = (unsigned long long)nblk;
      }
#endif
    }
    __syncwarp();
  } else {
    // ===================================== B-tile loader ==========================================
    if (kPlain ? elect_one() : lane == 0) {
      int s = 0; unsigned ph = 0;
      for (int it = 0; it < my_tiles; ++it) {
        for (int kb = 0; kb < KB; ++kb) {
          dcn_mbar_wait_relaxed(&bars.empty[s], ph ^ 1u, (unsigned)prm.ld_sleep);
          dcn_mbar_arrive_expect_tx(&bars.full[s], b_bytes);
          dcn_bulk_g2s(smem + (size_t)s * stage_bytes + a_bytes, prm.wk + (size_t)kb * Cout * kBK, b_bytes, &bars.full[s]);
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
      }
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == NP + 8) tmem_dealloc(tmem, prm.tmem_cols);
}

}  // namespace ctdet

static inline size_t dcn_align256(size_t v) { return (v + 255) / 256 * 256; }

extern "C" size_t dcnv2_workspace_bytes(int B, int Cin, int H, int W, int Cout) {
  if (B <= 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0) return 0;
  return dcn_align256((size_t)B * H * W * Cin * 2) + dcn_align256((size_t)Cout * Cin * 9 * 2);
}


extern "C" size_t dcnv2_packed_weight_bytes(int Cin, int Cout) {
  if (Cin <= 0 || Cout <= 0) return 0;
  return dcn_align256((size_t)Cout * Cin * 9 * 2);
}

extern "C" int dcnv2_pack_weight(const float* weight, int Cin, int Cout, void* packed, ctdet_stream_t stream) {
  using namespace ctdet;
  if (Cin <= 0 || Cout <= 0) return CTDET_EINVAL;
  if (Cin % kBK != 0 || Cout % 16 != 0 || Cout < 16 || Cout > 256) return CTDET_EUNSUPPORTED;
  if (!weight || !packed) return CTDET_ENULL;
  if (reinterpret_cast<uintptr_t>(packed) & 15u) return CTDET_EALIGN;
  dcn_pack_weight_kernel<<<kNumSMs, 256, 0, (cudaStream_t)stream>>>(weight, reinterpret_cast<__nv_bfloat16*>(packed), Cout, Cin, Cout);
  return check_launch();
}

// raw27 != 0: `offset` is the raw [B,27,Ho,Wo] output of conv_offset_mask (offsets = its channels 0..17 as they
// are, mask = sigmoid of channels 18..26), `om_bias` its bias or null; `mask` is ignored.
static int dcn_forward_impl(const float* input, const float* weight, const float* bias,
                            const float* offset, const float* mask, const float* om_bias, int raw27, float* output,
                            int B, int Cin, int H, int W, int Cout, int kernel_h, int kernel_w,
                            int stride_h, int stride_w, int pad_h, int pad_w, int dilation_h,
                            int dilation_w, int group, int deformable_group,
                            const float* scale, const float* shift, int flags,
                            void* workspace, size_t workspace_bytes, ctdet_stream_t stream, int input_is_packed = 0) {
  using namespace ctdet;
  if (B < 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0) return CTDET_EINVAL;
  if (flags & ~(DCN_FLAG_RELU | DCN_FLAG_FP32_BLEND | DCN_FLAG_PACKED_WEIGHT)) return CTDET_EINVAL;
  if ((scale == nullptr) != (shift == nullptr)) return CTDET_EINVAL;
  // what DLA-34's DeformConvV2 uses (deform_conv.py:505-513): 3x3, one group, one deformable group
  if (kernel_h != 3 || kernel_w != 3 || group != 1 || deformable_group != 1) return CTDET_EUNSUPPORTED;
  if (stride_h != stride_w || pad_h != pad_w || dilation_h != dilation_w || stride_h < 1 || dilation_h < 1 || pad_h < 0)
    return CTDET_EUNSUPPORTED;
  // tensor-core tiling: K blocks of 64 channels inside one tap, UMMA N = Cout
  if (Cin % kBK != 0 || Cout % 16 != 0 || Cout < 16 || Cout > 256) return CTDET_EUNSUPPORTED;
  const int Ho = (H + 2 * pad_h - (dilation_h * 2 + 1)) / stride_h + 1;
  const int Wo = (W + 2 * pad_w - (dilation_w * 2 + 1)) / stride_w + 1;
  if (Ho <= 0 || Wo <= 0) return CTDET_EINVAL;
  if (B == 0) return CTDET_OK;
  if (!input || !weight || !offset || (!mask && !raw27) || !output || !workspace) return CTDET_ENULL;
  if (workspace_bytes < dcnv2_workspace_bytes(B, Cin, H, W, Cout)) return CTDET_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15u) return CTDET_EALIGN;
  if ((size_t)H * W >= ((size_t)1 << 31) / (size_t)Cin) return CTDET_EINVAL;     // 32-bit element offsets per image
  cudaStream_t st = (cudaStream_t)stream;

  __nv_bfloat16* xh = reinterpret_cast<__nv_bfloat16*>(workspace);
  __nv_bfloat16* wk = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<unsigned char*>(workspace) +
                                                       dcn_align256((size_t)B * H * W * Cin * 2));
  const int Pin = H * W;
  if (B > 65535) return CTDET_EINVAL;
  const int vec = (Pin % 4 == 0 && aligned16(input)) ? 1 : 0;          // float4 loads of the NCHW input
  int rc = CTDET_OK;
  if (!input_is_packed) {                      // (dcnv2_forward_fused_offsets packed it for both of its kernels)
    dcn_pack_input_kernel<<<dim3((Pin + 63) / 64, (Cin + 63) / 64, B), 256, 0, st>>>(input, xh, Cin, Pin, vec, nullptr, 1, 0);
    rc = check_launch();
    if (rc) return rc;
  }
  if (flags & DCN_FLAG_PACKED_WEIGHT) {        // `weight` is the caller's dcnv2_pack_weight output for these weights
    if (reinterpret_cast<uintptr_t>(weight) & 15u) return CTDET_EALIGN;
    wk = reinterpret_cast<__nv_bfloat16*>(const_cast<float*>(weight));
  } else
  dcn_pack_weight_kernel<<<kNumSMs, 256, 0, st>>>(weight, wk, Cout, Cin, Cout);
  rc = check_launch();
  if (rc) return rc;

  DcnParams prm;
  prm.xh = xh; prm.offset = offset; prm.mask = mask; prm.wk = wk; prm.bias = bias;
  prm.off_bstride = 18 * Ho * Wo; prm.mask_bstride = 9 * Ho * Wo; prm.mask_logit = 0; prm.om_bias = nullptr;
  if (raw27) {
    if ((long long)27 * Ho * Wo > 0x7fffffffLL) return CTDET_EINVAL;
    prm.mask = offset + (size_t)18 * Ho * Wo;
    prm.off_bstride = prm.mask_bstride = 27 * Ho * Wo; prm.mask_logit = 1; prm.om_bias = om_bias;
  }
  prm.scale = scale; prm.shift = shift; prm.y = output;
  prm.Cin = Cin; prm.H = H; prm.W = W; prm.Ho = Ho; prm.Wo = Wo; prm.Cout = Cout;
  prm.stride = stride_h; prm.pad = pad_h; prm.dil = dilation_h; prm.relu = (flags & DCN_FLAG_RELU) ? 1 : 0;
  prm.xw = W; prm.ximg = Pin * Cin; prm.xorg = 0;
  prm.tiles_x = (Wo + kTileW - 1) / kTileW;
  prm.tiles_per_image = prm.tiles_x * ((Ho + kTileH - 1) / kTileH);
  const long long tiles = (long long)B * prm.tiles_per_image;
  if (tiles > 0x7fffffffLL) return CTDET_EINVAL;
  prm.num_tiles = (int)tiles;
  uint32_t cols = 32;
  while (cols < 2u * (uint32_t)Cout) cols <<= 1;             // two accumulators; TMEM allocations are powers of two >= 32
  prm.tmem_cols = cols;
  // pipeline depth: as many stages as fit beside the geometry buffers, capped -- every KB of shared
  // memory is a KB less L1 for the gathers
  const size_t stage_bytes = (size_t)kTileM * kBK * 2 + (size_t)Cout * kBK * 2;
  const size_t budget = 230 * 1024 - 1024 - kGeoBytes;
  int stages = (int)(budget / stage_bytes);
  const DcnTunables& tune = dcn_tunables();
  const int cap = tune.stages;
  if (stages > cap) stages = cap;
  if (stages > kMaxStages) stages = kMaxStages;
  if (stages < 2) return CTDET_EUNSUPPORTED;
  prm.stages = stages;
  prm.mma_sleep = tune.mma_sleep;
  prm.ld_sleep = tune.ld_sleep;
  const size_t smem = (size_t)stages * stage_bytes + kGeoBytes + 1024;
  const int np = tune.producer_warps;
  const bool f32 = (flags & DCN_FLAG_FP32_BLEND) != 0;
  void (*kern)(const DcnParams) = nullptr;
  if (np == 16) kern = f32 ? dcn_forward_kernel<true, 16> : dcn_forward_kernel<false, 16>;
  else          kern = f32 ? dcn_forward_kernel<true, 8> : dcn_forward_kernel<false, 8>;
  const int threads = ((np == 16 ? 16 : 8) + 10) * 32;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  const unsigned grid = (unsigned)(tiles < kNumSMs ? tiles : kNumSMs);
  kern<<<grid, threads, smem, st>>>(prm);
  return check_launch();
}

extern "C" int dcnv2_forward(const float* input, const float* weight, const float* bias,
                             const float* offset, const float* mask, float* output,
                             int B, int Cin, int H, int W, int Cout, int kernel_h, int kernel_w,
                             int stride_h, int stride_w, int pad_h, int pad_w, int dilation_h,
                             int dilation_w, int group, int deformable_group,
                             const float* scale, const float* shift, int flags,
                             void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  return dcn_forward_impl(input, weight, bias, offset, mask, nullptr, 0, output, B, Cin, H, W, Cout, kernel_h, kernel_w,
                          stride_h, stride_w, pad_h, pad_w, dilation_h, dilation_w, group, deformable_group,
                          scale, shift, flags, workspace, workspace_bytes, stream);
}

extern "C" int dcnv2_forward_raw27(const float* input, const float* weight, const float* bias,
                                   const float* offset_mask, const float* offset_mask_bias, float* output,
                                   int B, int Cin, int H, int W, int Cout, int stride, int pad, int dilation,
                                   const float* scale, const float* shift, int flags,
                                   void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  return dcn_forward_impl(input, weight, bias, offset_mask, nullptr, offset_mask_bias, 1, output, B, Cin, H, W, Cout, 3, 3,
                          stride, stride, pad, pad, dilation, dilation, 1, 1, scale, shift, flags,
                          workspace, workspace_bytes, stream);
}

// ------------------------------------------------------------------------------------------------
// conv_offset_mask on the tensor cores, feeding the DCN that follows: the 27-channel 3x3
// convolution of deform_conv.py:505-513 (convention resnet.py:308-313) is the main kernel in its `plain` mode -- zero
// offsets, mask 1, ONE corner gathered per (pixel, tap), N = 32 (27 real rows) -- over the same NHWC bf16 copy of x the DCN
// samples, so x is packed once, cuDNN's NCHW <-> NHWC conversions disappear and the raw 27 channels go straight into the
// DCN kernel's geometry stage (which adds the bias and applies the mask's sigmoid).
extern "C" size_t dcnv2_fused_offsets_workspace_bytes(int B, int Cin, int H, int W, int Cout) {
  if (B <= 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0) return 0;
  return dcnv2_workspace_bytes(B, Cin, H, W, Cout) + dcn_align256((size_t)32 * Cin * 9 * 2) +
         dcn_align256((size_t)B * 27 * H * W * 4);
}

extern "C" int dcnv2_forward_fused_offsets(const float* input, const float* weight, const float* bias,
                                           const float* om_weight, const float* om_bias, float* output,
                                           float* offset_mask_out, int B, int Cin, int H, int W, int Cout,
                                           const float* scale, const float* shift, int flags,
                                           void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0) return CTDET_EINVAL;
  if (flags & ~(DCN_FLAG_RELU | DCN_FLAG_FP32_BLEND | DCN_FLAG_PACKED_WEIGHT)) return CTDET_EINVAL;
  if (Cin % kBK != 0 || Cout % 16 != 0 || Cout < 16 || Cout > 256) return CTDET_EUNSUPPORTED;
  if (B == 0) return CTDET_OK;
  if (!input || !weight || !om_weight || !output || !workspace) return CTDET_ENULL;
  if (workspace_bytes < dcnv2_fused_offsets_workspace_bytes(B, Cin, H, W, Cout)) return CTDET_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15u) return CTDET_EALIGN;
  if ((size_t)H * W >= ((size_t)1 << 31) / (size_t)Cin || B > 65535 || (long long)27 * H * W > 0x7fffffffLL) return CTDET_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t base_bytes = dcnv2_workspace_bytes(B, Cin, H, W, Cout);
  unsigned char* wsp = reinterpret_cast<unsigned char*>(workspace);
  __nv_bfloat16* xh = reinterpret_cast<__nv_bfloat16*>(wsp);
  __nv_bfloat16* wk_om = reinterpret_cast<__nv_bfloat16*>(wsp + base_bytes);
  float* raw = offset_mask_out ? offset_mask_out
                               : reinterpret_cast<float*>(wsp + base_bytes + dcn_align256((size_t)32 * Cin * 9 * 2));
  const int Pin = H * W;
  const int vec = (Pin % 4 == 0 && aligned16(input)) ? 1 : 0;
  dcn_pack_input_kernel<<<dim3((Pin + 63) / 64, (Cin + 63) / 64, B), 256, 0, st>>>(input, xh, Cin, Pin, vec, nullptr, 1, 0);
  int rc = check_launch();
  if (rc) return rc;
  dcn_pack_weight_kernel<<<kNumSMs, 256, 0, st>>>(om_weight, wk_om, 32, Cin, 27);
  rc = check_launch();
  if (rc) return rc;

  DcnParams prm;
  prm.xh = xh; prm.offset = raw; prm.mask = raw; prm.wk = wk_om; prm.bias = nullptr;       // (offset / mask are not read)
  prm.off_bstride = 0; prm.mask_bstride = 0; prm.mask_logit = 0; prm.om_bias = nullptr;
  prm.scale = nullptr; prm.shift = nullptr; prm.y = raw;
  prm.Cin = Cin; prm.H = H; prm.W = W; prm.Ho = H; prm.Wo = W; prm.Cout = 32;
  prm.stride = 1; prm.pad = 1; prm.dil = 1; prm.relu = 0;
  prm.xw = W; prm.ximg = Pin * Cin; prm.xorg = 0;
  prm.tiles_x = (W + kTileW - 1) / kTileW;
  prm.tiles_per_image = prm.tiles_x * ((H + kTileH - 1) / kTileH);
  const long long tiles = (long long)B * prm.tiles_per_image;
  if (tiles > 0x7fffffffLL) return CTDET_EINVAL;
  prm.num_tiles = (int)tiles;
  prm.tmem_cols = 64;
  prm.plain = 1; prm.cout_store = 27;
  const size_t stage_bytes = (size_t)kTileM * kBK * 2 + (size_t)32 * kBK * 2;
  const DcnTunables& tune = dcn_tunables();
  static int plain_stages = 0;                  // CTDET_DCN_PLAIN_STAGES (read once): 8 (default) or 4
  if (plain_stages == 0) {
    const char* ev = getenv("CTDET_DCN_PLAIN_STAGES");
    plain_stages = ev ? atoi(ev) : 8;
    if (plain_stages != 4 && plain_stages != 8) plain_stages = 8;
  }
  const int stages = plain_stages;
  prm.stages = stages;
  prm.mma_sleep = tune.mma_sleep;
  // the weight loader is one lane that serves every K block in turn: a __nanosleep between its polls (20 ns asked for, ~0.45 us
  // slept) was free next to the 1568-cycle K blocks of the deformable kernel, but here it WAS the K block time (950 cycles
  // whatever the producers did: ncu showed them waiting for empty stages)
  prm.ld_sleep = 0;
  const size_t smem = (size_t)stages * stage_bytes + 1024;        // (plain mode: no geometry table)
  auto kern = dcn_forward_kernel<false, 16, true>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  kern<<<(unsigned)(tiles < kNumSMs ? tiles : kNumSMs), (16 + 10) * 32, smem, st>>>(prm);
  rc = check_launch();
  if (rc) return rc;
  return dcn_forward_impl(input, weight, bias, raw, nullptr, om_bias, 1, output, B, Cin, H, W, Cout, 3, 3, 1, 1, 1, 1, 1, 1, 1, 1,
                          scale, shift, flags, workspace, base_bytes, stream, 1);
}

// ------------------------------------------------------------------------------------------------
// An ordinary 3x3 convolution (stride 1, padding 1) + bias + optional ReLU on the tensor cores: the plain mode of the
// forward kernel as an entry point of its own, for the first layer of the ctdet heads (centernet.py:112-121:
// Conv2d(64, head_conv, 3, padding=1) + ReLU per branch -- 3.2 of the 3.5 ms the heads cost on cuDNN at batch 32).
// Output channels are produced in chunks of <= 128 (two TMEM accumulators of N columns and a 4-stage ring must fit).
extern "C" size_t ctdet_conv3x3_workspace_bytes(int B, int Cin, int H, int W, int Cout) {
  if (B <= 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0) return 0;
  return dcn_align256((size_t)B * H * W * Cin * 2) + dcn_align256((size_t)256 * Cin * 9 * 2);
}

// `n` convolutions of ONE input (the ctdet heads' first layers: hm, wh, reg): x is packed once.
extern "C" int ctdet_conv3x3_multi(const float* input, int n, const float* const* weights, const float* const* biases,
                                   float* const* outputs, const int* couts, int B, int Cin, int H, int W, int flags,
                                   void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || Cin <= 0 || H <= 0 || W <= 0 || n <= 0 || (flags & ~DCN_FLAG_RELU)) return CTDET_EINVAL;
  if (Cin % kBK != 0) return CTDET_EUNSUPPORTED;
  if (!input || !weights || !outputs || !couts || !workspace) return CTDET_ENULL;
  int cmax = 0;
  for (int i = 0; i < n; ++i) {
    if (couts[i] <= 0) return CTDET_EINVAL;
    if (!weights[i] || !outputs[i]) return CTDET_ENULL;
    if (couts[i] > cmax) cmax = couts[i];
  }
  if (B == 0) return CTDET_OK;
  if (workspace_bytes < ctdet_conv3x3_workspace_bytes(B, Cin, H, W, cmax)) return CTDET_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15u) return CTDET_EALIGN;
  if ((size_t)H * W >= ((size_t)1 << 31) / (size_t)Cin || B > 65535) return CTDET_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  unsigned char* wsp = reinterpret_cast<unsigned char*>(workspace);
  __nv_bfloat16* xh = reinterpret_cast<__nv_bfloat16*>(wsp);
  __nv_bfloat16* wk = reinterpret_cast<__nv_bfloat16*>(wsp + dcn_align256((size_t)B * H * W * Cin * 2));
  const int Pin = H * W;
  const int vec = (Pin % 4 == 0 && aligned16(input)) ? 1 : 0;
  dcn_pack_input_kernel<<<dim3((Pin + 63) / 64, (Cin + 63) / 64, B), 256, 0, st>>>(input, xh, Cin, Pin, vec, nullptr, 1, 0);
  int rc = check_launch();
  if (rc) return rc;
  for (int i = 0; i < n; ++i) {
    const int Cout = couts[i];
    // output channels in chunks of <= 256: two TMEM accumulators of N columns (512 in all) and a 4-stage ring of
    // 16 KB + N * 128 B stages (192 KB at N = 256; the plain mode has no geometry table)
    for (int c0 = 0; c0 < Cout; c0 += 256) {
      const int nc = Cout - c0 < 256 ? Cout - c0 : 256;
      const int n_pad = (nc + 15) / 16 * 16;
      dcn_pack_weight_kernel<<<kNumSMs, 256, 0, st>>>(weights[i] + (size_t)c0 * Cin * 9, wk, n_pad, Cin, nc);
      rc = check_launch();
      if (rc) return rc;
      DcnParams prm;
      prm.xh = xh; prm.offset = outputs[i]; prm.mask = outputs[i]; prm.wk = wk;            // (offset / mask: not read)
      prm.bias = (biases && biases[i]) ? biases[i] + c0 : nullptr;
      prm.off_bstride = 0; prm.mask_bstride = 0; prm.mask_logit = 0; prm.om_bias = nullptr;
      prm.scale = nullptr; prm.shift = nullptr; prm.y = outputs[i] + (size_t)c0 * Pin;
      prm.Cin = Cin; prm.H = H; prm.W = W; prm.Ho = H; prm.Wo = W; prm.Cout = n_pad;
      prm.stride = 1; prm.pad = 1; prm.dil = 1; prm.relu = (flags & DCN_FLAG_RELU) ? 1 : 0;
      prm.xw = W; prm.ximg = Pin * Cin; prm.xorg = 0;
      prm.tiles_x = (W + kTileW - 1) / kTileW;
      prm.tiles_per_image = prm.tiles_x * ((H + kTileH - 1) / kTileH);
      const long long tiles = (long long)B * prm.tiles_per_image;
      if (tiles > 0x7fffffffLL) return CTDET_EINVAL;
      prm.num_tiles = (int)tiles;
      uint32_t cols = 32;
      while (cols < 2u * (uint32_t)n_pad) cols <<= 1;
      prm.tmem_cols = cols;
      prm.plain = 1; prm.cout_store = nc; prm.y_bstride = (long long)Cout * Pin;
      const size_t stage_bytes = (size_t)kTileM * kBK * 2 + (size_t)n_pad * kBK * 2;
      prm.stages = stage_bytes <= 24 * 1024 ? 8 : 4;
      prm.mma_sleep = 0; prm.ld_sleep = 0;
      const size_t smem = (size_t)prm.stages * stage_bytes + 1024;   // (plain mode: no geometry table)
      auto kern = dcn_forward_kernel<false, 16, true>;
      cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
      kern<<<(unsigned)(tiles < kNumSMs ? tiles : kNumSMs), (16 + 10) * 32, smem, st>>>(prm);
      rc = check_launch();
      if (rc) return rc;
    }
  }
  return CTDET_OK;
}

extern "C" int ctdet_conv3x3(const float* input, const float* weight, const float* bias, float* output,
                             int B, int Cin, int H, int W, int Cout, int flags,
                             void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  if (Cout <= 0) return CTDET_EINVAL;
  return ctdet_conv3x3_multi(input, 1, &weight, &bias, &output, &Cout, B, Cin, H, W, flags, workspace, workspace_bytes, stream);
}

#ifdef CTDET_PLAIN_TIMING
extern "C" int ctdet_debug_plain_times(unsigned long long* out /* [16] host */) {
  return cudaMemcpyFromSymbol(out, ctdet::g_pl_ts, 16 * sizeof(unsigned long long)) == cudaSuccess ? 0 : 5;
}
#endif
