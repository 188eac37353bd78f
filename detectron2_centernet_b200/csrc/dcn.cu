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
#include <cuda_bf16.h>
#include "common.cuh"

namespace ctdet {

constexpr int kDcnThreads = 256;      // 8 warps per CTA
constexpr int kTileM = 128;           // pixels per tile = TMEM lanes = UMMA M
constexpr int kTileH = 8, kTileW = 16; // 2-D pixel tile
constexpr int kBK = 64;               // K block: 64 bf16 = one 128-byte swizzle row
constexpr int kStages = 2;

// ------------------------------------------------------------------------------------------------
// pre-pass 1: NCHW fp32 -> NHWC bf16.  One CTA transposes a 64-channel x 64-pixel block through
// shared memory: float4 loads along the pixel axis, 16-byte (8-channel) stores along the channel
// axis (C % 64 == 0 on this path).
__global__ void __launch_bounds__(256)
dcn_pack_input_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ xh, int C, int P, int vec) {
  __shared__ __nv_bfloat16 t[64][64 + 8];                 // [pixel][channel], padded rows (144 B)
  const int b = blockIdx.z, c0 = blockIdx.y * 64, p0 = blockIdx.x * 64;
  const float* src = x + (size_t)b * C * P;
  __nv_bfloat16* dst = xh + (size_t)b * P * C;
  const int tid = threadIdx.x;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int idx = tid + k * 256;                        // 64 channels x 16 float4
    const int c = idx >> 4, q = idx & 15;
    const int p = p0 + q * 4;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (c0 + c < C && p < P) {
      const float* sp = src + (size_t)(c0 + c) * P + p;
      if (vec) {
        v = ldg_stream4(sp);
      } else {                                            // P % 4 != 0 or unaligned base: scalar loads
        v.x = __ldg(sp);
        if (p + 1 < P) v.y = __ldg(sp + 1);
        if (p + 2 < P) v.z = __ldg(sp + 2);
        if (p + 3 < P) v.w = __ldg(sp + 3);
      }
    }
    t[q * 4 + 0][c] = __float2bfloat16_rn(v.x);
    t[q * 4 + 1][c] = __float2bfloat16_rn(v.y);
    t[q * 4 + 2][c] = __float2bfloat16_rn(v.z);
    t[q * 4 + 3][c] = __float2bfloat16_rn(v.w);
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 2; ++k) {
    const int idx = tid + k * 256;                        // 64 pixels x 8 chunks of 8 channels
    const int pp = idx >> 3, ch = idx & 7;
    if (p0 + pp < P && c0 + ch * 8 < C)
      *reinterpret_cast<uint4*>(dst + (size_t)(p0 + pp) * C + c0 + ch * 8) =
          *reinterpret_cast<const uint4*>(&t[pp][ch * 8]);
  }
}

// pre-pass 2: weight [Cout][Cin][9] fp32 -> Wk [Cout][9][Cin] bf16
__global__ void dcn_pack_weight_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wk,
                                       int Cout, int Cin) {
  const size_t n = (size_t)Cout * Cin * 9;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x) {
    const int ci = (int)(i % Cin);
    const int t = (int)((i / Cin) % 9);
    const int co = (int)(i / ((size_t)Cin * 9));
    wk[i] = __float2bfloat16_rn(w[((size_t)co * Cin + ci) * 9 + t]);
  }
}

// ------------------------------------------------------------------------------------------------
// tcgen05 / mbarrier plumbing (raw PTX; SASS: UTCHMMA, LDTM, UTCBAR)
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void dcn_mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void dcn_mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "WAIT_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE_%=;\n\t"
      "bra WAIT_%=;\n\t"
      "DONE_%=:\n\t}"
      :: "r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* slot, uint32_t ncols) {      // whole warp
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" :: "r"(smem_u32(slot)), "r"(ncols) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t addr, uint32_t ncols) {     // whole warp
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" :: "r"(addr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_bf16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t acc) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      :: "r"(d_tmem), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(acc) : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {                      // arrives when all prior MMAs retire
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld16(uint32_t taddr, uint32_t (&v)[16]) {     // 32 lanes x 16 columns
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]),
        "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr));
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

// K-major, 128-byte-swizzled operand tile (rows of 64 bf16 = 128 B, 8-row groups 1024 B apart):
// start address >> 4 | LBO (unused for swizzled K-major, 1) | SBO = 1024 B | version 1 | SWIZZLE_128B
__device__ __forceinline__ uint64_t make_sw128_desc(uint32_t smem_addr) {
  return (uint64_t)((smem_addr >> 4) & 0x3fffu) | (1ull << 16) | ((uint64_t)(1024 >> 4) << 32) |
         (1ull << 46) | (2ull << 61);
}
// kind::f16 instruction descriptor: D = fp32, A = B = bf16, both K-major, N>>3 at [17,23), M>>4 at [24,29)
__device__ __forceinline__ uint32_t make_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// ------------------------------------------------------------------------------------------------
struct DcnParams {
  const __nv_bfloat16* xh;     // [B][H*W][Cin]
  const float* offset;         // [B][18][Ho*Wo]   channel 2t = dh, 2t+1 = dw (deform_conv_cuda_kernel.cu:838-846)
  const float* mask;           // [B][9][Ho*Wo]
  const __nv_bfloat16* wk;     // [Cout][9*Cin]
  const float* bias;           // [Cout] or null
  const float* scale;          // [Cout] or null   y = (conv + bias) * scale + shift
  const float* shift;          // [Cout] or null
  float* y;                    // [B][Cout][Ho*Wo]
  int Cin, H, W, Ho, Wo, Cout, stride, pad, dil, relu, tiles_per_image, tiles_x;
  uint32_t tmem_cols;
};

__device__ __forceinline__ uint32_t blend_bf16x2(uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                                 __nv_bfloat162 w1, __nv_bfloat162 w2, __nv_bfloat162 w3, __nv_bfloat162 w4) {
  __nv_bfloat162 acc = __hmul2(w1, *reinterpret_cast<__nv_bfloat162*>(&a));
  acc = __hfma2(w2, *reinterpret_cast<__nv_bfloat162*>(&b), acc);
  acc = __hfma2(w3, *reinterpret_cast<__nv_bfloat162*>(&c), acc);
  acc = __hfma2(w4, *reinterpret_cast<__nv_bfloat162*>(&d), acc);
  return *reinterpret_cast<uint32_t*>(&acc);
}
__device__ __forceinline__ uint32_t blend_fp32(uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                               float w1, float w2, float w3, float w4) {
  const float lo = w1 * __uint_as_float(a << 16) + w2 * __uint_as_float(b << 16) +
                   w3 * __uint_as_float(c << 16) + w4 * __uint_as_float(d << 16);
  const float hi = w1 * __uint_as_float(a & 0xffff0000u) + w2 * __uint_as_float(b & 0xffff0000u) +
                   w3 * __uint_as_float(c & 0xffff0000u) + w4 * __uint_as_float(d & 0xffff0000u);
  const __nv_bfloat162 r = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&r);
}

// Sampling geometry of output pixel (ho, wo) for tap t (dmcn_im2col_bilinear, :666-699): corner
// pixel indices clamped into the image and bilinear weights * mask (0 for corners outside).
__device__ __forceinline__ void tap_geometry(const DcnParams& prm, const float* offb, const float* mskb,
                                             int P, int p, int ho, int wo, bool valid, int t, int4& o, float4& wq) {
  o = make_int4(0, 0, 0, 0);
  wq = make_float4(0.f, 0.f, 0.f, 0.f);
  if (!valid) return;
  const int H = prm.H, W = prm.W;
  const int i = t / 3, j = t - 3 * i;
  const float h_im = (float)(ho * prm.stride - prm.pad + i * prm.dil) + __ldg(offb + (size_t)(2 * t) * P + p);
  const float w_im = (float)(wo * prm.stride - prm.pad + j * prm.dil) + __ldg(offb + (size_t)(2 * t + 1) * P + p);
  const float m = __ldg(mskb + (size_t)t * P + p);
  if (h_im > -1.f && w_im > -1.f && h_im < (float)H && w_im < (float)W) {       // :852
    const int h_low = (int)floorf(h_im), w_low = (int)floorf(w_im);
    const int h_high = h_low + 1, w_high = w_low + 1;
    const float lh = h_im - (float)h_low, lw = w_im - (float)w_low;
    const float hh = 1.f - lh, hw = 1.f - lw;
    const bool t_ok = h_low >= 0, b_ok = h_high <= H - 1, l_ok = w_low >= 0, r_ok = w_high <= W - 1;
    const int hl = max(h_low, 0), hh_i = min(h_high, H - 1), wl = max(w_low, 0), wh_i = min(w_high, W - 1);
    o.x = hl * W + wl;     wq.x = (t_ok && l_ok) ? hh * hw * m : 0.f;
    o.y = hl * W + wh_i;   wq.y = (t_ok && r_ok) ? hh * lw * m : 0.f;
    o.z = hh_i * W + wl;   wq.z = (b_ok && l_ok) ? lh * hw * m : 0.f;
    o.w = hh_i * W + wh_i; wq.w = (b_ok && r_ok) ? lh * lw * m : 0.f;
  }
}

// One CTA = one 8 x 16 tile of output pixels (128 = UMMA M) of one image, 8 warps.
// 2-D tiles keep the sampled neighbourhood of the 9 taps (~14 x 22 pixels x Cin bf16) resident in
// L1 between K blocks; a row-of-128 tile touches three times as many lines.
template <bool kFp32Blend>
__global__ void __launch_bounds__(kDcnThreads)
dcn_forward_kernel(const DcnParams prm) {
  extern __shared__ __align__(1024) uint8_t dcn_smem[];
  __shared__ __align__(8) uint64_t s_bar[kStages];
  __shared__ uint32_t s_tmem;
  // per-pixel geometry of the current / next tap (double buffered by tap parity): the gather is
  // done by 8 lanes per pixel (one 16-byte chunk each), so one warp-wide load touches 4 full
  // 128-byte lines instead of 32 partial ones
  __shared__ int4 s_geo_o[2][kTileM];
  __shared__ float4 s_geo_w[2][kTileM];

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Cin = prm.Cin, Cout = prm.Cout;
  const int P = prm.Ho * prm.Wo, Pin = prm.H * prm.W;
  int tile = blockIdx.x;
  const int b = tile / prm.tiles_per_image;
  tile -= b * prm.tiles_per_image;
  const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
  // geometry role: thread r and thread r+128 both describe tile row r (pixel ty = r/16, tx = r%16)
  const int r = tid & (kTileM - 1);
  const int ho = tile_y * kTileH + (r >> 4), wo = tile_x * kTileW + (r & 15);
  const bool valid = ho < prm.Ho && wo < prm.Wo;
  const int p = valid ? ho * prm.Wo + wo : 0;
  const uint32_t stage_bytes = (uint32_t)kTileM * kBK * 2u + (uint32_t)Cout * kBK * 2u;   // A tile + Wk tile
  // dynamic shared memory is only guaranteed 16-byte aligned: round up to the 1024-byte swizzle atom
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(dcn_smem) + 1023) & ~(uintptr_t)1023);

  const int cpb = Cin / kBK;                     // K blocks per tap
  const int KB = 9 * cpb;
  const size_t Ktot = (size_t)9 * Cin;
  const __nv_bfloat16* xb = prm.xh + (size_t)b * Pin * Cin;
  const float* offb = prm.offset + (size_t)b * 18 * P;
  const float* mskb = prm.mask + (size_t)b * 9 * P;

  if (tid == 0) {
    for (int s = 0; s < kStages; ++s) dcn_mbar_init(&s_bar[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 0) tmem_alloc(&s_tmem, prm.tmem_cols);
  if (tid < kTileM) {                            // tap 0
    int4 o; float4 wq;
    tap_geometry(prm, offb, mskb, P, p, ho, wo, valid, 0, o, wq);
    s_geo_o[0][r] = o; s_geo_w[0][r] = wq;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const uint32_t idesc = make_idesc_bf16(kTileM, Cout);
  const int sub = lane >> 3, chunk = lane & 7;   // 4 pixels x 8 chunks per warp-wide load

  int t = 0, ci0 = 0;                            // tap and first channel of the current K block
  for (int kb = 0; kb < KB; ++kb, ci0 += kBK) {
    if (ci0 == Cin) { ci0 = 0; ++t; }
    const int s = kb & (kStages - 1), use = kb / kStages;
    if (kb >= kStages) dcn_mbar_wait(&s_bar[s], (unsigned)((use - 1) & 1));   // MMAs that read stage s retired
    // the upper half of the CTA prepares the next tap's geometry (published by this block's barrier)
    if (ci0 == 0 && t + 1 < 9 && tid >= kTileM) {
      int4 o; float4 wq;
      tap_geometry(prm, offb, mskb, P, p, ho, wo, valid, t + 1, o, wq);
      s_geo_o[(t + 1) & 1][r] = o; s_geo_w[(t + 1) & 1][r] = wq;
    }
    uint8_t* a_tile = smem + (size_t)s * stage_bytes;
    uint8_t* b_tile = a_tile + kTileM * kBK * 2;

    // ---- A rows: warp w builds tile rows 16w .. 16w+15 in 4 passes of (4 pixels x 8 chunks) ------
    {
      const __nv_bfloat16* xc = xb + ci0 + chunk * 8;
      const int4* go = s_geo_o[t & 1];
      const float4* gw = s_geo_w[t & 1];
#pragma unroll
      for (int it = 0; it < 4; ++it) {
        const int row = (warp << 4) + (it << 2) + sub;
        const int4 o = go[row];
        const float4 wq = gw[row];
        const uint4 a = __ldg(reinterpret_cast<const uint4*>(xc + (size_t)o.x * Cin));
        const uint4 bq = __ldg(reinterpret_cast<const uint4*>(xc + (size_t)o.y * Cin));
        const uint4 cq = __ldg(reinterpret_cast<const uint4*>(xc + (size_t)o.z * Cin));
        const uint4 d = __ldg(reinterpret_cast<const uint4*>(xc + (size_t)o.w * Cin));
        uint4 q;
        if (kFp32Blend) {
          q.x = blend_fp32(a.x, bq.x, cq.x, d.x, wq.x, wq.y, wq.z, wq.w);
          q.y = blend_fp32(a.y, bq.y, cq.y, d.y, wq.x, wq.y, wq.z, wq.w);
          q.z = blend_fp32(a.z, bq.z, cq.z, d.z, wq.x, wq.y, wq.z, wq.w);
          q.w = blend_fp32(a.w, bq.w, cq.w, d.w, wq.x, wq.y, wq.z, wq.w);
        } else {
          const __nv_bfloat162 q1 = __float2bfloat162_rn(wq.x), q2 = __float2bfloat162_rn(wq.y),
                               q3 = __float2bfloat162_rn(wq.z), q4 = __float2bfloat162_rn(wq.w);
          q.x = blend_bf16x2(a.x, bq.x, cq.x, d.x, q1, q2, q3, q4);
          q.y = blend_bf16x2(a.y, bq.y, cq.y, d.y, q1, q2, q3, q4);
          q.z = blend_bf16x2(a.z, bq.z, cq.z, d.z, q1, q2, q3, q4);
          q.w = blend_bf16x2(a.w, bq.w, cq.w, d.w, q1, q2, q3, q4);
        }
        *reinterpret_cast<uint4*>(a_tile + row * 128 + ((chunk ^ (row & 7)) << 4)) = q;
      }
    }
    // ---- Wk block [Cout][64], same swizzle --------------------------------------------------------
    {
      const __nv_bfloat16* wsrc = prm.wk + (size_t)kb * kBK;
      for (int idx = tid; idx < Cout * 8; idx += kDcnThreads) {
        const int n = idx >> 3, ch = idx & 7;
        const uint4 v = __ldg(reinterpret_cast<const uint4*>(wsrc + (size_t)n * Ktot) + ch);
        *reinterpret_cast<uint4*>(b_tile + n * 128 + ((ch ^ (n & 7)) << 4)) = v;
      }
    }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> tensor core reads
    __syncthreads();
    if (tid == 0) {
      tc_fence_after();
      const uint64_t adesc = make_sw128_desc(smem_u32(a_tile));
      const uint64_t bdesc = make_sw128_desc(smem_u32(b_tile));
#pragma unroll
      for (int k4 = 0; k4 < kBK / 16; ++k4)      // +32 bytes (16 bf16) along K inside the swizzle row
        umma_bf16(tmem, adesc + 2 * k4, bdesc + 2 * k4, idesc, (uint32_t)((kb | k4) != 0));
      umma_commit(&s_bar[s]);
    }
  }
  // ---- drain: the last commit covers every MMA issued before it --------------------------------
  {
    const int last = KB - 1;
    dcn_mbar_wait(&s_bar[last % kStages], (unsigned)((last / kStages) & 1));
    tc_fence_after();
  }
  // ---- epilogue: TMEM -> registers -> bias / affine / ReLU -> NCHW fp32 ---------------------------
  // warp w may only read TMEM lanes 32*(w%4) .. +31; the two warps of a lane quarter split the
  // 16-column groups between them
  {
    const int quarter = warp & 3, half = warp >> 2;
    const int er = (quarter << 5) + lane;                       // tile row = TMEM lane of this thread
    const int eho = tile_y * kTileH + (er >> 4), ewo = tile_x * kTileW + (er & 15);
    const bool evalid = eho < prm.Ho && ewo < prm.Wo;
    float* yb = prm.y + (size_t)b * Cout * P + (evalid ? eho * prm.Wo + ewo : 0);
    for (int c0 = half * 16; c0 < Cout; c0 += 32) {
      uint32_t v[16];
      tmem_ld16(tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        float q = __uint_as_float(v[j]);
        if (prm.bias) q += __ldg(prm.bias + c0 + j);
        if (prm.scale) q = q * __ldg(prm.scale + c0 + j) + __ldg(prm.shift + c0 + j);
        if (prm.relu) q = fmaxf(q, 0.f);
        if (evalid) yb[(size_t)(c0 + j) * P] = q;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 0) tmem_dealloc(tmem, prm.tmem_cols);
}

}  // namespace ctdet

static inline size_t dcn_align256(size_t v) { return (v + 255) / 256 * 256; }

extern "C" size_t dcnv2_workspace_bytes(int B, int Cin, int H, int W, int Cout) {
  if (B <= 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0) return 0;
  return dcn_align256((size_t)B * H * W * Cin * 2) + dcn_align256((size_t)Cout * Cin * 9 * 2);
}

extern "C" int dcnv2_forward(const float* input, const float* weight, const float* bias,
                             const float* offset, const float* mask, float* output,
                             int B, int Cin, int H, int W, int Cout, int kernel_h, int kernel_w,
                             int stride_h, int stride_w, int pad_h, int pad_w, int dilation_h,
                             int dilation_w, int group, int deformable_group,
                             const float* scale, const float* shift, int flags,
                             void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0) return CTDET_EINVAL;
  if (flags & ~(DCN_FLAG_RELU | DCN_FLAG_FP32_BLEND)) return CTDET_EINVAL;
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
  if (!input || !weight || !offset || !mask || !output || !workspace) return CTDET_ENULL;
  if (workspace_bytes < dcnv2_workspace_bytes(B, Cin, H, W, Cout)) return CTDET_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15u) return CTDET_EALIGN;
  if ((size_t)H * W >= ((size_t)1 << 31) / (size_t)Cin) return CTDET_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;

  __nv_bfloat16* xh = reinterpret_cast<__nv_bfloat16*>(workspace);
  __nv_bfloat16* wk = reinterpret_cast<__nv_bfloat16*>(reinterpret_cast<unsigned char*>(workspace) +
                                                       dcn_align256((size_t)B * H * W * Cin * 2));
  const int Pin = H * W;
  if (B > 65535) return CTDET_EINVAL;
  const int vec = (Pin % 4 == 0 && aligned16(input)) ? 1 : 0;          // float4 loads of the NCHW input
  dcn_pack_input_kernel<<<dim3((Pin + 63) / 64, (Cin + 63) / 64, B), 256, 0, st>>>(input, xh, Cin, Pin, vec);
  int rc = check_launch();
  if (rc) return rc;
  dcn_pack_weight_kernel<<<kNumSMs, 256, 0, st>>>(weight, wk, Cout, Cin);
  rc = check_launch();
  if (rc) return rc;

  DcnParams prm;
  prm.xh = xh; prm.offset = offset; prm.mask = mask; prm.wk = wk; prm.bias = bias;
  prm.scale = scale; prm.shift = shift; prm.y = output;
  prm.Cin = Cin; prm.H = H; prm.W = W; prm.Ho = Ho; prm.Wo = Wo; prm.Cout = Cout;
  prm.stride = stride_h; prm.pad = pad_h; prm.dil = dilation_h; prm.relu = (flags & DCN_FLAG_RELU) ? 1 : 0;
  prm.tiles_x = (Wo + kTileW - 1) / kTileW;
  prm.tiles_per_image = prm.tiles_x * ((Ho + kTileH - 1) / kTileH);
  uint32_t cols = 32;
  while (cols < (uint32_t)Cout) cols <<= 1;                  // TMEM allocations are powers of two >= 32
  prm.tmem_cols = cols;
  const size_t smem = (size_t)kStages * (kTileM * kBK * 2 + (size_t)Cout * kBK * 2) + 1024;
  auto kern = (flags & DCN_FLAG_FP32_BLEND) ? dcn_forward_kernel<true> : dcn_forward_kernel<false>;
  cudaError_t e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  const long long tiles = (long long)B * prm.tiles_per_image;
  if (tiles > 0x7fffffffLL) return CTDET_EINVAL;
  kern<<<(unsigned)tiles, kDcnThreads, smem, st>>>(prm);
  return check_launch();
}
