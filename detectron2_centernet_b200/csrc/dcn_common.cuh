// Shared device helpers of the DCNv2 kernels (dcn.cu forward, dcn_bwd.cu backward): tcgen05 / TMEM /
// mbarrier / bulk-copy PTX wrappers, UMMA descriptors, packed blends.  sm_100a only.
#pragma once
#include <cuda_bf16.h>
#include <cstdlib>
#include "common.cuh"

namespace ctdet {

constexpr int kTileM = 128;           // pixels per tile = TMEM lanes = UMMA M
constexpr int kTileH = 8, kTileW = 16; // 2-D pixel tile
constexpr int kBK = 64;               // K block: 64 bf16 = one 128-byte swizzle row


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
// One thread of a CONVERGED warp (elect.sync): unlike `lane == 0`, the compiler knows the region runs on exactly one thread
// and emits the uniform-datapath instructions inside it (UTCHMMA, UTCBAR, UBLKCP) straight-line instead of wrapping each in
// an ELECT / BRA.U.ANY loop -- ~85 -> ? cycles per tcgen05.mma issue, which is what bounds a kernel whose K blocks are cheap.
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile("{\n\t.reg .pred p;\n\telect.sync _|p, 0xffffffff;\n\tselp.u32 %0, 1, 0, p;\n\t}" : "=r"(pred));
  return pred != 0;
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
// MN-major, 128-byte-swizzled operand: rows of 64 contiguous M/N elements (128 B) per K index, 8-row
// (K) groups 1024 B apart (SBO); LBO = distance between 64-element atoms along M/N.
__device__ __forceinline__ uint64_t make_sw128_desc_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
         ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

// kind::f16 instruction descriptor: D = fp32, A = B = bf16, both K-major, N>>3 at [17,23), M>>4 at [24,29)
__device__ __forceinline__ uint32_t make_idesc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

// waiters that are idle most of the time back off between polls: every poll is a shared-memory
// pipe transaction that competes with the producers' A-tile stores
__device__ __forceinline__ void dcn_mbar_wait_relaxed(uint64_t* bar, unsigned parity, unsigned ns = 64) {
  const uint32_t addr = smem_u32(bar);
  for (;;) {
    uint32_t ok;
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok) : "r"(addr), "r"(parity) : "memory");
    if (ok) break;
    if (ns) __nanosleep(ns);
  }
}
__device__ __forceinline__ void dcn_mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" :: "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void dcn_mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(smem_u32(bar)), "r"(bytes) : "memory");
}
// 1-D bulk async copy global -> shared (TMA engine, no LSU instructions), completion on an mbarrier
__device__ __forceinline__ void dcn_bulk_g2s(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

__device__ __forceinline__ uint32_t blend_bf16x2(uint32_t a, uint32_t b, uint32_t c, uint32_t d,
                                                 uint32_t w1, uint32_t w2, uint32_t w3, uint32_t w4) {
  __nv_bfloat162 acc = __hmul2(*reinterpret_cast<__nv_bfloat162*>(&w1), *reinterpret_cast<__nv_bfloat162*>(&a));
  acc = __hfma2(*reinterpret_cast<__nv_bfloat162*>(&w2), *reinterpret_cast<__nv_bfloat162*>(&b), acc);
  acc = __hfma2(*reinterpret_cast<__nv_bfloat162*>(&w3), *reinterpret_cast<__nv_bfloat162*>(&c), acc);
  acc = __hfma2(*reinterpret_cast<__nv_bfloat162*>(&w4), *reinterpret_cast<__nv_bfloat162*>(&d), acc);
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


// 16-byte gather of 8 bf16 channels through L1 (volatile: keeps every gather of a K block ahead of
// the blends, so 4 * ITERS loads are in flight per thread)
__device__ __forceinline__ uint4 ldg_gather16(const __nv_bfloat16* p) {
  uint4 v;
  asm volatile("ld.global.nc.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

__device__ __forceinline__ float ldg_stream1(const float* p) {      // read-once scalar, volatile = kept in order
  float v;
  asm volatile("ld.global.nc.L1::no_allocate.f32 %0, [%1];" : "=f"(v) : "l"(p));
  return v;
}

// ------------------------------------------------------------------------------------------------
struct DcnParams {
  const __nv_bfloat16* xh;     // [B][ximg / Cin pixels][Cin]: rows of xw pixels, image pixel (0,0) at element xorg
                               // (forward: xw = W, xorg = 0; backward: a zero border of one pixel, xw = W + 2)
  const float* offset;         // [B][18][Ho*Wo]   channel 2t = dh, 2t+1 = dw (deform_conv_cuda_kernel.cu:838-846)
  const float* mask;           // [B][9][Ho*Wo]
  const __nv_bfloat16* wk;     // [9*Cin/64][Cout][64] swizzled B tiles
  const float* bias;           // [Cout] or null
  const float* scale;          // [Cout] or null   y = (conv + bias) * scale + shift
  const float* shift;          // [Cout] or null
  float* y;                    // [B][Cout][Ho*Wo]
  int Cin, H, W, Ho, Wo, Cout, stride, pad, dil, relu, tiles_per_image, tiles_x, num_tiles, stages, mma_sleep, ld_sleep;
  int xw, ximg, xorg;
  int off_bstride, mask_bstride;   // elements between images in offset / mask (18*P / 9*P, or 27*P for the raw conv output)
  int mask_logit;                  // 1: mask holds logits (raw conv_offset_mask output): sigmoid applied here
  const float* om_bias;            // [27] bias of conv_offset_mask to add (offset channels 0..17, mask 18..26) or null
  uint32_t tmem_cols;
  int plain = 0;                   // 1: an ordinary 3x3 convolution (zero offsets, mask 1): conv_offset_mask itself on the
                                   //    tensor cores -- the geometry stage loads nothing, the producers gather one corner
  int cout_store = 0;              // output channels written (0 = Cout; 27 of the 32 computed for conv_offset_mask)
  long long y_bstride = 0;         // elements between images in y (0 = cout_store * Ho * Wo): a chunk of a wider output
};


// Sampling geometry of output pixel (ho, wo) for tap t (dmcn_im2col_bilinear, :666-699).
//   g  = element offset of the clamped top-left corner pixel in the NHWC image (pixel * Cin, a
//        multiple of 64) | bit0: the right corner is a different pixel | bit1: the bottom corner is
//   wq = the four bilinear weights * mask (0 for corners outside the image); either fp32 or, for
//        the packed blend, each rounded to bf16 and broadcast into both halves of a bf16x2
template <bool kFp32Blend>
__device__ __forceinline__ void tap_geometry(const DcnParams& prm, float off_h, float off_w, float m,
                                             int ho, int wo, bool valid, int t, uint32_t& g, uint4& wout) {
  g = 0u;
  float4 wq = make_float4(0.f, 0.f, 0.f, 0.f);
  if (valid) {
    const int H = prm.H, W = prm.W;
    const int i = t / 3, j = t - 3 * i;
    const float h_im = (float)(ho * prm.stride - prm.pad + i * prm.dil) + off_h;
    const float w_im = (float)(wo * prm.stride - prm.pad + j * prm.dil) + off_w;
    if (h_im > -1.f && w_im > -1.f && h_im < (float)H && w_im < (float)W) {       // :852
      const int h_low = (int)floorf(h_im), w_low = (int)floorf(w_im);
      const int h_high = h_low + 1, w_high = w_low + 1;
      const float lh = h_im - (float)h_low, lw = w_im - (float)w_low;
      const float hh = 1.f - lh, hw = 1.f - lw;
      const bool t_ok = h_low >= 0, b_ok = h_high <= H - 1, l_ok = w_low >= 0, r_ok = w_high <= W - 1;
      const int hl = max(h_low, 0), hh_i = min(h_high, H - 1), wl = max(w_low, 0), wh_i = min(w_high, W - 1);
      g = (uint32_t)(hl * prm.xw + wl) * (uint32_t)prm.Cin | (wh_i != wl ? 1u : 0u) | (hh_i != hl ? 2u : 0u);
      wq.x = (t_ok && l_ok) ? hh * hw * m : 0.f;
      wq.y = (t_ok && r_ok) ? hh * lw * m : 0.f;
      wq.z = (b_ok && l_ok) ? lh * hw * m : 0.f;
      wq.w = (b_ok && r_ok) ? lh * lw * m : 0.f;
    }
  }
  if (kFp32Blend) {
    wout = make_uint4(__float_as_uint(wq.x), __float_as_uint(wq.y), __float_as_uint(wq.z), __float_as_uint(wq.w));
  } else {
    const __nv_bfloat162 q1 = __float2bfloat162_rn(wq.x), q2 = __float2bfloat162_rn(wq.y),
                         q3 = __float2bfloat162_rn(wq.z), q4 = __float2bfloat162_rn(wq.w);
    wout = make_uint4(*reinterpret_cast<const uint32_t*>(&q1), *reinterpret_cast<const uint32_t*>(&q2),
                      *reinterpret_cast<const uint32_t*>(&q3), *reinterpret_cast<const uint32_t*>(&q4));
  }
}

constexpr int kMaxStages = 8;
constexpr int kGeoBytes = 2 * 9 * kTileM * (16 + 4);      // double-buffered {uint4 weights, u32 g} per (tap, pixel)

struct DcnBarriers {
  uint64_t full[kMaxStages];     // A rows written by the producer warps + B tile landed  -> MMA may read stage
  uint64_t empty[kMaxStages];    // the MMAs that read the stage retired                  -> stage may be refilled
  uint64_t geo_full[2], geo_empty[2];   // per-tile sampling geometry (all 9 taps) in shared memory
  uint64_t acc_full[2], acc_empty[2];   // TMEM accumulator buffers
};



// ------------------------------------------------------------------------------------------------
// Producer warps (shared by the forward and the weight-gradient kernel): for tiles
// [tile_first, tile_first + my_tiles) and K blocks [kb0, kb1) (kb = cb * 9 + t) build the
// bilinear-sampled, mask-modulated im2col tile [128 pixels][64 channels] bf16 in the 128-byte
// swizzled layout, one pipeline stage per K block.  8 lanes x 16 B gather one pixel's 64 channels
// (one warp-wide load = 4 full 128-byte lines); the gathers of K block n+1 are issued while block n
// is blended (register double-buffering across K blocks AND tiles).
template <bool kFp32Blend, int NP, bool kPadOdd = false, bool kPlain = false>
__device__ __forceinline__ void dcn_producer_loop(const DcnParams& prm, uint8_t* smem, uint32_t stage_bytes, int NS,
                                                  const uint4* geo_w, const uint32_t* geo_g, DcnBarriers& bars,
                                                  int warp, int lane, int tile_first, int my_tiles, int kb0, int kb1) {
  constexpr int ITERS = 32 / NP;               // 4-row groups per warp per K block
  const int Cin = prm.Cin;
  const int sub = lane >> 3, chunk = lane & 7;
  const uint32_t swz = (uint32_t)((chunk ^ (((warp & 1) << 2) + sub)) << 4);      // row & 7 is fixed per thread
  const uint32_t row0 = (uint32_t)(warp << 2) + sub;
  const uint32_t ex = (uint32_t)Cin, ey = (uint32_t)prm.xw * (uint32_t)Cin;
  auto tile_x = [&](int it) {
    const int tile = tile_first + it;
    return prm.xh + (size_t)(tile / prm.tiles_per_image) * prm.ximg + prm.xorg + chunk * 8;
  };
  uint4 va[ITERS], vb[ITERS], vc[ITERS], vd[ITERS];
  auto gather = [&](int i, const __nv_bfloat16* xc, uint32_t g) {
    const uint32_t e0 = g & ~3u;
    const uint32_t dx = (g & 1u) ? ex : 0u;
    const uint32_t e2 = e0 + ((g & 2u) ? ey : 0u);
    va[i] = ldg_gather16(xc + e0);
    if (!kPlain) {                               // (plain convolution: integer positions, the top-left corner is the sample)
      vb[i] = ldg_gather16(xc + e0 + dx);
      vc[i] = ldg_gather16(xc + e2);
      vd[i] = ldg_gather16(xc + e2 + dx);
    }
  };
  const int total = my_tiles * (kb1 - kb0);
  const int cb0 = kb0 / 9, t0 = kb0 - 9 * cb0;
  int l_it = 0, l_kb = kb0, l_cb = cb0, l_t = t0;   // load cursor: the K block whose gathers are in flight
  const __nv_bfloat16* l_x = tile_x(0);
  if (total > 0) {
    dcn_mbar_wait(&bars.geo_full[0], 0u);
#pragma unroll
    for (int i = 0; i < ITERS; ++i) gather(i, l_x + l_cb * kBK, geo_g[l_t * kTileM + row0 + i * NP * 4]);
  }
  int s = 0; unsigned ph = 0;
  for (int n = 0; n < total; ++n) {
    const int c_it = l_it, c_t = l_t;          // consume cursor = the block that is in flight now
    const bool have_next = n + 1 < total;
    if (++l_kb == kb1) {
      l_kb = kb0; l_cb = cb0; l_t = t0; ++l_it;
      if (have_next) {
        l_x = tile_x(l_it);
        dcn_mbar_wait(&bars.geo_full[l_it & 1], (unsigned)((l_it >> 1) & 1));
      }
    } else if (++l_t == 9) {
      l_t = 0; ++l_cb;
    }
    const __nv_bfloat16* l_xc = l_x + l_cb * kBK;
    const uint32_t* gg = geo_g + ((l_it & 1) * 9 + l_t) * kTileM + row0;
    const uint4* gw = geo_w + ((c_it & 1) * 9 + c_t) * kTileM + row0;
    dcn_mbar_wait(&bars.empty[s], ph ^ 1u);
    uint8_t* a_tile = smem + (size_t)s * stage_bytes;
#pragma unroll
    for (int i = 0; i < ITERS; ++i) {
      const uint4 wq = gw[i * NP * 4];
      uint4 q;
      if (kPlain) {                              // weight of the one corner: 1, or 0 outside the image
        q = wq.x != 0u ? va[i] : make_uint4(0u, 0u, 0u, 0u);
      } else if (kFp32Blend) {
        const float w1 = __uint_as_float(wq.x), w2 = __uint_as_float(wq.y),
                    w3 = __uint_as_float(wq.z), w4 = __uint_as_float(wq.w);
        q.x = blend_fp32(va[i].x, vb[i].x, vc[i].x, vd[i].x, w1, w2, w3, w4);
        q.y = blend_fp32(va[i].y, vb[i].y, vc[i].y, vd[i].y, w1, w2, w3, w4);
        q.z = blend_fp32(va[i].z, vb[i].z, vc[i].z, vd[i].z, w1, w2, w3, w4);
        q.w = blend_fp32(va[i].w, vb[i].w, vc[i].w, vd[i].w, w1, w2, w3, w4);
      } else {
        q.x = blend_bf16x2(va[i].x, vb[i].x, vc[i].x, vd[i].x, wq.x, wq.y, wq.z, wq.w);
        q.y = blend_bf16x2(va[i].y, vb[i].y, vc[i].y, vd[i].y, wq.x, wq.y, wq.z, wq.w);
        q.z = blend_bf16x2(va[i].z, vb[i].z, vc[i].z, vd[i].z, wq.x, wq.y, wq.z, wq.w);
        q.w = blend_bf16x2(va[i].w, vb[i].w, vc[i].w, vd[i].w, wq.x, wq.y, wq.z, wq.w);
      }
      *reinterpret_cast<uint4*>(a_tile + (row0 + (uint32_t)(i * NP * 4)) * 128u + swz) = q;
    }
    // The proxy fence compiles to MEMBAR.ALL.CTA + FENCE.VIEW.ASYNC: it waits for every memory operation this warp
    // has in flight.  It therefore comes BEFORE the gathers of the next K block are issued (it used to follow them
    // and stalled every warp for a full gather round trip per K block, which defeated the register double-buffering).
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> tensor-core reads
    __syncwarp();
    if (lane == 0) {
      dcn_mbar_arrive(&bars.full[s]);
      if (l_it != c_it) dcn_mbar_arrive(&bars.geo_empty[c_it & 1]);   // tile c_it fully consumed by this warp
    }
    if (have_next) {
#pragma unroll
      for (int i = 0; i < ITERS; ++i) gather(i, l_xc, gg[i * NP * 4]);   // refill: K block n+1, same row groups
    }
    if (++s == NS) { s = 0; ph ^= 1u; }
    if (kPadOdd && l_it != c_it && ((kb1 - kb0) & 1)) {
      // wgrad pairs two K blocks per MMA (two adjacent stages): a tile with an odd number of K blocks hands over
      // one empty stage so that every tile starts on an even stage
      dcn_mbar_wait(&bars.empty[s], ph ^ 1u);
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.full[s]);
      if (++s == NS) { s = 0; ph ^= 1u; }
    }
  }
}

// Producer warps of the PLAIN 3x3 convolution (conv_offset_mask and the heads' first layer on the tensor cores; stride 1,
// padding 1, dilation 1): the A tile of a K block is a copy -- one 128-byte row of 64 channels per pixel from the integer
// tap position, zeros outside the image.  The first form went through the deformable machinery (geometry table in shared
// memory, 16-byte gathers into registers, swizzled stores): 950 cycles per K block whatever the warp layout, because the
// SM's L1 / shared-memory data pipe was 62 % busy with it (ncu: per K block 174 wavefronts of table reads incl. 96 bank
// conflicts, 127 of global loads, 168 of stores, + the tensor core's own operand reads).  This form has no table -- the
// positions are integers, computed from the row index -- and no registers in the path: `cp.async.ca` moves 16 bytes from
// global memory (through L1: the nine taps of a tile read the same pixels) straight to their swizzled place in the stage
// (src-size 0 = zero fill outside the image), one pass over
// the data pipe instead of two.  The NP warps form NS groups; group s owns pipeline stage s and builds blocks s, s + NS, ...
// (a stage has ONE group of writers: parity waits on its `empty` barrier only tell consecutive phases apart).
template <int NP>
__device__ __forceinline__ void dcn_plain_producer_loop(const DcnParams& prm, uint8_t* smem, uint32_t stage_bytes, int NS,
                                                        DcnBarriers& bars, int warp, int lane, int tile_first, int my_tiles,
                                                        int KB) {
  const int Cin = prm.Cin, H = prm.H, W = prm.W;
  const int sub = lane >> 3, chunk = lane & 7;
  const int total = my_tiles * KB;
  const int wps = NP / NS;                        // warps per stage (NP % NS == 0)
  const int s = warp / wps, part = warp - s * wps;
  const int rows = kTileM / wps;                  // rows of a block built by this warp (a multiple of 4)
  for (int n = s; n < total; n += NS) {
    const int it = n / KB, kb = n - it * KB;
    const int cb = kb / 9, t = kb - 9 * cb;
    const int dy = t / 3 - 1, dx = t - 3 * (t / 3) - 1;
    int tile = tile_first + it;
    const int b = tile / prm.tiles_per_image;
    tile -= b * prm.tiles_per_image;
    const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
    const __nv_bfloat16* xb = prm.xh + (size_t)b * prm.ximg + prm.xorg + chunk * 8 + cb * kBK;
    const uint32_t a_addr = smem_u32(smem + (size_t)s * stage_bytes);
    dcn_mbar_wait(&bars.empty[s], (unsigned)(((n / NS) & 1) ^ 1));
#pragma unroll 4
    for (int r = part * rows + sub; r < (part + 1) * rows; r += 4) {
      const int hi = tile_y * kTileH + (r >> 4) + dy, wi = tile_x * kTileW + (r & 15) + dx;
      // (a pixel of a ragged tile beyond the map: its output is never stored, whatever lands here)
      const bool inside = hi >= 0 && hi < H && wi >= 0 && wi < W;
      const __nv_bfloat16* src = xb + (inside ? (size_t)(hi * prm.xw + wi) * Cin : 0);
      const uint32_t dst = a_addr + (uint32_t)r * 128u + ((uint32_t)(chunk ^ (r & 7)) << 4);
      asm volatile("cp.async.ca.shared.global [%0], [%1], 16, %2;" :: "r"(dst), "l"(src), "r"(inside ? 16 : 0) : "memory");
    }
    asm volatile("cp.async.commit_group;" ::: "memory");
    asm volatile("cp.async.wait_group 0;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // generic-proxy writes -> tensor-core reads
    __syncwarp();
    if (lane == 0) dcn_mbar_arrive(&bars.full[s]);
  }
}

// Geometry of one tile (thread r = pixel of the tile, all 9 taps) into geometry buffer it & 1.
// Waits for the buffer to be free; the caller arrives on geo_full afterwards.
template <bool kFp32Blend>
__device__ __forceinline__ void dcn_geometry_tile(const DcnParams& prm, int tile, int it, int r,
                                                  uint4* geo_w, uint32_t* geo_g, DcnBarriers& bars) {
  const int P = prm.Ho * prm.Wo;
  const int b = tile / prm.tiles_per_image;
  tile -= b * prm.tiles_per_image;
  const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
  const int ho = tile_y * kTileH + (r >> 4), wo = tile_x * kTileW + (r & 15);
  const bool valid = ho < prm.Ho && wo < prm.Wo;
  const int p = valid ? ho * prm.Wo + wo : 0;
  const float* offb = prm.offset + (size_t)b * prm.off_bstride + p;
  const float* mskb = prm.mask + (size_t)b * prm.mask_bstride + p;
  const int buf = it & 1;
  // all 27 loads first (one DRAM round trip, issued before the buffer is even free)
  float ov[18], mv[9];
  if (prm.plain) {
#pragma unroll
    for (int k = 0; k < 18; ++k) ov[k] = 0.f;
#pragma unroll
    for (int k = 0; k < 9; ++k) mv[k] = 1.f;
  } else {
#pragma unroll
    for (int k = 0; k < 18; ++k) ov[k] = ldg_stream1(offb + (size_t)k * P);
#pragma unroll
    for (int k = 0; k < 9; ++k) mv[k] = ldg_stream1(mskb + (size_t)k * P);
  }
  if (prm.om_bias != nullptr) {          // conv_offset_mask's bias, not added by the caller
#pragma unroll
    for (int k = 0; k < 18; ++k) ov[k] += __ldg(prm.om_bias + k);
#pragma unroll
    for (int k = 0; k < 9; ++k) mv[k] += __ldg(prm.om_bias + 18 + k);
  }
  if (prm.mask_logit) {                  // mask = sigmoid(third chunk), resnet.py:311
#pragma unroll
    for (int k = 0; k < 9; ++k) mv[k] = 1.f / (1.f + expf(-mv[k]));
  }
  dcn_mbar_wait_relaxed(&bars.geo_empty[buf], (unsigned)(((it >> 1) & 1) ^ 1));
#pragma unroll
  for (int t = 0; t < 9; ++t) {
    uint32_t g; uint4 wq;
    tap_geometry<kFp32Blend>(prm, ov[2 * t], ov[2 * t + 1], mv[t], ho, wo, valid, t, g, wq);
    geo_w[(buf * 9 + t) * kTileM + r] = wq;
    geo_g[(buf * 9 + t) * kTileM + r] = g;
  }
}

// pre-passes defined in dcn.cu
__global__ void dcn_pack_input_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ xh, int C, int P, int vec,
                                      float* __restrict__ chan_sum, int chunks, int border_w);

// Tuning knobs (experiments only; the defaults are the shipped configuration).  Read from the environment ONCE per
// process -- function-local static, thread-safe initialisation -- never on the launch path.
struct DcnTunables {
  int stages;            // CTDET_DCN_STAGES          A/B ring depth cap of the forward kernel (4)
  int mma_sleep;         // CTDET_DCN_MMA_SLEEP       ns back-off of the MMA warp's barrier polls (0)
  int ld_sleep;          // CTDET_DCN_LD_SLEEP        ns back-off of the weight loader's polls (20)
  int producer_warps;    // CTDET_DCN_PRODUCER_WARPS  gather warps of the forward kernel, 16 or 8 (16)
  int dgrad_warps;       // CTDET_DCN_DGRAD_WARPS     consumer warps of the dgrad kernel, 16 or 8 (16)
  int wgrad_stages;      // CTDET_DCN_WGRAD_STAGES    im2col ring depth cap of the wgrad kernel, 2..8 (4)
  int wgrad_paired;      // CTDET_DCN_WGRAD_PAIRED    wgrad at Cout = 64: two K blocks per MMA (1)
  int dgrad_gemm;        // CTDET_DCN_DGRAD_GEMM      1: grad_input scatter as a tensor-core GEMM where supported (1)
  int dgrad_gb;          // CTDET_DCN_DGRAD_GB        grad_output tile buffers of that kernel at Cout = 64: 1 or 2 (1)
};
static inline const DcnTunables& dcn_tunables() {
  static const DcnTunables t = [] {
    auto env = [](const char* name, int dflt) { const char* v = getenv(name); return (v && *v) ? atoi(v) : dflt; };
    DcnTunables r;
    r.stages = env("CTDET_DCN_STAGES", 4);
    r.mma_sleep = env("CTDET_DCN_MMA_SLEEP", 0);
    r.ld_sleep = env("CTDET_DCN_LD_SLEEP", 20);
    r.producer_warps = env("CTDET_DCN_PRODUCER_WARPS", 16) == 8 ? 8 : 16;
    r.dgrad_warps = env("CTDET_DCN_DGRAD_WARPS", 16) == 8 ? 8 : 16;
    r.wgrad_stages = env("CTDET_DCN_WGRAD_STAGES", 4); if (r.wgrad_stages < 2) r.wgrad_stages = 2; if (r.wgrad_stages > kMaxStages) r.wgrad_stages = kMaxStages;
    r.wgrad_paired = env("CTDET_DCN_WGRAD_PAIRED", 1) != 0;
    r.dgrad_gemm = env("CTDET_DCN_DGRAD_GEMM", 1) != 0;
    r.dgrad_gb = env("CTDET_DCN_DGRAD_GB", 1) == 2 ? 2 : 1;
    return r;
  }();
  return t;
}

}  // namespace ctdet
