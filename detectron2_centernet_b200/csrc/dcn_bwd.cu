// DCNv2 (modulated deformable 3x3 convolution) BACKWARD on the 5th-generation tensor cores.
//
// Replaces modulated_deform_conv_cuda_backward (detectron2/layers/csrc/deformable/deform_conv_cuda.cu:
// 929-1129), which per IMAGE runs an fp32 SGEMM into columns[9*Cin, H*W] in HBM (:1003-1008), the
// col2im_coord kernel (deform_conv_cuda_kernel.cu:952-1060: grad_offset / grad_mask), the col2im
// kernel (:871-949: grad_input, one scalar atomicAdd per corner), a second im2col into the same HBM
// buffer (:1055) and two more SGEMMs for grad_weight / grad_bias (:1085-1100).  Here nothing of
// size 9*Cin*H*W ever reaches HBM:
//
//   pre-passes : x and grad_output NCHW fp32 -> NHWC bf16 (grad_bias = per-channel sums of grad_output
//                falls out of the same pass); weights -> bf16 transposed tiles
//   wgrad      : dW[co, (t,ci)] = sum_p gout[p, co] * col[p, (t,ci)].  The SAME producer as the
//                forward builds the sampled im2col tile in shared memory; it is consumed as an
//                MN-major UMMA operand (K = the 128 pixels of the tile), the grad_output tile as the
//                other MN-major operand.  Accumulators for a whole group of K blocks stay in TMEM
//                across every tile a CTA owns and are flushed once with fp32 atomics.
//   dgrad      : G[p, (t,ci)] = sum_co gout[p, co] * W[co, ci, t] per K block on the tensor cores
//                (TMEM), drained to shared memory as bf16; consumer warps gather the four corners of
//                every (pixel, tap) exactly like the forward (8 lanes x 16 B per pixel) and produce
//                  grad_mask   = sum_ci G * bilinear(x)                       (:1019-1030)
//                  grad_offset = sum_ci G * mask * d bilinear / d(h,w)        (:1031-1040)
//                  grad_input += G * mask * corner weight                     (:944, vector fp32 RED)
#include "dcn_common.cuh"

namespace ctdet {

struct DcnBwdParams {
  const __nv_bfloat16* gh;     // grad_output NHWC bf16 [B][Ho*Wo][Cout]
  const __nv_bfloat16* wt;     // dgrad B tiles: per K block kb, Cout/64 slabs of [64 ci][64 co] (swizzled)
  float* gw;                   // grad_weight [Cout][Cin][3][3] fp32 (zeroed before the launch)
  float* gx;                   // grad_input NHWC fp32 [B][(H+2)*(W+2)][Cin], bordered like xh (zeroed before the launch)
  float* goff;                 // grad_offset [B][18][Ho*Wo]
  float* gmask;                // grad_mask   [B][9][Ho*Wo]
  int n_groups, kb_per_group;  // wgrad: K blocks are split into groups that fit TMEM
  int paired;                  // wgrad, Cout = 64: D[(2 K blocks x 64 ci), co] -- two K blocks per MMA, rows fully used
  int g_stages, w_stages, gout_bufs;   // dgrad ring depths
  int s_bufs;                          // dgrad (GEMM scatter): S chunk ring depth
};

// grad_output tile [128 pixels][Cout] bf16 -> shared memory, Cout/64 slabs of [128][64] in the
// 128-byte swizzled layout (read K-major by dgrad, MN-major by wgrad).  thread r = pixel of the tile.
__device__ __forceinline__ void dcn_load_gout_tile(const DcnParams& prm, const __nv_bfloat16* gh, int tile, int r,
                                                   uint8_t* dst) {
  const int P = prm.Ho * prm.Wo, Cout = prm.Cout;
  const int b = tile / prm.tiles_per_image;
  tile -= b * prm.tiles_per_image;
  const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
  const int ho = tile_y * kTileH + (r >> 4), wo = tile_x * kTileW + (r & 15);
  const bool valid = ho < prm.Ho && wo < prm.Wo;
  const uint4* src = reinterpret_cast<const uint4*>(gh + ((size_t)b * P + (valid ? ho * prm.Wo + wo : 0)) * Cout);
  uint8_t* row = dst + r * 128;
  const int sw = r & 7;
  for (int j0 = 0; j0 < Cout / 8; j0 += 8) {            // one 64-channel slab per iteration
    uint4 v[8];
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = valid ? __ldg(src + j0 + j) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
    for (int j = 0; j < 8; ++j)
      *reinterpret_cast<uint4*>(row + (j0 >> 3) * (kTileM * 128) + ((j ^ sw) << 4)) = v[j];
  }
}

// ================================================================================================
// wgrad
// ================================================================================================
template <int NP>
__global__ void __launch_bounds__((NP + 10) * 32, 1)
dcn_wgrad_kernel(const DcnParams prm, const DcnBwdParams bw) {
  extern __shared__ __align__(1024) uint8_t dcn_smem[];
  __shared__ DcnBarriers bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Cin = prm.Cin, Cout = prm.Cout, NS = prm.stages;
  const uint32_t stage_bytes = (uint32_t)kTileM * kBK * 2u;                // im2col tile only
  uint8_t* smem = dcn_smem + ((1024u - (smem_u32(dcn_smem) & 1023u)) & 1023u);
  uint4* geo_w = reinterpret_cast<uint4*>(smem + (size_t)NS * stage_bytes);          // [2][9][128]
  uint32_t* geo_g = reinterpret_cast<uint32_t*>(geo_w + 2 * 9 * kTileM);             // [2][9][128]
  uint8_t* gout = smem + (size_t)NS * stage_bytes + kGeoBytes;                       // [2][Cout/64][128][128 B]
  const uint32_t gout_bytes = (uint32_t)(Cout / 64) * kTileM * 128u;
  const int KB = 9 * (Cin / kBK);
  const int grp = (int)blockIdx.x % bw.n_groups, chunk = (int)blockIdx.x / bw.n_groups;
  const int chunks = (int)gridDim.x / bw.n_groups;
  const int tile_first = (int)(((long long)chunk * prm.num_tiles) / chunks);
  const int my_tiles = (int)(((long long)(chunk + 1) * prm.num_tiles) / chunks) - tile_first;
  const int kb0 = grp * bw.kb_per_group;
  const int kb1 = min(KB, kb0 + bw.kb_per_group);
  const int nkb = kb1 - kb0;
  const int MT = Cout > 128 ? 2 : 1;             // 128-row M tiles over the output channels

  if (tid == 0) {
    for (int s = 0; s < NS; ++s) { dcn_mbar_init(&bars.full[s], NP); dcn_mbar_init(&bars.empty[s], 1); }
    for (int i = 0; i < 2; ++i) {
      dcn_mbar_init(&bars.geo_full[i], 4);  dcn_mbar_init(&bars.geo_empty[i], NP + 1);
      dcn_mbar_init(&bars.acc_full[i], 1);  dcn_mbar_init(&bars.acc_empty[i], 4);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == NP + 8) tmem_alloc(&s_tmem, prm.tmem_cols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;
  const bool work = my_tiles > 0 && nkb > 0;

  if (warp < NP) {
    if (nkb > 0) {
      if (bw.paired) dcn_producer_loop<false, NP, true>(prm, smem, stage_bytes, NS, geo_w, geo_g, bars, warp, lane, tile_first, my_tiles, kb0, kb1);
      else dcn_producer_loop<false, NP>(prm, smem, stage_bytes, NS, geo_w, geo_g, bars, warp, lane, tile_first, my_tiles, kb0, kb1);
    }
  } else if (warp < NP + 4) {
    // ============ final flush: TMEM -> fp32 atomics into grad_weight[co][ci][t] ====================
    if (work) {
      const int quarter = warp - NP;
      const int r = (quarter << 5) + lane;
      dcn_mbar_wait_relaxed(&bars.acc_full[0], 0u);
      tc_fence_after();
      if (bw.paired) {
        // rows = (K block of the pair, ci), columns = (pair, co)
        const int h = r >> 6, ci = r & 63;
        for (int j = 0; 2 * j < nkb; ++j) {
          const int kbl = 2 * j + h;
          const bool ok = kbl < nkb;                 // the odd last K block has no partner: rows 64..127 hold garbage
          const int kb = kb0 + (ok ? kbl : 0), cb = kb / 9, t = kb - 9 * cb;
          float* dst = bw.gw + ((size_t)cb * kBK + ci) * 9 + t;
          for (int c0 = 0; c0 < 64; c0 += 16) {
            uint32_t v[16];
            tmem_ld16(tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(j * 64 + c0), v);
            if (ok) {
#pragma unroll
              for (int k = 0; k < 16; ++k) atomicAdd(dst + (size_t)(c0 + k) * Cin * 9, __uint_as_float(v[k]));
            }
          }
        }
      } else
      for (int mt = 0; mt < MT; ++mt) {
        const int co = mt * 128 + r;
        const bool ok = co < Cout;                 // Cout = 64: rows 64..127 duplicate rows 0..63
        for (int kbl = 0; kbl < nkb; ++kbl) {
          const int kb = kb0 + kbl, cb = kb / 9, t = kb - 9 * cb;
          float* dst = bw.gw + ((size_t)(ok ? co : 0) * Cin + cb * kBK) * 9 + t;
          for (int c0 = 0; c0 < kBK; c0 += 16) {
            uint32_t v[16];
            tmem_ld16(tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)((mt * nkb + kbl) * kBK + c0), v);
            if (ok) {
#pragma unroll
              for (int k = 0; k < 16; ++k) atomicAdd(dst + (size_t)(c0 + k) * 9, __uint_as_float(v[k]));
            }
          }
        }
      }
      tc_fence_before();
    }
  } else if (warp < NP + 8) {
    // ============ geometry + grad_output tile of the tiles ahead ===================================
    const int r = ((warp - NP - 4) << 5) + lane;
    if (nkb > 0) {
      for (int it = 0; it < my_tiles; ++it) {
        dcn_geometry_tile<false>(prm, tile_first + it, it, r, geo_w, geo_g, bars);   // waits for buffer it & 1
        dcn_load_gout_tile(prm, bw.gh, tile_first + it, r, gout + (size_t)(it & 1) * gout_bytes);
        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
        __syncwarp();
        if (lane == 0) dcn_mbar_arrive(&bars.geo_full[it & 1]);
      }
    }
  } else if (warp == NP + 8) {
    // ============ MMA issuer: D[co, ci] += gout^T[co, p] * col[p, ci] ===============================
    if (lane == 0 && work) {
      const uint32_t idesc = make_idesc_bf16(kTileM, kBK) | (1u << 15) | (1u << 16);   // A and B MN-major
      const uint32_t lbo = Cout >= 128 ? (uint32_t)kTileM * 128u : 0u;   // second 64-channel atom of the M tile
      int s = 0; unsigned ph = 0;
      if (bw.paired) {
        // D[(h, ci), co] += col_{2j+h}^T[ci, p] * gout[p, co]: A = two adjacent stages (M atoms one stage apart),
        // B = the grad_output tile, both MN-major; a lone last K block is paired with the (empty) pad stage
        for (int it = 0; it < my_tiles; ++it) {
          const int buf = it & 1;
          dcn_mbar_wait_relaxed(&bars.geo_full[buf], (unsigned)((it >> 1) & 1), 0u);
          tc_fence_after();
          const uint32_t g_addr = smem_u32(gout + (size_t)buf * gout_bytes);
          for (int j = 0; 2 * j < nkb; ++j) {
            dcn_mbar_wait_relaxed(&bars.full[s], ph, 0u);
            dcn_mbar_wait_relaxed(&bars.full[s + 1], ph, 0u);
            tc_fence_after();
            const uint32_t c_addr = smem_u32(smem + (size_t)s * stage_bytes);
            const uint32_t d_tmem = tmem + (uint32_t)(j * 64);
#pragma unroll
            for (int k8 = 0; k8 < kTileM / 16; ++k8) {
              const uint64_t adesc = make_sw128_desc_mn(c_addr + k8 * 2048u, stage_bytes);
              const uint64_t bdesc = make_sw128_desc_mn(g_addr + k8 * 2048u, 1024u);
              umma_bf16(d_tmem, adesc, bdesc, idesc, (uint32_t)((it | k8) != 0));
            }
            umma_commit(&bars.empty[s]);
            umma_commit(&bars.empty[s + 1]);
            s += 2;
            if (s == NS) { s = 0; ph ^= 1u; }
          }
          umma_commit(&bars.geo_empty[buf]);
        }
      } else
      for (int it = 0; it < my_tiles; ++it) {
        const int buf = it & 1;
        dcn_mbar_wait_relaxed(&bars.geo_full[buf], (unsigned)((it >> 1) & 1), 0u);
        tc_fence_after();
        const uint32_t g_addr = smem_u32(gout + (size_t)buf * gout_bytes);
        for (int kbl = 0; kbl < nkb; ++kbl) {
          dcn_mbar_wait_relaxed(&bars.full[s], ph, 0u);
          tc_fence_after();
          const uint32_t c_addr = smem_u32(smem + (size_t)s * stage_bytes);
          for (int mt = 0; mt < MT; ++mt) {
            const uint32_t d_tmem = tmem + (uint32_t)((mt * nkb + kbl) * kBK);
#pragma unroll
            for (int k8 = 0; k8 < kTileM / 16; ++k8) {     // K = 16 pixels = 2 swizzle atoms = 2048 B
              const uint64_t adesc = make_sw128_desc_mn(g_addr + (uint32_t)(2 * mt) * kTileM * 128u + k8 * 2048u, lbo);
              const uint64_t bdesc = make_sw128_desc_mn(c_addr + k8 * 2048u, 1024u);
              umma_bf16(d_tmem, adesc, bdesc, idesc, (uint32_t)((it | k8) != 0));
            }
          }
          umma_commit(&bars.empty[s]);
          if (++s == NS) { s = 0; ph ^= 1u; }
        }
        umma_commit(&bars.geo_empty[buf]);           // grad_output tile + geometry buffer may be refilled
      }
      umma_commit(&bars.acc_full[0]);
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == NP + 8) tmem_dealloc(tmem, prm.tmem_cols);
}


// ================================================================================================
// dgrad: grad_input, grad_offset, grad_mask
// ================================================================================================
// weight [Cout][Cin][9] fp32 -> per K block kb = cb*9 + t, Cout/64 slabs of [64 ci][64 co] bf16
// (K-major over co) in the 128-byte swizzled shared-memory image, moved by 1-D bulk copies.
__global__ void dcn_pack_weight_t_kernel(const float* __restrict__ w, __nv_bfloat16* __restrict__ wt,
                                         int Cout, int Cin) {
  const size_t n_el = (size_t)Cout * Cin * 9;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n_el; i += (size_t)gridDim.x * blockDim.x) {
    const int k = (int)(i & 63);                          // co inside the slab
    const int n = (int)((i >> 6) & 63);                   // ci inside the block
    const int j = (int)((i >> 12) % (Cout / 64));         // slab = 64-channel block of co
    const int kb = (int)(i / ((size_t)Cout * 64));
    const int cb = kb / 9, t = kb - 9 * cb;
    wt[(size_t)kb * Cout * 64 + (size_t)j * 4096 + n * 64 + (((k >> 3) ^ (n & 7)) << 3) + (k & 7)] =
        __float2bfloat16_rn(w[((size_t)(64 * j + k) * Cin + cb * 64 + n) * 9 + t]);
  }
}

// grad_input bordered NHWC fp32 (pixel (y, x) at (y+1)*(W+2) + x+1) -> NCHW fp32, 64 x 64 blocks through shared memory
__global__ void __launch_bounds__(256)
dcn_unpack_grad_input_kernel(const float* __restrict__ gx, float* __restrict__ out, int C, int P, int W) {
  __shared__ float t[64][65];                             // [pixel][channel]
  const int b = blockIdx.z, c0 = blockIdx.y * 64, p0 = blockIdx.x * 64;
  const float* src = gx + (size_t)b * (size_t)(P / W + 2) * (W + 2) * C;   // bordered rows of W + 2 pixels
  float* dst = out + (size_t)b * C * P;
  const int tid = threadIdx.x;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int idx = tid + k * 256;                        // 64 pixels x 16 float4 of channels
    const int pp = idx >> 4, q = idx & 15;
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (p0 + pp < P) {
      const int p = p0 + pp, y = p / W;
      v = *reinterpret_cast<const float4*>(src + (size_t)(p + 2 * y + W + 3) * C + c0 + q * 4);
    }
    t[pp][q * 4 + 0] = v.x; t[pp][q * 4 + 1] = v.y; t[pp][q * 4 + 2] = v.z; t[pp][q * 4 + 3] = v.w;
  }
  __syncthreads();
#pragma unroll
  for (int k = 0; k < 16; ++k) {
    const int idx = tid + k * 256;                        // 64 channels x 64 pixels
    const int c = idx >> 6, pp = idx & 63;
    if (p0 + pp < P) dst[(size_t)(c0 + c) * P + p0 + pp] = t[pp][c];
  }
}

// Sampling geometry for dgrad: {e, lh, lw, mask}.  The backward's copies of x and of grad_input carry a
// one-pixel zero border (rows of W + 2 pixels), so the four corners of any tap that the reference samples
// (h_im in (-1, H), w_im in (-1, W), deform_conv_cuda_kernel.cu:852) are always addressable and the corners
// outside the image read zeros / collect gradients nobody reads: no clamping, no validity flags.
//   e = element offset of the top-left corner inside the bordered image ((h_low+1)*(W+2) + w_low+1) * Cin;
// the other corners are at +Cin, +(W+2)*Cin, +(W+3)*Cin.  A tap that falls outside altogether has all four
// fields zero (mask 0 -> no gradient, e = 0 -> border pixels).
__device__ __forceinline__ uint4 tap_geometry_raw(const DcnParams& prm, float off_h, float off_w, float m,
                                                  int ho, int wo, bool valid, int t) {
  uint4 o = make_uint4(0u, 0u, 0u, 0u);
  if (valid) {
    const int i = t / 3, j = t - 3 * i;
    const float h_im = (float)(ho * prm.stride - prm.pad + i * prm.dil) + off_h;
    const float w_im = (float)(wo * prm.stride - prm.pad + j * prm.dil) + off_w;
    if (h_im > -1.f && w_im > -1.f && h_im < (float)prm.H && w_im < (float)prm.W) {
      const int h_low = (int)floorf(h_im), w_low = (int)floorf(w_im);
      o.x = (uint32_t)((h_low + 1) * prm.xw + w_low + 1) * (uint32_t)prm.Cin;
      o.y = __float_as_uint(h_im - (float)h_low);
      o.z = __float_as_uint(w_im - (float)w_low);
      o.w = __float_as_uint(m);
    }
  }
  return o;
}

__device__ __forceinline__ void red_add_v4(float* p, float a, float b, float c, float d) {
  asm volatile("red.global.add.v4.f32 [%0], {%1,%2,%3,%4};" :: "l"(p), "f"(a), "f"(b), "f"(c), "f"(d) : "memory");
}
__device__ __forceinline__ float bf_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
// packed fp32 pairs (fma.rn.f32x2): the consumers are issue-bound, one instruction per two channels
typedef unsigned long long pk2;
__device__ __forceinline__ pk2 pk(float lo, float hi) { pk2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi)); return r; }
__device__ __forceinline__ void upk(pk2 v, float& lo, float& hi) { asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v)); }
__device__ __forceinline__ pk2 mul2(pk2 a, pk2 b) { pk2 r; asm("mul.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b)); return r; }
__device__ __forceinline__ pk2 fma2(pk2 a, pk2 b, pk2 c) { pk2 r; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c)); return r; }
__device__ __forceinline__ pk2 bf2(uint32_t u) { return pk(bf_lo(u), bf_hi(u)); }      // a bf16 pair as two fp32
__device__ __forceinline__ float dot8(const pk2 (&g)[4], uint4 v) {
  pk2 a = mul2(g[0], bf2(v.x));
  a = fma2(g[1], bf2(v.y), a); a = fma2(g[2], bf2(v.z), a); a = fma2(g[3], bf2(v.w), a);
  float lo, hi;
  upk(a, lo, hi);
  return lo + hi;
}

// dgrad lane layout: lane `chunk` (0..7) of a pixel owns channels [4c, 4c+4) and [32+4c, 32+4c+4) of the
// 64-channel block, so that the 8 lanes' vector REDs into the fp32 NHWC gradient cover whole 32-byte sectors
// (8 x 16 B contiguous) -- with 8 consecutive channels per lane every RED would touch half a sector.
__device__ __forceinline__ uint4 ldg_gather8x2(const __nv_bfloat16* p) {
  uint4 v;
  asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(p));
  asm volatile("ld.global.nc.v2.u32 {%0,%1}, [%2];" : "=r"(v.z), "=r"(v.w) : "l"(p + 32));
  return v;
}

constexpr int kDgMaxRing = 4;
struct DgradBarriers {
  uint64_t geo_full[2], geo_empty[2];       // raw geometry of a tile               (geometry warps -> consumers)
  uint64_t gout_full[2], gout_empty[2];     // grad_output tile                      (geometry warps -> MMA)
  uint64_t wt_full[kDgMaxRing], wt_empty[kDgMaxRing];   // transposed weight tiles   (bulk copies   -> MMA)
  uint64_t acc_full[2], acc_empty[2];       // TMEM column-gradient accumulators     (MMA -> transfer warps)
  uint64_t g_full[kDgMaxRing], g_empty[kDgMaxRing];     // bf16 column-gradient tile (transfer -> consumers)
};
constexpr int kDgGeoBytes = 2 * 9 * kTileM * 16;
constexpr int kDgAccBytes = 27 * kTileM * 4;

// Persistent, warp-specialised dgrad kernel, one CTA per SM over 8 x 16 pixel tiles:
//   warps [0, NP)        consumers: per K block read the bf16 column gradient G[128 px][64 ci] from shared
//                        memory and the four corner vectors of x (8 lanes x 16 B per pixel, prefetched one
//                        K block ahead like the forward's producer): 4 dot products -> grad_mask /
//                        grad_offset (8-lane shuffle reduce), G * mask * corner weight -> grad_input (RED.v4)
//   warps [NP, NP+4)     transfer: TMEM accumulator -> bf16 -> swizzled shared-memory G tile
//   warps [NP+4, NP+8)   grad_output tile + raw sampling geometry of the tiles ahead (thread = pixel)
//   warp  NP+8           one lane issues tcgen05.mma: G = gout[128 px, Cout] * Wt_kb[Cout, 64 ci]
//   warp  NP+9           one lane streams the transposed weight tiles with 1-D bulk copies
template <int NP>
__global__ void __launch_bounds__((NP + 10) * 32, 1)
dcn_dgrad_kernel(const DcnParams prm, const DcnBwdParams bw) {
  static_assert(NP % 4 == 0 && 32 % NP == 0, "consumer warps must tile the 32 row groups");
  extern __shared__ __align__(1024) uint8_t dcn_smem[];
  __shared__ DgradBarriers bars;
  __shared__ uint32_t s_tmem;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int Cin = prm.Cin, Cout = prm.Cout;
  const int P = prm.Ho * prm.Wo;
  const int NG = bw.g_stages, NW = bw.w_stages, GB = bw.gout_bufs;
  const uint32_t gout_bytes = (uint32_t)(Cout / 64) * kTileM * 128u, wt_bytes = (uint32_t)Cout * 128u;
  const uint32_t g_bytes = (uint32_t)kTileM * 128u;
  uint8_t* smem = dcn_smem + ((1024u - (smem_u32(dcn_smem) & 1023u)) & 1023u);
  uint8_t* gout = smem;                                               // [GB][Cout/64][128][128 B]
  uint8_t* wtr = gout + (size_t)GB * gout_bytes;                      // [NW][Cout/64][64][128 B]
  uint8_t* gring = wtr + (size_t)NW * wt_bytes;                       // [NG][128][128 B]
  uint4* geo = reinterpret_cast<uint4*>(gring + (size_t)NG * g_bytes);   // [2][9][128] {g, lh, lw, mask}
  float* acc = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(geo) + kDgGeoBytes);   // [27][128]
  const int cpb = Cin / kBK, KB = 9 * cpb;
  const int tile_first = (int)(((long long)blockIdx.x * prm.num_tiles) / gridDim.x);
  const int my_tiles = (int)(((long long)(blockIdx.x + 1) * prm.num_tiles) / gridDim.x) - tile_first;
  const int total = my_tiles * KB;

  if (tid == 0) {
    for (int i = 0; i < 2; ++i) {
      dcn_mbar_init(&bars.geo_full[i], 4);   dcn_mbar_init(&bars.geo_empty[i], NP);
      dcn_mbar_init(&bars.gout_full[i], 4);  dcn_mbar_init(&bars.gout_empty[i], 1);
      dcn_mbar_init(&bars.acc_full[i], 1);   dcn_mbar_init(&bars.acc_empty[i], 4);
    }
    for (int i = 0; i < kDgMaxRing; ++i) {
      dcn_mbar_init(&bars.wt_full[i], 1);    dcn_mbar_init(&bars.wt_empty[i], 1);
      dcn_mbar_init(&bars.g_full[i], 4);     dcn_mbar_init(&bars.g_empty[i], NP);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == NP + 8) tmem_alloc(&s_tmem, 128);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = s_tmem;

  if (warp < NP) {
    // ===================================== consumers ==============================================
    // the consumers are the register-hungry role: take the registers the 4 geometry warps give back
    // (CTA-local pool: 4 warps x 32 released = 16 warps x 8 acquired, or the inc would block forever)
    if (NP == 16) asm volatile("setmaxnreg.inc.sync.aligned.u32 80;");
    constexpr int ITERS = 32 / NP;
    const int sub = lane >> 3, chunk = lane & 7;
    const uint32_t r7 = (uint32_t)(((warp & 1) << 2) + sub);                       // row & 7 is fixed per thread
    const uint32_t g_half = (uint32_t)(chunk & 1) * 8u;                             // 8-byte half of a 16-byte chunk
    const uint32_t g_sw_lo = (((uint32_t)(chunk >> 1)) ^ r7) << 4, g_sw_hi = ((4u + (uint32_t)(chunk >> 1)) ^ r7) << 4;
    const uint32_t row0 = (uint32_t)(warp << 2) + sub;
    const ptrdiff_t ex = Cin, ey = (ptrdiff_t)prm.xw * Cin;                        // bordered rows
    uint4 va[ITERS], vb[ITERS], vc[ITERS], vd[ITERS];
    auto gather = [&](int i, const __nv_bfloat16* xc, uint32_t e0) {
      const __nv_bfloat16* p = xc + e0;            // one 64-bit multiply-add, the other corners are adds
      va[i] = ldg_gather8x2(p);
      vb[i] = ldg_gather8x2(p + ex);
      vc[i] = ldg_gather8x2(p + ey);
      vd[i] = ldg_gather8x2(p + ey + ex);
    };
    int c_img = 0, c_pix[ITERS];                 // image and output pixel (or -1) of this thread's rows
    auto set_tile = [&](int it) {
      int tile = tile_first + it;
      c_img = tile / prm.tiles_per_image;
      tile -= c_img * prm.tiles_per_image;
      const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
#pragma unroll
      for (int i = 0; i < ITERS; ++i) {
        const int r = (int)row0 + i * NP * 4;
        const int ho = tile_y * kTileH + (r >> 4), wo = tile_x * kTileW + (r & 15);
        c_pix[i] = (ho < prm.Ho && wo < prm.Wo) ? ho * prm.Wo + wo : -1;
      }
    };
    int l_it = 0, l_cb = 0, l_t = 0;             // load cursor: the K block whose gathers are in flight
    const __nv_bfloat16* l_x = prm.xh + chunk * 4;
    if (total > 0) {
      set_tile(0);
      l_x = prm.xh + (size_t)c_img * prm.ximg + chunk * 4;
      dcn_mbar_wait(&bars.geo_full[0], 0u);
#pragma unroll
      for (int i = 0; i < ITERS; ++i) gather(i, l_x, geo[row0 + i * NP * 4].x);
    }
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
            l_x = prm.xh + (size_t)(tile / prm.tiles_per_image) * prm.ximg + chunk * 4;
            dcn_mbar_wait(&bars.geo_full[l_it & 1], (unsigned)((l_it >> 1) & 1));
          }
        }
      }
      const __nv_bfloat16* l_xc = l_x + l_cb * kBK;
      const uint4* geo_n = geo + ((l_it & 1) * 9 + l_t) * kTileM + row0;
      const uint4* geo_c = geo + ((c_it & 1) * 9 + c_t) * kTileM + row0;
      float* gxp = bw.gx + (size_t)c_img * prm.ximg + c_cb * kBK + chunk * 4;
      asm volatile("" : "+l"(gxp));                // keep the base in registers (else it is re-derived per corner)
      dcn_mbar_wait(&bars.g_full[gs], gph);
      const uint8_t* g_tile = gring + (size_t)gs * g_bytes;
#pragma unroll
      for (int i = 0; i < ITERS; ++i) {
        const uint4 q4 = geo_c[i * NP * 4];
        uint4 gq;                                  // G channels [4c,4c+4) and [32+4c,32+4c+4) of this row
        {
          const uint8_t* grow = g_tile + (row0 + (uint32_t)(i * NP * 4)) * 128u + g_half;
          const uint2 lo = *reinterpret_cast<const uint2*>(grow + g_sw_lo);
          const uint2 hi = *reinterpret_cast<const uint2*>(grow + g_sw_hi);
          gq = make_uint4(lo.x, lo.y, hi.x, hi.y);
        }
        const uint32_t e0 = q4.x;
        const float lh = __uint_as_float(q4.y), lw = __uint_as_float(q4.z), m = __uint_as_float(q4.w);
        const pk2 g2[4] = {bf2(gq.x), bf2(gq.y), bf2(gq.z), bf2(gq.w)};
        const float d1 = dot8(g2, va[i]), d2 = dot8(g2, vb[i]), d3 = dot8(g2, vc[i]), d4 = dot8(g2, vd[i]);
        const float hh = 1.f - lh, hw = 1.f - lw;
        const float w1 = hh * hw, w2 = hh * lw, w3 = lh * hw, w4 = lh * lw;
        {                                        // grad_input: G * mask * corner weight (deform_conv_cuda_kernel.cu:944)
          float* gp = gxp + e0;
          asm volatile("" : "+l"(gp));
          auto scat = [&](float* q, float c) {
            if (c != 0.f) {
              const pk2 cc = pk(c, c);
              float a0, a1, a2, a3;
              upk(mul2(g2[0], cc), a0, a1); upk(mul2(g2[1], cc), a2, a3);
              red_add_v4(q, a0, a1, a2, a3);
              upk(mul2(g2[2], cc), a0, a1); upk(mul2(g2[3], cc), a2, a3);
              red_add_v4(q + 32, a0, a1, a2, a3);
            }
          };
          scat(gp, m * w1); scat(gp + ex, m * w2); scat(gp + ey, m * w3); scat(gp + ey + ex, m * w4);
        }
        if (have_next) gather(i, l_xc, geo_n[i * NP * 4].x);   // refill the slot: K block n+1, same row group
        // this lane's share of the three per-(pixel, tap) sums, then one 8-lane reduction each
        float vm = w1 * d1 + w2 * d2 + w3 * d3 + w4 * d4;             // :1019-1030
        float vh = hw * (d3 - d1) + lw * (d4 - d2);                   // dmcn_get_coordinate_weight, bp_dir 0 (x mask below)
        float vw = hh * (d2 - d1) + lh * (d4 - d3);                   // bp_dir 1
#pragma unroll
        for (int o = 1; o < 8; o <<= 1) {
          vm += __shfl_xor_sync(0xffffffffu, vm, o);
          vh += __shfl_xor_sync(0xffffffffu, vh, o);
          vw += __shfl_xor_sync(0xffffffffu, vw, o);
        }
        if (chunk == 0) {
          vh *= m; vw *= m;
          float* a = acc + (3 * c_t) * kTileM + row0 + i * NP * 4;      // partial sums over channel blocks
          if (c_cb > 0) { vm += a[0]; vh += a[kTileM]; vw += a[2 * kTileM]; }
          if (c_cb == cpb - 1) {
            if (c_pix[i] >= 0) {
              bw.goff[((size_t)c_img * 18 + 2 * c_t) * P + c_pix[i]] = vh;
              bw.goff[((size_t)c_img * 18 + 2 * c_t + 1) * P + c_pix[i]] = vw;
              bw.gmask[((size_t)c_img * 9 + c_t) * P + c_pix[i]] = vm;
            }
          } else {
            a[0] = vm; a[kTileM] = vh; a[2 * kTileM] = vw;
          }
        }
      }
      __syncwarp();
      if (lane == 0) {
        dcn_mbar_arrive(&bars.g_empty[gs]);
        if (l_it != c_it) dcn_mbar_arrive(&bars.geo_empty[c_it & 1]);
      }
      if (++gs == NG) { gs = 0; gph ^= 1u; }
      if (l_it != c_it && have_next) set_tile(l_it);
    }
  } else if (warp < NP + 4) {
    // ===================================== transfer: TMEM -> bf16 G tile ==========================
    const int quarter = warp - NP;
    const int r = (quarter << 5) + lane;
    int gs = 0; unsigned gph = 0;
    for (int n = 0; n < total; ++n) {
      const int ab = n & 1;
      dcn_mbar_wait(&bars.acc_full[ab], (unsigned)((n >> 1) & 1));
      tc_fence_after();
      uint32_t pk[32];
#pragma unroll
      for (int c0 = 0; c0 < kBK; c0 += 16) {
        uint32_t v[16];
        tmem_ld16(tmem + ((uint32_t)(quarter * 32) << 16) + (uint32_t)(ab * kBK + c0), v);
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const __nv_bfloat162 h = __floats2bfloat162_rn(__uint_as_float(v[2 * k]), __uint_as_float(v[2 * k + 1]));
          pk[c0 / 2 + k] = *reinterpret_cast<const uint32_t*>(&h);
        }
      }
      tc_fence_before();
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.acc_empty[ab]);            // TMEM buffer may be overwritten
      dcn_mbar_wait(&bars.g_empty[gs], gph ^ 1u);
      uint8_t* dst = gring + (size_t)gs * g_bytes + r * 128;
#pragma unroll
      for (int c = 0; c < 8; ++c)
        *reinterpret_cast<uint4*>(dst + ((c ^ (r & 7)) << 4)) = make_uint4(pk[4 * c], pk[4 * c + 1], pk[4 * c + 2], pk[4 * c + 3]);
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.g_full[gs]);
      if (++gs == NG) { gs = 0; gph ^= 1u; }
    }
  } else if (warp < NP + 8) {
    // ===================================== grad_output tile + geometry ============================
    if (NP == 16) asm volatile("setmaxnreg.dec.sync.aligned.u32 40;");
    const int r = ((warp - NP - 4) << 5) + lane;
    for (int it = 0; it < my_tiles; ++it) {
      int tile = tile_first + it;
      const int gb = it % GB;
      dcn_mbar_wait_relaxed(&bars.gout_empty[gb], (unsigned)(((it / GB) & 1) ^ 1));
      dcn_load_gout_tile(prm, bw.gh, tile, r, gout + (size_t)gb * gout_bytes);
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.gout_full[gb]);

      const int b = tile / prm.tiles_per_image;
      tile -= b * prm.tiles_per_image;
      const int tile_y = tile / prm.tiles_x, tile_x = tile - tile_y * prm.tiles_x;
      const int ho = tile_y * kTileH + (r >> 4), wo = tile_x * kTileW + (r & 15);
      const bool valid = ho < prm.Ho && wo < prm.Wo;
      const int p = valid ? ho * prm.Wo + wo : 0;
      const float* offb = prm.offset + (size_t)b * 18 * P + p;
      const float* mskb = prm.mask + (size_t)b * 9 * P + p;
      float ov[18], mv[9];
#pragma unroll
      for (int k = 0; k < 18; ++k) ov[k] = ldg_stream1(offb + (size_t)k * P);
#pragma unroll
      for (int k = 0; k < 9; ++k) mv[k] = ldg_stream1(mskb + (size_t)k * P);
      const int buf = it & 1;
      dcn_mbar_wait_relaxed(&bars.geo_empty[buf], (unsigned)(((it >> 1) & 1) ^ 1));
#pragma unroll
      for (int t = 0; t < 9; ++t)
        geo[(buf * 9 + t) * kTileM + r] = tap_geometry_raw(prm, ov[2 * t], ov[2 * t + 1], mv[t], ho, wo, valid, t);
      __syncwarp();
      if (lane == 0) dcn_mbar_arrive(&bars.geo_full[buf]);
    }
  } else if (warp == NP + 8) {
    // ===================================== MMA issuer =============================================
    if (lane == 0) {
      const uint32_t idesc = make_idesc_bf16(kTileM, kBK);
      int ws = 0; unsigned wph = 0;
      int n = 0;
      for (int it = 0; it < my_tiles; ++it) {
        const int gb = it % GB;
        dcn_mbar_wait_relaxed(&bars.gout_full[gb], (unsigned)((it / GB) & 1), 0u);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(gout + (size_t)gb * gout_bytes);
        for (int kb = 0; kb < KB; ++kb, ++n) {
          const int ab = n & 1;
          dcn_mbar_wait_relaxed(&bars.acc_empty[ab], (unsigned)(((n >> 1) & 1) ^ 1), 0u);
          dcn_mbar_wait_relaxed(&bars.wt_full[ws], wph, 0u);
          tc_fence_after();
          const uint32_t b_addr = smem_u32(wtr + (size_t)ws * wt_bytes);
          for (int j = 0; j < Cout / 64; ++j) {
            const uint64_t adesc = make_sw128_desc(a_addr + (uint32_t)j * kTileM * 128u);
            const uint64_t bdesc = make_sw128_desc(b_addr + (uint32_t)j * 64u * 128u);
#pragma unroll
            for (int k4 = 0; k4 < 4; ++k4)
              umma_bf16(tmem + (uint32_t)(ab * kBK), adesc + 2 * k4, bdesc + 2 * k4, idesc, (uint32_t)((j | k4) != 0));
          }
          umma_commit(&bars.wt_empty[ws]);
          umma_commit(&bars.acc_full[ab]);
          if (++ws == NW) { ws = 0; wph ^= 1u; }
        }
        umma_commit(&bars.gout_empty[gb]);
      }
    }
    __syncwarp();
  } else {
    // ===================================== transposed-weight tile loader ==========================
    if (lane == 0) {
      int ws = 0; unsigned wph = 0;
      for (int it = 0; it < my_tiles; ++it) {
        for (int kb = 0; kb < KB; ++kb) {
          dcn_mbar_wait_relaxed(&bars.wt_empty[ws], wph ^ 1u, 20u);
          dcn_mbar_arrive_expect_tx(&bars.wt_full[ws], wt_bytes);
          dcn_bulk_g2s(wtr + (size_t)ws * wt_bytes, bw.wt + (size_t)kb * Cout * kBK, wt_bytes, &bars.wt_full[ws]);
          if (++ws == NW) { ws = 0; wph ^= 1u; }
        }
      }
    }
    __syncwarp();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == NP + 8) tmem_dealloc(tmem, 128);
}
#include "dcn_dgrad2.cuh"

}  // namespace ctdet

static inline size_t dcnb_align256(size_t v) { return (v + 255) / 256 * 256; }

extern "C" size_t dcnv2_backward_workspace_bytes(int B, int Cin, int H, int W, int Cout, int Ho, int Wo) {
  if (B <= 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0 || Ho <= 0 || Wo <= 0) return 0;
  const size_t bordered = (size_t)B * (H + 2) * (W + 2) * Cin;       // x (bf16) and grad_input (fp32) with a zero border
  return dcnb_align256(bordered * 2) + dcnb_align256((size_t)B * Ho * Wo * Cout * 2) +
         dcnb_align256((size_t)Cout * Cin * 9 * 2) + dcnb_align256(bordered * 4);
}

extern "C" int dcnv2_backward(const float* input, const float* weight, const float* offset, const float* mask,
                              const float* grad_output, float* grad_input, float* grad_weight, float* grad_bias,
                              float* grad_offset, float* grad_mask,
                              int B, int Cin, int H, int W, int Cout, int kernel_h, int kernel_w,
                              int stride_h, int stride_w, int pad_h, int pad_w, int dilation_h, int dilation_w,
                              int group, int deformable_group, int flags,
                              void* workspace, size_t workspace_bytes, ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0 || flags != 0) return CTDET_EINVAL;
  if (kernel_h != 3 || kernel_w != 3 || group != 1 || deformable_group != 1) return CTDET_EUNSUPPORTED;
  if (stride_h != stride_w || pad_h != pad_w || dilation_h != dilation_w || stride_h < 1 || dilation_h < 1 || pad_h < 0)
    return CTDET_EUNSUPPORTED;
  if (Cin % kBK != 0 || Cout % 64 != 0 || Cout > 256) return CTDET_EUNSUPPORTED;
  const int Ho = (H + 2 * pad_h - (dilation_h * 2 + 1)) / stride_h + 1;
  const int Wo = (W + 2 * pad_w - (dilation_w * 2 + 1)) / stride_w + 1;
  if (Ho <= 0 || Wo <= 0) return CTDET_EINVAL;
  const bool want_d = grad_input || grad_offset || grad_mask;
  if (want_d && !(grad_input && grad_offset && grad_mask)) return CTDET_ENULL;   // the three come from one kernel
  if (B == 0) return CTDET_OK;
  if (!input || !weight || !offset || !mask || !grad_output || !workspace) return CTDET_ENULL;
  if (workspace_bytes < dcnv2_backward_workspace_bytes(B, Cin, H, W, Cout, Ho, Wo)) return CTDET_EWORKSPACE;
  if (reinterpret_cast<uintptr_t>(workspace) & 15u) return CTDET_EALIGN;
  if ((size_t)(H + 2) * (W + 2) >= ((size_t)1 << 31) / (size_t)Cin || B > 65535) return CTDET_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  const int Pin = H * W, P = Ho * Wo;
  const size_t ximg = (size_t)(H + 2) * (W + 2) * Cin;                // elements of one bordered image

  unsigned char* wsp = reinterpret_cast<unsigned char*>(workspace);
  __nv_bfloat16* xh = reinterpret_cast<__nv_bfloat16*>(wsp);
  wsp += dcnb_align256((size_t)B * ximg * 2);
  __nv_bfloat16* gh = reinterpret_cast<__nv_bfloat16*>(wsp);
  wsp += dcnb_align256((size_t)B * P * Cout * 2);
  __nv_bfloat16* wt = reinterpret_cast<__nv_bfloat16*>(wsp);
  wsp += dcnb_align256((size_t)Cout * Cin * 9 * 2);
  float* gx = reinterpret_cast<float*>(wsp);

  cudaError_t e;
  if (grad_bias) {
    e = cudaMemsetAsync(grad_bias, 0, (size_t)Cout * 4, st);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  }
  e = cudaMemsetAsync(xh, 0, (size_t)B * ximg * 2, st);               // the border (the pack fills the interior)
  if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  int vec = (Pin % 4 == 0 && aligned16(input)) ? 1 : 0;
  dcn_pack_input_kernel<<<dim3((Pin + 63) / 64, (Cin + 63) / 64, B), 256, 0, st>>>(input, xh, Cin, Pin, vec, nullptr, 1, W);
  int rc = check_launch();
  if (rc) return rc;
  vec = (P % 4 == 0 && aligned16(grad_output)) ? 1 : 0;     // grad_bias = per-channel sums, same pass
  {
    // with grad_bias a CTA covers several 64-pixel blocks (fewer same-address atomics), but the grid
    // keeps at least ~4 CTAs per SM
    const int blocks = (P + 63) / 64;
    int chunks = 1;
    if (grad_bias)
      while (chunks < 16 && (long long)(blocks / (chunks * 2)) * ((Cout + 63) / 64) * B >= 2LL * kNumSMs) chunks *= 2;
    dcn_pack_input_kernel<<<dim3((blocks + chunks - 1) / chunks, (Cout + 63) / 64, B), 256, 0, st>>>(grad_output, gh, Cout, P, vec,
                                                                                                    grad_bias, chunks, 0);
  }
  rc = check_launch();
  if (rc) return rc;

  DcnParams prm;
  prm.xh = xh; prm.offset = offset; prm.mask = mask; prm.wk = nullptr; prm.bias = nullptr;
  prm.off_bstride = 18 * Ho * Wo; prm.mask_bstride = 9 * Ho * Wo; prm.mask_logit = 0; prm.om_bias = nullptr;
  prm.scale = nullptr; prm.shift = nullptr; prm.y = nullptr;
  prm.Cin = Cin; prm.H = H; prm.W = W; prm.Ho = Ho; prm.Wo = Wo; prm.Cout = Cout;
  prm.stride = stride_h; prm.pad = pad_h; prm.dil = dilation_h; prm.relu = 0;
  prm.xw = W + 2; prm.ximg = (int)ximg; prm.xorg = (W + 3) * Cin;
  prm.tiles_x = (Wo + kTileW - 1) / kTileW;
  prm.tiles_per_image = prm.tiles_x * ((Ho + kTileH - 1) / kTileH);
  const long long tiles = (long long)B * prm.tiles_per_image;
  if (tiles > 0x7fffffffLL) return CTDET_EINVAL;
  prm.num_tiles = (int)tiles;
  prm.mma_sleep = 0; prm.ld_sleep = 20;
  DcnBwdParams bw;
  bw.gh = gh; bw.wt = wt; bw.gw = grad_weight; bw.gx = gx; bw.goff = grad_offset; bw.gmask = grad_mask;
  const int KB = 9 * (Cin / kBK);
  constexpr int NP = 16;

  if (grad_weight) {
    e = cudaMemsetAsync(grad_weight, 0, (size_t)Cout * Cin * 9 * 4, st);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
    const int MT = Cout > 128 ? 2 : 1;
    bw.paired = (Cout == 64 && dcn_tunables().wgrad_paired) ? 1 : 0;
    const int max_kb = bw.paired ? 16 : 512 / (kBK * MT);      // accumulators that fit TMEM
    bw.n_groups = (KB + max_kb - 1) / max_kb;
    bw.kb_per_group = (KB + bw.n_groups - 1) / bw.n_groups;
    uint32_t cols = 32;
    while (cols < (uint32_t)(bw.paired ? (bw.kb_per_group + 1) / 2 * 64 : bw.kb_per_group * kBK * MT)) cols <<= 1;
    prm.tmem_cols = cols;
    const size_t gout_bytes = (size_t)(Cout / 64) * kTileM * 128;
    const size_t fixed = 1024 + kGeoBytes + 2 * gout_bytes;
    int stages = (int)((230 * 1024 - fixed) / (kTileM * kBK * 2));
    const int stage_cap = dcn_tunables().wgrad_stages;
    if (stages > stage_cap) stages = stage_cap;
    if (stages < 2) return CTDET_EUNSUPPORTED;
    if (bw.paired) stages &= ~1;                               // stage pairs
    prm.stages = stages;
    const size_t smem = fixed + (size_t)stages * kTileM * kBK * 2;
    int chunks = kNumSMs / bw.n_groups;
    if (chunks < 1) chunks = 1;
    if ((long long)chunks > tiles) chunks = (int)tiles;
    auto kern = dcn_wgrad_kernel<NP>;
    e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
    kern<<<(unsigned)(chunks * bw.n_groups), (NP + 10) * 32, smem, st>>>(prm, bw);
    rc = check_launch();
    if (rc) return rc;
  }
  if (want_d) {
    e = cudaMemsetAsync(gx, 0, (size_t)B * ximg * 4, st);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
    dcn_pack_weight_t_kernel<<<kNumSMs, 256, 0, st>>>(weight, wt, Cout, Cin);
    rc = check_launch();
    if (rc) return rc;
    const size_t gout_bytes = (size_t)(Cout / 64) * kTileM * 128;
    const unsigned grid = (unsigned)(tiles < kNumSMs ? tiles : kNumSMs);
    const int npd = dcn_tunables().dgrad_warps;
    bool launched = false;
    if (dcn_tunables().dgrad_gemm && npd == 16 && stride_h == 1 && dilation_h == 1 && Cout <= 128 &&
        (size_t)(H + 2) * (W + 2) <= ((size_t)1 << 23)) {
      // second form: the grad_input scatter as a tensor-core GEMM (dcn_dgrad2.cuh)
      bw.gout_bufs = Cout <= 64 ? dcn_tunables().dgrad_gb : 1;
      bw.s_bufs = 2;                                     // (the builders' un-write bookkeeping assumes two buffers)
      const size_t fixed = 1024 + (size_t)bw.gout_bufs * gout_bytes + (size_t)bw.s_bufs * kSChunkBytes + kDgGeoBytes +
                           2 * kDgAccBytes;
      const size_t unit = (size_t)Cout * 128 + (size_t)kTileM * 128;
      int ring = fixed < 226 * 1024 ? (int)((226 * 1024 - fixed) / unit) : 0;
      if (ring > kDgMaxRing) ring = kDgMaxRing;
      if (ring >= 2) {
        bw.g_stages = bw.w_stages = ring;
        const size_t smem = fixed + (size_t)ring * unit;
        auto kern = dcn_dgrad2_kernel<16>;
        e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
        kern<<<grid, (16 + 10) * 32, smem, st>>>(prm, bw);
        rc = check_launch();
        if (rc) return rc;
        launched = true;
      }
    }
    if (!launched) {
      bw.gout_bufs = Cout <= 128 ? 2 : 1;
      const size_t fixed = 1024 + (size_t)bw.gout_bufs * gout_bytes + kDgGeoBytes + kDgAccBytes;
      const size_t unit = (size_t)Cout * 128 + (size_t)kTileM * 128;
      int ring = (int)((226 * 1024 - fixed) / unit);
      if (ring > kDgMaxRing) ring = kDgMaxRing;
      if (ring < 2) return CTDET_EUNSUPPORTED;
      bw.g_stages = bw.w_stages = ring;
      const size_t smem = fixed + (size_t)ring * unit;
      void (*kern)(const DcnParams, const DcnBwdParams) = npd == 8 ? dcn_dgrad_kernel<8> : dcn_dgrad_kernel<16>;
      e = cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
      kern<<<grid, (npd + 10) * 32, smem, st>>>(prm, bw);
      rc = check_launch();
      if (rc) return rc;
    }
    dcn_unpack_grad_input_kernel<<<dim3((Pin + 63) / 64, Cin / 64, B), 256, 0, st>>>(gx, grad_input, Cin, Pin, W);
    rc = check_launch();
    if (rc) return rc;
  }
  return CTDET_OK;
}

#ifdef CTDET_DGRAD_TIMING
extern "C" int ctdet_debug_dgrad_times(unsigned long long* out /* [16][64] host */) {
  return cudaMemcpyFromSymbol(out, ctdet::g_dg2_ts, sizeof(unsigned long long) * 16 * 64) == cudaSuccess ? 0 : 1;
}
#endif
