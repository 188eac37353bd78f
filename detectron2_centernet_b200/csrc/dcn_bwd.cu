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
  float* gx;                   // grad_input NHWC fp32 [B][H*W][Cin] (zeroed before the launch)
  float* goff;                 // grad_offset [B][18][Ho*Wo]
  float* gmask;                // grad_mask   [B][9][Ho*Wo]
  int n_groups, kb_per_group;  // wgrad: K blocks are split into groups that fit TMEM
  int g_stages, w_stages, gout_bufs;   // dgrad ring depths
};

// MN-major, 128-byte-swizzled operand: rows of 64 contiguous M/N elements (128 B) per K index, 8-row
// (K) groups 1024 B apart (SBO); LBO = distance between 64-element atoms along M/N.
__device__ __forceinline__ uint64_t make_sw128_desc_mn(uint32_t smem_addr, uint32_t lbo_bytes) {
  return (uint64_t)((smem_addr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
         ((uint64_t)(1024 >> 4) << 32) | (1ull << 46) | (2ull << 61);
}

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
    if (nkb > 0)
      dcn_producer_loop<false, NP>(prm, smem, stage_bytes, NS, geo_w, geo_g, bars, warp, lane, tile_first, my_tiles, kb0, kb1);
  } else if (warp < NP + 4) {
    // ============ final flush: TMEM -> fp32 atomics into grad_weight[co][ci][t] ====================
    if (work) {
      const int quarter = warp - NP;
      const int r = (quarter << 5) + lane;
      dcn_mbar_wait_relaxed(&bars.acc_full[0], 0u);
      tc_fence_after();
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


// HOST_PLACEHOLDER_DGRAD
}  // namespace ctdet

static inline size_t dcnb_align256(size_t v) { return (v + 255) / 256 * 256; }

extern "C" size_t dcnv2_backward_workspace_bytes(int B, int Cin, int H, int W, int Cout, int Ho, int Wo) {
  if (B <= 0 || Cin <= 0 || H <= 0 || W <= 0 || Cout <= 0 || Ho <= 0 || Wo <= 0) return 0;
  return dcnb_align256((size_t)B * H * W * Cin * 2) + dcnb_align256((size_t)B * Ho * Wo * Cout * 2) +
         dcnb_align256((size_t)Cout * Cin * 9 * 2) + dcnb_align256((size_t)B * H * W * Cin * 4);
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
  if ((size_t)H * W >= ((size_t)1 << 31) / (size_t)Cin || B > 65535) return CTDET_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  const int Pin = H * W, P = Ho * Wo;

  unsigned char* wsp = reinterpret_cast<unsigned char*>(workspace);
  __nv_bfloat16* xh = reinterpret_cast<__nv_bfloat16*>(wsp);
  wsp += dcnb_align256((size_t)B * Pin * Cin * 2);
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
  int vec = (Pin % 4 == 0 && aligned16(input)) ? 1 : 0;
  dcn_pack_input_kernel<<<dim3((Pin + 63) / 64, (Cin + 63) / 64, B), 256, 0, st>>>(input, xh, Cin, Pin, vec, nullptr);
  int rc = check_launch();
  if (rc) return rc;
  vec = (P % 4 == 0 && aligned16(grad_output)) ? 1 : 0;     // grad_bias = per-channel sums, same pass
  dcn_pack_input_kernel<<<dim3((P + 63) / 64, (Cout + 63) / 64, B), 256, 0, st>>>(grad_output, gh, Cout, P, vec, grad_bias);
  rc = check_launch();
  if (rc) return rc;

  DcnParams prm;
  prm.xh = xh; prm.offset = offset; prm.mask = mask; prm.wk = nullptr; prm.bias = nullptr;
  prm.scale = nullptr; prm.shift = nullptr; prm.y = nullptr;
  prm.Cin = Cin; prm.H = H; prm.W = W; prm.Ho = Ho; prm.Wo = Wo; prm.Cout = Cout;
  prm.stride = stride_h; prm.pad = pad_h; prm.dil = dilation_h; prm.relu = 0;
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
    const int max_kb = 512 / (kBK * MT);                       // accumulators that fit TMEM
    bw.n_groups = (KB + max_kb - 1) / max_kb;
    bw.kb_per_group = (KB + bw.n_groups - 1) / bw.n_groups;
    uint32_t cols = 32;
    while (cols < (uint32_t)(bw.kb_per_group * kBK * MT)) cols <<= 1;
    prm.tmem_cols = cols;
    const size_t gout_bytes = (size_t)(Cout / 64) * kTileM * 128;
    const size_t fixed = 1024 + kGeoBytes + 2 * gout_bytes;
    int stages = (int)((230 * 1024 - fixed) / (kTileM * kBK * 2));
    if (stages > 4) stages = 4;
    if (stages < 2) return CTDET_EUNSUPPORTED;
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
    // DGRAD_LAUNCH_PLACEHOLDER
    return CTDET_EUNSUPPORTED;
  }
  return CTDET_OK;
}
