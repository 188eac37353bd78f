// Gaussian target generation for a whole batch in ONE launch.
//
// Replaces the host-side numpy loop of detectron2/data/detection_utils.py:600-705
// (gen_heatmap / gaussian_radius / gaussian2D / draw_umich_gaussian) and the torch.cat + H2D of
// detectron2/modeling/meta_arch/centernet.py:197-201.
//
// Grid = (row bands, classes, images).  One CTA owns one band of one class plane: it bins the
// image's objects of that class into shared memory, zero-fills the band in shared memory, splats
// every binned Gaussian with a max-reduce (atomicMax on the int view: all values are >= 0),
// and writes the band to HBM exactly once with a single bulk async copy (cp.async.bulk
// shared -> global, "UBLKCP" in SASS).  Algorithmic traffic: 4 B per heatmap cell.
#include "target_common.cuh"

namespace ctdet {

constexpr int kTgThreads = 256;
constexpr int kBandBytes = 32 * 1024;

__global__ void __launch_bounds__(kTgThreads)
target_gen_kernel(const float* __restrict__ boxes, const int64_t* __restrict__ classes,
                  const int32_t* __restrict__ offsets, int C, int H, int W, int band_rows,
                  float* __restrict__ hm, uint8_t* __restrict__ reg_mask,
                  int64_t* __restrict__ ind, float* __restrict__ wh, float* __restrict__ reg,
                  int32_t* __restrict__ num_pos, int bulk_ok) {
  extern __shared__ __align__(128) float s_band[];
  __shared__ int s_cx[CTDET_MAX_OBJS], s_cy[CTDET_MAX_OBJS], s_r[CTDET_MAX_OBJS];
  __shared__ double s_ninv[CTDET_MAX_OBJS];
  __shared__ uint32_t s_key[CTDET_MAX_OBJS];
  __shared__ int s_nbin, s_npos;

  const int band = blockIdx.x, c = blockIdx.y, b = blockIdx.z;
  const int tid = threadIdx.x;
  const int y0 = band * band_rows;
  const int y1 = min(H, y0 + band_rows);
  const int o0 = offsets[b];
  const int n = min(offsets[b + 1] - o0, CTDET_MAX_OBJS);  // extras dropped (:615-616)
  const bool slot_cta = (band == 0 && c == 0);

  if (tid == 0) { s_nbin = 0; s_npos = 0; }
  __syncthreads();

  // ---- bin the objects of class c whose window touches this band; slot outputs ----------
  if (tid < CTDET_MAX_OBJS) {
    const int k = tid;
    bool have = k < n;
    int64_t cls = have ? classes[o0 + k] : -1;
    ObjGeom g;
    g.valid = false;
    if (have && (slot_cta || cls == c)) g = object_geometry(boxes + 4 * (size_t)(o0 + k));
    const bool inside = g.valid && g.cxi >= 0 && g.cxi < W && g.cyi >= 0 && g.cyi < H;
    if (inside && cls == c) {
      const int r = gaussian_radius_int((int)ceilf(g.bh), (int)ceilf(g.bw));
      if (g.cyi + r >= y0 && g.cyi - r < y1) {
        const int slot = atomicAdd(&s_nbin, 1);
        const double sigma = (double)(2 * r + 1) / 6.0;   // :692
        s_cx[slot] = g.cxi; s_cy[slot] = g.cyi; s_r[slot] = r;
        s_ninv[slot] = -1.0 / (2.0 * sigma * sigma);
      }
    }
    if (slot_cta) {
      const size_t o = (size_t)b * CTDET_MAX_OBJS + k;
      const bool v = g.valid;
      wh[2 * o] = v ? g.bw : 0.f;                          // :629
      wh[2 * o + 1] = v ? g.bh : 0.f;
      reg[2 * o] = v ? __fsub_rn(g.ctx, (float)g.cxi) : 0.f;   // :631
      reg[2 * o + 1] = v ? __fsub_rn(g.cty, (float)g.cyi) : 0.f;
      ind[o] = v ? (int64_t)g.cyi * W + g.cxi : 0;         // :630
      reg_mask[o] = v ? 1 : 0;                             // :632
      // a cell is a positive (hm == 1) iff it is the centre of a splatted object
      s_key[k] = (inside && cls >= 0 && cls < C)
                     ? (uint32_t)((int)cls * H * W + g.cyi * W + g.cxi) : 0xffffffffu;
    }
  }

  // ---- zero-fill the band in shared memory -----------------------------------------------
  const int band_elems = (y1 - y0) * W;
  for (int i = tid * 4; i < band_elems; i += kTgThreads * 4) {
    if (i + 3 < band_elems) {
      *reinterpret_cast<float4*>(s_band + i) = make_float4(0.f, 0.f, 0.f, 0.f);
    } else {
      for (int j = i; j < band_elems; ++j) s_band[j] = 0.f;
    }
  }
  __syncthreads();

  if (slot_cta && num_pos != nullptr) {
    if (tid < CTDET_MAX_OBJS) {
      const uint32_t key = s_key[tid];
      bool first = key != 0xffffffffu;
      for (int j = 0; j < tid && first; ++j) first = s_key[j] != key;
      if (first) atomicAdd(&s_npos, 1);
    }
  }

  // ---- splat: one warp per object, lanes over the clipped window ---------------------------
  const int nbin = s_nbin;
  const int warp = tid >> 5, lane = tid & 31;
  for (int o = warp; o < nbin; o += kTgThreads / 32) {
    const int cx = s_cx[o], cy = s_cy[o], r = s_r[o];
    const double ninv = s_ninv[o];
    const int wx0 = max(cx - r, 0), wx1 = min(cx + r, W - 1);     // :698
    const int wy0 = max(cy - r, y0), wy1 = min(cy + r, y1 - 1);   // :699, clipped to band
    const int ncols = wx1 - wx0 + 1;
    const int total = ncols * (wy1 - wy0 + 1);
    for (int i = lane; i < total; i += 32) {
      const int yy = wy0 + i / ncols, xx = wx0 + i % ncols;
      const int dx = xx - cx, dy = yy - cy;
      const float t = (float)((double)(dx * dx + dy * dy) * ninv);   // :686, exponent in fp64
      const float v = expf(t);                                       // exactly 1.0f at the centre
      atomicMax(reinterpret_cast<int*>(s_band + (yy - y0) * W + xx), __float_as_int(v));  // :704
    }
  }

  // ---- write the band once ---------------------------------------------------------------
  float* dst = hm + ((size_t)(b * C + c) * H + y0) * W;
  if (bulk_ok) {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (tid == 0) {
      const uint32_t src = (uint32_t)__cvta_generic_to_shared(s_band);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;"
                   :: "l"(dst), "r"(src), "r"((uint32_t)(band_elems * 4)) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      if (slot_cta && num_pos != nullptr) num_pos[b] = s_npos;
      asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
    }
  } else {
    __syncthreads();
    for (int i = tid; i < band_elems; i += kTgThreads) dst[i] = s_band[i];
    if (tid == 0 && slot_cta && num_pos != nullptr) num_pos[b] = s_npos;
  }
}

// Slot outputs and num_pos only (one CTA of 128 threads per image): the part of gen_heatmap that
// the fused target + loss kernel needs before it starts (it never materialises the heatmap).
__global__ void __launch_bounds__(CTDET_MAX_OBJS)
target_slots_kernel(const float* __restrict__ boxes, const int64_t* __restrict__ classes,
                    const int32_t* __restrict__ offsets, int C, int H, int W,
                    uint8_t* __restrict__ reg_mask, int64_t* __restrict__ ind,
                    float* __restrict__ wh, float* __restrict__ reg, int32_t* __restrict__ num_pos,
                    int4* __restrict__ obj_geo, double* __restrict__ obj_ninv, int32_t* __restrict__ cls_count) {
  __shared__ uint32_t s_key[CTDET_MAX_OBJS];
  __shared__ int s_npos;
  const int b = blockIdx.x, k = threadIdx.x;
  const int o0 = offsets[b];
  const int n = min(offsets[b + 1] - o0, CTDET_MAX_OBJS);
  if (k == 0) s_npos = 0;
  if (cls_count != nullptr) {
    for (int c = k; c < C; c += CTDET_MAX_OBJS) cls_count[(size_t)b * C + c] = 0;
    __syncthreads();
  }
  const bool have = k < n;
  const int64_t cls = have ? classes[o0 + k] : -1;
  ObjGeom g;
  g.valid = false;
  if (have) g = object_geometry(boxes + 4 * (size_t)(o0 + k));
  const bool inside = g.valid && g.cxi >= 0 && g.cxi < W && g.cyi >= 0 && g.cyi < H;
  const size_t o = (size_t)b * CTDET_MAX_OBJS + k;
  const bool v = g.valid;
  wh[2 * o] = v ? g.bw : 0.f;
  wh[2 * o + 1] = v ? g.bh : 0.f;
  reg[2 * o] = v ? __fsub_rn(g.ctx, (float)g.cxi) : 0.f;
  reg[2 * o + 1] = v ? __fsub_rn(g.cty, (float)g.cyi) : 0.f;
  ind[o] = v ? (int64_t)g.cyi * W + g.cxi : 0;
  reg_mask[o] = v ? 1 : 0;
  s_key[k] = (inside && cls >= 0 && cls < C) ? (uint32_t)((int)cls * H * W + g.cyi * W + g.cxi) : 0xffffffffu;
  if (obj_geo != nullptr) {
    // splat table of the fused target + loss kernel: centre, radius, class (-1 = not splatted), -1/(2 sigma^2)
    int4 q = make_int4(0, 0, 0, -1);
    double ninv = 0.0;
    if (inside && cls >= 0 && cls < C) {
      const int r = gaussian_radius_int((int)ceilf(g.bh), (int)ceilf(g.bw));
      const double sigma = (double)(2 * r + 1) / 6.0;        // :692
      q = make_int4(g.cxi, g.cyi, r, (int)cls);
      ninv = -1.0 / (2.0 * sigma * sigma);
      atomicAdd(&cls_count[(size_t)b * C + (int)cls], 1);
    }
    obj_geo[o] = q;
    obj_ninv[o] = ninv;
  }
  __syncthreads();
  const uint32_t key = s_key[k];
  // distinct (class, centre) cells: an object counts when no earlier slot has its key.  No early exit:
  // the broadcast LDS pipeline instead of forming a dependent chain of up to 127 compares.
  bool dup = false;
#pragma unroll 8
  for (int j = 0; j < CTDET_MAX_OBJS; ++j) dup |= (j < k) & (s_key[j] == key);
  const bool first = key != 0xffffffffu && !dup;
  const unsigned bal = __ballot_sync(0xffffffffu, first);
  if ((k & 31) == 0 && bal) atomicAdd(&s_npos, __popc(bal));
  __syncthreads();
  if (k == 0 && num_pos != nullptr) num_pos[b] = s_npos;
}

}  // namespace ctdet

extern "C" int ctdet_target_gen(const float* boxes, const int64_t* classes, const int32_t* offsets,
                                int B, int C, int H, int W, float* hm, uint8_t* reg_mask,
                                int64_t* ind, float* wh, float* reg, int32_t* num_pos,
                                ctdet_stream_t stream) {
  using namespace ctdet;
  if (B < 0 || C <= 0 || H <= 0 || W <= 0) return CTDET_EINVAL;
  if ((int64_t)C * H * W >= (int64_t)1 << 31) return CTDET_EINVAL;
  if (C > 65535 || B > 65535) return CTDET_EINVAL;
  if (B == 0) return CTDET_OK;
  if (!offsets || !hm || !reg_mask || !ind || !wh || !reg) return CTDET_ENULL;
  // boxes/classes may be NULL only when there are no objects at all (offsets all zero); the
  // kernel never dereferences them in that case.
  if (!aligned16(hm)) return CTDET_EALIGN;
  if ((size_t)W * 4 > 200 * 1024) return CTDET_EUNSUPPORTED;
  int band_rows = kBandBytes / (W * 4);
  if (band_rows < 1) band_rows = 1;
  if (band_rows > H) band_rows = H;
  const int bands = (H + band_rows - 1) / band_rows;
  const size_t smem = (size_t)band_rows * W * 4;
  // cp.async.bulk needs 16-byte aligned addresses and sizes for every band
  const int bulk_ok = (W % 4) == 0;
  if (smem > 48 * 1024) {
    cudaError_t e = cudaFuncSetAttribute(target_gen_kernel,
                                         cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (e != cudaSuccess) { set_last_cuda_error(e); return CTDET_ECUDA; }
  }
  dim3 grid(bands, C, B);
  target_gen_kernel<<<grid, kTgThreads, smem, (cudaStream_t)stream>>>(
      boxes, classes, offsets, C, H, W, band_rows, hm, reg_mask, ind, wh, reg, num_pos, bulk_ok);
  return check_launch();
}
