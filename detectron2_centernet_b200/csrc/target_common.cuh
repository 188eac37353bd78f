// Object geometry shared by the target-generation kernel (target_gen.cu) and the fused
// target + loss kernel (loss.cu): the reference's box arithmetic and Gaussian radius, bit for bit.
#pragma once
#include "common.cuh"

namespace ctdet {

// gaussian_radius (detection_utils.py:654-680) in fp64 with the reference's exact operation
// order; every product/sum that is not exact is an explicit round-to-nearest op so that ptxas
// cannot contract it into an FMA.  height/width are the math.ceil'ed box sides (:623).
__device__ __forceinline__ int gaussian_radius_int(int hh, int ww) {
  const double mo = 0.7;
  const double height = (double)hh, width = (double)ww;
  const double s = height + width;                       // exact (integers)
  const double area = width * height;                    // exact (integers)
  const double c1 = __ddiv_rn(__dmul_rn(area, 1 - mo), 1 + mo);
  const double sq1 = __dsqrt_rn(__dsub_rn(s * s, 4.0 * c1));
  const double r1 = __dadd_rn(s, sq1) * 0.5;
  const double b2 = 2.0 * s;
  const double c2 = __dmul_rn(__dmul_rn(1 - mo, width), height);
  const double sq2 = __dsqrt_rn(__dsub_rn(b2 * b2, 16.0 * c2));
  const double r2 = __dadd_rn(b2, sq2) * 0.5;
  const double a3 = 4 * mo;
  const double b3 = __dmul_rn(-2 * mo, s);
  const double c3 = __dmul_rn(__dmul_rn(mo - 1, width), height);
  const double sq3 = __dsqrt_rn(__dsub_rn(__dmul_rn(b3, b3), __dmul_rn(4 * a3, c3)));
  const double r3 = __dadd_rn(b3, sq3) * 0.5;
  const double r = fmin(r1, fmin(r2, r3));
  const int ri = (int)r;                                 // int(): truncate toward zero (:624)
  return ri > 0 ? ri : 0;
}

struct ObjGeom {
  float bw, bh, ctx, cty;
  int cxi, cyi;
  bool valid;
};

// fp32 box arithmetic exactly as torch evaluates detection_utils.py:618-627.
__device__ __forceinline__ ObjGeom object_geometry(const float* __restrict__ box) {
  ObjGeom g;
  const float x1 = box[0] * 0.25f, y1 = box[1] * 0.25f;   // "/ 4" (:618), exact
  const float x2 = box[2] * 0.25f, y2 = box[3] * 0.25f;
  g.bw = __fsub_rn(x2, x1);
  g.bh = __fsub_rn(y2, y1);
  g.valid = (g.bh > 0.f) && (g.bw > 0.f);                 // :621
  g.ctx = __fadd_rn(x1, x2) * 0.5f;                       // :625-626
  g.cty = __fadd_rn(y1, y2) * 0.5f;
  g.cxi = (int)g.ctx;                                     // astype(int32): truncation (:627)
  g.cyi = (int)g.cty;
  return g;
}


// Slot targets (wh / reg / ind / reg_mask, detection_utils.py:629-632) and num_pos for image
// blockIdx.x without touching the heatmap; defined in target_gen.cu.
__global__ void target_slots_kernel(const float* __restrict__ boxes, const int64_t* __restrict__ classes,
                                    const int32_t* __restrict__ offsets, int C, int H, int W,
                                    uint8_t* __restrict__ reg_mask, int64_t* __restrict__ ind,
                                    float* __restrict__ wh, float* __restrict__ reg, int32_t* __restrict__ num_pos,
                                    int4* __restrict__ obj_geo, double* __restrict__ obj_ninv, int32_t* __restrict__ cls_count);

}  // namespace ctdet
