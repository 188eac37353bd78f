// Micro-benchmark: bilinear RGBA16F fetches from a pitch-linear 2-D texture (the access pattern of a DCN gather:
// per output pixel 9 taps x 16 channel groups of 4), against the same bytes gathered with LDG.128 from NHWC bf16.
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>
#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s failed: %s\n", #x, cudaGetErrorString(e)); return 1; } } while (0)

__global__ void tex_kernel(cudaTextureObject_t tex, const float* __restrict__ off, float* out, int H, int W, int planes, int reps) {
  // thread = (pixel, channel group): 16 consecutive lanes = the 16 groups of one pixel
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int cg = gid & 15, pix = gid >> 4;
  if (pix >= H * W) return;
  const int y = pix / W, x = pix - y * W;
  float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
  for (int r = 0; r < reps; ++r) {
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float oy = off[(size_t)(2 * t) * H * W + pix], ox = off[(size_t)(2 * t + 1) * H * W + pix];
      const float fy = (float)(y + t / 3 - 1) + oy + 0.5f + (float)(cg * (H + 1)), fx = (float)(x + t % 3 - 1) + ox + 0.5f;
      const float4 v = tex2D<float4>(tex, fx, fy);
      acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
    }
  }
  if (acc.x == 12345.f) out[gid] = acc.x + acc.y + acc.z + acc.w;
}

__global__ void ldg_kernel(const uint4* __restrict__ xh, const float* __restrict__ off, float* out, int H, int W, int reps) {
  // thread = (pixel, 16-byte chunk of 8 channels): 8 lanes per pixel, 4 corners per tap (what dcn_producer_loop does)
  const int gid = blockIdx.x * blockDim.x + threadIdx.x;
  const int ch = gid & 7, pix = gid >> 3;
  if (pix >= H * W) return;
  const int y = pix / W, x = pix - y * W;
  unsigned acc = 0;
  for (int r = 0; r < reps; ++r) {
#pragma unroll
    for (int t = 0; t < 9; ++t) {
      const float oy = off[(size_t)(2 * t) * H * W + pix], ox = off[(size_t)(2 * t + 1) * H * W + pix];
      int y0 = (int)floorf((float)(y + t / 3 - 1) + oy), x0 = (int)floorf((float)(x + t % 3 - 1) + ox);
      y0 = min(max(y0, 0), H - 2); x0 = min(max(x0, 0), W - 2);
      const uint4* p = xh + ((size_t)y0 * W + x0) * 8 + ch;
      const uint4 a = __ldg(p), b = __ldg(p + 8), c = __ldg(p + (size_t)W * 8), d = __ldg(p + (size_t)W * 8 + 8);
      acc += a.x ^ b.y ^ c.z ^ d.w;
    }
  }
  if (acc == 12345u) out[gid] = (float)acc;
}

int main() {
  const int H = 128, W = 128, planes = 16, B = 1, reps = 8;
  const int rows = planes * (H + 1);
  __half* d; size_t pitch;
  CK(cudaMallocPitch(&d, &pitch, (size_t)W * 4 * sizeof(__half), rows));
  CK(cudaMemset2D(d, pitch, 0x3c, (size_t)W * 8, rows));
  cudaResourceDesc rd = {}; rd.resType = cudaResourceTypePitch2D;
  rd.res.pitch2D.devPtr = d; rd.res.pitch2D.desc = cudaCreateChannelDescHalf4();
  rd.res.pitch2D.width = W; rd.res.pitch2D.height = rows; rd.res.pitch2D.pitchInBytes = pitch;
  cudaTextureDesc td = {}; td.addressMode[0] = td.addressMode[1] = cudaAddressModeBorder;
  td.filterMode = cudaFilterModeLinear; td.readMode = cudaReadModeElementType; td.normalizedCoords = 0;
  cudaTextureObject_t tex; CK(cudaCreateTextureObject(&tex, &rd, &td, nullptr));
  float *off, *out; uint4* xh;
  CK(cudaMalloc(&off, (size_t)18 * H * W * 4)); CK(cudaMalloc(&out, (size_t)H * W * 16 * 4)); CK(cudaMalloc(&xh, (size_t)H * W * 128));
  CK(cudaMemset(xh, 0x11, (size_t)H * W * 128));
  { float* h = (float*)malloc((size_t)18 * H * W * 4); unsigned s = 1;
    for (size_t i = 0; i < (size_t)18 * H * W; ++i) { s = s * 1664525u + 1013904223u; h[i] = ((s >> 8) / 16777216.f - 0.5f) * 4.f; }
    CK(cudaMemcpy(off, h, (size_t)18 * H * W * 4, cudaMemcpyHostToDevice)); free(h); }
  cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
  int sms = 148, clk = 0; cudaDeviceGetAttribute(&clk, cudaDevAttrClockRate, 0);
  for (int it = 0; it < 2; ++it) {
    const int n = H * W * 16, nb = (n + 255) / 256;
    // many images' worth of work: launch the same map 64 times in one grid by making the grid bigger via reps
    cudaEventRecord(e0);
    for (int k = 0; k < 32; ++k) tex_kernel<<<nb, 256>>>(tex, off, out, H, W, planes, reps);
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double fetches = 32.0 * (double)n * 9 * reps;
    if (it) printf("tex  : %.3f ms, %.1f G fetches/s = %.2f fetches/clk/SM (RGBA16F bilinear), %.1f G channel-samples/s\n", ms,
                   fetches / ms / 1e6, fetches / (ms * 1e-3) / sms / (clk * 1e3), 4 * fetches / ms / 1e6);
  }
  for (int it = 0; it < 2; ++it) {
    const int n = H * W * 8, nb = (n + 255) / 256;
    cudaEventRecord(e0);
    for (int k = 0; k < 32; ++k) ldg_kernel<<<nb, 256>>>(xh, off, out, H, W, reps);
    cudaEventRecord(e1); CK(cudaEventSynchronize(e1));
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    const double taps = 32.0 * (double)H * W * 9 * reps;     // (pixel, tap) samples of 64 channels
    if (it) printf("ldg  : %.3f ms, %.1f G (pixel,tap)x64ch/s = %.1f G channel-samples/s (4 corners x LDG.128, no blend)\n", ms,
                   taps / ms / 1e6, 64 * taps / ms / 1e6);
  }
  printf("clock %d kHz\n", clk);
  return 0;
}
