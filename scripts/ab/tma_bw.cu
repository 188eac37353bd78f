// Microbenchmark (development tool): achievable HBM read bandwidth of a persistent ring of 1-D bulk
// async copies (cp.async.bulk global->shared, mbarrier completion) as a function of chunk size,
// ring depth and CTAs per SM.  nvcc -arch=sm_100a -o tma_bw tma_bw.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" :: "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile("{\n\t.reg .pred p;\n\tW_%=:\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t@p bra D_%=;\n\tbra W_%=;\n\tD_%=:\n\t}"
               :: "r"((uint32_t)__cvta_generic_to_shared(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_g2s(void* dst, const void* src, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               :: "r"((uint32_t)__cvta_generic_to_shared(dst)), "l"(src), "r"(bytes), "r"((uint32_t)__cvta_generic_to_shared(bar)) : "memory");
}
template <int S>
__global__ void ring(const float* __restrict__ src, size_t n_chunks, unsigned chunk, int split, int consume, float* out) {
  extern __shared__ __align__(128) unsigned char smem[];
  __shared__ __align__(8) uint64_t full[S];
  const int tid = threadIdx.x, G = gridDim.x;
  auto issue = [&](long long k) {
    const long long i = (long long)blockIdx.x + k * G;
    if (i >= (long long)n_chunks) return;
    const int s = (int)(k % S);
    mbar_expect_tx(&full[s], chunk);
    const unsigned part = chunk / split;
    for (int q = 0; q < split; ++q)
      bulk_g2s(smem + (size_t)s * chunk + (size_t)q * part, (const unsigned char*)src + (size_t)i * chunk + (size_t)q * part, part, &full[s]);
  };
  if (tid == 0) { for (int s = 0; s < S; ++s) mbar_init(&full[s], 1); asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory"); for (int k = 0; k < S; ++k) issue(k); }
  __syncthreads();
  float acc = 0.f;
  long long k = 0;
  for (long long i = blockIdx.x; i < (long long)n_chunks; i += G, ++k) {
    const int s = (int)(k % S);
    mbar_wait(&full[s], (unsigned)((k / S) & 1));
    if (consume) {
      const float4* p = reinterpret_cast<const float4*>(smem + (size_t)s * chunk);
      for (unsigned j = tid; j < chunk / 16; j += blockDim.x) { const float4 q = p[j]; acc += fmaxf(fmaxf(q.x, q.y), fmaxf(q.z, q.w)); }
    }
    __syncthreads();
    if (tid == 0) issue(k + S);
  }
  if (acc == 12345.678f) out[0] = acc;
}
template <int S>
float run(const float* d, size_t bytes, unsigned chunk, int occ, int threads, int split, int consume, float* out) {
  size_t smem = (size_t)S * chunk;
  cudaFuncSetAttribute(ring<S>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  int o = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&o, ring<S>, threads, smem);
  if (o < occ) return -1.f;
  size_t n = bytes / chunk;
  cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
  float best = 1e9f;
  for (int it = 0; it < 5; ++it) {
    cudaEventRecord(a);
    ring<S><<<148 * occ, threads, smem>>>(d, n, chunk, split, consume, out);
    cudaEventRecord(b); cudaEventSynchronize(b);
    float ms; cudaEventElapsedTime(&ms, a, b); if (ms < best) best = ms;
  }
  return (float)(n * (size_t)chunk) / best / 1e6f;   // GB/s
}
int main() {
  const size_t bytes = 1ull << 30;   // 1 GiB (larger than L2)
  float *d, *out; cudaMalloc(&d, bytes); cudaMalloc(&out, 4); cudaMemset(d, 0, bytes);
  unsigned chunks[] = {4096, 8192, 16384, 17408, 32768, 65536};
  printf("chunk stages occ threads split consume GB/s\n");
  for (unsigned c : chunks) for (int occ : {1, 2, 3, 4}) for (int split : {1, 4}) {
    float r2 = run<2>(d, bytes, c, occ, 256, split, 1, out);
    float r3 = run<3>(d, bytes, c, occ, 256, split, 1, out);
    float r4 = run<4>(d, bytes, c, occ, 256, split, 1, out);
    float r6 = run<6>(d, bytes, c, occ, 256, split, 1, out);
    printf("%6u occ=%d split=%d  S2 %7.0f  S3 %7.0f  S4 %7.0f  S6 %7.0f\n", c, occ, split, r2, r3, r4, r6);
  }
  cudaError_t e = cudaDeviceSynchronize(); printf("%s\n", cudaGetErrorString(e));
  return 0;
}
