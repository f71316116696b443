// TMA streaming probe (developer tool): how fast can 148 persistent CTAs stream a row-major [N, K] byte
// matrix from HBM with 2-D tiled TMA boxes of `rows` x `inner` bytes, `stages` deep, no compute?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/tma_probe tools/tma_probe.cu && /tmp/tma_probe
// Answers the question behind gemm_skinny.cu's tile shape: is a [128 x 128 B] box at a K-byte row pitch
// (one 128-byte line per DRAM page visit) able to reach the HBM roofline, and what does?
#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("CUDA error %s at %d\n", cudaGetErrorString(e), __LINE__); exit(1); } } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint64_t* b, uint32_t c) { asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(b)), "r"(c) : "memory"); }
__device__ __forceinline__ void mbar_expect(uint64_t* b, uint32_t bytes) { asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(b)), "r"(bytes) : "memory"); }
__device__ __forceinline__ void mbar_wait(uint64_t* b, uint32_t ph) {
  uint32_t ok = 0;
  while (!ok) asm volatile("{\n\t.reg .pred p;\n\tmbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.u32 %0, 1, 0, p;\n\t}\n" : "=r"(ok) : "r"(smem_u32(b)), "r"(ph) : "memory");
}
__device__ __forceinline__ void tma2d(void* dst, const CUtensorMap* m, int c0, int c1, uint64_t* bar) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(smem_u32(dst)), "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1) : "memory");
}
__device__ __forceinline__ void bulk1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)), "l"(src), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}

// mode 0: tiled 2-D boxes; tiles laid out tile-major over (row-tile, k-block) like gemm_skinny
// mode 1: 1-D bulk copies of `inner` contiguous bytes per row, `rows` rows per stage (one elected thread per row)
__global__ void __launch_bounds__(128, 1) probe(const __grid_constant__ CUtensorMap tm, const uint8_t* base, long long N, long long K,
                                                int rows, int inner, int stages, int mode, unsigned long long* sink) {
  extern __shared__ __align__(1024) uint8_t smem[];
  const int stage_bytes = rows * inner;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + (size_t)stages * stage_bytes);
  const long long kb_total = K / inner, tiles = N / rows, U = tiles * kb_total;
  const long long u0 = blockIdx.x * U / gridDim.x, u1 = (blockIdx.x + 1) * U / gridDim.x;
  if (threadIdx.x == 0) {
    for (int s = 0; s < stages; ++s) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (mode >= 10) {
    // nw issuing warps, warp w owns units u0 + w, u0 + w + nw, ... and stages w, w + nw, ... (32-bit counters only)
    const int nw = mode - 10 + 1, w = threadIdx.x >> 5;
    if (w < nw && (threadIdx.x & 31) == 0) {
      const int n = (int)(u1 - u0), depth = stages / nw;  // ops in flight per warp
      int tile = (int)(u0 / kb_total), kb = (int)(u0 % kb_total);
      for (int i = 0; i < w; ++i) { if (++kb == kb_total) { kb = 0; ++tile; } }
      int wt = tile, wk = kb;  // position of the next op to ISSUE
      int issued = w, waited = w;
      unsigned long long acc = 0;
      while (waited < n) {
        while (issued < n && (issued - waited) / nw < depth) {
          const int s = issued % (depth * nw);
          mbar_expect(&full[s], stage_bytes);
          tma2d(smem + (size_t)s * stage_bytes, &tm, wk * inner, wt * rows, &full[s]);
          issued += nw;
          for (int i = 0; i < nw; ++i) { if (++wk == kb_total) { wk = 0; ++wt; } }
        }
        const int s = waited % (depth * nw);
        mbar_wait(&full[s], (uint32_t)((waited / (depth * nw)) & 1));
        acc += smem[(size_t)s * stage_bytes];
        waited += nw;
      }
      if (acc == 0x12345678abcdefULL) sink[0] = acc;
    }
    return;
  }
  if (threadIdx.x < 32) {
    // single warp: issue ahead `stages`, wait in order, re-issue (no consumer: the wait IS the consumption)
    long long issued = u0, waited = u0;
    unsigned long long acc = 0;
    while (waited < u1) {
      while (issued < u1 && issued - waited < stages) {
        const int s = (int)((issued - u0) % stages);
        const long long tile = issued / kb_total, kb = issued % kb_total;
        if (mode == 0) {
          if (threadIdx.x == 0) {
            mbar_expect(&full[s], stage_bytes);
            tma2d(smem + (size_t)s * stage_bytes, &tm, (int)(kb * inner), (int)(tile * rows), &full[s]);
          }
        } else {
          if (threadIdx.x == 0) mbar_expect(&full[s], stage_bytes);
          __syncwarp();
          for (int r = threadIdx.x; r < rows; r += 32)
            bulk1d(smem + (size_t)s * stage_bytes + (size_t)r * inner, base + (tile * rows + r) * K + kb * inner, inner, &full[s]);
        }
        ++issued;
      }
      const int s = (int)((waited - u0) % stages);
      mbar_wait(&full[s], (uint32_t)(((waited - u0) / stages) & 1));
      acc += smem[(size_t)s * stage_bytes + threadIdx.x];
      __syncwarp();
      ++waited;
    }
    if (acc == 0x12345678abcdefULL) sink[0] = acc;
  }
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main() {
  void* fp = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fp, cudaEnableDefault, &q));
  PFN_enc enc = (PFN_enc)fp;
  struct Shape { long long N, K; } shapes[] = {{28672, 4096}, {4096, 14336}, {6144, 4096}};
  unsigned long long* sink;
  CK(cudaMalloc(&sink, 8));
  CK(cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  for (auto sh : shapes) {
    uint8_t* w[16];
    int nbuf = (int)((400ll << 20) / (sh.N * sh.K)) + 1;
    if (nbuf < 3) nbuf = 3;
    if (nbuf > 16) nbuf = 16;
    for (int i = 0; i < nbuf; ++i) { CK(cudaMalloc(&w[i], sh.N * sh.K)); CK(cudaMemset(w[i], i + 1, sh.N * sh.K)); }
    struct Cfg { int rows, inner, mode, sw, promo; } cfgs[] = {
        {128, 128, 0, 1, 2}, {128, 128, 0, 1, 0}, {128, 128, 0, 1, 1}, {256, 128, 0, 1, 2}, {64, 128, 0, 1, 2},
        {128, 128, 10, 1, 2}, {128, 128, 11, 1, 2}, {128, 128, 13, 1, 2}, {64, 128, 10, 1, 2}, {64, 128, 11, 1, 2}, {64, 128, 13, 1, 2},
        {256, 128, 10, 1, 2}, {256, 128, 11, 1, 2}, {32, 128, 13, 1, 2}};
    for (auto c : cfgs) {
      if (sh.K % c.inner || sh.N % c.rows) continue;
      const int stage_bytes = c.rows * c.inner;
      int stages = (220 * 1024) / stage_bytes;
      if (stages > 24) stages = 24;
      if (c.mode >= 10) stages = stages / (c.mode - 9) * (c.mode - 9);
      CUtensorMap tm[16];
      for (int i = 0; i < nbuf; ++i) {
        cuuint64_t dims[2] = {(cuuint64_t)sh.K, (cuuint64_t)sh.N};
        cuuint64_t strides[1] = {(cuuint64_t)sh.K};
        cuuint32_t box[2] = {(cuuint32_t)c.inner, (cuuint32_t)c.rows};
        cuuint32_t es[2] = {1, 1};
        CUresult r = enc(&tm[i], CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, w[i], dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                         c.sw ? CU_TENSOR_MAP_SWIZZLE_128B : CU_TENSOR_MAP_SWIZZLE_NONE,
                         c.promo == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_256B : c.promo == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B : CU_TENSOR_MAP_L2_PROMOTION_NONE,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        if (r != CUDA_SUCCESS) { printf("encode failed %d for rows %d inner %d\n", (int)r, c.rows, c.inner); goto next; }
      }
      {
        const int smem = stages * stage_bytes + 1024;
        cudaEvent_t e0, e1;
        CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
        for (int i = 0; i < nbuf; ++i) probe<<<148, 128, smem>>>(tm[i], w[i], sh.N, sh.K, c.rows, c.inner, stages, c.mode, sink);
        CK(cudaDeviceSynchronize());
        const int reps = 2 * nbuf;
        CK(cudaEventRecord(e0));
        for (int i = 0; i < reps; ++i) probe<<<148, 128, smem>>>(tm[i % nbuf], w[i % nbuf], sh.N, sh.K, c.rows, c.inner, stages, c.mode, sink);
        CK(cudaEventRecord(e1));
        CK(cudaDeviceSynchronize());
        float ms;
        CK(cudaEventElapsedTime(&ms, e0, e1));
        ms /= reps;
        printf("N %6lld K %6lld  %s rows %4d inner %5d B  stages %2d promo %d: %8.1f us  %7.1f GB/s\n", sh.N, sh.K,
               c.mode == 0 ? (c.sw ? "tiled-sw128" : "tiled-none ") : c.mode == 1 ? "bulk-1d    " : c.mode == 10 ? "tiled 1warp" : c.mode == 11 ? "tiled 2warp" : "tiled 4warp", c.rows, c.inner, stages, c.promo, ms * 1e3,
               sh.N * sh.K / ms / 1e6);
      }
    next:;
    }
    for (int i = 0; i < nbuf; ++i) CK(cudaFree(w[i]));
  }
  return 0;
}
