// AWQ INT4 W4A16 linear: python/sglang/srt/layers/quantization/awq.py:186-204
//   out = x @ awq_dequantize(qweight, scales, qzeros) (+ bias)
// Decode shapes (M <= 256, K % 64 == 0, N % 32 == 0): ONE fused kernel, gemm_skinny_kernel<MODE 2> (gemm_skinny.cu):
// the packed int4 tile is dequantised in registers into the tcgen05 operand tile, 0.5 B of HBM traffic per weight.
// Other shapes keep the reference's two-phase structure, entirely in our own sm_100a kernels and in the
// layout the tcgen05 GEMM wants: the weight is dequantised TRANSPOSED into a K-major [N, K] half
// buffer in the caller's workspace (so both GEMM operands are K-major / SWIZZLE_128B TMA tiles),
// then gemm_tc_kernel<MODE 2> (tcgen05.mma kind::f16, fp32 accumulate in TMEM) does x . W^T.
// The dequantised values are bit-identical to awq_dequantize (awq_kernel.cu:126-221).
#include <stdlib.h>

#include "gemm_skinny.h"

namespace b200 {

int dense_half_gemm(const void* a, int64_t lda, const void* w, int64_t ldw, const void* bias, void* out, int64_t ldo,
                    int dtype, int M, int N, int K, void* workspace, size_t workspace_bytes, cudaStream_t st);
size_t dense_half_gemm_workspace(int M, int N, int K);
// awq_fused.cu: NT sub-tiles per CTA, separate operand ring (N % 128 == 0)
bool awq_fused_applicable(int M, int N, int K, int group);
int launch_awq_fused(const void* x, int64_t ldx, const int32_t* qweight, const void* scales, const int32_t* qzeros,
                     const void* bias, void* out, int64_t ldo, int dtype, int M, int N, int K, int group, cudaStream_t st);

// block = (32, 8): tile of 64 K-rows x 32 packed columns (= 256 N); out_t[n, k]
template <bool BF16>
__global__ void __launch_bounds__(256) awq_dequant_t_kernel(const int32_t* __restrict__ qw,
                                                            const uint16_t* __restrict__ sc,
                                                            const int32_t* __restrict__ qz, uint16_t* __restrict__ out_t,
                                                            int K, int cols, int group) {
  __shared__ uint16_t tile[256][64 + 8];  // [n_local][k_local], +8 halves padding
  const int c0 = blockIdx.x * 32, k0 = blockIdx.y * 64;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c = c0 + tx;
  for (int kk = ty; kk < 64; kk += 8) {
    const int k = k0 + kk;
    if (k < K && c < cols) {
      const int gi = k / group;
      const uint32_t w = (uint32_t)qw[(int64_t)k * cols + c], z = (uint32_t)qz[(int64_t)gi * cols + c];
      const uint4 sraw = *reinterpret_cast<const uint4*>(sc + ((int64_t)gi * cols + c) * 8);
      const uint32_t sw[4] = {sraw.x, sraw.y, sraw.z, sraw.w};
#pragma unroll
      for (int p = 0; p < 4; ++p) {
        const int d0 = (int)((w >> (4 * p)) & 15u) - (int)((z >> (4 * p)) & 15u);
        const int d1 = (int)((w >> (4 * p + 16)) & 15u) - (int)((z >> (4 * p + 16)) & 15u);
        const float2 s2 = unpack_half2<BF16>(sw[p]);
        tile[tx * 8 + 2 * p][kk] = float_to_half<BF16>(__fmul_rn((float)d0, s2.x));
        tile[tx * 8 + 2 * p + 1][kk] = float_to_half<BF16>(__fmul_rn((float)d1, s2.y));
      }
    }
  }
  __syncthreads();
  // write 256 rows (n) x 64 k: each thread writes 8 halves (16 B) at a time
  for (int e = threadIdx.x; e < 256 * 8; e += 256) {
    const int nl = e >> 3, kc = (e & 7) * 8;
    const int n = c0 * 8 + nl, k = k0 + kc;
    if (n < cols * 8 && k < K) {
      const uint4 v = *reinterpret_cast<const uint4*>(&tile[nl][kc]);
      *reinterpret_cast<uint4*>(out_t + (int64_t)n * K + k) = v;
    }
  }
}

static bool awq_fused_shape(int M, int N, int K) { return skinny_applicable(M) && K % 64 == 0 && N % 32 == 0; }

}  // namespace b200

using namespace b200;

extern "C" int b200_awq_w4a16_mm(const void* x, int64_t ldx, const int32_t* qweight, const void* scales,
                                 const int32_t* qzeros, const void* bias, void* out, int64_t ldo, int dtype, int M, int N,
                                 int K, int group, void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(x && qweight && scales && qzeros && out, B200_ERR_ARG, "awq_w4a16_mm: null tensor");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "awq_w4a16_mm: activations must be f16/bf16");
  B200_REQUIRE(M >= 0 && N > 0 && K > 0 && N % 8 == 0 && K % 8 == 0 && group > 0 && K % group == 0, B200_ERR_ARG,
               "awq_w4a16_mm: need N %% 8 == 0, K %% 8 == 0, K %% group == 0 (N=%d K=%d group=%d)", N, K, group);
  if (M == 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "awq_w4a16_mm: requires an sm_100 device (got sm_%d); no fallback", arch);
  cudaStream_t st = (cudaStream_t)stream;
  if (awq_fused_shape(M, N, K) && group % 8 == 0 && ldx % 8 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)qweight & 15) == 0 &&
      ((uintptr_t)scales & 15) == 0) {
    const bool use_v2 = tune(TUNE_AWQ_V1) == 0;
    if (use_v2 && awq_fused_applicable(M, N, K, group))
      return launch_awq_fused(x, ldx, qweight, scales, qzeros, bias, out, ldo, dtype, M, N, K, group, st);
    SkinnyParams sp{};
    sp.bias = bias; sp.out = out; sp.ldo = ldo; sp.out_bf16 = dtype == B200_BF16;
    sp.qzeros = qzeros; sp.scales = (const uint16_t*)scales; sp.group = group;
    sp.M = M; sp.N = N; sp.K = K;
    return launch_skinny<2>(x, ldx * 2, qweight, (int64_t)N / 2, sp, workspace, workspace_bytes, st);
  }
  const size_t wt_bytes = ((size_t)N * K * 2 + 255) & ~(size_t)255;
  const size_t gemm_ws = (dense_half_gemm_workspace(M, N, K) + 255) & ~(size_t)255;
  B200_REQUIRE(workspace && workspace_bytes >= wt_bytes + gemm_ws && ((uintptr_t)workspace & 255) == 0, B200_ERR_ARG,
               "awq_w4a16_mm: workspace too small or not 256-byte aligned (need %zu bytes, have %zu)", wt_bytes + gemm_ws,
               workspace_bytes);
  // the split-K ticket counters live at the START of the workspace (zeroed once by the caller)
  uint8_t* gemm_ws_ptr = reinterpret_cast<uint8_t*>(workspace);
  uint16_t* wt = reinterpret_cast<uint16_t*>(gemm_ws_ptr + gemm_ws);
  const int cols = N / 8;
  dim3 grid((cols + 31) / 32, (K + 63) / 64);
  if (dtype == B200_BF16)
    awq_dequant_t_kernel<true><<<grid, 256, 0, st>>>(qweight, (const uint16_t*)scales, qzeros, wt, K, cols, group);
  else
    awq_dequant_t_kernel<false><<<grid, 256, 0, st>>>(qweight, (const uint16_t*)scales, qzeros, wt, K, cols, group);
  if (int rc = launched("awq_dequant_t_kernel launch")) return rc;
  return dense_half_gemm(x, ldx, wt, K, bias, out, ldo, dtype, M, N, K, gemm_ws_ptr, gemm_ws, st);
}

extern "C" size_t b200_awq_w4a16_workspace_size(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 256;
  // (the fused path needs none; a call that cannot take it -- unaligned operands, group % 8 != 0 -- is told so by the
  // workspace check of b200_awq_w4a16_mm)
  if (awq_fused_shape(M, N, K)) return (skinny_workspace_bytes(M, N, K) + 255) & ~(size_t)255;
  const size_t wt_bytes = ((size_t)N * K * 2 + 255) & ~(size_t)255;
  return ((dense_half_gemm_workspace(M, N, K) + 255) & ~(size_t)255) + wt_bytes;
}
