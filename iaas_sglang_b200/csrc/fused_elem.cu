// Steps either side of the quantised linears, fused: the producer of an FP8 linear's input also
// writes the per-token-group quantisation, so the half-precision activation never makes an HBM round trip.
//
//   b200_rmsnorm_quant_fp8      = fused_add_rmsnorm (sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu,
//                                 python/sglang/srt/layers/layernorm.py) + sgl_per_token_group_quant_fp8
//                                 (sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:25-115)
//   b200_silu_mul_quant_fp8     = silu_and_mul (sgl-kernel/csrc/elementwise/activation.cu) + the same quantiser
// Both round exactly where the separate kernels round (the normalised / activated value goes through the 16-bit
// type before it is quantised), so q and the scales are BIT-IDENTICAL to rmsnorm -> per_token_group_quant_fp8 and
// silu_and_mul -> per_token_group_quant_fp8 (tests/test_gpu_quant_gemm.py checks that).
#include "common.cuh"
#include "rmsnorm_row.cuh"

namespace b200 {

// one CTA per token row, 256 threads (rmsnorm_row.cuh)
template <bool BF16, int NV>
__global__ void __launch_bounds__(256) rmsnorm_quant_fp8_kernel(const uint16_t* __restrict__ x, const RowNormArgs a) {
  griddep_wait();    // programmatic dependent launch: everything below reads what the previous kernel wrote
  griddep_launch();  // the next kernel may take its blocks and run its prologue
  const int64_t m = blockIdx.x;
  __shared__ float red[8];
  const uint16_t* xrow = x + m * a.H;
  rmsnorm_quant_row<BF16, NV>(a, m, [&](int vec) { return *reinterpret_cast<const uint4*>(xrow + vec * 8); }, red);
}

// x [M, 2d] = (gate | up); one thread per 8 outputs, 16 threads per 128-group.  d % 128 == 0.
template <bool BF16>
__global__ void __launch_bounds__(256) silu_mul_quant_fp8_kernel(const uint16_t* __restrict__ x, uint8_t* __restrict__ q,
                                                                 int64_t q_sm, float* __restrict__ s, int64_t s_sm,
                                                                 int64_t s_sg, int64_t M, int d, float q_eps, float fmin_,
                                                                 float fmax_) {
  griddep_wait();    // programmatic dependent launch: everything below reads what the previous kernel wrote
  griddep_launch();  // the next kernel may take its blocks and run its prologue
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  const int dv = d / 8;
  if (i >= M * dv) return;  // (whole 16-thread groups leave together: dv % 16 == 0)
  const int64_t m = i / dv;
  const int c = (int)(i % dv);
  const uint4 ga = *reinterpret_cast<const uint4*>(x + m * 2 * d + c * 8);
  const uint4 ua = *reinterpret_cast<const uint4*>(x + m * 2 * d + d + c * 8);
  const uint32_t gw[4] = {ga.x, ga.y, ga.z, ga.w}, uw[4] = {ua.x, ua.y, ua.z, ua.w};
  float y[8];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const float2 g = unpack_half2<BF16>(gw[p]), u = unpack_half2<BF16>(uw[p]);
    const float s0 = half_to_float<BF16>(float_to_half<BF16>(g.x / (1.f + expf(-g.x))));
    const float s1 = half_to_float<BF16>(float_to_half<BF16>(g.y / (1.f + expf(-g.y))));
    const float2 f = unpack_half2<BF16>(pack_half2<BF16>(s0 * u.x, s1 * u.y));
    y[2 * p] = f.x;
    y[2 * p + 1] = f.y;
  }
  const uint2 pk = quant_group8(y, q_eps, fmin_, fmax_, s + m * s_sm + (c >> 4) * s_sg, (c & 15) == 0);
  *reinterpret_cast<uint2*>(q + m * q_sm + c * 8) = pk;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_rmsnorm_quant_fp8(const void* x, void* residual, const void* weight, void* out, void* q, int64_t q_sm,
                                      float* s, int64_t s_sm, int64_t s_sg, int dtype, int M, int H, float eps, float q_eps,
                                      float fp8_min, float fp8_max, void* stream) {
  B200_REQUIRE(x && weight && ((q && s) || (!q && out)) && M >= 0 && H > 0, B200_ERR_ARG, "rmsnorm_quant_fp8: bad argument");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "rmsnorm_quant_fp8: dtype");
  B200_REQUIRE(H % 128 == 0 && H <= 8192, B200_ERR_UNSUPPORTED, "rmsnorm_quant_fp8: H must be a multiple of 128, at most 8192 (got %d)", H);
  B200_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)weight & 15) == 0 && ((uintptr_t)q & 7) == 0 && q_sm % 8 == 0 &&
                   (!residual || ((uintptr_t)residual & 15) == 0) && (!out || ((uintptr_t)out & 15) == 0),
               B200_ERR_ARG, "rmsnorm_quant_fp8: operands must be 16-byte aligned");
  if (M == 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "rmsnorm_quant_fp8: requires an sm_100 device (got sm_%d); no fallback", arch);
  cudaStream_t st = (cudaStream_t)stream;
  const int nv = (H / 8 + 255) / 256;
  RowNormArgs ra;
  ra.residual = (uint16_t*)residual; ra.w = (const uint16_t*)weight; ra.out = (uint16_t*)out; ra.q = (uint8_t*)q; ra.q_sm = q_sm;
  ra.s = s; ra.s_sm = s_sm; ra.s_sg = s_sg; ra.H = H; ra.eps = eps; ra.q_eps = q_eps; ra.fmin_ = fp8_min; ra.fmax_ = fp8_max;
#define B200_RQ(BF, NV) launch_pdl(rmsnorm_quant_fp8_kernel<BF, NV>, dim3(M), dim3(256), 0, st, (const uint16_t*)x, ra)
  if (dtype == B200_BF16) {
    if (nv <= 1) B200_RQ(true, 1); else if (nv <= 2) B200_RQ(true, 2); else B200_RQ(true, 4);
  } else {
    if (nv <= 1) B200_RQ(false, 1); else if (nv <= 2) B200_RQ(false, 2); else B200_RQ(false, 4);
  }
#undef B200_RQ
  return launched("rmsnorm_quant_fp8_kernel launch");
}

extern "C" int b200_silu_mul_quant_fp8(const void* x, void* q, int64_t q_sm, float* s, int64_t s_sm, int64_t s_sg, int dtype,
                                       int M, int d, float q_eps, float fp8_min, float fp8_max, void* stream) {
  B200_REQUIRE(x && q && s && M >= 0 && d > 0, B200_ERR_ARG, "silu_mul_quant_fp8: bad argument");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "silu_mul_quant_fp8: dtype");
  B200_REQUIRE(d % 128 == 0, B200_ERR_UNSUPPORTED, "silu_mul_quant_fp8: d must be a multiple of 128 (got %d)", d);
  B200_REQUIRE(((uintptr_t)x & 15) == 0 && ((uintptr_t)q & 7) == 0 && q_sm % 8 == 0, B200_ERR_ARG,
               "silu_mul_quant_fp8: operands must be 16-byte aligned");
  if (M == 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "silu_mul_quant_fp8: requires an sm_100 device (got sm_%d); no fallback", arch);
  const int64_t n = (int64_t)M * (d / 8);
  const unsigned blocks = (unsigned)((n + 255) / 256);
  if (dtype == B200_BF16)
    launch_pdl(silu_mul_quant_fp8_kernel<true>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const uint16_t*)x, (uint8_t*)q, q_sm, s, s_sm, s_sg, M, d, q_eps, fp8_min, fp8_max);
  else
    launch_pdl(silu_mul_quant_fp8_kernel<false>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const uint16_t*)x, (uint8_t*)q, q_sm, s, s_sm, s_sg, M, d, q_eps, fp8_min, fp8_max);
  return launched("silu_mul_quant_fp8_kernel launch");
}
