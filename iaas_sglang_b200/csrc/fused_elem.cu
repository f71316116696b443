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

namespace b200 {

// quantise 8 values owned by one of the 16 threads of a 128-group; returns the packed fp8 bytes, writes the scale
__device__ __forceinline__ uint2 quant_group8(const float (&v)[8], float eps, float fmin_, float fmax_, float* s_out,
                                              bool write_scale) {
  float amax = eps;
#pragma unroll
  for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(v[i]));
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o, 16));
  const float ys = __fdiv_rn(amax, fmax_);
  if (write_scale) *s_out = ys;
  uint32_t packed[2];
#pragma unroll
  for (int i = 0; i < 2; ++i) {
    uint32_t pk = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float qv = fminf(fmaxf(__fdiv_rn(v[4 * i + j], ys), fmin_), fmax_);
      pk |= (uint32_t)__nv_cvt_float_to_fp8(qv, __NV_SATFINITE, __NV_E4M3) << (8 * j);
    }
    packed[i] = pk;
  }
  return make_uint2(packed[0], packed[1]);
}

// one CTA per token row, 256 threads, thread t owns the 8-element vectors t, t + 256, ... (NV of them): the row
// stays in registers between the sum of squares and the quantisation.  H % 128 == 0, H <= 256 * 8 * NV.
template <bool BF16, int NV>
__global__ void __launch_bounds__(256) rmsnorm_quant_fp8_kernel(const uint16_t* __restrict__ x, uint16_t* __restrict__ residual,
                                                                const uint16_t* __restrict__ w, uint16_t* __restrict__ out,
                                                                uint8_t* __restrict__ q, int64_t q_sm, float* __restrict__ s,
                                                                int64_t s_sm, int64_t s_sg, int H, float eps, float q_eps,
                                                                float fmin_, float fmax_) {
  griddep_wait();    // programmatic dependent launch: everything below reads what the previous kernel wrote
  griddep_launch();  // the next kernel may take its blocks and run its prologue
  const int64_t m = blockIdx.x;
  const int nvec = H >> 3;
  __shared__ float red[8];
  float v[NV][8];
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int vec = threadIdx.x + j * 256;
    if (vec < nvec) {
      const uint4 xr = *reinterpret_cast<const uint4*>(x + m * H + vec * 8);
      const uint32_t xw[4] = {xr.x, xr.y, xr.z, xr.w};
      if (residual) {
        const uint4 rr = *reinterpret_cast<const uint4*>(residual + m * H + vec * 8);
        const uint32_t rw[4] = {rr.x, rr.y, rr.z, rr.w};
        uint32_t ow[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 a = unpack_half2<BF16>(xw[i]), b = unpack_half2<BF16>(rw[i]);
          ow[i] = pack_half2<BF16>(a.x + b.x, a.y + b.y);
          const float2 f = unpack_half2<BF16>(ow[i]);
          v[j][2 * i] = f.x;
          v[j][2 * i + 1] = f.y;
        }
        *reinterpret_cast<uint4*>(residual + m * H + vec * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 f = unpack_half2<BF16>(xw[i]);
          v[j][2 * i] = f.x;
          v[j][2 * i + 1] = f.y;
        }
      }
#pragma unroll
      for (int i = 0; i < 8; ++i) ss += v[j][i] * v[j][i];
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  // (b200_rmsnorm runs this same kernel for aligned rows, so the fused and the two-step path share the summation order)
  ss = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) ss += red[i];
  const float r = rsqrtf(ss / (float)H + eps);
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int vec = threadIdx.x + j * 256;
    if (vec < nvec) {
      const uint4 wr = *reinterpret_cast<const uint4*>(w + vec * 8);
      const uint32_t ww[4] = {wr.x, wr.y, wr.z, wr.w};
      float y[8];
      uint32_t ow[4];
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float2 wf = unpack_half2<BF16>(ww[i]);
        const float n0 = half_to_float<BF16>(float_to_half<BF16>(v[j][2 * i] * r));
        const float n1 = half_to_float<BF16>(float_to_half<BF16>(v[j][2 * i + 1] * r));
        ow[i] = pack_half2<BF16>(n0 * wf.x, n1 * wf.y);
        const float2 f = unpack_half2<BF16>(ow[i]);
        y[2 * i] = f.x;
        y[2 * i + 1] = f.y;
      }
      if (out) *reinterpret_cast<uint4*>(out + m * H + vec * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      if (q) {  // (uniform over the grid; b200_rmsnorm's fast path runs this kernel without the quantiser)
        const int g = vec >> 4;
        const uint2 pk = quant_group8(y, q_eps, fmin_, fmax_, s + m * s_sm + g * s_sg, (vec & 15) == 0);
        *reinterpret_cast<uint2*>(q + m * q_sm + vec * 8) = pk;
      }
    }
  }
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
#define B200_RQ(BF, NV)                                                                                                   \
  launch_pdl(rmsnorm_quant_fp8_kernel<BF, NV>, dim3(M), dim3(256), 0, st, (const uint16_t*)x, (uint16_t*)residual, (const uint16_t*)weight, \
                                                      (uint16_t*)out, (uint8_t*)q, q_sm, s, s_sm, s_sg, H, eps, q_eps,   \
                                                      fp8_min, fp8_max)
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
