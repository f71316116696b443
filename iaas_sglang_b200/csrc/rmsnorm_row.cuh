// One token row of fused-add + RMSNorm (+ group-128 FP8 quantisation) on a 256-thread CTA -- shared by
// rmsnorm_quant_fp8_kernel (fused_elem.cu) and the all-reduce epilogue (allreduce.cu), so that
// all-reduce -> rmsnorm_quant_fp8 and the fused all-reduce produce the same bits (same rounding points, same
// summation order).  References: sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu,
// python/sglang/srt/layers/layernorm.py, sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:25-115.
#pragma once
#include "common.cuh"

namespace b200 {

// quantise 8 values owned by one of the 16 threads of a 128-group; returns the packed fp8 bytes, writes the scale
__device__ __forceinline__ uint2 quant_group8(const float (&v)[8], float eps, float fmin_, float fmax_, float* s_out,
                                              bool write_scale) {
  float amax = eps;
#pragma unroll
  for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(v[i]));
  // only the 16 lanes of this group are named: the other half of the warp may have left (H or M*d an odd multiple of 128)
  const unsigned half_mask = 0xffffu << (threadIdx.x & 16);
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(half_mask, amax, o, 16));
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

struct RowNormArgs {
  uint16_t* residual;   // [M, H] in/out or nullptr
  const uint16_t* w;    // [H]
  uint16_t* out;        // [M, H] normalised half result or nullptr
  uint8_t* q;           // [M, H] e4m3 (row stride q_sm) or nullptr
  int64_t q_sm;
  float* s;             // scales s[m * s_sm + g * s_sg]
  int64_t s_sm, s_sg;
  int H;
  float eps, q_eps, fmin_, fmax_;
};

// Row m: thread t of 256 owns the 8-element vectors t, t + 256, ... (NV of them); load_x(vec) returns vector `vec`
// of the row's input x as 8 packed halves.  The row stays in registers between the sum of squares and the
// quantisation.  H % 128 == 0, H <= 256 * 8 * NV.  `red` = 8 floats of shared memory; every thread of the CTA must call.
template <bool BF16, int NV, class LoadX>
__device__ __forceinline__ void rmsnorm_quant_row(const RowNormArgs& a, int64_t m, LoadX load_x, float* red) {
  const int H = a.H;
  const int nvec = H >> 3;
  float v[NV][8];
  float ss = 0.f;
  // the norm weight is requested together with the row (it is only needed after the reduction, but a load issued there
  // puts one more L2 round trip on a kernel that is nothing but a latency chain)
  // (rows of up to 4096 elements: 7.2 -> 5.7 us at M = 512; wider rows keep the late load -- sixteen more live registers
  // cost them more than the round trip, 6.5 -> 7.0 us at 256 x 8192)
  constexpr bool W_EARLY = NV <= 2;
  uint4 wr[NV];
  if constexpr (W_EARLY) {
#pragma unroll
    for (int j = 0; j < NV; ++j) {
      const int vec = threadIdx.x + j * 256;
      wr[j] = vec < nvec ? *reinterpret_cast<const uint4*>(a.w + vec * 8) : make_uint4(0u, 0u, 0u, 0u);
    }
  }
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int vec = threadIdx.x + j * 256;
    if (vec < nvec) {
      const uint4 xr = load_x(vec);
      const uint32_t xw[4] = {xr.x, xr.y, xr.z, xr.w};
      if (a.residual) {
        const uint4 rr = *reinterpret_cast<const uint4*>(a.residual + m * H + vec * 8);
        const uint32_t rw[4] = {rr.x, rr.y, rr.z, rr.w};
        uint32_t ow[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
          const float2 x2 = unpack_half2<BF16>(xw[i]), b = unpack_half2<BF16>(rw[i]);
          ow[i] = pack_half2<BF16>(x2.x + b.x, x2.y + b.y);
          const float2 f = unpack_half2<BF16>(ow[i]);
          v[j][2 * i] = f.x;
          v[j][2 * i + 1] = f.y;
        }
        *reinterpret_cast<uint4*>(a.residual + m * H + vec * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
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
  // (b200_rmsnorm runs this same code for aligned rows, so the fused and the two-step path share the summation order)
  ss = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) ss += red[i];
  const float r = rsqrtf(ss / (float)H + a.eps);
#pragma unroll
  for (int j = 0; j < NV; ++j) {
    const int vec = threadIdx.x + j * 256;
    if (vec < nvec) {
      if constexpr (!W_EARLY) wr[j] = *reinterpret_cast<const uint4*>(a.w + vec * 8);
      const uint32_t ww[4] = {wr[j].x, wr[j].y, wr[j].z, wr[j].w};
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
      if (a.out) *reinterpret_cast<uint4*>(a.out + m * H + vec * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
      if (a.q) {  // (uniform over the grid)
        const int g = vec >> 4;
        const uint2 pk = quant_group8(y, a.q_eps, a.fmin_, a.fmax_, a.s + m * a.s_sm + g * a.s_sg, (vec & 15) == 0);
        *reinterpret_cast<uint2*>(a.q + m * a.q_sm + vec * 8) = pk;
      }
    }
  }
  __syncthreads();  // `red` may be reused for the CTA's next row
}

}  // namespace b200
