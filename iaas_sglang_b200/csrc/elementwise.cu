// HBM-bound helper kernels of the decode path: page-table index builders (bit-exact integer work),
// activation quantisation (FP8 group-128, INT8 per-token), AWQ dequantisation, and the steps either
// side of the path (RMSNorm, SiLU-and-mul, RoPE).  All are coalesced 16-byte-vector SIMT kernels.
#include <math.h>

#include "common.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// python/sglang/srt/layers/attention/utils.py:5-41 create_flashinfer_kv_indices_triton
// grid = (B, chunks): each block copies up to 4096 entries of one request's req_to_token row.
// ------------------------------------------------------------------------------------------------
__global__ void kv_indices_kernel(const void* __restrict__ r2t, int idx64, int64_t r2t_stride,
                                  const int64_t* __restrict__ rpi, const int64_t* __restrict__ lens,
                                  const int64_t* __restrict__ indptr, const int64_t* __restrict__ kv_start,
                                  int32_t* __restrict__ out) {
  const int b = blockIdx.x;
  const int64_t len = lens[b], start = kv_start ? kv_start[b] : 0, off = indptr[b], req = rpi[b];
  for (int64_t j = (int64_t)blockIdx.y * blockDim.x + threadIdx.x; j < len; j += (int64_t)gridDim.y * blockDim.x)
    out[off + j] = (int32_t)load_index(r2t, req * r2t_stride + start + j, idx64);
}

// utils.py:44-94 create_flashmla_kv_indices_triton
__global__ void page_table_kernel(const void* __restrict__ r2t, int idx64, int64_t r2t_stride,
                                  const int64_t* __restrict__ rpi, const int64_t* __restrict__ lens,
                                  const int64_t* __restrict__ kv_start, int32_t* __restrict__ out, int64_t out_stride,
                                  int page_size, int max_pages) {
  const int b = blockIdx.x;
  const int64_t len = lens[b], start = kv_start ? kv_start[b] : 0, req = rpi[b];
  int64_t np = (len + page_size - 1) / page_size;
  if (np > max_pages) np = max_pages;
  for (int64_t pg = threadIdx.x; pg < np; pg += blockDim.x)
    out[b * out_stride + pg] = (int32_t)(load_index(r2t, req * r2t_stride + start + pg * page_size, idx64) / page_size);
}

// ------------------------------------------------------------------------------------------------
// sgl_per_token_group_quant_fp8 (per_token_group_quant_8bit.cu:25-115): 16 threads per 128-group,
// each thread owns 8 contiguous elements (one 16-byte load), 16 groups per 256-thread block.
// ------------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(256) group_quant_fp8_kernel(const uint16_t* __restrict__ x, int64_t x_sm,
                                                              uint8_t* __restrict__ q, int64_t q_sm,
                                                              float* __restrict__ s, int64_t s_sm, int64_t s_sg, int M,
                                                              int groups_per_row, float eps, float fmin_, float fmax_) {
  griddep_wait();    // programmatic dependent launch: everything below reads what the previous kernel wrote
  griddep_launch();  // the next kernel may take its blocks and run its prologue
  const int64_t gid = (int64_t)blockIdx.x * 16 + (threadIdx.x >> 4);
  const int sub = threadIdx.x & 15;
  if (gid >= (int64_t)M * groups_per_row) return;
  const int m = (int)(gid / groups_per_row), g = (int)(gid % groups_per_row);
  const uint4 raw = *reinterpret_cast<const uint4*>(x + m * x_sm + g * 128 + sub * 8);
  const uint32_t w[4] = {raw.x, raw.y, raw.z, raw.w};
  float v[8];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = unpack_half2<BF16>(w[i]);
    v[2 * i] = f.x;
    v[2 * i + 1] = f.y;
  }
  float amax = eps;
#pragma unroll
  for (int i = 0; i < 8; ++i) amax = fmaxf(amax, fabsf(v[i]));
#pragma unroll
  for (int o = 8; o > 0; o >>= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffu << (threadIdx.x & 16), amax, o, 16));
  const float ys = __fdiv_rn(amax, fmax_);
  if (sub == 0) s[m * s_sm + g * s_sg] = ys;
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
  *reinterpret_cast<uint2*>(q + m * q_sm + g * 128 + sub * 8) = make_uint2(packed[0], packed[1]);
}

// generic group size (multiple of 8) fallback: one warp per group
template <bool BF16>
__global__ void group_quant_fp8_generic_kernel(const uint16_t* __restrict__ x, int64_t x_sm, uint8_t* __restrict__ q,
                                               int64_t q_sm, float* __restrict__ s, int64_t s_sm, int64_t s_sg, int M,
                                               int groups_per_row, int group, float eps, float fmin_, float fmax_) {
  const int64_t gid = (int64_t)blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (gid >= (int64_t)M * groups_per_row) return;
  const int m = (int)(gid / groups_per_row), g = (int)(gid % groups_per_row);
  const uint16_t* xr = x + m * x_sm + (int64_t)g * group;
  float amax = eps;
  for (int i = lane; i < group; i += 32) amax = fmaxf(amax, fabsf(half_to_float<BF16>(xr[i])));
  amax = warp_max(amax);
  const float ys = __fdiv_rn(amax, fmax_);
  if (lane == 0) s[m * s_sm + g * s_sg] = ys;
  for (int i = lane; i < group; i += 32) {
    const float qv = fminf(fmaxf(__fdiv_rn(half_to_float<BF16>(xr[i]), ys), fmin_), fmax_);
    q[m * q_sm + (int64_t)g * group + i] = (uint8_t)__nv_cvt_float_to_fp8(qv, __NV_SATFINITE, __NV_E4M3);
  }
}

// per_token_quant_int8 (int8_kernel.py:20-81): one 256-thread block per row
template <bool BF16>
__global__ void __launch_bounds__(256) token_quant_int8_kernel(const uint16_t* __restrict__ x, int64_t x_sm,
                                                               int8_t* __restrict__ q, int64_t q_sm,
                                                               float* __restrict__ s, int K) {
  const int m = blockIdx.x;
  const uint16_t* xr = x + m * x_sm;
  __shared__ float red[8];
  float amax = 0.f;
  for (int i = threadIdx.x; i < K; i += 256) amax = fmaxf(amax, fabsf(half_to_float<BF16>(xr[i])));
  amax = warp_max(amax);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = amax;
  __syncthreads();
  amax = red[0];
#pragma unroll
  for (int i = 1; i < 8; ++i) amax = fmaxf(amax, red[i]);
  amax = fmaxf(amax, 1e-10f);
  if (threadIdx.x == 0) s[m] = __fdiv_rn(amax, 127.f);
  const float r = __fdiv_rn(127.f, amax);
  for (int i = threadIdx.x; i < K; i += 256) {
    const float v = roundf(__fmul_rn(half_to_float<BF16>(xr[i]), r));  // round half away from zero (libdevice.round)
    q[m * q_sm + i] = (int8_t)v;
  }
}

// ------------------------------------------------------------------------------------------------
// awq_dequantize (awq_kernel.cu:126-221): one thread per packed int32 -> 8 outputs (one 16-byte store)
// ------------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(256) awq_dequant_kernel(const int32_t* __restrict__ qw, const uint16_t* __restrict__ sc,
                                                          const int32_t* __restrict__ qz, uint16_t* __restrict__ out,
                                                          int rows, int cols, int group) {
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= (int64_t)rows * cols) return;
  const int r = (int)(i / cols), c = (int)(i % cols);
  const int gi = r / group;
  const uint32_t w = (uint32_t)qw[i], z = (uint32_t)qz[(int64_t)gi * cols + c];
  const uint4 sraw = *reinterpret_cast<const uint4*>(sc + ((int64_t)gi * cols + c) * 8);
  const uint32_t sw[4] = {sraw.x, sraw.y, sraw.z, sraw.w};
  uint32_t ow[4];
  // output j takes nibble ord[j], ord = {0,4,1,5,2,6,3,7}: pair (2i, 2i+1) = nibbles (i, i+4)
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const int d0 = (int)((w >> (4 * p)) & 15u) - (int)((z >> (4 * p)) & 15u);
    const int d1 = (int)((w >> (4 * p + 16)) & 15u) - (int)((z >> (4 * p + 16)) & 15u);
    const float2 s2 = unpack_half2<BF16>(sw[p]);
    // (w - z) is exact in half precision; the product is rounded once (mul.rn.f16x2 / __hmul2)
    ow[p] = pack_half2<BF16>(__fmul_rn((float)d0, s2.x), __fmul_rn((float)d1, s2.y));
  }
  *reinterpret_cast<uint4*>(out + i * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
}

// ------------------------------------------------------------------------------------------------
// RMSNorm (+ residual), SiLU-and-mul, neox RoPE
// ------------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(256) rmsnorm_kernel(const uint16_t* __restrict__ x, uint16_t* __restrict__ residual,
                                                      const uint16_t* __restrict__ w, uint16_t* __restrict__ out, int H,
                                                      float eps) {
  griddep_wait();    // programmatic dependent launch: everything below reads what the previous kernel wrote
  griddep_launch();  // the next kernel may take its blocks and run its prologue
  const int64_t m = blockIdx.x;
  extern __shared__ float row[];
  __shared__ float red[8];
  float ss = 0.f;
  for (int i = threadIdx.x; i < H; i += 256) {
    float v = half_to_float<BF16>(x[m * H + i]);
    if (residual) {
      v += half_to_float<BF16>(residual[m * H + i]);
      const uint16_t hv = float_to_half<BF16>(v);
      residual[m * H + i] = hv;
      v = half_to_float<BF16>(hv);
    }
    row[i] = v;
    ss += v * v;
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  ss = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) ss += red[i];
  const float r = rsqrtf(ss / (float)H + eps);
  for (int i = threadIdx.x; i < H; i += 256) {
    const float v = half_to_float<BF16>(float_to_half<BF16>(row[i] * r));
    out[m * H + i] = float_to_half<BF16>(v * half_to_float<BF16>(w[i]));
  }
}

template <bool BF16>
__global__ void __launch_bounds__(256) silu_mul_kernel(const uint16_t* __restrict__ x, uint16_t* __restrict__ out,
                                                       int64_t M, int d) {
  griddep_wait();    // programmatic dependent launch: everything below reads what the previous kernel wrote
  griddep_launch();  // the next kernel may take its blocks and run its prologue
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;  // over M * d/8 vectors
  const int dv = d / 8;
  if (i >= M * dv) return;
  const int64_t m = i / dv;
  const int c = (int)(i % dv);
  const uint4 ga = *reinterpret_cast<const uint4*>(x + m * 2 * d + c * 8);
  const uint4 ua = *reinterpret_cast<const uint4*>(x + m * 2 * d + d + c * 8);
  const uint32_t gw[4] = {ga.x, ga.y, ga.z, ga.w}, uw[4] = {ua.x, ua.y, ua.z, ua.w};
  uint32_t ow[4];
#pragma unroll
  for (int p = 0; p < 4; ++p) {
    const float2 g = unpack_half2<BF16>(gw[p]), u = unpack_half2<BF16>(uw[p]);
    const float s0 = half_to_float<BF16>(float_to_half<BF16>(g.x / (1.f + expf(-g.x))));
    const float s1 = half_to_float<BF16>(float_to_half<BF16>(g.y / (1.f + expf(-g.y))));
    ow[p] = pack_half2<BF16>(s0 * u.x, s1 * u.y);
  }
  *reinterpret_cast<uint4*>(out + m * d + c * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
}

// one CTA per token; a thread owns ONE frequency i (sin/cos evaluated once) and walks the heads with that column
template <bool BF16>
__global__ void rope_neox_kernel(uint16_t* __restrict__ q, int64_t q_sm, uint16_t* __restrict__ k, int64_t k_sm,
                                 const int64_t* __restrict__ pos, int Hq, int Hkv, int D, float log2_theta) {
  griddep_wait();    // programmatic dependent launch: everything below reads what the previous kernel wrote
  griddep_launch();  // the next kernel may take its blocks and run its prologue
  const int m = blockIdx.x;
  const int half = D / 2;
  const float ps = (float)pos[m];
  const int slots = blockDim.x / half;  // heads processed per pass (blockDim.x is a multiple of half, see the launcher)
  if (slots == 0) {  // D / 2 > blockDim.x: element-wise walk
    const int total = (Hq + Hkv) * half;
    for (int e = threadIdx.x; e < total; e += blockDim.x) {
      const int h = e / half, i = e % half;
      uint16_t* base = h < Hq ? q + (int64_t)m * q_sm + h * D : k + (int64_t)m * k_sm + (h - Hq) * D;
      float sn, cs;
      rope_sincos(ps, i, D, log2_theta, sn, cs);
      const float x1 = half_to_float<BF16>(base[i]), x2 = half_to_float<BF16>(base[i + half]);
      base[i] = float_to_half<BF16>(rope_lo(x1, x2, cs, sn));
      base[i + half] = float_to_half<BF16>(rope_hi(x1, x2, cs, sn));
    }
    return;
  }
  const int i = threadIdx.x % half, slot = threadIdx.x / half;
  if (slot >= slots) return;
  float sn, cs;
  rope_sincos(ps, i, D, log2_theta, sn, cs);
  // blockIdx.y splits the heads so that a decode batch (M = 64 tokens) still fills the SMs; two heads per trip keep
  // four independent loads in flight per thread
  const int step = slots * gridDim.y;
  auto row = [&](int h) { return h < Hq ? q + (int64_t)m * q_sm + h * D : k + (int64_t)m * k_sm + (h - Hq) * D; };
  int h = slot + blockIdx.y * slots;
  for (; h + step < Hq + Hkv; h += 2 * step) {
    uint16_t *ba = row(h), *bb = row(h + step);
    const float a1 = half_to_float<BF16>(ba[i]), a2 = half_to_float<BF16>(ba[i + half]);
    const float b1 = half_to_float<BF16>(bb[i]), b2 = half_to_float<BF16>(bb[i + half]);
    ba[i] = float_to_half<BF16>(rope_lo(a1, a2, cs, sn));
    ba[i + half] = float_to_half<BF16>(rope_hi(a1, a2, cs, sn));
    bb[i] = float_to_half<BF16>(rope_lo(b1, b2, cs, sn));
    bb[i + half] = float_to_half<BF16>(rope_hi(b1, b2, cs, sn));
  }
  if (h < Hq + Hkv) {
    uint16_t* base = row(h);
    const float x1 = half_to_float<BF16>(base[i]), x2 = half_to_float<BF16>(base[i + half]);
    base[i] = float_to_half<BF16>(rope_lo(x1, x2, cs, sn));
    base[i + half] = float_to_half<BF16>(rope_hi(x1, x2, cs, sn));
  }
}

}  // namespace b200

using namespace b200;

static int require_sm100(const char* who) {
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "%s: requires an sm_100 device (got sm_%d); no fallback", who,
               arch);
  return 0;
}

extern "C" int b200_create_kv_indices(const void* r2t, int idx64, int64_t r2t_stride, const int64_t* rpi,
                                      const int64_t* lens, const int64_t* indptr, const int64_t* kv_start,
                                      int32_t* out, int B, void* stream) {
  B200_REQUIRE(r2t && rpi && lens && indptr && out && B >= 0, B200_ERR_ARG, "create_kv_indices: bad argument");
  if (B == 0) return B200_OK;
  if (int rc = require_sm100("create_kv_indices")) return rc;
  kv_indices_kernel<<<dim3(B, 8), 512, 0, (cudaStream_t)stream>>>(r2t, idx64, r2t_stride, rpi, lens, indptr, kv_start, out);
  return launched("kv_indices_kernel launch");
}

extern "C" int b200_create_page_table(const void* r2t, int idx64, int64_t r2t_stride, const int64_t* rpi,
                                      const int64_t* lens, const int64_t* kv_start, int32_t* out, int64_t out_stride,
                                      int B, int page_size, int max_pages, void* stream) {
  B200_REQUIRE(r2t && rpi && lens && out && B >= 0 && page_size > 0, B200_ERR_ARG, "create_page_table: bad argument");
  if (B == 0) return B200_OK;
  if (int rc = require_sm100("create_page_table")) return rc;
  page_table_kernel<<<B, 128, 0, (cudaStream_t)stream>>>(r2t, idx64, r2t_stride, rpi, lens, kv_start, out, out_stride,
                                                        page_size, max_pages);
  return launched("page_table_kernel launch");
}

extern "C" int b200_per_token_group_quant_fp8(const void* x, int64_t x_sm, int x_dtype, void* q, int64_t q_sm, float* s,
                                              int64_t s_sm, int64_t s_sg, int M, int K, int group, float eps,
                                              float fp8_min, float fp8_max, void* stream) {
  B200_REQUIRE(x && q && s && M >= 0 && K > 0 && group > 0 && K % group == 0, B200_ERR_ARG,
               "per_token_group_quant_fp8: K (%d) must be a multiple of group (%d)", K, group);
  B200_REQUIRE(x_dtype == B200_BF16 || x_dtype == B200_F16, B200_ERR_UNSUPPORTED, "per_token_group_quant_fp8: input must be bf16/f16");
  if (M == 0) return B200_OK;
  if (int rc = require_sm100("per_token_group_quant_fp8")) return rc;
  cudaStream_t st = (cudaStream_t)stream;
  const int gpr = K / group;
  const int64_t ngroups = (int64_t)M * gpr;
  const bool fast = group == 128 && ((uintptr_t)x & 15) == 0 && x_sm % 8 == 0 && ((uintptr_t)q & 7) == 0 && q_sm % 8 == 0;
  if (fast) {
    const unsigned blocks = (unsigned)((ngroups + 15) / 16);
    if (x_dtype == B200_BF16)
      launch_pdl(group_quant_fp8_kernel<true>, dim3(blocks), dim3(256), 0, st, (const uint16_t*)x, x_sm, (uint8_t*)q, q_sm, s, s_sm, s_sg, M, gpr, eps, fp8_min, fp8_max);
    else
      launch_pdl(group_quant_fp8_kernel<false>, dim3(blocks), dim3(256), 0, st, (const uint16_t*)x, x_sm, (uint8_t*)q, q_sm, s, s_sm, s_sg, M, gpr, eps, fp8_min, fp8_max);
  } else {
    const unsigned blocks = (unsigned)((ngroups + 7) / 8);
    if (x_dtype == B200_BF16)
      group_quant_fp8_generic_kernel<true><<<blocks, 256, 0, st>>>((const uint16_t*)x, x_sm, (uint8_t*)q, q_sm, s, s_sm, s_sg, M, gpr, group, eps, fp8_min, fp8_max);
    else
      group_quant_fp8_generic_kernel<false><<<blocks, 256, 0, st>>>((const uint16_t*)x, x_sm, (uint8_t*)q, q_sm, s, s_sm, s_sg, M, gpr, group, eps, fp8_min, fp8_max);
  }
  return launched("group_quant_fp8_kernel launch");
}

extern "C" int b200_per_token_quant_int8(const void* x, int64_t x_sm, int x_dtype, int8_t* q, int64_t q_sm, float* s, int M,
                                         int K, void* stream) {
  B200_REQUIRE(x && q && s && M >= 0 && K > 0, B200_ERR_ARG, "per_token_quant_int8: bad argument");
  B200_REQUIRE(x_dtype == B200_BF16 || x_dtype == B200_F16, B200_ERR_UNSUPPORTED, "per_token_quant_int8: input must be bf16/f16");
  if (M == 0) return B200_OK;
  if (int rc = require_sm100("per_token_quant_int8")) return rc;
  if (x_dtype == B200_BF16)
    token_quant_int8_kernel<true><<<M, 256, 0, (cudaStream_t)stream>>>((const uint16_t*)x, x_sm, q, q_sm, s, K);
  else
    token_quant_int8_kernel<false><<<M, 256, 0, (cudaStream_t)stream>>>((const uint16_t*)x, x_sm, q, q_sm, s, K);
  return launched("token_quant_int8_kernel launch");
}

extern "C" int b200_awq_dequantize(const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out, int dtype,
                                   int rows, int cols, int group, void* stream) {
  B200_REQUIRE(qweight && scales && qzeros && out && rows > 0 && cols > 0 && group > 0, B200_ERR_ARG, "awq_dequantize: bad argument");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "awq_dequantize: scales must be f16/bf16");
  B200_REQUIRE(((uintptr_t)scales & 15) == 0 && ((uintptr_t)out & 15) == 0, B200_ERR_ARG, "awq_dequantize: scales/out must be 16-byte aligned");
  if (int rc = require_sm100("awq_dequantize")) return rc;
  const int64_t n = (int64_t)rows * cols;
  const unsigned blocks = (unsigned)((n + 255) / 256);
  if (dtype == B200_BF16)
    awq_dequant_kernel<true><<<blocks, 256, 0, (cudaStream_t)stream>>>(qweight, (const uint16_t*)scales, qzeros, (uint16_t*)out, rows, cols, group);
  else
    awq_dequant_kernel<false><<<blocks, 256, 0, (cudaStream_t)stream>>>(qweight, (const uint16_t*)scales, qzeros, (uint16_t*)out, rows, cols, group);
  return launched("awq_dequant_kernel launch");
}

extern "C" int b200_rmsnorm(const void* x, void* residual, const void* weight, void* out, int dtype, int M, int H, float eps,
                            void* stream) {
  B200_REQUIRE(x && weight && out && M >= 0 && H > 0 && H <= 16384, B200_ERR_ARG, "rmsnorm: bad argument");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "rmsnorm: dtype");
  if (M == 0) return B200_OK;
  if (int rc = require_sm100("rmsnorm")) return rc;
  // aligned rows: the register-resident vector kernel of fused_elem.cu without its quantiser
  if (H % 128 == 0 && H <= 8192 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)weight & 15) == 0 && ((uintptr_t)out & 15) == 0 &&
      (!residual || ((uintptr_t)residual & 15) == 0))
    return b200_rmsnorm_quant_fp8(x, residual, weight, out, nullptr, 0, nullptr, 0, 0, dtype, M, H, eps, 0.f, 0.f, 0.f, stream);
  static bool cfg = false;
  if (!cfg) {
    cudaFuncSetAttribute(rmsnorm_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    cudaFuncSetAttribute(rmsnorm_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
    cfg = true;
  }
  if (dtype == B200_BF16)
    launch_pdl(rmsnorm_kernel<true>, dim3(M), dim3(256), H * 4, (cudaStream_t)stream, (const uint16_t*)x, (uint16_t*)residual, (const uint16_t*)weight, (uint16_t*)out, H, eps);
  else
    launch_pdl(rmsnorm_kernel<false>, dim3(M), dim3(256), H * 4, (cudaStream_t)stream, (const uint16_t*)x, (uint16_t*)residual, (const uint16_t*)weight, (uint16_t*)out, H, eps);
  return launched("rmsnorm_kernel launch");
}

extern "C" int b200_silu_and_mul(const void* x, void* out, int dtype, int M, int d, void* stream) {
  B200_REQUIRE(x && out && M >= 0 && d > 0 && d % 8 == 0, B200_ERR_ARG, "silu_and_mul: d must be a multiple of 8");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "silu_and_mul: dtype");
  if (M == 0) return B200_OK;
  if (int rc = require_sm100("silu_and_mul")) return rc;
  const int64_t n = (int64_t)M * (d / 8);
  const unsigned blocks = (unsigned)((n + 255) / 256);
  if (dtype == B200_BF16)
    launch_pdl(silu_mul_kernel<true>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const uint16_t*)x, (uint16_t*)out, M, d);
  else
    launch_pdl(silu_mul_kernel<false>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const uint16_t*)x, (uint16_t*)out, M, d);
  return launched("silu_mul_kernel launch");
}

// cos_sin[m][i] = (cos, sin)(positions[m] * theta^(-2 i / D)), i < D / 2: the values rope_neox_kernel evaluates per layer,
// once per decode step (positions are the same for every layer) for the qkv GEMM's fused RoPE epilogue
__global__ void rope_cos_sin_kernel(const int64_t* __restrict__ pos, int M, int half, float log2_theta, float2* __restrict__ out) {
  griddep_wait();
  griddep_launch();
  const int e = blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= M * half) return;
  const int m = e / half, i = e - m * half;
  float sn, cs;
  rope_sincos((float)pos[m], i, 2 * half, log2_theta, sn, cs);
  out[e] = make_float2(cs, sn);
}

extern "C" int b200_rope_cos_sin(const int64_t* positions, int M, int head_dim, float theta, float* cos_sin, void* stream) {
  B200_REQUIRE(positions && cos_sin && M >= 0 && head_dim >= 2 && head_dim % 2 == 0 && theta > 0.f, B200_ERR_ARG, "rope_cos_sin: bad argument");
  if (M == 0) return B200_OK;
  if (int rc = require_sm100("rope_cos_sin")) return rc;
  const int n = M * (head_dim / 2);
  launch_pdl(rope_cos_sin_kernel, dim3((n + 255) / 256), dim3(256), 0, (cudaStream_t)stream, positions, M, head_dim / 2, log2f(theta),
             reinterpret_cast<float2*>(cos_sin));
  return launched("rope_cos_sin_kernel launch");
}

extern "C" int b200_rope_neox(void* q, int64_t q_sm, void* k, int64_t k_sm, const int64_t* positions, int dtype, int M, int Hq, int Hkv, int D,
                              float theta, void* stream) {
  B200_REQUIRE(q && k && positions && M >= 0 && D >= 2 && D % 2 == 0, B200_ERR_ARG, "rope_neox: bad argument");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "rope_neox: dtype");
  if (M == 0) return B200_OK;
  if (int rc = require_sm100("rope_neox")) return rc;
  // heads are split over blockIdx.y (two per thread) while the batch alone would leave SMs idle
  const int slots = 256 / (D / 2);
  int gy = 1;
  if (slots > 0 && M < 1024) gy = ((Hq + Hkv + slots - 1) / slots + 1) / 2;
  if (gy < 1) gy = 1;
  if (dtype == B200_BF16)
    launch_pdl(rope_neox_kernel<true>, dim3(M, gy), dim3(256), 0, (cudaStream_t)stream, (uint16_t*)q, q_sm, (uint16_t*)k, k_sm, positions, Hq, Hkv, D, log2f(theta));
  else
    launch_pdl(rope_neox_kernel<false>, dim3(M, gy), dim3(256), 0, (cudaStream_t)stream, (uint16_t*)q, q_sm, (uint16_t*)k, k_sm, positions, Hq, Hkv, D, log2f(theta));
  return launched("rope_neox_kernel launch");
}
