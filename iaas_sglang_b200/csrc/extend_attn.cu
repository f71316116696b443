// Extend (prefill) attention over the paged prefix + the new tokens: AttentionBackend.forward_extend
// (python/sglang/srt/layers/attention/base_attn_backend.py:101-111).  Replaces extend_attention_cpu
// (sgl-kernel/csrc/cpu/extend.cpp; schema csrc/cpu/torch_extension_cpu.cpp:271-275) and
// extend_attention_fwd (python/sglang/srt/layers/attention/triton_ops/extend_attention.py).
//
// v1: correct for every shape, not roofline-tuned ("surface must exist", tuning is
// 8f).  grid = (extend position, q head, request); 4 warps stride over the keys with an online
// softmax per warp and merge through shared memory.
#include <math.h>

#include "common.cuh"

namespace b200 {

struct ExtendKParams {
  const uint16_t* q;
  int64_t q_st, q_sh;
  const uint16_t* k_ext;
  int64_t ke_st, ke_sh;
  const uint16_t* v_ext;
  int64_t ve_st, ve_sh;
  uint16_t* out;
  const uint16_t* k_buf;
  int64_t kb_st, kb_sh;
  const uint16_t* v_buf;
  int64_t vb_st, vb_sh;
  const void* r2t;
  int64_t r2t_stride;
  const int64_t *rpi, *seq_lens, *ext_lens, *ext_start;
  int idx64, Hq, G, D, Dv;
  float sm_scale, logit_cap;
};

template <int MAXE, bool BF16>
__global__ void __launch_bounds__(128) extend_attn_kernel(const __grid_constant__ ExtendKParams p) {
  const int i = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int64_t ext = p.ext_lens[b];
  if (i >= ext) return;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int64_t start = p.ext_start[b], pre = p.seq_lens[b] - ext, req = p.rpi[b];
  const int64_t nkv = pre + i + 1;
  const int kvh = h / p.G;
  const int D = p.D, Dv = p.Dv;
  __shared__ float sm_o[4][32 * MAXE];
  __shared__ float sm_ml[4][2];
  float qr[MAXE], acc[MAXE];
#pragma unroll
  for (int e = 0; e < MAXE; ++e) {
    const int d = e * 32 + lane;
    qr[e] = d < D ? half_to_float<BF16>(p.q[(start + i) * p.q_st + (int64_t)h * p.q_sh + d]) : 0.f;
    acc[e] = 0.f;
  }
  float m_run = -INFINITY, l_run = 0.f;
  const bool has_cap = p.logit_cap > 0.f;
  constexpr float kLog2e = 1.4426950408889634f;
  for (int64_t j = warp; j < nkv; j += 4) {
    const uint16_t *kr, *vr;
    if (j < pre) {
      const int64_t idx = load_index(p.r2t, req * p.r2t_stride + j, p.idx64);
      kr = p.k_buf + idx * p.kb_st + kvh * p.kb_sh;
      vr = p.v_buf + idx * p.vb_st + kvh * p.vb_sh;
    } else {
      kr = p.k_ext + (start + j - pre) * p.ke_st + kvh * p.ke_sh;
      vr = p.v_ext + (start + j - pre) * p.ve_st + kvh * p.ve_sh;
    }
    float s = 0.f;
#pragma unroll
    for (int e = 0; e < MAXE; ++e) {
      const int d = e * 32 + lane;
      if (d < D) s += qr[e] * half_to_float<BF16>(kr[d]);
    }
    s = warp_sum(s) * p.sm_scale;
    if (has_cap) s = p.logit_cap * tanhf(s / p.logit_cap);
    s *= kLog2e;
    const float m_new = fmaxf(m_run, s);
    const float alpha = exp2f(m_run - m_new);
    const float pr = exp2f(s - m_new);
    l_run = l_run * alpha + pr;
    m_run = m_new;
#pragma unroll
    for (int e = 0; e < MAXE; ++e) {
      const int d = e * 32 + lane;
      if (d < Dv) acc[e] = acc[e] * alpha + pr * half_to_float<BF16>(vr[d]);
    }
  }
#pragma unroll
  for (int e = 0; e < MAXE; ++e) sm_o[warp][e * 32 + lane] = acc[e];
  if (lane == 0) {
    sm_ml[warp][0] = m_run;
    sm_ml[warp][1] = l_run;
  }
  __syncthreads();
  const float M = fmaxf(fmaxf(sm_ml[0][0], sm_ml[1][0]), fmaxf(sm_ml[2][0], sm_ml[3][0]));
  const float Mu = (M == -INFINITY) ? 0.f : M;
  float w[4], L = 0.f;
#pragma unroll
  for (int x = 0; x < 4; ++x) {
    w[x] = exp2f(sm_ml[x][0] - Mu);
    L += sm_ml[x][1] * w[x];
  }
  const float inv = L > 0.f ? 1.f / L : 0.f;
  for (int d = threadIdx.x; d < Dv; d += 128) {
    const float O = sm_o[0][d] * w[0] + sm_o[1][d] * w[1] + sm_o[2][d] * w[2] + sm_o[3][d] * w[3];
    p.out[((start + i) * p.Hq + h) * (int64_t)Dv + d] = float_to_half<BF16>(O * inv);
  }
}

template <bool BF16>
static int launch_extend(const ExtendKParams& kp, dim3 grid, cudaStream_t st) {
  const int mx = kp.D > kp.Dv ? kp.D : kp.Dv;
  if (mx <= 128) extend_attn_kernel<4, BF16><<<grid, 128, 0, st>>>(kp);
  else if (mx <= 256) extend_attn_kernel<8, BF16><<<grid, 128, 0, st>>>(kp);
  else if (mx <= 640) extend_attn_kernel<20, BF16><<<grid, 128, 0, st>>>(kp);
  else extend_attn_kernel<32, BF16><<<grid, 128, 0, st>>>(kp);
  return launched("extend_attn_kernel launch");
}

}  // namespace b200

using namespace b200;

extern "C" int b200_extend_attn(const b200_extend_attn_params* a, void* stream) {
  B200_REQUIRE(a != nullptr, B200_ERR_ARG, "extend_attn: null params");
  B200_REQUIRE(a->B >= 0 && a->Hq > 0 && a->Hkv > 0 && a->D > 0 && a->Dv > 0 && a->Hq % a->Hkv == 0, B200_ERR_ARG,
               "extend_attn: bad sizes");
  B200_REQUIRE(a->D <= 1024 && a->Dv <= 1024, B200_ERR_UNSUPPORTED, "extend_attn: head dim > 1024");
  B200_REQUIRE(a->dtype == B200_BF16 || a->dtype == B200_F16, B200_ERR_UNSUPPORTED, "extend_attn: dtype must be bf16/f16");
  B200_REQUIRE(a->q && a->k_ext && a->v_ext && a->out && a->k_buf && a->v_buf && a->req_to_token && a->req_pool_indices &&
                   a->seq_lens && a->extend_seq_lens && a->extend_start_loc,
               B200_ERR_ARG, "extend_attn: null tensor");
  if (a->B == 0 || a->max_len_extend <= 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "extend_attn: requires an sm_100 device (got sm_%d); no fallback", arch);
  ExtendKParams kp;
  kp.q = (const uint16_t*)a->q; kp.q_st = a->q_stride_t; kp.q_sh = a->q_stride_h;
  kp.k_ext = (const uint16_t*)a->k_ext; kp.ke_st = a->kext_stride_t; kp.ke_sh = a->kext_stride_h;
  kp.v_ext = (const uint16_t*)a->v_ext; kp.ve_st = a->vext_stride_t; kp.ve_sh = a->vext_stride_h;
  kp.out = (uint16_t*)a->out;
  kp.k_buf = (const uint16_t*)a->k_buf; kp.kb_st = a->kbuf_stride_t; kp.kb_sh = a->kbuf_stride_h;
  kp.v_buf = (const uint16_t*)a->v_buf; kp.vb_st = a->vbuf_stride_t; kp.vb_sh = a->vbuf_stride_h;
  kp.r2t = a->req_to_token; kp.r2t_stride = a->req_to_token_stride; kp.idx64 = a->idx_is_i64;
  kp.rpi = a->req_pool_indices; kp.seq_lens = a->seq_lens; kp.ext_lens = a->extend_seq_lens; kp.ext_start = a->extend_start_loc;
  kp.Hq = a->Hq; kp.G = a->Hq / a->Hkv; kp.D = a->D; kp.Dv = a->Dv; kp.sm_scale = a->sm_scale; kp.logit_cap = a->logit_cap;
  dim3 grid(a->max_len_extend, a->Hq, a->B);
  return a->dtype == B200_BF16 ? launch_extend<true>(kp, grid, (cudaStream_t)stream)
                               : launch_extend<false>(kp, grid, (cudaStream_t)stream);
}
