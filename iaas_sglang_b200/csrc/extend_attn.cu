// Extend (prefill) attention over the paged prefix + the new tokens: AttentionBackend.forward_extend
// (python/sglang/srt/layers/attention/base_attn_backend.py:101-111).  Replaces extend_attention_cpu
// (sgl-kernel/csrc/cpu/extend.cpp; schema csrc/cpu/torch_extension_cpu.cpp:271-275) and
// extend_attention_fwd (python/sglang/srt/layers/attention/triton_ops/extend_attention.py).
//
// Two kernels:
//   extend_attn_mma_kernel (D = Dv in {64, 128}): the tuned one.  A CTA owns 64 query ROWS of one
//     request and one kv head -- a row is a (new token, q head of the kv head's GQA group) pair, so the G heads that
//     share a kv head share its staged K/V tiles -- and walks the keys in 64-token tiles: page-table lookup + 16-byte
//     cp.async into a double-buffered, bank-conflict-free (row pitch + 16 B) shared-memory tile, QK^T and PV on the warp
//     tensor cores (mma.sync m16n8k16, fp32 accumulate), online softmax in registers, causal mask only on the tiles that
//     straddle the diagonal.  Keys come from the paged pool for the prefix and from k_extend / v_extend for the new tokens
//     (as the reference's Triton kernel does, triton_ops/extend_attention.py).  K/V are read ceil(rows / 64) times per
//     kv head instead of once per (token, q head).  Optional: prefix_only (every row sees exactly the prefix: the shared-
//     prefix half of cascade decode) and a per-row log-sum-exp output for merge_state.
//   extend_attn_kernel: SIMT, any head size (MLA 576/512, odd sizes), one (token, q head) per CTA.
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
  int prefix_only;   // tuned kernel: every row attends to exactly the prefix (no new tokens, no causal mask)
  float* lse_out;    // tuned kernel: optional [T_ext, Hq] natural-log LSE
  float* out_f32;    // tuned kernel: optional [T_ext, Hq, Dv] fp32 output instead of `out`; P is then kept exact (hi + lo)
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

// ---------------------------------------------------------------------------------------------------------------
// tuned kernel
// ---------------------------------------------------------------------------------------------------------------
constexpr int EX_BM = 64, EX_BN = 64;
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <int D, bool BF16>
__global__ void __launch_bounds__(128) extend_attn_mma_kernel(const __grid_constant__ ExtendKParams p) {
  constexpr int PITCH = D * 2 + 16;              // bytes per staged row: 16 B of padding keeps ldmatrix conflict-free
  constexpr int TILE_BYTES = EX_BN * PITCH;
  constexpr int CPR = D * 2 / 16;                // 16-byte chunks per row
  constexpr int RPP = 128 / CPR;                 // rows per pass of the 128 threads
  constexpr float kLog2e = 1.4426950408889634f;
  extern __shared__ __align__(16) uint8_t ex_smem[];   // [2 stages][K tile | V tile]

  const int b = blockIdx.z, kvh = blockIdx.y;
  const int G = p.G;
  const int64_t ext = p.ext_lens[b];
  const int64_t rows_total = ext * G;
  const int64_t row0 = (int64_t)blockIdx.x * EX_BM;
  if (row0 >= rows_total) return;
  const int64_t start = p.ext_start[b], pre = p.seq_lens[b] - ext, req = p.rpi[b];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int g = lane >> 2, t = lane & 3;
  const bool prefix_only = p.prefix_only != 0;

  // this thread's two rows (MMA rows g and g + 8 of the warp's 16)
  int64_t rr[2], tok[2], nkv_row[2];
  int head[2];
  bool rvalid[2];
#pragma unroll
  for (int x = 0; x < 2; ++x) {
    rr[x] = row0 + warp * 16 + g + 8 * x;
    rvalid[x] = rr[x] < rows_total;
    const int64_t r = rvalid[x] ? rr[x] : rows_total - 1;
    tok[x] = r / G;
    head[x] = kvh * G + (int)(r - tok[x] * G);
    nkv_row[x] = rvalid[x] ? (prefix_only ? pre : pre + tok[x] + 1) : 0;
  }
  int64_t last_row = row0 + EX_BM - 1;
  if (last_row > rows_total - 1) last_row = rows_total - 1;
  const int64_t nkv_cta = prefix_only ? pre : pre + last_row / G + 1;    // keys any row of the CTA may see
  const int64_t first_full = prefix_only ? pre : pre + row0 / G + 1;     // keys EVERY row of the CTA sees
  const int n_tiles = (int)((nkv_cta + EX_BN - 1) / EX_BN);

  // ---- Q fragments (16 rows x D per warp), straight from global memory ----
  uint32_t qf[D / 16][4];
  {
    const uint16_t* q0 = p.q + (start + tok[0]) * p.q_st + (int64_t)head[0] * p.q_sh + 2 * t;
    const uint16_t* q1 = p.q + (start + tok[1]) * p.q_st + (int64_t)head[1] * p.q_sh + 2 * t;
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
      qf[ks][0] = rvalid[0] ? *reinterpret_cast<const uint32_t*>(q0 + ks * 16) : 0u;
      qf[ks][1] = rvalid[1] ? *reinterpret_cast<const uint32_t*>(q1 + ks * 16) : 0u;
      qf[ks][2] = rvalid[0] ? *reinterpret_cast<const uint32_t*>(q0 + ks * 16 + 8) : 0u;
      qf[ks][3] = rvalid[1] ? *reinterpret_cast<const uint32_t*>(q1 + ks * 16 + 8) : 0u;
    }
  }

  // ---- tile loader: 128 threads, thread = (row in pass, 16-byte chunk) ----
  const int lrow = threadIdx.x / CPR, chunk = threadIdx.x % CPR;
  const uint8_t* r2t_row = reinterpret_cast<const uint8_t*>(p.r2t) + (req * p.r2t_stride) * (p.idx64 ? 8 : 4);
  auto load_tile = [&](int tile, int stage) {
    const uint32_t kdst = smem_u32(ex_smem + stage * 2 * TILE_BYTES) + chunk * 16;
#pragma unroll
    for (int ps = 0; ps < EX_BN / RPP; ++ps) {
      const int row = ps * RPP + lrow;
      int64_t pos = (int64_t)tile * EX_BN + row;
      if (pos > nkv_cta - 1) pos = nkv_cta - 1;   // tail: a valid row again (its scores are masked)
      const uint16_t *ks, *vs;
      if (pos < pre) {
        const int64_t idx = p.idx64 ? *reinterpret_cast<const int64_t*>(r2t_row + pos * 8)
                                    : (int64_t) * reinterpret_cast<const int32_t*>(r2t_row + pos * 4);
        ks = p.k_buf + idx * p.kb_st + kvh * p.kb_sh;
        vs = p.v_buf + idx * p.vb_st + kvh * p.vb_sh;
      } else {
        ks = p.k_ext + (start + pos - pre) * p.ke_st + kvh * p.ke_sh;
        vs = p.v_ext + (start + pos - pre) * p.ve_st + kvh * p.ve_sh;
      }
      cp_async_16(kdst + row * PITCH, ks + chunk * 8);
      cp_async_16(kdst + TILE_BYTES + row * PITCH, vs + chunk * 8);
    }
    cp_async_commit();
  };

  const bool has_cap = p.logit_cap > 0.f;
  const float scale2 = p.sm_scale * kLog2e;
  const int lm = lane >> 3, lr = lane & 7;
  const uint32_t k_lane_off = (uint32_t)(((lm >> 1) * 8 + lr) * PITCH + (lm & 1) * 16);
  const uint32_t v_lane_off = (uint32_t)(((lm & 1) * 8 + lr) * PITCH + (lm >> 1) * 16);
  float o[D / 8][4];
#pragma unroll
  for (int j = 0; j < D / 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};

  if (n_tiles > 0) load_tile(0, 0);
  for (int it = 0; it < n_tiles; ++it) {
    if (it + 1 < n_tiles) {
      load_tile(it + 1, (it + 1) & 1);
      cp_async_wait<1>();
    } else {
      cp_async_wait<0>();
    }
    __syncthreads();
    const uint32_t kbase = smem_u32(ex_smem + (it & 1) * 2 * TILE_BYTES);
    const uint32_t vbase = kbase + TILE_BYTES;
    // ---- S = Q K^T: 16 rows x 64 keys per warp ----
    float sc[EX_BN / 8][4];
#pragma unroll
    for (int nt = 0; nt < EX_BN / 8; ++nt) sc[nt][0] = sc[nt][1] = sc[nt][2] = sc[nt][3] = 0.f;
#pragma unroll
    for (int ks = 0; ks < D / 16; ++ks) {
#pragma unroll
      for (int np = 0; np < EX_BN / 16; ++np) {
        uint32_t r0, r1, r2, r3;
        ldsm_x4(kbase + np * 16 * PITCH + k_lane_off + ks * 32, r0, r1, r2, r3);
        mma_16816<BF16>(sc[2 * np], qf[ks], r0, r1);
        mma_16816<BF16>(sc[2 * np + 1], qf[ks], r2, r3);
      }
    }
    // ---- scale, cap, mask, online softmax ----
    const int64_t tile0 = (int64_t)it * EX_BN;
    const bool need_mask = tile0 + EX_BN > first_full;
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < EX_BN / 8; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = sc[nt][e];
        if (has_cap) {
          v *= p.sm_scale;
          v = p.logit_cap * tanhf(v / p.logit_cap) * kLog2e;
        } else {
          v *= scale2;
        }
        if (need_mask) {
          const int64_t pos = tile0 + nt * 8 + 2 * t + (e & 1);
          if (pos >= nkv_row[e >> 1]) v = -INFINITY;
        }
        sc[nt][e] = v;
        mx[e >> 1] = fmaxf(mx[e >> 1], v);
      }
    float alpha[2], m_use[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float m_new = fmaxf(m_run[r], mx[r]);
      m_use[r] = (m_new == -INFINITY) ? 0.f : m_new;
      alpha[r] = exp2f(m_run[r] - m_use[r]);
      m_run[r] = m_new;
    }
    // P is rounded to the KV dtype before P V, as the reference kernels do (extend_attention.py: p.to(v.dtype))
    // (exact mode -- the fp32 state for cascade decode, which is held to the decode tolerance: P = hi + lo halves and
    // a second PV pass, as the decode kernel does)
    const bool exact = p.out_f32 != nullptr;
    uint32_t pa[EX_BN / 16][4], pb[EX_BN / 16][4];
    float ps[2] = {0.f, 0.f};
#pragma unroll
    for (int nt = 0; nt < EX_BN / 8; ++nt) {
      const float p0 = exp2f(sc[nt][0] - m_use[0]), p1 = exp2f(sc[nt][1] - m_use[0]);
      const float p2 = exp2f(sc[nt][2] - m_use[1]), p3 = exp2f(sc[nt][3] - m_use[1]);
      const uint32_t h01 = pack_half2<BF16>(p0, p1), h23 = pack_half2<BF16>(p2, p3);
      pa[nt >> 1][(nt & 1) * 2 + 0] = h01;
      pa[nt >> 1][(nt & 1) * 2 + 1] = h23;
      const float2 f01 = unpack_half2<BF16>(h01), f23 = unpack_half2<BF16>(h23);
      pb[nt >> 1][(nt & 1) * 2 + 0] = pack_half2<BF16>(p0 - f01.x, p1 - f01.y);
      pb[nt >> 1][(nt & 1) * 2 + 1] = pack_half2<BF16>(p2 - f23.x, p3 - f23.y);
      ps[0] += p0 + p1;
      ps[1] += p2 + p3;
    }
    l_run[0] = l_run[0] * alpha[0] + ps[0];
    l_run[1] = l_run[1] * alpha[1] + ps[1];
    if (alpha[0] != 1.f || alpha[1] != 1.f) {
#pragma unroll
      for (int j = 0; j < D / 8; ++j) {
        o[j][0] *= alpha[0];
        o[j][1] *= alpha[0];
        o[j][2] *= alpha[1];
        o[j][3] *= alpha[1];
      }
    }
    // ---- O += P V: 16 x 64 times 64 x D ----
#pragma unroll
    for (int kk = 0; kk < EX_BN / 16; ++kk) {
#pragma unroll
      for (int jp = 0; jp < D / 16; ++jp) {
        uint32_t r0, r1, r2, r3;
        ldsm_x4_t(vbase + kk * 16 * PITCH + v_lane_off + jp * 32, r0, r1, r2, r3);
        mma_16816<BF16>(o[2 * jp], pa[kk], r0, r1);
        mma_16816<BF16>(o[2 * jp + 1], pa[kk], r2, r3);
        if (exact) {
          mma_16816<BF16>(o[2 * jp], pb[kk], r0, r1);
          mma_16816<BF16>(o[2 * jp + 1], pb[kk], r2, r3);
        }
      }
    }
    __syncthreads();   // everyone is done with this stage before the tile after next lands in it
  }

  // ---- normalise and store ----
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
#pragma unroll
  for (int x = 0; x < 2; ++x) {
    if (!rvalid[x]) continue;
    const float inv = l_run[x] > 0.f ? 1.f / l_run[x] : 0.f;
    if (p.out_f32) {
      float* frow = p.out_f32 + ((start + tok[x]) * p.Hq + head[x]) * (int64_t)D + 2 * t;
#pragma unroll
      for (int j = 0; j < D / 8; ++j) *reinterpret_cast<float2*>(frow + j * 8) = make_float2(o[j][2 * x] * inv, o[j][2 * x + 1] * inv);
    } else {
      uint16_t* orow = p.out + ((start + tok[x]) * p.Hq + head[x]) * (int64_t)D + 2 * t;
#pragma unroll
      for (int j = 0; j < D / 8; ++j)
        *reinterpret_cast<uint32_t*>(orow + j * 8) = pack_half2<BF16>(o[j][2 * x] * inv, o[j][2 * x + 1] * inv);
    }
    if (p.lse_out && t == 0)   // natural-log LSE of the row (-inf for an empty key set), for merge_state
      p.lse_out[(start + tok[x]) * p.Hq + head[x]] = l_run[x] > 0.f ? (m_run[x] + log2f(l_run[x])) / kLog2e : -INFINITY;
  }
}

template <int D, bool BF16>
static int launch_extend_mma(const ExtendKParams& kp, int B, int Hkv, int max_len_extend, cudaStream_t st) {
  constexpr int SMEM = 2 * 2 * EX_BN * (D * 2 + 16);
  auto kern = extend_attn_mma_kernel<D, BF16>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SMEM));
    configured = true;
  }
  const int64_t rows = (int64_t)max_len_extend * kp.G;
  dim3 grid((unsigned)((rows + EX_BM - 1) / EX_BM), Hkv, B);
  kern<<<grid, 128, SMEM, st>>>(kp);
  return launched("extend_attn_mma_kernel launch");
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
  B200_REQUIRE(a->q && a->k_ext && a->v_ext && (a->out || a->out_f32) && a->k_buf && a->v_buf && a->req_to_token && a->req_pool_indices &&
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
  kp.prefix_only = a->prefix_only; kp.lse_out = (float*)a->lse_out; kp.out_f32 = (float*)a->out_f32;
  const bool bf = a->dtype == B200_BF16;
  auto al16 = [](const void* ptr, int64_t s0, int64_t s1) { return ((uintptr_t)ptr & 15) == 0 && s0 % 8 == 0 && s1 % 8 == 0; };
  const bool tuned = a->D == a->Dv && (a->D == 64 || a->D == 128) && al16(a->k_ext, a->kext_stride_t, a->kext_stride_h) &&
                     al16(a->v_ext, a->vext_stride_t, a->vext_stride_h) && al16(a->k_buf, a->kbuf_stride_t, a->kbuf_stride_h) &&
                     al16(a->v_buf, a->vbuf_stride_t, a->vbuf_stride_h) && ((uintptr_t)a->q & 3) == 0 && a->q_stride_t % 2 == 0 &&
                     a->q_stride_h % 2 == 0;
  if (tuned) {
    cudaStream_t st = (cudaStream_t)stream;
    if (a->D == 128) return bf ? launch_extend_mma<128, true>(kp, a->B, a->Hkv, a->max_len_extend, st)
                               : launch_extend_mma<128, false>(kp, a->B, a->Hkv, a->max_len_extend, st);
    return bf ? launch_extend_mma<64, true>(kp, a->B, a->Hkv, a->max_len_extend, st)
              : launch_extend_mma<64, false>(kp, a->B, a->Hkv, a->max_len_extend, st);
  }
  B200_REQUIRE(!a->prefix_only && !a->lse_out && !a->out_f32, B200_ERR_UNSUPPORTED,
               "extend_attn: prefix_only / lse_out / out_f32 need the tuned kernel (D == Dv in {64, 128}, 16-byte aligned rows)");
  dim3 grid(a->max_len_extend, a->Hq, a->B);
  return a->dtype == B200_BF16 ? launch_extend<true>(kp, grid, (cudaStream_t)stream)
                               : launch_extend<false>(kp, grid, (cudaStream_t)stream);
}
