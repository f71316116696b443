// MLA absorbed decode (kv_lora_rank 512 + rope 64) for sm_100a -- included by decode_attn.cu inside namespace b200.
//
// Replaces cutlass_mla_decode (sgl-kernel/csrc/attention/cutlass_mla_kernel.cu:210-246) and the MLA branch of
// decode_attention_cpu (sgl-kernel/csrc/cpu/decode.cpp:1040-1230, detected as in :1455): one latent row of 576
// 16-bit values per token is both the key (all 576 columns) and the value (its first 512 columns) of EVERY query
// head, so a token is read from HBM once (1152 B) and feeds QK^T and PV of a 16-head block.  At 16 heads per GPU
// (DeepSeek-V3 under TP8) that is 30-45 flop per byte: HBM-bound with the math on the warp tensor-core path.
//
//   * same persistent one-CTA-per-SM launch and on-device work partition as decode_attn_mma_kernel (a unit is
//     (request, 16-head q block), a tile is 64 tokens); partial results of split units go through the same
//     two-slots-per-CTA workspace and decode_merge_part_kernel;
//   * four producer warps read req_to_token in place and issue one 1152-byte cp.async.bulk per token row into a
//     5-stage ring of 32-row stages (a tile is two consecutive stages);
//   * eight consumer warps.  QK^T: warp (tw, kh) multiplies the 16-token slice tw by half kh of the 576 columns
//     (Q fragments of that half stay in registers), the two halves swap one accumulator tile through shared memory;
//     softmax: the tile maximum is agreed through a 8x16 table, every warp exponentiates its own 16x8 block once
//     and publishes it as hi + lo 16-bit parts (P.V then equals an fp32-P product);
//     PV: warp w owns value columns [64w, 64w+64) for all 16 heads and all 64 tokens -- no cross-warp merge of O.

struct MlaCfg {
  static constexpr int D = 576, Dv = 512;
  static constexpr int TILE_T = 64;  // tokens per consumer tile (the unit of the work partition)
  static constexpr int SUB_T = 32;   // token rows per ring stage
  static constexpr int ROW_BYTES = D * 2;
  static constexpr int ROW_PITCH = ROW_BYTES + 16;  // 73 x 16 B: ldmatrix rows fall in distinct bank groups
  static constexpr int STAGE_BYTES = SUB_T * ROW_PITCH;
  static constexpr int STAGES = 5;
  static constexpr int RING_BYTES = STAGES * STAGE_BYTES;
  static constexpr int X_BYTES = kConsumerWarps * 32 * 16;  // score exchange: one float4 per lane
  static constexpr int P_PITCH = TILE_T * 2 + 16;
  static constexpr int P_BYTES = 16 * P_PITCH;              // one of (hi, lo)
  static constexpr int ML_BYTES = kConsumerWarps * 16 * 4;  // per-warp row maxima / row sums
  static constexpr int TABLE_BYTES = kMaxBatchTable * 8 + (kMaxBatchTable + 4) * 4 + 2 * kMaxBatchTable * 4;
  static constexpr int SCRATCH_OFF = RING_BYTES;
  static constexpr int TABLE_OFF = SCRATCH_OFF + X_BYTES + 2 * P_BYTES + 2 * ML_BYTES;
  static constexpr int SMEM_BYTES = TABLE_OFF + TABLE_BYTES + 2 * STAGES * 8 + 16;
  static_assert(STAGE_BYTES % 128 == 0 && TABLE_OFF % 16 == 0, "alignment");
  static_assert(SMEM_BYTES <= kMaxSmem, "shared memory");
};

template <bool BF16>
__global__ void __launch_bounds__(kAttnThreads, 1) decode_mla_mma_kernel(const __grid_constant__ DecodeKParams p) {
  using C = MlaCfg;
  constexpr int KH_STEPS = C::D / 32;  // k-steps (of 16 columns) per half: 18
  extern __shared__ __align__(128) uint8_t smem[];
  float4* s_x = reinterpret_cast<float4*>(smem + C::SCRATCH_OFF);
  uint8_t* s_phi = smem + C::SCRATCH_OFF + C::X_BYTES;
  uint8_t* s_plo = s_phi + C::P_BYTES;
  float* s_max = reinterpret_cast<float*>(s_plo + C::P_BYTES);  // [warp][16]
  float* s_sum = s_max + kConsumerWarps * 16;                   // [warp][16]
  int64_t* s_loc = reinterpret_cast<int64_t*>(smem + C::TABLE_OFF);
  int32_t* s_P = reinterpret_cast<int32_t*>(s_loc + kMaxBatchTable);
  int32_t* s_len = s_P + kMaxBatchTable + 4;
  int32_t* s_rpi = s_len + kMaxBatchTable;
  uint64_t* full = reinterpret_cast<uint64_t*>(s_rpi + kMaxBatchTable);
  uint64_t* empty = full + C::STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_qblk = (p.Hq + 15) >> 4;
  const int NG = n_qblk;

  for (int b = threadIdx.x; b < p.B; b += kAttnThreads) {
    int64_t len = seq_len_at(p, b);
    if (len < 0) len = 0;
    s_len[b] = (int32_t)len;
    s_rpi[b] = (int32_t)req_at(p, b);
    s_loc[b] = p.fused_new ? loc_at(p, b) : -1;
    s_P[b + 1] = (int32_t)((len + C::TILE_T - 1) / C::TILE_T);
  }
  if (threadIdx.x == 0) {
    s_P[0] = 0;
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], kProducerWarps);
      mbar_init(&empty[s], kConsumerWarps);
    }
    fence_mbar_init();
  }
  __syncthreads();
  if (warp == 0) {
    const int chunk = (p.B + 31) / 32;
    const int b0 = lane * chunk, b1 = min(p.B, b0 + chunk);
    int sum = 0;
    for (int b = b0; b < b1; ++b) sum += s_P[b + 1];
    int incl = sum;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    int run = incl - sum;
    for (int b = b0; b < b1; ++b) {
      run += s_P[b + 1];
      s_P[b + 1] = run;
    }
  }
  __syncthreads();

  const int64_t W = (int64_t)NG * s_P[p.B];
  const int64_t t_begin = part_start(blockIdx.x, W, gridDim.x), t_end = part_start(blockIdx.x + 1, W, gridDim.x);
  if (t_begin >= t_end) return;
  SegCursor cur;
  cur.init(s_P, p.B, NG, t_begin, t_end);

  if (warp >= kConsumerWarps) {
    // ================================ producer warps ================================
    const int pw = warp - kConsumerWarps;
    constexpr int RW = C::SUB_T / kProducerWarps;  // 8 rows per producer warp per stage, one per lane
    uint32_t gs = 0;                               // stages issued so far
    for (; !cur.done();) {
      const int n = cur.seg_len();
      const int b = cur.b, qblk = cur.g;
      const int64_t len = s_len[b], req = s_rpi[b], loc_b = s_loc[b];
      const int64_t kv0 = (int64_t)cur.tile * C::TILE_T;
      const int64_t kv_end = (int64_t)(cur.tile + n) * C::TILE_T;
      const int64_t kv1 = kv_end < len ? kv_end : len;
      const uint16_t* knew_row = p.k_new + b * p.kn_sb;
      const uint16_t* kpe_row = p.k_pe_new ? p.k_pe_new + b * p.kpe_sb : nullptr;
      if (p.fused_new && pw == 0 && qblk == 0 && cur.tile + n == cur.tiles_b) {
        // fused set_mla_kv_buffer (memory_pool.py:880-905): the attention reads the new row from k_new
        const uint4* ks = reinterpret_cast<const uint4*>(knew_row);
        uint4* kd = reinterpret_cast<uint4*>(p.k_buf + loc_b * p.kb_st);
        if (kpe_row) {  // set_mla_kv_buffer(layer, loc, k_nope, k_rope): the two parts of the latent row arrive apart
          const uint4* kr = reinterpret_cast<const uint4*>(kpe_row);
          for (int i = lane; i < C::Dv * 2 / 16; i += 32) kd[i] = ks[i];
          if (lane < (C::D - C::Dv) * 2 / 16) kd[C::Dv * 2 / 16 + lane] = kr[lane];
        } else {
          for (int i = lane; i < C::ROW_BYTES / 16; i += 32) kd[i] = ks[i];
        }
      }
      const uint8_t* r2t_row = reinterpret_cast<const uint8_t*>(p.r2t) + (req * p.r2t_stride) * (p.idx64 ? 8 : 4);
      int32_t lo = 0, hi = 0, nlo = 0, nhi = 0;
      auto load_idx = [&](int j, int32_t& dlo, int32_t& dhi) {
        int64_t pos = kv0 + (int64_t)j * C::SUB_T + pw * RW + lane;
        if (pos > kv1 - 1) pos = kv1 - 1;  // tail: duplicate the last valid row (masked by the consumers)
        pos >>= p.page_shift;              // page tables (cutlass_mla_decode): one entry per page
        if (lane < RW) {
          if (p.idx64) {
            const int2 w = *reinterpret_cast<const int2*>(r2t_row + pos * 8);
            dlo = w.x; dhi = w.y;
          } else {
            dlo = *reinterpret_cast<const int32_t*>(r2t_row + pos * 4);
          }
        }
      };
      const int64_t page_mask = ((int64_t)1 << p.page_shift) - 1;
      load_idx(0, lo, hi);
      const int nsub = 2 * n;
      for (int j = 0; j < nsub; ++j, ++gs) {
        const int s = gs % C::STAGES;
        const uint32_t ph = (gs / C::STAGES) & 1;
        if (j + 1 < nsub) load_idx(j + 1, nlo, nhi);
        mbar_wait(&empty[s], ph ^ 1);
        if (lane == 0) mbar_arrive_expect_tx(&full[s], RW * C::ROW_BYTES);
        __syncwarp();
        if (lane < RW) {
          int64_t idx = p.idx64 ? (((int64_t)hi << 32) | (uint32_t)lo) : (int64_t)lo;
          if (p.page_shift) {
            int64_t pos = kv0 + (int64_t)j * C::SUB_T + pw * RW + lane;
            if (pos > kv1 - 1) pos = kv1 - 1;
            idx = (idx << p.page_shift) + (pos & page_mask);
          }
          const uint16_t* src = p.k_buf + idx * p.kb_st;
          uint8_t* dst = smem + s * C::STAGE_BYTES + (pw * RW + lane) * C::ROW_PITCH;
          if (idx == loc_b && kpe_row) {  // the new row in two pieces: same transaction bytes
            bulk_g2s(dst, knew_row, C::Dv * 2, &full[s]);
            bulk_g2s(dst + C::Dv * 2, kpe_row, (C::D - C::Dv) * 2, &full[s]);
          } else {
            if (idx == loc_b) src = knew_row;
            bulk_g2s(dst, src, C::ROW_BYTES, &full[s]);
          }
        }
        lo = nlo;
        hi = nhi;
      }
      cur.advance(n);
    }
    return;
  }

  // ================================ consumer warps ================================
  const int tw = warp & 3, kh = warp >> 2;
  const int g = lane >> 2, t = lane & 3;
  const float scale2 = p.sm_scale * kLog2e;
  const int lm = lane >> 3, lr = lane & 7;
  const uint32_t k_lane_off = (uint32_t)(((lm >> 1) * 8 + lr) * C::ROW_PITCH + (lm & 1) * 16);
  const uint32_t v_lane_off = (uint32_t)(((lm & 1) * 8 + lr) * C::ROW_PITCH + (lm >> 1) * 16);
  const uint32_t p_lane_off = (uint32_t)(((lm & 1) * 8 + lr) * C::P_PITCH + (lm >> 1) * 16);
  const uint32_t smem_base = smem_u32(smem);
  const uint32_t phi_base = smem_u32(s_phi) + p_lane_off, plo_base = smem_u32(s_plo) + p_lane_off;
  // where this lane publishes its four softmax weights: rows g / g+8, tokens tw*16 + kh*8 + 2t, +1
  const uint32_t p_st_off = (uint32_t)(g * C::P_PITCH + (tw * 16 + kh * 8 + 2 * t) * 2);
  uint32_t gs = 0;

  for (; !cur.done();) {
    const int n = cur.seg_len();
    const int b = cur.b, qblk = cur.g;
    const int row0 = qblk * 16;
    const int rows_valid = min(16, p.Hq - row0);
    const int64_t len = s_len[b];
    const int64_t kv0 = (int64_t)cur.tile * C::TILE_T;
    const int64_t kv_end = (int64_t)(cur.tile + n) * C::TILE_T;
    const int64_t kv1 = kv_end < len ? kv_end : len;
    const bool whole_unit = (cur.tile == 0 && n == cur.tiles_b);
    const bool first_seg = (cur.t == t_begin);

    uint32_t qf[KH_STEPS][4];
    {
      const uint16_t* q0 = p.q + b * p.q_sb + (int64_t)(row0 + g) * p.q_sh + kh * (C::D / 2) + 2 * t;
      const uint16_t* q1 = q0 + 8 * p.q_sh;
      // q_nope / q_pe given apart (cutlass_mla_backend.py:252-262): columns >= 512 come from q_pe, no concat
      const uint16_t* e0 = p.q_pe ? p.q_pe + b * p.qpe_sb + (int64_t)(row0 + g) * p.qpe_sh + 2 * t - C::Dv : nullptr;
      const uint16_t* e1 = e0 ? e0 + 8 * p.qpe_sh : nullptr;
      const bool v0 = g < rows_valid, v1 = g + 8 < rows_valid;
#pragma unroll
      for (int ks = 0; ks < KH_STEPS; ++ks) {
        const int col = kh * (C::D / 2) + ks * 16;  // 16-column k-steps never straddle column 512
        const bool pe = p.q_pe != nullptr && col >= C::Dv;
        const uint16_t* a0 = pe ? e0 + col : q0 + ks * 16;
        const uint16_t* a1 = pe ? e1 + col : q1 + ks * 16;
        qf[ks][0] = v0 ? *reinterpret_cast<const uint32_t*>(a0) : 0u;
        qf[ks][1] = v1 ? *reinterpret_cast<const uint32_t*>(a1) : 0u;
        qf[ks][2] = v0 ? *reinterpret_cast<const uint32_t*>(a0 + 8) : 0u;
        qf[ks][3] = v1 ? *reinterpret_cast<const uint32_t*>(a1 + 8) : 0u;
      }
    }
    float o[8][4];
#pragma unroll
    for (int j = 0; j < 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_part[2] = {0.f, 0.f};

    for (int it = 0; it < n; ++it, gs += 2) {
      const int sA = gs % C::STAGES, sB = (gs + 1) % C::STAGES;
      const uint32_t phA = (gs / C::STAGES) & 1, phB = ((gs + 1) / C::STAGES) & 1;
      const uint32_t stA = smem_base + sA * C::STAGE_BYTES, stB = smem_base + sB * C::STAGE_BYTES;
      // ---- QK^T: 16 heads x 16 tokens (slice tw) over half kh of the 576 columns ----
      if (tw < 2) mbar_wait(&full[sA], phA); else mbar_wait(&full[sB], phB);
      const uint32_t kbase = (tw < 2 ? stA : stB) + ((tw & 1) * 16) * C::ROW_PITCH + kh * C::D + k_lane_off;
      float sc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
#pragma unroll
      for (int ks = 0; ks < KH_STEPS; ++ks) {
        uint32_t r0, r1, r2, r3;
        ldsm_x4(kbase + ks * 32, r0, r1, r2, r3);
        mma_16816<BF16>(sc[0], qf[ks], r0, r1);
        mma_16816<BF16>(sc[1], qf[ks], r2, r3);
      }
      // the partner warp (other half of the columns, same tokens) gets the 8-token block it will finish
      s_x[warp * 32 + lane] = kh ? make_float4(sc[0][0], sc[0][1], sc[0][2], sc[0][3])
                                 : make_float4(sc[1][0], sc[1][1], sc[1][2], sc[1][3]);
      consumer_bar();
      float sv[4];
      {
        const float4 o4 = s_x[(warp ^ 4) * 32 + lane];
        sv[0] = (kh ? sc[1][0] : sc[0][0]) + o4.x;
        sv[1] = (kh ? sc[1][1] : sc[0][1]) + o4.y;
        sv[2] = (kh ? sc[1][2] : sc[0][2]) + o4.z;
        sv[3] = (kh ? sc[1][3] : sc[0][3]) + o4.w;
      }
      const int64_t tok0 = kv0 + (int64_t)it * C::TILE_T + tw * 16 + kh * 8 + 2 * t;
      float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        float v = sv[e] * scale2;
        if (tok0 + (e & 1) >= kv1) v = -INFINITY;
        sv[e] = v;
        mx[e >> 1] = fmaxf(mx[e >> 1], v);
      }
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
        mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      }
      if (t == 0) {
        s_max[warp * 16 + g] = mx[0];
        s_max[warp * 16 + g + 8] = mx[1];
      }
      consumer_bar();
      float alpha[2], m_use[2];
#pragma unroll
      for (int r = 0; r < 2; ++r) {
        float m_new = m_run[r];
#pragma unroll
        for (int w = 0; w < kConsumerWarps; ++w) m_new = fmaxf(m_new, s_max[w * 16 + g + 8 * r]);
        m_use[r] = (m_new == -INFINITY) ? 0.f : m_new;
        alpha[r] = exp2f(m_run[r] - m_use[r]);
        m_run[r] = m_new;
      }
      {
        const float p0 = exp2f(sv[0] - m_use[0]), p1 = exp2f(sv[1] - m_use[0]);
        const float p2 = exp2f(sv[2] - m_use[1]), p3 = exp2f(sv[3] - m_use[1]);
        const uint32_t h01 = pack_half2<BF16>(p0, p1), h23 = pack_half2<BF16>(p2, p3);
        const float2 f01 = unpack_half2<BF16>(h01), f23 = unpack_half2<BF16>(h23);
        *reinterpret_cast<uint32_t*>(s_phi + p_st_off) = h01;
        *reinterpret_cast<uint32_t*>(s_phi + p_st_off + 8 * C::P_PITCH) = h23;
        *reinterpret_cast<uint32_t*>(s_plo + p_st_off) = pack_half2<BF16>(p0 - f01.x, p1 - f01.y);
        *reinterpret_cast<uint32_t*>(s_plo + p_st_off + 8 * C::P_PITCH) = pack_half2<BF16>(p2 - f23.x, p3 - f23.y);
        l_part[0] = l_part[0] * alpha[0] + (p0 + p1);
        l_part[1] = l_part[1] * alpha[1] + (p2 + p3);
      }
      if (alpha[0] != 1.f || alpha[1] != 1.f) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          o[j][0] *= alpha[0];
          o[j][1] *= alpha[0];
          o[j][2] *= alpha[1];
          o[j][3] *= alpha[1];
        }
      }
      consumer_bar();
      // ---- PV: all 16 heads x 64 tokens into value columns [64 warp, 64 warp + 64) ----
      if (tw < 2) mbar_wait(&full[sB], phB); else mbar_wait(&full[sA], phA);
#pragma unroll
      for (int ks = 0; ks < 4; ++ks) {
        uint32_t pa[4], pb[4];
        ldsm_x4(phi_base + ks * 32, pa[0], pa[1], pa[2], pa[3]);
        ldsm_x4(plo_base + ks * 32, pb[0], pb[1], pb[2], pb[3]);
        const uint32_t vbase = (ks < 2 ? stA : stB) + ((ks & 1) * 16) * C::ROW_PITCH + warp * 128 + v_lane_off;
#pragma unroll
        for (int jp = 0; jp < 4; ++jp) {
          uint32_t r0, r1, r2, r3;
          ldsm_x4_t(vbase + jp * 32, r0, r1, r2, r3);
          mma_16816<BF16>(o[2 * jp], pa, r0, r1);
          mma_16816<BF16>(o[2 * jp + 1], pa, r2, r3);
          mma_16816<BF16>(o[2 * jp], pb, r0, r1);
          mma_16816<BF16>(o[2 * jp + 1], pb, r2, r3);
        }
      }
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&empty[sA]);
        mbar_arrive(&empty[sB]);
      }
    }

    // ---- segment end: row sums over the quad and the eight warps; the maxima are already common ----
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      l_part[r] += __shfl_xor_sync(0xffffffffu, l_part[r], 1);
      l_part[r] += __shfl_xor_sync(0xffffffffu, l_part[r], 2);
    }
    if (t == 0) {
      s_sum[warp * 16 + g] = l_part[0];
      s_sum[warp * 16 + g + 8] = l_part[1];
    }
    consumer_bar();
    float L[2] = {0.f, 0.f};
#pragma unroll
    for (int w = 0; w < kConsumerWarps; ++w) {
      L[0] += s_sum[w * 16 + g];
      L[1] += s_sum[w * 16 + g + 8];
    }
    const float inv0 = L[0] > 0.f ? 1.f / L[0] : 0.f, inv1 = L[1] > 0.f ? 1.f / L[1] : 0.f;
    if (whole_unit) {
      uint16_t* o0 = p.out + ((int64_t)b * p.Hq + row0 + g) * C::Dv + warp * 64 + 2 * t;
      uint16_t* o1 = o0 + 8 * C::Dv;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        if (g < rows_valid) *reinterpret_cast<uint32_t*>(o0 + j * 8) = pack_half2<BF16>(o[j][0] * inv0, o[j][1] * inv0);
        if (g + 8 < rows_valid) *reinterpret_cast<uint32_t*>(o1 + j * 8) = pack_half2<BF16>(o[j][2] * inv1, o[j][3] * inv1);
      }
    } else {
      const int prow = ((int)blockIdx.x * 2 + (first_seg ? 0 : 1)) * 16;
      float* o0 = p.part_o + (int64_t)(prow + g) * C::Dv + warp * 64 + 2 * t;
      float* o1 = o0 + 8 * C::Dv;
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        *reinterpret_cast<float2*>(o0 + j * 8) = make_float2(o[j][0] * inv0, o[j][1] * inv0);
        *reinterpret_cast<float2*>(o1 + j * 8) = make_float2(o[j][2] * inv1, o[j][3] * inv1);
      }
      if (warp == 0 && t == 0) {
        p.part_lse[prow + g] = L[0] > 0.f ? m_run[0] + log2f(L[0]) : -INFINITY;
        p.part_lse[prow + g + 8] = L[1] > 0.f ? m_run[1] + log2f(L[1]) : -INFINITY;
      }
    }
    cur.advance(n);
  }
}

static bool mla_path_ok(int Hkv, int D, int Dv, bool is_mla) { return is_mla && Hkv == 1 && D == MlaCfg::D && Dv == MlaCfg::Dv; }

static size_t mla_workspace_bytes() { return (size_t)attn_grid() * 2 * 16 * (MlaCfg::Dv + 1) * sizeof(float) + 256; }

template <bool BF16>
static int launch_mla(const DecodeKParams& kp, cudaStream_t st) {
  using C = MlaCfg;
  auto kern = decode_mla_mma_kernel<BF16>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    configured = true;
  }
  const int grid = attn_grid();
  kern<<<grid, kAttnThreads, C::SMEM_BYTES, st>>>(kp);
  if (int rc = launched("decode_mla_mma_kernel launch")) return rc;
  const int n_qblk = (kp.Hq + 15) / 16;
  return launch_merge_part<BF16>(kp, kp.Hq, 1, C::TILE_T, n_qblk, n_qblk, grid, st);
}
