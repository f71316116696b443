// Decode-shape ("skinny", M <= 256) quantised GEMMs for sm_100a: swap-AB tcgen05 + stream-K.
//
//   MODE 0  fp8_blockwise_scaled_mm  (sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:345-423)
//   MODE 1  int8_scaled_mm           (sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-747)
//
// At decode M (64..256 tokens) these GEMMs stream the weight once and are bound by HBM, not by the tensor
// core, so the kernel is built like the attention kernel: keep the copy engine saturated on every SM.
//   * swap-AB: the WEIGHT tile [128 n x 128 k-bytes] is the UMMA "A" operand (M = 128 lanes of TMEM), the
//     activations [mt_pad x 128] are the "B" operand (UMMA N = M rounded up to 16): no MMA rows are wasted on
//     padding and the accumulator is only mt_pad columns wide;
//   * stream-K: the (n-tile, k-block) units are laid out tile-major and CTA c of `grid` (one per SM) owns units
//     [c*U/grid, (c+1)*U/grid): every SM streams the same number of weight bytes for ANY (N, K); a tile whose
//     k-blocks are spread over several CTAs meets in an fp32 workspace (two partial slots per CTA) and the last
//     CTA to arrive (ticket) sums the partials in fixed CTA order (bit-reproducible) and writes the tile;
//   * warp 0 = TMA producer (weights evict-first), warp 1 = single-thread tcgen05.mma issuer, warp 2 = TMEM
//     allocator, warps 4.. = epilogue (one TMEM lane = one output column n per thread).
//   FP8: every 128-deep k-block lands in its own TMEM buffer (double buffered) and is promoted
//     acc[m] += tmem[m] * (sa[m,kb] * sb[kb,nb]) in fp32 registers (UMMA has no fp32 block scaling);
//   INT8: int32 accumulation in TMEM over the CTA's k-range of the tile (exact), scales in the epilogue.
#include <stdlib.h>

#include "gemm_skinny.h"

namespace b200 {

constexpr int SK_BN = 128;  // weight rows per tile (UMMA M)
constexpr int SK_BK = 128;  // K bytes per pipeline stage (one 128-byte swizzle row, one fp8 scale block)
constexpr int SK_MAX_STAGES = 12;
constexpr int SK_SMEM_LIMIT = 232448;


__device__ __forceinline__ long long sk_start(long long c, long long U, long long G) { return c * U / G; }
__device__ __forceinline__ long long sk_owner(long long u, long long U, long long G) { return ((u + 1) * G - 1) / U; }

template <int GROUPS>
__device__ __forceinline__ void sk_epi_bar_all() {
  asm volatile("bar.sync 1, %0;" ::"n"(GROUPS * 128) : "memory");
}
__device__ __forceinline__ void sk_epi_bar_group(int grp) {
  asm volatile("bar.sync %0, 128;" ::"r"(2 + grp) : "memory");
}

template <int MODE, int COLS, int GROUPS>
__global__ void __launch_bounds__((4 + 4 * GROUPS) * 32, 1)
gemm_skinny_kernel(const __grid_constant__ CUtensorMap tma_w, const __grid_constant__ CUtensorMap tma_x,
                   const __grid_constant__ SkinnyParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  const int x_bytes = p.mt_pad * SK_BK;
  const int stage_bytes = SK_BN * SK_BK + x_bytes;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.stages * stage_bytes);
  uint64_t* empty = full + SK_MAX_STAGES;
  uint64_t* tfull = empty + SK_MAX_STAGES;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  int* flag = reinterpret_cast<int*>(tmem_slot + 1);
  float* sbuf = reinterpret_cast<float*>(tmem_slot + 4);  // MODE 0: [GROUPS][2][COLS]; MODE 1: sa[256]
  int* s_off = reinterpret_cast<int*>(sbuf + 512);        // fix-up: partial-slot index of every contributor (<= 256)

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  constexpr int BUFCOLS = GROUPS * COLS;
  constexpr uint32_t kTmemCols = 2 * BUFCOLS;
  const long long U = p.units, G = gridDim.x;
  const long long u0 = sk_start(blockIdx.x, U, G), u1 = sk_start(blockIdx.x + 1, U, G);
  const int n_units = (int)(u1 - u0);

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_w);
    tma_prefetch_desc(&tma_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], GROUPS * 128);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();  // let the next kernel of the stream start its prologue (no-op without PDL)

  if (warp != 1 && warp < 4) {
    // ================================ TMA producers ================================
    // One thread sustains only about one tensor-map copy per ~0.5 us (tools/tma_probe.cu: 148 CTAs streaming
    // 8 KB boxes reach 2.4 TB/s with one issuing warp, 4.3 with two, 5.1 with four), so three warps issue:
    // unit i belongs to producer i % 3.
    constexpr int NP = 3;
    const int pidx = warp == 0 ? 0 : warp - 1;
    if (lane == 0) {
      const uint64_t pol_w = make_policy_evict_first();
      // weights do not depend on the previous kernel: fill the ring with them before waiting for it
      const int pre = n_units < p.stages ? n_units : p.stages;
      for (int i = pidx; i < pre; i += NP) {
        const long long u = u0 + i;
        const int tile = (int)(u / p.kb_total), kb = (int)(u - (long long)tile * p.kb_total);
        mbar_arrive_expect_tx(&full[i], stage_bytes);
        tma_load_2d_hint(smem + i * stage_bytes, &tma_w, kb * SK_BK, tile * SK_BN, &full[i], pol_w);
      }
      griddep_wait();
      for (int i = pidx; i < pre; i += NP) {
        const int kb = (int)((u0 + i) % p.kb_total);
        tma_load_2d(smem + i * stage_bytes + SK_BN * SK_BK, &tma_x, kb * SK_BK, 0, &full[i]);
      }
      int i = pidx;
      while (i < pre) i += NP;
      long long u = u0 + i;
      int tile = (int)(u / p.kb_total), kb = (int)(u - (long long)tile * p.kb_total);
      for (; i < n_units; i += NP) {
        const int s = i % p.stages;
        const uint32_t ph = (i / p.stages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_arrive_expect_tx(&full[s], stage_bytes);
        uint8_t* st = smem + s * stage_bytes;
        tma_load_2d_hint(st, &tma_w, kb * SK_BK, tile * SK_BN, &full[s], pol_w);
        tma_load_2d(st + SK_BN * SK_BK, &tma_x, kb * SK_BK, 0, &full[s]);
        kb += NP;
        while (kb >= p.kb_total) {
          kb -= p.kb_total;
          ++tile;
        }
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (lane == 0) {
      const uint32_t idesc = MODE == 0 ? umma_idesc(1 /*f32*/, 0 /*e4m3*/, 0, SK_BN, p.mt_pad)
                                       : umma_idesc(2 /*s32*/, 1 /*s8*/, 1, SK_BN, p.mt_pad);
      int seg = -1;
      int kb = (int)(u0 % p.kb_total) - 1;
      for (int i = 0; i < n_units; ++i) {
        const int s = i % p.stages;
        const uint32_t ph = (i / p.stages) & 1;
        if (++kb == p.kb_total) kb = 0;
        const bool seg_first = (i == 0 || kb == 0), seg_last = (i == n_units - 1 || kb == p.kb_total - 1);
        if (seg_first) ++seg;
        const int a = MODE == 0 ? (i & 1) : (seg & 1);
        if (MODE == 0) mbar_wait(&tempty[a], ((i >> 1) & 1) ^ 1);
        else if (seg_first) mbar_wait(&tempty[a], ((seg >> 1) & 1) ^ 1);
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t w_addr = smem_u32(smem + s * stage_bytes);
        const uint64_t adesc = umma_desc_k_sw128(w_addr);
        const uint64_t bdesc = umma_desc_k_sw128(w_addr + SK_BN * SK_BK);
        const uint32_t d = tmem_base + a * BUFCOLS;
#pragma unroll
        for (int k = 0; k < SK_BK / 32; ++k) {
          const uint32_t accum = MODE == 0 ? (k > 0) : (!seg_first || k > 0);
          if (MODE == 0) umma_f8f6f4(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
          else umma_i8(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
        }
        tc_commit(&empty[s]);
        if (MODE == 0 || seg_last) tc_commit(&tfull[a]);
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ================================ epilogue warps ================================
    const int ew = warp - 4, grp = ew >> 2, quad = ew & 3;  // quad == warp % 4: TMEM lanes [32*quad, +32)
    const int et = (ew & 3) * 32 + lane;                    // thread index inside the group, 0..127
    const int row = quad * 32 + lane;                       // output column n inside the tile
    const int col0 = grp * COLS;                            // first token handled by this group
    int ncols = p.mt_pad - col0;
    ncols = ncols < 0 ? 0 : (ncols > COLS ? COLS : ncols);
    const int nchunk = (ncols + 31) / 32;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + col0;
    float acc[COLS];
    griddep_wait();  // activation scales, workspace and the output buffer belong to the stream order

    if (MODE == 1) {  // per-token scales once
      for (int m = threadIdx.x - 128; m < 256; m += GROUPS * 128) sbuf[m] = m < p.M ? __ldg(p.sa + m) : 0.f;
      sk_epi_bar_all<GROUPS>();
    }

    // out[m, n] for this thread's column n of tile t from acc[] (MODE 1: int32 bits -> scales, bias)
    auto write_tile = [&](int t) {
      const int n = t * SK_BN + row;
      if (n >= p.N) return;
      uint16_t* ocol = reinterpret_cast<uint16_t*>(p.out) + n;
      float sbn = 1.f, bv = 0.f;
      if (MODE == 1) {
        sbn = __ldg(p.sb + n);
        if (p.bias) {
          const uint16_t bb = __ldg(reinterpret_cast<const uint16_t*>(p.bias) + n);
          bv = p.out_bf16 ? half_to_float<true>(bb) : half_to_float<false>(bb);
        }
      }
#pragma unroll
      for (int j = 0; j < COLS; ++j) {
        const int m = col0 + j;
        if (j < ncols && m < p.M) {
          float v = acc[j];
          if (MODE == 1) {
            v = __fmul_rn(__fmul_rn((float)__float_as_int(acc[j]), sbuf[m]), sbn);
            if (p.bias) v = __fadd_rn(v, bv);
          }
          ocol[(size_t)m * p.ldo] = p.out_bf16 ? float_to_half<true>(v) : float_to_half<false>(v);
        }
      }
    };
    int pending[2], n_pending = 0;  // tiles this CTA holds a partial of (head and/or tail of its unit range)

    int seg = -1;
    long long seg_u0 = u0;
    int tile = (int)(u0 / p.kb_total), kb = (int)(u0 - (long long)tile * p.kb_total);
    // fp8: the scale of the NEXT unit is loaded one iteration ahead (two dependent L2 reads otherwise sit on the
    // critical path of every k-block)
    auto load_scale = [&](int t_, int kb_) -> float {
      const int m = col0 + et;
      if (et >= COLS || m >= p.M) return 0.f;
      return __ldg(p.sa + (int64_t)m * p.sa_sm + (int64_t)kb_ * p.sa_sk) * __ldg(p.sb + (int64_t)kb_ * p.sb_sk + (int64_t)t_ * p.sb_sn);
    };
    float s_cur = 0.f;
    if (MODE == 0 && n_units > 0) s_cur = load_scale(tile, kb);
    for (int i = 0; i < n_units; ++i) {
      const long long u = u0 + i;
      if (i > 0 && ++kb == p.kb_total) {
        kb = 0;
        ++tile;
      }
      const bool seg_first = (i == 0 || kb == 0), seg_last = (i == n_units - 1 || kb == p.kb_total - 1);
      if (seg_first) {
        ++seg;
        seg_u0 = u;
        if (MODE == 0) {
#pragma unroll
          for (int j = 0; j < COLS; ++j) acc[j] = 0.f;
        }
      }
      if (MODE == 0) {
        // scale vector of this k-block for the group's tokens: s[m] = sa[m,kb] * sb[kb,tile]
        float* sv = sbuf + (grp * 2 + (i & 1)) * COLS;
        if (et < COLS) sv[et] = s_cur;
        if (i + 1 < n_units) {
          const bool wrap = (kb + 1 == p.kb_total);
          s_cur = load_scale(wrap ? tile + 1 : tile, wrap ? 0 : kb + 1);
        }
        sk_epi_bar_group(grp);
        const int a = i & 1;
        mbar_wait(&tfull[a], (i >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < COLS / 32; ++c) {
          if (c < nchunk && !(p.debug & 2)) {
            uint32_t r[32];
            tmem_ld_32x32(lane_addr + a * BUFCOLS + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 s4 = *reinterpret_cast<const float4*>(sv + c * 32 + j);
              acc[c * 32 + j + 0] = fmaf(__uint_as_float(r[j + 0]), s4.x, acc[c * 32 + j + 0]);
              acc[c * 32 + j + 1] = fmaf(__uint_as_float(r[j + 1]), s4.y, acc[c * 32 + j + 1]);
              acc[c * 32 + j + 2] = fmaf(__uint_as_float(r[j + 2]), s4.z, acc[c * 32 + j + 2]);
              acc[c * 32 + j + 3] = fmaf(__uint_as_float(r[j + 3]), s4.w, acc[c * 32 + j + 3]);
            }
          }
        }
        tc_fence_before();
        mbar_arrive(&tempty[a]);
      }
      if (!seg_last) continue;

      if (MODE == 1) {
        const int a = seg & 1;
        mbar_wait(&tfull[a], (seg >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < COLS / 32; ++c) {
          if (c < nchunk) {
            uint32_t r[32];
            tmem_ld_32x32(lane_addr + a * BUFCOLS + c * 32, r);
            tmem_ld_wait();
#pragma unroll
            for (int j = 0; j < 32; ++j) acc[c * 32 + j] = __uint_as_float(r[j]);  // raw int32 bits
          }
        }
        tc_fence_before();
        mbar_arrive(&tempty[a]);
      }

      // ---- segment done: a whole tile is written now; a partial one goes to this CTA's workspace slot and
      //      its ticket / reduction is deferred to the end of the CTA's work, so that the (several dependent
      //      L2 round trips of the) fix-up never stall the k-loop: measured 27-37 us -> see DESIGN.md ----
      const bool whole = (seg_u0 == (long long)tile * p.kb_total) && (kb == p.kb_total - 1);
      if (whole) {
        write_tile(tile);
      } else if (!(p.debug & 1)) {
        const int slot = (seg_u0 == u0) ? 0 : 1;
        float* mine = p.ws_partial + ((size_t)(blockIdx.x * 2 + slot) * p.mt_pad + col0) * SK_BN + row;
        if (!(p.debug & 8)) {
#pragma unroll
          for (int j = 0; j < COLS; ++j)
            if (j < ncols) mine[(size_t)j * SK_BN] = acc[j];
        }
        pending[n_pending++] = tile;
      }
    }

    if (n_pending > 0) {
      if (!(p.debug & 4)) __threadfence();
      sk_epi_bar_all<GROUPS>();
      if (threadIdx.x == 128) {
        int n_reduce = 0;
        for (int q = 0; q < n_pending; ++q) {
          const int t = pending[q];
          const long long uf = (long long)t * p.kb_total;
          const int contributors = (int)(sk_owner(uf + p.kb_total - 1, U, G) - sk_owner(uf, U, G)) + 1;
          const int old = atomicAdd(p.ws_counter + t, 1);
          if (old == contributors - 1) {   // last to arrive: this CTA reduces the tile
            p.ws_counter[t] = 0;           // self-reset for the next launch
            flag[1 + n_reduce++] = t;
          }
        }
        flag[0] = n_reduce;
        if (n_reduce) __threadfence();
      }
    } else if (threadIdx.x == 128) {
      flag[0] = 0;
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }

  // ---- fix-up of the tiles this CTA arrived last on: ALL threads reduce the tile as a flat [mt_pad][128]
  //      array, 4 output columns per thread and step, 4 x 4 independent 16-byte loads in flight per thread.
  //      (Summing into the row-owner registers issued the loads one by one at 255 registers: 28 us.) ----
  const int n_reduce = flag[0];
  constexpr int NT = (4 + 4 * GROUPS) * 32;
  for (int q = 0; q < n_reduce && !(p.debug & 16); ++q) {
    const int t = flag[1 + q];
    const long long uf = (long long)t * p.kb_total;
    const int c_first = (int)sk_owner(uf, U, G), c_last = (int)sk_owner(uf + p.kb_total - 1, U, G);
    const int nc = c_last - c_first + 1;
    __syncthreads();  // s_off is reused between tiles
    for (int i = threadIdx.x; i < nc; i += NT) {
      const int c = c_first + i;
      s_off[i] = c * 2 + ((uf <= sk_start(c, U, G)) ? 0 : 1);
    }
    __syncthreads();
    const int nvec = p.mt_pad * (SK_BN / 4);
    const bool vec_ok = (p.ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 7) == 0);
    for (int vb = threadIdx.x; vb < nvec; vb += 4 * NT) {
      float4 sum[4];
#pragma unroll
      for (int k = 0; k < 4; ++k) sum[k] = make_float4(0.f, 0.f, 0.f, 0.f);
      for (int i0 = 0; i0 < nc; i0 += 4) {  // contributors in fixed order: the sum is bit-reproducible
        float4 x[4][4];
#pragma unroll
        for (int ii = 0; ii < 4; ++ii) {
          const size_t base = (size_t)s_off[i0 + ii < nc ? i0 + ii : nc - 1] * p.mt_pad * SK_BN;
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            const int v = vb + k * NT;
            x[k][ii] = (v < nvec && i0 + ii < nc)
                           ? __ldcg(reinterpret_cast<const float4*>(p.ws_partial + base + (size_t)(v >> 5) * SK_BN + (v & 31) * 4))
                           : make_float4(0.f, 0.f, 0.f, 0.f);
          }
        }
#pragma unroll
        for (int ii = 0; ii < 4; ++ii)
#pragma unroll
          for (int k = 0; k < 4; ++k) {
            if (MODE == 0) {
              sum[k].x += x[k][ii].x; sum[k].y += x[k][ii].y; sum[k].z += x[k][ii].z; sum[k].w += x[k][ii].w;
            } else {
              sum[k].x = __int_as_float(__float_as_int(sum[k].x) + __float_as_int(x[k][ii].x));
              sum[k].y = __int_as_float(__float_as_int(sum[k].y) + __float_as_int(x[k][ii].y));
              sum[k].z = __int_as_float(__float_as_int(sum[k].z) + __float_as_int(x[k][ii].z));
              sum[k].w = __int_as_float(__float_as_int(sum[k].w) + __float_as_int(x[k][ii].w));
            }
          }
      }
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        const int v = vb + k * NT;
        if (v >= nvec) continue;
        const int m = v >> 5, n = t * SK_BN + (v & 31) * 4;
        if (m >= p.M || n >= p.N) continue;
        float o4[4] = {sum[k].x, sum[k].y, sum[k].z, sum[k].w};
        if (MODE == 1) {
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int ne = n + e < p.N ? n + e : p.N - 1;
            float val = __fmul_rn(__fmul_rn((float)__float_as_int(o4[e]), sbuf[m]), __ldg(p.sb + ne));
            if (p.bias) {
              const uint16_t bb = __ldg(reinterpret_cast<const uint16_t*>(p.bias) + ne);
              val = __fadd_rn(val, p.out_bf16 ? half_to_float<true>(bb) : half_to_float<false>(bb));
            }
            o4[e] = val;
          }
        }
        uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + (size_t)m * p.ldo + n;
        if (vec_ok && n + 4 <= p.N) {
          uint2 pk;
          pk.x = p.out_bf16 ? pack_half2<true>(o4[0], o4[1]) : pack_half2<false>(o4[0], o4[1]);
          pk.y = p.out_bf16 ? pack_half2<true>(o4[2], o4[3]) : pack_half2<false>(o4[2], o4[3]);
          *reinterpret_cast<uint2*>(dst) = pk;
        } else {
#pragma unroll
          for (int e = 0; e < 4; ++e)
            if (n + e < p.N) dst[e] = p.out_bf16 ? float_to_half<true>(o4[e]) : float_to_half<false>(o4[e]);
        }
      }
    }
  }
}

}  // namespace b200

namespace b200 {

static int skinny_grid(long long units) {
  int g = device_sm_count() > 0 ? device_sm_count() : 148;
  // at least ~4 k-blocks per CTA: below that the fix-up traffic outweighs the extra parallelism
  const long long cap = units / 4 > 0 ? units / 4 : 1;
  if (g > cap) g = (int)cap;
  return g;
}

static int skinny_mt_pad(int M) { return (M + 15) / 16 * 16; }

size_t skinny_workspace_bytes(int M, int N, int K) {
  const int mt = skinny_mt_pad(M);
  const int n_tiles = (N + SK_BN - 1) / SK_BN;
  const long long units = (long long)n_tiles * ((K + SK_BK - 1) / SK_BK);
  const size_t counters = ((size_t)n_tiles * 4 + 255) & ~(size_t)255;
  return counters + (size_t)skinny_grid(units) * 2 * mt * SK_BN * 4;
}

bool skinny_applicable(int M) { return M <= 256; }

template <int MODE, int COLS, int GROUPS>
static int launch_skinny_t(const CUtensorMap& tw, const CUtensorMap& tx, const SkinnyParams& sp, int grid, int smem,
                           cudaStream_t st) {
  auto kern = gemm_skinny_kernel<MODE, COLS, GROUPS>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, SK_SMEM_LIMIT));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3((4 + 4 * GROUPS) * 32);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, sp));
  return launched("gemm_skinny_kernel launch");
}

// a [M,K] activations, w [N,K] weight (both K-major, 1-byte elements)
template <int MODE>
int launch_skinny(const void* a, int64_t lda, const void* w, int64_t ldw, SkinnyParams& sp, void* workspace,
                  size_t workspace_bytes, cudaStream_t st) {
  sp.mt_pad = skinny_mt_pad(sp.M);
  sp.n_tiles = (sp.N + SK_BN - 1) / SK_BN;
  sp.kb_total = (sp.K + SK_BK - 1) / SK_BK;
  sp.units = (long long)sp.n_tiles * sp.kb_total;
  const int grid = skinny_grid(sp.units);
  const int stage_bytes = SK_BN * SK_BK + sp.mt_pad * SK_BK;
  int stages = (SK_SMEM_LIMIT - 1024 - 4096) / stage_bytes;
  if (stages > SK_MAX_STAGES) stages = SK_MAX_STAGES;
  sp.stages = stages;
  {
    static int dbg = -1;
    if (dbg < 0) { const char* e = getenv("B200_SK_DEBUG"); dbg = e ? atoi(e) : 0; }
    sp.debug = dbg;
  }
  const int smem = stages * stage_bytes + 1024 + 4096;
  const size_t need = skinny_workspace_bytes(sp.M, sp.N, sp.K);
  B200_REQUIRE(workspace && workspace_bytes >= need && ((uintptr_t)workspace & 15) == 0, B200_ERR_ARG,
               "gemm: workspace too small or misaligned (need %zu bytes, have %zu)", need, workspace_bytes);
  sp.ws_counter = reinterpret_cast<int*>(workspace);
  sp.ws_partial = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + (((size_t)sp.n_tiles * 4 + 255) & ~(size_t)255));
  CUtensorMap tw, tx;
  if (int rc = make_tmap_2d(&tw, w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, sp.N, sp.K, ldw, SK_BN, SK_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  if (int rc = make_tmap_2d(&tx, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, sp.M, sp.K, lda, sp.mt_pad, SK_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  if (sp.mt_pad <= 64) return launch_skinny_t<MODE, 64, 1>(tw, tx, sp, grid, smem, st);
  if (sp.mt_pad <= 128) return launch_skinny_t<MODE, 128, 1>(tw, tx, sp, grid, smem, st);
  return launch_skinny_t<MODE, 64, 4>(tw, tx, sp, grid, smem, st);
}

template int launch_skinny<0>(const void*, int64_t, const void*, int64_t, SkinnyParams&, void*, size_t, cudaStream_t);
template int launch_skinny<1>(const void*, int64_t, const void*, int64_t, SkinnyParams&, void*, size_t, cudaStream_t);

}  // namespace b200
