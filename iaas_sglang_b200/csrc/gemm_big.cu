// Large-M (prefill / big-batch) FP8 block-scaled and INT8 GEMMs for sm_100a: persistent, warp-specialised,
// 256-wide tcgen05 tiles, optional CTA pairs (cta_group::2).
//
//   MODE 0  fp8_blockwise_scaled_mm  (sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:345-423)
//           out[m,n] = sum_kb sa[m,kb]*sb[kb,nb] * (A[m, kb*128:(kb+1)*128] . W[n, kb*128:(kb+1)*128])
//   MODE 1  int8_scaled_mm           (sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-747)
//           out[m,n] = ((float)sum_k A[m,k]*W[n,k] * sa[m]) * sb[n] + bias[n]
//
// Tile = (128 x PAIR) rows of A x 256 rows of W per CTA (pair).  Both operands are K-major with 128-byte
// (= one fp8 scale block) k-steps in the SWIZZLE_128B layout that tcgen05.mma reads directly.
//   PAIR = 1: one CTA per tile; it stages A [128 x 128 B] and W [256 x 128 B] per k-block.
//   PAIR = 2: two CTAs of a cluster share a 256 x 256 tile: each stages its own 128 rows of A and HALF of the W
//             rows, the leader issues tcgen05.mma.cta_group::2 (M = 256) which reads both halves; per CTA and
//             k-block 32 KB come from L2 instead of 48 KB.  (L2 -> SM bandwidth, not the tensor core, is what
//             bounds 128-wide tiles: 1.0 PFLOP/s measured with 128 x 128.)
// Warps: 0 = TMA producer, 1 = MMA issuer (one thread), 2 = TMEM allocator, 3 idle, 4..19 = epilogue.
// TMEM: two accumulator buffers of 256 fp32 columns.
//   FP8: UMMA has no fp32 128-block scaling, so every k-block is accumulated into its own buffer and the sixteen
//   epilogue warps promote it  acc += tmem * (sa[m,kb] * sb[kb,nb])  with packed FFMA2 while the tensor core
//   works on the next k-block (each warp: 32 rows x 64 columns).
//   INT8: int32 accumulation stays in TMEM for the whole K loop (exact); the two buffers alternate between tiles,
//   so the epilogue of tile i overlaps the main loop of tile i+1.
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

constexpr int GB_BM = 128;   // rows of A per CTA
constexpr int GB_BK = 128;   // bytes of K per k-block
// BN = output columns per tile (UMMA N): 256 (two TMEM buffers) or 128 (four); 4 * BN / 64 epilogue warps
constexpr int GB_MAX_STAGES = 8;

struct BigParams {
  const float* sa;
  int64_t sa_sm, sa_sk;
  const float* sb;
  int64_t sb_sk, sb_sn;
  const void* bias;
  void* out;
  int64_t ldo;
  int out_bf16;
  int M, N, K;
  int kb_total, tiles_m, tiles_n, n_tiles, group_m, workers;
  long long* prof;  // developer (B200_GB_PROF=1): per-CTA clock64 sums of the pipeline waits
};

// ---- cluster / cta_group::2 helpers ----
__device__ __forceinline__ uint32_t gb_cluster_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void gb_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t gb_mapa(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void gb_remote_arrive(uint32_t cluster_bar) {
  asm volatile("mbarrier.arrive.release.cluster.shared::cluster.b64 _, [%0];" ::"r"(cluster_bar) : "memory");
}
template <int PAIR>
__device__ __forceinline__ void gb_tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  if constexpr (PAIR == 1) {
    tmem_alloc(dst_smem, ncols);
  } else {
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
}
template <int PAIR>
__device__ __forceinline__ void gb_tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  if constexpr (PAIR == 1) tmem_dealloc(taddr, ncols);
  else asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
// TMA load whose completion is signalled on the mbarrier at cluster address `bar_cluster` (the pair leader's)
__device__ __forceinline__ void gb_tma_load_2d_pair(void* dst_smem, const CUtensorMap* m, int c0, int c1, uint32_t bar_cluster) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];" ::"r"(
          smem_u32(dst_smem)),
      "l"(m), "r"(bar_cluster), "r"(c0), "r"(c1)
      : "memory");
}
template <int PAIR, int MODE>
__device__ __forceinline__ void gb_umma(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  if constexpr (PAIR == 1) {
    if constexpr (MODE == 0) umma_f8f6f4(d_tmem, a_desc, b_desc, idesc, accumulate);
    else umma_i8(d_tmem, a_desc, b_desc, idesc, accumulate);
  } else if constexpr (MODE == 0) {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  } else {
    asm volatile(
        "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
        "tcgen05.mma.cta_group::2.kind::i8 [%0], %1, %2, %3, p;\n\t}\n" ::"r"(d_tmem),
        "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
  }
}
// arrive on `bar` (same offset in every CTA of the pair) once the MMAs issued so far have completed
template <int PAIR>
__device__ __forceinline__ void gb_commit(uint64_t* bar) {
  if constexpr (PAIR == 1) {
    tc_commit(bar);
  } else {
    asm volatile("tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(smem_u32(bar)),
                 "h"((uint16_t)3)
                 : "memory");
  }
}
// wait for this thread's tcgen05.ld results; the registers are operands so that no use can be scheduled above it
__device__ __forceinline__ void gb_ld_wait32(uint32_t (&r)[32]) {
  asm volatile("tcgen05.wait::ld.sync.aligned;"
               : "+r"(r[0]), "+r"(r[1]), "+r"(r[2]), "+r"(r[3]), "+r"(r[4]), "+r"(r[5]), "+r"(r[6]), "+r"(r[7]), "+r"(r[8]),
                 "+r"(r[9]), "+r"(r[10]), "+r"(r[11]), "+r"(r[12]), "+r"(r[13]), "+r"(r[14]), "+r"(r[15]), "+r"(r[16]),
                 "+r"(r[17]), "+r"(r[18]), "+r"(r[19]), "+r"(r[20]), "+r"(r[21]), "+r"(r[22]), "+r"(r[23]), "+r"(r[24]),
                 "+r"(r[25]), "+r"(r[26]), "+r"(r[27]), "+r"(r[28]), "+r"(r[29]), "+r"(r[30]), "+r"(r[31])::"memory");
}
__device__ __forceinline__ void gb_fma2(float2& acc, uint32_t x0, uint32_t x1, float s) {
#ifdef GB_PLAIN_FMA
  acc.x = fmaf(__uint_as_float(x0), s, acc.x);
  acc.y = fmaf(__uint_as_float(x1), s, acc.y);
  return;
#endif
  uint64_t x, sc;
  asm("mov.b64 %0, {%1, %2};" : "=l"(x) : "r"(x0), "r"(x1));
  asm("mov.b64 %0, {%1, %1};" : "=l"(sc) : "f"(s));
  asm("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(reinterpret_cast<uint64_t&>(acc)) : "l"(x), "l"(sc));
}

// tile index -> (m tile, n tile): m fastest inside groups of `group_m` tile rows (L2 reuse of the W panel)
__device__ __forceinline__ void gb_tile_coord(const BigParams& p, int t, int& tm, int& tn) {
  const int per_group = p.group_m * p.tiles_n;
  const int g = t / per_group, r = t - g * per_group;
  const int first_m = g * p.group_m;
  const int gm = p.tiles_m - first_m < p.group_m ? p.tiles_m - first_m : p.group_m;
  tn = r / gm;
  tm = first_m + (r - tn * gm);
}

template <int MODE, int PAIR, int BN>
__global__ void __launch_bounds__((4 + BN / 16) * 32, 1)
gemm_big_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                const __grid_constant__ BigParams p, const int stages) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int A_BYTES = GB_BM * GB_BK;                   // 16 KB
  constexpr int GB_BN = BN, GB_EPI_WARPS = BN / 16, NBUF = 512 / BN;
  constexpr int B_ROWS = GB_BN / PAIR;                     // W rows staged by this CTA
  constexpr int B_BYTES = B_ROWS * GB_BK;                  // 32 KB / 16 KB
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t kTmemCols = 512;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + stages * STAGE_BYTES);
  uint64_t* empty = full + GB_MAX_STAGES;
  uint64_t* tfull = empty + GB_MAX_STAGES;
  uint64_t* tempty = tfull + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 4);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int rank = PAIR == 2 ? (int)gb_cluster_rank() : 0;
  const int worker = blockIdx.x / PAIR;
  const bool leader = rank == 0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < NBUF; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], GB_EPI_WARPS * PAIR);
    }
    fence_mbar_init();
  }
  if (warp == 2) gb_tmem_alloc<PAIR>(tmem_slot, kTmemCols);
  tc_fence_before();
  if constexpr (PAIR == 2) gb_cluster_sync(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0 && lane == 0) {
    // ================================ TMA producer ================================
    const uint32_t full0_cluster = PAIR == 2 ? gb_mapa(smem_u32(&full[0]), 0) : 0;
    int it = 0;
    for (int t = worker; t < p.n_tiles; t += p.workers) {
      int tm, tn;
      gb_tile_coord(p, t, tm, tn);
      const int m0 = (tm * PAIR + rank) * GB_BM, n0 = tn * GB_BN + rank * B_ROWS;
      for (int kb = 0; kb < p.kb_total; ++kb, ++it) {
        const int s = it % stages;
        const uint32_t ph = (it / stages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        uint8_t* st = smem + s * STAGE_BYTES;
        if constexpr (PAIR == 1) {
          mbar_arrive_expect_tx(&full[s], STAGE_BYTES);
          tma_load_2d(st, &tma_a, kb * GB_BK, m0, &full[s]);
          tma_load_2d(st + A_BYTES, &tma_b, kb * GB_BK, n0, &full[s]);
        } else {
          // both CTAs' copies complete on the LEADER's barrier (it expects the bytes of the pair)
          if (leader) mbar_arrive_expect_tx(&full[s], 2 * STAGE_BYTES);
          const uint32_t bar = full0_cluster + (uint32_t)(s * 8);
          gb_tma_load_2d_pair(st, &tma_a, kb * GB_BK, m0, bar);
          gb_tma_load_2d_pair(st + A_BYTES, &tma_b, kb * GB_BK, n0, bar);
        }
      }
    }
  } else if (warp == 1 && lane == 0 && leader) {
    // ================================ MMA issuer (pair leader only) ================================
    const uint32_t idesc = MODE == 0 ? umma_idesc(1 /*f32*/, 0 /*e4m3*/, 0, GB_BM * PAIR, GB_BN)
                                     : umma_idesc(2 /*s32*/, 1 /*s8*/, 1, GB_BM * PAIR, GB_BN);
    int it = 0, tile_it = 0;
    long long pm_tempty = 0, pm_full = 0;
    const long long pm_t0 = clock64();
    for (int t = worker; t < p.n_tiles; t += p.workers, ++tile_it) {
      for (int kb = 0; kb < p.kb_total; ++kb, ++it) {
        const int s = it % stages;
        const uint32_t ph = (it / stages) & 1;
        const int a = MODE == 0 ? (it % NBUF) : (tile_it % NBUF);
        const long long c0 = p.prof ? clock64() : 0;
        if (MODE == 0) mbar_wait(&tempty[a], ((it / NBUF) & 1) ^ 1);
        else if (kb == 0) mbar_wait(&tempty[a], ((tile_it / NBUF) & 1) ^ 1);
        const long long c1 = p.prof ? clock64() : 0;
        mbar_wait(&full[s], ph);
        tc_fence_after();
        if (p.prof) {
          const long long c2 = clock64();
          pm_tempty += c1 - c0;
          pm_full += c2 - c1;
        }
        const uint32_t a_addr = smem_u32(smem + s * STAGE_BYTES);
        const uint64_t adesc = umma_desc_k_sw128(a_addr);
        const uint64_t bdesc = umma_desc_k_sw128(a_addr + A_BYTES);
        const uint32_t d = tmem_base + a * GB_BN;
#pragma unroll
        for (int k = 0; k < GB_BK / 32; ++k) {
          const uint32_t accum = MODE == 0 ? (k > 0) : (kb > 0 || k > 0);
          gb_umma<PAIR, MODE>(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
        }
        gb_commit<PAIR>(&empty[s]);
        if (MODE == 0 || kb == p.kb_total - 1) gb_commit<PAIR>(&tfull[a]);
      }
    }
    if (p.prof) {
      p.prof[blockIdx.x * 8 + 0] = clock64() - pm_t0;
      p.prof[blockIdx.x * 8 + 1] = pm_tempty;
      p.prof[blockIdx.x * 8 + 2] = pm_full;
      p.prof[blockIdx.x * 8 + 3] = it;
    }
  } else if (warp >= 4) {
    // ================================ epilogue warps ================================
    const int ew = warp - 4, quad = ew & 3, cg = ew >> 2;  // TMEM lanes [32*quad, +32), columns [64*cg, +64)
    const int row_in_tile = quad * 32 + lane;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + cg * 64;
    const uint32_t tempty0_leader = PAIR == 2 ? gb_mapa(smem_u32(&tempty[0]), 0) : 0;
    auto release = [&](int a) {  // this warp is done reading TMEM buffer a
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        if constexpr (PAIR == 1) mbar_arrive(&tempty[a]);
        else gb_remote_arrive(tempty0_leader + (uint32_t)(a * 8));
      }
    };
    int it = 0, tile_it = 0;
    long long pe_wait = 0, pe_ld = 0;
    const long long pe_t0 = clock64();
    for (int t = worker; t < p.n_tiles; t += p.workers, ++tile_it) {
      int tm, tn;
      gb_tile_coord(p, t, tm, tn);
      const int row = (tm * PAIR + rank) * GB_BM + row_in_tile;
      const int n0 = tn * GB_BN + cg * 64;
      float2 acc[32];
      if (MODE == 0) {
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = make_float2(0.f, 0.f);
        const int nb = n0 / 128;
        const bool nb_ok = n0 < p.N;
        // (raw values are kept and multiplied at the point of use: the product would wait for the loads right here)
        auto load_sa = [&](int kb) -> float { return row < p.M ? __ldg(p.sa + (int64_t)row * p.sa_sm + (int64_t)kb * p.sa_sk) : 0.f; };
        auto load_sb = [&](int kb) -> float { return nb_ok ? __ldg(p.sb + (int64_t)kb * p.sb_sk + (int64_t)nb * p.sb_sn) : 0.f; };
        // the scale loads come from an L2 that the TMA ring keeps saturated: they are issued four k-blocks ahead
        float saq[4], sbq[4];
#pragma unroll
        for (int d = 0; d < 4; ++d) {
          saq[d] = d < p.kb_total ? load_sa(d) : 0.f;
          sbq[d] = d < p.kb_total ? load_sb(d) : 0.f;
        }
        for (int kbq = 0; kbq < p.kb_total; kbq += 4) {
#pragma unroll
          for (int d = 0; d < 4; ++d) {
            const int kb = kbq + d;
            if (kb < p.kb_total) {
              const int a = it % NBUF;
              const float sc = saq[d] * sbq[d];
              if (kb + 4 < p.kb_total) {
                saq[d] = load_sa(kb + 4);
                sbq[d] = load_sb(kb + 4);
              }
#ifdef GB_PROF_EPI
              const long long e0 = p.prof ? clock64() : 0;
#endif
              mbar_wait(&tfull[a], (it / NBUF) & 1);
              tc_fence_after();
#ifdef GB_PROF_EPI
              const long long e1 = p.prof ? clock64() : 0;
#endif
              uint32_t r0[32], r1[32];
#ifdef GB_SKIP_LD
#pragma unroll
              for (int j = 0; j < 32; ++j) r0[j] = r1[j] = (uint32_t)(it + j);
#else
              tmem_ld_32x32(lane_addr + a * GB_BN, r0);
              tmem_ld_32x32(lane_addr + a * GB_BN + 32, r1);
              gb_ld_wait32(r0);
              gb_ld_wait32(r1);
#endif
              release(a);
#ifdef GB_PROF_EPI
              if (p.prof) {
                const long long e2 = clock64();
                pe_wait += e1 - e0;
                pe_ld += e2 - e1;
              }
#endif
#pragma unroll
              for (int j = 0; j < 16; ++j) gb_fma2(acc[j], r0[2 * j], r0[2 * j + 1], sc);
#pragma unroll
              for (int j = 0; j < 16; ++j) gb_fma2(acc[16 + j], r1[2 * j], r1[2 * j + 1], sc);
              ++it;
            }
          }
        }
      } else {
        const int a = tile_it % NBUF;
        mbar_wait(&tfull[a], (tile_it / NBUF) & 1);
        tc_fence_after();
        uint32_t r0[32], r1[32];
        tmem_ld_32x32(lane_addr + a * GB_BN, r0);
        tmem_ld_32x32(lane_addr + a * GB_BN + 32, r1);
        gb_ld_wait32(r0);
        gb_ld_wait32(r1);
        release(a);
        const float sa_row = row < p.M ? __ldg(p.sa + row) : 0.f;
#pragma unroll
        for (int j = 0; j < 32; ++j) {
          const uint32_t x0 = j < 16 ? r0[2 * j] : r1[2 * j - 32], x1 = j < 16 ? r0[2 * j + 1] : r1[2 * j - 31];
          const int n = n0 + 2 * j;
          const float s0 = n < p.N ? __ldg(p.sb + n) : 0.f, s1 = n + 1 < p.N ? __ldg(p.sb + n + 1) : 0.f;
          float v0 = __fmul_rn(__fmul_rn((float)(int)x0, sa_row), s0), v1 = __fmul_rn(__fmul_rn((float)(int)x1, sa_row), s1);
          if (p.bias) {
            const uint16_t* bp = reinterpret_cast<const uint16_t*>(p.bias);
            const uint16_t u0 = n < p.N ? __ldg(bp + n) : (uint16_t)0, u1 = n + 1 < p.N ? __ldg(bp + n + 1) : (uint16_t)0;
            v0 = __fadd_rn(v0, p.out_bf16 ? half_to_float<true>(u0) : half_to_float<false>(u0));
            v1 = __fadd_rn(v1, p.out_bf16 ? half_to_float<true>(u1) : half_to_float<false>(u1));
          }
          acc[j] = make_float2(v0, v1);
        }
      }
      // ---- write 64 output columns of this thread's row ----
      if (row < p.M && n0 < p.N) {
        uint16_t* orow = reinterpret_cast<uint16_t*>(p.out) + (size_t)row * p.ldo + n0;
        const bool vec_ok = (n0 + 64 <= p.N) && (p.ldo % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
        if (vec_ok) {
#pragma unroll
          for (int j = 0; j < 32; j += 4) {
            uint4 pk;
            if (p.out_bf16) {
              pk.x = pack_half2<true>(acc[j].x, acc[j].y); pk.y = pack_half2<true>(acc[j + 1].x, acc[j + 1].y);
              pk.z = pack_half2<true>(acc[j + 2].x, acc[j + 2].y); pk.w = pack_half2<true>(acc[j + 3].x, acc[j + 3].y);
            } else {
              pk.x = pack_half2<false>(acc[j].x, acc[j].y); pk.y = pack_half2<false>(acc[j + 1].x, acc[j + 1].y);
              pk.z = pack_half2<false>(acc[j + 2].x, acc[j + 2].y); pk.w = pack_half2<false>(acc[j + 3].x, acc[j + 3].y);
            }
            *reinterpret_cast<uint4*>(orow + 2 * j) = pk;
          }
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) {
            if (n0 + 2 * j < p.N) orow[2 * j] = p.out_bf16 ? float_to_half<true>(acc[j].x) : float_to_half<false>(acc[j].x);
            if (n0 + 2 * j + 1 < p.N) orow[2 * j + 1] = p.out_bf16 ? float_to_half<true>(acc[j].y) : float_to_half<false>(acc[j].y);
          }
        }
      }
      if (p.prof && threadIdx.x == 128 && t + p.workers >= p.n_tiles) {
        p.prof[blockIdx.x * 8 + 4] = clock64() - pe_t0;
        p.prof[blockIdx.x * 8 + 5] = pe_wait;
        p.prof[blockIdx.x * 8 + 6] = pe_ld;
      }
    }
  }

  tc_fence_before();
  __syncwarp();
  if constexpr (PAIR == 2) gb_cluster_sync(); else __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    gb_tmem_dealloc<PAIR>(tmem_base, kTmemCols);
  }
}

template <int MODE, int PAIR, int BN>
static int launch_big_t(const void* a, int64_t lda, const void* w, int64_t ldw, BigParams& bp, cudaStream_t st) {
  constexpr int GB_BN = BN, GB_THREADS = (4 + BN / 16) * 32;
  constexpr int STAGE_BYTES = GB_BM * GB_BK + (GB_BN / PAIR) * GB_BK;
  const int stages = (226 * 1024 - 1024 - 512) / STAGE_BYTES > GB_MAX_STAGES ? GB_MAX_STAGES : (226 * 1024 - 1024 - 512) / STAGE_BYTES;
  const int smem = stages * STAGE_BYTES + 1024 + 512;
  bp.kb_total = (bp.K + GB_BK - 1) / GB_BK;
  bp.tiles_m = (bp.M + GB_BM * PAIR - 1) / (GB_BM * PAIR);
  bp.tiles_n = (bp.N + GB_BN - 1) / GB_BN;
  bp.n_tiles = bp.tiles_m * bp.tiles_n;
  bp.group_m = 16 / PAIR;
  const int sms = device_sm_count() > 0 ? device_sm_count() : 148;
  int workers = sms / PAIR;
  if (workers > bp.n_tiles) workers = bp.n_tiles;
  bp.workers = workers;
  CUtensorMap ta, tb;
  if (int rc = make_tmap_2d(&ta, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, bp.M, bp.K, lda, GB_BM, GB_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  if (int rc = make_tmap_2d(&tb, w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, bp.N, bp.K, ldw, GB_BN / PAIR, GB_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  auto kern = gemm_big_kernel<MODE, PAIR, BN>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(workers * PAIR);
  cfg.blockDim = dim3(GB_THREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeClusterDimension;
  attr[0].val.clusterDim.x = PAIR;
  attr[0].val.clusterDim.y = 1;
  attr[0].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = PAIR > 1 ? 1 : 0;
  const int prof = tune(TUNE_GB_PROF);
  static long long* dprof = nullptr;
  if (prof && !dprof) cudaMalloc(&dprof, 160 * 64);
  bp.prof = prof ? dprof : nullptr;
  B200_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, bp, stages));
  if (prof) {
    static int n = 0;
    cudaStreamSynchronize(st);
    long long h[160 * 8];
    cudaMemcpy(h, dprof, sizeof(h), cudaMemcpyDeviceToHost);
    if (n++ % 8 == 3) {
      double s[8] = {0};
      const int g = workers * PAIR;
      for (int c = 0; c < g; ++c) for (int i = 0; i < 8; ++i) s[i] += (double)h[c * 8 + i] / g;
      fprintf(stderr, "[gemm_big prof] M=%d N=%d K=%d BN=%d: MMA thread total %.0f cyc, wait tempty %.0f, wait full %.0f, k-blocks %.0f | epi warp total %.0f, wait tfull %.0f, ld+release %.0f\n",
              bp.M, bp.N, bp.K, BN, s[0], s[1], s[2], s[3], s[4], s[5], s[6]);
    }
  }
  return launched("gemm_big_kernel launch");
}

// a [M,K] activations, w [N,K] weight (both K-major, 1-byte elements); mode 0 fp8 block-scaled, 1 int8
int launch_big(int mode, const void* a, int64_t lda, const void* w, int64_t ldw, const float* sa, int64_t sa_sm, int64_t sa_sk,
               const float* sb, int64_t sb_sk, int64_t sb_sn, const void* bias, void* out, int64_t ldo, int out_bf16, int M,
               int N, int K, cudaStream_t st) {
  BigParams bp{};
  bp.sa = sa; bp.sa_sm = sa_sm; bp.sa_sk = sa_sk;
  bp.sb = sb; bp.sb_sk = sb_sk; bp.sb_sn = sb_sn;
  bp.bias = bias; bp.out = out; bp.ldo = ldo; bp.out_bf16 = out_bf16;
  bp.M = M; bp.N = N; bp.K = K;
  const int pair = tune(TUNE_GB_PAIR), bn = tune(TUNE_GB_BN);
  if (mode == 0) {
    // fp8: the per-k-block TMEM hand-shake with remote CTAs is too slow for pairs (464 vs 1190 TFLOP/s measured).
    // 128-wide tiles have four TMEM buffers, so the tensor core never waits for the promotion (the MMA thread is then
    // issue-bound: a 128x128x32 instruction costs the same ~175 cycles as a 128x256x32 one); 256-wide tiles have two
    // buffers and wait ~980 cycles per k-block for the promotion, whose tcgen05.ld shares the TMEM port with the
    // running MMA (tools/gb_prof.py).  1.25 vs 1.19 PFLOP/s.
    if (bn == 128) return launch_big_t<0, 1, 128>(a, lda, w, ldw, bp, st);
    return launch_big_t<0, 1, 256>(a, lda, w, ldw, bp, st);
  }
  if (pair != 1 && M > GB_BM) return launch_big_t<1, 2, 256>(a, lda, w, ldw, bp, st);
  return launch_big_t<1, 1, 256>(a, lda, w, ldw, bp, st);
}

}  // namespace b200
