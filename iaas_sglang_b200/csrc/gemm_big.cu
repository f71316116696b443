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
// TMEM: 512 / BN accumulator buffers of BN fp32 columns.
//   FP8: UMMA has no fp32 128-block scaling, so every k-block is accumulated into its own buffer and the epilogue
//   warps promote it  acc += tmem * (sa[m,kb] * sb[kb,nb])  with packed FFMA2 while the tensor core works on the next
//   k-block (each warp: 32 rows x 64 columns, in two 32-column rounds).
//   INT8: int32 accumulation stays in TMEM for the whole K loop (exact); the buffers alternate between tiles, so the
//   epilogue of tile i overlaps the main loop of tile i+1.
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"

#ifndef GB_PROBE
#define GB_PROBE 0   // developer probe builds (tools/build_probe.sh): 2 no tcgen05.ld, 4 no FFMA2, 8 release before the loads
#endif

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
  // split-K (fp8, 128-wide tiles, few tiles): n_units = n_tiles * ksplit <= workers, unit u = (tile u / ksplit, K slice
  // u % ksplit); the slices of a tile meet in `part` (fp32, [unit][thread][32 float2]) behind two self-resetting counters
  // per tile in `tickets` (arrived, departed)
  int ksplit, n_units;
  int* tickets;
  float* part;
  long long* prof;  // developer (-DGB_PROF + knob gb_prof): per-CTA clock64 sums of the MMA thread's pipeline waits
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

// Warp roles.  The hardware warp scheduler prefers the HIGHEST warp id of a sub-partition among its eligible warps
// (B300_MICROARCH.md, "arbiter priority"), and the tensor core is fed by ONE thread: with the issuer in warp 1 the
// busy epilogue warps of its sub-partition (fp8 promotion: tcgen05.ld + FFMA2 every k-block) starved it -- 700 issue
// cycles per k-block whatever the tile width (profiles/r02_gemm_notes.md).  So: epilogue warps first (their TMEM lane
// quarter is warp id % 4), then one warpgroup of service warps with the MMA issuer LAST:
//   warps 0 .. EW-1        epilogue (EW = BN / 16; warp w: TMEM lanes [32 (w%4), +32), columns [64 (w/4), +64))
//   warp  EW               TMEM allocator       warp EW+1  idle
//   warp  EW+2             TMA producer         warp EW+3  MMA issuer (one thread)
// Registers: the service warpgroup gives its registers back (setmaxnreg.dec), the epilogue warpgroups take them
// (setmaxnreg.inc) -- 64 fp32 accumulators + a 32-column staging block per thread do not fit the 102-register share
// of a 640-thread CTA otherwise (the first version spilled 344 bytes per thread into the k-loop).
template <int N>
__device__ __forceinline__ void gb_setmaxnreg_inc() { asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(N)); }
template <int N>
__device__ __forceinline__ void gb_setmaxnreg_dec() { asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(N)); }

template <int MODE, int PAIR, int BN>
__global__ void __launch_bounds__((4 + BN / 16) * 32, 1)
gemm_big_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
                const __grid_constant__ BigParams p, const int stages) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int A_BYTES = GB_BM * GB_BK;                   // 16 KB
  constexpr int GB_BN = BN, EW = BN / 16, NBUF = 512 / BN;
  constexpr int B_ROWS = GB_BN / PAIR;                     // W rows staged by this CTA
  constexpr int B_BYTES = B_ROWS * GB_BK;                  // 32 KB / 16 KB
  constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  constexpr uint32_t kTmemCols = 512;
  // register budget after re-allocation.  setmaxnreg.inc only hands out what setmaxnreg.dec returned inside the SAME
  // CTA (asking for more than that waits for ever -- measured): 640 threads x 96: 128 x (96 - 32) = 512 x (112 - 96);
  // 384 threads x 168: 128 x (168 - 40) >= 256 x (200 - 168)
  constexpr int EPI_REGS = EW == 16 ? 112 : 200, SVC_REGS = EW == 16 ? 32 : 40;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + stages * STAGE_BYTES);
  uint64_t* empty = full + GB_MAX_STAGES;
  uint64_t* tfull = empty + GB_MAX_STAGES;
  uint64_t* tempty = tfull + 4;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 4);

  // (the shuffle makes the warp index provably warp-uniform for the compiler: the service roles below then keep their
  // loop state, barrier addresses and UMMA / TMA descriptors in UNIFORM registers.  A role entered through a per-thread
  // `lane == 0` branch makes ptxas wrap every UTCQMMA / UTCBAR / UTMALDG in an ELECT + R2UR.BROADCAST loop -- about
  // twenty extra instructions per tensor-core instruction on the one thread that feeds the tensor core.)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int rank = PAIR == 2 ? (int)gb_cluster_rank() : 0;
  const int worker = blockIdx.x / PAIR;
  const bool leader = rank == 0;

  if (warp == EW + 2 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == EW + 3 && lane == 0) {
    for (int s = 0; s < stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < NBUF; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], EW * PAIR);
    }
    fence_mbar_init();
  }
  if (warp == EW) gb_tmem_alloc<PAIR>(tmem_slot, kTmemCols);
  tc_fence_before();
  if constexpr (PAIR == 2) gb_cluster_sync(); else __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();  // programmatic dependent launch: the next kernel of the stream may start its prologue

  if (warp >= EW) {
    gb_setmaxnreg_dec<SVC_REGS>();
    if (warp == EW + 2) {
      // ================================ TMA producer (whole warp in the loop, one elected lane issues) ============
      const uint32_t full0_cluster = PAIR == 2 ? gb_mapa(smem_u32(&full[0]), 0) : 0;
      int s = 0;
      uint32_t ph = 0;
      // weights do not depend on the previous kernel of the stream: the first ring-full of W tiles is requested before
      // griddepcontrol.wait, the activations after it (decode-shape calls: the ramp of this GEMM overlaps its predecessor)
      int pre = 0;
      const int S = p.ksplit;
      if constexpr (PAIR == 1) {
        if (worker < p.n_units) {
          int tm, tn;
          const int t = worker / S, si = worker - t * S;
          gb_tile_coord(p, t, tm, tn);
          const int kb_lo = si * p.kb_total / S, kb_hi = (si + 1) * p.kb_total / S;
          pre = kb_hi - kb_lo < stages ? kb_hi - kb_lo : stages;
          if (elect_one()) {
            for (int i = 0; i < pre; ++i) {
              mbar_arrive_expect_tx(&full[i], STAGE_BYTES);
              tma_load_2d(smem + i * STAGE_BYTES + A_BYTES, &tma_b, (kb_lo + i) * GB_BK, tn * GB_BN, &full[i]);
            }
          }
          __syncwarp();
        }
      }
      griddep_wait();
      for (int u = worker; u < p.n_units; u += p.workers) {
        int tm, tn;
        const int t = u / S, si = u - t * S;
        gb_tile_coord(p, t, tm, tn);
        const int m0 = (tm * PAIR + rank) * GB_BM, n0 = tn * GB_BN + rank * B_ROWS;
        const int kb_lo = si * p.kb_total / S, kb_hi = (si + 1) * p.kb_total / S;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          if (pre > 0) {   // this stage's W tile is already in flight (first tile only): add the activations
            --pre;
            if (elect_one()) tma_load_2d(smem + s * STAGE_BYTES, &tma_a, kb * GB_BK, m0, &full[s]);
            __syncwarp();
            if (++s == stages) { s = 0; ph ^= 1; }
            continue;
          }
          mbar_wait(&empty[s], ph ^ 1);
          if (elect_one()) {
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
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
        }
      }
    } else if (warp == EW + 3 && leader) {
      // ================================ MMA issuer (pair leader only; one elected lane issues) ====================
      constexpr uint32_t idesc = MODE == 0 ? umma_idesc(1 /*f32*/, 0 /*e4m3*/, 0, GB_BM * PAIR, GB_BN)
                                           : umma_idesc(2 /*s32*/, 1 /*s8*/, 1, GB_BM * PAIR, GB_BN);
      int s = 0, a = 0;
      uint32_t ph = 0, aph = 0;   // smem ring stage / phase, TMEM buffer / phase
      const uint64_t desc0 = umma_desc_k_sw128(smem_u32(smem));
#ifdef GB_PROF
      long long pm_tempty = 0, pm_full = 0, pm_n = 0;
      const long long pm_t0 = clock64();
#endif
      const int S = p.ksplit;
      for (int u = worker; u < p.n_units; u += p.workers) {
        const int si = u % S;
        const int kb_lo = si * p.kb_total / S, kb_hi = (si + 1) * p.kb_total / S;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
#ifdef GB_PROF
          const long long c0 = clock64();
#endif
          if (MODE == 0 || kb == kb_lo) mbar_wait(&tempty[a], aph ^ 1);
#ifdef GB_PROF
          const long long c1 = clock64();
#endif
          mbar_wait(&full[s], ph);
          tc_fence_after();
#ifdef GB_PROF
          pm_tempty += c1 - c0;
          pm_full += clock64() - c1;
          ++pm_n;
#endif
          // (the 14-bit start-address field counts 16-byte units: stage s starts s * STAGE_BYTES / 16 further on)
          const uint64_t adesc = desc0 + (uint64_t)((uint32_t)(s * STAGE_BYTES) >> 4);
          const uint64_t bdesc = adesc + (A_BYTES >> 4);
          const uint32_t d = tmem_base + a * GB_BN;
          const bool last_kb = kb == kb_hi - 1;
          if (elect_one()) {
#pragma unroll
            for (int k = 0; k < GB_BK / 32; ++k) {
              const uint32_t accum = MODE == 0 ? (k > 0) : (kb > kb_lo || k > 0);
              gb_umma<PAIR, MODE>(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
            }
            gb_commit<PAIR>(&empty[s]);
            if (MODE == 0 || last_kb) gb_commit<PAIR>(&tfull[a]);
          }
          __syncwarp();
          if (++s == stages) { s = 0; ph ^= 1; }
          if (MODE == 0 || last_kb) {
            if (++a == NBUF) { a = 0; aph ^= 1; }
          }
        }
      }
#ifdef GB_PROF
      if (p.prof && lane == 0) {
        p.prof[blockIdx.x * 8 + 0] = clock64() - pm_t0;
        p.prof[blockIdx.x * 8 + 1] = pm_tempty;
        p.prof[blockIdx.x * 8 + 2] = pm_full;
        p.prof[blockIdx.x * 8 + 3] = pm_n;
      }
#endif
    }
  } else {
    // ================================ epilogue warps ================================
    gb_setmaxnreg_inc<EPI_REGS>();
    griddep_wait();   // the activation scales (and the output buffer) belong to the stream order
    const int quad = warp & 3, cg = warp >> 2;  // TMEM lanes [32*quad, +32), columns [64*cg, +64)
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
    int a = 0;
    uint32_t aph = 0;   // TMEM buffer / phase, advanced exactly as the MMA thread advances its own
    const int S = p.ksplit;
    for (int u = worker; u < p.n_units; u += p.workers) {
      int tm, tn;
      const int t = u / S, si = u - t * S;
      const int kb_lo = si * p.kb_total / S, kb_hi = (si + 1) * p.kb_total / S;
      gb_tile_coord(p, t, tm, tn);
      const int row = (tm * PAIR + rank) * GB_BM + row_in_tile;
      const int n0 = tn * GB_BN + cg * 64;
      float2 acc[32];
      if constexpr (MODE == 0) {
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[j] = make_float2(0.f, 0.f);
        const int nb = n0 / 128;
        const bool nb_ok = n0 < p.N;
        const float* sa_p = p.sa + (int64_t)(row < p.M ? row : 0) * p.sa_sm;
        const float* sb_p = p.sb + (int64_t)(nb_ok ? nb : 0) * p.sb_sn;
        // the scale loads come from an L2 that the TMA ring keeps busy: they are issued two k-blocks ahead
        float sa0 = __ldg(sa_p + (int64_t)kb_lo * p.sa_sk), sb0 = __ldg(sb_p + (int64_t)kb_lo * p.sb_sk);
        float sa1 = kb_lo + 1 < kb_hi ? __ldg(sa_p + (int64_t)(kb_lo + 1) * p.sa_sk) : 0.f;
        float sb1 = kb_lo + 1 < kb_hi ? __ldg(sb_p + (int64_t)(kb_lo + 1) * p.sb_sk) : 0.f;
        for (int kb = kb_lo; kb < kb_hi; ++kb) {
          const float sc = sa0 * sb0;
          sa0 = sa1;
          sb0 = sb1;
          if (kb + 2 < kb_hi) {
            sa1 = __ldg(sa_p + (int64_t)(kb + 2) * p.sa_sk);
            sb1 = __ldg(sb_p + (int64_t)(kb + 2) * p.sb_sk);
          }
          mbar_wait(&tfull[a], aph);
          tc_fence_after();
          uint32_t r[32];
          if constexpr (GB_PROBE != 0) {
            if (GB_PROBE & 8) release(a);
#pragma unroll
            for (int h = 0; h < 2; ++h) {
              if (GB_PROBE & 2) {
#pragma unroll
                for (int j = 0; j < 32; ++j) r[j] = (uint32_t)(kb + j);
              } else {
                tmem_ld_32x32(lane_addr + a * GB_BN + 32 * h, r);
                gb_ld_wait32(r);
              }
              if (h == 1 && !(GB_PROBE & 8)) release(a);
              if (GB_PROBE & 4) {
                acc[h].x += __uint_as_float(r[0]) + __uint_as_float(r[31]);
              } else {
#pragma unroll
                for (int j = 0; j < 16; ++j) gb_fma2(acc[16 * h + j], r[2 * j], r[2 * j + 1], sc);
              }
            }
          } else if constexpr (EW == 16) {
            // 112 registers: one 32-column block in flight; the buffer is released after the second load
            tmem_ld_32x32(lane_addr + a * GB_BN, r);
            gb_ld_wait32(r);
#pragma unroll
            for (int j = 0; j < 16; ++j) gb_fma2(acc[j], r[2 * j], r[2 * j + 1], sc);
            tmem_ld_32x32(lane_addr + a * GB_BN + 32, r);
            gb_ld_wait32(r);
            release(a);
#pragma unroll
            for (int j = 0; j < 16; ++j) gb_fma2(acc[16 + j], r[2 * j], r[2 * j + 1], sc);
          } else {
            uint32_t r1[32];
            tmem_ld_32x32(lane_addr + a * GB_BN, r);
            tmem_ld_32x32(lane_addr + a * GB_BN + 32, r1);
            gb_ld_wait32(r);
            gb_ld_wait32(r1);
            release(a);
#pragma unroll
            for (int j = 0; j < 16; ++j) gb_fma2(acc[j], r[2 * j], r[2 * j + 1], sc);
#pragma unroll
            for (int j = 0; j < 16; ++j) gb_fma2(acc[16 + j], r1[2 * j], r1[2 * j + 1], sc);
          }
          if (++a == NBUF) { a = 0; aph ^= 1; }
        }
      } else {
        mbar_wait(&tfull[a], aph);
        tc_fence_after();
        const float sa_row = row < p.M ? __ldg(p.sa + row) : 0.f;
        const uint16_t* bp = reinterpret_cast<const uint16_t*>(p.bias);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
          uint32_t r[32];
          tmem_ld_32x32(lane_addr + a * GB_BN + 32 * h, r);
          gb_ld_wait32(r);
          if (h == 1) release(a);
#pragma unroll
          for (int j = 0; j < 16; ++j) {
            const int n = n0 + 32 * h + 2 * j;
            const float s0 = n < p.N ? __ldg(p.sb + n) : 0.f, s1 = n + 1 < p.N ? __ldg(p.sb + n + 1) : 0.f;
            float v0 = __fmul_rn(__fmul_rn((float)(int)r[2 * j], sa_row), s0), v1 = __fmul_rn(__fmul_rn((float)(int)r[2 * j + 1], sa_row), s1);
            if (bp) {
              const uint16_t u0 = n < p.N ? __ldg(bp + n) : (uint16_t)0, u1 = n + 1 < p.N ? __ldg(bp + n + 1) : (uint16_t)0;
              v0 = __fadd_rn(v0, p.out_bf16 ? half_to_float<true>(u0) : half_to_float<false>(u0));
              v1 = __fadd_rn(v1, p.out_bf16 ? half_to_float<true>(u1) : half_to_float<false>(u1));
            }
            acc[16 * h + j] = make_float2(v0, v1);
          }
        }
        if (++a == NBUF) { a = 0; aph ^= 1; }
      }
      if constexpr (MODE == 0 && EW == 8 && PAIR == 1) {
        if (S > 1) {
          // ---- split-K: the S slices of tile t meet in global memory; every CTA of the tile then reduces 1/S of it ----
          // (all S CTAs are resident: n_units <= workers, one unit per CTA -- the wait below cannot deadlock)
          const int tid = threadIdx.x;   // 0 .. 255: the epilogue warps are warps 0 .. 7
          float4* part4 = reinterpret_cast<float4*>(p.part);
          if (row < p.M && n0 < p.N) {
            float4* mine = part4 + ((size_t)u * 256 + tid) * 16;
#pragma unroll
            for (int j4 = 0; j4 < 16; ++j4) mine[j4] = make_float4(acc[2 * j4].x, acc[2 * j4].y, acc[2 * j4 + 1].x, acc[2 * j4 + 1].y);
          }
          __threadfence();
          asm volatile("bar.sync 1, 256;" ::: "memory");
          if (tid == 0) {
            int* arrived = p.tickets + 2 * t;
            asm volatile("red.release.gpu.global.add.s32 [%0], 1;" ::"l"(arrived) : "memory");   // no return value: one L2 round trip less
            int seen;
            do {
              asm volatile("ld.acquire.gpu.global.s32 %0, [%1];" : "=r"(seen) : "l"(arrived) : "memory");
            } while (seen < S);
          }
          asm volatile("bar.sync 1, 256;" ::: "memory");
          __threadfence();
          const int lo = si * 256 / S, hi = (si + 1) * 256 / S;
          const int row_base = (tm * PAIR + rank) * GB_BM, col_base = tn * GB_BN;
          uint16_t* outp = reinterpret_cast<uint16_t*>(p.out);
          const bool vec_out = (p.ldo % 4 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 7) == 0);
          // all S partials of two output chunks are requested before the first add (a load -> add chain per slice costs an
          // L2 round trip each: measured 9 us for S = 6)
          const int nf = (hi - lo) * 16;
          for (int f0 = tid; f0 < nf; f0 += 512) {
            float4 v[2][8];
            bool ok[2];
            int orow[2], ocol[2];
#pragma unroll
            for (int q = 0; q < 2; ++q) {
              const int f = f0 + q * 256;
              const int owner = lo + (f >> 4), j4 = f & 15;
              const int ow = owner >> 5;   // owner's warp: TMEM lane quarter ow & 3, column group ow >> 2
              orow[q] = row_base + (ow & 3) * 32 + (owner & 31);
              ocol[q] = col_base + (ow >> 2) * 64 + 4 * j4;
              ok[q] = f < nf && orow[q] < p.M && col_base + (ow >> 2) * 64 < p.N;
#pragma unroll
              for (int s2 = 0; s2 < 8; ++s2)
                if (ok[q] && s2 < S) v[q][s2] = __ldcg(part4 + ((size_t)(t * S + s2) * 256 + owner) * 16 + j4);
            }
#pragma unroll
            for (int q = 0; q < 2; ++q) {
              if (!ok[q]) continue;
              float4 sum = v[q][0];
#pragma unroll
              for (int s2 = 1; s2 < 8; ++s2)   // fixed order: bit-reproducible whoever arrives last
                if (s2 < S) { sum.x += v[q][s2].x; sum.y += v[q][s2].y; sum.z += v[q][s2].z; sum.w += v[q][s2].w; }
              uint16_t* dst = outp + (size_t)orow[q] * p.ldo + ocol[q];
              if (vec_out && ocol[q] + 4 <= p.N) {
                uint2 pk;
                if (p.out_bf16) { pk.x = pack_half2<true>(sum.x, sum.y); pk.y = pack_half2<true>(sum.z, sum.w); }
                else { pk.x = pack_half2<false>(sum.x, sum.y); pk.y = pack_half2<false>(sum.z, sum.w); }
                *reinterpret_cast<uint2*>(dst) = pk;
              } else {
                const float o4[4] = {sum.x, sum.y, sum.z, sum.w};
#pragma unroll
                for (int e = 0; e < 4; ++e)
                  if (ocol[q] + e < p.N) dst[e] = p.out_bf16 ? float_to_half<true>(o4[e]) : float_to_half<false>(o4[e]);
              }
            }
          }
          asm volatile("bar.sync 1, 256;" ::: "memory");
          if (tid == 0) {   // the last CTA to leave re-arms the tile's counters for the next call
            if (atomicAdd(p.tickets + 2 * t + 1, 1) == S - 1) {
              p.tickets[2 * t] = 0;
              p.tickets[2 * t + 1] = 0;
            }
          }
          continue;
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
    }
  }

  tc_fence_before();
  __syncwarp();
  if constexpr (PAIR == 2) gb_cluster_sync(); else __syncthreads();
  if (warp == EW) {
    tc_fence_after();
    gb_tmem_dealloc<PAIR>(tmem_base, kTmemCols);
  }
}

constexpr size_t GB_TICKET_BYTES = 4096;            // two counters per tile, <= 512 tiles
constexpr size_t GB_PART_BYTES = 256 * 32 * 8;      // one unit's fp32 partial tile: 256 threads x 32 float2

template <int MODE, int PAIR, int BN>
static int launch_big_t(const void* a, int64_t lda, const void* w, int64_t ldw, BigParams& bp, void* workspace,
                        size_t workspace_bytes, cudaStream_t st) {
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
  // split-K where the tiles fill less than half of the SMs (fp8, 128-wide tiles, large-M kernel chosen: M > 128 or a short
  // K loop): a CTA streams at most ~100 GB/s, so 24 tiles of a 1536 x 4096 weight keep 24 SMs busy for 13 us whatever the
  // HBM could do; S slices per tile put S times the SMs on the same bytes.  A function of the shape only (the TP
  // emulation of bench.py and the fused / unfused parity tests rely on that); knob gb_split: 0 auto, 1 off, n forces n.
  int S = 1;
  if (MODE == 0 && PAIR == 1 && BN == 128 && workspace && ((uintptr_t)workspace & 15) == 0 && bp.n_tiles * 2 <= sms &&
      (size_t)bp.n_tiles * 8 <= GB_TICKET_BYTES) {
    const int knob = tune(TUNE_GB_SPLIT);
    S = sms / bp.n_tiles;
    if (S > 8) S = 8;
    if (S > bp.kb_total / 4) S = bp.kb_total / 4;   // >= 4 k-blocks per slice
    // the slices meet through L2 (partial tile, fence, counter, wait, reload: ~5 us of round trips, tools/gb_split_sweep.py),
    // a CTA spends ~0.3 us per k-block: worth it only if the k-loop shrinks by more than that -- measured: K = 8192 with 20
    // tiles 22.0 -> 13.0 us, K = 7168 with 34 tiles 20.3 -> 14.1, K = 4096 with 24 tiles 13.0 -> 11.7; K <= 3584 loses
    if (S > 1 && bp.kb_total * (S - 1) < 24 * S) S = 1;
    if (knob > 0) S = knob;
    if (S > bp.kb_total) S = bp.kb_total;
    while (S > 1 && (bp.n_tiles * S > sms || GB_TICKET_BYTES + (size_t)bp.n_tiles * S * GB_PART_BYTES > workspace_bytes)) --S;
    if (S < 1) S = 1;
  }
  bp.ksplit = S;
  bp.n_units = bp.n_tiles * S;
  bp.tickets = S > 1 ? reinterpret_cast<int*>(workspace) : nullptr;
  bp.part = S > 1 ? reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + GB_TICKET_BYTES) : nullptr;
  int workers = sms / PAIR;
  if (workers > bp.n_units) workers = bp.n_units;
  bp.workers = workers;   // (S > 1: n_units <= sms, every unit has its own resident CTA)
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
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = PAIR;
  attr[1].val.clusterDim.y = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = PAIR > 1 ? 2 : 1;
#ifdef GB_PROF
  const int prof = tune(TUNE_GB_PROF);
  static long long* dprof = nullptr;
  if (prof && !dprof) cudaMalloc(&dprof, 160 * 64);
  bp.prof = prof ? dprof : nullptr;
#endif
  B200_CUDA(cudaLaunchKernelEx(&cfg, kern, ta, tb, bp, stages));
#ifdef GB_PROF
  if (prof) {
    static int n = 0;
    cudaStreamSynchronize(st);
    long long h[160 * 8];
    cudaMemcpy(h, dprof, sizeof(h), cudaMemcpyDeviceToHost);
    if (n++ % 8 == 3) {
      double s[8] = {0};
      const int g = workers * PAIR;
      for (int c = 0; c < g; ++c) for (int i = 0; i < 8; ++i) s[i] += (double)h[c * 8 + i] / g;
      fprintf(stderr, "[gemm_big prof] M=%d N=%d K=%d BN=%d: MMA thread total %.0f cyc, wait tempty %.0f, wait full %.0f, k-blocks %.0f\n",
              bp.M, bp.N, bp.K, BN, s[0], s[1], s[2], s[3]);
    }
  }
#endif
  return launched("gemm_big_kernel launch");
}

// a [M,K] activations, w [N,K] weight (both K-major, 1-byte elements); mode 0 fp8 block-scaled, 1 int8
size_t big_workspace_bytes() {
  const int sms = device_sm_count() > 0 ? device_sm_count() : 148;
  return GB_TICKET_BYTES + (size_t)sms * GB_PART_BYTES;
}

int launch_big(int mode, const void* a, int64_t lda, const void* w, int64_t ldw, const float* sa, int64_t sa_sm, int64_t sa_sk,
               const float* sb, int64_t sb_sk, int64_t sb_sn, const void* bias, void* out, int64_t ldo, int out_bf16, int M,
               int N, int K, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  BigParams bp{};
  bp.sa = sa; bp.sa_sm = sa_sm; bp.sa_sk = sa_sk;
  bp.sb = sb; bp.sb_sk = sb_sk; bp.sb_sn = sb_sn;
  bp.bias = bias; bp.out = out; bp.ldo = ldo; bp.out_bf16 = out_bf16;
  bp.M = M; bp.N = N; bp.K = K;
  const int pair = tune(TUNE_GB_PAIR);
  int bn = tune(TUNE_GB_BN);   // 0 = by shape
  if (mode == 0) {
    // 256-wide tiles are the faster pipeline (2112 vs 1880 TFLOP/s at 8192^3) but a skinny problem has too few of them
    // to put every SM's copy engine to work: below ~3/4 of a wave of 256-wide tiles take 128-wide ones
    // (one row of tiles, M <= 128: HBM-bound, and a wave of 112 .. 147 wide tiles leaves SMs without a stream --
    // 28672 x 4096 at M = 64: 25.3 us with 224 tiles of 128 against 26.4 with 112 of 256)
    if (bn == 0) {
      const int wide = ((M + GB_BM - 1) / GB_BM) * ((N + 255) / 256);
      bn = wide >= 112 && !(M <= 64 && wide < 148) ? 256 : 128;
    }
    // knob gb_pair = 4 selects cta_group::2 pairs for fp8 too (measured slower: the per-k-block release crosses the pair)
    if (pair == 4 && M > GB_BM) {
      if (bn == 128) return launch_big_t<0, 2, 128>(a, lda, w, ldw, bp, workspace, workspace_bytes, st);
      return launch_big_t<0, 2, 256>(a, lda, w, ldw, bp, workspace, workspace_bytes, st);
    }
    if (bn == 128) return launch_big_t<0, 1, 128>(a, lda, w, ldw, bp, workspace, workspace_bytes, st);
    return launch_big_t<0, 1, 256>(a, lda, w, ldw, bp, workspace, workspace_bytes, st);
  }
  if (pair != 1 && M > GB_BM) return launch_big_t<1, 2, 256>(a, lda, w, ldw, bp, workspace, workspace_bytes, st);
  return launch_big_t<1, 1, 256>(a, lda, w, ldw, bp, workspace, workspace_bytes, st);
}

}  // namespace b200
