// AWQ INT4 W4A16 decode-shape GEMM (M <= 256, N % 128 == 0, K % 64 == 0), fused: int4 box -> registers -> tcgen05.
//
// Replaces AWQLinearMethod.apply (python/sglang/srt/layers/quantization/awq.py:186-204), i.e. awq_dequantize
// (sgl-kernel/csrc/gemm/awq_kernel.cu:126-221) followed by a half GEMM, with ONE kernel whose only weight traffic is
// the packed tensor (0.5 B per weight).  Same skeleton as gemm_skinny_kernel (swap-AB tcgen05, cluster split-K with a
// fixed-order reduce-scatter through distributed shared memory) with two differences that the first fused version
// (gemm_skinny.cu MODE 2, still the fallback for N % 128 != 0) showed to matter at M = 128:
//   * a CTA owns NT (1..3) adjacent 128-column sub-tiles, so one activation tile [M x 64 k] (16 KB at M 128, four
//     times the 4 KB of int4 weights it meets) is staged once per NT sub-tiles instead of once per sub-tile;
//   * the dequantised operand tiles live in their own two-deep ring, so the TMA ring holds only raw boxes and
//     activations and is 2-3x deeper for the same shared memory (the stage lifetime -- TMA, wait for a dequantiser
//     group, dequantise, MMA -- was what bounded the first version: 6 stages / ~2.5 us).
// Dequantisation (bit-identical to awq_dequantize): a thread takes 8 k-rows x one packed word (8 columns), forms
// ((nib | magic) - (zero | magic)) * scale in half2 / bfloat162 (magic = 1024 / 128: exact integers, one rounding in
// the multiply, as awq_kernel.cu:151-175), transposes n-pairs into k-pairs with PRMT and stores eight 16-byte K-major
// chunks into the SWIZZLE_128B operand tile.  UMMA row i holds sub-tile column n = 8 (i % 16) + i / 16: a quarter
// warp then writes eight different bank groups; the write-out un-permutes.
#include <stdio.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

constexpr int AQ_BN = 128;   // columns per sub-tile (UMMA M)
constexpr int AQ_BK = 128;   // operand bytes of k per stage (64 half elements)
constexpr int AQ_RAW = 4096; // raw box of one sub-tile: 64 k-rows x 16 packed words
constexpr int AQ_MAX_STAGES = 12;
constexpr int AQ_MAX_SPLITS = 8;
constexpr int AQ_NONEPI = 4;  // warp 0/3 TMA producers, warp 1 MMA issuer, warp 2 TMEM allocator
constexpr int AQ_NA = 2;      // dequantised operand buffers

struct AwqParams {
  const int32_t* qzeros;   // [K/group, N/8]
  const uint16_t* scales;  // [K/group, N]
  const void* bias;        // [N] or null
  void* out;
  int64_t ldo;
  int bf16;
  int M, N, K, mt_pad, group;
  int n_tiles;   // CTAs along N (each NT sub-tiles)
  int kb_total;  // K / 64
  int splits, stages;
};

__device__ __forceinline__ uint32_t aq_cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void aq_cluster_sync() {
  // .relaxed: these barriers order execution only (see gemm_skinny.cu cluster_arrive: .release is a MEMBAR.ALL.GPU)
  asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t aq_mapa(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void aq_bulk_s2c(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster),
               "r"(src_cta), "r"(bytes), "r"(bar_cluster)
               : "memory");
}
__device__ __forceinline__ void tma_load_3d_hint(void* dst_smem, const CUtensorMap* m, int c0, int c1, int c2, uint64_t* bar,
                                                 uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, "
      "%4, %5}], [%2], %6;" ::"r"(smem_u32(dst_smem)),
      "l"(m), "r"(smem_u32(bar)), "r"(c0), "r"(c1), "r"(c2), "l"(policy)
      : "memory");
}
template <int GROUPS>
__device__ __forceinline__ void aq_epi_bar() {
  asm volatile("bar.sync 1, %0;" ::"n"(GROUPS * 128) : "memory");
}

// one raw sub-tile box [64 k][16 words] -> operand tile [128 rows x 128 B], this thread's (c, ko) item
template <bool BF>
__device__ __forceinline__ void aq_dequant_item(const uint32_t* raw, uint8_t* a_tile, int c, int ko, int h, const uint32_t (&zp)[4],
                                                const uint32_t (&s2)[4]) {
  constexpr uint32_t magic = BF ? 0x43004300u : 0x64006400u;
  uint32_t tmp[8], w[8];
  // the two half-warps read rows 512 B apart: opposite row parity keeps them on different banks
#pragma unroll
  for (int r = 0; r < 8; ++r) tmp[r] = raw[(ko * 8 + (r ^ h)) * 16 + c];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    w[2 * q] = h ? tmp[2 * q + 1] : tmp[2 * q];
    w[2 * q + 1] = h ? tmp[2 * q] : tmp[2 * q + 1];
  }
  uint8_t* arow = a_tile + c * 128 + ((ko ^ (c & 7)) << 4);
#pragma unroll
  for (int q = 0; q < 4; ++q) {  // columns 2q (low halves) and 2q + 1 (high halves) of the packed word's 8
    uint32_t v[8];
#pragma unroll
    for (int r = 0; r < 8; ++r) {
      const uint32_t t = ((w[r] >> (4 * q)) & 0x000F000Fu) | magic;
      if constexpr (BF) {
        const __nv_bfloat162 d = __hsub2(*reinterpret_cast<const __nv_bfloat162*>(&t), *reinterpret_cast<const __nv_bfloat162*>(&zp[q]));
        const __nv_bfloat162 m = __hmul2(d, *reinterpret_cast<const __nv_bfloat162*>(&s2[q]));
        v[r] = *reinterpret_cast<const uint32_t*>(&m);
      } else {
        const __half2 d = __hsub2(*reinterpret_cast<const __half2*>(&t), *reinterpret_cast<const __half2*>(&zp[q]));
        const __half2 m = __hmul2(d, *reinterpret_cast<const __half2*>(&s2[q]));
        v[r] = *reinterpret_cast<const uint32_t*>(&m);
      }
    }
    const uint4 lo = make_uint4(__byte_perm(v[0], v[1], 0x5410), __byte_perm(v[2], v[3], 0x5410),
                                __byte_perm(v[4], v[5], 0x5410), __byte_perm(v[6], v[7], 0x5410));
    const uint4 hi = make_uint4(__byte_perm(v[0], v[1], 0x7632), __byte_perm(v[2], v[3], 0x7632),
                                __byte_perm(v[4], v[5], 0x7632), __byte_perm(v[6], v[7], 0x7632));
    *reinterpret_cast<uint4*>(arow + (16 * (2 * q)) * 128) = lo;
    *reinterpret_cast<uint4*>(arow + (16 * (2 * q + 1)) * 128) = hi;
  }
}

template <int COLS, int GROUPS, int NT, bool BF>
__global__ void __launch_bounds__((AQ_NONEPI + 4 * GROUPS) * 32, 1)
awq_fused_kernel(const __grid_constant__ CUtensorMap tma_w, const __grid_constant__ CUtensorMap tma_x,
                 const __grid_constant__ AwqParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int BUFCOLS = GROUPS * COLS;  // TMEM columns of one sub-tile accumulator (>= mt_pad)
  constexpr uint32_t kTmemCols = NT * BUFCOLS <= 32 ? 32 : NT * BUFCOLS <= 64 ? 64 : NT * BUFCOLS <= 128 ? 128 : NT * BUFCOLS <= 256 ? 256 : 512;
  static_assert(NT * BUFCOLS <= 512, "tensor memory");
  constexpr int NE = GROUPS * 128;
  constexpr int TEAM = NT * 128 < NE ? NT * 128 : NE;    // dequantiser threads per stage
  constexpr int NTEAMS = NE / TEAM;                      // teams take alternate stages
  constexpr int ITEMS = (NT * 128 + TEAM - 1) / TEAM;    // items per thread per stage
  constexpr int RB = NTEAMS > AQ_NA ? NTEAMS : AQ_NA;    // "operand buffer free" barriers (see gfree)
  static_assert(RB <= 4 && RB % NTEAMS == 0 && RB % AQ_NA == 0, "free-barrier rotation");
  constexpr int A_RING = AQ_NA * NT * AQ_BN * AQ_BK;
  const int x_bytes = p.mt_pad * AQ_BK;
  const int stage_bytes = x_bytes + NT * AQ_RAW;  // [activation tile][NT raw boxes]
  uint8_t* ring = smem + A_RING;
  uint64_t* full = reinterpret_cast<uint64_t*>(ring + p.stages * stage_bytes);
  uint64_t* empty = full + AQ_MAX_STAGES;
  uint64_t* afull = empty + AQ_MAX_STAGES;
  // "operand buffer free": the MMA thread signals barrier (j + NA) % RB after stage j, stage i waits on barrier i % RB
  // (RB = max(teams, buffers)).  A per-buffer barrier alone could run two phases ahead of a waiting team when there are
  // more teams than buffers, which a parity wait cannot tell from zero phases (seen as a hang at M = 128).
  uint64_t* gfree = afull + AQ_NA;
  uint64_t* tfull = gfree + 4;
  uint64_t* xbar = tfull + 1;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(xbar + 1);
  float* send = reinterpret_cast<float*>(smem);  // after the k-loop everything below the barriers is scratch

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  const int S = p.splits;
  const int rank = S > 1 ? (int)aq_cluster_ctarank() : 0;
  const int kb0 = (int)((long long)rank * p.kb_total / S), kb1 = (int)((long long)(rank + 1) * p.kb_total / S);
  const int nkb = kb1 - kb0;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_w);
    tma_prefetch_desc(&tma_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < AQ_NA; ++a) mbar_init(&afull[a], TEAM / 32);  // one arrival per dequantiser warp
    for (int g = 0; g < RB; ++g) mbar_init(&gfree[g], 1);
    mbar_init(tfull, 1);
    mbar_init(xbar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();

  if ((warp == 0 || warp == 3) && lane == 0) {
    // ================================ TMA producers (stage i belongs to producer i % 2) ================================
    const int pidx = warp == 0 ? 0 : 1;
    const uint64_t pol_w = make_policy_evict_first();
    const uint32_t tx_bytes = (uint32_t)stage_bytes;
    const int pre = nkb < p.stages ? nkb : p.stages;
    // the packed weights do not depend on the previous kernel of the stream: start them before waiting for it
    for (int i = pidx; i < pre; i += 2) {
      mbar_arrive_expect_tx(&full[i], tx_bytes);
      tma_load_3d_hint(ring + i * stage_bytes + x_bytes, &tma_w, 0, (kb0 + i) * 64, tile * NT, &full[i], pol_w);
    }
    griddep_wait();
    for (int i = pidx; i < pre; i += 2) tma_load_2d(ring + i * stage_bytes, &tma_x, (kb0 + i) * AQ_BK, 0, &full[i]);
    int i = pidx;
    while (i < pre) i += 2;
    for (; i < nkb; i += 2) {
      const int s = i % p.stages;
      const uint32_t ph = (i / p.stages) & 1;
      mbar_wait(&empty[s], ph ^ 1);
      mbar_arrive_expect_tx(&full[s], tx_bytes);
      uint8_t* st = ring + s * stage_bytes;
      tma_load_3d_hint(st + x_bytes, &tma_w, 0, (kb0 + i) * 64, tile * NT, &full[s], pol_w);
      tma_load_2d(st, &tma_x, (kb0 + i) * AQ_BK, 0, &full[s]);
    }
  } else if (warp == 1 && lane == 0) {
    // ================================ MMA issuer ================================
    const uint32_t hfmt = BF ? 1u : 0u;
    const uint32_t idesc = umma_idesc(1 /*f32*/, hfmt, hfmt, AQ_BN, p.mt_pad);
    for (int i = 0; i < nkb; ++i) {
      const int s = i % p.stages, a = i % AQ_NA;
      mbar_wait(&full[s], (i / p.stages) & 1);   // the activation tile (async proxy) is visible to this thread
      mbar_wait(&afull[a], (i / AQ_NA) & 1);
      tc_fence_after();
      const uint64_t bdesc = umma_desc_k_sw128(smem_u32(ring + s * stage_bytes));
#pragma unroll
      for (int u = 0; u < NT; ++u) {
        const uint64_t adesc = umma_desc_k_sw128(smem_u32(smem + (a * NT + u) * (AQ_BN * AQ_BK)));
#pragma unroll
        for (int k = 0; k < AQ_BK / 32; ++k) umma_f16(tmem_base + u * BUFCOLS, adesc + 2 * k, bdesc + 2 * k, idesc, (i > 0 || k > 0));
      }
      tc_commit(&empty[s]);
      tc_commit(&gfree[(i + AQ_NA) % RB]);  // stage i + NA reuses this operand buffer
      if (i == nkb - 1) tc_commit(tfull);
    }
  }

  const int ew = warp - AQ_NONEPI, grp = ew >> 2, quad = ew & 3;
  const int row = quad * 32 + lane;   // thread index inside its 128-thread group == TMEM lane
  const int col0 = grp * COLS;        // first token handled by this group in the write-out
  const int et_all = threadIdx.x - AQ_NONEPI * 32;
  int ncols = p.mt_pad - col0;
  ncols = ncols < 0 ? 0 : (ncols > COLS ? COLS : ncols);
  const int nchunk = (ncols + 31) / 32;

  if (warp >= AQ_NONEPI) {
    // ================================ dequantiser teams ================================
    // A stage holds NT x 128 items (one per (sub-tile, packed word column, k-octet)).  TEAM = min(NE, NT x 128) threads
    // work on ONE stage together, so a stage is dequantised in about one item's latency and the two operand buffers
    // are enough to keep the tensor core fed; the NE / TEAM teams take alternate stages.
    griddep_wait();
    const int team = et_all / TEAM, tt = et_all - team * TEAM;
    constexpr uint32_t magic = BF ? 0x43004300u : 0x64006400u;
    const int c = tt & 15, ko = (tt >> 4) & 7, h = ko & 1;  // (TEAM and 128 are multiples of 128: same for all items)
    int cur_g = -1;
    uint32_t zp[ITEMS][4], s2[ITEMS][4];
    if (team < NTEAMS) {
      for (int i = team; i < nkb; i += NTEAMS) {
        const int s = i % p.stages, a = i % AQ_NA;
        const int gi = ((kb0 + i) * 64 + ko * 8) / p.group;  // this thread's 8 k-rows share a group (group % 8 == 0)
        if (gi != cur_g) {
          cur_g = gi;
#pragma unroll
          for (int it = 0; it < ITEMS; ++it) {
            const int u = (tt + it * TEAM) >> 7;
            const int ncol = (tile * NT + u) * 16 + c;
            uint4 sraw = make_uint4(0u, 0u, 0u, 0u);
            uint32_t z = 0;
            if (u < NT && ncol * 8 < p.N) {
              sraw = __ldg(reinterpret_cast<const uint4*>(p.scales + ((int64_t)gi * p.N + (int64_t)ncol * 8)));
              z = (uint32_t)__ldg(p.qzeros + (int64_t)gi * (p.N / 8) + ncol);
            }
            s2[it][0] = sraw.x; s2[it][1] = sraw.y; s2[it][2] = sraw.z; s2[it][3] = sraw.w;
#pragma unroll
            for (int q = 0; q < 4; ++q) zp[it][q] = ((z >> (4 * q)) & 0x000F000Fu) | magic;
          }
        }
        // the MMAs of stage i - NA, which read this operand buffer, have completed: barrier i % R, its n-th completion
        if (i >= AQ_NA) mbar_wait(&gfree[i % RB], (uint32_t)((i / RB - ((i % RB) < AQ_NA ? 1 : 0)) & 1));
        mbar_wait(&full[s], (i / p.stages) & 1);
        const uint8_t* st = ring + s * stage_bytes + x_bytes;
#pragma unroll
        for (int it = 0; it < ITEMS; ++it) {
          const int u = (tt + it * TEAM) >> 7;
          if (u < NT)
            aq_dequant_item<BF>(reinterpret_cast<const uint32_t*>(st + u * AQ_RAW), smem + (a * NT + u) * (AQ_BN * AQ_BK), c, ko, h, zp[it], s2[it]);
        }
        fence_proxy_async();  // generic-proxy stores -> read by the tensor core through the async proxy
        __syncwarp();
        if (lane == 0) mbar_arrive(&afull[a]);
      }
    }
    mbar_wait(tfull, 0);
    tc_fence_after();
  }

  // ================================ write-out, one sub-tile at a time ================================
  // [split-K exchange through DSMEM,] sum in CTA order (bit-reproducible), bias, convert, coalesced store
  const int P = p.mt_pad < 128 ? p.mt_pad : 128;
  const int passes = (p.mt_pad + P - 1) / P;
  const int slice = (P + S - 1) / S;
  float* recv = send + P * AQ_BN;  // [S-1][slice][128]: slot (src - me - 1) mod S holds the partial of CTA src
  // sub-tile column n held by this TMEM lane, stored at physical column n ^ (((n >> 5) & 3) << 1): without the XOR a
  // warp's 32 columns {8a + b} fall on four banks (eight-way conflict on every one of the COLS stores per thread)
  const int scol = 8 * (row & 15) + (row >> 4);
  const int srow = scol ^ (((scol >> 5) & 3) << 1);
  uint32_t xph = 0;
  for (int u = 0; u < NT; ++u) {
    const int tile_n = tile * NT + u;
    if (tile_n * AQ_BN >= p.N) break;  // (uniform over the cluster)
    float acc[COLS];
    if (warp >= AQ_NONEPI) {
      const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + u * BUFCOLS + col0;
#pragma unroll
      for (int cc = 0; cc < COLS / 32; ++cc) {
        if (cc < nchunk) {
          uint32_t r[32];
          tmem_ld_32x32(lane_addr + cc * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[cc * 32 + j] = __uint_as_float(r[j]);
        }
      }
    }
    for (int ps = 0; ps < passes; ++ps) {
      const int t0 = ps * P;
      const int pt = p.mt_pad - t0 < P ? p.mt_pad - t0 : P;
      const int my0 = rank * slice;
      int mine = pt - my0;
      mine = mine < 0 ? 0 : (mine > slice ? slice : mine);
      if (warp >= AQ_NONEPI && col0 >= t0 && col0 < t0 + pt) {
#pragma unroll
        for (int j = 0; j < COLS; ++j)
          if (j < ncols) send[(col0 - t0 + j) * AQ_BN + srow] = acc[j];
        if (S > 1) fence_proxy_async();
      }
      if (S > 1) {
        if (warp >= AQ_NONEPI) aq_epi_bar<GROUPS>();   // send[] complete at CTA scope (the cluster barrier is relaxed)
        if (threadIdx.x == AQ_NONEPI * 32) mbar_arrive_expect_tx(xbar, (uint32_t)((S - 1) * mine * AQ_BN * 4));
        __syncwarp();
        aq_cluster_sync();  // A: partials are in place and every barrier is armed
        if (warp == AQ_NONEPI) {   // one lane per destination CTA
          const int o = lane;
          if (o < S && o != rank) {
            const int o0 = o * slice;
            int on = pt - o0;
            on = on < 0 ? 0 : (on > slice ? slice : on);
            if (on > 0) {
              int slot = rank - o - 1;
              if (slot < 0) slot += S;
              aq_bulk_s2c(aq_mapa(smem_u32(recv + slot * slice * AQ_BN), (uint32_t)o), smem_u32(send + o0 * AQ_BN),
                          (uint32_t)(on * AQ_BN * 4), aq_mapa(smem_u32(xbar), (uint32_t)o));
            }
          }
          __syncwarp();
        }
        if (warp >= AQ_NONEPI) mbar_wait(xbar, xph & 1);
        ++xph;
      } else if (warp >= AQ_NONEPI) {
        aq_epi_bar<GROUPS>();
      }
      if (warp >= AQ_NONEPI) {
        const int pi = et_all & 63, tq = et_all >> 6;
        const int n_in = 2 * pi, n = tile_n * AQ_BN + n_in;
        const bool pair_ok = (n + 1 < p.N) && (p.ldo % 2 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 3) == 0);
        float b0 = 0.f, b1 = 0.f;
        if (p.bias) {
          const uint16_t* bp = reinterpret_cast<const uint16_t*>(p.bias);
          const uint16_t u0 = n < p.N ? __ldg(bp + n) : (uint16_t)0, u1 = n + 1 < p.N ? __ldg(bp + n + 1) : (uint16_t)0;
          b0 = half_to_float<BF>(u0);
          b1 = half_to_float<BF>(u1);
        }
        int poff[AQ_MAX_SPLITS];   // partial of CTA s for my first token (floats from `send`): no division in the token loop
#pragma unroll
        for (int s2i = 0; s2i < AQ_MAX_SPLITS; ++s2i) {
          int slot = s2i - rank - 1;
          if (slot < 0) slot += S;
          poff[s2i] = (s2i == rank ? my0 * AQ_BN : (int)(recv - send) + slot * slice * AQ_BN) + (n_in ^ (((n_in >> 5) & 3) << 1));
        }
#pragma unroll 2
        for (int tk = tq; tk < mine; tk += NE / 64) {
          const int m = t0 + my0 + tk;
          if (m >= p.M || n >= p.N) continue;
          float o0 = 0.f, o1 = 0.f;
#pragma unroll
          for (int s2i = 0; s2i < AQ_MAX_SPLITS; ++s2i) {
            if (s2i < S) {
              const float2 v = *reinterpret_cast<const float2*>(send + poff[s2i] + tk * AQ_BN);
              o0 = s2i == 0 ? v.x : o0 + v.x;
              o1 = s2i == 0 ? v.y : o1 + v.y;
            }
          }
          if (p.bias) {
            o0 += b0;
            o1 += b1;
          }
          uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + (size_t)m * p.ldo + n;
          if (pair_ok) {
            *reinterpret_cast<uint32_t*>(dst) = pack_half2<BF>(o0, o1);
          } else {
            dst[0] = float_to_half<BF>(o0);
            if (n + 1 < p.N) dst[1] = float_to_half<BF>(o1);
          }
        }
      }
      if (S > 1) {
        __syncwarp();
        aq_cluster_sync();  // B: my outgoing copies have been consumed, nobody reads my shared memory any more
      } else if (warp >= AQ_NONEPI) {
        aq_epi_bar<GROUPS>();
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
struct AwqPlan {
  int nt, splits, stages, smem;
};

static int aq_stage_bytes(int mt_pad, int nt) { return mt_pad * AQ_BK + nt * AQ_RAW; }

static int aq_stages(int mt_pad, int nt) {
  const int budget = 226 * 1024 - 1024 /*align*/ - 1024 /*barriers*/ - AQ_NA * nt * AQ_BN * AQ_BK;
  int st = budget / aq_stage_bytes(mt_pad, nt);
  return st > AQ_MAX_STAGES ? AQ_MAX_STAGES : st;
}

// CTAs that can be resident at once when launched as clusters of S (one CTA per SM; a cluster lives inside one GPC, so
// S = 5 or 7 strands SMs: asking for 28 x 5 CTAs ran as two waves before this was queried)
static int aq_cluster_capacity(int S) {
  static int cap[AQ_MAX_SPLITS + 1] = {0};
  static bool done = false;
  if (!done) {
    auto kern = awq_fused_kernel<32, 4, 1, true>;
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    const int sms = device_sm_count() > 0 ? device_sm_count() : 148;
    for (int s = 1; s <= AQ_MAX_SPLITS; ++s) {
      cudaLaunchConfig_t cfg = {};
      cfg.blockDim = dim3((AQ_NONEPI + 16) * 32);
      cfg.gridDim = dim3(1, s);
      cfg.dynamicSmemBytes = 215 * 1024;
      cudaLaunchAttribute attr[1];
      attr[0].id = cudaLaunchAttributeClusterDimension;
      attr[0].val.clusterDim.x = 1;
      attr[0].val.clusterDim.y = s;
      attr[0].val.clusterDim.z = 1;
      cfg.attrs = attr;
      cfg.numAttrs = 1;
      int ncl = 0;
      if (cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg) != cudaSuccess || ncl <= 0) {
        cudaGetLastError();
        ncl = (sms / 8 / s) * 8;  // eight GPCs, conservative
      }
      cap[s] = ncl * s;
    }
    done = true;
  }
  return cap[S];
}

// pick (NT, S): one wave of (n-tiles x S) CTAs where possible; cost in "sub-tile stages" per CTA, with the activation
// traffic (re-staged once per CTA and stage, whatever NT) as a surcharge that shrinks with NT
static AwqPlan aq_plan(int M, int N, int K, int mt_pad) {
  const int sms = device_sm_count() > 0 ? device_sm_count() : 148;
  const int sub = N / AQ_BN, kb = K / 64;
  const int force_nt = tune(TUNE_AWQ_NT), force_s = tune(TUNE_AWQ_SPLITS);
  AwqPlan best{1, 1, 2, 0};
  double best_cost = 1e30;
  const int nt_max = mt_pad <= 128 ? 3 : (mt_pad <= 256 ? 2 : 1);  // NT * accumulator columns <= 512
  for (int nt = 1; nt <= nt_max; ++nt) {
    if (force_nt > 0 && nt != force_nt) continue;
    const int stages = aq_stages(mt_pad, nt);
    if (stages < 2) continue;
    const int tiles = (sub + nt - 1) / nt;
    const int P = mt_pad < 128 ? mt_pad : 128;
    const int scratch = AQ_NA * nt * AQ_BN * AQ_BK + stages * aq_stage_bytes(mt_pad, nt);
    for (int s = 1; s <= AQ_MAX_SPLITS; ++s) {
      if (force_s > 0 && s != force_s) continue;
      if (s > 1 && (tiles * s > aq_cluster_capacity(s) || kb / s < 2)) continue;
      const int slice = (P + s - 1) / s;
      if ((P + (s - 1) * slice) * AQ_BN * 4 > scratch) continue;
      const int waves = (tiles * s + sms - 1) / sms;
      // measured (tools/kbench.py awq, B200_AWQ_NT sweep): ~0.3 us per sub-tile stage, ~0.1 us of activation staging per
      // stage, and a write-out of ~2 us per sub-tile, ~6 us when it goes through the cluster exchange
      const double cost = (double)waves * ((kb + s - 1) / s) * (0.3 * nt + 0.1) + nt * (s > 1 ? 5.0 + 0.25 * s : 2.0);
      if (cost < best_cost) {
        best_cost = cost;
        best = AwqPlan{nt, s, stages, 0};
      }
    }
  }
  best.smem = 1024 + AQ_NA * best.nt * AQ_BN * AQ_BK + best.stages * aq_stage_bytes(mt_pad, best.nt) + 1024;
  return best;
}

template <int COLS, int GROUPS, int NT, bool BF>
static int aq_launch_t(const CUtensorMap& tw, const CUtensorMap& tx, const AwqParams& ap, int smem, cudaStream_t st) {
  auto kern = awq_fused_kernel<COLS, GROUPS, NT, BF>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeNonPortableClusterSizeAllowed, 0));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3((AQ_NONEPI + 4 * GROUPS) * 32);
  cfg.gridDim = dim3(ap.n_tiles, ap.splits);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = 1;
  attr[1].val.clusterDim.y = ap.splits;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  cfg.numAttrs = ap.splits > 1 ? 2 : 1;
  B200_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, ap));
  return launched("awq_fused_kernel launch");
}

template <int COLS, int GROUPS, bool BF>
static int aq_launch_nt(const CUtensorMap& tw, const CUtensorMap& tx, const AwqParams& ap, int nt, int smem, cudaStream_t st) {
  if (nt == 1) return aq_launch_t<COLS, GROUPS, 1, BF>(tw, tx, ap, smem, st);
  if constexpr (COLS * GROUPS * 2 <= 512) {
    if (nt == 2) return aq_launch_t<COLS, GROUPS, 2, BF>(tw, tx, ap, smem, st);
  }
  if constexpr (COLS * GROUPS * 3 <= 512) {
    if (nt == 3) return aq_launch_t<COLS, GROUPS, 3, BF>(tw, tx, ap, smem, st);
  }
  set_error("awq_fused: no instantiation for NT=%d", nt);
  return B200_ERR_UNSUPPORTED;
}

bool awq_fused_applicable(int M, int N, int K, int group) {
  return M >= 1 && M <= 256 && N % AQ_BN == 0 && K % 64 == 0 && K >= 128 && group % 8 == 0;
}

// x half [M, K] (ldx elements), qweight int32 [K, N/8] contiguous
int launch_awq_fused(const void* x, int64_t ldx, const int32_t* qweight, const void* scales, const int32_t* qzeros,
                     const void* bias, void* out, int64_t ldo, int dtype, int M, int N, int K, int group, cudaStream_t st) {
  AwqParams ap{};
  ap.qzeros = qzeros; ap.scales = (const uint16_t*)scales; ap.bias = bias; ap.out = out; ap.ldo = ldo;
  ap.bf16 = dtype == B200_BF16;
  ap.M = M; ap.N = N; ap.K = K; ap.group = group;
  ap.mt_pad = (M + 15) / 16 * 16;
  ap.kb_total = K / 64;
  const AwqPlan pl = aq_plan(M, N, K, ap.mt_pad);
  B200_REQUIRE(pl.stages >= 2, B200_ERR_UNSUPPORTED, "awq_fused: shared-memory plan does not fit (M=%d)", M);
  ap.splits = pl.splits;
  ap.stages = pl.stages;
  ap.n_tiles = (N / AQ_BN + pl.nt - 1) / pl.nt;
  {
    const int dbg = tune(TUNE_AWQ_DEBUG);
    if (dbg & 1)
      fprintf(stderr, "[awq_fused] M=%d N=%d K=%d NT=%d S=%d tiles=%d stages=%d smem=%d\n", M, N, K, pl.nt, pl.splits, ap.n_tiles, pl.stages, pl.smem);
  }
  // qweight as [N/128 sub-tiles][K rows][64 bytes]: one 3-D box of NT sub-tiles x 64 rows lands as NT dense 4 KB boxes
  CUtensorMap tw, tx;
  {
    PFN_encodeTiled enc = get_encode_tiled();
    B200_REQUIRE(enc != nullptr, B200_ERR_CUDA, "cuTensorMapEncodeTiled is not available from the driver");
    cuuint64_t dims[3] = {64, (cuuint64_t)K, (cuuint64_t)(N / AQ_BN)};
    cuuint64_t strides[2] = {(cuuint64_t)N / 2, 64};
    cuuint32_t box[3] = {64, 64, (cuuint32_t)pl.nt};
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = enc(&tw, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, const_cast<int32_t*>(qweight), dims, strides, box, estr,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    B200_REQUIRE(r == CUDA_SUCCESS, B200_ERR_CUDA, "cuTensorMapEncodeTiled (awq qweight) failed with CUresult %d", (int)r);
  }
  if (int rc = make_tmap_2d(&tx, x, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, M, (uint64_t)K * 2, ldx * 2, ap.mt_pad, AQ_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
#define AQ_GO(COLS, GROUPS)                                                                          \
  return ap.bf16 ? aq_launch_nt<COLS, GROUPS, true>(tw, tx, ap, pl.nt, pl.smem, st)                  \
                 : aq_launch_nt<COLS, GROUPS, false>(tw, tx, ap, pl.nt, pl.smem, st)
  if (ap.mt_pad <= 64) { AQ_GO(32, 2); }
  if (ap.mt_pad <= 128) { AQ_GO(32, 4); }
  AQ_GO(64, 4);
#undef AQ_GO
}

}  // namespace b200
