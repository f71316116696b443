// Decode-shape ("skinny", M <= 256) quantised GEMMs for sm_100a: swap-AB tcgen05 + cluster split-K.
//
//   MODE 0  fp8_blockwise_scaled_mm  (sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:345-423)
//   MODE 1  int8_scaled_mm           (sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-747)
//   MODE 2  AWQ W4A16 linear         (python/sglang/srt/layers/quantization/awq.py:186-204 = awq_dequantize,
//           sgl-kernel/csrc/gemm/awq_kernel.cu:126-221, followed by a half GEMM) fused: the packed int4 tile is the
//           only weight traffic (0.5 B per weight instead of 4.5 B through a dequantised copy)
//
// At decode M (64..256 tokens) these GEMMs stream the weight once and are bound by HBM, not by the tensor
// core, so the kernel is built to keep the copy engine busy on every SM and to keep everything else short:
//   * swap-AB: the WEIGHT tile [128 n x 128 k-bytes] is the UMMA "A" operand (128 TMEM lanes), the activations
//     [mt_pad x 128] are the "B" operand (UMMA N = M rounded up to 16): no MMA rows are spent on padding;
//   * grid = (n-tiles, S): the S CTAs of one n-tile form a thread-block CLUSTER and split K between them; each
//     streams its K range through a TMA ring, and the partial tiles meet in DISTRIBUTED SHARED MEMORY as a
//     reduce-scatter (CTA r sums tokens [r*mt/S, (r+1)*mt/S) of all S partials in fixed order, bit-reproducible,
//     and writes them) -- no global workspace, no tickets, no L2 round trips.  (The first version was a
//     persistent stream-K kernel whose partial tiles met in an fp32 global workspace: that fix-up alone cost
//     10-50 us per launch, profiles/r01_gemm_notes.md.)
//   * up to two CTAs are resident per SM at M <= 64, so 296 CTA slots take whatever (tiles x S) the shape gives;
//   * warps 0/3 = TMA producers (weights evict-first, prefetched before the programmatic-launch wait), warp 1 =
//     single-thread tcgen05.mma issuer, warp 2 = TMEM allocator, warps 4.. = epilogue (one TMEM lane = one
//     output column n per thread).
//   FP8: every 128-deep k-block lands in its own TMEM buffer (ring of 2-4) and is promoted
//     acc[m] += tmem[m] * (sa[m,kb] * sb[kb,nb]) in fp32 registers (UMMA has no fp32 block scaling); the scale
//     products of the CTA's K range are tabulated in shared memory once, so the k-loop has no global loads and
//     no CTA barriers;
//   INT8: int32 accumulation in TMEM over the CTA's K range (exact), integer reduce-scatter, scales at the end.
//   AWQ: TMA brings the raw [64 k x 16 words] int4 box (4 KB) and the activation tile; the epilogue warps, idle during
//     the k-loop, are the dequantisers: group g takes every GROUPS-th stage, a thread turns 8 k-rows x one packed word
//     into eight 16-byte K-major chunks ((nib | magic) - (zero | magic)) * scale in half2 -- the values awq_dequantize
//     produces, bit for bit -- and stores them into the SWIZZLE_128B operand tile (UMMA row i holds column
//     n = 8 (i % 16) + i / 16 of the tile, which makes the stores bank-conflict free); fp32 accumulation in TMEM.
#include <stdio.h>
#include <stdlib.h>

#include "gemm_skinny.h"

namespace b200 {

constexpr int SK_BN = 128;  // weight rows per tile (UMMA M)
constexpr int SK_BK = 128;  // K bytes per pipeline stage (one 128-byte swizzle row, one fp8 scale block)
constexpr int SK_MAX_STAGES = 12;
constexpr int SK_MAX_SPLITS = 8;
constexpr int SK_NONEPI = 4;    // warps 0..3: warp 1 issues the MMAs, warp 2 allocates tensor memory, warps 0 and 3 are TMA producers
constexpr int SK_NPROD = 2;
constexpr int SK_FIXED_SMEM = 1024 /*align*/ + 512 /*barriers*/;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
// .relaxed: the two cluster barriers of the write-out order EXECUTION only (a peer's ring is drained / my outgoing bulk
// copies have landed); the data itself is ordered by fence.proxy.async + the mbarrier transaction count.  The .release form
// compiles to MEMBAR.ALL.GPU, which waits for every outstanding global store of the CTA -- measured with the time stamps
// below: 1.5 us at barrier A and 3 us at barrier B (behind the output stores) of an 11 us kernel.
__device__ __forceinline__ void cluster_arrive() { asm volatile("barrier.cluster.arrive.relaxed.aligned;" ::: "memory"); }
__device__ __forceinline__ void cluster_wait() { asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory"); }
__device__ __forceinline__ uint32_t mapa_u32(uint32_t smem_addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(smem_addr), "r"(rank));
  return r;
}
// shared::cta -> shared::cluster bulk copy, completes (complete_tx) on an mbarrier of the destination CTA
__device__ __forceinline__ void bulk_s2c(uint32_t dst_cluster, uint32_t src_cta, uint32_t bytes, uint32_t bar_cluster) {
  asm volatile("cp.async.bulk.shared::cluster.shared::cta.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst_cluster),
               "r"(src_cta), "r"(bytes), "r"(bar_cluster)
               : "memory");
}

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}
#define SK_STAMP(i)                                                                               \
  do {                                                                                            \
    if (p.stamps && threadIdx.x == SK_NONEPI * 32) p.stamps[(blockIdx.y * gridDim.x + blockIdx.x) * 8 + (i)] = gtimer(); \
  } while (0)

// Write-out of a qkv tile with fused neox RoPE (this tile is one q / k head: head_dim == SK_BN).  A thread also sums its
// PARTNER columns (n_in ^ 64), rounds both pairs to the output type -- what the plain GEMM would have stored and
// rope_neox_kernel read back -- and rotates with the helpers that kernel uses: same bits, one launch less.  cos / sin come
// from a per-step table (rope_cos_sin_kernel: the same rope_sincos, once per decode step instead of once per layer); with
// sincosf inlined here its large-argument path cost a stack frame and spills, and the GEMM lost half its speed.
__device__ __forceinline__ void sk_rope_store(const SkinnyParams& p, const float* send, int rank, int S, int my0, int P, int slice,
                                           int mine, int t0, int tq, int tstep, int n_in, int n, bool pair_ok) {
  constexpr int half = SK_BN / 2;
  const int i0 = n_in & (half - 1), pdelta = (n_in ^ half) - n_in;
  for (int tk = tq; tk < mine; tk += tstep) {
    const int m = t0 + my0 + tk;
    if (m >= p.M || n >= p.N) continue;
    // partial of CTA s2 (same slots and the same summation order as the plain write-out; no arrays: this instantiation must
    // stay free of spills -- a stack frame cost the whole kernel 10 us)
    float o0 = 0.f, o1 = 0.f, q0 = 0.f, q1 = 0.f;
#pragma unroll
    for (int s2 = 0; s2 < SK_MAX_SPLITS; ++s2) {
      if (s2 < S) {
        int slot = s2 - rank - 1;
        if (slot < 0) slot += S;
        const float* src = send + (s2 == rank ? my0 * SK_BN : (P + slot * slice) * SK_BN) + n_in + tk * SK_BN;
        const float2 v = *reinterpret_cast<const float2*>(src), w = *reinterpret_cast<const float2*>(src + pdelta);
        if (s2 == 0) { o0 = v.x; o1 = v.y; q0 = w.x; q1 = w.y; }
        else { o0 += v.x; o1 += v.y; q0 += w.x; q1 += w.y; }
      }
    }
    auto rt = [&](float x) { return p.out_bf16 ? half_to_float<true>(float_to_half<true>(x)) : half_to_float<false>(float_to_half<false>(x)); };
    o0 = rt(o0); o1 = rt(o1); q0 = rt(q0); q1 = rt(q1);
    const float4 cs4 = __ldg(reinterpret_cast<const float4*>(p.rope_cs + (size_t)m * half + i0));   // i0 is even
    const float cs0 = cs4.x, sn0 = cs4.y, cs1 = cs4.z, sn1 = cs4.w;
    float r0, r1;
    if (n_in < half) { r0 = rope_lo(o0, q0, cs0, sn0); r1 = rope_lo(o1, q1, cs1, sn1); }
    else { r0 = rope_hi(q0, o0, cs0, sn0); r1 = rope_hi(q1, o1, cs1, sn1); }
    uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + (size_t)m * p.ldo + n;
    if (pair_ok) {
      *reinterpret_cast<uint32_t*>(dst) = p.out_bf16 ? pack_half2<true>(r0, r1) : pack_half2<false>(r0, r1);
    } else {
      dst[0] = p.out_bf16 ? float_to_half<true>(r0) : float_to_half<false>(r0);
      if (n + 1 < p.N) dst[1] = p.out_bf16 ? float_to_half<true>(r1) : float_to_half<false>(r1);
    }
  }
}

template <int GROUPS>
__device__ __forceinline__ void sk_epi_bar_all() {
  asm volatile("bar.sync 1, %0;" ::"n"(GROUPS * 128) : "memory");
}

// MODE 2: dequantise the raw int4 boxes of stages grp, grp + GROUPS, ... into the K-major SWIZZLE_128B operand tile.
// `row` = thread index inside its 128-thread group: packed word column c = row % 16, k-octet ko = row / 16.
template <bool BF, int GROUPS>
__device__ __forceinline__ void awq_dequant_stages(const SkinnyParams& p, uint8_t* smem, uint64_t* full, uint64_t* afull,
                                                   int stage_bytes, int x_bytes, int tile, int kb0, int nkb, int rot,
                                                   int grp, int row) {
  const int c = row & 15, ko = row >> 4, h = ko & 1;
  constexpr uint32_t magic = BF ? 0x43004300u : 0x64006400u;  // 128 + nibble (bf16) / 1024 + nibble (f16), exact
  const int ncol = tile * 16 + c;
  const bool col_ok = ncol * 8 < p.N;
  int cur_g = -1;
  uint32_t zp[4] = {0, 0, 0, 0}, s2[4] = {0, 0, 0, 0};
  for (int i = grp; i < nkb; i += GROUPS) {
    const int s = i % p.stages;
    const uint32_t ph = (i / p.stages) & 1;
    const int kb = kb0 + (i + rot >= nkb ? i + rot - nkb : i + rot);
    const int gi = (kb * 64 + ko * 8) / p.group;  // a thread's 8 k-rows share a group (group % 8 == 0)
    if (gi != cur_g) {
      cur_g = gi;
      uint4 sraw = make_uint4(0u, 0u, 0u, 0u);
      uint32_t z = 0;
      if (col_ok) {
        sraw = __ldg(reinterpret_cast<const uint4*>(p.scales + ((int64_t)gi * p.N + (int64_t)ncol * 8)));
        z = (uint32_t)__ldg(p.qzeros + (int64_t)gi * (p.N / 8) + ncol);
      }
      s2[0] = sraw.x; s2[1] = sraw.y; s2[2] = sraw.z; s2[3] = sraw.w;
#pragma unroll
      for (int q = 0; q < 4; ++q) zp[q] = ((z >> (4 * q)) & 0x000F000Fu) | magic;
    }
    mbar_wait(&full[s], ph);
    if (p.debug & 32) {  // developer probe: no dequantisation (wrong results), what is left is the pipeline
      mbar_arrive(&afull[s]);
      continue;
    }
    uint8_t* st = smem + s * stage_bytes;
    const uint32_t* raw = reinterpret_cast<const uint32_t*>(st + SK_BN * SK_BK + x_bytes);
    uint32_t tmp[8], w[8];
    // the two half-warps read rows 512 B apart: opposite row parity keeps them on different banks
#pragma unroll
    for (int r = 0; r < 8; ++r) tmp[r] = raw[(ko * 8 + (r ^ h)) * 16 + c];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      w[2 * q] = h ? tmp[2 * q + 1] : tmp[2 * q];
      w[2 * q + 1] = h ? tmp[2 * q] : tmp[2 * q + 1];
    }
    uint8_t* arow = st + c * 128 + ((ko ^ (c & 7)) << 4);
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
    fence_proxy_async();  // generic-proxy stores -> read by the tensor core through the async proxy
    mbar_arrive(&afull[s]);
  }
}

template <int MODE, int COLS, int GROUPS, bool ROPE = false>
// (M <= 64: small enough in registers and shared memory for two CTAs per SM -- the prologue and write-out of one
// overlap the streaming of the other, also across consecutive launches)
__global__ void __launch_bounds__((SK_NONEPI + 4 * GROUPS) * 32, (COLS * GROUPS <= 64) ? 2 : 1)
gemm_skinny_kernel(const __grid_constant__ CUtensorMap tma_w, const __grid_constant__ CUtensorMap tma_x,
                   const __grid_constant__ SkinnyParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  // (pointer arithmetic, not an integer round trip: the compiler keeps the shared address space and emits LDS/STS)
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  constexpr int BUFCOLS = GROUPS * COLS;
  constexpr int NBUF = MODE == 0 ? (BUFCOLS <= 128 ? 4 : 2) : 1;
  constexpr uint32_t kTmemCols = NBUF * BUFCOLS < 32 ? 32 : NBUF * BUFCOLS;
  constexpr int NE = GROUPS * 128;  // epilogue threads
  constexpr int RAW_BYTES = MODE == 2 ? 64 * 64 : 0;  // AWQ: 64 k-rows x 16 packed words
  const int x_bytes = p.mt_pad * SK_BK;
  const int stage_bytes = SK_BN * SK_BK + x_bytes + RAW_BYTES;
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.stages * stage_bytes);
  uint64_t* empty = full + SK_MAX_STAGES;
  uint64_t* tfull = empty + SK_MAX_STAGES;
  uint64_t* tempty = tfull + 4;
  uint64_t* xbar = tempty + 4;  // split-K exchange: incoming partial slices (one phase per pass)
  uint64_t* afull = xbar + 1;  // MODE 2: the dequantised operand tile of the stage is complete
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(afull + SK_MAX_STAGES);
  float* tab = reinterpret_cast<float*>(smem + p.stages * stage_bytes + 512);  // MODE 0: [tab_cap][mt_pad]; MODE 1: sa[256]
  // after the k-loop the (drained) stage ring is reused: send[P][128] fp32 then recv[S-1][slice][128]
  float* send = reinterpret_cast<float*>(smem);

  // (shuffle: provably warp-uniform for the compiler -> the service roles keep their state in uniform registers and
  // issue UTMALDG / UTCxMMA / UTCBAR without an ELECT + R2UR.BROADCAST loop around each one; see gemm_big.cu)
  const int warp = __shfl_sync(0xffffffffu, (int)(threadIdx.x >> 5), 0), lane = threadIdx.x & 31;
  const int tile = blockIdx.x;
  SK_STAMP(0);
  const int S = p.splits;
  const int rank = S > 1 ? (int)cluster_ctarank() : 0;
  const int kb0 = (int)((long long)rank * p.kb_total / S), kb1 = (int)((long long)(rank + 1) * p.kb_total / S);
  const int nkb = kb1 - kb0;
  // k-blocks are visited in a per-tile rotated order so that the CTAs do not all pull the same activation lines from L2 at once
  const int rot = (p.debug & 16) ? (int)((blockIdx.x * 7u) % (unsigned)nkb) : 0;
#define SK_KB(i) (kb0 + ((i) + rot >= nkb ? (i) + rot - nkb : (i) + rot))

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_w);
    tma_prefetch_desc(&tma_x);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.stages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
      mbar_init(&afull[s], 128);
    }
    for (int a = 0; a < NBUF; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], NE / 32);  // one arrival per epilogue WARP (512 per-thread arrivals on one barrier serialise)
    }
    mbar_init(xbar, 1);
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  griddep_launch();  // let the next kernel of the stream start its prologue (no-op without PDL)

  if (warp == 0 || warp == 3) {
    // ================================ TMA producers (k-block i belongs to producer i % SK_NPROD) ================================
    // whole warp in the loop (uniform control flow), one elected lane issues
    const int pidx = warp == 0 ? 0 : 1;
    const uint64_t pol_w = make_policy_evict_first();
    // weights do not depend on the previous kernel: fill the ring with them before waiting for it
    const int pre = nkb < p.stages ? nkb : p.stages;
    const uint32_t tx_bytes = MODE == 2 ? RAW_BYTES + ((p.debug & 8) ? 0 : x_bytes) : ((p.debug & 8) ? SK_BN * SK_BK : stage_bytes);
    const bool issuer = elect_one();
    int i = pidx;
    for (; i < pre; i += SK_NPROD) {
      if (issuer) {
        mbar_arrive_expect_tx(&full[i], tx_bytes);
        if (MODE == 2) tma_load_2d_hint(smem + i * stage_bytes + SK_BN * SK_BK + x_bytes, &tma_w, tile * 64, SK_KB(i) * 64, &full[i], pol_w);
        else tma_load_2d_hint(smem + i * stage_bytes, &tma_w, SK_KB(i) * SK_BK, tile * SK_BN, &full[i], pol_w);
      }
    }
    griddep_wait();
    if (!(p.debug & 8)) {
      for (int j = pidx; j < pre; j += SK_NPROD)
        if (issuer) tma_load_2d(smem + j * stage_bytes + SK_BN * SK_BK, &tma_x, SK_KB(j) * SK_BK, 0, &full[j]);
    }
    // i >= stages from here on: stage / phase kept incrementally (no division by the run-time ring depth)
    int sidx = i - p.stages;
    uint32_t ph = 1;
    while (sidx >= p.stages) { sidx -= p.stages; ph ^= 1; }
    for (; i < nkb; i += SK_NPROD) {
      mbar_wait(&empty[sidx], ph ^ 1);
      if (issuer) {
        mbar_arrive_expect_tx(&full[sidx], tx_bytes);
        uint8_t* st = smem + sidx * stage_bytes;
        if (MODE == 2) tma_load_2d_hint(st + SK_BN * SK_BK + x_bytes, &tma_w, tile * 64, SK_KB(i) * 64, &full[sidx], pol_w);
        else tma_load_2d_hint(st, &tma_w, SK_KB(i) * SK_BK, tile * SK_BN, &full[sidx], pol_w);
        if (!(p.debug & 8)) tma_load_2d(st + SK_BN * SK_BK, &tma_x, SK_KB(i) * SK_BK, 0, &full[sidx]);
      }
      __syncwarp();
      sidx += SK_NPROD;
      if (sidx >= p.stages) { sidx -= p.stages; ph ^= 1; }
    }
  } else if (warp == 1) {
    // ================================ MMA issuer (one elected lane) ================================
    const uint32_t hfmt = p.out_bf16 ? 1u : 0u;  // kind::f16 operand format: 0 = f16, 1 = bf16
    const uint32_t idesc = MODE == 0   ? umma_idesc(1 /*f32*/, 0 /*e4m3*/, 0, SK_BN, p.mt_pad)
                           : MODE == 1 ? umma_idesc(2 /*s32*/, 1 /*s8*/, 1, SK_BN, p.mt_pad)
                                       : umma_idesc(1 /*f32*/, hfmt, hfmt, SK_BN, p.mt_pad);
    int sidx = 0, a = 0;
    uint32_t ph = 0, aph = 0;
    const uint64_t desc0 = umma_desc_k_sw128(smem_u32(smem));
    for (int i = 0; i < nkb; ++i) {
      if (MODE == 0) mbar_wait(&tempty[a], aph ^ 1);
      mbar_wait(&full[sidx], ph);
      if (MODE == 2) mbar_wait(&afull[sidx], ph);
      tc_fence_after();
      const uint64_t adesc = desc0 + (uint64_t)((uint32_t)(sidx * stage_bytes) >> 4);
      const uint64_t bdesc = adesc + ((SK_BN * SK_BK) >> 4);
      const uint32_t d = tmem_base + a * BUFCOLS;
      if (elect_one()) {
#pragma unroll
        for (int k = 0; k < SK_BK / 32; ++k) {
          const uint32_t accum = MODE == 0 ? (k > 0) : (i > 0 || k > 0);
          if (MODE == 0) umma_f8f6f4(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
          else if (MODE == 1) umma_i8(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
          else umma_f16(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
        }
        tc_commit(&empty[sidx]);
        if (MODE == 0 || i == nkb - 1) tc_commit(&tfull[a]);
      }
      __syncwarp();
      if (++sidx == p.stages) { sidx = 0; ph ^= 1; }
      if (MODE == 0) {
        if (++a == NBUF) { a = 0; aph ^= 1; }
      }
    }
  }

  // ================================ epilogue warps: promote / collect the accumulator ================================
  const int ew = warp - SK_NONEPI, grp = ew >> 2, quad = ew & 3;  // quad == warp % 4: TMEM lanes [32*quad, +32)
  const int row = quad * 32 + lane;                       // output column n inside the tile
  const int col0 = grp * COLS;                            // first token handled by this group
  const int et_all = threadIdx.x - SK_NONEPI * 32;                   // 0 .. NE-1 for the epilogue threads
  int ncols = p.mt_pad - col0;
  ncols = ncols < 0 ? 0 : (ncols > COLS ? COLS : ncols);
  const int nchunk = (ncols + 31) / 32;
  float acc[COLS];

  if (warp >= SK_NONEPI) {
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quad * 32) << 16) + col0;
    SK_STAMP(1);
    griddep_wait();  // activation scales and the output buffer belong to the stream order
#pragma unroll
    for (int j = 0; j < COLS; ++j) acc[j] = 0.f;

    if (MODE == 0) {
      for (int c0 = 0; c0 < nkb; c0 += p.tab_cap) {
        const int cn = nkb - c0 < p.tab_cap ? nkb - c0 : p.tab_cap;
        if (c0 > 0) sk_epi_bar_all<GROUPS>();  // everyone is done reading the previous chunk of the table
        // scale products of this chunk: tab[kbl][m] = sa[m, kb] * sb[kb, tile]
        for (int e = et_all; e < cn * p.mt_pad; e += NE) {
          const int kbl = e / p.mt_pad, m = e - kbl * p.mt_pad;
          const int kb = SK_KB(c0 + kbl);
          float v = 0.f;
          if (m < p.M)
            v = __ldg(p.sa + (int64_t)m * p.sa_sm + (int64_t)kb * p.sa_sk) * __ldg(p.sb + (int64_t)kb * p.sb_sk + (int64_t)tile * p.sb_sn);
          tab[e] = v;
        }
        sk_epi_bar_all<GROUPS>();
        for (int i = c0; i < c0 + cn; ++i) {
          const int a = i % NBUF;
          const float* sv = tab + (i - c0) * p.mt_pad + col0;
          mbar_wait(&tfull[a], (i / NBUF) & 1);
          tc_fence_after();
          if constexpr (COLS == 32) {
            uint32_t r[32];
            tmem_ld_32x32(lane_addr + a * BUFCOLS, r);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[a]);  // the values are in registers: the tensor core may refill the buffer
#pragma unroll
            for (int j = 0; j < 32; j += 4) {
              const float4 s4 = *reinterpret_cast<const float4*>(sv + j);
              acc[j + 0] = fmaf(__uint_as_float(r[j + 0]), s4.x, acc[j + 0]);
              acc[j + 1] = fmaf(__uint_as_float(r[j + 1]), s4.y, acc[j + 1]);
              acc[j + 2] = fmaf(__uint_as_float(r[j + 2]), s4.z, acc[j + 2]);
              acc[j + 3] = fmaf(__uint_as_float(r[j + 3]), s4.w, acc[j + 3]);
            }
          } else {
#pragma unroll
            for (int c = 0; c < COLS / 32; ++c) {
              if (c < nchunk) {
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
            __syncwarp();
            if (lane == 0) mbar_arrive(&tempty[a]);
          }
        }
      }
    } else {
      if (MODE == 1) {
        for (int m = et_all; m < 256; m += NE) tab[m] = m < p.M ? __ldg(p.sa + m) : 0.f;  // per-token scales
      } else {
        if (p.out_bf16) awq_dequant_stages<true, GROUPS>(p, smem, full, afull, stage_bytes, x_bytes, tile, kb0, nkb, rot, grp, row);
        else awq_dequant_stages<false, GROUPS>(p, smem, full, afull, stage_bytes, x_bytes, tile, kb0, nkb, rot, grp, row);
      }
      mbar_wait(&tfull[0], 0);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < COLS / 32; ++c) {
        if (c < nchunk) {
          uint32_t r[32];
          tmem_ld_32x32(lane_addr + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[c * 32 + j] = __uint_as_float(r[j]);  // raw int32 bits (MODE 1) / fp32 (MODE 2)
        }
      }
      tc_fence_before();
    }
    SK_STAMP(2);
  }

  // ================================ write-out: [split-K exchange through DSMEM,] sum, coalesced store ================================
  // pass = up to 128 tokens; within a pass CTA r of the cluster owns tokens [r*slice, (r+1)*slice)
  const int P = p.mt_pad < 128 ? p.mt_pad : 128;
  const int passes = (p.mt_pad + P - 1) / P;
  const int slice = (P + S - 1) / S;
  float* recv = send + P * SK_BN;  // [S-1][slice][128]: slot (src - me - 1) mod S holds the partial of CTA src
  const int srow = MODE == 2 ? 8 * (row & 15) + (row >> 4) : row;  // tile column held by this TMEM lane
  for (int ps = 0; ps < passes; ++ps) {
    const int t0 = ps * P;
    const int pt = p.mt_pad - t0 < P ? p.mt_pad - t0 : P;
    const int my0 = rank * slice;
    int mine = pt - my0;
    mine = mine < 0 ? 0 : (mine > slice ? slice : mine);
    // 1. this CTA's partial accumulators of the pass -> send[token][n]  (the ring is drained: the epilogue has seen the last MMA)
    if (warp >= SK_NONEPI && col0 >= t0 && col0 < t0 + pt) {
#pragma unroll
      for (int j = 0; j < COLS; ++j)
        if (j < ncols) send[(col0 - t0 + j) * SK_BN + srow] = acc[j];
      if (S > 1) fence_proxy_async();  // the bulk copies below read these through the async proxy
    }
    if (S > 1) {
      // send[] is complete for every thread of this CTA (CTA-scope ordering: the cluster barrier below is relaxed)
      if (warp >= SK_NONEPI) sk_epi_bar_all<GROUPS>();
      if (threadIdx.x == SK_NONEPI * 32) mbar_arrive_expect_tx(xbar, (uint32_t)((S - 1) * mine * SK_BN * 4));  // armed before anyone can send
      __syncwarp();
      cluster_arrive();  // A: every CTA of the cluster has drained its ring and armed its barrier
      cluster_wait();
      SK_STAMP(3);
      // one lane per destination CTA (a single thread looping over the peers took 1.2 us at S = 6)
      if (warp == SK_NONEPI && !(p.debug & 1)) {
        const int o = lane;
        if (o < S && o != rank) {
          const int o0 = o * slice;
          int on = pt - o0;
          on = on < 0 ? 0 : (on > slice ? slice : on);
          if (on > 0) {
            int slot = rank - o - 1;
            if (slot < 0) slot += S;
            bulk_s2c(mapa_u32(smem_u32(recv + slot * slice * SK_BN), (uint32_t)o), smem_u32(send + o0 * SK_BN),
                     (uint32_t)(on * SK_BN * 4), mapa_u32(smem_u32(xbar), (uint32_t)o));
          }
        }
        __syncwarp();
      }
      SK_STAMP(4);
      if (warp >= SK_NONEPI && !(p.debug & 1)) mbar_wait(xbar, ps & 1);
      SK_STAMP(5);
    } else if (warp >= SK_NONEPI) {
      SK_STAMP(3);
      sk_epi_bar_all<GROUPS>();
      SK_STAMP(4);
    }
    // 2. sum the S partials of the owned tokens in CTA order (bit-reproducible), scale (int8), convert, store:
    //    a thread owns two adjacent output columns, a warp writes whole 128-byte lines
    if (warp >= SK_NONEPI) {
      const int pi = et_all & 63, tq = et_all >> 6;
      const int n_in = 2 * pi, n = tile * SK_BN + n_in;
      const bool pair_ok = (n + 1 < p.N) && (p.ldo % 2 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 3) == 0);
      float sb0 = 1.f, sb1 = 1.f, b0 = 0.f, b1 = 0.f;
      if (MODE != 0) {
        if (MODE == 1) {
          sb0 = n < p.N ? __ldg(p.sb + n) : 0.f;
          sb1 = n + 1 < p.N ? __ldg(p.sb + n + 1) : 0.f;
        }
        if (p.bias) {
          const uint16_t* bp = reinterpret_cast<const uint16_t*>(p.bias);
          const uint16_t u0 = n < p.N ? __ldg(bp + n) : (uint16_t)0, u1 = n + 1 < p.N ? __ldg(bp + n + 1) : (uint16_t)0;
          b0 = p.out_bf16 ? half_to_float<true>(u0) : half_to_float<false>(u0);
          b1 = p.out_bf16 ? half_to_float<true>(u1) : half_to_float<false>(u1);
        }
      }
      // where the partial of CTA s2 for my first token sits (floats from `send`): no division inside the token loop
      int poff[SK_MAX_SPLITS];
#pragma unroll
      for (int s2 = 0; s2 < SK_MAX_SPLITS; ++s2) {
        int slot = s2 - rank - 1;
        if (slot < 0) slot += S;
        poff[s2] = (s2 == rank ? my0 * SK_BN : (P + slot * slice) * SK_BN) + n_in;
      }
      bool roped = false;
      if constexpr (ROPE) {
        if (tile * SK_BN < p.rope_cols) {
          sk_rope_store(p, send, rank, S, my0, P, slice, mine, t0, tq, NE / 64, n_in, n, pair_ok);
          roped = true;
        }
      }
      if (!roped) {
#pragma unroll 2
      for (int tk = tq; tk < mine; tk += NE / 64) {
        const int m = t0 + my0 + tk;
        if (m >= p.M || n >= p.N) continue;
        float2 v[SK_MAX_SPLITS];
#pragma unroll
        for (int s2 = 0; s2 < SK_MAX_SPLITS; ++s2)
          if (s2 < S) v[s2] = *reinterpret_cast<const float2*>(send + poff[s2] + tk * SK_BN);
        float o0, o1;
        if (MODE != 1) {
          o0 = v[0].x;
          o1 = v[0].y;
#pragma unroll
          for (int s2 = 1; s2 < SK_MAX_SPLITS; ++s2)
            if (s2 < S) {
              o0 += v[s2].x;
              o1 += v[s2].y;
            }
          if (MODE == 2 && p.bias) {
            o0 += b0;
            o1 += b1;
          }
        } else {
          int i0 = __float_as_int(v[0].x), i1 = __float_as_int(v[0].y);
#pragma unroll
          for (int s2 = 1; s2 < SK_MAX_SPLITS; ++s2)
            if (s2 < S) {
              i0 += __float_as_int(v[s2].x);
              i1 += __float_as_int(v[s2].y);
            }
          const float sam = tab[m];
          o0 = __fmul_rn(__fmul_rn((float)i0, sam), sb0);
          o1 = __fmul_rn(__fmul_rn((float)i1, sam), sb1);
          if (p.bias) {
            o0 = __fadd_rn(o0, b0);
            o1 = __fadd_rn(o1, b1);
          }
        }
        uint16_t* dst = reinterpret_cast<uint16_t*>(p.out) + (size_t)m * p.ldo + n;
        if (pair_ok) {
          *reinterpret_cast<uint32_t*>(dst) = p.out_bf16 ? pack_half2<true>(o0, o1) : pack_half2<false>(o0, o1);
        } else {
          dst[0] = p.out_bf16 ? float_to_half<true>(o0) : float_to_half<false>(o0);
          if (n + 1 < p.N) dst[1] = p.out_bf16 ? float_to_half<true>(o1) : float_to_half<false>(o1);
        }
      }
      }
    }
    if (S > 1) {
      __syncwarp();
      cluster_arrive();  // B: my outgoing copies have been consumed, nobody reads my shared memory any more
      cluster_wait();
    } else if (ps + 1 < passes && warp >= SK_NONEPI) {
      sk_epi_bar_all<GROUPS>();
    }
  }

  SK_STAMP(6);
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, kTmemCols);
  }
}

}  // namespace b200

namespace b200 {

static int skinny_mt_pad(int M) { return (M + 15) / 16 * 16; }

// the kernel needs no global workspace any more (the K slices meet in distributed shared memory)
size_t skinny_workspace_bytes(int, int, int) { return 2048 * 8 * 8; }  // (developer time stamps only)

bool skinny_applicable(int M) { return M <= 256; }

// FP8 block-scaled: skinny (swap-AB, cluster split-K) or the persistent large-M kernel?  Measured policy
// (tools/kb_cross.py over the TP-local shapes of cfg 2/4/5 at M 32 / 64 / 128, profiles/r02_kb_cross.txt, second table:
// after the write-out of this kernel lost its GPU-scope fences and its divisions): the large-M kernel wins at every
// M > 128, on short K loops (K <= 2304 from M = 49; K < 2048 from M = 33; K <= 1024 from M = 17), on the widest weights at
// K <= 4096 from M = 49 (gate_up: 224 tiles of 128 columns stream better from the persistent kernel) and, at
// 64 < M <= 128, below K = 3584 or where K = 3584 meets N > 4096; the skinny kernel wins everywhere else -- its split-K
// puts every SM to work on few-tile shapes (third table of r02b_kb_cross.txt: M 16 / 48 / 96 / 128).
// Knob sk_max_m >= 0 forces "skinny iff M <= sk_max_m".
bool skinny_applicable_fp8(int M, int N, int K) {
  const int max_m = tune(TUNE_SK_MAX_M);
  if (max_m >= 0) return M <= (max_m > 256 ? 256 : max_m);
  if (M > 128) return false;
  if (M <= 16) return true;
  if (M <= 32) return K >= 1536;
  if (M <= 48) return K >= 2048;                       // (gate_up too: 28672 x 4096 at M = 48 23.8 us against 26.7)
  if (K <= 2304) return false;
  if (M <= 64) return !(K <= 4096 && N >= 16384);
  if (K >= 7168) return true;                          // 64 < M <= 128 from here
  if (K >= 4096) return N < 16384;
  return K >= 3584 && N <= 4096;
}

// INT8 (per-token / per-channel scales, exact int32 accumulation in TMEM over the whole K range: no per-k-block promotion,
// so the large-M kernel has less to lose than in FP8): measured crossover in tools/kb_cross.py --int8.
bool skinny_applicable_int8(int M, int N, int K) {
  const int max_m = tune(TUNE_SK_MAX_M);
  if (max_m >= 0) return M <= (max_m > 256 ? 256 : max_m);
  // profiles/r02b_kb_cross_int8.txt: the decode-shape kernel wins at every M <= 128 except the widest weight at K <= 4096
  // (28672 x 4096 at M = 128: 31.3 us against 28.4), and at M = 129 .. 256 on long K or few tiles (4096 x 14336: 27 against
  // 42; 1280 x 8192: 19.5 against 27.7); the large-M kernel takes the rest (4096 x 1024 at M = 256: 12.3 against 16.7)
  if (M > 256) return false;   // (the decode-shape kernel's limit)
  if (M <= 64) return true;
  if (M <= 128) return !(K <= 4096 && N >= 16384);
  return K >= 7168 || (K >= 4096 && N <= 4096);
}

struct SkinnyPlan {
  int stages, tab_cap, smem;
};

// shared-memory plan: the scale table, the barriers and as many ring stages as fit (two CTAs per SM at M <= 64)
static SkinnyPlan skinny_plan(int mode, int mt_pad) {
  SkinnyPlan pl;
  // (kernels that allocate tensor memory are scheduled one CTA per SM: cudaOccupancyMaxActiveBlocksPerMultiprocessor
  // answers 1 whatever the shared-memory and register footprint, measured)
  const int budget = mt_pad <= 64 ? 112 * 1024 : 226 * 1024;
  const int stage_bytes = SK_BN * SK_BK + mt_pad * SK_BK + (mode == 2 ? 4096 : 0);
  const int tab_bytes = mode == 0 ? (mt_pad <= 64 ? 8192 : mt_pad <= 128 ? 16384 : 32768) : 1024;
  pl.tab_cap = mode == 0 ? tab_bytes / (mt_pad * 4) : 0;
  int stages = (budget - SK_FIXED_SMEM - tab_bytes) / stage_bytes;
  if (stages > SK_MAX_STAGES) stages = SK_MAX_STAGES;
  if (const int cap = tune(TUNE_SK_STAGES); cap > 0 && stages > cap) stages = cap;
  pl.stages = stages;
  pl.smem = stages * stage_bytes + SK_FIXED_SMEM + tab_bytes;
  return pl;
}

template <int MODE, int COLS, int GROUPS, bool ROPE = false>
static int launch_skinny_t(const CUtensorMap& tw, const CUtensorMap& tx, SkinnyParams& sp, int smem, cudaStream_t st) {
  auto kern = gemm_skinny_kernel<MODE, COLS, GROUPS, ROPE>;
  constexpr int kThreads = (SK_NONEPI + 4 * GROUPS) * 32;
  static bool configured = false;
  static int capacity[SK_MAX_SPLITS + 1] = {0};  // CTAs that can be resident at once with clusters of S = 1..8
  static int capacity_smem = -1;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    configured = true;
  }
  cudaLaunchConfig_t cfg = {};
  cfg.blockDim = dim3(kThreads);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute attr[2];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = pdl_enabled() ? 1 : 0;
  attr[1].id = cudaLaunchAttributeClusterDimension;
  attr[1].val.clusterDim.x = 1;
  attr[1].val.clusterDim.z = 1;
  cfg.attrs = attr;
  if (capacity_smem != smem) {  // (the plan depends on mt_pad only through smem)
    for (int S = 1; S <= SK_MAX_SPLITS; ++S) {
      cfg.gridDim = dim3(1, S);
      attr[1].val.clusterDim.y = S;
      cfg.numAttrs = 2;
      int ncl = 0;
      B200_CUDA(cudaOccupancyMaxActiveClusters(&ncl, kern, &cfg));
      // (the occupancy API answers one CTA per SM for kernels that allocate tensor memory; two do co-reside when
      // registers, shared memory and TMEM columns allow it -- measured -- so the M <= 64 variants count double)
      capacity[S] = ncl * S * ((COLS * GROUPS <= 64) ? 2 : 1);
    }
    capacity_smem = smem;
  }
  // K slices per n-tile = cluster size: the S that puts the most CTAs to work while every CTA of (tiles x S) is
  // resident at once, owns at least two k-blocks, and the exchange buffers fit in the drained ring
  const int ring_bytes = sp.stages * (SK_BN * SK_BK + sp.mt_pad * SK_BK + (MODE == 2 ? 4096 : 0));
  const int P = sp.mt_pad < 128 ? sp.mt_pad : 128;
  int S = 1, best = sp.n_tiles < capacity[1] ? sp.n_tiles : capacity[1];
  for (int s2 = 2; s2 <= SK_MAX_SPLITS; ++s2) {
    const int slice = (P + s2 - 1) / s2;
    const bool fits = (P + (s2 - 1) * slice) * SK_BN * 4 <= ring_bytes;
    // (capacity + 1/15: the occupancy API under-counts clusters of 8 -- 32 tiles x 8 = 256 CTAs against an answer of 240
    // run in one wave, 14.1 us against 16.0 for S = 6 on 4096 x 14336; 48 x 6 = 288 against 264 do not: 14.2 against 9.7 for S = 5)
    // (only the two-CTAs-per-SM variants: at M > 64 the API's answer is exact -- 48 x 3 = 144 CTAs against 138 ran in two
    // waves, 19.5 us instead of 13.1)
    const int cap2 = capacity[s2] + ((COLS * GROUPS <= 64) ? capacity[s2] / 15 : 0);
    if (sp.n_tiles * s2 <= cap2 && sp.kb_total / s2 >= 2 && fits && sp.n_tiles * s2 > best) {
      S = s2;
      best = sp.n_tiles * s2;
    }
  }
  const int force = tune(TUNE_SK_SPLITS);
  if (force > 0 && force <= SK_MAX_SPLITS && sp.kb_total >= force) S = force;
  sp.splits = S;
  cfg.gridDim = dim3(sp.n_tiles, S);
  attr[1].val.clusterDim.y = S;
  cfg.numAttrs = S > 1 ? 2 : 1;
  if (sp.debug & 128) {
    for (int kb = 48; kb <= 116; kb += 4) {
      int nblk = -1;
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nblk, kern, kThreads, kb * 1024);
      fprintf(stderr, "[skinny] threads %d smem %d KB -> blocks/SM %d\n", kThreads, kb, nblk);
    }
    cudaFuncAttributes fa;
    cudaFuncGetAttributes(&fa, kern);
    fprintf(stderr, "[skinny] regs %d static smem %zu maxdyn %d carveout %d\n", fa.numRegs, fa.sharedSizeBytes, fa.maxDynamicSharedSizeBytes, fa.preferredShmemCarveout);
  }
  if (sp.debug & 64)
    fprintf(stderr, "[skinny] M=%d N=%d K=%d tiles=%d S=%d smem=%d stages=%d capacity %d/%d/%d/%d/%d/%d/%d/%d\n", sp.M, sp.N, sp.K,
            sp.n_tiles, S, smem, sp.stages, capacity[1], capacity[2], capacity[3], capacity[4], capacity[5], capacity[6], capacity[7], capacity[8]);
  B200_CUDA(cudaLaunchKernelEx(&cfg, kern, tw, tx, sp));
  if (sp.stamps) {
    static int printed = 0;
    cudaStreamSynchronize(st);
    const int n = sp.n_tiles * S;
    unsigned long long* h = (unsigned long long*)malloc((size_t)n * 64);
    cudaMemcpy(h, sp.stamps, (size_t)n * 64, cudaMemcpyDeviceToHost);
    unsigned long long t0 = ~0ULL;
    for (int c = 0; c < n; ++c) if (h[c * 8] < t0) t0 = h[c * 8];
    if (printed++ % 8 == 7) {
      fprintf(stderr, "[skinny stamps] M=%d N=%d K=%d S=%d (ns since first CTA start; min/avg/max over %d CTAs)\n", sp.M, sp.N, sp.K, S, n);
      for (int i = 0; i < 7; ++i) {
        unsigned long long mn = ~0ULL, mx = 0, sum = 0;
        for (int c = 0; c < n; ++c) { unsigned long long v = h[c * 8 + i] - t0; if (v < mn) mn = v; if (v > mx) mx = v; sum += v; }
        fprintf(stderr, "   stamp %d: %6llu %6llu %6llu\n", i, mn, sum / n, mx);
      }
    }
    free(h);
  }
  return launched("gemm_skinny_kernel launch");
}

// a [M,K] activations, w [N,K] weight (both K-major, 1-byte elements)
template <int MODE>
int launch_skinny(const void* a, int64_t lda, const void* w, int64_t ldw, SkinnyParams& sp, void* workspace,
                  size_t workspace_bytes, cudaStream_t st) {
  sp.stamps = nullptr;
  sp.mt_pad = skinny_mt_pad(sp.M);
  sp.n_tiles = (sp.N + SK_BN - 1) / SK_BN;
  sp.kb_total = ((MODE == 2 ? 2 * sp.K : sp.K) + SK_BK - 1) / SK_BK;  // 128 operand bytes per k-block
  const SkinnyPlan pl = skinny_plan(MODE, sp.mt_pad);
  sp.stages = pl.stages;
  sp.tab_cap = pl.tab_cap;
  {
    const int dbg = tune(TUNE_SK_DEBUG);
    sp.debug = dbg;
    if ((dbg & 256) && workspace && workspace_bytes >= (size_t)sp.n_tiles * 8 * 64) sp.stamps = (unsigned long long*)workspace;
  }
  B200_REQUIRE(pl.stages >= 2 && (sp.mt_pad < 128 ? sp.mt_pad : 128) * 512 <= pl.stages * (SK_BN * SK_BK + sp.mt_pad * SK_BK + (MODE == 2 ? 4096 : 0)), B200_ERR_UNSUPPORTED,
               "gemm_skinny: shared-memory plan does not fit (M=%d)", sp.M);
  CUtensorMap tw, tx;
  if (MODE == 2) {
    // w = qweight int32 [K, N/8] (row pitch ldw bytes): box of 64 k-rows x 64 bytes (16 words = 128 columns), dense;
    // a = half activations [M, K], addressed in bytes like the 1-byte modes (128 bytes = 64 k per block)
    if (int rc = make_tmap_2d(&tw, w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, sp.K, (uint64_t)sp.N / 2, ldw, 64, 64, CU_TENSOR_MAP_SWIZZLE_NONE)) return rc;
    if (int rc = make_tmap_2d(&tx, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, sp.M, (uint64_t)sp.K * 2, lda, sp.mt_pad, SK_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  } else {
    if (int rc = make_tmap_2d(&tw, w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, sp.N, sp.K, ldw, SK_BN, SK_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
    if (int rc = make_tmap_2d(&tx, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, sp.M, sp.K, lda, sp.mt_pad, SK_BK, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  }
  if constexpr (MODE == 0) {
    if (sp.rope_cs) {   // qkv projection with fused RoPE: its own instantiations
      if (sp.mt_pad <= 64) return launch_skinny_t<MODE, 32, 2, true>(tw, tx, sp, pl.smem, st);
      if (sp.mt_pad <= 128) return launch_skinny_t<MODE, 32, 4, true>(tw, tx, sp, pl.smem, st);
      return launch_skinny_t<MODE, 64, 4, true>(tw, tx, sp, pl.smem, st);
    }
  }
  if (sp.mt_pad <= 64) return launch_skinny_t<MODE, 32, 2>(tw, tx, sp, pl.smem, st);
  if (sp.mt_pad <= 128) return launch_skinny_t<MODE, 32, 4>(tw, tx, sp, pl.smem, st);
  return launch_skinny_t<MODE, 64, 4>(tw, tx, sp, pl.smem, st);
}

template int launch_skinny<0>(const void*, int64_t, const void*, int64_t, SkinnyParams&, void*, size_t, cudaStream_t);
template int launch_skinny<1>(const void*, int64_t, const void*, int64_t, SkinnyParams&, void*, size_t, cudaStream_t);
template int launch_skinny<2>(const void*, int64_t, const void*, int64_t, SkinnyParams&, void*, size_t, cudaStream_t);

}  // namespace b200
