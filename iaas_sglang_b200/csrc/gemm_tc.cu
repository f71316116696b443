// tcgen05 / TMEM / TMA GEMMs for the quantised linear layers of the decode path (sm_100a).
//
//   MODE 0  fp8_blockwise_scaled_mm  (sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:345-423)
//           out[m,n] = sum_kb sa[m,kb]*sb[kb,nb] * (A[m, kb*128:(kb+1)*128] . W[n, kb*128:(kb+1)*128])
//   MODE 1  int8_scaled_mm           (sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-747; no sm100 path in the
//           reference, :724-743)     out[m,n] = ((float)sum_k A[m,k]*W[n,k] * sa[m]) * sb[n] + bias[n]
//   MODE 2  dense bf16/f16 GEMM      out[m,n] = sum_k A[m,k]*W[n,k] + bias[n]   (second half of the AWQ
//           W4A16 linear, python/sglang/srt/layers/quantization/awq.py:199-200: dequantise, then matmul)
//
// Both operands are K-major (A [M,K] row-major, B = the [N,K] row-major weight), so one 128-byte TMA
// row is one 128-element K block: a stage is a 128x128B A tile + a 128x128B B tile in the canonical
// SWIZZLE_128B layout that tcgen05.mma reads directly.
//
// CTA = 8 warps: warp 0 TMA producer (one lane), warp 1 MMA issuer (one lane), warp 2 TMEM allocator,
// warps 4-7 epilogue (one TMEM lane = one output row per thread).
//   FP8: UMMA has no fp32 128-block scaling, so every K block is accumulated into its own TMEM
//   buffer (2 x 128 columns, double buffered) and the epilogue warps promote it into fp32 registers
//   with acc += tmem * (sa[m,kb]*sb[kb,nb]) while the tensor core works on the next block.
//   INT8: the int32 accumulator stays in TMEM for the whole K loop (exact), scales in the epilogue.
// Decode-shape GEMMs (M = 64..256) are weight-bandwidth bound and have fewer 128x128 tiles than SMs,
// so K is split across CTAs (grid.z); partial tiles meet in an fp32 workspace and the last CTA to
// arrive (atomic ticket) reduces and writes the tile.
#include "common.cuh"
#include "gemm_skinny.h"

namespace b200 {

// gemm_big.cu: persistent 256-wide-tile kernel for M > 256 (mode 0 fp8 block-scaled, 1 int8)
int launch_big(int mode, const void* a, int64_t lda, const void* w, int64_t ldw, const float* sa, int64_t sa_sm, int64_t sa_sk,
               const float* sb, int64_t sb_sk, int64_t sb_sn, const void* bias, void* out, int64_t ldo, int out_bf16, int M,
               int N, int K, void* workspace, size_t workspace_bytes, cudaStream_t st);
size_t big_workspace_bytes();

constexpr int BM = 128, BN = 128, BKB = 128;  // BKB: bytes (== elements) of K per stage
constexpr int kStages = 6;
constexpr int kStageBytes = (BM + BN) * BKB;  // 32 KB
constexpr int kGemmThreads = 256;
constexpr int kGemmSmem = kStages * kStageBytes + 1024 /*align slack*/ + 256 /*barriers*/;
constexpr int kCounterBytes = 4096;

struct GemmParams {
  const float* sa;
  int64_t sa_sm, sa_sk;
  const float* sb;
  int64_t sb_sk, sb_sn;
  const void* bias;
  void* out;
  int64_t ldo;
  int out_bf16;
  int M, N, K;
  int kb_total, kb_per_split, splits;
  uint32_t* ws_partial;  // [tile][split][128][128] 32-bit
  int* ws_counter;       // [tile]
};

__device__ __forceinline__ void epi_bar() { asm volatile("bar.sync 2, 128;" ::: "memory"); }

template <int MODE>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_tc_kernel(const __grid_constant__ CUtensorMap tma_a, const __grid_constant__ CUtensorMap tma_b,
               const __grid_constant__ GemmParams p) {
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + kStages * kStageBytes);
  uint64_t* empty = full + kStages;
  uint64_t* tfull = empty + kStages;
  uint64_t* tempty = tfull + 2;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(tempty + 2);
  int* last_flag = reinterpret_cast<int*>(tmem_slot + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n0 = blockIdx.x * BN, m0 = blockIdx.y * BM, split = blockIdx.z;
  const int kb0 = split * p.kb_per_split;
  const int nkb = min(p.kb_per_split, p.kb_total - kb0);
  constexpr uint32_t kTmemCols = MODE == 0 ? 256 : 128;

  if (warp == 0 && lane == 0) {
    tma_prefetch_desc(&tma_a);
    tma_prefetch_desc(&tma_b);
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&full[s], 1);
      mbar_init(&empty[s], 1);
    }
    for (int a = 0; a < 2; ++a) {
      mbar_init(&tfull[a], 1);
      mbar_init(&tempty[a], 128);
    }
    fence_mbar_init();
  }
  if (warp == 2) tmem_alloc(tmem_slot, kTmemCols);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ================================ TMA producer ================================
    if (lane == 0) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % kStages;
        const uint32_t ph = (i / kStages) & 1;
        mbar_wait(&empty[s], ph ^ 1);
        mbar_arrive_expect_tx(&full[s], kStageBytes);
        uint8_t* a_s = smem + s * kStageBytes;
        tma_load_2d(a_s, &tma_a, (kb0 + i) * BKB, m0, &full[s]);
        tma_load_2d(a_s + BM * BKB, &tma_b, (kb0 + i) * BKB, n0, &full[s]);
      }
    }
    __syncwarp();
  } else if (warp == 1) {
    // ================================ MMA issuer ================================
    if (lane == 0) {
      const uint32_t idesc = MODE == 0   ? umma_idesc(1 /*f32*/, 0 /*e4m3*/, 0, BM, BN)
                             : MODE == 1 ? umma_idesc(2 /*s32*/, 1 /*s8*/, 1, BM, BN)
                                         : umma_idesc(1 /*f32*/, p.out_bf16 ? 1 : 0, p.out_bf16 ? 1 : 0, BM, BN);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % kStages;
        const uint32_t ph = (i / kStages) & 1;
        const int a = MODE == 0 ? (i & 1) : 0;
        if (MODE == 0) mbar_wait(&tempty[a], ((i >> 1) & 1) ^ 1);
        mbar_wait(&full[s], ph);
        tc_fence_after();
        const uint32_t a_addr = smem_u32(smem + s * kStageBytes);
        const uint64_t adesc = umma_desc_k_sw128(a_addr);
        const uint64_t bdesc = umma_desc_k_sw128(a_addr + BM * BKB);
        const uint32_t d = tmem_base + a * BN;
#pragma unroll
        for (int k = 0; k < BKB / 32; ++k) {
          const uint32_t accum = MODE == 0 ? (k > 0) : (i > 0 || k > 0);
          if (MODE == 0) umma_f8f6f4(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
          else if (MODE == 1) umma_i8(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
          else umma_f16(d, adesc + 2 * k, bdesc + 2 * k, idesc, accum);
        }
        tc_commit(&empty[s]);
        if (MODE == 0 || i == nkb - 1) tc_commit(&tfull[a]);
      }
    }
    __syncwarp();
  } else if (warp >= 4) {
    // ================================ epilogue warps ================================
    const int w = warp - 4;              // == warp % 4: TMEM lanes [32w, 32w+32)
    const int row_in_tile = w * 32 + lane;
    const int row = m0 + row_in_tile;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(w * 32) << 16);
    const int tile_id = blockIdx.y * gridDim.x + blockIdx.x;
    float acc[BN];

    if (MODE == 0) {
#pragma unroll
      for (int j = 0; j < BN; ++j) acc[j] = 0.f;
      const int nb = n0 / 128;
      float sc_next = 0.f;
      {
        const float sa_v = row < p.M ? __ldg(p.sa + row * p.sa_sm + (int64_t)kb0 * p.sa_sk) : 0.f;
        sc_next = sa_v * __ldg(p.sb + (int64_t)kb0 * p.sb_sk + nb * p.sb_sn);
      }
      for (int i = 0; i < nkb; ++i) {
        const int a = i & 1;
        const float sc = sc_next;
        if (i + 1 < nkb) {
          const float sa_v = row < p.M ? __ldg(p.sa + row * p.sa_sm + (int64_t)(kb0 + i + 1) * p.sa_sk) : 0.f;
          sc_next = sa_v * __ldg(p.sb + (int64_t)(kb0 + i + 1) * p.sb_sk + nb * p.sb_sn);
        }
        mbar_wait(&tfull[a], (i >> 1) & 1);
        tc_fence_after();
#pragma unroll
        for (int c = 0; c < BN / 32; ++c) {
          uint32_t r[32];
          tmem_ld_32x32(lane_addr + a * BN + c * 32, r);
          tmem_ld_wait();
#pragma unroll
          for (int j = 0; j < 32; ++j) acc[c * 32 + j] = fmaf(__uint_as_float(r[j]), sc, acc[c * 32 + j]);
        }
        tc_fence_before();
        mbar_arrive(&tempty[a]);
      }
    } else {
      mbar_wait(&tfull[0], 0);
      tc_fence_after();
#pragma unroll
      for (int c = 0; c < BN / 32; ++c) {
        uint32_t r[32];
        tmem_ld_32x32(lane_addr + c * 32, r);
        tmem_ld_wait();
#pragma unroll
        for (int j = 0; j < 32; ++j) acc[c * 32 + j] = __uint_as_float(r[j]);  // raw int32 bits
      }
      tc_fence_before();
    }

    // ---- split-K: meet the other K slices of this tile in the workspace ----
    bool writer = true;
    if (p.splits > 1) {
      uint32_t* mine = p.ws_partial + ((size_t)tile_id * p.splits + split) * (BM * BN) + (size_t)row_in_tile * BN;
#pragma unroll
      for (int j = 0; j < BN; j += 4)
        *reinterpret_cast<uint4*>(mine + j) = make_uint4(__float_as_uint(acc[j]), __float_as_uint(acc[j + 1]),
                                                         __float_as_uint(acc[j + 2]), __float_as_uint(acc[j + 3]));
      __threadfence();
      epi_bar();
      if (threadIdx.x == 128) {
        const int old = atomicAdd(p.ws_counter + tile_id, 1);
        *last_flag = (old == p.splits - 1);
        if (old == p.splits - 1) p.ws_counter[tile_id] = 0;  // self-reset for the next launch
      }
      epi_bar();
      writer = *last_flag != 0;
      if (writer) {
        __threadfence();
        for (int s2 = 0; s2 < p.splits; ++s2) {
          if (s2 == split) continue;
          const uint32_t* other = p.ws_partial + ((size_t)tile_id * p.splits + s2) * (BM * BN) + (size_t)row_in_tile * BN;
#pragma unroll
          for (int j = 0; j < BN; j += 4) {
            const uint4 v = __ldcg(reinterpret_cast<const uint4*>(other + j));
            if (MODE != 1) {
              acc[j] += __uint_as_float(v.x);
              acc[j + 1] += __uint_as_float(v.y);
              acc[j + 2] += __uint_as_float(v.z);
              acc[j + 3] += __uint_as_float(v.w);
            } else {
              acc[j] = __int_as_float(__float_as_int(acc[j]) + (int)v.x);
              acc[j + 1] = __int_as_float(__float_as_int(acc[j + 1]) + (int)v.y);
              acc[j + 2] = __int_as_float(__float_as_int(acc[j + 2]) + (int)v.z);
              acc[j + 3] = __int_as_float(__float_as_int(acc[j + 3]) + (int)v.w);
            }
          }
        }
      }
    }

    if (writer && row < p.M) {
      float sa_row = 1.f;
      if (MODE == 1) sa_row = __ldg(p.sa + row);
      uint16_t* orow = reinterpret_cast<uint16_t*>(p.out) + (size_t)row * p.ldo + n0;
      const bool vec_ok = (n0 + BN <= p.N) && (p.ldo % 8 == 0) && ((reinterpret_cast<uintptr_t>(p.out) & 15) == 0);
#pragma unroll
      for (int j = 0; j < BN; j += 8) {
        float v[8];
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          if (MODE == 0) {
            v[e] = acc[j + e];
          } else if (MODE == 2) {
            float t = acc[j + e];
            if (p.bias) {
              const int n = n0 + j + e;
              const uint16_t bb = n < p.N ? __ldg(reinterpret_cast<const uint16_t*>(p.bias) + n) : (uint16_t)0;
              t += p.out_bf16 ? half_to_float<true>(bb) : half_to_float<false>(bb);
            }
            v[e] = t;
          } else {
            const int n = n0 + j + e;
            const float sbn = n < p.N ? __ldg(p.sb + n) : 0.f;
            float t = __fmul_rn(__fmul_rn((float)__float_as_int(acc[j + e]), sa_row), sbn);
            if (p.bias) {
              const uint16_t bb = n < p.N ? __ldg(reinterpret_cast<const uint16_t*>(p.bias) + n) : (uint16_t)0;
              t = __fadd_rn(t, p.out_bf16 ? half_to_float<true>(bb) : half_to_float<false>(bb));
            }
            v[e] = t;
          }
        }
        if (vec_ok) {
          uint4 pk;
          if (p.out_bf16) {
            pk.x = pack_half2<true>(v[0], v[1]); pk.y = pack_half2<true>(v[2], v[3]);
            pk.z = pack_half2<true>(v[4], v[5]); pk.w = pack_half2<true>(v[6], v[7]);
          } else {
            pk.x = pack_half2<false>(v[0], v[1]); pk.y = pack_half2<false>(v[2], v[3]);
            pk.z = pack_half2<false>(v[4], v[5]); pk.w = pack_half2<false>(v[6], v[7]);
          }
          *reinterpret_cast<uint4*>(orow + j) = pk;
        } else {
#pragma unroll
          for (int e = 0; e < 8; ++e)
            if (n0 + j + e < p.N) orow[j + e] = p.out_bf16 ? float_to_half<true>(v[e]) : float_to_half<false>(v[e]);
        }
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

struct SplitPlan {
  int tiles_m, tiles_n, kb_total, splits, kb_per_split;
};

static SplitPlan plan_split(int M, int N, int K) {
  SplitPlan sp;
  sp.tiles_m = (M + BM - 1) / BM;
  sp.tiles_n = (N + BN - 1) / BN;
  sp.kb_total = (K + BKB - 1) / BKB;  // K is in BYTES along the contraction dimension
  const int tiles = sp.tiles_m * sp.tiles_n;
  const int sms = device_sm_count() > 0 ? device_sm_count() : 148;
  int splits = 1;
  if (tiles < (sms * 4) / 5) {
    splits = sms / tiles;
    const int max_by_k = sp.kb_total / 2 > 0 ? sp.kb_total / 2 : 1;  // >= 2 K blocks per slice
    if (splits > max_by_k) splits = max_by_k;
    if (splits > 16) splits = 16;
    if (splits < 1) splits = 1;
  }
  sp.kb_per_split = (sp.kb_total + splits - 1) / splits;
  sp.splits = (sp.kb_total + sp.kb_per_split - 1) / sp.kb_per_split;
  return sp;
}

template <int MODE>
static int launch_gemm(const void* a, int64_t lda, const void* b, int64_t ldb, GemmParams& gp, void* workspace,
                       size_t workspace_bytes, cudaStream_t st) {
  // gp.K, lda, ldb are in bytes here (1-byte operands for MODE 0/1, 2-byte for MODE 2)
  const SplitPlan sp = plan_split(gp.M, gp.N, gp.K);
  gp.kb_total = sp.kb_total;
  gp.kb_per_split = sp.kb_per_split;
  gp.splits = sp.splits;
  gp.ws_partial = nullptr;
  gp.ws_counter = nullptr;
  if (sp.splits > 1) {
    const size_t need = kCounterBytes + (size_t)sp.tiles_m * sp.tiles_n * sp.splits * BM * BN * 4;
    B200_REQUIRE(workspace && workspace_bytes >= need && ((uintptr_t)workspace & 15) == 0, B200_ERR_ARG,
                 "gemm: workspace too small or misaligned (need %zu bytes, have %zu)", need, workspace_bytes);
    B200_REQUIRE((size_t)sp.tiles_m * sp.tiles_n * 4 <= kCounterBytes, B200_ERR_UNSUPPORTED, "gemm: too many split tiles");
    gp.ws_counter = reinterpret_cast<int*>(workspace);
    gp.ws_partial = reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(workspace) + kCounterBytes);
  }
  CUtensorMap ta, tb;
  if (int rc = make_tmap_2d(&ta, a, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, gp.M, gp.K, lda, BM, BKB, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  if (int rc = make_tmap_2d(&tb, b, CU_TENSOR_MAP_DATA_TYPE_UINT8, 1, gp.N, gp.K, ldb, BN, BKB, CU_TENSOR_MAP_SWIZZLE_128B)) return rc;
  auto kern = gemm_tc_kernel<MODE>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, kGemmSmem));
    configured = true;
  }
  dim3 grid(sp.tiles_n, sp.tiles_m, sp.splits);
  kern<<<grid, kGemmThreads, kGemmSmem, st>>>(ta, tb, gp);
  return launched("gemm_tc_kernel launch");
}

}  // namespace b200

using namespace b200;

extern "C" size_t b200_gemm_workspace_size(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0) return 16;
  // the large-M fp8 kernel splits K on few-tile shapes (gemm_big.cu): tickets + one fp32 partial tile per SM
  const size_t big = big_workspace_bytes();
  if (skinny_applicable(M)) return skinny_workspace_bytes(M, N, K) > big ? skinny_workspace_bytes(M, N, K) : big;
  const SplitPlan sp = plan_split(M, N, K);
  const size_t old_plan = sp.splits <= 1 ? 16 : kCounterBytes + (size_t)sp.tiles_m * sp.tiles_n * sp.splits * BM * BN * 4;
  return old_plan > big ? old_plan : big;
}

extern "C" int b200_fp8_blockwise_scaled_mm(const void* a, int64_t lda, const void* b, int64_t ldb, const float* sa,
                                            int64_t sa_sm, int64_t sa_sk, const float* sb, int64_t sb_sk, int64_t sb_sn,
                                            void* out, int64_t ldo, int out_dtype, int M, int N, int K, void* workspace,
                                            size_t workspace_bytes, void* stream) {
  B200_REQUIRE(a && b && sa && sb && out, B200_ERR_ARG, "fp8_blockwise_scaled_mm: null tensor");
  B200_REQUIRE(M >= 0 && N > 0 && K > 0, B200_ERR_ARG, "fp8_blockwise_scaled_mm: bad sizes");
  B200_REQUIRE(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_ARG, "fp8_blockwise_scaled_mm: out_dtype must be Half or BFloat16");
  // fp8_blockwise_gemm_kernel.cu:359-362,383: K and N*sizeof(out) must be multiples of 16 bytes
  B200_REQUIRE(K % 16 == 0, B200_ERR_ARG, "fp8_blockwise_scaled_mm: mat_a must be multiple of 16 bytes for memory alignment");
  B200_REQUIRE((N * 2) % 16 == 0, B200_ERR_ARG, "fp8_blockwise_scaled_mm: mat_b/out rows must be multiple of 16 bytes");
  B200_REQUIRE(lda % 16 == 0 && ldb % 16 == 0, B200_ERR_ARG, "fp8_blockwise_scaled_mm: operand pitches must be multiples of 16 bytes");
  if (M == 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED,
               "fp8_blockwise_scaled_mm: requires an sm_100 device (got sm_%d); no fallback", arch);
  if (skinny_applicable_fp8(M, N, K)) {
    SkinnyParams sp{};
    sp.sa = sa; sp.sa_sm = sa_sm; sp.sa_sk = sa_sk;
    sp.sb = sb; sp.sb_sk = sb_sk; sp.sb_sn = sb_sn;
    sp.bias = nullptr; sp.out = out; sp.ldo = ldo; sp.out_bf16 = out_dtype == B200_BF16;
    sp.M = M; sp.N = N; sp.K = K;
    return launch_skinny<0>(a, lda, b, ldb, sp, workspace, workspace_bytes, (cudaStream_t)stream);
  }
  return launch_big(0, a, lda, b, ldb, sa, sa_sm, sa_sk, sb, sb_sk, sb_sn, nullptr, out, ldo, out_dtype == B200_BF16, M, N, K,
                    workspace, workspace_bytes, (cudaStream_t)stream);
}

// qkv projection + neox RoPE in one launch (fused steps either side of the hot kernels): the first
// rope_cols output columns are q and k heads of head_dim columns each, rotated by token m's cos / sin row (b200_rope_cos_sin)
// exactly as b200_rope_neox would rotate the stored GEMM output (same helpers, same bits).  Served by the decode-shape kernel's write-out, where a
// head is one 128-column tile: B200_ERR_UNSUPPORTED for any other head_dim, or where the measured policy gives this
// (M, N, K) to the large-M kernel -- the caller then runs the two launches.
extern "C" int b200_fp8_blockwise_scaled_mm_rope(const void* a, int64_t lda, const void* b, int64_t ldb, const float* sa,
                                                 int64_t sa_sm, int64_t sa_sk, const float* sb, int64_t sb_sk, int64_t sb_sn,
                                                 void* out, int64_t ldo, int out_dtype, int M, int N, int K,
                                                 const float* cos_sin, int rope_cols, int head_dim,
                                                 void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(a && b && sa && sb && out && cos_sin && ((uintptr_t)cos_sin & 15) == 0, B200_ERR_ARG, "fp8_blockwise_scaled_mm_rope: null or misaligned tensor");
  B200_REQUIRE(M >= 0 && N > 0 && K > 0 && rope_cols >= 0 && rope_cols <= N, B200_ERR_ARG, "fp8_blockwise_scaled_mm_rope: bad sizes");
  B200_REQUIRE(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_ARG, "fp8_blockwise_scaled_mm_rope: out_dtype must be Half or BFloat16");
  B200_REQUIRE(K % 16 == 0 && (N * 2) % 16 == 0 && lda % 16 == 0 && ldb % 16 == 0, B200_ERR_ARG, "fp8_blockwise_scaled_mm_rope: 16-byte alignment of K, N and the pitches");
  B200_REQUIRE(head_dim == 128 && rope_cols % 128 == 0, B200_ERR_UNSUPPORTED,
               "fp8_blockwise_scaled_mm_rope: the fused epilogue needs head_dim 128 (got %d); run the GEMM and b200_rope_neox", head_dim);
  if (M == 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED,
               "fp8_blockwise_scaled_mm_rope: requires an sm_100 device (got sm_%d); no fallback", arch);
  B200_REQUIRE(skinny_applicable_fp8(M, N, K), B200_ERR_UNSUPPORTED,
               "fp8_blockwise_scaled_mm_rope: M=%d N=%d K=%d belongs to the large-M kernel; run the GEMM and b200_rope_neox", M, N, K);
  SkinnyParams sp{};
  sp.sa = sa; sp.sa_sm = sa_sm; sp.sa_sk = sa_sk;
  sp.sb = sb; sp.sb_sk = sb_sk; sp.sb_sn = sb_sn;
  sp.bias = nullptr; sp.out = out; sp.ldo = ldo; sp.out_bf16 = out_dtype == B200_BF16;
  sp.M = M; sp.N = N; sp.K = K;
  sp.rope_cs = reinterpret_cast<const float2*>(cos_sin); sp.rope_cols = rope_cols;
  return launch_skinny<0>(a, lda, b, ldb, sp, workspace, workspace_bytes, (cudaStream_t)stream);
}

namespace b200 {
// dense half-precision GEMM out = A[M,K] . W[N,K]^T (+bias); used by the AWQ W4A16 linear
int dense_half_gemm(const void* a, int64_t lda, const void* w, int64_t ldw, const void* bias, void* out, int64_t ldo,
                    int dtype, int M, int N, int K, void* workspace, size_t workspace_bytes, cudaStream_t st) {
  B200_REQUIRE(K % 8 == 0 && lda % 8 == 0 && ldw % 8 == 0, B200_ERR_ARG, "dense gemm: K and pitches must be multiples of 8 elements");
  GemmParams gp{};
  gp.bias = bias; gp.out = out; gp.ldo = ldo; gp.out_bf16 = dtype == B200_BF16;
  gp.M = M; gp.N = N; gp.K = K * 2;
  return launch_gemm<2>(a, lda * 2, w, ldw * 2, gp, workspace, workspace_bytes, st);
}
size_t dense_half_gemm_workspace(int M, int N, int K) {
  const SplitPlan sp = plan_split(M, N, K * 2);
  if (sp.splits <= 1) return 16;
  return kCounterBytes + (size_t)sp.tiles_m * sp.tiles_n * sp.splits * BM * BN * 4;
}
}  // namespace b200

extern "C" int b200_int8_scaled_mm(const void* a, int64_t lda, const void* b, int64_t ldb, const float* sa, const float* sb,
                                   const void* bias, void* out, int64_t ldo, int out_dtype, int M, int N, int K,
                                   void* workspace, size_t workspace_bytes, void* stream) {
  B200_REQUIRE(a && b && sa && sb && out, B200_ERR_ARG, "int8_scaled_mm: null tensor");
  B200_REQUIRE(M >= 0 && N > 0 && K > 0, B200_ERR_ARG, "int8_scaled_mm: bad sizes");
  B200_REQUIRE(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_ARG, "int8_scaled_mm: out_dtype must be Half or BFloat16");
  // int8_gemm_kernel.cu:677-679
  B200_REQUIRE(K % 16 == 0, B200_ERR_ARG, "int8_scaled_mm: mat_a.size(1) must be multiple of 16 for memory alignment");
  B200_REQUIRE(N % 8 == 0, B200_ERR_ARG, "int8_scaled_mm: mat_b.size(1) must be multiple of 8 for memory alignment");
  B200_REQUIRE(lda % 16 == 0 && ldb % 16 == 0, B200_ERR_ARG, "int8_scaled_mm: operand pitches must be multiples of 16 bytes");
  if (M == 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "int8_scaled_mm: requires an sm_100 device (got sm_%d); no fallback", arch);
  if (skinny_applicable_int8(M, N, K)) {
    SkinnyParams sp{};
    sp.sa = sa; sp.sb = sb; sp.bias = bias; sp.out = out; sp.ldo = ldo; sp.out_bf16 = out_dtype == B200_BF16;
    sp.M = M; sp.N = N; sp.K = K;
    return launch_skinny<1>(a, lda, b, ldb, sp, workspace, workspace_bytes, (cudaStream_t)stream);
  }
  return launch_big(1, a, lda, b, ldb, sa, 0, 0, sb, 0, 0, bias, out, ldo, out_dtype == B200_BF16, M, N, K, nullptr, 0,
                    (cudaStream_t)stream);
}
