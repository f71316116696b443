// Paged-KV decode attention for sm_100a.
//
// Replaces (see include/b200_sgl.h): decode_attention_cpu (sgl-kernel/csrc/cpu/decode.cpp:1375-1575),
// decode_attention_fwd (python/sglang/srt/layers/attention/triton_ops/decode_attention.py:677-728) and the
// fused KVCache.set_kv_buffer (python/sglang/srt/mem_cache/memory_pool.py:418-456).
//
// Design (DESIGN.md "decode attention"): the op is HBM-bound (G..2G flop per KV byte), so the kernel is
// built around keeping ~200 KB of scattered KV rows in flight per SM:
//   * grid = (kv split, kv-head group x q-block, request); one CTA per SM (smem-limited);
//   * a producer warp reads the page table IN PLACE (req_to_token row of the request; no flattened
//     kv_indices copy) and issues one TMA bulk copy (cp.async.bulk, UBLKCP) per token row per K/V into
//     a 3..6-stage shared-memory ring; completion is tracked by mbarrier transaction counts;
//   * 8 consumer warps read the rows with ldmatrix (16-byte vectors), run QK^T and PV on the warp
//     tensor-core path (mma.sync m16n8k16, fp32 accumulate) with the GQA group padded to the 16-row
//     MMA tile, and keep an online softmax in registers (quad shuffles for the row max / sum);
//   * per-warp states are merged through shared memory, per-split states through a small fp32
//     workspace and an LSE-merge kernel (decode.cpp:812-860 / decode_attention.py:492-552).
// A generic SIMT kernel covers every other head dimension (incl. MLA 576/512) so that the op has no
// unsupported shapes and no CPU fallback.
#include <math.h>
#include <stdio.h>

#include <pthread.h>

#include "common.cuh"

namespace b200 {

constexpr int kConsumerWarps = 8;
// UBLKCP takes its addresses from uniform registers, so a warp issues its lanes' bulk copies one after the
// other (ncu: a single producer warp was the bottleneck, consumers idle on the full barrier 44 % of the
// samples).  Four producer warps split every tile's rows and prefetch the page-table entries one tile ahead.
constexpr int kProducerWarps = 4;
constexpr int kAttnThreads = (kConsumerWarps + kProducerWarps) * 32;
constexpr int kTokPerWarp = 16;
constexpr float kLog2e = 1.4426950408889634f;

// optional epilogue of the split merge: per_token_group_quant_fp8(out.view(B, Hq * Dv), 128) for the o_proj linear
// (fp8_utils.py:202-204), bit-identical to the separate quantiser kernel; q == nullptr: off
struct MergeQuant {
  uint8_t* q;
  int64_t q_sm;
  float* s;
  int64_t s_sm, s_sg;
  float eps, fmin_, fmax_;
};

struct DecodeKParams {
  const uint16_t* q;
  int64_t q_sb, q_sh;
  const uint16_t* k_new;
  int64_t kn_sb, kn_sh;
  const uint16_t* v_new;
  int64_t vn_sb, vn_sh;
  uint16_t* k_buf;
  int64_t kb_st, kb_sh;
  uint16_t* v_buf;
  int64_t vb_st, vb_sh;
  uint16_t* out;
  float* part_o;    // generic path: [B, Hq, S, Dv];  tuned path: [grid, 2, HPC*16, Dv]
  float* part_lse;  // generic path: [B, Hq, S];      tuned path: [grid, 2, HPC*16]   (log2 domain)
  const void* r2t;
  int64_t r2t_stride;
  const void* rpi;       // [B] int64 or int32 (meta32 bit 1); nullptr = identity (row b of the table is request b)
  const void* seq_lens;  // [B] int64 or int32 (meta32 bit 0)
  const void* loc;       // [B] int64 or int32 (meta32 bit 2)
  int meta32;
  int page_shift;        // table entries are page ids of 1 << page_shift tokens (0 = token slots, req_to_token)
  const uint16_t* q_pe;  // MLA: rope part of q [B, Hq, 64] when q holds only the 512 latent columns (cutlass_mla_decode)
  int64_t qpe_sb, qpe_sh;
  const uint16_t* k_pe_new;  // MLA: rope part of the new latent row when k_new holds only the 512 latent columns
  int64_t kpe_sb;
  int idx64;
  int B, Hq, Hkv, G, D, Dv;
  int num_splits;
  int fused_new;  // substitute k_new/v_new for the slot out_cache_loc[b] and write it back in-kernel
  int is_mla;
  float sm_scale, logit_cap;
  MergeQuant mq;
  // FP8 KV cache (memory_pool.py:116-121,431-440; layers/quantization/kv_cache.py): the pool holds e4m3 codes,
  // K = code * k_scale, V = code * v_scale (tuned GQA path, D = 128; strides in ELEMENTS of the pool dtype)
  int kv_fp8;
  float k_scale, v_scale;
  // cascade / shared-prefix decode (tuned GQA path): the kernel attends to positions [*kv_start, seq_len) only, every
  // (request, head) state goes through the merge kernel, which adds the state over the shared prefix [0, *kv_start)
  // computed once for the whole batch (extend kernel, prefix_only)
  const int32_t* kv_start;   // device scalar or nullptr
  float* casc_o;             // [B, Hq, Dv] fp32: normalised suffix output of the units one CTA saw whole
  float* casc_lse;           // [B, Hq] fp32, log2 domain
  const float* pre_o;        // [B, Hq, Dv] fp32 prefix state
  const float* pre_lse;      // [B, Hq] natural-log LSE of the prefix state (-inf: empty)
};

// seq_lens / req_pool_indices / out_cache_loc arrive as int64 (schedule_batch.py) or int32 (cuda_graph_runner.py:273-276)
__device__ __forceinline__ int64_t meta_at(const void* p, int64_t i, bool is32) {
  return is32 ? (int64_t) reinterpret_cast<const int32_t*>(p)[i] : reinterpret_cast<const int64_t*>(p)[i];
}
__device__ __forceinline__ int64_t seq_len_at(const DecodeKParams& p, int b) { return meta_at(p.seq_lens, b, p.meta32 & 1); }
__device__ __forceinline__ int64_t req_at(const DecodeKParams& p, int b) { return p.rpi ? meta_at(p.rpi, b, p.meta32 & 2) : (int64_t)b; }
__device__ __forceinline__ int64_t loc_at(const DecodeKParams& p, int b) { return meta_at(p.loc, b, p.meta32 & 4); }

constexpr int kMaxSmem = 232448;  // 227 KB: the sm_100 per-block opt-in limit
constexpr int kMaxBatchTable = 1024;
constexpr int kMaxPartials = 256;  // >= launch grid of the tuned kernel (one CTA per SM)

template <int D, int HPC, int EB = 2>   // EB = bytes per KV element: 2 (bf16 / f16 pool) or 1 (e4m3 pool, FP8 KV cache)
struct AttnCfg {
  static constexpr int WT = kConsumerWarps / HPC;
  static constexpr int TILE_T = WT * kTokPerWarp;
  static constexpr int ROW_BYTES = HPC * D * EB;
  static constexpr int ROW_PITCH = ROW_BYTES + 16;
  // rows >= 1 KB go through the TMA bulk-copy engine (one UBLKCP per row); shorter rows (few KV heads per
  // GPU under TP) are limited by the per-copy cost of that engine (measured: 256 B rows -> 47 % of HBM), so
  // they are staged with 16-byte cp.async (LDGSTS) instead, tracked by the same mbarriers.
  static constexpr bool BULK = ROW_BYTES >= 1024;
  static constexpr int FULL_COUNT = BULK ? kProducerWarps : kProducerWarps * 32;
  static constexpr int KV_BYTES = TILE_T * ROW_PITCH;
  // a finished stage doubles as the scratch for the cross-warp merge at the end of a segment
  static constexpr int SCRATCH_BYTES = kConsumerWarps * 16 * (D + 2) * 4;
  static constexpr int STAGE_RAW = 2 * KV_BYTES > SCRATCH_BYTES ? 2 * KV_BYTES : SCRATCH_BYTES;
  static constexpr int STAGE_BYTES = (STAGE_RAW + 127) / 128 * 128;
  // per-CTA copy of the batch: tiles prefix, seq_len, request row, new-token slot
  static constexpr int TABLE_BYTES = kMaxBatchTable * 8 + (kMaxBatchTable + 4) * 4 + 2 * kMaxBatchTable * 4;
  static constexpr int STAGES_FIT = (kMaxSmem - TABLE_BYTES - 256) / STAGE_BYTES;
  static constexpr int STAGES = STAGES_FIT > 8 ? 8 : STAGES_FIT;
  static constexpr int RING_BYTES = STAGES * STAGE_BYTES;
  static constexpr int SMEM_BYTES = RING_BYTES + TABLE_BYTES + 2 * STAGES * 8 + 16;
  static_assert(STAGES >= 2, "ring too small");
};

__device__ __forceinline__ void consumer_bar() { asm volatile("bar.sync 1, %0;" ::"n"(kConsumerWarps * 32) : "memory"); }

// Work partition of the tuned kernel (shared with the merge kernel).  A "unit" is one (request b, kv-head
// group x q-block g) pair and costs tiles(b) = ceil(seq_len[b] / TILE_T) tiles; units are laid out b-major.
// CTA c of `grid` owns the flattened tile range [c*W/grid, (c+1)*W/grid), W = NG * sum_b tiles(b): every SM
// streams the same number of KV bytes whatever the length distribution, and there is one pipeline ramp-up per
// SM per launch instead of one per (request, split) CTA.
// (32-bit division whenever W * grid fits: a 64-bit division is ~100 dependent instructions, and the merge kernel's warps
// each do four of them before their first load)
__device__ __forceinline__ int64_t part_start(int64_t c, int64_t W, int64_t grid) {
  if ((((uint64_t)W * (uint64_t)grid) >> 32) == 0) return (int64_t)(((uint32_t)c * (uint32_t)W) / (uint32_t)grid);
  return c * W / grid;
}
__device__ __forceinline__ int64_t part_owner(int64_t t, int64_t W, int64_t grid) {
  if ((((uint64_t)W * (uint64_t)grid) >> 32) == 0) return (int64_t)((((uint32_t)t + 1u) * (uint32_t)grid - 1u) / (uint32_t)W);
  return ((t + 1) * grid - 1) / W;
}

struct SegCursor {
  // walks the (b, g, tile) space from flattened tile t; tables live in shared memory
  const int32_t* P;
  int B, NG;
  int b, g, tile, tiles_b;
  int64_t t, t_end;
  __device__ __forceinline__ void init(const int32_t* P_, int B_, int NG_, int64_t t0, int64_t t1) {
    P = P_; B = B_; NG = NG_; t = t0; t_end = t1;
    int lo = 0, hi = B - 1;
    while (lo < hi) {
      const int mid = (lo + hi) >> 1;
      if ((int64_t)NG * P[mid + 1] > t0) hi = mid; else lo = mid + 1;
    }
    b = lo;
    tiles_b = P[b + 1] - P[b];
    const int64_t rem = t0 - (int64_t)NG * P[b];
    g = tiles_b > 0 ? (int)(rem / tiles_b) : 0;
    tile = tiles_b > 0 ? (int)(rem - (int64_t)g * tiles_b) : 0;
  }
  __device__ __forceinline__ bool done() const { return t >= t_end; }
  // tiles of the current segment
  __device__ __forceinline__ int seg_len() const {
    const int64_t left = t_end - t;
    const int in_unit = tiles_b - tile;
    return left < in_unit ? (int)left : in_unit;
  }
  __device__ __forceinline__ void advance(int n) {
    t += n;
    tile += n;
    if (tile == tiles_b) {
      tile = 0;
      if (++g == NG) {
        g = 0;
        do { ++b; } while (b < B && P[b + 1] == P[b]);
        tiles_b = b < B ? P[b + 1] - P[b] : 0;
      }
    }
  }
};

// two e4m3 codes (low / high 16 bits of `w`) -> packed f16x2, exact
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint32_t w, bool hi) {
  uint32_t r;
  const uint16_t h = (uint16_t)(hi ? (w >> 16) : (w & 0xffffu));
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(r) : "h"(h));
  return r;
}
// bf16 / f16 pair -> f16x2 (Q for the FP8-KV path: the MMAs run in f16)
template <bool BF16>
__device__ __forceinline__ uint32_t to_f16x2(uint32_t w) {
  if constexpr (!BF16) return w;
  const float2 f = unpack_half2<true>(w);
  return pack_half2<false>(f.x, f.y);
}

template <int D, int HPC, bool BF16, bool KV8 = false>
__global__ void __launch_bounds__(kAttnThreads, 1) decode_attn_mma_kernel(const __grid_constant__ DecodeKParams p) {
  constexpr int EB = KV8 ? 1 : 2;
  using C = AttnCfg<D, HPC, EB>;
  constexpr bool MB = KV8 ? false : BF16;   // MMA operand type: e4m3 converts exactly to f16, so FP8 KV runs f16 MMAs
  constexpr int Dv = D;
  extern __shared__ __align__(128) uint8_t smem[];
  int64_t* s_loc = reinterpret_cast<int64_t*>(smem + C::RING_BYTES);
  int32_t* s_P = reinterpret_cast<int32_t*>(s_loc + kMaxBatchTable);  // [B + 1] exclusive prefix of tiles(b)
  int32_t* s_len = s_P + kMaxBatchTable + 4;
  int32_t* s_rpi = s_len + kMaxBatchTable;
  uint64_t* full = reinterpret_cast<uint64_t*>(s_rpi + kMaxBatchTable);
  uint64_t* empty = full + C::STAGES;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_qblk = (p.G + 15) >> 4;
  const int NG = (p.Hkv / HPC) * n_qblk;

  // ---- batch tables + barriers ----
  griddep_wait();    // programmatic dependent launch: q / k_new / the batch tensors belong to the stream order
  griddep_launch();  // the merge kernel may be scheduled as soon as SM resources free up
  const int64_t kv_start = p.kv_start ? (int64_t)*p.kv_start : 0;
  for (int b = threadIdx.x; b < p.B; b += kAttnThreads) {
    int64_t len = seq_len_at(p, b) - kv_start;
    if (len < 0) len = 0;
    s_len[b] = (int32_t)len;
    s_rpi[b] = (int32_t)req_at(p, b);
    s_loc[b] = p.fused_new ? loc_at(p, b) : -1;
    s_P[b + 1] = (int32_t)((len + C::TILE_T - 1) / C::TILE_T);
  }
  if (threadIdx.x == 0) {
    s_P[0] = 0;
    for (int s = 0; s < C::STAGES; ++s) {
      mbar_init(&full[s], C::FULL_COUNT);
      mbar_init(&empty[s], kConsumerWarps);
    }
    fence_mbar_init();
  }
  __syncthreads();
  if (warp == 0) {  // inclusive scan of tiles(b) into s_P[1..B]
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
    constexpr int RW = C::TILE_T / kProducerWarps;  // token rows per producer warp per tile
    uint32_t gt = 0;                                // tiles issued so far (ring position)
    for (; !cur.done();) {
      const int n = cur.seg_len();
      const int b = cur.b, hg = cur.g / n_qblk, qblk = cur.g - hg * n_qblk;
      const int kvh0 = hg * HPC;
      const int64_t len = s_len[b], req = s_rpi[b], loc_b = s_loc[b];
      const int64_t kv0 = (int64_t)cur.tile * C::TILE_T;
      const int64_t kv_end = (int64_t)(cur.tile + n) * C::TILE_T;
      const int64_t kv1 = kv_end < len ? kv_end : len;
      const uint16_t* knew_row = p.k_new + b * p.kn_sb + kvh0 * p.kn_sh;
      const uint16_t* vnew_row = p.v_new + b * p.vn_sb + kvh0 * p.vn_sh;
      if (p.fused_new && pw == 0 && qblk == 0 && cur.tile + n == cur.tiles_b) {
        // fused KVCache.set_kv_buffer for this unit's heads (memory_pool.py:418-456); the attention itself
        // reads the new row straight from k_new / v_new (substituted below), so no ordering is needed
        const uint4* ks = reinterpret_cast<const uint4*>(knew_row);
        const uint4* vs = reinterpret_cast<const uint4*>(vnew_row);
        uint4* kd = reinterpret_cast<uint4*>(p.k_buf + loc_b * p.kb_st + kvh0 * p.kb_sh);
        uint4* vd = reinterpret_cast<uint4*>(p.v_buf + loc_b * p.vb_st + kvh0 * p.vb_sh);
        for (int i = lane; i < C::ROW_BYTES / 16; i += 32) {
          kd[i] = ks[i];
          vd[i] = vs[i];
        }
      }
      const uint8_t* r2t_row = reinterpret_cast<const uint8_t*>(p.r2t) + (req * p.r2t_stride + kv_start) * (p.idx64 ? 8 : 4);
      if constexpr (C::BULK) {
        constexpr int ITEMS = 2 * RW;  // one K and one V bulk copy per row
        constexpr int R = (ITEMS + 31) / 32;
        // raw page-table words are kept as loaded (no arithmetic on them) so that the prefetch of the next
        // tile's entries does not stall on the load before this tile's copies are issued
        int32_t lo[R], hi[R], nlo[R], nhi[R];
        auto load_idx = [&](int it, int32_t(&dlo)[R], int32_t(&dhi)[R]) {
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const int item = lane + 32 * r;
            int64_t pos = kv0 + (int64_t)it * C::TILE_T + pw * RW + (item >> 1);
            if (pos > kv1 - 1) pos = kv1 - 1;  // tail: duplicate the last valid row (masked by the consumers)
            dlo[r] = 0; dhi[r] = 0;
            if (item < ITEMS) {
              if (p.idx64) {
                const int2 w = *reinterpret_cast<const int2*>(r2t_row + pos * 8);
                dlo[r] = w.x; dhi[r] = w.y;
              } else {
                dlo[r] = *reinterpret_cast<const int32_t*>(r2t_row + pos * 4);
              }
            }
          }
        };
        load_idx(0, lo, hi);
        for (int it = 0; it < n; ++it, ++gt) {
          const int s = gt % C::STAGES;
          const uint32_t ph = (gt / C::STAGES) & 1;
          if (it + 1 < n) load_idx(it + 1, nlo, nhi);
          mbar_wait(&empty[s], ph ^ 1);
          if (lane == 0) mbar_arrive_expect_tx(&full[s], ITEMS * C::ROW_BYTES);
          __syncwarp();
          uint8_t* kdst = smem + s * C::STAGE_BYTES;
#pragma unroll
          for (int r = 0; r < R; ++r) {
            const int item = lane + 32 * r;
            if (item < ITEMS) {
              const int64_t idx = p.idx64 ? (((int64_t)hi[r] << 32) | (uint32_t)lo[r]) : (int64_t)lo[r];
              const int row = pw * RW + (item >> 1);
              const bool is_v = item & 1;
              const void* src = is_v ? (const void*)(reinterpret_cast<const uint8_t*>(p.v_buf) + (idx * p.vb_st + kvh0 * p.vb_sh) * EB)
                                     : (const void*)(reinterpret_cast<const uint8_t*>(p.k_buf) + (idx * p.kb_st + kvh0 * p.kb_sh) * EB);
              if (!KV8 && idx == loc_b) src = is_v ? vnew_row : knew_row;
              bulk_g2s(kdst + (is_v ? C::KV_BYTES : 0) + row * C::ROW_PITCH, src, C::ROW_BYTES, &full[s]);
            }
          }
#pragma unroll
          for (int r = 0; r < R; ++r) {
            lo[r] = nlo[r];
            hi[r] = nhi[r];
          }
        }
      } else {
        // 16-byte cp.async: CPR lanes per row, 32 / CPR rows per warp instruction
        constexpr int CPR = C::ROW_BYTES / 16;
        static_assert(CPR <= 32 && 32 % CPR == 0 && RW <= 32, "short-row producer layout");
        constexpr int RPI = 32 / CPR;
        const int sub = lane / CPR, chunk = lane % CPR;
        int32_t lo = 0, hi = 0, nlo = 0, nhi = 0;
        auto load_idx = [&](int it, int32_t& dlo, int32_t& dhi) {
          int64_t pos = kv0 + (int64_t)it * C::TILE_T + pw * RW + lane;
          if (pos > kv1 - 1) pos = kv1 - 1;
          if (lane < RW) {
            if (p.idx64) {
              const int2 w = *reinterpret_cast<const int2*>(r2t_row + pos * 8);
              dlo = w.x; dhi = w.y;
            } else {
              dlo = *reinterpret_cast<const int32_t*>(r2t_row + pos * 4);
            }
          }
        };
        load_idx(0, lo, hi);
        // The issue rate of these warps is what bounds the short-row path (a longer instruction sequence per row showed
        // up one-to-one in the kernel time), so the loop is kept to a shuffle, two 32 x 32 -> 64-bit multiply-adds and
        // the two copies per pair of rows: int32 page-table entries (the usual req_to_token dtype) with byte strides
        // that fit 32 bits; the new token's substitution is only looked for in the tile that can hold it.
        const bool fast32 = !p.idx64 && p.kb_st * EB < (int64_t)1 << 31 && p.vb_st * EB < (int64_t)1 << 31;
        const uint32_t kst_b = (uint32_t)(p.kb_st * EB), vst_b = (uint32_t)(p.vb_st * EB);
        const char* kbase = reinterpret_cast<const char*>(p.k_buf) + kvh0 * p.kb_sh * EB + chunk * 16;
        const char* vbase = reinterpret_cast<const char*>(p.v_buf) + kvh0 * p.vb_sh * EB + chunk * 16;
        const bool unit_tail = p.fused_new && cur.tile + n == cur.tiles_b;   // the segment ends at the request's last token
        for (int it = 0; it < n; ++it, ++gt) {
          const int s = gt % C::STAGES;
          const uint32_t ph = (gt / C::STAGES) & 1;
          if (it + 1 < n) load_idx(it + 1, nlo, nhi);
          mbar_wait(&empty[s], ph ^ 1);
          const uint32_t kdst = smem_u32(smem + s * C::STAGE_BYTES) + (pw * RW + sub) * C::ROW_PITCH + chunk * 16;
          if (fast32 && !(unit_tail && it == n - 1)) {
#pragma unroll
            for (int r = 0; r < RW; r += RPI) {
              const uint32_t ri = (uint32_t)__shfl_sync(0xffffffffu, lo, r + sub);
              const char* ks = kbase + (uint64_t)ri * kst_b;
              const char* vs = vbase + (uint64_t)ri * vst_b;
              if constexpr (C::ROW_BYTES >= 256) {
                // L2::256B: the first chunk of a row pulls its whole aligned 256 bytes into L2 (one DRAM request per row)
                cp_async_16_l2_256(kdst + r * C::ROW_PITCH, ks);
                cp_async_16_l2_256(kdst + C::KV_BYTES + r * C::ROW_PITCH, vs);
              } else {
                cp_async_16(kdst + r * C::ROW_PITCH, ks);
                cp_async_16(kdst + C::KV_BYTES + r * C::ROW_PITCH, vs);
              }
            }
          } else {
            const int64_t idx = p.idx64 ? (((int64_t)hi << 32) | (uint32_t)lo) : (int64_t)lo;
#pragma unroll 4
            for (int r = 0; r < RW; r += RPI) {
              const int64_t ri = __shfl_sync(0xffffffffu, idx, r + sub);
              const char* ks = reinterpret_cast<const char*>(p.k_buf) + (ri * p.kb_st + kvh0 * p.kb_sh) * EB;
              const char* vs = reinterpret_cast<const char*>(p.v_buf) + (ri * p.vb_st + kvh0 * p.vb_sh) * EB;
              if (!KV8 && ri == loc_b) {
                ks = reinterpret_cast<const char*>(knew_row);
                vs = reinterpret_cast<const char*>(vnew_row);
              }
              cp_async_16(kdst + r * C::ROW_PITCH, ks + chunk * 16);
              cp_async_16(kdst + C::KV_BYTES + r * C::ROW_PITCH, vs + chunk * 16);
            }
          }
          cp_async_mbar_arrive_noinc(&full[s]);
          lo = nlo;
          hi = nhi;
        }
      }
      cur.advance(n);
    }
    return;
  }

  // ================================ consumer warps ================================
  const int wh = warp / C::WT, wt = warp - wh * C::WT;
  const int g = lane >> 2, t = lane & 3;
  const bool has_cap = p.logit_cap > 0.f;
  const float sm_scale_k = KV8 ? p.sm_scale * p.k_scale : p.sm_scale;   // logits = q . (code * k_scale)
  const float scale2 = sm_scale_k * kLog2e;
  const int lm = lane >> 3, lr = lane & 7;
  // ldmatrix lane addressing: K non-transposed, V transposed
  const uint32_t k_lane_off = (uint32_t)(((lm >> 1) * 8 + lr) * C::ROW_PITCH + (lm & 1) * 16);
  const uint32_t v_lane_off = (uint32_t)(((lm & 1) * 8 + lr) * C::ROW_PITCH + (lm >> 1) * 16);
  const uint32_t smem_base = smem_u32(smem);
  uint32_t gt = 0;

  for (; !cur.done();) {
    const int n = cur.seg_len();
    const int b = cur.b, hg = cur.g / n_qblk, qblk = cur.g - hg * n_qblk;
    const int kvh0 = hg * HPC, kvh = kvh0 + wh;
    const int row0 = qblk * 16;  // first row of the GQA group handled here
    const int rows_valid = min(16, p.G - row0);
    const int qh0 = kvh * p.G + row0;
    const int64_t len = s_len[b];
    const int64_t kv0 = (int64_t)cur.tile * C::TILE_T;
    const int64_t kv_end = (int64_t)(cur.tile + n) * C::TILE_T;
    const int64_t kv1 = kv_end < len ? kv_end : len;
    const bool whole_unit = (cur.tile == 0 && n == cur.tiles_b);
    const bool first_seg = (cur.t == t_begin);

    uint32_t qf[D / 16][4];
    {
      const uint16_t* q0 = p.q + b * p.q_sb + (int64_t)(qh0 + g) * p.q_sh + 2 * t;
      const uint16_t* q1 = q0 + 8 * p.q_sh;
      const bool v0 = g < rows_valid, v1 = g + 8 < rows_valid;
#pragma unroll
      for (int ks = 0; ks < D / 16; ++ks) {
        if constexpr (KV8) {
          // One ldmatrix of e4m3 rows hands thread (g, t) the four codes k = 16 ks + 4 t .. + 3 of a key, so the MMA's
          // k-slots (2t, 2t+1 | 2t+8, 2t+9) carry k = (4t, 4t+1 | 4t+2, 4t+3): Q is loaded in the same permutation.
          const uint16_t* a0 = p.q + b * p.q_sb + (int64_t)(qh0 + g) * p.q_sh + ks * 16 + 4 * t;
          const uint16_t* a1 = a0 + 8 * p.q_sh;
          qf[ks][0] = v0 ? to_f16x2<BF16>(*reinterpret_cast<const uint32_t*>(a0)) : 0u;
          qf[ks][1] = v1 ? to_f16x2<BF16>(*reinterpret_cast<const uint32_t*>(a1)) : 0u;
          qf[ks][2] = v0 ? to_f16x2<BF16>(*reinterpret_cast<const uint32_t*>(a0 + 2)) : 0u;
          qf[ks][3] = v1 ? to_f16x2<BF16>(*reinterpret_cast<const uint32_t*>(a1 + 2)) : 0u;
        } else {
          qf[ks][0] = v0 ? *reinterpret_cast<const uint32_t*>(q0 + ks * 16) : 0u;
          qf[ks][1] = v1 ? *reinterpret_cast<const uint32_t*>(q1 + ks * 16) : 0u;
          qf[ks][2] = v0 ? *reinterpret_cast<const uint32_t*>(q0 + ks * 16 + 8) : 0u;
          qf[ks][3] = v1 ? *reinterpret_cast<const uint32_t*>(q1 + ks * 16 + 8) : 0u;
        }
      }
    }
    float o[Dv / 8][4];
#pragma unroll
    for (int j = 0; j < Dv / 8; ++j) o[j][0] = o[j][1] = o[j][2] = o[j][3] = 0.f;
    float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
    int s_last = 0;

    for (int it = 0; it < n; ++it, ++gt) {
      const int s = gt % C::STAGES;
      const uint32_t ph = (gt / C::STAGES) & 1;
      s_last = s;
      mbar_wait(&full[s], ph);
      const uint32_t kbase = smem_base + s * C::STAGE_BYTES + (wt * kTokPerWarp) * C::ROW_PITCH + wh * D * EB;
      const uint32_t vbase = kbase + C::KV_BYTES;

      float sc[2][4] = {{0.f, 0.f, 0.f, 0.f}, {0.f, 0.f, 0.f, 0.f}};
      if constexpr (KV8) {
        // 32 bytes of a key row = 32 k: matrices (keys 0-7 | 8-15) x (k 0-15 | 16-31); a register = four codes
#pragma unroll
        for (int kq = 0; kq < D / 32; ++kq) {
          uint32_t r0, r1, r2, r3;
          ldsm_x4(kbase + k_lane_off + kq * 32, r0, r1, r2, r3);
          mma_16816<false>(sc[0], qf[2 * kq], e4m3x2_to_f16x2(r0, false), e4m3x2_to_f16x2(r0, true));
          mma_16816<false>(sc[0], qf[2 * kq + 1], e4m3x2_to_f16x2(r1, false), e4m3x2_to_f16x2(r1, true));
          mma_16816<false>(sc[1], qf[2 * kq], e4m3x2_to_f16x2(r2, false), e4m3x2_to_f16x2(r2, true));
          mma_16816<false>(sc[1], qf[2 * kq + 1], e4m3x2_to_f16x2(r3, false), e4m3x2_to_f16x2(r3, true));
        }
      } else {
#pragma unroll
        for (int ks = 0; ks < D / 16; ++ks) {
          uint32_t r0, r1, r2, r3;
          ldsm_x4(kbase + k_lane_off + ks * 32, r0, r1, r2, r3);
          mma_16816<BF16>(sc[0], qf[ks], r0, r1);
          mma_16816<BF16>(sc[1], qf[ks], r2, r3);
        }
      }
      const int64_t tok0 = kv0 + (int64_t)it * C::TILE_T + wt * kTokPerWarp;
      float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
      for (int nt = 0; nt < 2; ++nt)
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          float v = sc[nt][e];
          if (has_cap) {
            v *= sm_scale_k;
            v = p.logit_cap * tanhf(v / p.logit_cap) * kLog2e;
          } else {
            v *= scale2;
          }
          const int64_t pos = tok0 + nt * 8 + 2 * t + (e & 1);
          if (pos >= kv1) v = -INFINITY;
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
      // P is split into hi + lo half-precision parts (p = hi + lo up to 2^-17) and PV is issued twice,
      // so the result matches an fp32-P reference (decode.cpp:972-984 keeps P in fp32 for MHA/GQA);
      // the extra MMAs are free for an HBM-bound kernel.
      uint32_t pa[4], pb[4];
      float ps[2] = {0.f, 0.f};
#pragma unroll
      for (int nt = 0; nt < 2; ++nt) {
        const float p0 = exp2f(sc[nt][0] - m_use[0]), p1 = exp2f(sc[nt][1] - m_use[0]);
        const float p2 = exp2f(sc[nt][2] - m_use[1]), p3 = exp2f(sc[nt][3] - m_use[1]);
        pa[nt * 2 + 0] = pack_half2<MB>(p0, p1);
        pa[nt * 2 + 1] = pack_half2<MB>(p2, p3);
        const float2 h01 = unpack_half2<MB>(pa[nt * 2 + 0]), h23 = unpack_half2<MB>(pa[nt * 2 + 1]);
        pb[nt * 2 + 0] = pack_half2<MB>(p0 - h01.x, p1 - h01.y);
        pb[nt * 2 + 1] = pack_half2<MB>(p2 - h23.x, p3 - h23.y);
        ps[0] += p0 + p1;
        ps[1] += p2 + p3;
      }
      l_run[0] = l_run[0] * alpha[0] + ps[0];
      l_run[1] = l_run[1] * alpha[1] + ps[1];
      if (alpha[0] != 1.f || alpha[1] != 1.f) {
#pragma unroll
        for (int j = 0; j < Dv / 8; ++j) {
          o[j][0] *= alpha[0];
          o[j][1] *= alpha[0];
          o[j][2] *= alpha[1];
          o[j][3] *= alpha[1];
        }
      }
      if constexpr (KV8) {
        // 32 bytes of the value rows = 32 columns.  The transposing ldmatrix moves byte PAIRS: thread (g, t) receives
        // the codes V[tok 2t | 2t+1][col 2g | 2g+1]; a byte permute separates the even and the odd column, each goes
        // through its own MMA, and the thread ends up owning the four adjacent columns 16 q + 4 t .. + 3 of block q:
        // o[2q] holds columns +0 / +2 (c0 / c1), o[2q+1] columns +1 / +3 (the scratch write below knows).
#pragma unroll
        for (int jq = 0; jq < Dv / 32; ++jq) {
          uint32_t r0, r1, r2, r3;
          ldsm_x4_t(vbase + v_lane_off + jq * 32, r0, r1, r2, r3);
          uint32_t be0 = e4m3x2_to_f16x2(__byte_perm(r0, 0, 0x4420), false), be1 = e4m3x2_to_f16x2(__byte_perm(r1, 0, 0x4420), false);
          uint32_t bo0 = e4m3x2_to_f16x2(__byte_perm(r0, 0, 0x4431), false), bo1 = e4m3x2_to_f16x2(__byte_perm(r1, 0, 0x4431), false);
          // (P goes in as f16 -- 11 mantissa bits against V's 3 -- so the hi + lo second pass of the bf16 path is not needed)
          mma_16816<false>(o[4 * jq], pa, be0, be1);
          mma_16816<false>(o[4 * jq + 1], pa, bo0, bo1);
          be0 = e4m3x2_to_f16x2(__byte_perm(r2, 0, 0x4420), false); be1 = e4m3x2_to_f16x2(__byte_perm(r3, 0, 0x4420), false);
          bo0 = e4m3x2_to_f16x2(__byte_perm(r2, 0, 0x4431), false); bo1 = e4m3x2_to_f16x2(__byte_perm(r3, 0, 0x4431), false);
          mma_16816<false>(o[4 * jq + 2], pa, be0, be1);
          mma_16816<false>(o[4 * jq + 3], pa, bo0, bo1);
        }
      } else {
#pragma unroll
        for (int jp = 0; jp < Dv / 16; ++jp) {
          uint32_t r0, r1, r2, r3;
          ldsm_x4_t(vbase + v_lane_off + jp * 32, r0, r1, r2, r3);
          mma_16816<BF16>(o[2 * jp], pa, r0, r1);
          mma_16816<BF16>(o[2 * jp + 1], pa, r2, r3);
          mma_16816<BF16>(o[2 * jp], pb, r0, r1);
          mma_16816<BF16>(o[2 * jp + 1], pb, r2, r3);
        }
      }
      if (it + 1 < n) {  // the segment's last stage is released after it has served as merge scratch
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty[s]);
      }
    }

    // quad-reduce the row sums
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
      l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
    }

    // ---- merge the WT token-warps of each head through the stage the consumers still own ----
    consumer_bar();  // every warp is done reading K/V of the last tile
    float* scratch = reinterpret_cast<float*>(smem + s_last * C::STAGE_BYTES);  // [warp][16][Dv + 2]
    {
      float* my = scratch + (size_t)warp * 16 * (Dv + 2);
      if constexpr (KV8) {   // columns 16 q + 4 t .. + 3: (even tile c0, odd tile c0, even tile c1, odd tile c1)
#pragma unroll
        for (int q = 0; q < Dv / 16; ++q) {
          float* d0 = my + g * (Dv + 2) + q * 16 + 4 * t;
          float* d1 = my + (g + 8) * (Dv + 2) + q * 16 + 4 * t;
          *reinterpret_cast<float2*>(d0) = make_float2(o[2 * q][0], o[2 * q + 1][0]);
          *reinterpret_cast<float2*>(d0 + 2) = make_float2(o[2 * q][1], o[2 * q + 1][1]);
          *reinterpret_cast<float2*>(d1) = make_float2(o[2 * q][2], o[2 * q + 1][2]);
          *reinterpret_cast<float2*>(d1 + 2) = make_float2(o[2 * q][3], o[2 * q + 1][3]);
        }
      } else {
#pragma unroll
        for (int j = 0; j < Dv / 8; ++j) {
          *reinterpret_cast<float2*>(my + g * (Dv + 2) + j * 8 + 2 * t) = make_float2(o[j][0], o[j][1]);
          *reinterpret_cast<float2*>(my + (g + 8) * (Dv + 2) + j * 8 + 2 * t) = make_float2(o[j][2], o[j][3]);
        }
      }
      if (t == 0) {
        my[g * (Dv + 2) + Dv] = m_run[0];
        my[g * (Dv + 2) + Dv + 1] = l_run[0];
        my[(g + 8) * (Dv + 2) + Dv] = m_run[1];
        my[(g + 8) * (Dv + 2) + Dv + 1] = l_run[1];
      }
    }
    consumer_bar();
    // every consumer thread produces elements of (head h, row r, dv): HPC*rows_valid*Dv elements
    const int slot_row0 = ((int)blockIdx.x * 2 + (first_seg ? 0 : 1)) * (HPC * 16);
    const int total = HPC * rows_valid * Dv;
    for (int e = threadIdx.x; e < total; e += kConsumerWarps * 32) {
      const int dv = e % Dv;
      const int r = (e / Dv) % rows_valid;
      const int h = e / (Dv * rows_valid);
      const float* base = scratch + (size_t)(h * C::WT) * 16 * (Dv + 2) + r * (Dv + 2);
      float M = -INFINITY;
#pragma unroll
      for (int w = 0; w < C::WT; ++w) M = fmaxf(M, base[(size_t)w * 16 * (Dv + 2) + Dv]);
      float L = 0.f, O = 0.f;
      const float Mu = (M == -INFINITY) ? 0.f : M;
#pragma unroll
      for (int w = 0; w < C::WT; ++w) {
        const float* bw = base + (size_t)w * 16 * (Dv + 2);
        const float sc_w = exp2f(bw[Dv] - Mu);
        L += bw[Dv + 1] * sc_w;
        O += bw[dv] * sc_w;
      }
      const float inv = (L > 0.f ? 1.f / L : 0.f) * (KV8 ? p.v_scale : 1.f);   // V = code * v_scale
      if (whole_unit) {
        const int head = (kvh0 + h) * p.G + row0 + r;
        if (p.casc_o) {   // cascade: the merge kernel still has to add the shared-prefix state -> fp32 state, not the output
          p.casc_o[((int64_t)b * p.Hq + head) * Dv + dv] = O * inv;
          if (dv == 0) p.casc_lse[(int64_t)b * p.Hq + head] = L > 0.f ? M + log2f(L) : -INFINITY;
        } else {
          p.out[((int64_t)b * p.Hq + head) * Dv + dv] = float_to_half<BF16>(O * inv);
        }
      } else {
        const int prow = slot_row0 + h * 16 + r;
        p.part_o[(int64_t)prow * Dv + dv] = O * inv;
        if (dv == 0) p.part_lse[prow] = L > 0.f ? M + log2f(L) : -INFINITY;
      }
    }
    fence_proxy_async();  // the stage goes back to the async-proxy writers (TMA) after generic writes
    consumer_bar();
    if (lane == 0) mbar_arrive(&empty[s_last]);
    cur.advance(n);
  }
}

// merge of the units that more than one CTA worked on (decode.cpp:812-860 decode_accumulate_kv_splits;
// decode_attention.py:492-552 _fwd_kernel_stage2).  One WARP per (request, head): grid = B * ceil(Hq / 8) CTAs of
// 8 warps, a lane owns 4 consecutive output columns of every 128.  Warps are independent -- no shared memory, no
// block barrier: each one recomputes the partition from seq_lens, keeps its (row, weight) pairs in registers and
// hands them round with shuffles.  NV = ceil(Dv / 128).
// QUANT (Dv == 128, one quantisation group per head): every (request, head) row is also emitted as e4m3 + scale;
// rows that needed no merge are read back from `out` (written by the attention kernel, which the grid dependency
// orders before this kernel).
template <bool BF16>
__device__ __forceinline__ void merge_quant_row(const MergeQuant& mq, uint2 pk, int b, int h, int lane) {
  const float2 f0 = unpack_half2<BF16>(pk.x), f1 = unpack_half2<BF16>(pk.y);
  const float v[4] = {f0.x, f0.y, f1.x, f1.y};
  float amax = mq.eps;
#pragma unroll
  for (int i = 0; i < 4; ++i) amax = fmaxf(amax, fabsf(v[i]));
  amax = warp_max(amax);
  const float ys = __fdiv_rn(amax, mq.fmax_);
  if (lane == 0) mq.s[b * mq.s_sm + h * mq.s_sg] = ys;
  uint32_t q4 = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const float qv = fminf(fmaxf(__fdiv_rn(v[j], ys), mq.fmin_), mq.fmax_);
    q4 |= (uint32_t)__nv_cvt_float_to_fp8(qv, __NV_SATFINITE, __NV_E4M3) << (8 * j);
  }
  *reinterpret_cast<uint32_t*>(mq.q + b * mq.q_sm + h * 128 + lane * 4) = q4;
}

template <bool BF16, int NV, bool QUANT>
__global__ void __launch_bounds__(256) decode_merge_part_kernel(const float* __restrict__ part_o,
                                                                const float* __restrict__ part_lse,
                                                                uint16_t* out,
                                                                const void* __restrict__ seq_lens, int lens32, int B, int Hq,
                                                                int Ggrp, int HPC, int tile_t, int NG, int n_qblk,
                                                                int grid, int Dv, const MergeQuant mq,
                                                                const int32_t* __restrict__ kv_start_p,
                                                                const float* __restrict__ casc_o,
                                                                const float* __restrict__ casc_lse,
                                                                const float* __restrict__ pre_o,
                                                                const float* __restrict__ pre_lse) {
  // Everything up to the first read of a partial depends only on seq_lens (and the cascade prefix length), which the
  // attention kernel does not write: it runs BEFORE the grid-dependency wait, i.e. under the tail of the attention kernel
  // on the SMs its first finished CTAs have left.  (This kernel is a pure latency chain -- measured in place: 8.4 us after a
  // 174 us attention kernel at cfg 2, 10.3 us at cfg 4 TP8 -- so every round trip taken off it counts.)
  griddep_launch();
  const int lane = threadIdx.x & 31;
  const int hblocks = (Hq + 7) >> 3;
  const int b = blockIdx.x / hblocks, h = (blockIdx.x - b * hblocks) * 8 + (threadIdx.x >> 5);
  if (h >= Hq) return;
  const int64_t kv_start = kv_start_p ? (int64_t)*kv_start_p : 0;
  const bool cascade = casc_o != nullptr;
  const bool tile_pow2 = (tile_t & (tile_t - 1)) == 0;
  const int tile_sh = __ffs(tile_t) - 1;
  int all = 0, before = 0, tiles_b = 0;
  for (int i = lane; i < B; i += 32) {
    int64_t len = meta_at(seq_lens, i, lens32) - kv_start;
    if (len < 0) len = 0;
    const int tl = tile_pow2 ? (int)((len + tile_t - 1) >> tile_sh) : (int)((len + tile_t - 1) / tile_t);   // (no 64-bit division)
    all += tl;
    if (i < b) before += tl;
    if (i == b) tiles_b = tl;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    all += __shfl_xor_sync(0xffffffffu, all, o);
    before += __shfl_xor_sync(0xffffffffu, before, o);
    tiles_b += __shfl_xor_sync(0xffffffffu, tiles_b, o);
  }
  uint16_t* orow = out + ((int64_t)b * Hq + h) * Dv;
  if (tiles_b == 0 && !cascade) {  // seq_len == 0: defined output (zeros)
    griddep_wait();
    for (int d = lane; d < Dv; d += 32) orow[d] = 0;
    if constexpr (QUANT) merge_quant_row<BF16>(mq, make_uint2(0u, 0u), b, h, lane);
    return;
  }
  const int kvh = h / Ggrp, within = h - kvh * Ggrp;
  const int hg = kvh / HPC, hl = kvh - hg * HPC, qblk = within >> 4, r = within & 15;
  const int g = hg * n_qblk + qblk;
  const int64_t W = (int64_t)NG * all;
  const int64_t t0 = (int64_t)NG * before + (int64_t)g * tiles_b, t1 = t0 + tiles_b;
  const int c0 = tiles_b > 0 ? (int)part_owner(t0, W, grid) : 0, c1 = tiles_b > 0 ? (int)part_owner(t1 - 1, W, grid) : 0;
  const bool whole = c0 == c1;  // one CTA saw the whole unit
  griddep_wait();               // from here on: what the attention kernel wrote
  if (whole && !cascade) {      // ... and wrote the output itself
    if constexpr (QUANT) merge_quant_row<BF16>(mq, *reinterpret_cast<const uint2*>(orow + lane * 4), b, h, lane);
    return;
  }
  // partial (o, lse) pairs of the CTAs c0..c1 (CTAs whose range is empty wrote nothing)
  const int rows = HPC * 16;
  const int n = (whole || tiles_b == 0) ? 0 : c1 - c0 + 1;  // <= grid <= kMaxPartials
  constexpr int R = kMaxPartials / 32;
  int prow[R];
  float w[R];
  float M = -INFINITY;
#pragma unroll
  for (int k = 0; k < R; ++k) {
    const int i = lane + 32 * k;
    prow[k] = -1;
    w[k] = -INFINITY;
    if (i < n) {
      const int c = c0 + i;
      const int64_t cs = part_start(c, W, grid);
      if (cs < part_start(c + 1, W, grid)) {
        prow[k] = (c * 2 + (t0 <= cs ? 0 : 1)) * rows + hl * 16 + r;
        w[k] = part_lse[prow[k]];
      }
    }
    M = fmaxf(M, w[k]);
  }
  M = warp_max(M);
  // cascade: two more states of this (request, head) -- the suffix state of a whole unit (fp32, written by the attention
  // kernel instead of the output) and the state over the shared prefix (computed once for the batch; natural-log LSE)
  float lse_w = -INFINITY, lse_p = -INFINITY;
  if (cascade) {
    if (whole && tiles_b > 0) lse_w = casc_lse[(int64_t)b * Hq + h];
    lse_p = pre_lse[(int64_t)b * Hq + h] * kLog2e;
    M = fmaxf(M, fmaxf(lse_w, lse_p));
  }
  const float Mu = (M == -INFINITY) ? 0.f : M;
  float L = 0.f;
#pragma unroll
  for (int k = 0; k < R; ++k) {
    w[k] = prow[k] >= 0 ? exp2f(w[k] - Mu) : 0.f;
    L += w[k];
  }
  L = warp_sum(L);
  const float w_w = lse_w > -INFINITY ? exp2f(lse_w - Mu) : 0.f, w_p = lse_p > -INFINITY ? exp2f(lse_p - Mu) : 0.f;
  L += w_w + w_p;
  const float inv = L > 0.f ? 1.f / L : 0.f;
  float4 acc[NV];
#pragma unroll
  for (int v = 0; v < NV; ++v) acc[v] = make_float4(0.f, 0.f, 0.f, 0.f);
  if (cascade) {
#pragma unroll
    for (int v = 0; v < NV; ++v) {
      const int d4 = lane + 32 * v;
      if (d4 * 4 < Dv) {
        if (w_w > 0.f) {
          const float4 x = reinterpret_cast<const float4*>(casc_o + ((int64_t)b * Hq + h) * Dv)[d4];
          acc[v].x += w_w * x.x; acc[v].y += w_w * x.y; acc[v].z += w_w * x.z; acc[v].w += w_w * x.w;
        }
        if (w_p > 0.f) {
          const float4 x = reinterpret_cast<const float4*>(pre_o + ((int64_t)b * Hq + h) * Dv)[d4];
          acc[v].x += w_p * x.x; acc[v].y += w_p * x.y; acc[v].z += w_p * x.z; acc[v].w += w_p * x.w;
        }
      }
    }
  }
#pragma unroll
  for (int k = 0; k < R; ++k) {
    if (32 * k < n) {  // warp-uniform
      const int cnt = n - 32 * k < 32 ? n - 32 * k : 32;
#pragma unroll 4
      for (int j = 0; j < cnt; ++j) {
        const float wj = __shfl_sync(0xffffffffu, w[k], j);
        const int pj = __shfl_sync(0xffffffffu, prow[k], j);
        if (wj > 0.f) {
          const float4* src = reinterpret_cast<const float4*>(part_o + (int64_t)pj * Dv);
#pragma unroll
          for (int v = 0; v < NV; ++v) {
            const int d4 = lane + 32 * v;
            if (d4 * 4 < Dv) {
              const float4 x = src[d4];
              acc[v].x += wj * x.x;
              acc[v].y += wj * x.y;
              acc[v].z += wj * x.z;
              acc[v].w += wj * x.w;
            }
          }
        }
      }
    }
  }
  const bool vec_ok = (reinterpret_cast<uintptr_t>(orow) & 7) == 0;
#pragma unroll
  for (int v = 0; v < NV; ++v) {
    const int d = (lane + 32 * v) * 4;
    if (d < Dv) {
      if (vec_ok) {
        const uint2 pk = make_uint2(pack_half2<BF16>(acc[v].x * inv, acc[v].y * inv), pack_half2<BF16>(acc[v].z * inv, acc[v].w * inv));
        *reinterpret_cast<uint2*>(orow + d) = pk;
        if constexpr (QUANT) merge_quant_row<BF16>(mq, pk, b, h, lane);  // NV == 1, Dv == 128: every lane is here
      } else {
        orow[d] = float_to_half<BF16>(acc[v].x * inv);
        orow[d + 1] = float_to_half<BF16>(acc[v].y * inv);
        orow[d + 2] = float_to_half<BF16>(acc[v].z * inv);
        orow[d + 3] = float_to_half<BF16>(acc[v].w * inv);
      }
    }
  }
}

// launcher shared by the GQA and the MLA kernels (Dv % 4 == 0, Dv <= 512)
template <bool BF16>
static int launch_merge_part(const DecodeKParams& kp, int Ggrp, int HPC, int tile_t, int NG, int n_qblk, int grid, cudaStream_t st) {
  const dim3 g(kp.B * ((kp.Hq + 7) / 8)), blk(256);
  if (kp.mq.q)  // Dv == 128 and an 8-byte aligned `out` were checked by the entry point
    launch_pdl(decode_merge_part_kernel<BF16, 1, true>, g, blk, 0, st, kp.part_o, kp.part_lse, kp.out, kp.seq_lens, kp.meta32 & 1, kp.B, kp.Hq, Ggrp,
               HPC, tile_t, NG, n_qblk, grid, kp.Dv, kp.mq, kp.kv_start, (const float*)kp.casc_o, (const float*)kp.casc_lse, kp.pre_o, kp.pre_lse);
  else if (kp.Dv <= 128)
    launch_pdl(decode_merge_part_kernel<BF16, 1, false>, g, blk, 0, st, kp.part_o, kp.part_lse, kp.out, kp.seq_lens, kp.meta32 & 1, kp.B, kp.Hq, Ggrp,
               HPC, tile_t, NG, n_qblk, grid, kp.Dv, kp.mq, kp.kv_start, (const float*)kp.casc_o, (const float*)kp.casc_lse, kp.pre_o, kp.pre_lse);
  else
    launch_pdl(decode_merge_part_kernel<BF16, 4, false>, g, blk, 0, st, kp.part_o, kp.part_lse, kp.out, kp.seq_lens, kp.meta32 & 1, kp.B, kp.Hq, Ggrp,
               HPC, tile_t, NG, n_qblk, grid, kp.Dv, kp.mq, kp.kv_start, (const float*)kp.casc_o, (const float*)kp.casc_lse, kp.pre_o, kp.pre_lse);
  return launched("decode_merge_part_kernel launch");
}

// ------------------------------------------------------------------------------------------------
// generic SIMT kernel: any D, Dv <= 32*MAXE (MLA 576/512, odd head sizes of the reference tests).
// grid = (split, Hq, B), 4 warps; warp w handles tokens kv0 + w, kv0 + w + 4, ...
// ------------------------------------------------------------------------------------------------
template <int MAXE, bool BF16>
__global__ void __launch_bounds__(128) decode_attn_generic_kernel(const __grid_constant__ DecodeKParams p) {
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int split = blockIdx.x, h = blockIdx.y, b = blockIdx.z;
  const int kvh = h / p.G;
  const int D = p.D, Dv = p.Dv;
  const int64_t len = seq_len_at(p, b), req = req_at(p, b);
  const int64_t loc_b = p.fused_new ? loc_at(p, b) : -1;
  const int64_t page_mask = ((int64_t)1 << p.page_shift) - 1;
  const int64_t per = (len + p.num_splits - 1) / p.num_splits;
  const int64_t kv0 = (int64_t)split * per;
  const int64_t kv1 = (kv0 + per < len) ? kv0 + per : len;
  __shared__ float sm_o[4][32 * MAXE];
  __shared__ float sm_ml[4][2];

  if (p.fused_new && split == 0 && (h % p.G) == 0) {
    // fused set_kv_buffer for kv head kvh (MLA: k only, v aliases k)
    const uint16_t* ks = p.k_new + b * p.kn_sb + kvh * p.kn_sh;
    uint16_t* kd = p.k_buf + loc_b * p.kb_st + kvh * p.kb_sh;
    for (int i = threadIdx.x; i < D; i += 128) kd[i] = ks[i];
    if (!p.is_mla) {
      const uint16_t* vs = p.v_new + b * p.vn_sb + kvh * p.vn_sh;
      uint16_t* vd = p.v_buf + loc_b * p.vb_st + kvh * p.vb_sh;
      for (int i = threadIdx.x; i < Dv; i += 128) vd[i] = vs[i];
    }
  }

  float qr[MAXE], acc[MAXE];
#pragma unroll
  for (int e = 0; e < MAXE; ++e) {
    const int d = e * 32 + lane;
    qr[e] = 0.f;
    if (d < D) {
      const uint16_t* qp = (p.q_pe && d >= Dv) ? p.q_pe + b * p.qpe_sb + (int64_t)h * p.qpe_sh + (d - Dv)
                                               : p.q + b * p.q_sb + (int64_t)h * p.q_sh + d;
      qr[e] = half_to_float<BF16>(*qp);
    }
    acc[e] = 0.f;
  }
  float m_run = -INFINITY, l_run = 0.f;
  const bool has_cap = p.logit_cap > 0.f;
  for (int64_t pos = kv0 + warp; pos < kv1; pos += 4) {
    const int64_t idx = (load_index(p.r2t, req * p.r2t_stride + (pos >> p.page_shift), p.idx64) << p.page_shift) + (pos & page_mask);
    const uint16_t* kr = p.k_buf + idx * p.kb_st + kvh * p.kb_sh;
    const uint16_t* vr = p.v_buf + idx * p.vb_st + kvh * p.vb_sh;
    if (idx == loc_b) {
      kr = p.k_new + b * p.kn_sb + kvh * p.kn_sh;
      vr = p.is_mla ? kr : p.v_new + b * p.vn_sb + kvh * p.vn_sh;
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
    float pr = exp2f(s - m_new);
    l_run = l_run * alpha + pr;
    // (the reference's MLA path rounds pr to bf16 here, decode.cpp:1157-1159; we keep fp32 -- see DESIGN.md)
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
  float M = fmaxf(fmaxf(sm_ml[0][0], sm_ml[1][0]), fmaxf(sm_ml[2][0], sm_ml[3][0]));
  const float Mu = (M == -INFINITY) ? 0.f : M;
  float w[4], L = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    w[i] = exp2f(sm_ml[i][0] - Mu);
    L += sm_ml[i][1] * w[i];
  }
  const float inv = L > 0.f ? 1.f / L : 0.f;
  for (int d = threadIdx.x; d < Dv; d += 128) {
    const float O = sm_o[0][d] * w[0] + sm_o[1][d] * w[1] + sm_o[2][d] * w[2] + sm_o[3][d] * w[3];
    if (p.num_splits == 1)
      p.out[((int64_t)b * p.Hq + h) * Dv + d] = float_to_half<BF16>(O * inv);
    else
      p.part_o[(((int64_t)b * p.Hq + h) * p.num_splits + split) * Dv + d] = O * inv;
  }
  if (threadIdx.x == 0 && p.num_splits > 1)
    p.part_lse[((int64_t)b * p.Hq + h) * p.num_splits + split] = L > 0.f ? M + log2f(L) : -INFINITY;
}

// ------------------------------------------------------------------------------------------------
// LSE merge of the KV splits (decode.cpp:812-860 decode_accumulate_kv_splits;
// decode_attention.py:492-552 _fwd_kernel_stage2).  grid = B*Hq, block = 128.
// ------------------------------------------------------------------------------------------------
template <bool BF16>
__global__ void __launch_bounds__(128) decode_merge_kernel(const float* __restrict__ part_o,
                                                           const float* __restrict__ part_lse, uint16_t* __restrict__ out,
                                                           int S, int Dv) {
  const int64_t bh = blockIdx.x;
  const float* lse = part_lse + bh * S;
  float M = -INFINITY;
  for (int s = 0; s < S; ++s) M = fmaxf(M, lse[s]);
  const float Mu = (M == -INFINITY) ? 0.f : M;
  float L = 0.f;
  for (int s = 0; s < S; ++s) L += exp2f(lse[s] - Mu);
  const float inv = L > 0.f ? 1.f / L : 0.f;
  for (int d = threadIdx.x; d < Dv; d += 128) {
    float acc = 0.f;
    for (int s = 0; s < S; ++s) {
      const float w = exp2f(lse[s] - Mu);
      if (w > 0.f) acc += w * part_o[(bh * S + s) * Dv + d];
    }
    out[bh * Dv + d] = float_to_half<BF16>(acc * inv);
  }
}

// KVCache.set_kv_buffer (memory_pool.py:418-456 / decode.cpp:771-809): grid = (B, Hkv), 16-bit elements
// k_pe (MLA, set_mla_kv_buffer memory_pool.py:859-876): columns [D - Dpe, D) of the row come from k_pe [B, Dpe]
__global__ void set_kv_kernel(uint16_t* k_buf, int64_t kb_st, int64_t kb_sh, uint16_t* v_buf, int64_t vb_st,
                              int64_t vb_sh, const uint16_t* k_new, int64_t kn_sb, int64_t kn_sh, const uint16_t* v_new,
                              int64_t vn_sb, int64_t vn_sh, const void* loc, int loc32, int D, int Dv,
                              const uint16_t* k_pe, int64_t kpe_sb, int Dpe) {
  const int b = blockIdx.x, h = blockIdx.y;
  const int64_t l = meta_at(loc, b, loc32);
  if (k_new) {
    const uint16_t* s = k_new + b * kn_sb + h * kn_sh;
    uint16_t* d = k_buf + l * kb_st + h * kb_sh;
    const int Dn = k_pe ? D - Dpe : D;
    for (int i = threadIdx.x; i < Dn; i += blockDim.x) d[i] = s[i];
    if (k_pe)
      for (int i = threadIdx.x; i < Dpe; i += blockDim.x) d[Dn + i] = k_pe[b * kpe_sb + i];
  }
  if (v_new) {
    const uint16_t* s = v_new + b * vn_sb + h * vn_sh;
    uint16_t* d = v_buf + l * vb_st + h * vb_sh;
    for (int i = threadIdx.x; i < Dv; i += blockDim.x) d[i] = s[i];
  }
}

// FP8 KV cache append (memory_pool.py:431-440: cache_k.div_(k_scale); cache_k.to(float8_e4m3fn); store as uint8):
// code = e4m3(x / scale), round to nearest even, saturating -- what torch's .to(float8_e4m3fn) does for finite values in
// range (the oracle restates it).  grid (B, Hkv), strides of the pool in bytes (= elements).
template <bool BF16>
__global__ void set_kv_fp8_kernel(uint8_t* k_buf, int64_t kb_st, int64_t kb_sh, uint8_t* v_buf, int64_t vb_st, int64_t vb_sh,
                                  const uint16_t* k_new, int64_t kn_sb, int64_t kn_sh, const uint16_t* v_new, int64_t vn_sb,
                                  int64_t vn_sh, const void* loc, int loc32, int D, int Dv, float k_scale, float v_scale) {
  const int b = blockIdx.x, h = blockIdx.y;
  const int64_t l = meta_at(loc, b, loc32);
  for (int i = threadIdx.x; i < D; i += blockDim.x) {
    const float x = __fdiv_rn(half_to_float<BF16>(k_new[b * kn_sb + h * kn_sh + i]), k_scale);
    k_buf[l * kb_st + h * kb_sh + i] = (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
  }
  for (int i = threadIdx.x; i < Dv; i += blockDim.x) {
    const float x = __fdiv_rn(half_to_float<BF16>(v_new[b * vn_sb + h * vn_sh + i]), v_scale);
    v_buf[l * vb_st + h * vb_sh + i] = (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
  }
}

// ------------------------------------------------------------------------------------------------
// host side
// ------------------------------------------------------------------------------------------------
static int pick_hpc(int Hkv) {
  if (Hkv % 8 == 0) return 8;
  if (Hkv % 4 == 0) return 4;
  if (Hkv % 2 == 0) return 2;
  return 1;
}

static bool mma_path_ok(int Hq, int Hkv, int D, int Dv) { return D == Dv && (D == 64 || D == 128) && Hq % Hkv == 0; }

static int attn_grid() {
  const int sms = device_sm_count();
  const int g = sms > 0 ? sms : 148;
  return g > kMaxPartials ? kMaxPartials : g;
}

template <int D, int HPC, bool BF16, bool KV8 = false>
static int launch_mma(const DecodeKParams& kp, cudaStream_t st) {
  using C = AttnCfg<D, HPC, KV8 ? 1 : 2>;
  auto kern = decode_attn_mma_kernel<D, HPC, BF16, KV8>;
  static bool configured = false;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, C::SMEM_BYTES));
    configured = true;
  }
  const int grid = attn_grid();
  launch_pdl(kern, dim3(grid), dim3(kAttnThreads), C::SMEM_BYTES, st, kp);
  if (int rc = launched("decode_attn_mma_kernel launch")) return rc;
  const int n_qblk = (kp.G + 15) / 16;
  return launch_merge_part<BF16>(kp, kp.G, HPC, C::TILE_T, (kp.Hkv / HPC) * n_qblk, n_qblk, grid, st);
}

template <bool BF16>
static int dispatch_mma(const DecodeKParams& kp, int hpc, cudaStream_t st) {
#define B200_CASE(DD, HH) \
  if (kp.D == DD && hpc == HH) return launch_mma<DD, HH, BF16>(kp, st);
  B200_CASE(128, 8) B200_CASE(128, 4) B200_CASE(128, 2) B200_CASE(128, 1)
  B200_CASE(64, 8) B200_CASE(64, 4) B200_CASE(64, 2) B200_CASE(64, 1)
#undef B200_CASE
  set_error("decode_attn: no mma instantiation for D=%d hpc=%d", kp.D, hpc);
  return B200_ERR_UNSUPPORTED;
}

// FP8 KV cache: D = 128 only (Llama / Qwen head size)
template <bool BF16>
static int dispatch_mma_kv8(const DecodeKParams& kp, int hpc, cudaStream_t st) {
  if (kp.D == 128 && hpc == 8) return launch_mma<128, 8, BF16, true>(kp, st);
  if (kp.D == 128 && hpc == 4) return launch_mma<128, 4, BF16, true>(kp, st);
  if (kp.D == 128 && hpc == 2) return launch_mma<128, 2, BF16, true>(kp, st);
  if (kp.D == 128 && hpc == 1) return launch_mma<128, 1, BF16, true>(kp, st);
  set_error("decode_attn: the FP8 KV path needs head_dim 128 (got %d)", kp.D);
  return B200_ERR_UNSUPPORTED;
}

// scratch of the tuned path: two partial slots per CTA, each HPC*16 rows of (Dv + 1) floats
static size_t mma_workspace_bytes(int Dv) { return (size_t)attn_grid() * 2 * (8 * 16) * (Dv + 1) * sizeof(float) + 256; }

template <bool BF16>
static int launch_generic(const DecodeKParams& kp, cudaStream_t st) {
  dim3 grid(kp.num_splits, kp.Hq, kp.B);
  const int mx = kp.D > kp.Dv ? kp.D : kp.Dv;
  if (mx <= 128) decode_attn_generic_kernel<4, BF16><<<grid, 128, 0, st>>>(kp);
  else if (mx <= 256) decode_attn_generic_kernel<8, BF16><<<grid, 128, 0, st>>>(kp);
  else if (mx <= 640) decode_attn_generic_kernel<20, BF16><<<grid, 128, 0, st>>>(kp);
  else decode_attn_generic_kernel<32, BF16><<<grid, 128, 0, st>>>(kp);
  return launched("decode_attn_generic_kernel launch");
}

#include "decode_mla.cuh"

}  // namespace b200

using namespace b200;

extern "C" int b200_decode_attn_num_splits(int B, int Hq, int Hkv, int D, int Dv, int max_seq_len, int num_splits_hint) {
  if (B <= 0 || Hq <= 0 || Hkv <= 0 || Hq % Hkv != 0) return 1;
  // tuned path: the KV work is partitioned on the device from the actual seq_lens (no fixed split count)
  if (mma_path_ok(Hq, Hkv, D, Dv)) return 1;
  if (Hkv == 1 && D == MlaCfg::D && Dv == MlaCfg::Dv && num_splits_hint <= 1) return 1;  // tuned MLA path
  // generic path: 128-thread CTAs, many per SM -> aim for >= 8 CTAs per SM
  if (num_splits_hint > 0) return num_splits_hint > 64 ? 64 : num_splits_hint;
  if (max_seq_len <= 0) max_seq_len = 4096;
  const int sms = device_sm_count() > 0 ? device_sm_count() : 148;
  int s = (8 * sms + B * Hq - 1) / (B * Hq);
  int max_s = max_seq_len / 64;
  if (max_s < 1) max_s = 1;
  if (s > max_s) s = max_s;
  if (s > 32) s = 32;
  if (s < 1) s = 1;
  return s;
}

extern "C" size_t b200_decode_attn_workspace_size(int B, int Hq, int Dv, int num_splits) {
  const size_t tuned = (Dv == 64 || Dv == 128) ? mma_workspace_bytes(Dv) : (Dv == MlaCfg::Dv ? mla_workspace_bytes() : 16);
  const size_t generic = num_splits <= 1 ? 16 : ((size_t)B * Hq * num_splits * (Dv + 1)) * sizeof(float) + 256;
  return tuned > generic ? tuned : generic;
}

extern "C" int b200_set_kv_buffer(void* k_buf, int64_t kb_st, int64_t kb_sh, void* v_buf, int64_t vb_st, int64_t vb_sh,
                                  const void* k_new, int64_t kn_sb, int64_t kn_sh, const void* v_new, int64_t vn_sb,
                                  int64_t vn_sh, const int64_t* loc, int B, int Hkv, int D, int Dv, int elem_bytes,
                                  void* stream) {
  B200_REQUIRE(elem_bytes == 2, B200_ERR_UNSUPPORTED, "set_kv_buffer: only 16-bit KV dtypes are supported");
  B200_REQUIRE(B >= 0 && Hkv > 0 && D > 0, B200_ERR_ARG, "set_kv_buffer: bad sizes");
  if (B == 0) return B200_OK;
  set_kv_kernel<<<dim3(B, Hkv), 128, 0, (cudaStream_t)stream>>>(
      (uint16_t*)k_buf, kb_st, kb_sh, (uint16_t*)v_buf, vb_st, vb_sh, (const uint16_t*)k_new, kn_sb, kn_sh,
      (const uint16_t*)v_new, vn_sb, vn_sh, loc, 0, D, Dv, nullptr, 0, 0);
  return launched("set_kv_kernel launch");
}

extern "C" int b200_set_kv_buffer_fp8(void* k_buf, int64_t kb_st, int64_t kb_sh, void* v_buf, int64_t vb_st, int64_t vb_sh,
                                      const void* k_new, int64_t kn_sb, int64_t kn_sh, const void* v_new, int64_t vn_sb,
                                      int64_t vn_sh, const int64_t* loc, int B, int Hkv, int D, int Dv, int src_dtype,
                                      float k_scale, float v_scale, void* stream) {
  B200_REQUIRE(k_buf && v_buf && k_new && v_new && loc && B >= 0 && Hkv > 0 && D > 0 && Dv > 0, B200_ERR_ARG, "set_kv_buffer_fp8: bad argument");
  B200_REQUIRE(src_dtype == B200_BF16 || src_dtype == B200_F16, B200_ERR_UNSUPPORTED, "set_kv_buffer_fp8: source must be bf16/f16");
  if (B == 0) return B200_OK;
  const float ks = k_scale > 0.f ? k_scale : 1.f, vs = v_scale > 0.f ? v_scale : 1.f;
  if (src_dtype == B200_BF16)
    set_kv_fp8_kernel<true><<<dim3(B, Hkv), 128, 0, (cudaStream_t)stream>>>((uint8_t*)k_buf, kb_st, kb_sh, (uint8_t*)v_buf, vb_st, vb_sh,
        (const uint16_t*)k_new, kn_sb, kn_sh, (const uint16_t*)v_new, vn_sb, vn_sh, loc, 0, D, Dv, ks, vs);
  else
    set_kv_fp8_kernel<false><<<dim3(B, Hkv), 128, 0, (cudaStream_t)stream>>>((uint8_t*)k_buf, kb_st, kb_sh, (uint8_t*)v_buf, vb_st, vb_sh,
        (const uint16_t*)k_new, kn_sb, kn_sh, (const uint16_t*)v_new, vn_sb, vn_sh, loc, 0, D, Dv, ks, vs);
  return launched("set_kv_fp8_kernel launch");
}

static int decode_attn_impl(const b200_decode_attn_params* a, const MergeQuant* mq, void* stream) {
  B200_REQUIRE(a != nullptr, B200_ERR_ARG, "decode_attn: null params");
  B200_REQUIRE(a->B >= 0 && a->Hq > 0 && a->Hkv > 0 && a->D > 0 && a->Dv > 0, B200_ERR_ARG, "decode_attn: bad sizes");
  B200_REQUIRE(a->Hq % a->Hkv == 0, B200_ERR_ARG, "decode_attn: Hq %% Hkv != 0");
  B200_REQUIRE(a->dtype == B200_BF16 || a->dtype == B200_F16, B200_ERR_UNSUPPORTED, "decode_attn: dtype must be bf16/f16");
  B200_REQUIRE(a->D <= 1024 && a->Dv <= 1024, B200_ERR_UNSUPPORTED, "decode_attn: head dim > 1024");
  B200_REQUIRE(a->q && a->k_buf && a->v_buf && a->out && a->req_to_token && a->seq_lens, B200_ERR_ARG, "decode_attn: null tensor");
  int page_shift = 0;
  if (a->page_size > 1) {
    B200_REQUIRE((a->page_size & (a->page_size - 1)) == 0, B200_ERR_UNSUPPORTED, "decode_attn: page_size %d is not a power of two", a->page_size);
    while ((1 << page_shift) < a->page_size) ++page_shift;
  }
  if (a->B == 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "decode_attn: requires an sm_100 device (got sm_%d); no fallback", arch);
  cudaStream_t st = (cudaStream_t)stream;
  const bool is_mla = (a->k_buf == a->v_buf && a->Hkv == 1 && a->D == a->Dv + 64);  // decode.cpp:1455
  B200_REQUIRE(!(is_mla && a->logit_cap != 0.f), B200_ERR_ARG, "decode MLA: expect no logit_cap (decode.cpp:1042)");
  int write_kv = a->write_kv;
  if (write_kv) {
    B200_REQUIRE(a->k_new && (is_mla || a->v_new) && a->out_cache_loc, B200_ERR_ARG, "decode_attn: write_kv needs k_new/v_new/out_cache_loc");
  }

  DecodeKParams kp;
  kp.q = (const uint16_t*)a->q; kp.q_sb = a->q_stride_b; kp.q_sh = a->q_stride_h;
  kp.k_new = (const uint16_t*)a->k_new; kp.kn_sb = a->knew_stride_b; kp.kn_sh = a->knew_stride_h;
  kp.v_new = (const uint16_t*)a->v_new; kp.vn_sb = a->vnew_stride_b; kp.vn_sh = a->vnew_stride_h;
  kp.k_buf = (uint16_t*)a->k_buf; kp.kb_st = a->kbuf_stride_t; kp.kb_sh = a->kbuf_stride_h;
  kp.v_buf = (uint16_t*)a->v_buf; kp.vb_st = a->vbuf_stride_t; kp.vb_sh = a->vbuf_stride_h;
  kp.out = (uint16_t*)a->out;
  kp.r2t = a->req_to_token; kp.r2t_stride = a->req_to_token_stride; kp.idx64 = a->idx_is_i64;
  kp.rpi = a->req_pool_indices; kp.seq_lens = a->seq_lens; kp.loc = a->out_cache_loc;
  kp.meta32 = a->meta_i32_mask & 7; kp.page_shift = page_shift;
  kp.q_pe = (const uint16_t*)a->q_rope; kp.qpe_sb = a->qrope_stride_b; kp.qpe_sh = a->qrope_stride_h;
  kp.k_pe_new = (const uint16_t*)a->k_rope_new; kp.kpe_sb = a->krope_stride_b;
  B200_REQUIRE(is_mla || (!a->q_rope && !a->k_rope_new), B200_ERR_ARG, "decode_attn: q_rope / k_rope_new are MLA arguments");
  kp.B = a->B; kp.Hq = a->Hq; kp.Hkv = a->Hkv; kp.G = a->Hq / a->Hkv; kp.D = a->D; kp.Dv = a->Dv;
  kp.is_mla = is_mla; kp.sm_scale = a->sm_scale; kp.logit_cap = a->logit_cap;
  kp.mq = MergeQuant{};
  kp.kv_start = (const int32_t*)a->kv_start;
  kp.casc_o = (float*)a->cascade_o; kp.casc_lse = (float*)a->cascade_lse;
  kp.pre_o = (const float*)a->prefix_o; kp.pre_lse = (const float*)a->prefix_lse;
  kp.kv_fp8 = a->kv_dtype == B200_KV_E4M3 ? 1 : 0;
  kp.k_scale = a->k_scale > 0.f ? a->k_scale : 1.f;
  kp.v_scale = a->v_scale > 0.f ? a->v_scale : 1.f;
  B200_REQUIRE(a->kv_dtype == 0 || a->kv_dtype == B200_KV_E4M3, B200_ERR_ARG, "decode_attn: kv_dtype must be 0 (q's dtype) or B200_KV_E4M3");
  const bool cascade = a->kv_start || a->cascade_o || a->cascade_lse || a->prefix_o || a->prefix_lse;
  B200_REQUIRE(!cascade || (a->kv_start && a->cascade_o && a->cascade_lse && a->prefix_o && a->prefix_lse), B200_ERR_ARG,
               "decode_attn: cascade needs kv_start, cascade_o, cascade_lse, prefix_o and prefix_lse together");

  // tuned path needs 16-byte aligned, head-contiguous rows
  bool mma = mma_path_ok(a->Hq, a->Hkv, a->D, a->Dv) && page_shift == 0;
  int hpc = pick_hpc(a->Hkv);
  if (mma) {
    auto al16 = [](const void* p) { return ((uintptr_t)p & 15) == 0; };
    const int am = a->kv_dtype == B200_KV_E4M3 ? 16 : 8;   // elements per 16 bytes of the pool
    const bool rows_ok = al16(a->k_buf) && al16(a->v_buf) && a->kbuf_stride_t % am == 0 && a->vbuf_stride_t % am == 0 &&
                         a->kbuf_stride_h % am == 0 && a->vbuf_stride_h % am == 0;
    const bool q_ok = ((uintptr_t)a->q & 3) == 0 && a->q_stride_b % 2 == 0 && a->q_stride_h % 2 == 0;
    if (!rows_ok || !q_ok) mma = false;
    if (mma && hpc > 1 && (a->kbuf_stride_h != a->D || a->vbuf_stride_h != a->Dv)) hpc = 1;  // heads not adjacent
  }
  if (mq) {
    B200_REQUIRE(mq->q && mq->s, B200_ERR_ARG, "decode_attn_quant_fp8: null output");
    B200_REQUIRE(mma && a->Dv == 128 && ((uintptr_t)a->out & 7) == 0 && ((uintptr_t)mq->q & 3) == 0 && mq->q_sm % 4 == 0,
                 B200_ERR_UNSUPPORTED,
                 "decode_attn_quant_fp8: needs the tuned GQA kernel with Dv == 128 and aligned outputs (quantise separately otherwise)");
    kp.mq = *mq;
  }
  B200_REQUIRE(!cascade || (mma && a->Dv % 4 == 0 && ((uintptr_t)a->prefix_o & 15) == 0 && ((uintptr_t)a->cascade_o & 15) == 0),
               B200_ERR_UNSUPPORTED, "decode_attn: cascade needs the tuned GQA kernel (D == Dv in {64, 128}, aligned rows)");
  // tuned MLA path (decode_mla.cuh): 16-byte aligned latent rows, 4-byte aligned q rows; an explicit split-count
  // hint selects the generic kernel (the tuned kernel partitions the KV work itself)
  bool mla = mla_path_ok(a->Hkv, a->D, a->Dv, is_mla) && a->num_splits <= 1;
  if (mla) {
    const bool rows_ok = ((uintptr_t)a->k_buf & 15) == 0 && a->kbuf_stride_t % 8 == 0;
    bool q_ok = ((uintptr_t)a->q & 3) == 0 && a->q_stride_b % 2 == 0 && a->q_stride_h % 2 == 0;
    if (a->q_rope) q_ok = q_ok && ((uintptr_t)a->q_rope & 3) == 0 && a->qrope_stride_b % 2 == 0 && a->qrope_stride_h % 2 == 0;
    if (!rows_ok || !q_ok) mla = false;
  }
  // fused append (write_kv == 2) needs aligned, head-contiguous new rows on the tuned path
  bool fused = (write_kv == 2);
  if (fused && mla) {
    if (((uintptr_t)a->k_new & 15) != 0 || a->knew_stride_b % 8 != 0) fused = false;
    if (a->k_rope_new && (((uintptr_t)a->k_rope_new & 15) != 0 || a->krope_stride_b % 8 != 0)) fused = false;
  }
  if (fused && !mla && a->k_rope_new) fused = false;  // only the tuned MLA kernel substitutes a two-piece row
  if (fused && mma) {
    auto al16 = [](const void* p) { return ((uintptr_t)p & 15) == 0; };
    const bool new_ok = al16(a->k_new) && al16(a->v_new) && a->knew_stride_b % 8 == 0 && a->vnew_stride_b % 8 == 0 &&
                        a->knew_stride_h % 8 == 0 && a->vnew_stride_h % 8 == 0 &&
                        (hpc == 1 || (a->knew_stride_h == a->D && a->vnew_stride_h == a->Dv));
    if (!new_ok) fused = false;
  }
  if (kp.kv_fp8) {
    // e4m3 pool: tuned GQA kernel with head_dim 128 only; the append is its own (quantising) launch
    B200_REQUIRE(mma && !is_mla && a->D == 128 && !cascade, B200_ERR_UNSUPPORTED,
                 "decode_attn: the FP8 KV cache needs the tuned GQA kernel with head_dim 128 (D=%d Dv=%d), no cascade", a->D, a->Dv);
    fused = false;
    if (write_kv) {
      if (a->dtype == B200_BF16)
        set_kv_fp8_kernel<true><<<dim3(a->B, a->Hkv), 128, 0, st>>>((uint8_t*)a->k_buf, a->kbuf_stride_t, a->kbuf_stride_h, (uint8_t*)a->v_buf,
            a->vbuf_stride_t, a->vbuf_stride_h, (const uint16_t*)a->k_new, a->knew_stride_b, a->knew_stride_h, (const uint16_t*)a->v_new,
            a->vnew_stride_b, a->vnew_stride_h, a->out_cache_loc, (a->meta_i32_mask >> 2) & 1, a->D, a->Dv, kp.k_scale, kp.v_scale);
      else
        set_kv_fp8_kernel<false><<<dim3(a->B, a->Hkv), 128, 0, st>>>((uint8_t*)a->k_buf, a->kbuf_stride_t, a->kbuf_stride_h, (uint8_t*)a->v_buf,
            a->vbuf_stride_t, a->vbuf_stride_h, (const uint16_t*)a->k_new, a->knew_stride_b, a->knew_stride_h, (const uint16_t*)a->v_new,
            a->vnew_stride_b, a->vnew_stride_h, a->out_cache_loc, (a->meta_i32_mask >> 2) & 1, a->D, a->Dv, kp.k_scale, kp.v_scale);
      if (int rc = launched("set_kv_fp8_kernel launch")) return rc;
    }
  } else if (write_kv && !fused) {
    set_kv_kernel<<<dim3(a->B, a->Hkv), 128, 0, st>>>(
        (uint16_t*)a->k_buf, a->kbuf_stride_t, a->kbuf_stride_h, (uint16_t*)a->v_buf, a->vbuf_stride_t, a->vbuf_stride_h,
        (const uint16_t*)a->k_new, a->knew_stride_b, a->knew_stride_h, is_mla ? nullptr : (const uint16_t*)a->v_new,
        a->vnew_stride_b, a->vnew_stride_h, a->out_cache_loc, (a->meta_i32_mask >> 2) & 1, a->D, a->Dv,
        (const uint16_t*)a->k_rope_new, a->krope_stride_b, a->D - a->Dv);
    if (int rc = launched("set_kv_kernel launch")) return rc;
  }
  kp.fused_new = fused ? 1 : 0;

  if (mla && !mma) {
    const size_t need = mla_workspace_bytes();
    B200_REQUIRE(a->workspace && a->workspace_bytes >= need && ((uintptr_t)a->workspace & 15) == 0, B200_ERR_ARG,
                 "decode_attn: workspace too small or misaligned (need %zu bytes, have %zu)", need, a->workspace_bytes);
    kp.num_splits = 1;
    kp.part_o = (float*)a->workspace;
    kp.part_lse = kp.part_o + (size_t)attn_grid() * 2 * 16 * a->Dv;
    for (int b0 = 0; b0 < a->B; b0 += kMaxBatchTable) {
      DecodeKParams c = kp;
      c.B = a->B - b0 < kMaxBatchTable ? a->B - b0 : kMaxBatchTable;
      c.q = kp.q + (int64_t)b0 * kp.q_sb;
      if (kp.k_new) c.k_new = kp.k_new + (int64_t)b0 * kp.kn_sb;
      c.out = kp.out + (int64_t)b0 * kp.Hq * kp.Dv;
      if (kp.q_pe) c.q_pe = kp.q_pe + (int64_t)b0 * kp.qpe_sb;
      if (kp.k_pe_new) c.k_pe_new = kp.k_pe_new + (int64_t)b0 * kp.kpe_sb;
      if (b0 > 0) {
        B200_REQUIRE(kp.rpi != nullptr, B200_ERR_UNSUPPORTED, "decode_attn: B > %d needs req_pool_indices", kMaxBatchTable);
        c.rpi = (const uint8_t*)kp.rpi + (size_t)b0 * ((kp.meta32 & 2) ? 4 : 8);
        c.seq_lens = (const uint8_t*)kp.seq_lens + (size_t)b0 * ((kp.meta32 & 1) ? 4 : 8);
        if (kp.loc) c.loc = (const uint8_t*)kp.loc + (size_t)b0 * ((kp.meta32 & 4) ? 4 : 8);
      }
      const int rc = a->dtype == B200_BF16 ? launch_mla<true>(c, st) : launch_mla<false>(c, st);
      if (rc != 0) return rc;
    }
    return B200_OK;
  }
  if (mma) {
    const size_t need = mma_workspace_bytes(a->Dv);
    B200_REQUIRE(a->workspace && a->workspace_bytes >= need && ((uintptr_t)a->workspace & 15) == 0, B200_ERR_ARG,
                 "decode_attn: workspace too small or misaligned (need %zu bytes, have %zu)", need, a->workspace_bytes);
    kp.num_splits = 1;
    kp.part_o = (float*)a->workspace;
    kp.part_lse = kp.part_o + (size_t)attn_grid() * 2 * (8 * 16) * a->Dv;
    // the kernel keeps a per-CTA copy of the batch in shared memory: larger batches go in chunks
    for (int b0 = 0; b0 < a->B; b0 += kMaxBatchTable) {
      DecodeKParams c = kp;
      c.B = a->B - b0 < kMaxBatchTable ? a->B - b0 : kMaxBatchTable;
      c.q = kp.q + (int64_t)b0 * kp.q_sb;
      if (kp.k_new) c.k_new = kp.k_new + (int64_t)b0 * kp.kn_sb;
      if (kp.v_new) c.v_new = kp.v_new + (int64_t)b0 * kp.vn_sb;
      c.out = kp.out + (int64_t)b0 * kp.Hq * kp.Dv;
      if (b0 > 0) {
        B200_REQUIRE(kp.rpi != nullptr, B200_ERR_UNSUPPORTED, "decode_attn: B > %d needs req_pool_indices", kMaxBatchTable);
        c.rpi = (const uint8_t*)kp.rpi + (size_t)b0 * ((kp.meta32 & 2) ? 4 : 8);
        c.seq_lens = (const uint8_t*)kp.seq_lens + (size_t)b0 * ((kp.meta32 & 1) ? 4 : 8);
        if (kp.loc) c.loc = (const uint8_t*)kp.loc + (size_t)b0 * ((kp.meta32 & 4) ? 4 : 8);
      }
      if (kp.mq.q) {
        c.mq.q = kp.mq.q + (int64_t)b0 * kp.mq.q_sm;
        c.mq.s = kp.mq.s + (int64_t)b0 * kp.mq.s_sm;
      }
      const int rc = kp.kv_fp8 ? (a->dtype == B200_BF16 ? dispatch_mma_kv8<true>(c, hpc, st) : dispatch_mma_kv8<false>(c, hpc, st))
                               : (a->dtype == B200_BF16 ? dispatch_mma<true>(c, hpc, st) : dispatch_mma<false>(c, hpc, st));
      if (rc != 0) return rc;
    }
    return B200_OK;
  }
  const int S = b200_decode_attn_num_splits(a->B, a->Hq, a->Hkv, a->D, a->Dv, a->max_seq_len, a->num_splits);
  kp.num_splits = S;
  kp.part_o = nullptr; kp.part_lse = nullptr;
  if (S > 1) {
    const size_t need = ((size_t)a->B * a->Hq * S * (a->Dv + 1)) * sizeof(float) + 256;
    B200_REQUIRE(a->workspace && a->workspace_bytes >= need && ((uintptr_t)a->workspace & 15) == 0, B200_ERR_ARG,
                 "decode_attn: workspace too small or misaligned (need %zu bytes, have %zu)", need, a->workspace_bytes);
    kp.part_o = (float*)a->workspace;
    kp.part_lse = kp.part_o + (size_t)a->B * a->Hq * S * a->Dv;
  }
  int rc = a->dtype == B200_BF16 ? launch_generic<true>(kp, st) : launch_generic<false>(kp, st);
  if (rc != 0) return rc;
  if (S > 1) {
    if (a->dtype == B200_BF16)
      decode_merge_kernel<true><<<a->B * a->Hq, 128, 0, st>>>(kp.part_o, kp.part_lse, kp.out, S, a->Dv);
    else
      decode_merge_kernel<false><<<a->B * a->Hq, 128, 0, st>>>(kp.part_o, kp.part_lse, kp.out, S, a->Dv);
    return launched("decode_merge_kernel launch");
  }
  return B200_OK;
}

extern "C" int b200_decode_attn(const b200_decode_attn_params* a, void* stream) { return decode_attn_impl(a, nullptr, stream); }

extern "C" int b200_decode_attn_quant_fp8(const b200_decode_attn_params* a, void* q, int64_t q_sm, float* s, int64_t s_sm,
                                          int64_t s_sg, float eps, float fp8_min, float fp8_max, void* stream) {
  MergeQuant mq;
  mq.q = (uint8_t*)q; mq.q_sm = q_sm; mq.s = s; mq.s_sm = s_sm; mq.s_sg = s_sg;
  mq.eps = eps; mq.fmin_ = fp8_min; mq.fmax_ = fp8_max;
  return decode_attn_impl(a, &mq, stream);
}

// ---------------------------------------------------------------------------------------------------------------
// b200_decode_attn_static: b200_decode_attn for callers that must not allocate (CUDA-graph capture, tight serving
// loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
// The split-KV partial states go into ONE file-scope device array of
// B200_DECODE_ATTN_STATIC_SCRATCH_BYTES (20 MiB; every GPU that loads the library holds it until exit) instead of the
// caller's workspace.  Nothing in that scratch has to start at zero: every partial the merge reads is written by the
// same call (decode_merge_part_kernel / decode_merge_kernel only visit slots the attention kernel of this call filled).
// All calls share the array, so two calls that overlap on the GPU (different streams, or graphs replayed concurrently)
// race; calls on one stream are ordered and safe.  No cudaMalloc / cudaFree / allocating routine on this path.
// ---------------------------------------------------------------------------------------------------------------
namespace {

static __device__ __align__(256) unsigned char g_decode_attn_static_scratch[B200_DECODE_ATTN_STATIC_SCRATCH_BYTES];

constexpr int kMaxStaticDevices = 64;

// The array's name is a host-side stand-in in host code, not a device address: ask the runtime for the address, once
// per device, with that device current.
int decode_attn_static_scratch(void** out) {
  static void* ptrs[kMaxStaticDevices] = {};
  static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  B200_REQUIRE(dev >= 0 && dev < kMaxStaticDevices, B200_ERR_UNSUPPORTED, "decode_attn_static: device ordinal %d out of range", dev);
  pthread_mutex_lock(&mu);
  void* ptr = ptrs[dev];
  cudaError_t e = cudaSuccess;
  if (!ptr) {
    e = cudaGetSymbolAddress(&ptr, g_decode_attn_static_scratch);
    if (e == cudaSuccess) ptrs[dev] = ptr;
  }
  pthread_mutex_unlock(&mu);
  if (e != cudaSuccess) return check_cuda(e, "cudaGetSymbolAddress(decode_attn static scratch)");
  *out = ptr;
  return B200_OK;
}

}  // namespace

extern "C" int b200_decode_attn_static(const b200_decode_attn_params* a, void* stream) {
  B200_REQUIRE(a != nullptr, B200_ERR_ARG, "decode_attn_static: null params");
  B200_REQUIRE(a->B >= 0 && a->Hq > 0 && a->Hkv > 0 && a->D > 0 && a->Dv > 0, B200_ERR_ARG, "decode_attn_static: bad sizes");
  const int S = a->B > 0 ? b200_decode_attn_num_splits(a->B, a->Hq, a->Hkv, a->D, a->Dv, a->max_seq_len, a->num_splits) : 1;
  const size_t need = b200_decode_attn_workspace_size(a->B, a->Hq, a->Dv, S);
  B200_REQUIRE(need <= (size_t)B200_DECODE_ATTN_STATIC_SCRATCH_BYTES, B200_ERR_ARG,
               "decode_attn_static: B=%d Hq=%d Dv=%d with %d KV splits needs %zu bytes of scratch, the static array holds %zu; "
               "use b200_decode_attn with a caller-provided workspace",
               a->B, a->Hq, a->Dv, S, need, (size_t)B200_DECODE_ATTN_STATIC_SCRATCH_BYTES);
  void* scratch = nullptr;
  if (int rc = decode_attn_static_scratch(&scratch)) return rc;
  b200_decode_attn_params p = *a;
  p.workspace = scratch;
  p.workspace_bytes = (size_t)B200_DECODE_ATTN_STATIC_SCRATCH_BYTES;
  return decode_attn_impl(&p, nullptr, stream);
}
