// All-reduce(sum) of the row-parallel linears' [M, hidden] half outputs over NVLink peer memory / NVSwitch
// multicast, optionally fused with the residual-add + RMSNorm (+ FP8 group quantisation) that follows it.
//
// Replaces, for messages up to a few MiB, the custom P2P all-reduce the reference routes small messages to
// (sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:203-456, python/sglang/srt/distributed/device_communicators/
// custom_all_reduce.py:148,414-480; call site communication_op.py:11-13 after RowParallelLinear, linear.py:1375-1376)
// and the all-reduce + add + RMSNorm fusion hook (python/sglang/srt/layers/flashinfer_comm_fusion.py:122
// flashinfer_allreduce_add_rmsnorm, layers/communicator.py).  Larger messages and anything this path rejects stay
// on ncclAllReduce (tp.py).
//
// Every rank owns one region of the same layout; every rank can address every region (CUDA IPC mappings, or the
// peer pointers of a symmetric-memory allocation, which also gives the MULTICAST address of the same buffer):
//   [flags: MAX_BLOCKS x 8 uint32, written by PEERS] [call counter | finished-block counter | error word]
//   [in slot 0] [in slot 1] [out stage 0] [out stage 1]
// One kernel per all-reduce, no host synchronisation, CUDA-graph capturable.  Block b of a rank only ever depends
// on block b of its peers, through a per-block flag barrier (st.release.sys into the peer's flag word [b][my rank],
// ld.acquire.sys on my own); all blocks of a call use the same flag values 2n+1 / 2n+2 and the same data slot
// n & 1, where n is ONE call counter per rank in device memory: the last block to finish increments it, and the
// stream order (or griddepcontrol.wait) makes the next kernel read the new value -- so a replayed CUDA graph keeps
// counting and calls of different sizes, grids and kernels can be mixed freely.
//
// Why two slots need no end-of-call barrier: a rank starts call n + 2 (same slot as n) only after its call n + 1
// has completed, i.e. after at least its block 0 passed a barrier of call n + 1 with block 0 of EVERY peer; a peer
// signals in call n + 1 only after its whole call-n kernel has finished (stream order).  So when slot n & 1 is
// overwritten, nobody is still reading call n's data.  The same argument covers the out stages.
//
// Kernels:
//   allreduce_oneshot_kernel  flat message: copy-in, barrier, every rank pulls all W copies and sums in RANK ORDER in
//                             fp32 -> bit-identical results on every rank;
//   allreduce_twoshot_kernel  flat message, reduce-scatter + all-gather over peer loads (2 barriers);
//   allreduce_rows_kernel     row layout (block b owns whole rows of [M, H]): copy-in, barrier, then either
//                               PULL: sum the W copies of the block's rows in rank order (one-shot), or
//                               NVLS: rank c reduces sub-range c of the block's rows with ONE multimem.ld_reduce per
//                                     16 bytes (the switch adds the W copies) and multimem.st's the result into
//                                     every rank's out stage; barrier;
//                             and an epilogue over the block's full rows, read locally: plain copy to `out`, or
//                             residual += x; RMSNorm; optional group-128 e4m3 quantisation (rmsnorm_row.cuh: the same
//                             code, hence the same bits, as rmsnorm_quant_fp8_kernel after a plain all-reduce).
//   allreduce_lamport_kernel  row layout, PUSH protocol without flags or barriers (one NVLink one-way latency per phase
//                             instead of flag round trips): the sender stores its 16-byte vectors straight into the
//                             receivers' buffers, the receiver polls the DATA -- every 32-bit word of a receive buffer
//                             holds a sentinel (0x80000000, a pattern the sender never transmits: such a word is sent as
//                             +0.0/+0.0, which leaves every fp32 sum unchanged) until the payload lands.  Three receive
//                             buffers rotate by call number; call n re-arms the buffer of call n - 1 (a peer can only
//                             write into it again in call n + 2, which it cannot start before it has received this
//                             rank's call n + 1 data, sent after call n has completed).  One-shot: every rank receives
//                             all W copies and sums them in rank order (bit-identical to the pull kernels).  Two-shot:
//                             rank c receives sub-range c of every block's rows, reduces it in rank order and pushes the
//                             result to everybody (one multimem.st where the buffer has a multicast mapping).  Same
//                             epilogues as the rows kernel.
// A spin that lasts longer than ~20 s sets the error word and lets the kernel finish (wrong data, no hang); the
// host polls it with b200_ar_error() (tp.py does so once per step).
#include <stdlib.h>
#include <string.h>

#include "common.cuh"
#include "rmsnorm_row.cuh"

namespace b200 {

constexpr int AR_MAX_BLOCKS = 512;
constexpr int AR_MAX_RANKS = 8;
constexpr int AR_THREADS = 512;       // flat kernels
constexpr int AR_FLAT_BLOCKS = 128;   // flat kernels
constexpr int AR_ROW_THREADS = 256;   // rows kernel (= the RMSNorm row layout)
constexpr size_t AR_FLAG_BYTES = (size_t)AR_MAX_BLOCKS * AR_MAX_RANKS * 4;  // 16 KB
constexpr size_t AR_CALLS_OFF = AR_FLAG_BYTES;        // uint32: calls completed by this rank
constexpr size_t AR_DONE_OFF = AR_CALLS_OFF + 64;     // uint32: blocks of the running call that have finished
constexpr size_t AR_ERR_OFF = AR_DONE_OFF + 64;       // uint32: a peer never arrived
constexpr size_t AR_LCALLS_OFF = AR_ERR_OFF + 64;     // uint32: push-protocol calls completed by this rank
constexpr size_t AR_LDONE_OFF = AR_LCALLS_OFF + 64;   // uint32: blocks of the running push-protocol call that have finished
constexpr size_t AR_LPREV_OFF = AR_LDONE_OFF + 64;    // uint32: 16-byte vectors of the receive buffer the last push call used
constexpr size_t AR_DATA_OFF = 32768;
static_assert(AR_LPREV_OFF + 64 <= AR_DATA_OFF, "region header");
constexpr size_t AR_LAM_PAD = 65536;                  // AR_MAX_BLOCKS x (AR_MAX_RANKS - 1) vectors of sub-range rounding
constexpr uint32_t AR_SENTINEL = 0x80000000u;         // never transmitted; receive buffers are armed with it

struct ArParams {
  uint8_t* base[AR_MAX_RANKS];  // region of every rank (own = local pointer)
  uint8_t* mc;                  // multicast address of the same region (NVLS) or nullptr
  int rank, world;
  size_t slot_bytes;
  size_t lam_off, lam_bytes;    // three push-protocol receive buffers of lam_bytes each at base + lam_off
  const uint4* in;
  uint4* out;
  int64_t nvec;  // 16-byte vectors
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// peer / staged data is read around the L1 (a line of the same slot from two calls ago must not be served from it)
__device__ __forceinline__ uint4 ld_volatile_v4(const uint4* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
// NVSwitch multicast (NVLS): one load returns the fp32-accumulated sum of the same 16 bytes of every rank's copy;
// one store lands in every rank's copy
template <bool BF16>
__device__ __forceinline__ uint4 multimem_ld_reduce(const void* mc_addr) {
  uint4 v;
  if constexpr (BF16)
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc_addr) : "memory");
  else
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0, %1, %2, %3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(mc_addr) : "memory");
  return v;
}
__device__ __forceinline__ void multimem_st(void* mc_addr, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(mc_addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

template <bool BF16>
__device__ __forceinline__ void add8(float (&acc)[8], const uint4& v) {
  const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float2 f = unpack_half2<BF16>(w[i]);
    acc[2 * i] += f.x;
    acc[2 * i + 1] += f.y;
  }
}
template <bool BF16>
__device__ __forceinline__ uint4 pack8(const float (&acc)[8]) {
  return make_uint4(pack_half2<BF16>(acc[0], acc[1]), pack_half2<BF16>(acc[2], acc[3]), pack_half2<BF16>(acc[4], acc[5]),
                    pack_half2<BF16>(acc[6], acc[7]));
}

// this rank's call number (see the file header); read after griddepcontrol.wait
__device__ __forceinline__ uint32_t ar_call_number(const uint8_t* mine) {
  return *reinterpret_cast<const volatile uint32_t*>(mine + AR_CALLS_OFF);
}
// last block out bumps the call number
__device__ __forceinline__ void ar_finish(uint8_t* mine, uint32_t call) {
  __syncthreads();
  if (threadIdx.x == 0) {
    uint32_t* done = reinterpret_cast<uint32_t*>(mine + AR_DONE_OFF);
    __threadfence();
    if (atomicAdd(done, 1u) == gridDim.x - 1) {
      *done = 0;
      *reinterpret_cast<volatile uint32_t*>(mine + AR_CALLS_OFF) = call + 1;
    }
  }
}

// block b of this rank <-> block b of every peer; `value` is the same on every rank and grows with every barrier
template <int WORLD>
__device__ __forceinline__ void ar_block_barrier(const ArParams& p, uint8_t* mine, int b, uint32_t value) {
  __syncthreads();  // (the release below orders the whole block's stores: cumulativity through the barrier)
  if (threadIdx.x < WORLD && threadIdx.x != p.rank) {
    const int peer = threadIdx.x;
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(p.base[peer]) + b * AR_MAX_RANKS + p.rank, value);
    const uint32_t* flag = reinterpret_cast<const uint32_t*>(mine) + b * AR_MAX_RANKS + peer;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(flag) - value) < 0) {
      if (clock64() - t0 > 40000000000LL) {  // ~20 s: a peer is gone; report instead of hanging the GPU
        *reinterpret_cast<uint32_t*>(mine + AR_ERR_OFF) = 1;
        break;
      }
    }
  }
  __syncthreads();
}

template <bool BF16, int WORLD>
__global__ void __launch_bounds__(AR_THREADS) allreduce_oneshot_kernel(const __grid_constant__ ArParams p) {
  const int b = blockIdx.x;
  uint8_t* mine = p.base[p.rank];
  const uint32_t call = ar_call_number(mine);
  const int slot = call & 1;
  const int64_t per = (p.nvec + gridDim.x - 1) / gridDim.x;
  const int64_t v0 = (int64_t)b * per, v1 = v0 + per < p.nvec ? v0 + per : p.nvec;

  // 1. my chunk -> my slot
  uint4* my_slot = reinterpret_cast<uint4*>(mine + AR_DATA_OFF + slot * p.slot_bytes);
  for (int64_t i = v0 + threadIdx.x; i < v1; i += AR_THREADS) my_slot[i] = p.in[i];
  // 2. per-block barrier with the same block of every peer
  ar_block_barrier<WORLD>(p, mine, b, 2 * call + 1);
  // 3. reduce chunk b of every rank in rank order
  const uint4* src[WORLD];
#pragma unroll
  for (int r = 0; r < WORLD; ++r) src[r] = reinterpret_cast<const uint4*>(p.base[r] + AR_DATA_OFF + slot * p.slot_bytes);
  for (int64_t i = v0 + threadIdx.x; i < v1; i += AR_THREADS) {
    uint4 v[WORLD];
#pragma unroll
    for (int r = 0; r < WORLD; ++r) v[r] = ld_volatile_v4(src[r] + i);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < WORLD; ++r) add8<BF16>(acc, v[r]);
    p.out[i] = pack8<BF16>(acc);
  }
  ar_finish(mine, call);
}

// Two-shot variant for more ranks / larger messages (custom_all_reduce.cuh:300-400 picks the same way): rank r
// reduces only chunk r (1/W of the message) from every rank, publishes it in place in its own slot, and all ranks
// gather the W reduced chunks: 2 (W-1)/W of the message crosses NVLink per rank instead of (W-1) times the message.
// Block b owns sub-range b of EVERY chunk in all three phases, so the per-block flag barrier still covers exactly
// the data the block depends on.
template <bool BF16, int WORLD>
__global__ void __launch_bounds__(AR_THREADS) allreduce_twoshot_kernel(const __grid_constant__ ArParams p) {
  const int b = blockIdx.x;
  uint8_t* mine = p.base[p.rank];
  const uint32_t call = ar_call_number(mine);
  const int slot = call & 1;
  const int64_t cv = (p.nvec + WORLD - 1) / WORLD;          // vectors per chunk
  const int64_t sv = (cv + gridDim.x - 1) / gridDim.x;      // vectors per (chunk, block)
  const int64_t j0 = (int64_t)b * sv, j1 = j0 + sv < cv ? j0 + sv : cv;
  uint4* my_slot = reinterpret_cast<uint4*>(mine + AR_DATA_OFF + slot * p.slot_bytes);
  // 1. my input, sub-range b of every chunk -> my slot
#pragma unroll
  for (int c = 0; c < WORLD; ++c)
    for (int64_t j = j0 + threadIdx.x; j < j1; j += AR_THREADS) {
      const int64_t i = c * cv + j;
      if (i < p.nvec) my_slot[i] = p.in[i];
    }
  ar_block_barrier<WORLD>(p, mine, b, 2 * call + 1);
  // 2. reduce chunk `rank` in rank order, publish it in place
  const uint4* src[WORLD];
#pragma unroll
  for (int r = 0; r < WORLD; ++r) src[r] = reinterpret_cast<const uint4*>(p.base[r] + AR_DATA_OFF + slot * p.slot_bytes);
  for (int64_t j = j0 + threadIdx.x; j < j1; j += AR_THREADS) {
    const int64_t i = p.rank * cv + j;
    if (i >= p.nvec) break;
    uint4 v[WORLD];
#pragma unroll
    for (int r = 0; r < WORLD; ++r) v[r] = ld_volatile_v4(src[r] + i);
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < WORLD; ++r) add8<BF16>(acc, v[r]);
    const uint4 o = pack8<BF16>(acc);
    my_slot[i] = o;
    p.out[i] = o;
  }
  ar_block_barrier<WORLD>(p, mine, b, 2 * call + 2);
  // 3. gather the other ranks' reduced chunks: all W - 1 remote loads of a thread are issued before the first store
  for (int64_t j = j0 + threadIdx.x; j < j1; j += AR_THREADS) {
    uint4 v[WORLD];
#pragma unroll
    for (int c = 0; c < WORLD; ++c) {
      const int64_t i = c * cv + j;
      if (c != p.rank && i < p.nvec) v[c] = ld_volatile_v4(src[c] + i);
    }
#pragma unroll
    for (int c = 0; c < WORLD; ++c) {
      const int64_t i = c * cv + j;
      if (c != p.rank && i < p.nvec) p.out[i] = v[c];
    }
  }
  ar_finish(mine, call);
}

// ---------------------------------------------------------------------------------------------------------------
// row layout: the message is [M rows x rowvec vectors]; block b owns rows [b * rpb, (b + 1) * rpb)
// ---------------------------------------------------------------------------------------------------------------
struct ArRows {
  int M, rowvec, rpb;
  int epi;  // 0: out = sum;  1: fused add + RMSNorm (+ quant) epilogue (RowNormArgs)
};

template <bool BF16, int WORLD, bool NVLS, int NV>
__global__ void __launch_bounds__(AR_ROW_THREADS) allreduce_rows_kernel(const __grid_constant__ ArParams p, const ArRows r,
                                                                        const RowNormArgs e) {
  griddep_wait();
  griddep_launch();
  __shared__ float red[8];
  const int b = blockIdx.x;
  uint8_t* mine = p.base[p.rank];
  const uint32_t call = ar_call_number(mine);
  const int slot = call & 1;
  const int row0 = b * r.rpb, row1 = min(r.M, row0 + r.rpb);
  const int64_t v0 = (int64_t)row0 * r.rowvec;
  const int64_t v1 = min((int64_t)row1 * r.rowvec, p.nvec);
  const size_t slot_off = AR_DATA_OFF + (size_t)slot * p.slot_bytes;
  const size_t stage_off = AR_DATA_OFF + (size_t)(2 + slot) * p.slot_bytes;

  // 0. my rows -> my slot
  uint4* my_slot = reinterpret_cast<uint4*>(mine + slot_off);
  for (int64_t i = v0 + threadIdx.x; i < v1; i += AR_ROW_THREADS) my_slot[i] = p.in[i];
  ar_block_barrier<WORLD>(p, mine, b, 2 * call + 1);

  const uint4* x_src;  // where the epilogue finds the summed rows (NVLS), or nullptr (PULL: summed on the fly)
  if constexpr (NVLS) {
    // 1. sub-range `rank` of the block's rows: the switch reduces, the result is multicast into every out stage
    const int64_t len = v1 - v0;
    const int64_t sub = (len + WORLD - 1) / WORLD;
    const int64_t a0 = v0 + p.rank * sub, a1 = min(a0 + sub, v1);
    const uint4* mc_in = reinterpret_cast<const uint4*>(p.mc + slot_off);
    uint4* mc_out = reinterpret_cast<uint4*>(p.mc + stage_off);
    constexpr int U = 4;  // loads in flight per thread
    for (int64_t i = a0 + threadIdx.x; i < a1; i += U * AR_ROW_THREADS) {
      uint4 v[U];
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + u * AR_ROW_THREADS < a1) v[u] = multimem_ld_reduce<BF16>(mc_in + i + u * AR_ROW_THREADS);
#pragma unroll
      for (int u = 0; u < U; ++u)
        if (i + u * AR_ROW_THREADS < a1) multimem_st(mc_out + i + u * AR_ROW_THREADS, v[u]);
    }
    ar_block_barrier<WORLD>(p, mine, b, 2 * call + 2);
    x_src = reinterpret_cast<const uint4*>(mine + stage_off);
  } else {
    x_src = nullptr;
  }
  const uint4* src[WORLD];
#pragma unroll
  for (int q = 0; q < WORLD; ++q) src[q] = reinterpret_cast<const uint4*>(p.base[q] + slot_off);
  auto summed = [&](int64_t i) -> uint4 {
    if constexpr (NVLS) {
      return ld_volatile_v4(x_src + i);
    } else {
      uint4 v[WORLD];
#pragma unroll
      for (int q = 0; q < WORLD; ++q) v[q] = ld_volatile_v4(src[q] + i);
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int q = 0; q < WORLD; ++q) add8<BF16>(acc, v[q]);
      return pack8<BF16>(acc);
    }
  };

  // 2. epilogue over the block's full rows
  if (r.epi == 0) {
    for (int64_t i = v0 + threadIdx.x; i < v1; i += AR_ROW_THREADS) p.out[i] = summed(i);
  } else {
    for (int m = row0; m < row1; ++m) {
      const int64_t rv = (int64_t)m * r.rowvec;
      rmsnorm_quant_row<BF16, NV>(e, m, [&](int vec) { return summed(rv + vec); }, red);
    }
  }
  ar_finish(mine, call);
}


// ---------------------------------------------------------------------------------------------------------------
// push protocol (see the file header): no flags, the receiver polls the payload
// ---------------------------------------------------------------------------------------------------------------
__device__ __forceinline__ void st_volatile_v4(uint4* p, const uint4& v) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
__device__ __forceinline__ uint4 lam_clean(uint4 v) {   // a word that equals the sentinel travels as +0.0 / +0.0
  v.x = v.x == AR_SENTINEL ? 0u : v.x;
  v.y = v.y == AR_SENTINEL ? 0u : v.y;
  v.z = v.z == AR_SENTINEL ? 0u : v.z;
  v.w = v.w == AR_SENTINEL ? 0u : v.w;
  return v;
}
__device__ __forceinline__ bool lam_landed(const uint4& v) {
  return v.x != AR_SENTINEL && v.y != AR_SENTINEL && v.z != AR_SENTINEL && v.w != AR_SENTINEL;
}
__device__ __forceinline__ bool lam_timed_out(long long t0, uint8_t* mine) {
  if (clock64() - t0 > 40000000000LL) {  // ~20 s: a peer is gone; report instead of hanging the GPU
    *reinterpret_cast<uint32_t*>(mine + AR_ERR_OFF) = 1;
    return true;
  }
  return false;
}
__device__ __forceinline__ uint4 lam_poll(const uint4* p, uint8_t* mine) {
  uint4 v = ld_volatile_v4(p);
  if (lam_landed(v)) return v;
  const long long t0 = clock64();
  do {
    v = ld_volatile_v4(p);
  } while (!lam_landed(v) && !lam_timed_out(t0, mine));
  return v;
}

template <bool BF16, int WORLD, bool TWOSHOT, int NV>
__global__ void __launch_bounds__(AR_ROW_THREADS) allreduce_lamport_kernel(const __grid_constant__ ArParams p, const ArRows r,
                                                                           const RowNormArgs e) {
  griddep_wait();
  griddep_launch();
  __shared__ float red[8];
  const int b = blockIdx.x, t = threadIdx.x;
  uint8_t* mine = p.base[p.rank];
  const uint32_t call = *reinterpret_cast<const volatile uint32_t*>(mine + AR_LCALLS_OFF);
  const uint32_t prev_vec = *reinterpret_cast<const volatile uint32_t*>(mine + AR_LPREV_OFF);
  const size_t buf_off = p.lam_off + (size_t)(call % 3u) * p.lam_bytes;
  const size_t old_off = p.lam_off + (size_t)((call + 2u) % 3u) * p.lam_bytes;   // the buffer of call - 1
  const int row0 = b * r.rpb, row1 = min(r.M, row0 + r.rpb);
  const int64_t v0 = (int64_t)row0 * r.rowvec;
  const int64_t v1 = min((int64_t)row1 * r.rowvec, p.nvec);
  // two-shot geometry: sub-range c of the block's vectors belongs to rank c
  const int64_t sub = ((int64_t)r.rpb * r.rowvec + WORLD - 1) / WORLD;
  const int64_t a1s = (int64_t)gridDim.x * sub;            // vectors one source sends to one owner (all blocks)
  const int64_t red_off = TWOSHOT ? (int64_t)WORLD * a1s : 0;   // where the reduced message starts in the buffer
  uint32_t extent;                                          // vectors of the buffer this call writes

  // 1. push
  if constexpr (!TWOSHOT) {
    extent = (uint32_t)(WORLD * p.nvec);
    for (int64_t i = v0 + t; i < v1; i += AR_ROW_THREADS) {
      const uint4 v = lam_clean(p.in[i]);
#pragma unroll
      for (int q = 0; q < WORLD; ++q)
        if (q != p.rank) st_volatile_v4(reinterpret_cast<uint4*>(p.base[q] + buf_off) + (int64_t)p.rank * p.nvec + i, v);
    }
  } else {
    extent = (uint32_t)(red_off + p.nvec);
    for (int64_t i = v0 + t; i < v1; i += AR_ROW_THREADS) {
      const int c = (int)((i - v0) / sub);
      if (c == p.rank) continue;
      const int64_t k = (i - v0) - c * sub;
      st_volatile_v4(reinterpret_cast<uint4*>(p.base[c] + buf_off) + (int64_t)p.rank * a1s + (int64_t)b * sub + k, lam_clean(p.in[i]));
    }
  }
  // 2. re-arm the buffer of the previous call while the payload is in flight
  {
    uint4* old = reinterpret_cast<uint4*>(mine + old_off);
    const uint4 s4 = make_uint4(AR_SENTINEL, AR_SENTINEL, AR_SENTINEL, AR_SENTINEL);
    for (int64_t j = (int64_t)b * AR_ROW_THREADS + t; j < (int64_t)prev_vec; j += (int64_t)gridDim.x * AR_ROW_THREADS) old[j] = s4;
  }
  const uint4* rx = reinterpret_cast<const uint4*>(mine + buf_off);
  // 3. (two-shot) reduce my sub-range in rank order, push the result to everybody
  if constexpr (TWOSHOT) {
    const int64_t a0 = v0 + p.rank * sub, a1 = min(a0 + sub, v1);
    for (int64_t i = a0 + t; i < a1; i += AR_ROW_THREADS) {
      const int64_t k = (int64_t)b * sub + (i - a0);
      uint4 v[WORLD];
      bool all = true;
      long long t0 = 0;
      do {
        all = true;
#pragma unroll
        for (int q = 0; q < WORLD; ++q)
          if (q != p.rank) {
            v[q] = ld_volatile_v4(rx + (int64_t)q * a1s + k);
            all = all && lam_landed(v[q]);
          }
        if (!all) {
          if (t0 == 0) t0 = clock64();
          if (lam_timed_out(t0, mine)) break;
        }
      } while (!all);
      // (no v[p.rank] = ...: a dynamically indexed array lives in local memory -- a stack frame in a latency-bound kernel)
      const uint4 own = p.in[i];
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int q = 0; q < WORLD; ++q) add8<BF16>(acc, q == p.rank ? own : v[q]);
      const uint4 o = lam_clean(pack8<BF16>(acc));
      if (p.mc) {
        multimem_st(reinterpret_cast<uint4*>(p.mc + buf_off) + red_off + i, o);
      } else {
#pragma unroll
        for (int q = 0; q < WORLD; ++q) st_volatile_v4(reinterpret_cast<uint4*>(p.base[q] + buf_off) + red_off + i, o);
      }
    }
  }
  // 4. the summed vector i of this block's rows
  auto summed = [&](int64_t i) -> uint4 {
    if constexpr (TWOSHOT) {
      return lam_poll(rx + red_off + i, mine);
    } else {
      uint4 v[WORLD];
      bool all = true;
      long long t0 = 0;
      do {
        all = true;
#pragma unroll
        for (int q = 0; q < WORLD; ++q)
          if (q != p.rank) {
            v[q] = ld_volatile_v4(rx + (int64_t)q * p.nvec + i);
            all = all && lam_landed(v[q]);
          }
        if (!all) {
          if (t0 == 0) t0 = clock64();
          if (lam_timed_out(t0, mine)) break;
        }
      } while (!all);
      // (no v[p.rank] = ...: a dynamically indexed array lives in local memory -- a stack frame in a latency-bound kernel)
      const uint4 own = p.in[i];
      float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int q = 0; q < WORLD; ++q) add8<BF16>(acc, q == p.rank ? own : v[q]);
      return pack8<BF16>(acc);
    }
  };
  if (r.epi == 0) {
    for (int64_t i = v0 + t; i < v1; i += AR_ROW_THREADS) p.out[i] = summed(i);
  } else {
    for (int m = row0; m < row1; ++m) {
      const int64_t rv = (int64_t)m * r.rowvec;
      rmsnorm_quant_row<BF16, NV>(e, m, [&](int vec) { return summed(rv + vec); }, red);
    }
  }
  // last block out: next call number, and how much of this buffer the next call has to re-arm
  __syncthreads();
  if (t == 0) {
    uint32_t* done = reinterpret_cast<uint32_t*>(mine + AR_LDONE_OFF);
    __threadfence();
    if (atomicAdd(done, 1u) == gridDim.x - 1) {
      *done = 0;
      *reinterpret_cast<volatile uint32_t*>(mine + AR_LPREV_OFF) = extent;
      *reinterpret_cast<volatile uint32_t*>(mine + AR_LCALLS_OFF) = call + 1;
    }
  }
}

__global__ void ar_arm_kernel(uint4* buf, int64_t nvec) {
  const uint4 s4 = make_uint4(AR_SENTINEL, AR_SENTINEL, AR_SENTINEL, AR_SENTINEL);
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < nvec; i += (int64_t)gridDim.x * blockDim.x) buf[i] = s4;
}

struct ArContext {
  ArParams p;
  void* own;  // cudaMalloc'ed by b200_ar_alloc (freed by destroy) or nullptr when the caller owns the regions
  void* opened[AR_MAX_RANKS];
  int device;
  int mode;   // -1 = measured policy (b200_ar_mode); 1 .. 6 force a kernel (must be the same on every rank)
};

}  // namespace b200

using namespace b200;

extern "C" size_t b200_ar_region_bytes(size_t max_message_bytes) {
  const size_t slot = (max_message_bytes + 255) & ~(size_t)255;
  return AR_DATA_OFF + 4 * slot + 3 * (2 * slot + AR_LAM_PAD);   // 2 in slots, 2 out stages, 3 push-protocol receive buffers
}

namespace {
// arm this rank's three receive buffers with the sentinel (the caller synchronises every rank before the first call)
int ar_arm(ArContext* c) {
  c->p.lam_off = AR_DATA_OFF + 4 * c->p.slot_bytes;
  c->p.lam_bytes = 2 * c->p.slot_bytes + AR_LAM_PAD;   // two-shot: the scattered message + the reduced one (+ per-block rounding)
  ar_arm_kernel<<<296, 512>>>(reinterpret_cast<uint4*>(c->p.base[c->p.rank] + c->p.lam_off), (int64_t)(3 * c->p.lam_bytes / 16));
  if (int rc = launched("ar_arm_kernel launch")) return rc;
  return check_cuda(cudaDeviceSynchronize(), "ar_arm_kernel");
}
}  // namespace

// allocate this rank's region (zeroed) and export its 64-byte IPC handle
extern "C" int b200_ar_alloc(size_t max_message_bytes, void** region_out, void* ipc_handle_out64) {
  B200_REQUIRE(region_out && ipc_handle_out64 && max_message_bytes > 0, B200_ERR_ARG, "ar_alloc: bad argument");
  static_assert(sizeof(cudaIpcMemHandle_t) == 64, "IPC handle size");
  void* ptr = nullptr;
  const size_t bytes = b200_ar_region_bytes(max_message_bytes);
  B200_CUDA(cudaMalloc(&ptr, bytes));
  B200_CUDA(cudaMemset(ptr, 0, bytes));
  B200_CUDA(cudaDeviceSynchronize());
  cudaIpcMemHandle_t h;
  const cudaError_t e = cudaIpcGetMemHandle(&h, ptr);
  if (e != cudaSuccess) {
    cudaFree(ptr);
    return check_cuda(e, "cudaIpcGetMemHandle");
  }
  memcpy(ipc_handle_out64, &h, 64);
  *region_out = ptr;
  return B200_OK;
}

extern "C" void b200_ar_free_region(void* region) {
  if (region) cudaFree(region);
}

// handles = world x 64 bytes in rank order (own entry ignored); returns an opaque context that owns `own_region`
extern "C" int b200_ar_init(void* own_region, const void* handles, int rank, int world, size_t max_message_bytes,
                            void** ctx_out) {
  B200_REQUIRE(own_region && handles && ctx_out && world >= 2 && world <= AR_MAX_RANKS && rank >= 0 && rank < world, B200_ERR_ARG,
               "ar_init: need 2 <= world <= %d", AR_MAX_RANKS);
  ArContext* c = new ArContext();
  memset(c, 0, sizeof(*c));
  c->own = own_region;
  c->mode = -1;
  cudaGetDevice(&c->device);
  c->p.rank = rank;
  c->p.world = world;
  c->p.mc = nullptr;
  c->p.slot_bytes = (max_message_bytes + 255) & ~(size_t)255;
  for (int r = 0; r < world; ++r) {
    if (r == rank) {
      c->p.base[r] = (uint8_t*)own_region;
      continue;
    }
    cudaIpcMemHandle_t h;
    memcpy(&h, (const uint8_t*)handles + (size_t)r * 64, 64);
    void* peer = nullptr;
    const cudaError_t e = cudaIpcOpenMemHandle(&peer, h, cudaIpcMemLazyEnablePeerAccess);
    if (e != cudaSuccess) {
      set_error("ar_init: cudaIpcOpenMemHandle(rank %d) failed: %s", r, cudaGetErrorString(e));
      cudaGetLastError();
      for (int q = 0; q < r; ++q)
        if (c->opened[q]) cudaIpcCloseMemHandle(c->opened[q]);
      delete c;
      return B200_ERR_CUDA;
    }
    c->opened[r] = peer;
    c->p.base[r] = (uint8_t*)peer;
  }
  if (int rc = ar_arm(c)) {
    for (int q = 0; q < AR_MAX_RANKS; ++q)
      if (c->opened[q]) cudaIpcCloseMemHandle(c->opened[q]);
    delete c;
    return rc;
  }
  *ctx_out = c;
  return B200_OK;
}

// The caller already holds mappings of every rank's region (e.g. a torch symmetric-memory rendezvous): `regions` is a
// HOST array of `world` device pointers in rank order, each >= b200_ar_region_bytes(max_message_bytes) bytes and
// zero-filled before the first call; `multicast` is the NVSwitch multicast address of the same buffer or NULL.
// The context does not own the regions.
extern "C" int b200_ar_init_ptrs(const uint64_t* regions, uint64_t multicast, int rank, int world, size_t max_message_bytes,
                                 void** ctx_out) {
  B200_REQUIRE(regions && ctx_out && world >= 2 && world <= AR_MAX_RANKS && rank >= 0 && rank < world, B200_ERR_ARG,
               "ar_init_ptrs: need 2 <= world <= %d", AR_MAX_RANKS);
  for (int r = 0; r < world; ++r)
    B200_REQUIRE(regions[r] != 0 && (regions[r] & 255) == 0, B200_ERR_ARG, "ar_init_ptrs: region %d is null or misaligned", r);
  ArContext* c = new ArContext();
  memset(c, 0, sizeof(*c));
  c->mode = -1;
  cudaGetDevice(&c->device);
  c->p.rank = rank;
  c->p.world = world;
  c->p.mc = (uint8_t*)(uintptr_t)multicast;
  c->p.slot_bytes = (max_message_bytes + 255) & ~(size_t)255;
  for (int r = 0; r < world; ++r) c->p.base[r] = (uint8_t*)(uintptr_t)regions[r];
  if (int rc = ar_arm(c)) {
    delete c;
    return rc;
  }
  *ctx_out = c;
  return B200_OK;
}

extern "C" int b200_ar_has_multicast(void* ctx) { return ctx && ((ArContext*)ctx)->p.mc ? 1 : 0; }

extern "C" int b200_ar_set_mode(void* ctx, int mode) {
  B200_REQUIRE(ctx && mode >= -1 && mode <= 6, B200_ERR_ARG, "ar_set_mode: mode must be -1 (policy) or 1..6");
  ((ArContext*)ctx)->mode = mode;
  return B200_OK;
}

extern "C" int b200_ar_error(void* ctx) {
  ArContext* c = (ArContext*)ctx;
  if (!c) return -1;
  uint32_t v = 0;
  if (cudaMemcpy(&v, c->p.base[c->p.rank] + AR_ERR_OFF, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return (int)v;
}

extern "C" void b200_ar_destroy(void* ctx) {
  ArContext* c = (ArContext*)ctx;
  if (!c) return;
  for (int r = 0; r < AR_MAX_RANKS; ++r)
    if (c->opened[r]) cudaIpcCloseMemHandle(c->opened[r]);
  if (c->own) cudaFree(c->own);
  delete c;
}

// Which kernel is fastest, measured on 2 / 4 / 8 B200 of one NVSwitch box (tools/ar_check.py, AR_SWEEP=1;
// profiles/r02_allreduce.md; us, plain all-reduce in a graph of 50 calls):
//   1 one-shot pull (flat), 2 two-shot pull (flat), 3 rows kernel with NVLS multicast, 4 rows kernel one-shot pull,
//   5 push protocol one-shot, 6 push protocol two-shot.
// The push protocol (no flag round trips) wins at every size the receive buffers hold:
//   2 ranks: push-1 1 MB 8.0 / 4 MB 10.7 (pull one-shot 14.3 / 20.5, NCCL 21 / 44);
//   4 ranks: push-1 64 KB 5.5, 512 KB 7.2, 1 MB 9.4; push-2 1 MB 8.8, 2 MB 10.3, 4 MB 20.0, 8 MB 26.9
//            (pull one-shot 12.1 / 16.7 / 19.1, two-shot 27.8 @ 2 MB, 42.2 @ 8 MB; NCCL 22 .. 53);
//   8 ranks: push-1 64 KB 7.1, 256 KB 8.0, 512 KB 10.8, 1 MB 16.5; push-2 512 KB 11.6, 1 MB 12.1, 2 MB 13.0, 4 MB 16.7,
//            8 MB 30.5 (NVLS 20.7 .. 35.1 @ 4 MB, 59.1 @ 8 MB; pull two-shot 36.1 @ 4 MB; NCCL 33 .. 78).
// The reference makes the same kind of split (custom_all_reduce.cuh: world 2 one stage; 4: < 512 KB; 8: < 256 KB one
// stage, else two).  `fused`: the kernel also runs the add + RMSNorm (+ quant) epilogue -- real rows, so the one-shot
// form (more, smaller blocks) stays ahead a little longer.
static int ar_policy(int world, size_t bytes, bool fused) {
  if (world < 2 || world > AR_MAX_RANKS) return 0;
  if (world == 2) return 5;
  (void)fused;   // 4 ranks, fused 128x5120 (1.3 MB): push-1 15.7, push-2 13.5; 256x4096: 17.9 / 13.3 -- same threshold as plain
  if (world <= 4) return bytes <= ((size_t)768 << 10) ? 5 : 6;
  return bytes <= ((size_t)512 << 10) ? 5 : 6;
}
extern "C" int b200_ar_mode2(int world, size_t bytes, int multicast) {
  (void)multicast;
  return ar_policy(world, bytes, false);
}
extern "C" int b200_ar_mode(int world, size_t bytes) { return b200_ar_mode2(world, bytes, 0); }

namespace {

template <bool NVLS, int NV>
int launch_rows_w(const ArParams& p, const ArRows& r, const RowNormArgs& e, int nb, bool bf16, cudaStream_t st) {
#define B200_ARR(W)                                                                                                          \
  case W:                                                                                                                    \
    if (bf16) launch_pdl(allreduce_rows_kernel<true, W, NVLS, NV>, dim3(nb), dim3(AR_ROW_THREADS), 0, st, p, r, e);          \
    else launch_pdl(allreduce_rows_kernel<false, W, NVLS, NV>, dim3(nb), dim3(AR_ROW_THREADS), 0, st, p, r, e);              \
    break;
  switch (p.world) {
    B200_ARR(2) B200_ARR(3) B200_ARR(4) B200_ARR(5) B200_ARR(6) B200_ARR(7) B200_ARR(8)
    default:
      set_error("ar: world %d", p.world);
      return B200_ERR_UNSUPPORTED;
  }
#undef B200_ARR
  return launched("allreduce_rows_kernel launch");
}

// rows kernel: M rows of rowvec vectors (the last row may be short when epi == 0)
int launch_rows(const ArContext* c, ArParams p, int M, int rowvec, int epi, const RowNormArgs& e, bool nvls, bool bf16,
                cudaStream_t st) {
  // enough blocks that a block's share of a reduce-scatter (1 / world of its rows) still fills its 256 threads;
  // a function of the sizes only, so every rank launches the same grid
  int64_t per_block_vec = (int64_t)AR_ROW_THREADS * (nvls ? p.world : 1);
  int rpb = (int)((per_block_vec + rowvec - 1) / rowvec);
  if (rpb < 1) rpb = 1;
  int nb = (M + rpb - 1) / rpb;
  if (nb > AR_MAX_BLOCKS) {
    rpb = (M + AR_MAX_BLOCKS - 1) / AR_MAX_BLOCKS;
    nb = (M + rpb - 1) / rpb;
  }
  ArRows r;
  r.M = M; r.rowvec = rowvec; r.rpb = rpb; r.epi = epi;
  const int nv = epi ? (rowvec + 255) / 256 : 1;
  if (nvls) {
    if (nv <= 1) return launch_rows_w<true, 1>(p, r, e, nb, bf16, st);
    if (nv <= 2) return launch_rows_w<true, 2>(p, r, e, nb, bf16, st);
    return launch_rows_w<true, 4>(p, r, e, nb, bf16, st);
  }
  if (nv <= 1) return launch_rows_w<false, 1>(p, r, e, nb, bf16, st);
  if (nv <= 2) return launch_rows_w<false, 2>(p, r, e, nb, bf16, st);
  return launch_rows_w<false, 4>(p, r, e, nb, bf16, st);
}

template <bool TWOSHOT, int NV>
int launch_lamport_w(const ArParams& p, const ArRows& r, const RowNormArgs& e, int nb, bool bf16, cudaStream_t st) {
#define B200_ARL(W)                                                                                                             \
  case W:                                                                                                                       \
    if (bf16) launch_pdl(allreduce_lamport_kernel<true, W, TWOSHOT, NV>, dim3(nb), dim3(AR_ROW_THREADS), 0, st, p, r, e);       \
    else launch_pdl(allreduce_lamport_kernel<false, W, TWOSHOT, NV>, dim3(nb), dim3(AR_ROW_THREADS), 0, st, p, r, e);           \
    break;
  switch (p.world) {
    B200_ARL(2) B200_ARL(3) B200_ARL(4) B200_ARL(5) B200_ARL(6) B200_ARL(7) B200_ARL(8)
    default:
      set_error("ar: world %d", p.world);
      return B200_ERR_UNSUPPORTED;
  }
#undef B200_ARL
  return launched("allreduce_lamport_kernel launch");
}

// how many 16-byte vectors of a receive buffer a push-protocol call of this geometry writes
int64_t lamport_extent(int world, int64_t nvec, int M, int rowvec, bool twoshot, int* rpb_out, int* nb_out) {
  // one row (>= 256 vectors) per block: measured at 4 ranks (fused 256x4096 16.9 us with 1024 vectors per block, 13.3 with
  // 256 / 512, 25.7 with 2048) -- more, smaller blocks hide the poll latency of the row epilogue
  int64_t per_block_vec = AR_ROW_THREADS;
  if (tune(TUNE_AR_BLOCK_VEC) > 0) per_block_vec = tune(TUNE_AR_BLOCK_VEC);
  int rpb = (int)((per_block_vec + rowvec - 1) / rowvec);
  if (rpb < 1) rpb = 1;
  int nb = (M + rpb - 1) / rpb;
  if (nb > AR_MAX_BLOCKS) {
    rpb = (M + AR_MAX_BLOCKS - 1) / AR_MAX_BLOCKS;
    nb = (M + rpb - 1) / rpb;
  }
  *rpb_out = rpb;
  *nb_out = nb;
  if (!twoshot) return (int64_t)world * nvec;
  const int64_t sub = ((int64_t)rpb * rowvec + world - 1) / world;
  return (int64_t)world * nb * sub + nvec;
}

// push-protocol kernel: M rows of rowvec vectors; B200_ERR_UNSUPPORTED when the message does not fit a receive buffer
int launch_lamport(const ArContext* c, ArParams p, int M, int rowvec, int epi, const RowNormArgs& e, bool twoshot, bool bf16,
                   cudaStream_t st) {
  if (p.world == 2) twoshot = false;
  ArRows r;
  int nb;
  const int64_t extent = lamport_extent(p.world, p.nvec, M, rowvec, twoshot, &r.rpb, &nb);
  B200_REQUIRE((size_t)extent * 16 <= p.lam_bytes, B200_ERR_UNSUPPORTED,
               "ar: a %s push all-reduce of %lld bytes at %d ranks needs %lld bytes of receive buffer, have %zu", twoshot ? "two-shot" : "one-shot",
               (long long)p.nvec * 16, p.world, (long long)extent * 16, p.lam_bytes);
  r.M = M; r.rowvec = rowvec; r.epi = epi;
  const int nv = epi ? (rowvec + 255) / 256 : 1;
  if (twoshot) {
    if (nv <= 1) return launch_lamport_w<true, 1>(p, r, e, nb, bf16, st);
    if (nv <= 2) return launch_lamport_w<true, 2>(p, r, e, nb, bf16, st);
    return launch_lamport_w<true, 4>(p, r, e, nb, bf16, st);
  }
  if (nv <= 1) return launch_lamport_w<false, 1>(p, r, e, nb, bf16, st);
  if (nv <= 2) return launch_lamport_w<false, 2>(p, r, e, nb, bf16, st);
  return launch_lamport_w<false, 4>(p, r, e, nb, bf16, st);
}

}  // namespace

// in / out: n_elems 16-bit values, 16-byte aligned, n_elems % 8 == 0, n_elems * 2 <= max_message_bytes; out may alias in
extern "C" int b200_ar_allreduce(void* ctx, const void* in, void* out, int64_t n_elems, int dtype, void* stream) {
  ArContext* c = (ArContext*)ctx;
  B200_REQUIRE(c && in && out && n_elems > 0, B200_ERR_ARG, "ar_allreduce: bad argument");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "ar_allreduce: dtype must be bf16/f16");
  B200_REQUIRE(n_elems % 8 == 0 && ((uintptr_t)in & 15) == 0 && ((uintptr_t)out & 15) == 0, B200_ERR_ARG,
               "ar_allreduce: need 16-byte aligned buffers and a multiple of 8 elements");
  B200_REQUIRE((size_t)n_elems * 2 <= c->p.slot_bytes, B200_ERR_ARG, "ar_allreduce: message of %lld bytes exceeds the %zu-byte slots",
               (long long)n_elems * 2, c->p.slot_bytes);
  ArParams p = c->p;
  p.in = (const uint4*)in;
  p.out = (uint4*)out;
  p.nvec = n_elems / 8;
  cudaStream_t st = (cudaStream_t)stream;
  // a function of (world, size, multicast) only, so every rank picks the same kernel
  int mode = c->mode > 0 ? c->mode : b200_ar_mode2(p.world, (size_t)n_elems * 2, p.mc != nullptr);
  if (mode == 3 && !p.mc) mode = 2;
  if (mode == 2 && p.world == 2) mode = 1;
  if (mode == 0) mode = p.world <= 4 ? 1 : 2;  // the caller asked for this path although the policy prefers NCCL
  if (mode == 5 || mode == 6) {
    const int rowvec = 512;  // 8 KB pseudo-rows
    const int M = (int)((p.nvec + rowvec - 1) / rowvec);
    RowNormArgs e = {};
    if (c->mode <= 0 && mode == 5 && (size_t)p.world * p.nvec * 16 > p.lam_bytes) mode = 6;   // W copies do not fit: scatter
    return launch_lamport(c, p, M, rowvec, 0, e, mode == 6, dtype == B200_BF16, st);
  }
  if (mode == 3 || mode == 4) {
    const int rowvec = 512;  // 8 KB pseudo-rows
    const int M = (int)((p.nvec + rowvec - 1) / rowvec);
    RowNormArgs e = {};
    return launch_rows(c, p, M, rowvec, 0, e, mode == 3, dtype == B200_BF16, st);
  }
  if (mode == 2) {
    int64_t nb2 = ((p.nvec + p.world - 1) / p.world + AR_THREADS - 1) / AR_THREADS;
    if (nb2 > AR_FLAT_BLOCKS) nb2 = AR_FLAT_BLOCKS;
    if (nb2 < 1) nb2 = 1;
#define B200_AR2(W)                                                                                      \
  case W:                                                                                                \
    if (dtype == B200_BF16) allreduce_twoshot_kernel<true, W><<<(unsigned)nb2, AR_THREADS, 0, st>>>(p);  \
    else allreduce_twoshot_kernel<false, W><<<(unsigned)nb2, AR_THREADS, 0, st>>>(p);                    \
    break;
    switch (p.world) {
      B200_AR2(3) B200_AR2(4) B200_AR2(5) B200_AR2(6) B200_AR2(7) B200_AR2(8)
      default:
        set_error("ar_allreduce: world %d", p.world);
        return B200_ERR_UNSUPPORTED;
    }
#undef B200_AR2
    return launched("allreduce_twoshot_kernel launch");
  }
  // enough blocks to keep NVLink reads in flight; the block count must be the same on every rank (it is a function
  // of the message size only) because block b pairs with block b of the peers
  int64_t nb = (p.nvec + AR_THREADS - 1) / AR_THREADS;  // one 16-byte vector per thread up to 1 MiB: latency, not bandwidth, bounds these sizes
  if (nb > AR_FLAT_BLOCKS) nb = AR_FLAT_BLOCKS;
  if (nb < 1) nb = 1;
#define B200_AR(W)                                                                        \
  case W:                                                                                 \
    if (dtype == B200_BF16) allreduce_oneshot_kernel<true, W><<<(unsigned)nb, AR_THREADS, 0, st>>>(p); \
    else allreduce_oneshot_kernel<false, W><<<(unsigned)nb, AR_THREADS, 0, st>>>(p);      \
    break;
  switch (p.world) {
    B200_AR(2) B200_AR(3) B200_AR(4) B200_AR(5) B200_AR(6) B200_AR(7) B200_AR(8)
    default:
      set_error("ar_allreduce: world %d", p.world);
      return B200_ERR_UNSUPPORTED;
  }
#undef B200_AR
  return launched("allreduce_oneshot_kernel launch");
}

// all-reduce(sum) of in [M, H] fused with what follows a row-parallel linear in a decoder layer
// (flashinfer_comm_fusion.py:122 flashinfer_allreduce_add_rmsnorm; layernorm.py forward with residual):
//   x = sum over ranks (rounded to the half type, as the plain all-reduce writes it)
//   residual += x (stored back);  y = RMSNorm(residual) * weight
//   norm_out = y (optional);  (q, s) = per_token_group_quant_fp8(y, 128) (optional)
// Bit-identical to b200_ar_allreduce followed by b200_rmsnorm_quant_fp8 when the one-shot pull kernel runs; with NVLS
// the switch's fp32 accumulation order is not the rank order (results still identical on every rank).
extern "C" int b200_ar_allreduce_add_rmsnorm(void* ctx, const void* in, void* residual, const void* weight, void* norm_out,
                                             void* q, int64_t q_sm, float* s, int64_t s_sm, int64_t s_sg, int M, int H,
                                             int dtype, float eps, float q_eps, float fp8_min, float fp8_max, void* stream) {
  ArContext* c = (ArContext*)ctx;
  B200_REQUIRE(c && in && residual && weight && (norm_out || (q && s)) && M > 0 && H > 0, B200_ERR_ARG, "ar_allreduce_add_rmsnorm: bad argument");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "ar_allreduce_add_rmsnorm: dtype must be bf16/f16");
  B200_REQUIRE(H % 128 == 0 && H <= 8192, B200_ERR_UNSUPPORTED, "ar_allreduce_add_rmsnorm: H must be a multiple of 128, at most 8192 (got %d)", H);
  B200_REQUIRE((size_t)M * H * 2 <= c->p.slot_bytes, B200_ERR_UNSUPPORTED, "ar_allreduce_add_rmsnorm: message of %zu bytes exceeds the %zu-byte slots",
               (size_t)M * H * 2, c->p.slot_bytes);
  B200_REQUIRE(((uintptr_t)in & 15) == 0 && ((uintptr_t)residual & 15) == 0 && ((uintptr_t)weight & 15) == 0 &&
                   (!norm_out || ((uintptr_t)norm_out & 15) == 0) && (!q || (((uintptr_t)q & 7) == 0 && q_sm % 8 == 0)),
               B200_ERR_ARG, "ar_allreduce_add_rmsnorm: operands must be 16-byte aligned");
  ArParams p = c->p;
  p.in = (const uint4*)in;
  p.out = nullptr;
  p.nvec = (int64_t)M * H / 8;
  const size_t bytes = (size_t)M * H * 2;
  // the fused kernel is the faster plan at every (world, M, H) the push protocol serves (tools/ar_check.py: 2 ranks
  // 128x4096 9.8 us against 18.7 for all-reduce -> rmsnorm_quant_fp8; 4 ranks 256x4096 16.9 / 31.8, 256x8192 24.8 / 41.3)
  // 8 ranks (one row per block; tools/ar_check.py, us, fused against push all-reduce -> rmsnorm_quant_fp8): 512x4096 28.8
  // against 25.5, 128x5120 16.8 / 15.3, 256x8192 24.2 / 28.7 -- the fused form only pays on wide rows there (the separate
  // norm is one 16 KB row per CTA).  B200_ERR_UNSUPPORTED is the "two kernels" answer (tp.py runs them; every rank decides
  // from (world, M, H) alone)
  if (c->mode <= 0 && p.world > tune(TUNE_AR_FUSED_WORLD) && H < 8192) {
    set_error("ar_allreduce_add_rmsnorm: %d x %d at %d ranks: the two separate kernels are the faster plan", M, H, p.world);
    return B200_ERR_UNSUPPORTED;
  }
  int mode = c->mode > 0 ? c->mode : ar_policy(p.world, bytes, true);
  if (c->mode <= 0 && mode == 5 && (size_t)p.world * p.nvec * 16 > p.lam_bytes) mode = 6;
  const bool nvls = (mode == 3) && p.mc;
  RowNormArgs e;
  e.residual = (uint16_t*)residual; e.w = (const uint16_t*)weight; e.out = (uint16_t*)norm_out; e.q = (uint8_t*)q; e.q_sm = q_sm;
  e.s = s; e.s_sm = s_sm; e.s_sg = s_sg; e.H = H; e.eps = eps; e.q_eps = q_eps; e.fmin_ = fp8_min; e.fmax_ = fp8_max;
  if (mode == 5 || mode == 6) return launch_lamport(c, p, M, H / 8, 1, e, mode == 6, dtype == B200_BF16, (cudaStream_t)stream);
  return launch_rows(c, p, M, H / 8, 1, e, nvls, dtype == B200_BF16, (cudaStream_t)stream);
}
