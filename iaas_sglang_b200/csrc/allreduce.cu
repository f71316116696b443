// One-shot all-reduce(sum) of the row-parallel linears' [M, hidden] half outputs over NVLink peer memory.
//
// Replaces, for messages up to a few MiB, the custom P2P all-reduce the reference routes small messages to
// (sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:203-456, python/sglang/srt/distributed/device_communicators/
// custom_all_reduce.py:148,414-480; call site communication_op.py:11-13 after RowParallelLinear, linear.py:1375-1376).
// Larger messages and anything this path rejects stay on ncclAllReduce (tp.py).
//
// Every rank owns one cudaMalloc'ed region, opened by its peers through CUDA IPC:
//   [flags: MAX_BLOCKS x 8 uint32, written by PEERS] [flag sequence + call counters: per block, private] [err]
//   [data slot 0] [data slot 1]
// One kernel per all-reduce, grid = nblocks (<= MAX_BLOCKS), no host synchronisation, CUDA-graph capturable:
//   1. block b copies its chunk of the input into the rank's own data slot (slot = call parity);
//   2. block b tells block b of every peer "my chunk of call n is in place" (st.release.sys into the peer's flag
//      word [b][my rank]) and waits until every peer has said the same (ld.acquire.sys on its own flag words);
//   3. block b reads chunk b of all ranks' slots (its own from local HBM, the peers' over NVLink), sums them in
//      RANK ORDER in fp32 -- every rank computes bit-identical results -- and writes the output (may alias the input).
// The call number lives in device memory (per block), so a replayed CUDA graph keeps counting.  Two data slots make
// an end-of-call barrier unnecessary: a rank can only start call n + 2 (same slot as n) after it passed the barrier of
// call n + 1, which every peer signals only after finishing call n.  A spin that lasts longer than ~20 s sets an
// error word and lets the kernel finish (wrong data, no hang); the host checks it with b200_ar_error().
#include <stdlib.h>
#include <string.h>

#include "common.cuh"

namespace b200 {

constexpr int AR_MAX_BLOCKS = 128;
constexpr int AR_MAX_RANKS = 8;
constexpr int AR_THREADS = 512;
constexpr size_t AR_FLAG_BYTES = AR_MAX_BLOCKS * AR_MAX_RANKS * 4;  // 4 KB
constexpr size_t AR_CNT_OFF = AR_FLAG_BYTES;
constexpr size_t AR_ERR_OFF = AR_CNT_OFF + AR_MAX_BLOCKS * 4;
constexpr size_t AR_CALLS_OFF = AR_ERR_OFF + 64;  // per-block number of calls (its parity picks the data slot)
constexpr size_t AR_DATA_OFF = 8192;
static_assert(AR_CALLS_OFF + AR_MAX_BLOCKS * 4 <= AR_DATA_OFF, "region header");

struct ArParams {
  uint8_t* base[AR_MAX_RANKS];  // region of every rank (own = local pointer, peers = IPC mappings)
  int rank, world;
  size_t slot_bytes;
  const uint4* in;
  uint4* out;
  int64_t nvec;  // 16-byte vectors
};

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// peer data is read around the L1 (a line of the same slot from two calls ago must not be served from it)
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

template <bool BF16, int WORLD>
__global__ void __launch_bounds__(AR_THREADS) allreduce_oneshot_kernel(const __grid_constant__ ArParams p) {
  const int b = blockIdx.x;
  uint8_t* mine = p.base[p.rank];
  uint32_t* cnt = reinterpret_cast<uint32_t*>(mine + AR_CNT_OFF) + b;
  uint32_t* calls = reinterpret_cast<uint32_t*>(mine + AR_CALLS_OFF) + b;
  const uint32_t call = *cnt + 1;  // this block's flag sequence number (private to the block, bumped at the end)
  const int slot = *calls & 1;     // consecutive calls (of either kernel) alternate between the two data slots
  const int64_t per = (p.nvec + gridDim.x - 1) / gridDim.x;
  const int64_t v0 = (int64_t)b * per, v1 = v0 + per < p.nvec ? v0 + per : p.nvec;

  // 1. my chunk -> my slot
  uint4* my_slot = reinterpret_cast<uint4*>(mine + AR_DATA_OFF + slot * p.slot_bytes);
  for (int64_t i = v0 + threadIdx.x; i < v1; i += AR_THREADS) my_slot[i] = p.in[i];
  __syncthreads();  // (the release below orders the whole block's stores: cumulativity through the barrier)

  // 2. per-block barrier with the same block of every peer
  if (threadIdx.x < WORLD && threadIdx.x != p.rank) {
    const int peer = threadIdx.x;
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(p.base[peer]) + b * AR_MAX_RANKS + p.rank, call);
    const uint32_t* flag = reinterpret_cast<const uint32_t*>(mine) + b * AR_MAX_RANKS + peer;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(flag) - call) < 0) {
      if (clock64() - t0 > 40000000000LL) {  // ~20 s: a peer is gone; report instead of hanging the GPU
        *reinterpret_cast<uint32_t*>(mine + AR_ERR_OFF) = 1;
        break;
      }
    }
  }
  __syncthreads();

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
    p.out[i] = make_uint4(pack_half2<BF16>(acc[0], acc[1]), pack_half2<BF16>(acc[2], acc[3]), pack_half2<BF16>(acc[4], acc[5]),
                          pack_half2<BF16>(acc[6], acc[7]));
  }
  if (threadIdx.x == 0) {
    *cnt = call;
    *calls = *calls + 1;
  }
}

// Two-shot variant for more ranks / larger messages (custom_all_reduce.cuh:300-400 picks the same way): rank r
// reduces only chunk r (1/W of the message) from every rank, publishes it in place in its own slot, and all ranks
// gather the W reduced chunks: 2 (W-1)/W of the message crosses NVLink per rank instead of (W-1) times the message.
// Block b owns sub-range b of EVERY chunk in all three phases, so the per-block flag barrier still covers exactly
// the data the block depends on.  Two barriers per call, two consecutive flag sequence numbers.
template <int WORLD>
__device__ __forceinline__ void ar_block_barrier(const ArParams& p, uint8_t* mine, int b, uint32_t value) {
  if (threadIdx.x < WORLD && threadIdx.x != p.rank) {
    const int peer = threadIdx.x;
    __threadfence_system();
    st_release_sys(reinterpret_cast<uint32_t*>(p.base[peer]) + b * AR_MAX_RANKS + p.rank, value);
    const uint32_t* flag = reinterpret_cast<const uint32_t*>(mine) + b * AR_MAX_RANKS + peer;
    const long long t0 = clock64();
    while ((int32_t)(ld_acquire_sys(flag) - value) < 0) {
      if (clock64() - t0 > 40000000000LL) {
        *reinterpret_cast<uint32_t*>(mine + AR_ERR_OFF) = 1;
        break;
      }
    }
  }
  __syncthreads();
}

template <bool BF16, int WORLD>
__global__ void __launch_bounds__(AR_THREADS) allreduce_twoshot_kernel(const __grid_constant__ ArParams p) {
  const int b = blockIdx.x;
  uint8_t* mine = p.base[p.rank];
  uint32_t* cnt = reinterpret_cast<uint32_t*>(mine + AR_CNT_OFF) + b;
  uint32_t* calls = reinterpret_cast<uint32_t*>(mine + AR_CALLS_OFF) + b;
  const uint32_t call = *cnt + 1;  // two barriers per call: flag values call and call + 1
  const int slot = *calls & 1;
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
  __syncthreads();
  ar_block_barrier<WORLD>(p, mine, b, call);
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
    const uint4 o = make_uint4(pack_half2<BF16>(acc[0], acc[1]), pack_half2<BF16>(acc[2], acc[3]),
                               pack_half2<BF16>(acc[4], acc[5]), pack_half2<BF16>(acc[6], acc[7]));
    my_slot[i] = o;
    p.out[i] = o;
  }
  __syncthreads();
  ar_block_barrier<WORLD>(p, mine, b, call + 1);
  // 3. gather the other ranks' reduced chunks: all W - 1 remote loads of a thread are issued before the first store
  //    (volatile loads are not reordered across the stores by the compiler: chunk by chunk they were W - 1 serial
  //    NVLink round trips)
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
  if (threadIdx.x == 0) {
    *cnt = call + 1;
    *calls = *calls + 1;
  }
}

struct ArContext {
  ArParams p;
  void* own;
  void* opened[AR_MAX_RANKS];
  int device;
};

}  // namespace b200

using namespace b200;

extern "C" size_t b200_ar_region_bytes(size_t max_message_bytes) {
  const size_t slot = (max_message_bytes + 255) & ~(size_t)255;
  return AR_DATA_OFF + 2 * slot;
}

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
  B200_CUDA(cudaIpcGetMemHandle(&h, ptr));
  memcpy(ipc_handle_out64, &h, 64);
  *region_out = ptr;
  return B200_OK;
}

// handles = world x 64 bytes in rank order (own entry ignored); returns an opaque context
extern "C" int b200_ar_init(void* own_region, const void* handles, int rank, int world, size_t max_message_bytes,
                            void** ctx_out) {
  B200_REQUIRE(own_region && handles && ctx_out && world >= 2 && world <= AR_MAX_RANKS && rank >= 0 && rank < world, B200_ERR_ARG,
               "ar_init: need 2 <= world <= %d", AR_MAX_RANKS);
  ArContext* c = new ArContext();
  memset(c, 0, sizeof(*c));
  c->own = own_region;
  cudaGetDevice(&c->device);
  c->p.rank = rank;
  c->p.world = world;
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
  *ctx_out = c;
  return B200_OK;
}

extern "C" int b200_ar_error(void* ctx) {
  ArContext* c = (ArContext*)ctx;
  if (!c) return -1;
  uint32_t v = 0;
  if (cudaMemcpy(&v, (uint8_t*)c->own + AR_ERR_OFF, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
  return (int)v;
}

extern "C" void b200_ar_destroy(void* ctx) {
  ArContext* c = (ArContext*)ctx;
  if (!c) return;
  for (int r = 0; r < AR_MAX_RANKS; ++r)
    if (c->opened[r]) cudaIpcCloseMemHandle(c->opened[r]);
  cudaFree(c->own);
  delete c;
}

// Which path is fastest for a message, measured on 2 / 4 / 8 B200 of one NVSwitch box (tools/ar_check.py AR_SWEEP=1,
// profiles/r01b_allreduce.md): 1 = one-shot, 2 = two-shot, 0 = leave it to ncclAllReduce.  The reference makes the same
// kind of split (custom_all_reduce.cuh: world 2 one stage; 4: < 512 KB; 8: < 256 KB one stage, else two).
extern "C" int b200_ar_mode(int world, size_t bytes) {
  if (world < 2 || world > AR_MAX_RANKS) return 0;
  if (world == 2) return 1;
  if (world <= 4) return bytes <= ((size_t)6 << 20) ? 1 : 2;
  if (bytes <= (size_t)1536 * 1024) return 1;
  if (bytes < ((size_t)3 << 20)) return 0;
  return 2;
}

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
  // enough blocks to keep NVLink reads in flight; the block count must be the same on every rank (it is a function
  // of the message size only) because block b pairs with block b of the peers
  int64_t nb = (p.nvec + AR_THREADS - 1) / AR_THREADS;  // one 16-byte vector per thread up to 1 MiB: latency, not bandwidth, bounds these sizes
  if (nb > AR_MAX_BLOCKS) nb = AR_MAX_BLOCKS;
  if (nb < 1) nb = 1;
  cudaStream_t st = (cudaStream_t)stream;
  // a function of (world, size) only, so every rank picks the same kernel
  bool two_shot = b200_ar_mode(p.world, (size_t)n_elems * 2) == 2;
  {
    static int mode = -1;  // developer override: B200_AR_MODE=1 one-shot, 2 two-shot (must be the same on every rank)
    if (mode < 0) { const char* e = getenv("B200_AR_MODE"); mode = e ? atoi(e) : 0; }
    if (mode == 1) two_shot = false;
    if (mode == 2 && p.world > 2) two_shot = true;
  }
  if (two_shot) {
    int64_t nb2 = ((p.nvec + p.world - 1) / p.world + AR_THREADS - 1) / AR_THREADS;
    if (nb2 > AR_MAX_BLOCKS) nb2 = AR_MAX_BLOCKS;
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
