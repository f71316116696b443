// Cascade (shared-prefix) support for paged decode attention:
//   b200_merge_state          merge_state / merge_state_v2 of the reference (sgl-kernel/csrc/attention/cascade.cu:12-55,
//                             merge_attn_states.cu:30-120; python/sgl_kernel/attention.py:12-52): two attention states
//                             (V [T, H, D] half, S = natural-log LSE [T, H] fp32) of the same queries over disjoint key
//                             sets -> the state over the union.
//   b200_shared_prefix_len    how many leading page-table entries ALL requests of a decode batch share (RadixAttention
//                             hands requests with a common prompt prefix the same KV slots, SURVEY 8d cfg 4).  The result
//                             stays on the device -- the decode kernels read it there, so a captured CUDA graph follows the
//                             batch -- together with the two one-element tensors that present the whole batch to the extend
//                             kernel as ONE request whose prefix is the shared one.
#include <limits.h>

#include "common.cuh"

namespace b200 {

// one thread per (token, head, 8 columns): out = (e^{sa} va + e^{sb} vb) / (e^{sa} + e^{sb}), s = log(e^{sa} + e^{sb})
template <bool BF16>
__global__ void __launch_bounds__(256) merge_state_kernel(const uint16_t* __restrict__ va, const float* __restrict__ sa,
                                                          const uint16_t* __restrict__ vb, const float* __restrict__ sb,
                                                          uint16_t* __restrict__ vo, float* __restrict__ so, int64_t TH, int D) {
  griddep_wait();
  griddep_launch();
  const int per = D / 8;
  const int64_t i = (int64_t)blockIdx.x * 256 + threadIdx.x;
  if (i >= TH * per) return;
  const int64_t th = i / per;
  const int c = (int)(i - th * per);
  float la = sa[th], lb = sb[th];
  // merge_attn_states.cu:66-67: an infinite LSE (of either sign) marks an empty state
  if (isinf(la)) la = -INFINITY;
  if (isinf(lb)) lb = -INFINITY;
  const float mx = fmaxf(la, lb);
  float wa = 0.f, wb = 0.f, lse = -INFINITY;
  if (mx != -INFINITY) {
    const float ea = expf(la - mx), eb = expf(lb - mx);
    const float den = ea + eb;
    wa = ea / den;
    wb = eb / den;
    lse = logf(den) + mx;
  }
  const uint4 a = *reinterpret_cast<const uint4*>(va + th * D + c * 8);
  const uint4 b = *reinterpret_cast<const uint4*>(vb + th * D + c * 8);
  const uint32_t aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
  uint32_t ow[4];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const float2 fa = unpack_half2<BF16>(aw[k]), fb = unpack_half2<BF16>(bw[k]);
    ow[k] = pack_half2<BF16>(fa.x * wa + fb.x * wb, fa.y * wa + fb.y * wb);
  }
  *reinterpret_cast<uint4*>(vo + th * D + c * 8) = make_uint4(ow[0], ow[1], ow[2], ow[3]);
  if (so && c == 0) so[th] = lse;
}

__device__ __forceinline__ int64_t casc_meta(const void* p, int64_t i, bool is32) {
  return is32 ? (int64_t) reinterpret_cast<const int32_t*>(p)[i] : reinterpret_cast<const int64_t*>(p)[i];
}

// block b >= 1 compares the page-table row of request b with the row of request 0 over [0, min(len_0, len_b) - 1) and
// lowers *raw to the first difference; block 0 only contributes its own length bound
__global__ void __launch_bounds__(256) shared_prefix_scan_kernel(const void* __restrict__ r2t, int idx64, int64_t r2t_stride,
                                                                 const void* __restrict__ rpi, const void* __restrict__ seq_lens,
                                                                 int meta32, int B, int* __restrict__ raw) {
  const int b = blockIdx.x;
  const int64_t req0 = rpi ? casc_meta(rpi, 0, meta32 & 2) : 0, reqb = rpi ? casc_meta(rpi, b, meta32 & 2) : b;
  // the new token (position len - 1) is never part of the prefix: its slot is written by this very step
  int64_t lim = casc_meta(seq_lens, b, meta32 & 1) - 1;
  const int64_t lim0 = casc_meta(seq_lens, 0, meta32 & 1) - 1;
  if (lim0 < lim) lim = lim0;
  if (lim < 0) lim = 0;
  __shared__ int first;
  if (threadIdx.x == 0) first = (int)(lim > INT_MAX ? INT_MAX : lim);
  __syncthreads();
  if (b > 0) {
    const int esz = idx64 ? 8 : 4;
    const uint8_t* row0 = reinterpret_cast<const uint8_t*>(r2t) + req0 * r2t_stride * esz;
    const uint8_t* rowb = reinterpret_cast<const uint8_t*>(r2t) + reqb * r2t_stride * esz;
    for (int64_t base = 0; base < lim; base += 256 * 8) {
      const int f = first;
      __syncthreads();            // every thread has read the same value: writes only happen between this barrier and the next
      if (base >= f) break;       // (block-uniform)
      int64_t miss = INT_MAX;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int64_t pos = base + u * 256 + threadIdx.x;
        if (pos < lim) {
          const bool same = idx64 ? reinterpret_cast<const int64_t*>(row0)[pos] == reinterpret_cast<const int64_t*>(rowb)[pos]
                                  : reinterpret_cast<const int32_t*>(row0)[pos] == reinterpret_cast<const int32_t*>(rowb)[pos];
          if (!same && pos < miss) miss = pos;
        }
      }
      if (miss != INT_MAX) atomicMin(&first, (int)miss);
      __syncthreads();
    }
  }
  __syncthreads();
  if (threadIdx.x == 0) atomicMin(raw, first);
}

// P = raw rounded down to `align`, 0 below `min_len` or for a batch of one; plus the one-request view of the batch
__global__ void shared_prefix_finish_kernel(int* raw, int B, int align, int min_len, const void* rpi, int meta32,
                                            int32_t* prefix_len, int64_t* one_seq_len, int64_t* one_req) {
  int p = *raw;
  p = p / align * align;
  if (p < min_len || B < 2) p = 0;
  *prefix_len = p;
  *one_seq_len = (int64_t)p + B;   // "one request": B new tokens (the batch's queries) after p shared prefix tokens
  *one_req = rpi ? casc_meta(rpi, 0, meta32 & 2) : 0;
  *raw = INT_MAX;                  // armed for the next call
}

}  // namespace b200

using namespace b200;

extern "C" int b200_merge_state(const void* v_a, const float* s_a, const void* v_b, const float* s_b, void* v_merged,
                                float* s_merged, int64_t T, int H, int D, int dtype, void* stream) {
  B200_REQUIRE(v_a && s_a && v_b && s_b && v_merged && T >= 0 && H > 0 && D > 0, B200_ERR_ARG, "merge_state: bad argument");
  B200_REQUIRE(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED, "merge_state: dtype must be bf16/f16");
  B200_REQUIRE(D % 8 == 0 && ((uintptr_t)v_a & 15) == 0 && ((uintptr_t)v_b & 15) == 0 && ((uintptr_t)v_merged & 15) == 0, B200_ERR_UNSUPPORTED,
               "merge_state: head_dim must be a multiple of 8 and the states 16-byte aligned");
  if (T == 0) return B200_OK;
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "merge_state: requires an sm_100 device (got sm_%d); no fallback", arch);
  const int64_t n = T * H * (D / 8);
  const unsigned blocks = (unsigned)((n + 255) / 256);
  if (dtype == B200_BF16)
    launch_pdl(merge_state_kernel<true>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const uint16_t*)v_a, s_a, (const uint16_t*)v_b, s_b, (uint16_t*)v_merged, s_merged, T * H, D);
  else
    launch_pdl(merge_state_kernel<false>, dim3(blocks), dim3(256), 0, (cudaStream_t)stream, (const uint16_t*)v_a, s_a, (const uint16_t*)v_b, s_b, (uint16_t*)v_merged, s_merged, T * H, D);
  return launched("merge_state_kernel launch");
}

// scratch: one int32 that holds INT_MAX between calls (the caller fills it with INT_MAX once; 0x7f bytes will do too)
extern "C" int b200_shared_prefix_len(const void* req_to_token, int idx_is_i64, int64_t req_to_token_stride,
                                      const void* req_pool_indices, const void* seq_lens, int meta_i32_mask, int B, int align,
                                      int min_len, int32_t* scratch, int32_t* prefix_len, int64_t* one_seq_len,
                                      int64_t* one_req, void* stream) {
  B200_REQUIRE(req_to_token && seq_lens && scratch && prefix_len && one_seq_len && one_req && B >= 1 && align >= 1, B200_ERR_ARG,
               "shared_prefix_len: bad argument");
  const int arch = b200_sm_arch();
  B200_REQUIRE(arch >= 100 && arch < 110, B200_ERR_UNSUPPORTED, "shared_prefix_len: requires an sm_100 device (got sm_%d); no fallback", arch);
  cudaStream_t st = (cudaStream_t)stream;
  shared_prefix_scan_kernel<<<B, 256, 0, st>>>(req_to_token, idx_is_i64, req_to_token_stride, req_pool_indices, seq_lens,
                                                meta_i32_mask, B, scratch);
  if (int rc = launched("shared_prefix_scan_kernel launch")) return rc;
  shared_prefix_finish_kernel<<<1, 1, 0, st>>>(scratch, B, align, min_len, req_pool_indices, meta_i32_mask, prefix_len,
                                               one_seq_len, one_req);
  return launched("shared_prefix_finish_kernel launch");
}
