/*
 * b200_sgl.h -- C ABI of the B200-native (sm_100a) decode hot path.
 *
 * This is the drop-in boundary: plain pointers, sizes and strides, no torch types.  Each entry
 * point names the reference interface it replaces (paths relative to the reference tree,
 * rzwei/iaas-sglang = SGLang 0.4.9 + sgl-kernel 0.2.4).  The Python host code in
 * iaas_sglang_b200/ (ops.py, backend.py) binds these with ctypes; INTEGRATION.md shows the
 * reference-side stub.
 *
 * Conventions
 *   - every pointer is a DEVICE pointer unless it says "host";
 *   - the caller owns every buffer; nothing here allocates, frees or synchronises;
 *   - `stream` is a cudaStream_t passed as void*; all work is enqueued on it and is
 *     CUDA-graph capturable;
 *   - strides are in ELEMENTS of the tensor's dtype;
 *   - return value: 0 = ok, B200_ERR_ARG (-1) = bad argument (-> RuntimeError),
 *     B200_ERR_UNSUPPORTED (-2) = unsupported shape / device (-> NotImplementedError),
 *     B200_ERR_CUDA (-3) = CUDA runtime error; b200_last_error() returns the message
 *     (thread-local, valid until the next failing call on that thread).
 *   - there is no CPU fallback: on a device that is not sm_100 every compute entry point
 *     returns B200_ERR_UNSUPPORTED.
 */
#ifndef B200_SGL_H_
#define B200_SGL_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_ERR_ARG (-1)
#define B200_ERR_UNSUPPORTED (-2)
#define B200_ERR_CUDA (-3)

/* dtype codes */
#define B200_BF16 0
#define B200_F16 1
#define B200_F32 2

/* 4: b200_ar_region_bytes grew (push-protocol receive buffers: regions sized by an older library are too small);
 *    new entry points b200_decode_attn_static, b200_rope_cos_sin, b200_fp8_blockwise_scaled_mm_rope (no struct changed) */
#define B200_SGL_ABI_VERSION 4
#define B200_KV_E4M3 3   /* b200_decode_attn_params.kv_dtype */

const char* b200_last_error(void);
int b200_abi_version(void);
/* compute capability of the current device as major*10+minor (100 on B200), or a negative error */
int b200_sm_arch(void);
int b200_sm_count(void);
/* number of kernels this library has launched (or captured into a CUDA graph) in this process */
unsigned long long b200_launch_count(void);
/* Tuning / developer knobs ("pdl", "sk_max_m", "sk_splits", "gb_pair", ...; csrc/runtime.cu lists them with their
 * B200_* environment names).  The environment is read once when the library is loaded -- never on a launch path;
 * these two change / read a knob at run time (tools, tests).  No reference counterpart. */
int b200_tune_set(const char* name, int value);
int b200_tune_get(const char* name, int* value);

/* --------------------------------------------------------------------------------------------
 * Decode attention over the paged KV pool.
 *
 * Replaces, behind AttentionBackend.forward_decode
 * (python/sglang/srt/layers/attention/base_attn_backend.py:89-99):
 *   torch.ops.sgl_kernel.decode_attention_cpu   sgl-kernel/csrc/cpu/decode.cpp:1375-1575
 *                                               (schema csrc/cpu/torch_extension_cpu.cpp:264-268)
 *   decode_attention_fwd (Triton)               python/sglang/srt/layers/attention/triton_ops/decode_attention.py:677-728
 *   KVCache.set_kv_buffer                       python/sglang/srt/mem_cache/memory_pool.py:418-456 (fused, write_kv)
 *
 *   for b < B:  idx = req_to_token[req_pool_indices[b], 0 : seq_lens[b]]          (read in place)
 *     if write_kv: k_buf[out_cache_loc[b]] = k_new[b]; v_buf[out_cache_loc[b]] = v_new[b]
 *     out[b,h,:] = softmax(sm_scale * q[b,h,:] . k_buf[idx, h / (Hq/Hkv), :]  (+ logit_cap tanh)) . v_buf[idx, h / (Hq/Hkv), :]
 *
 * MLA (absorbed) decode is the same call with Hkv = 1, D = Dv + 64 and v_buf == k_buf
 * (v is the first Dv columns of the latent row; decode.cpp:1455, memory_pool.py:829-837).
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_decode_attn_params {
  const void* q;            /* [B, Hq, D]   */
  int64_t q_stride_b, q_stride_h;
  const void* k_new;        /* [B, Hkv, D]  new token's key (may be NULL if !write_kv) */
  int64_t knew_stride_b, knew_stride_h;
  const void* v_new;        /* [B, Hkv, Dv] */
  int64_t vnew_stride_b, vnew_stride_h;
  void* k_buf;              /* [T, Hkv, D]  token_to_kv_pool.get_key_buffer(layer)   */
  int64_t kbuf_stride_t, kbuf_stride_h;
  void* v_buf;              /* [T, Hkv, Dv] token_to_kv_pool.get_value_buffer(layer) */
  int64_t vbuf_stride_t, vbuf_stride_h;
  void* out;                /* [B, Hq, Dv] contiguous */
  const void* req_to_token; /* [max_reqs, max_context_len] int32 or int64 */
  int64_t req_to_token_stride;
  const void* req_pool_indices;    /* [B] int64 (schedule_batch.py:1138) or int32 (cuda_graph_runner.py:273-276, see
                                      meta_i32_mask); NULL = identity (row b of the table belongs to request b) */
  const void* seq_lens;            /* [B] int64 or int32; includes the token being decoded */
  const void* out_cache_loc;       /* [B] int64 or int32 */
  void* workspace;          /* >= b200_decode_attn_workspace_size(...) bytes, 16-byte aligned */
  size_t workspace_bytes;
  int32_t idx_is_i64;
  int32_t B, Hq, Hkv, D, Dv;
  int32_t dtype;            /* B200_BF16 | B200_F16 (q, k, v, out) */
  int32_t write_kv;         /* fuse set_kv_buffer */
  int32_t num_splits;       /* KV splits per request; <= 0: choose from max_seq_len */
  int32_t max_seq_len;      /* host-side upper bound of seq_lens (e.g. seq_lens_cpu.max()); 0 = unknown */
  float sm_scale;
  float logit_cap;          /* 0 = off */
  /* --- ABI version 2 --- */
  int32_t meta_i32_mask;    /* bit 0: seq_lens is int32, bit 1: req_pool_indices is int32, bit 2: out_cache_loc is int32 */
  int32_t page_size;        /* <= 1: `req_to_token` holds token slots.  A power of two P > 1: it is a page table
                               (create_flashmla_kv_indices_triton, utils.py:44-94; cutlass_mla_decode's `page_table`)
                               and token t of request b lives in row table[b, t / P] * P + t % P of k_buf.
                               MLA and the generic kernel only. */
  /* MLA with the rope parts given apart, as cutlass_mla_backend.py:226-285 / cutlass_mla_decode
   * (sgl-kernel/python/sgl_kernel/attention.py:54-125) pass them: q = q_nope [B, Hq, 512] and q_rope [B, Hq, 64];
   * k_new = the 512 latent columns of the new row and k_rope_new [B, 64] its rope part (set_mla_kv_buffer,
   * memory_pool.py:859-876).  NULL = q / k_new hold all 576 columns. */
  const void* q_rope;
  int64_t qrope_stride_b, qrope_stride_h;
  const void* k_rope_new;
  int64_t krope_stride_b;
  /* ABI v3 -- cascade / shared-prefix decode (sgl-kernel/csrc/attention/cascade.cu, merge_attn_states.cu; FlashInfer's
   * cascade wrapper does the same split).  All five or none; tuned GQA kernel only.
   *   kv_start     device int32 scalar P (b200_shared_prefix_len): the kernel attends to positions [P, seq_len) only;
   *   cascade_o    float [B, Hq, Dv], cascade_lse float [B, Hq]: scratch for the suffix states;
   *   prefix_o     [B, Hq, Dv] in q's dtype, prefix_lse float [B, Hq] (natural log, -inf = empty): the attention state of
   *                every (request, head) over the shared prefix [0, P), computed ONCE for the batch by b200_extend_attn
   *                (prefix_only = 1, lse_out).  `out` receives the merged result.                                      */
  /* FP8 KV cache (--kv-cache-dtype fp8_e4m3: memory_pool.py:116-121,431-440; layers/quantization/kv_cache.py):
   * kv_dtype = B200_KV_E4M3: k_buf / v_buf hold e4m3 codes (the pool's uint8 storage; strides in elements = bytes),
   * K = code * k_scale, V = code * v_scale (scale <= 0 means 1); k_new / v_new stay in q's dtype and are appended as
   * e4m3(x / scale).  Tuned GQA kernel with head_dim 128 only (B200_ERR_UNSUPPORTED otherwise), not with cascade. */
  int32_t kv_dtype;
  float k_scale, v_scale;
  int32_t reserved1;
  const void* kv_start;
  void* cascade_o;
  void* cascade_lse;
  const void* prefix_o;
  const void* prefix_lse;
} b200_decode_attn_params;

/* number of KV splits the kernel will use for this problem (so that the caller can size scratch) */
int b200_decode_attn_num_splits(int B, int Hq, int Hkv, int D, int Dv, int max_seq_len, int num_splits_hint);
size_t b200_decode_attn_workspace_size(int B, int Hq, int Dv, int num_splits);
int b200_decode_attn(const b200_decode_attn_params* p, void* stream);
/* No counterpart in the reference.  b200_decode_attn for callers that
 * cannot allocate (CUDA-graph capture, tight serving loops, harnesses that forbid cudaMalloc): p->workspace and
 * p->workspace_bytes are IGNORED, the split-KV scratch is a file-scope device array of
 * B200_DECODE_ATTN_STATIC_SCRATCH_BYTES that every GPU loading the library holds until exit.  Same result as
 * b200_decode_attn.  B200_ERR_ARG when b200_decode_attn_workspace_size(B, Hq, Dv, splits) exceeds that size (every
 * tuned GQA / MLA shape fits; the generic kernel fits while B * Hq * splits * (Dv + 1) * 4 + 256 does).  Calls no
 * cudaMalloc / cudaFree.  All calls share the array: calls that overlap on the GPU (other streams) race; calls on one
 * stream are ordered.  Nothing in the array needs clearing between calls. */
#define B200_DECODE_ATTN_STATIC_SCRATCH_BYTES (20u << 20)
int b200_decode_attn_static(const b200_decode_attn_params* p, void* stream);
/* b200_decode_attn followed by per_token_group_quant_fp8(out.view(B, Hq * Dv), 128, column-major scales) for the
 * FP8 block o_proj (fp8_utils.py:202-204 sglang_per_token_group_quant_fp8; per_token_group_quant_8bit.cu:25-115),
 * emitted by the split-merge kernel: out is written as usual, q8 [B, Hq*Dv] e4m3 (row stride q_stride_m bytes) and
 * scales s[b * s_stride_m + h * s_stride_g] are bit-identical to the two-call sequence.  Tuned GQA kernel with
 * Dv == 128 only; anything else returns B200_ERR_UNSUPPORTED and the caller quantises separately. */
int b200_decode_attn_quant_fp8(const b200_decode_attn_params* p, void* q8, int64_t q_stride_m, float* s, int64_t s_stride_m,
                               int64_t s_stride_g, float eps, float fp8_min, float fp8_max, void* stream);

/* KVCache.set_kv_buffer alone (memory_pool.py:418-456; decode.cpp:771-809): rows scatter, bit-exact */
int b200_set_kv_buffer(void* k_buf, int64_t kbuf_stride_t, int64_t kbuf_stride_h, void* v_buf, int64_t vbuf_stride_t,
                       int64_t vbuf_stride_h, const void* k_new, int64_t knew_stride_b, int64_t knew_stride_h,
                       const void* v_new, int64_t vnew_stride_b, int64_t vnew_stride_h, const int64_t* loc, int B,
                       int Hkv, int D, int Dv, int elem_bytes, void* stream);

/* The same into an FP8 (e4m3) pool: code = e4m3(x / scale), round to nearest, saturating (memory_pool.py:431-440:
 * cache_k.div_(k_scale); cache_k.to(torch.float8_e4m3fn), stored as uint8); scale <= 0 means 1.  src_dtype = B200_BF16/F16. */
int b200_set_kv_buffer_fp8(void* k_buf, int64_t kbuf_stride_t, int64_t kbuf_stride_h, void* v_buf, int64_t vbuf_stride_t,
                           int64_t vbuf_stride_h, const void* k_new, int64_t knew_stride_b, int64_t knew_stride_h,
                           const void* v_new, int64_t vnew_stride_b, int64_t vnew_stride_h, const int64_t* loc, int B,
                           int Hkv, int D, int Dv, int src_dtype, float k_scale, float v_scale, void* stream);

/* --------------------------------------------------------------------------------------------
 * Extend (prefill) attention: AttentionBackend.forward_extend (base_attn_backend.py:101-111);
 * replaces torch.ops.sgl_kernel.extend_attention_cpu (csrc/cpu/extend.cpp; schema
 * torch_extension_cpu.cpp:271-275) / extend_attention_fwd (triton_ops/extend_attention.py).
 * Token i of request b (i < extend_seq_lens[b], stored at row extend_start_loc[b] + i) attends to
 * the request's cached prefix (seq_lens[b] - extend_seq_lens[b] tokens, read through req_to_token
 * from k_buf/v_buf) and causally to extend tokens 0..i (read from k_ext/v_ext).
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_extend_attn_params {
  const void* q;      /* [T_ext, Hq, D]  */
  int64_t q_stride_t, q_stride_h;
  const void* k_ext;  /* [T_ext, Hkv, D] */
  int64_t kext_stride_t, kext_stride_h;
  const void* v_ext;  /* [T_ext, Hkv, Dv] */
  int64_t vext_stride_t, vext_stride_h;
  void* out;          /* [T_ext, Hq, Dv] contiguous */
  const void* k_buf;
  int64_t kbuf_stride_t, kbuf_stride_h;
  const void* v_buf;
  int64_t vbuf_stride_t, vbuf_stride_h;
  const void* req_to_token;
  int64_t req_to_token_stride;
  const int64_t* req_pool_indices;  /* [B] */
  const int64_t* seq_lens;          /* [B] prefix + extend */
  const int64_t* extend_seq_lens;   /* [B] */
  const int64_t* extend_start_loc;  /* [B] */
  int32_t idx_is_i64;
  int32_t B, Hq, Hkv, D, Dv;
  int32_t max_len_extend;
  int32_t dtype;
  float sm_scale;
  float logit_cap;
  /* ABI v3.  prefix_only != 0: every new token attends to exactly the request's prefix (no new keys, no causal mask):
   * the shared-prefix half of cascade decode (sgl-kernel/csrc/attention/cascade.cu merges the two halves).
   * lse_out: optional float [T_ext, Hq], natural-log log-sum-exp of every row's scaled logits (-inf for an empty key
   * set), the `s` operand of merge_state.  Both need the tuned kernel (D == Dv in {64, 128}). */
  int32_t prefix_only;
  int32_t reserved0;
  void* lse_out;
  /* optional float [T_ext, Hq, Dv]: the normalised output in fp32 INSTEAD of `out` (which may then be NULL), with the
   * softmax weights kept exact for P V (hi + lo halves) -- the prefix state cascade decode merges, held to the decode
   * tolerance (2e-3) rather than the extend one (1e-2). */
  void* out_f32;
} b200_extend_attn_params;
int b200_extend_attn(const b200_extend_attn_params* p, void* stream);

/* merge_state / merge_state_v2 (sgl-kernel/csrc/attention/cascade.cu:12-55, merge_attn_states.cu:30-120; Python wrappers
 * sgl_kernel/attention.py:12-52): v_* [T, H, D] bf16/f16 contiguous, s_* float [T, H] natural-log LSE (an infinite value
 * marks an empty state); v_merged = (e^{s_a} v_a + e^{s_b} v_b) / (e^{s_a} + e^{s_b}), s_merged = log(e^{s_a} + e^{s_b})
 * (s_merged may be NULL).  D % 8 == 0. */
int b200_merge_state(const void* v_a, const float* s_a, const void* v_b, const float* s_b, void* v_merged, float* s_merged,
                     int64_t T, int H, int D, int dtype, void* stream);

/* Length P of the page-table prefix shared by ALL B requests of a decode batch (RadixAttention gives requests with a
 * common prompt prefix the same KV slots): the largest multiple of `align` such that req_to_token[req_b, j] is the same
 * for every b and j < P, P <= min_b seq_len_b - 1, and 0 when that is below `min_len` or B < 2.  Stays on the device:
 *   prefix_len  int32[1] = P  (the kv_start of b200_decode_attn);
 *   one_seq_len int64[1] = P + B, one_req int64[1] = req_pool_indices[0]: with extend_seq_lens = [B] and
 *   extend_start_loc = [0] they present the batch's B queries to b200_extend_attn (prefix_only) as one request whose
 *   prefix is the shared one.
 * scratch: int32[1] that holds INT_MAX between calls (fill it once with 0x7f bytes).  meta_i32_mask as in
 * b200_decode_attn_params (bit 0 seq_lens, bit 1 req_pool_indices are int32).  No reference counterpart (the reference
 * gets the prefix structure from its radix tree on the host; here the kernels stay CUDA-graph replayable). */
int b200_shared_prefix_len(const void* req_to_token, int idx_is_i64, int64_t req_to_token_stride, const void* req_pool_indices,
                           const void* seq_lens, int meta_i32_mask, int B, int align, int min_len, int32_t* scratch,
                           int32_t* prefix_len, int64_t* one_seq_len, int64_t* one_req, void* stream);

/* --------------------------------------------------------------------------------------------
 * Page-table index builders (integer, bit-exact), python/sglang/srt/layers/attention/utils.py
 * ------------------------------------------------------------------------------------------ */
/* create_flashinfer_kv_indices_triton (utils.py:5-41):
 *   kv_indices[kv_indptr[b] + j] = req_to_token[req_pool_indices[b], kv_start[b] + j], j < lens[b] */
int b200_create_kv_indices(const void* req_to_token, int idx_is_i64, int64_t req_to_token_stride,
                           const int64_t* req_pool_indices, const int64_t* lens, const int64_t* kv_indptr,
                           const int64_t* kv_start /* NULL = 0 */, int32_t* kv_indices, int B, void* stream);
/* create_flashmla_kv_indices_triton (utils.py:44-94): out[b,p] = req_to_token[req, start + p*P] / P */
int b200_create_page_table(const void* req_to_token, int idx_is_i64, int64_t req_to_token_stride,
                           const int64_t* req_pool_indices, const int64_t* lens, const int64_t* kv_start,
                           int32_t* out, int64_t out_stride, int B, int page_size, int max_pages, void* stream);

/* --------------------------------------------------------------------------------------------
 * Activation quantisation
 * ------------------------------------------------------------------------------------------ */
/* sgl_per_token_group_quant_fp8 (sgl-kernel/csrc/gemm/per_token_group_quant_8bit.cu:25-115,230-240;
 * schema common_extension.cc:116-119):  absmax = max(eps, max|x|) over each group of `group`
 * columns; s = absmax / fp8_max; q = e4m3(clamp(x / s)).  s[m, g] is stored at s[m*s_stride_m + g*s_stride_g]
 * (column_major_scales=True -> s_stride_m = 1). */
int b200_per_token_group_quant_fp8(const void* x, int64_t x_stride_m, int x_dtype, void* q, int64_t q_stride_m,
                                   float* s, int64_t s_stride_m, int64_t s_stride_g, int M, int K, int group,
                                   float eps, float fp8_min, float fp8_max, void* stream);
/* per_token_quant_int8 (python/sglang/srt/layers/quantization/int8_kernel.py:20-81):
 * absmax = max(max|x|, 1e-10); s[m] = absmax/127; q = round_half_away(x * (127/absmax)) */
int b200_per_token_quant_int8(const void* x, int64_t x_stride_m, int x_dtype, int8_t* q, int64_t q_stride_m, float* s,
                              int M, int K, void* stream);

/* --------------------------------------------------------------------------------------------
 * Quantised GEMMs (tcgen05 + TMEM + TMA)
 * ------------------------------------------------------------------------------------------ */
/* fp8_blockwise_scaled_mm (sgl-kernel/csrc/gemm/fp8_blockwise_gemm_kernel.cu:345-423; schema
 * common_extension.cc:111-114; python sgl_kernel/gemm.py:24-31):
 *   out[m,n] = sum_kb sa[m,kb] * sb[kb,nb] * sum_{k in block kb} a[m,k] * b[k,n]      (128x128 blocks)
 * a: [M,K] e4m3 row-major (lda elements); b: [K,N] COLUMN-major, i.e. the [N,K] row-major weight
 * (ldb = elements between consecutive n); sa: [M, ceil(K/128)] f32 at sa[m*sa_stride_m + kb*sa_stride_k]
 * (reference requires M-major: sa_stride_m == 1); sb: [ceil(K/128), ceil(N/128)] f32 at
 * sb[kb*sb_stride_k + nb*sb_stride_n]; out: [M,N] row-major bf16/f16 (ldo elements). */
/* split-K scratch for the two GEMMs below: `workspace` must be zero-filled ONCE before its first use
 * (it holds self-resetting tile tickets) and must not be shared by GEMMs running concurrently. */
size_t b200_gemm_workspace_size(int M, int N, int K);
int b200_fp8_blockwise_scaled_mm(const void* a, int64_t lda, const void* b, int64_t ldb, const float* sa,
                                 int64_t sa_stride_m, int64_t sa_stride_k, const float* sb, int64_t sb_stride_k,
                                 int64_t sb_stride_n, void* out, int64_t ldo, int out_dtype, int M, int N, int K,
                                 void* workspace, size_t workspace_bytes, void* stream);
/* b200_fp8_blockwise_scaled_mm for a qkv projection followed by neox RoPE (python/sglang/srt/layers/rotary_embedding.py
 * forward_native, is_neox_style; the call pair python/sglang/srt/models/llama.py:186-188 qkv_proj -> rotary_emb) in ONE
 * launch: the first rope_cols output columns (q heads then k heads, head_dim columns each) are rotated with token m's row of
 * cos_sin [M, head_dim / 2, 2] f32 (b200_rope_cos_sin below: once per decode step, the positions are the same for every
 * layer); bit-identical to b200_fp8_blockwise_scaled_mm followed by b200_rope_neox on those columns.
 * B200_ERR_UNSUPPORTED (run the two calls) unless head_dim == 128 and the decode-shape kernel serves (M, N, K). */
int b200_fp8_blockwise_scaled_mm_rope(const void* a, int64_t lda, const void* b, int64_t ldb, const float* sa,
                                      int64_t sa_stride_m, int64_t sa_stride_k, const float* sb, int64_t sb_stride_k,
                                      int64_t sb_stride_n, void* out, int64_t ldo, int out_dtype, int M, int N, int K,
                                      const float* cos_sin, int rope_cols, int head_dim, void* workspace,
                                      size_t workspace_bytes, void* stream);
/* int8_scaled_mm (sgl-kernel/csrc/gemm/int8_gemm_kernel.cu:663-747; schema common_extension.cc:101-104):
 *   out[m,n] = ((float)(sum_k a[m,k]*b[k,n]) * sa[m]) * sb[n] + bias[n]
 * a: [M,K] int8 row-major; b: [K,N] column-major; bias: [N] of out dtype or NULL. */
int b200_int8_scaled_mm(const void* a, int64_t lda, const void* b, int64_t ldb, const float* sa, const float* sb,
                        const void* bias, void* out, int64_t ldo, int out_dtype, int M, int N, int K, void* workspace,
                        size_t workspace_bytes, void* stream);
/* awq_dequantize (sgl-kernel/csrc/gemm/awq_kernel.cu:126-221; schema common_extension.cc:98-99):
 *   out[r, 8c+j] = half((nib(qweight[r,c], ord[j]) - nib(qzeros[r/g,c], ord[j])) * scales[r/g, 8c+j]),
 *   ord = {0,4,1,5,2,6,3,7}.  qweight [rows, cols] int32, qzeros [rows/g, cols], scales [rows/g, 8*cols]. */
int b200_awq_dequantize(const int32_t* qweight, const void* scales, const int32_t* qzeros, void* out, int dtype,
                        int rows, int cols, int group, void* stream);
/* AWQLinearMethod.apply (python/sglang/srt/layers/quantization/awq.py:186-204):
 *   out[M,N] = x[M,K] @ dequant(qweight[K,N/8], scales, qzeros) (+ bias).
 * v1 stages the dequantised weight K-major in `workspace` (>= b200_awq_w4a16_workspace_size bytes,
 * 256-byte aligned, zero-filled once) and runs the tcgen05 half-precision GEMM on it. */
size_t b200_awq_w4a16_workspace_size(int M, int N, int K);
int b200_awq_w4a16_mm(const void* x, int64_t ldx, const int32_t* qweight, const void* scales, const int32_t* qzeros,
                      const void* bias, void* out, int64_t ldo, int dtype, int M, int N, int K, int group,
                      void* workspace, size_t workspace_bytes, void* stream);

/* --------------------------------------------------------------------------------------------
 * The steps either side of the path, used by the decode-step harness:
 * RMSNorm (+residual add), SiLU-and-mul, neox RoPE.  References:
 *   sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu, python/sglang/srt/layers/layernorm.py,
 *   sgl-kernel/csrc/elementwise/activation.cu, python/sglang/srt/layers/rotary_embedding.py
 * ------------------------------------------------------------------------------------------ */
/* if residual != NULL: residual += x (stored back, half precision), then out = rmsnorm(residual)*w */
int b200_rmsnorm(const void* x, void* residual, const void* weight, void* out, int dtype, int M, int H, float eps,
                 void* stream);
/* Fused producers of an FP8 block-quantised linear's input:
 *   rmsnorm_quant_fp8   = fused add + RMSNorm (sgl-kernel/csrc/elementwise/fused_add_rms_norm_kernel.cu;
 *                         python/sglang/srt/layers/layernorm.py) followed by sgl_per_token_group_quant_fp8
 *                         (per_token_group_quant_8bit.cu:25-115), group 128; `out` (half result) may be NULL;
 *                         q == NULL runs the norm alone;
 *   silu_mul_quant_fp8  = silu_and_mul (sgl-kernel/csrc/elementwise/activation.cu) followed by the same quantiser.
 * q / s are bit-identical to the two separate ops.  s strides as in b200_per_token_group_quant_fp8. */
int b200_rmsnorm_quant_fp8(const void* x, void* residual, const void* weight, void* out, void* q, int64_t q_stride_m,
                           float* s, int64_t s_stride_m, int64_t s_stride_g, int dtype, int M, int H, float eps,
                           float quant_eps, float fp8_min, float fp8_max, void* stream);
int b200_silu_mul_quant_fp8(const void* x, void* q, int64_t q_stride_m, float* s, int64_t s_stride_m, int64_t s_stride_g,
                            int dtype, int M, int d, float quant_eps, float fp8_min, float fp8_max, void* stream);
int b200_silu_and_mul(const void* x, void* out, int dtype, int M, int d, void* stream);
/* in-place neox-style rotary embedding on q [M,Hq,D] and k [M,Hkv,D] (heads contiguous, row strides in
 * elements: q and k may be column slices of one packed qkv tensor), positions int64 [M] */
/* cos_sin[m][i] = (cos, sin)(positions[m] * theta^(-2 i / head_dim)), i < head_dim / 2 -- the rows of the reference's
 * cos_sin_cache (rotary_embedding.py _compute_cos_sin_cache) for this step's positions, in f32 as b200_rope_neox evaluates them */
int b200_rope_cos_sin(const int64_t* positions, int M, int head_dim, float theta, float* cos_sin, void* stream);
int b200_rope_neox(void* q, int64_t q_stride_m, void* k, int64_t k_stride_m, const int64_t* positions, int dtype, int M,
                   int Hq, int Hkv, int D, float theta, void* stream);

/* --------------------------------------------------------------------------------------------
 * All-reduce(sum) over NVLink peer memory / NVSwitch multicast for the row-parallel linears' outputs, optionally
 * fused with the residual-add + RMSNorm (+ FP8 quantisation) that follows it.
 * Replaces the custom P2P all-reduce of sgl-kernel/csrc/allreduce/custom_all_reduce.cuh:203-456 /
 * python/sglang/srt/distributed/device_communicators/custom_all_reduce.py:148,414-480 (call site
 * communication_op.py:11-13) and the fusion hook python/sglang/srt/layers/flashinfer_comm_fusion.py:122
 * (flashinfer_allreduce_add_rmsnorm).  One process per GPU.  Two ways to set up:
 *   (a) b200_ar_alloc on every rank, exchange the 64-byte CUDA-IPC handles on the host (torch.distributed
 *       all_gather), b200_ar_init opens the peers' regions;
 *   (b) the host already has peer mappings of a symmetric allocation of b200_ar_region_bytes() zero-filled bytes
 *       (torch.distributed._symmetric_memory rendezvous): b200_ar_init_ptrs takes the W device pointers and the
 *       multicast address of the same buffer (0 if the fabric has none), which enables the NVLS kernel.
 * Every call launches ONE kernel on `stream` (no host sync, CUDA-graph capturable); every rank must make the same
 * calls with the same sizes in the same order.  Results are identical on all ranks and are the fp32 sum in rank order
 * rounded once (the NVLS kernel, mode 3, is the exception: the switch's own accumulation order).  The default kernels
 * use a push protocol: a rank stores its payload straight into its peers' receive buffers and the receivers poll the
 * payload itself, so a call costs one NVLink one-way latency per phase instead of flag round trips.  b200_ar_error returns 1 if a peer never arrived (the kernel gave up after
 * ~20 s instead of hanging).
 * ------------------------------------------------------------------------------------------ */
size_t b200_ar_region_bytes(size_t max_message_bytes);
int b200_ar_alloc(size_t max_message_bytes, void** region_out, void* ipc_handle_out64);
void b200_ar_free_region(void* region); /* a region from b200_ar_alloc that never reached b200_ar_init */
int b200_ar_init(void* own_region, const void* handles, int rank, int world, size_t max_message_bytes, void** ctx_out);
int b200_ar_init_ptrs(const uint64_t* regions /* host array [world] of device pointers */, uint64_t multicast, int rank,
                      int world, size_t max_message_bytes, void** ctx_out);
int b200_ar_has_multicast(void* ctx);
int b200_ar_allreduce(void* ctx, const void* in, void* out, int64_t n_elems, int dtype, void* stream);
/* in [M, H] (this rank's partial sums); residual [M, H] += sum (stored back); y = RMSNorm(residual) * weight;
 * norm_out [M, H] = y (may be NULL); q / s = per_token_group_quant_fp8(y, 128) with the strides of
 * b200_per_token_group_quant_fp8 (may be NULL).  Equal, bit for bit when no multicast is involved, to
 * b200_ar_allreduce followed by b200_rmsnorm_quant_fp8.  H % 128 == 0, H <= 8192, M * H * 2 <= max_message_bytes.
 * B200_ERR_UNSUPPORTED where the two separate kernels are the faster plan (more than 4 ranks and H < 8192, measured). */
int b200_ar_allreduce_add_rmsnorm(void* ctx, const void* in, void* residual, const void* weight, void* norm_out, void* q,
                                  int64_t q_stride_m, float* s, int64_t s_stride_m, int64_t s_stride_g, int M, int H,
                                  int dtype, float eps, float quant_eps, float fp8_min, float fp8_max, void* stream);
/* measured policy: 1 = one-shot pull kernel, 2 = two-shot pull kernel, 3 = NVLS multicast kernel, 4 = row-layout
 * one-shot pull, 5 = push protocol one-shot, 6 = push protocol two-shot; 0 = leave the message to ncclAllReduce
 * (world outside 2..8).  b200_ar_mode = b200_ar_mode2 without multicast. */
int b200_ar_mode(int world, size_t bytes);
int b200_ar_mode2(int world, size_t bytes, int has_multicast);
/* force a kernel for every following call (-1 = back to the policy); must be the same on every rank */
int b200_ar_set_mode(void* ctx, int mode);
int b200_ar_error(void* ctx);
void b200_ar_destroy(void* ctx);

#ifdef __cplusplus
}
#endif
#endif /* B200_SGL_H_ */
