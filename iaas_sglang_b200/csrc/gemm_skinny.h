// shared between gemm_tc.cu (entry points) and gemm_skinny.cu (decode-shape kernel)
#pragma once
#include "common.cuh"

namespace b200 {

struct SkinnyParams {
  const float* sa;
  int64_t sa_sm, sa_sk;
  const float* sb;
  int64_t sb_sk, sb_sn;
  const void* bias;
  // MODE 2 (AWQ W4A16): zero points int32 [K/group, N/8], scales half [K/group, N], group size (multiple of 8)
  const int32_t* qzeros;
  const uint16_t* scales;
  int group;
  void* out;
  int64_t ldo;
  int out_bf16;
  int M, N, K, mt_pad;
  int n_tiles, kb_total;
  int splits;   // CTAs per cluster = K slices per n-tile (1, 2, 4 or 8)
  int stages;   // depth of the TMA ring
  int tab_cap;  // MODE 0: k-blocks per chunk of the shared-memory scale table
  // fused neox RoPE on the first rope_cols output columns (q and k heads of a qkv projection; head_dim == the 128-column
  // tile): rope_cs[m][i] = (cos, sin) of token m's position times frequency i (b200_rope_cos_sin), nullptr = plain GEMM
  const float2* rope_cs;
  int rope_cols;
  unsigned long long* stamps;  // developer: per-CTA %globaltimer stamps (B200_SK_DEBUG & 256)
  int debug;    // developer knobs (B200_SK_DEBUG): 1 skip the DSMEM scatter, 2 skip the whole exchange, 4 skip the final sum
};

size_t skinny_workspace_bytes(int M, int N, int K);
bool skinny_applicable(int M);      // int8 / AWQ decode shapes and workspace sizing
bool skinny_applicable_int8(int M, int N, int K); // int8: measured policy
bool skinny_applicable_fp8(int M, int N, int K);  // fp8 block-scaled: measured policy skinny vs large-M kernel
// a [M,K] activations, w [N,K] weight (both K-major, 1-byte elements); MODE 0 fp8 block-scaled, 1 int8;
// MODE 2: a = half activations [M,K] (lda in BYTES), w = AWQ qweight int32 [K, N/8] (ldw in BYTES)
template <int MODE>
int launch_skinny(const void* a, int64_t lda, const void* w, int64_t ldw, SkinnyParams& sp, void* workspace,
                  size_t workspace_bytes, cudaStream_t st);

}  // namespace b200
