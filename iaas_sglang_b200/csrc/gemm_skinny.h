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
  void* out;
  int64_t ldo;
  int out_bf16;
  int M, N, K, mt_pad;
  int n_tiles, kb_total;
  long long units;
  int stages;
  int debug;  // developer knobs (B200_SK_DEBUG): 1 = skip the split-tile fix-up, 2 = skip the fp8 promotion
  float* ws_partial;  // [grid][2][mt_pad][128] fp32 (int32 bits in MODE 1)
  int* ws_counter;    // [n_tiles] tickets, zero on entry, self-resetting
};

size_t skinny_workspace_bytes(int M, int N, int K);
bool skinny_applicable(int M);
// a [M,K] activations, w [N,K] weight (both K-major, 1-byte elements); MODE 0 fp8 block-scaled, 1 int8
template <int MODE>
int launch_skinny(const void* a, int64_t lda, const void* w, int64_t ldw, SkinnyParams& sp, void* workspace,
                  size_t workspace_bytes, cudaStream_t st);

}  // namespace b200
