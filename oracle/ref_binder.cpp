// TEST INFRASTRUCTURE ONLY -- binder for the reference's own CPU kernels.
//
// The reference CPU sources (sgl-kernel/csrc/cpu/{decode,extend,gemm,gemm_fp8,gemm_int8}.cpp)
// are compiled unmodified from /root/reference by oracle/build_ref.py; this file only
// exposes their entry points to torch.ops under the namespace `sgl_ref` so that the
// tests and bench.py's cpu_baseline leg can call the reference itself.
// Entry points bound here are the ones declared in
// sgl-kernel/csrc/cpu/torch_extension_cpu.cpp:65-133 (reference).
#include <ATen/ATen.h>
#include <torch/library.h>

#include <optional>
#include <tuple>
#include <vector>

using at::Tensor;
using OptTensor = const std::optional<at::Tensor>&;

void decode_attention_cpu(Tensor&, Tensor&, Tensor&, Tensor&, Tensor&, Tensor&, Tensor&, Tensor&,
                          Tensor&, Tensor&, Tensor&, double, double);
void extend_attention_cpu(Tensor&, Tensor&, Tensor&, Tensor&, Tensor&, Tensor&, Tensor&, Tensor&,
                          Tensor&, Tensor&, Tensor&, int64_t, double, double);
Tensor convert_weight_packed(Tensor&);
std::tuple<Tensor, Tensor> per_token_quant_int8_cpu(Tensor&);
Tensor weight_packed_linear(Tensor&, Tensor&, OptTensor, bool);
Tensor int8_scaled_mm_cpu(Tensor&, Tensor&, Tensor&, Tensor&, OptTensor, at::ScalarType, bool);
Tensor fp8_scaled_mm_cpu(Tensor&, Tensor&, Tensor&, std::vector<int64_t>, OptTensor, at::ScalarType, bool);
Tensor int8_scaled_mm_with_quant(Tensor&, Tensor&, Tensor&, OptTensor, at::ScalarType, bool);

TORCH_LIBRARY_FRAGMENT(sgl_ref, m) {
  m.def(
      "decode_attention_cpu(Tensor query, Tensor k_cache, Tensor v_cache, Tensor(a!) output, Tensor key, "
      "Tensor value, Tensor loc, Tensor(b!) attn_logits, Tensor req_to_token, Tensor req_pool_indices, "
      "Tensor seq_lens, float sm_scale, float logit_cap) -> ()");
  m.impl("decode_attention_cpu", c10::kCPU, &decode_attention_cpu);
  m.def(
      "extend_attention_cpu(Tensor q_extend, Tensor k_extend, Tensor v_extend, Tensor(a!) o_extend, "
      "Tensor k_buffer, Tensor v_buffer, Tensor req_to_token, Tensor req_pool_indices, Tensor seq_lens, "
      "Tensor extend_seq_lens, Tensor extend_start_loc, int max_len_extend, float sm_scale, "
      "float logit_cap) -> ()");
  m.impl("extend_attention_cpu", c10::kCPU, &extend_attention_cpu);
  m.def("convert_weight_packed(Tensor weight) -> Tensor");
  m.impl("convert_weight_packed", c10::kCPU, &convert_weight_packed);
  m.def("per_token_quant_int8_cpu(Tensor A) -> (Tensor, Tensor)");
  m.impl("per_token_quant_int8_cpu", c10::kCPU, &per_token_quant_int8_cpu);
  m.def("weight_packed_linear(Tensor mat1, Tensor mat2, Tensor? bias, bool is_vnni) -> Tensor");
  m.impl("weight_packed_linear", c10::kCPU, &weight_packed_linear);
  m.def(
      "int8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales1, Tensor scales2, Tensor? bias, "
      "ScalarType out_dtype, bool is_vnni) -> Tensor");
  m.impl("int8_scaled_mm_cpu", c10::kCPU, &int8_scaled_mm_cpu);
  m.def(
      "fp8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales2, int[] block_size, Tensor? bias, "
      "ScalarType out_dtype, bool is_vnni) -> Tensor");
  m.impl("fp8_scaled_mm_cpu", c10::kCPU, &fp8_scaled_mm_cpu);
  m.def(
      "int8_scaled_mm_with_quant(Tensor mat1, Tensor mat2, Tensor scales2, Tensor? bias, "
      "ScalarType out_dtype, bool is_vnni) -> Tensor");
  m.impl("int8_scaled_mm_with_quant", c10::kCPU, &int8_scaled_mm_with_quant);
}
