// TEST INFRASTRUCTURE — op registration shim for the reference's own CPU kernels.
//
// The kernels themselves are compiled straight from /root/reference/sgl-kernel/csrc/cpu/
// {gemm,gemm_int8,gemm_fp8,decode,extend}.cpp (see oracle/build_ref.py; nothing is copied into this repo).
// This file only declares their entry points (signatures as in the reference's
// torch_extension_cpu.cpp:65-133) and registers them under the `b200_ref` torch library so the
// tests / the CPU-baseline leg of bench.py can call torch.ops.b200_ref.*.
#include <ATen/ATen.h>
#include <torch/library.h>

#include <optional>
#include <tuple>
#include <vector>

void decode_attention_cpu(at::Tensor& query, at::Tensor& k_cache, at::Tensor& v_cache, at::Tensor& output,
                          at::Tensor& key, at::Tensor& value, at::Tensor& loc, at::Tensor& attn_logits,
                          at::Tensor& req_to_token, at::Tensor& req_pool_indices, at::Tensor& seq_lens,
                          double sm_scale, double logit_cap);
void extend_attention_cpu(at::Tensor& q_extend, at::Tensor& k_extend, at::Tensor& v_extend, at::Tensor& o_extend,
                          at::Tensor& k_buffer, at::Tensor& v_buffer, at::Tensor& req_to_token,
                          at::Tensor& req_pool_indices, at::Tensor& seq_lens, at::Tensor& extend_seq_lens,
                          at::Tensor& extend_start_loc, int64_t max_len_extend, double sm_scale, double logit_cap);
at::Tensor convert_weight_packed(at::Tensor& weight);
std::tuple<at::Tensor, at::Tensor> per_token_quant_int8_cpu(at::Tensor& A);
at::Tensor int8_scaled_mm_cpu(at::Tensor& mat1, at::Tensor& mat2, at::Tensor& scales1, at::Tensor& scales2,
                              const std::optional<at::Tensor>& bias, at::ScalarType out_dtype, bool is_vnni);
at::Tensor int8_scaled_mm_with_quant(at::Tensor& mat1, at::Tensor& mat2, at::Tensor& scales2,
                                     const std::optional<at::Tensor>& bias, at::ScalarType out_dtype, bool is_vnni);

// block-scaled FP8 linear of the CPU backend (W8A16: bf16 activations, e4m3 weights; gemm_fp8.cpp:468-548)
at::Tensor fp8_scaled_mm_cpu(at::Tensor& mat1, at::Tensor& mat2, at::Tensor& scales2, std::vector<int64_t> block_size,
                             const std::optional<at::Tensor>& bias, at::ScalarType out_dtype, bool is_vnni);

TORCH_LIBRARY_FRAGMENT(b200_ref, m) {
  m.def(
      "fp8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales2, int[] block_size, Tensor? bias, ScalarType "
      "out_dtype, bool is_vnni) -> Tensor");
  m.impl("fp8_scaled_mm_cpu", c10::kCPU, &fp8_scaled_mm_cpu);
  m.def(
      "decode_attention_cpu(Tensor query, Tensor k_cache, Tensor v_cache, Tensor output, Tensor key, Tensor value, "
      "Tensor loc, Tensor attn_logits, Tensor req_to_token, Tensor req_pool_indices, Tensor seq_lens, float sm_scale, "
      "float logit_cap) -> ()");
  m.impl("decode_attention_cpu", c10::kCPU, &decode_attention_cpu);
  m.def(
      "extend_attention_cpu(Tensor q_extend, Tensor k_extend, Tensor v_extend, Tensor o_extend, Tensor k_buffer, "
      "Tensor v_buffer, Tensor req_to_token, Tensor req_pool_indices, Tensor seq_lens, Tensor extend_seq_lens, Tensor "
      "extend_start_loc, int max_len_extend, float sm_scale, float logit_cap) -> ()");
  m.impl("extend_attention_cpu", c10::kCPU, &extend_attention_cpu);
  m.def("convert_weight_packed(Tensor weight) -> Tensor");
  m.impl("convert_weight_packed", c10::kCPU, &convert_weight_packed);
  m.def("per_token_quant_int8_cpu(Tensor A) -> (Tensor, Tensor)");
  m.impl("per_token_quant_int8_cpu", c10::kCPU, &per_token_quant_int8_cpu);
  m.def(
      "int8_scaled_mm_cpu(Tensor mat1, Tensor mat2, Tensor scales1, Tensor scales2, Tensor? bias, ScalarType "
      "out_dtype, bool is_vnni) -> Tensor");
  m.impl("int8_scaled_mm_cpu", c10::kCPU, &int8_scaled_mm_cpu);
  m.def(
      "int8_scaled_mm_with_quant(Tensor mat1, Tensor mat2, Tensor scales2, Tensor? bias, ScalarType out_dtype, bool "
      "is_vnni) -> Tensor");
  m.impl("int8_scaled_mm_with_quant", c10::kCPU, &int8_scaled_mm_with_quant);
}
