// TEST INFRASTRUCTURE ONLY -- not part of the shipped product path.
//
// Binding shim for the UNMODIFIED reference CPU paged-attention sources
// (/root/reference/csrc/cpu/cpu_attn.cpp + utils.cpp).  It re-declares the three entry points that
// the reference registers in csrc/cpu/torch_bindings.cpp:577-600 and exposes them as
// torch.ops._b200_refcpu.* so that tests / bench can use the reference's own CPU implementation as
// an oracle and as the timed CPU baseline.  No reference source is copied: the .cpp files are
// compiled where they lie (see oracle/build_ref_cpu_attn.sh).
#include <torch/all.h>
#include <torch/library.h>

#include <optional>
#include <string>

torch::Tensor get_scheduler_metadata(const int64_t num_req, const int64_t num_heads_q,
                                     const int64_t num_heads_kv, const int64_t head_dim,
                                     const torch::Tensor& seq_lens, at::ScalarType dtype,
                                     const torch::Tensor& query_start_loc, const bool causal,
                                     const int64_t window_size, const std::string& isa_hint,
                                     const bool enable_kv_split,
                                     const std::optional<torch::Tensor>& dynamic_causal,
                                     const std::string& kv_cache_dtype);

void cpu_attn_reshape_and_cache(const torch::Tensor& key, const torch::Tensor& value,
                                torch::Tensor& key_cache, torch::Tensor& value_cache,
                                const torch::Tensor& slot_mapping, const std::string& isa,
                                const double k_scale, const double v_scale,
                                const std::string& kv_cache_dtype);

void cpu_attention_with_kv_cache(const torch::Tensor& query, const torch::Tensor& key_cache,
                                 const torch::Tensor& value_cache, torch::Tensor& output,
                                 const torch::Tensor& query_start_loc, const torch::Tensor& seq_lens,
                                 const double scale, const bool causal,
                                 const std::optional<torch::Tensor>& alibi_slopes,
                                 const int64_t sliding_window, const torch::Tensor& block_table,
                                 const double softcap, const torch::Tensor& scheduler_metadata,
                                 const std::optional<torch::Tensor>& s_aux,
                                 const std::optional<torch::Tensor>& dynamic_causal,
                                 const double k_scale, const double v_scale,
                                 const std::string& kv_cache_dtype);

TORCH_LIBRARY(_b200_refcpu, ops) {
  ops.def(
      "get_scheduler_metadata(int num_req, int num_heads_q, int num_heads_kv, int head_dim, "
      "Tensor seq_lens, ScalarType dtype, Tensor query_start_loc, bool casual, int window_size, "
      "str isa_hint, bool enable_kv_split, Tensor? dynamic_causal, str kv_cache_dtype=\"auto\") "
      "-> Tensor",
      &get_scheduler_metadata);
  ops.def(
      "cpu_attn_reshape_and_cache(Tensor key, Tensor value, Tensor(a2!) key_cache, "
      "Tensor(a3!) value_cache, Tensor slot_mapping, str isa, float k_scale=1.0, "
      "float v_scale=1.0, str kv_cache_dtype=\"auto\") -> ()",
      &cpu_attn_reshape_and_cache);
  ops.def(
      "cpu_attention_with_kv_cache(Tensor query, Tensor key_cache, Tensor value_cache, "
      "Tensor(a3!) output, Tensor query_start_loc, Tensor seq_lens, float scale, bool causal, "
      "Tensor? alibi_slopes, SymInt sliding_window_size, Tensor block_table, float softcap, "
      "Tensor scheduler_metadata, Tensor? s_aux, Tensor? dynamic_causal, float k_scale=1.0, "
      "float v_scale=1.0, str kv_cache_dtype=\"auto\") -> ()",
      &cpu_attention_with_kv_cache);
}
