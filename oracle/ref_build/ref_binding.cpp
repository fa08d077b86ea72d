// TEST INFRASTRUCTURE ONLY -- not part of the product path.
//
// Minimal pybind11 front for the *unmodified* reference CUDA kernels so that the GPU parity
// tests and bench.py can run the reference's own paged_attention / cache ops next to ours on
// identical inputs.  The kernels themselves are compiled straight from
// /root/reference/csrc/{attention/attention_kernels.cu,cache_kernels.cu,cuda_utils_kernels.cu,
// pos_encoding_kernels.cu, custom_all_reduce.cu}
// by oracle/build_ref.sh; nothing from the reference is copied into this repository.  The
// prototypes below restate csrc/ops.h:5-34, csrc/cache.h:8-29 and csrc/cuda_utils.h.
#include <torch/extension.h>

#include <map>
#include <string>
#include <vector>

void paged_attention_v1(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
                        torch::Tensor& value_cache, int num_kv_heads, float scale,
                        torch::Tensor& block_tables, torch::Tensor& context_lens, int block_size,
                        int max_context_len, const c10::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype);

void paged_attention_v2(torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits,
                        torch::Tensor& tmp_out, torch::Tensor& query, torch::Tensor& key_cache,
                        torch::Tensor& value_cache, int num_kv_heads, float scale,
                        torch::Tensor& block_tables, torch::Tensor& context_lens, int block_size,
                        int max_context_len, const c10::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype);

void swap_blocks(torch::Tensor& src, torch::Tensor& dst,
                 const std::map<int64_t, int64_t>& block_mapping);
void copy_blocks(std::vector<torch::Tensor>& key_caches, std::vector<torch::Tensor>& value_caches,
                 const std::map<int64_t, std::vector<int64_t>>& block_mapping);
void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache,
                       torch::Tensor& value_cache, torch::Tensor& slot_mapping,
                       const std::string& kv_cache_dtype);
void convert_fp8_e5m2(torch::Tensor& src_cache, torch::Tensor& dst_cache);

// csrc/ops.h:48-63 (compiled from csrc/pos_encoding_kernels.cu)
void rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int head_size,
                      torch::Tensor& cos_sin_cache, bool is_neox);
void batched_rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int head_size,
                              torch::Tensor& cos_sin_cache, bool is_neox, int rot_dim,
                              torch::Tensor& cos_sin_cache_offsets);

// csrc/ops.h:96-117 (compiled from csrc/custom_all_reduce.cu): the reference's P2P all-reduce, SURVEY 8f row N4
using fptr_t = uint64_t;
fptr_t init_custom_ar(torch::Tensor& meta, torch::Tensor& rank_data, const std::vector<std::string>& handles,
                      const std::vector<int64_t>& offsets, int rank, bool full_nvlink);
bool should_custom_ar(torch::Tensor& inp, int max_size, int world_size, bool full_nvlink);
void all_reduce_reg(fptr_t _fa, torch::Tensor& inp, torch::Tensor& out);
void all_reduce_unreg(fptr_t _fa, torch::Tensor& inp, torch::Tensor& reg_buffer, torch::Tensor& out);
void dispose(fptr_t _fa);
int meta_size();
void register_buffer(fptr_t _fa, torch::Tensor& t, const std::vector<std::string>& handles,
                     const std::vector<int64_t>& offsets);

int get_device_attribute(int attribute, int device_id);
int get_max_shared_memory_per_block_device_attribute(int device_id);

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  pybind11::module ops = m.def_submodule("ops");
  ops.def("paged_attention_v1", &paged_attention_v1);
  ops.def("paged_attention_v2", &paged_attention_v2);
  ops.def("rotary_embedding", &rotary_embedding);
  ops.def("batched_rotary_embedding", &batched_rotary_embedding);
  pybind11::module cache_ops = m.def_submodule("cache_ops");
  cache_ops.def("swap_blocks", &swap_blocks);
  cache_ops.def("copy_blocks", &copy_blocks);
  cache_ops.def("reshape_and_cache", &reshape_and_cache);
  cache_ops.def("convert_fp8_e5m2", &convert_fp8_e5m2);
  pybind11::module custom_ar = m.def_submodule("custom_ar");
  custom_ar.def("init_custom_ar", &init_custom_ar);
  custom_ar.def("should_custom_ar", &should_custom_ar);
  custom_ar.def("all_reduce_reg", &all_reduce_reg);
  custom_ar.def("all_reduce_unreg", &all_reduce_unreg);
  custom_ar.def("dispose", &dispose);
  custom_ar.def("meta_size", &meta_size);
  custom_ar.def("register_buffer", &register_buffer);
  pybind11::module cuda_utils = m.def_submodule("cuda_utils");
  cuda_utils.def("get_device_attribute", &get_device_attribute);
  cuda_utils.def("get_max_shared_memory_per_block_device_attribute",
                 &get_max_shared_memory_per_block_device_attribute);
}
