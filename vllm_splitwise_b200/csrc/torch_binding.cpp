// Thin PyTorch front of libb200kv.so: a pybind11 module with the reference's `vllm._C` shape
// (csrc/pybind.cpp:6-129): submodules `ops`, `cache_ops`, `cuda_utils`, the same positional signatures.
// Nothing here computes: each function checks the tensors the way the reference launchers do
// (dtype / device / density), reads pointers, sizes and strides, and calls the C ABI of
// include/b200_kv.h on torch's current stream.  Status codes become c10::Error (Python RuntimeError)
// with the reference's message text.  The ctypes layer in vllm_splitwise_b200/_C/*.py is the same
// binding written in Python; this compiled one exists because a decode step calls two ops per layer and
// the interpreter-level marshalling (~10 us per call) is then comparable to the kernels themselves.
#include <torch/extension.h>
#include <ATen/cuda/CUDAContext.h>
#include <c10/cuda/CUDAGraphsC10Utils.h>
#include <c10/cuda/CUDAGuard.h>

#include <map>
#include <mutex>
#include <tuple>
#include <vector>

#include "../../include/b200_kv.h"

namespace {

int dtype_code(const torch::Tensor& t, const char* what) {
  switch (t.scalar_type()) {
    case at::ScalarType::Float: return B200_F32;
    case at::ScalarType::Half: return B200_F16;
    case at::ScalarType::BFloat16: return B200_BF16;
    case at::ScalarType::Byte: return B200_U8_E5M2;
    default: TORCH_CHECK(false, "Unsupported data type: ", t.scalar_type(), " (", what, ")");
  }
  return -1;
}

void check_status(int st, const char* what, int64_t detail = -1) {
  if (st == B200_OK) return;
  std::string msg = b200_kv_strerror(st);
  if (st == B200_ERR_CUDA) msg += std::string(": ") + b200_kv_last_cuda_error();
  if (detail >= 0 && (st == B200_ERR_HEAD_SIZE || st == B200_ERR_BLOCK_SIZE)) msg += ": " + std::to_string(detail);
  TORCH_CHECK(false, what, ": ", msg);
}

void require_cuda(const torch::Tensor& t, const char* name) {
  TORCH_CHECK(t.is_cuda(), name, " must be a CUDA tensor (there is no CPU path)");
}

int cache_dtype_code(const std::string& kv_cache_dtype, const torch::Tensor& query, const torch::Tensor& cache) {
  if (kv_cache_dtype == "auto") {
    TORCH_CHECK(cache.scalar_type() == query.scalar_type(), "Unsupported data type: cache ", cache.scalar_type(),
                " with query ", query.scalar_type());
    return dtype_code(query, "query");
  }
  if (kv_cache_dtype == "fp8_e5m2") {
    TORCH_CHECK(cache.scalar_type() == at::ScalarType::Byte, "Unsupported data type: fp8_e5m2 cache must be uint8");
    return B200_U8_E5M2;
  }
  TORCH_CHECK(false, "Unsupported data type of kv cache: ", kv_cache_dtype);
  return -1;
}

b200_paged_attn_args_t fill_args(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache,
                                 torch::Tensor& value_cache, int num_kv_heads, float scale,
                                 torch::Tensor& block_tables, torch::Tensor& context_lens, int block_size,
                                 int max_context_len, const c10::optional<torch::Tensor>& alibi_slopes,
                                 const std::string& kv_cache_dtype) {
  require_cuda(out, "out"); require_cuda(query, "query"); require_cuda(key_cache, "key_cache");
  require_cuda(value_cache, "value_cache"); require_cuda(block_tables, "block_tables");
  require_cuda(context_lens, "context_lens");
  TORCH_CHECK(block_tables.scalar_type() == at::ScalarType::Int && context_lens.scalar_type() == at::ScalarType::Int,
              "block_tables and context_lens must be int32");
  TORCH_CHECK(query.dim() == 3 && query.stride(2) == 1 && query.stride(1) == query.size(2) && out.is_contiguous(),
              "query must be dense in its last two dims and out contiguous");
  b200_paged_attn_args_t a{};
  a.out = out.data_ptr();
  a.query = query.data_ptr();
  a.key_cache = key_cache.data_ptr();
  a.value_cache = value_cache.data_ptr();
  a.block_tables = block_tables.data_ptr<int32_t>();
  a.context_lens = context_lens.data_ptr<int32_t>();
  if (alibi_slopes.has_value() && alibi_slopes->defined()) {
    TORCH_CHECK(alibi_slopes->scalar_type() == at::ScalarType::Float, "alibi_slopes must be float32");
    a.alibi_slopes = alibi_slopes->data_ptr<float>();
  }
  a.scale = scale;
  a.num_seqs = (int32_t)query.size(0);
  a.num_heads = (int32_t)query.size(1);
  a.head_size = (int32_t)query.size(2);
  a.num_kv_heads = num_kv_heads;
  a.block_size = block_size;
  a.max_context_len = max_context_len;
  a.max_num_blocks_per_seq = (int32_t)block_tables.size(1);
  a.q_stride = query.stride(0);
  a.kv_block_stride = key_cache.stride(0);
  a.kv_head_stride = key_cache.stride(1);
  a.query_dtype = dtype_code(query, "query");
  a.cache_dtype = cache_dtype_code(kv_cache_dtype, query, key_cache);
  a.device = query.get_device();
  return a;
}

// Scratch for the internal split of v1 calls (partials + reduce tickets; b200_kv.h).  Two calls that may execute
// concurrently must not share it, so there is one buffer per (device, stream) for eager calls and one per
// (device, stream-capture id) for calls recorded into a CUDA graph: a graph replay can then overlap any eager
// call or any other graph.  b200_paged_attention_workspace_bytes() is exact for the call's arguments (0 when the
// planner does not split; a few MB at most).  A buffer that must grow is replaced by a larger zeroed one; the
// old one goes back to torch's caching allocator, which only reuses it in stream order.  Buffers recorded into a
// graph are never replaced (the capture id is unique), so a captured pointer stays valid while the graph lives;
// allocation during capture comes from the graph's private pool, as every other tensor the capture creates.
struct WsKey {
  int device;
  int captured;
  unsigned long long id;  // stream handle (eager) or capture id
  bool operator<(const WsKey& o) const {
    return std::tie(device, captured, id) < std::tie(o.device, o.captured, o.id);
  }
};
std::mutex g_ws_mutex;
std::map<WsKey, torch::Tensor>& workspaces() {
  static auto* m = new std::map<WsKey, torch::Tensor>();  // never destroyed: tensors must not outlive-race CUDA teardown
  return *m;
}

void attach_workspace(b200_paged_attn_args_t& a, const torch::Tensor& like) {
  if (a.max_context_len < 256 || a.exp_sums != nullptr) return;
  const int64_t need = b200_paged_attention_workspace_bytes(&a);
  if (need <= 0) return;
  const cudaStream_t stream = at::cuda::getCurrentCUDAStream().stream();
  WsKey key{a.device, 0, (unsigned long long)reinterpret_cast<uintptr_t>(stream)};
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  unsigned long long cap_id = 0;
  if (cudaStreamGetCaptureInfo(stream, &cs, &cap_id) == cudaSuccess && cs == cudaStreamCaptureStatusActive)
    key = WsKey{a.device, 1, cap_id};
  else
    (void)cudaGetLastError();
  std::lock_guard<std::mutex> lock(g_ws_mutex);
  torch::Tensor& ws = workspaces()[key];
  if (!ws.defined() || ws.numel() < need)
    ws = torch::zeros({std::max<int64_t>(need, 1 << 20)}, like.options().dtype(torch::kUInt8));
  a.workspace = ws.data_ptr();
  a.workspace_bytes = ws.numel();
}

void paged_attention_v1(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key_cache, torch::Tensor& value_cache,
                        int num_kv_heads, float scale, torch::Tensor& block_tables, torch::Tensor& context_lens,
                        int block_size, int max_context_len, const c10::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype) {
  b200_paged_attn_args_t a = fill_args(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                                       context_lens, block_size, max_context_len, alibi_slopes, kv_cache_dtype);
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  attach_workspace(a, query);
  const int st = b200_paged_attention_v1(&a, at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "paged_attention_v1", st == B200_ERR_HEAD_SIZE ? a.head_size : a.block_size);
}

void paged_attention_v2(torch::Tensor& out, torch::Tensor& exp_sums, torch::Tensor& max_logits, torch::Tensor& tmp_out,
                        torch::Tensor& query, torch::Tensor& key_cache, torch::Tensor& value_cache, int num_kv_heads,
                        float scale, torch::Tensor& block_tables, torch::Tensor& context_lens, int block_size,
                        int max_context_len, const c10::optional<torch::Tensor>& alibi_slopes,
                        const std::string& kv_cache_dtype) {
  b200_paged_attn_args_t a = fill_args(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables,
                                       context_lens, block_size, max_context_len, alibi_slopes, kv_cache_dtype);
  TORCH_CHECK(exp_sums.scalar_type() == at::ScalarType::Float && max_logits.scalar_type() == at::ScalarType::Float,
              "exp_sums and max_logits must be float32");
  TORCH_CHECK(tmp_out.scalar_type() == query.scalar_type(), "tmp_out must have the query dtype");
  TORCH_CHECK(exp_sums.is_contiguous() && max_logits.is_contiguous() && tmp_out.is_contiguous(),
              "v2 scratch tensors must be contiguous");
  a.exp_sums = exp_sums.data_ptr<float>();
  a.max_logits = max_logits.data_ptr<float>();
  a.tmp_out = tmp_out.data_ptr();
  a.max_num_partitions = (int32_t)exp_sums.size(2);
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const int st = b200_paged_attention_v2(&a, at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "paged_attention_v2", st == B200_ERR_HEAD_SIZE ? a.head_size : a.block_size);
}

void check_kv_source(const torch::Tensor& t, const char* name) {
  require_cuda(t, name);
  TORCH_CHECK(t.dim() == 3 && t.stride(2) == 1 && t.stride(1) == t.size(2), "key/value must be dense in their last two dims");
}

void reshape_and_cache(torch::Tensor& key, torch::Tensor& value, torch::Tensor& key_cache, torch::Tensor& value_cache,
                       torch::Tensor& slot_mapping, const std::string& kv_cache_dtype) {
  check_kv_source(key, "key"); check_kv_source(value, "value");
  require_cuda(key_cache, "key_cache"); require_cuda(value_cache, "value_cache"); require_cuda(slot_mapping, "slot_mapping");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous(),
              "slot_mapping must be a contiguous int64 tensor");
  const int cache_code = cache_dtype_code(kv_cache_dtype, key, key_cache);
  TORCH_CHECK(key_cache.is_contiguous() && value_cache.is_contiguous(), "key_cache/value_cache must be contiguous");
  const at::cuda::OptionalCUDAGuard guard(device_of(key));
  const int st = b200_reshape_and_cache(
      key.data_ptr(), value.data_ptr(), key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr<int64_t>(),
      (int32_t)key.size(0), (int32_t)key.size(1), (int32_t)key.size(2), (int32_t)key_cache.size(3),
      (int32_t)key_cache.size(4), key.stride(0), value.stride(0), key_cache.size(0), dtype_code(key, "key"), cache_code,
      at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "reshape_and_cache");
}

void copy_blocks(std::vector<torch::Tensor>& key_caches, std::vector<torch::Tensor>& value_caches,
                 const std::map<int64_t, std::vector<int64_t>>& block_mapping) {
  const int num_layers = (int)key_caches.size();
  TORCH_CHECK(num_layers == (int)value_caches.size(), "key_caches and value_caches must have the same length");
  if (num_layers == 0) return;
  require_cuda(key_caches[0], "key_caches[0]");
  std::vector<int64_t> kp(num_layers), vp(num_layers), pairs;
  for (int l = 0; l < num_layers; ++l) {
    kp[l] = reinterpret_cast<int64_t>(key_caches[l].data_ptr());
    vp[l] = reinterpret_cast<int64_t>(value_caches[l].data_ptr());
  }
  for (const auto& kv : block_mapping)  // std::map: ascending source block (reference cache_kernels.cu:119-126)
    for (int64_t dst : kv.second) {
      pairs.push_back(kv.first);
      pairs.push_back(dst);
    }
  if (pairs.empty()) return;
  const int64_t block_bytes = key_caches[0][0].numel() * key_caches[0].element_size();
  const at::cuda::OptionalCUDAGuard guard(device_of(key_caches[0]));
  const int st = b200_copy_blocks(kp.data(), vp.data(), num_layers, pairs.data(), (int32_t)(pairs.size() / 2), block_bytes,
                                  at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "copy_blocks");
}

void swap_blocks(torch::Tensor& src, torch::Tensor& dst, const std::map<int64_t, int64_t>& block_mapping) {
  int kind;
  torch::Device dev = src.device();
  if (src.is_cuda() && dst.is_cuda()) {
    TORCH_CHECK(src.get_device() == dst.get_device(), "src and dst must be on the same GPU");
    kind = B200_SWAP_D2D;
  } else if (src.is_cuda() && dst.is_cpu()) {
    kind = B200_SWAP_D2H;
  } else if (src.is_cpu() && dst.is_cuda()) {
    kind = B200_SWAP_H2D;
    dev = dst.device();
  } else {
    TORCH_CHECK(false, "Invalid device combination");
  }
  std::vector<int64_t> pairs;
  pairs.reserve(2 * block_mapping.size());
  for (const auto& kv : block_mapping) {
    pairs.push_back(kv.first);
    pairs.push_back(kv.second);
  }
  if (pairs.empty()) return;
  const int64_t block_bytes = src.element_size() * src[0].numel();
  const at::cuda::CUDAGuard guard(dev);
  const int st = b200_swap_blocks(src.data_ptr(), dst.data_ptr(), block_bytes, pairs.data(), (int32_t)(pairs.size() / 2),
                                  kind, src.is_cuda() ? src.get_device() : -1, dst.is_cuda() ? dst.get_device() : -1,
                                  at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "swap_blocks");
}

void convert_fp8_e5m2(torch::Tensor& src_cache, torch::Tensor& dst_cache) {
  require_cuda(src_cache, "src_cache"); require_cuda(dst_cache, "dst_cache");
  TORCH_CHECK(src_cache.numel() == dst_cache.numel(), "src_cache and dst_cache must have the same number of elements");
  TORCH_CHECK(src_cache.is_contiguous() && dst_cache.is_contiguous(), "convert_fp8_e5m2 needs contiguous tensors");
  const at::cuda::OptionalCUDAGuard guard(device_of(src_cache));
  const int st = b200_convert_fp8_e5m2(src_cache.data_ptr(), dst_cache.data_ptr(), src_cache.numel(),
                                       dtype_code(src_cache, "src_cache"), dtype_code(dst_cache, "dst_cache"),
                                       at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "convert_fp8_e5m2");
}

// ---- rotary embedding (reference csrc/ops.h:48-63) and its fusion with the cache write ----
struct RotaryShape {
  int64_t num_tokens, query_stride, key_stride;
  int num_heads, num_kv_heads, rot_dim;
};

RotaryShape rotary_shape(const torch::Tensor& positions, const torch::Tensor& query, const torch::Tensor& key,
                         int head_size, const torch::Tensor& cos_sin_cache) {
  require_cuda(positions, "positions"); require_cuda(query, "query"); require_cuda(key, "key");
  require_cuda(cos_sin_cache, "cos_sin_cache");
  TORCH_CHECK(positions.scalar_type() == at::ScalarType::Long && positions.is_contiguous(),
              "positions must be a contiguous int64 tensor");
  TORCH_CHECK(query.dim() >= 2 && key.dim() >= 2 && head_size > 0 && query.size(-1) % head_size == 0 &&
                  key.size(-1) % head_size == 0 && query.stride(-1) == 1 && key.stride(-1) == 1,
              "query/key must be [..., num_heads * head_size] with a dense last dim");
  TORCH_CHECK(key.scalar_type() == query.scalar_type() && cos_sin_cache.scalar_type() == query.scalar_type(),
              "query, key and cos_sin_cache must have the same dtype");
  TORCH_CHECK(cos_sin_cache.dim() == 2 && cos_sin_cache.is_contiguous(), "cos_sin_cache must be [max_position, rot_dim]");
  RotaryShape r;
  r.num_tokens = query.numel() / query.size(-1);  // reference pos_encoding_kernels.cu:133
  r.num_heads = (int)(query.size(-1) / head_size);
  r.num_kv_heads = (int)(key.size(-1) / head_size);
  r.rot_dim = (int)cos_sin_cache.size(1);
  r.query_stride = query.stride(-2);
  r.key_stride = key.stride(-2);
  TORCH_CHECK(positions.numel() == r.num_tokens && key.numel() / key.size(-1) == r.num_tokens,
              "positions, query and key must describe the same number of tokens");
  return r;
}

void rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int head_size,
                      torch::Tensor& cos_sin_cache, bool is_neox) {
  const RotaryShape r = rotary_shape(positions, query, key, head_size, cos_sin_cache);
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const int st = b200_rotary_embedding(positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(),
                                       cos_sin_cache.data_ptr(), nullptr, r.num_tokens, r.num_heads, r.num_kv_heads,
                                       head_size, r.rot_dim, r.query_stride, r.key_stride, is_neox ? 1 : 0,
                                       dtype_code(query, "query"), at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "rotary_embedding");
}

void batched_rotary_embedding(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key, int head_size,
                              torch::Tensor& cos_sin_cache, bool is_neox, int rot_dim,
                              torch::Tensor& cos_sin_cache_offsets) {
  const RotaryShape r = rotary_shape(positions, query, key, head_size, cos_sin_cache);
  require_cuda(cos_sin_cache_offsets, "cos_sin_cache_offsets");
  TORCH_CHECK(cos_sin_cache_offsets.scalar_type() == at::ScalarType::Long && cos_sin_cache_offsets.is_contiguous() &&
                  cos_sin_cache_offsets.numel() == r.num_tokens,
              "cos_sin_cache_offsets must be a contiguous int64 tensor with one entry per token");
  TORCH_CHECK(rot_dim == r.rot_dim, "rot_dim must equal cos_sin_cache.size(1)");
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const int st = b200_rotary_embedding(positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(),
                                       cos_sin_cache.data_ptr(), cos_sin_cache_offsets.data_ptr<int64_t>(), r.num_tokens,
                                       r.num_heads, r.num_kv_heads, head_size, r.rot_dim, r.query_stride, r.key_stride,
                                       is_neox ? 1 : 0, dtype_code(query, "query"),
                                       at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "batched_rotary_embedding");
}

// rotary_embedding(positions, query, key, ...) then reshape_and_cache(key, value, ...) in one call (one launch for
// decode-sized batches).  query/key/value are [..., heads * head_size] as the model holds them after qkv.split
// (vllm/model_executor/models/llama.py:155-158).
void rotary_embedding_reshape_and_cache(torch::Tensor& positions, torch::Tensor& query, torch::Tensor& key,
                                        torch::Tensor& value, int head_size, torch::Tensor& cos_sin_cache, bool is_neox,
                                        torch::Tensor& key_cache, torch::Tensor& value_cache, torch::Tensor& slot_mapping,
                                        const std::string& kv_cache_dtype) {
  const RotaryShape r = rotary_shape(positions, query, key, head_size, cos_sin_cache);
  require_cuda(value, "value"); require_cuda(key_cache, "key_cache"); require_cuda(value_cache, "value_cache");
  require_cuda(slot_mapping, "slot_mapping");
  TORCH_CHECK(value.dim() == key.dim() && value.size(-1) == key.size(-1) && value.stride(-1) == 1 &&
                  value.scalar_type() == key.scalar_type() && value.numel() == key.numel(),
              "value must have the shape and dtype of key");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous() &&
                  slot_mapping.numel() == r.num_tokens,
              "slot_mapping must be a contiguous int64 tensor with one entry per token");
  TORCH_CHECK(key_cache.is_contiguous() && value_cache.is_contiguous(), "key_cache/value_cache must be contiguous");
  TORCH_CHECK(key_cache.dim() == 5 && key_cache.size(1) == r.num_kv_heads, "key_cache must be [blocks, kv_heads, D/x, bs, x]");
  const int cache_code = cache_dtype_code(kv_cache_dtype, key, key_cache);
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  const int st = b200_rotary_embedding_reshape_and_cache(
      positions.data_ptr<int64_t>(), query.data_ptr(), key.data_ptr(), value.data_ptr(), cos_sin_cache.data_ptr(),
      r.num_tokens, r.num_heads, r.num_kv_heads, head_size, r.rot_dim, r.query_stride, r.key_stride, value.stride(-2),
      is_neox ? 1 : 0, key_cache.data_ptr(), value_cache.data_ptr(), slot_mapping.data_ptr<int64_t>(),
      (int32_t)key_cache.size(3), (int32_t)key_cache.size(4), key_cache.size(0), dtype_code(key, "key"), cache_code,
      at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "rotary_embedding_reshape_and_cache");
}

// One decode step of one attention layer in one call (b200_decode_step): [metadata ->] RoPE + cache write ->
// attention.  query [S, Hq, D], key / value [S, Hkv, D] are the views the model holds after qkv.split
// (vllm/model_executor/models/llama.py:155): dense in their last two dims, any row stride.
void decode_step(torch::Tensor& out, torch::Tensor& query, torch::Tensor& key, torch::Tensor& value,
                 torch::Tensor& key_cache, torch::Tensor& value_cache, int num_kv_heads, float scale,
                 torch::Tensor& block_tables, torch::Tensor& context_lens, int block_size, int max_context_len,
                 const c10::optional<torch::Tensor>& alibi_slopes, const std::string& kv_cache_dtype,
                 torch::Tensor& positions, torch::Tensor& slot_mapping,
                 const c10::optional<torch::Tensor>& cos_sin_cache, bool is_neox,
                 const c10::optional<torch::Tensor>& state_block_tables, const c10::optional<torch::Tensor>& state_seq_lens,
                 const c10::optional<torch::Tensor>& batch_rows, int sliding_window,
                 const c10::optional<torch::Tensor>& exp_sums, const c10::optional<torch::Tensor>& max_logits,
                 const c10::optional<torch::Tensor>& tmp_out) {
  b200_decode_step_args_t d{};
  d.attn = fill_args(out, query, key_cache, value_cache, num_kv_heads, scale, block_tables, context_lens, block_size,
                     max_context_len, alibi_slopes, kv_cache_dtype);
  check_kv_source(key, "key"); check_kv_source(value, "value");
  TORCH_CHECK(key.scalar_type() == query.scalar_type() && value.scalar_type() == query.scalar_type(),
              "query, key and value must have the same dtype");
  TORCH_CHECK(key.size(0) == query.size(0) && value.size(0) == query.size(0) && key.size(1) == num_kv_heads &&
                  value.size(1) == num_kv_heads && key.size(2) == query.size(2) && value.size(2) == query.size(2),
              "key/value must be [num_seqs, num_kv_heads, head_size]");
  require_cuda(positions, "positions"); require_cuda(slot_mapping, "slot_mapping");
  TORCH_CHECK(positions.scalar_type() == at::ScalarType::Long && positions.is_contiguous() &&
                  positions.numel() >= query.size(0),
              "positions must be a contiguous int64 tensor with one entry per sequence");
  TORCH_CHECK(slot_mapping.scalar_type() == at::ScalarType::Long && slot_mapping.is_contiguous() &&
                  slot_mapping.numel() >= query.size(0),
              "slot_mapping must be a contiguous int64 tensor with one entry per sequence");
  TORCH_CHECK(key_cache.is_contiguous() && value_cache.is_contiguous() && block_tables.is_contiguous() &&
                  context_lens.is_contiguous(),
              "key_cache / value_cache / block_tables / context_lens must be contiguous");
  d.positions = positions.data_ptr<int64_t>();
  d.slot_mapping = slot_mapping.data_ptr<int64_t>();
  d.key = key.data_ptr();
  d.value = value.data_ptr();
  d.key_stride = key.stride(0);
  d.value_stride = value.stride(0);
  d.num_blocks = key_cache.size(0);
  d.is_neox = is_neox ? 1 : 0;
  if (cos_sin_cache.has_value() && cos_sin_cache->defined()) {
    require_cuda(*cos_sin_cache, "cos_sin_cache");
    TORCH_CHECK(cos_sin_cache->dim() == 2 && cos_sin_cache->is_contiguous() &&
                    cos_sin_cache->scalar_type() == query.scalar_type(),
                "cos_sin_cache must be a contiguous [max_position, rot_dim] tensor of the query dtype");
    d.cos_sin_cache = cos_sin_cache->data_ptr();
    d.rot_dim = (int32_t)cos_sin_cache->size(1);
  }
  if (state_block_tables.has_value() && state_block_tables->defined()) {
    TORCH_CHECK(state_seq_lens.has_value() && batch_rows.has_value(), "state_seq_lens and batch_rows are required with state_block_tables");
    require_cuda(*state_block_tables, "state_block_tables"); require_cuda(*state_seq_lens, "state_seq_lens");
    require_cuda(*batch_rows, "batch_rows");
    TORCH_CHECK(state_block_tables->scalar_type() == at::ScalarType::Int && state_block_tables->dim() == 2 &&
                    state_block_tables->is_contiguous() && state_seq_lens->scalar_type() == at::ScalarType::Int &&
                    state_seq_lens->is_contiguous() && batch_rows->scalar_type() == at::ScalarType::Int &&
                    batch_rows->is_contiguous() && batch_rows->numel() >= query.size(0),
                "decode state tensors must be contiguous int32");
    d.state_block_tables = state_block_tables->data_ptr<int32_t>();
    d.state_seq_lens = state_seq_lens->data_ptr<int32_t>();
    d.batch_rows = batch_rows->data_ptr<int32_t>();
    d.state_max_blocks_per_seq = (int32_t)state_block_tables->size(1);
    d.sliding_window = sliding_window;
  }
  if (exp_sums.has_value() && exp_sums->defined()) {
    TORCH_CHECK(max_logits.has_value() && tmp_out.has_value(), "v2 scratch: exp_sums, max_logits and tmp_out go together");
    TORCH_CHECK(exp_sums->scalar_type() == at::ScalarType::Float && max_logits->scalar_type() == at::ScalarType::Float &&
                    tmp_out->scalar_type() == query.scalar_type() && exp_sums->is_contiguous() &&
                    max_logits->is_contiguous() && tmp_out->is_contiguous(),
                "v2 scratch tensors must be contiguous (f32, f32, query dtype)");
    d.attn.exp_sums = exp_sums->data_ptr<float>();
    d.attn.max_logits = max_logits->data_ptr<float>();
    d.attn.tmp_out = tmp_out->data_ptr();
    d.attn.max_num_partitions = (int32_t)exp_sums->size(2);
  }
  const at::cuda::OptionalCUDAGuard guard(device_of(query));
  attach_workspace(d.attn, query);
  const int st = b200_decode_step(&d, at::cuda::getCurrentCUDAStream().stream());
  check_status(st, "decode_step", st == B200_ERR_HEAD_SIZE ? d.attn.head_size : d.attn.block_size);
}

}  // namespace

PYBIND11_MODULE(TORCH_EXTENSION_NAME, m) {
  pybind11::module ops = m.def_submodule("ops", "paged attention (reference csrc/ops.h:5-34)");
  ops.def("paged_attention_v1", &paged_attention_v1,
          "Compute the attention between an input query and the cached keys/values using PagedAttention.");
  ops.def("paged_attention_v2", &paged_attention_v2, "PagedAttention V2.");
  ops.def("decode_step", &decode_step,
          "One decode step of one attention layer: [decode metadata ->] RoPE + cache write -> paged attention",
          pybind11::arg("out"), pybind11::arg("query"), pybind11::arg("key"), pybind11::arg("value"),
          pybind11::arg("key_cache"), pybind11::arg("value_cache"), pybind11::arg("num_kv_heads"), pybind11::arg("scale"),
          pybind11::arg("block_tables"), pybind11::arg("context_lens"), pybind11::arg("block_size"),
          pybind11::arg("max_context_len"), pybind11::arg("alibi_slopes"), pybind11::arg("kv_cache_dtype"),
          pybind11::arg("positions"), pybind11::arg("slot_mapping"), pybind11::arg("cos_sin_cache") = pybind11::none(),
          pybind11::arg("is_neox") = true, pybind11::arg("state_block_tables") = pybind11::none(),
          pybind11::arg("state_seq_lens") = pybind11::none(), pybind11::arg("batch_rows") = pybind11::none(),
          pybind11::arg("sliding_window") = 0, pybind11::arg("exp_sums") = pybind11::none(),
          pybind11::arg("max_logits") = pybind11::none(), pybind11::arg("tmp_out") = pybind11::none());
  ops.def("rotary_embedding", &rotary_embedding, "Apply GPT-NeoX or GPT-J style rotary embedding to query and key");
  ops.def("batched_rotary_embedding", &batched_rotary_embedding,
          "Apply GPT-NeoX or GPT-J style rotary embedding to query and key (supports multiple loras)");
  pybind11::module cache_ops = m.def_submodule("cache_ops", "cache writes and block moves (reference csrc/cache.h:8-29)");
  cache_ops.def("swap_blocks", &swap_blocks, "Swap in (out) the cache blocks from src to dst");
  cache_ops.def("copy_blocks", &copy_blocks, "Copy the cache blocks from src to dst");
  cache_ops.def("reshape_and_cache", &reshape_and_cache, "Reshape the key and value tensors and cache them");
  cache_ops.def("convert_fp8_e5m2", &convert_fp8_e5m2, "Convert the key and value cache to fp8_e5m2 data type");
  cache_ops.def("rotary_embedding_reshape_and_cache", &rotary_embedding_reshape_and_cache,
                "rotary_embedding followed by reshape_and_cache in one call (one launch for decode batches)");
  pybind11::module cuda_utils = m.def_submodule("cuda_utils", "device attributes (reference csrc/pybind.cpp:101-111)");
  cuda_utils.def("get_device_attribute",
                 [](int attribute, int device_id) { return b200_get_device_attribute(attribute, device_id); },
                 "Gets the specified device attribute.");
  cuda_utils.def("get_max_shared_memory_per_block_device_attribute",
                 [](int device_id) { return b200_get_max_shared_memory_per_block_device_attribute(device_id); },
                 "Gets the maximum shared memory per block device attribute.");
}
