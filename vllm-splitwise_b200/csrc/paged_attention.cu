// Host side of b200_paged_attention_v1/v2: argument checks, grid/partition planning, dispatch.
// Replaces the reference launchers (csrc/attention/attention_kernels.cu:606-747, 784-940).
#include <algorithm>

#include <stdlib.h>

#include "paged_attention_mma.cuh"

namespace b200 {
namespace {

constexpr int kV2PartitionTokens = 512;  // reference PARTITION_SIZE (attention_kernels.cu:791, paged_attn.py:12)

int check_common(const b200_paged_attn_args_t* a) {
  if (!a) return B200_ERR_ARG;
  if (a->num_seqs == 0) return B200_OK;
  if (!a->out || !a->query || !a->key_cache || !a->value_cache || !a->block_tables || !a->context_lens)
    return B200_ERR_ARG;
  if (a->num_seqs < 0 || a->num_heads <= 0 || a->num_kv_heads <= 0 || a->num_heads % a->num_kv_heads != 0)
    return B200_ERR_ARG;
  switch (a->head_size) {
    case 64: case 80: case 96: case 112: case 128: case 256: break;
    default: return B200_ERR_HEAD_SIZE;
  }
  switch (a->block_size) {
    case 8: case 16: case 32: break;
    default: return B200_ERR_BLOCK_SIZE;
  }
  if (a->query_dtype != B200_F32 && a->query_dtype != B200_F16 && a->query_dtype != B200_BF16)
    return B200_ERR_DTYPE;
  if (a->cache_dtype != a->query_dtype && a->cache_dtype != B200_U8_E5M2) return B200_ERR_DTYPE;
  const int cs = dtype_size(a->cache_dtype);
  // bulk copies need 16-byte aligned chunks
  if ((reinterpret_cast<uintptr_t>(a->key_cache) % 16) || (reinterpret_cast<uintptr_t>(a->value_cache) % 16) ||
      ((a->kv_block_stride * cs) % 16) || ((a->kv_head_stride * cs) % 16))
    return B200_ERR_ALIGN;
  return B200_OK;
}

AttnParams base_params(const b200_paged_attn_args_t* a) {
  AttnParams p{};
  p.out = a->out;
  p.q = a->query;
  p.k_cache = a->key_cache;
  p.v_cache = a->value_cache;
  p.block_tables = a->block_tables;
  p.context_lens = a->context_lens;
  p.alibi_slopes = a->alibi_slopes;
  p.scale = a->scale;
  p.num_heads = a->num_heads;
  p.num_kv_heads = a->num_kv_heads;
  p.group = a->num_heads / a->num_kv_heads;
  p.max_num_blocks_per_seq = a->max_num_blocks_per_seq;
  p.q_stride = a->q_stride;
  p.kv_block_stride = a->kv_block_stride;
  p.kv_head_stride = a->kv_head_stride;
  return p;
}

// Which consumer path a call takes.  Grouped queries (more than one query head per KV head) in f16/bf16
// with 16- or 32-token blocks go to the tensor-core kernel; everything else (MHA, fp32, block_size 8)
// stays on the SIMT kernel, which is already at the HBM ceiling for MHA.
struct TilePlan {
  bool mma;
  bool hi;   // mma: tile rows 8..15 in use (more than 8 query heads per KV head)
  int gt;    // query heads per CTA
};

TilePlan plan_tile(const b200_paged_attn_args_t* a) {
  const int group = a->num_heads / a->num_kv_heads;
  TilePlan t{false, false, group == 1 ? 1 : 8};
  static const bool force_simt = getenv("B200_ATTN_FORCE_SIMT") != nullptr;  // A/B knob for benchmarks and tests
  if (!force_simt && group > 1 && (a->query_dtype == B200_F16 || a->query_dtype == B200_BF16) &&
      (a->block_size == 16 || a->block_size == 32)) {
    t.mma = true;
    t.hi = group > 8;
    t.gt = t.hi ? 16 : 8;
  }
  return t;
}

int dispatch_mma(const b200_paged_attn_args_t* a, const AttnParams& p, bool hi, dim3 grid, int smem_budget,
                 cudaStream_t stream) {
  const bool fp8 = a->cache_dtype == B200_U8_E5M2;
  if (a->query_dtype == B200_F16)
    return fp8 ? launch_mma<__half, uint8_t>(p, a->head_size, a->block_size, hi, grid, smem_budget, stream)
               : launch_mma<__half, __half>(p, a->head_size, a->block_size, hi, grid, smem_budget, stream);
  if (a->query_dtype == B200_BF16)
    return fp8 ? launch_mma<__nv_bfloat16, uint8_t>(p, a->head_size, a->block_size, hi, grid, smem_budget, stream)
               : launch_mma<__nv_bfloat16, __nv_bfloat16>(p, a->head_size, a->block_size, hi, grid, smem_budget, stream);
  return B200_ERR_DTYPE;
}

int dispatch_simt(const b200_paged_attn_args_t* a, const AttnParams& p, int gt, dim3 grid,
                  int smem_budget, cudaStream_t stream) {
  const bool fp8 = a->cache_dtype == B200_U8_E5M2;
  switch (a->query_dtype) {
    case B200_F32:
      return fp8 ? launch_simt<float, uint8_t>(p, a->head_size, a->block_size, gt, grid, smem_budget, stream)
                 : launch_simt<float, float>(p, a->head_size, a->block_size, gt, grid, smem_budget, stream);
    case B200_F16:
      return fp8 ? launch_simt<__half, uint8_t>(p, a->head_size, a->block_size, gt, grid, smem_budget, stream)
                 : launch_simt<__half, __half>(p, a->head_size, a->block_size, gt, grid, smem_budget, stream);
    case B200_BF16:
      return fp8 ? launch_simt<__nv_bfloat16, uint8_t>(p, a->head_size, a->block_size, gt, grid, smem_budget, stream)
                 : launch_simt<__nv_bfloat16, __nv_bfloat16>(p, a->head_size, a->block_size, gt, grid, smem_budget, stream);
    default:
      return B200_ERR_DTYPE;
  }
}

int launch_reduce(const b200_paged_attn_args_t* a, void* out, const float* exp_sums, const float* max_logits,
                  const void* tmp_out, int max_parts, int partition_tokens, cudaStream_t stream) {
  dim3 grid(a->num_heads, a->num_seqs);
  const size_t smem = (size_t)max_parts * sizeof(float);
  switch (a->query_dtype) {
    case B200_F32:
      paged_attn_reduce_kernel<float><<<grid, 128, smem, stream>>>((float*)out, exp_sums, max_logits, (const float*)tmp_out,
                                                                   a->context_lens, a->head_size, max_parts, partition_tokens);
      break;
    case B200_F16:
      paged_attn_reduce_kernel<__half><<<grid, 128, smem, stream>>>((__half*)out, exp_sums, max_logits, (const __half*)tmp_out,
                                                                    a->context_lens, a->head_size, max_parts, partition_tokens);
      break;
    case B200_BF16:
      paged_attn_reduce_kernel<__nv_bfloat16><<<grid, 128, smem, stream>>>((__nv_bfloat16*)out, exp_sums, max_logits,
                                                                           (const __nv_bfloat16*)tmp_out, a->context_lens,
                                                                           a->head_size, max_parts, partition_tokens);
      break;
    default:
      return B200_ERR_DTYPE;
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// Shared-memory budget per CTA: aim for two resident CTAs per SM (2 x ~110 KiB of the 227 KiB),
// fall back to one when a single ring stage is too large for that (fp32, D=256, bs=32).
int pick_smem_budget(const b200_paged_attn_args_t* a, const TilePlan& t) {
  const int stage = 2 * a->head_size * a->block_size * dtype_size(a->cache_dtype);
  const int gt = t.gt;
  const int fixed = t.mma ? kConsumerWarps * kMmaHeads * 8 + 512
                          : gt * a->head_size * 4 + kConsumerWarps * gt * a->block_size * 4 +
                                kConsumerWarps * gt * a->head_size * 4 + kConsumerWarps * gt * 8 + 512;
  const int two = 112 * 1024;
  const int merge = t.mma ? kConsumerWarps * kMmaHeads * a->head_size * 4 : 0;  // aliased onto the ring
  if ((two - fixed) / stage >= 4 && (two - fixed) / stage / 4 * 4 * stage >= merge) return two;
  return 226 * 1024;
}

int dispatch(const b200_paged_attn_args_t* a, const AttnParams& p, const TilePlan& t, dim3 grid, cudaStream_t stream) {
  const int budget = pick_smem_budget(a, t);
  return t.mma ? dispatch_mma(a, p, t.hi, grid, budget, stream) : dispatch_simt(a, p, t.gt, grid, budget, stream);
}

// Internal split of a v1 call (needs caller workspace): choose the partition length so that the
// grid covers the SMs a few times over; 0 = do not split.
int plan_v1_split(const b200_paged_attn_args_t* a, int gtiles, int sms, int* partition_blocks, int* num_parts) {
  *partition_blocks = 0;
  *num_parts = 1;
  const long long ctas = (long long)a->num_seqs * a->num_kv_heads * gtiles;
  const int max_blocks = (a->max_context_len + a->block_size - 1) / a->block_size;
  if (!a->workspace || ctas >= 2LL * sms * 2 || max_blocks < 32) return 0;
  int want = (int)std::min<long long>((4LL * sms + ctas - 1) / ctas, 64);
  int pb = std::max(16, (max_blocks + want - 1) / want);
  int np = (max_blocks + pb - 1) / pb;
  if (np <= 1) return 0;
  const long long need = (long long)a->num_seqs * a->num_heads * np *
                         ((long long)a->head_size * dtype_size(a->query_dtype) + 8);
  if (need > a->workspace_bytes) return 0;
  *partition_blocks = pb;
  *num_parts = np;
  return 1;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

B200_API int64_t b200_paged_attention_workspace_bytes(const b200_paged_attn_args_t* a) {
  if (!a || a->num_seqs <= 0 || a->num_heads <= 0) return 0;
  // upper bound used by plan_v1_split: at most 64 partitions
  return (int64_t)a->num_seqs * a->num_heads * 64 * ((int64_t)a->head_size * dtype_size(a->query_dtype) + 8);
}

B200_API int b200_paged_attention_v1(const b200_paged_attn_args_t* a, b200_stream_t stream_) {
  int st = check_common(a);
  if (st != B200_OK || a->num_seqs == 0) return st;
  DeviceGuard guard;
  B200_CUDA_TRY(guard.set(a->device));
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AttnParams p = base_params(a);
  const TilePlan tile = plan_tile(a);
  const int gt = tile.gt;
  p.gtiles = (p.group + gt - 1) / gt;
  int pb = 0, np = 1;
  if (plan_v1_split(a, p.gtiles, num_sms(a->device), &pb, &np)) {
    // carve the workspace: tmp_out | exp_sums | max_logits
    const long long rows = (long long)a->num_seqs * a->num_heads * np;
    unsigned char* ws = static_cast<unsigned char*>(a->workspace);
    p.tmp_out = ws;
    long long off = rows * a->head_size * dtype_size(a->query_dtype);
    off = (off + 15) / 16 * 16;
    p.exp_sums = reinterpret_cast<float*>(ws + off);
    p.max_logits = p.exp_sums + rows;
    p.partition_blocks = pb;
    p.max_num_partitions = np;
    p.write_partials = 1;
    dim3 grid(a->num_kv_heads * p.gtiles, a->num_seqs, np);
    st = dispatch(a, p, tile, grid, stream);
    if (st != B200_OK) return st;
    return launch_reduce(a, a->out, p.exp_sums, p.max_logits, p.tmp_out, np, pb * a->block_size, stream);
  }
  p.partition_blocks = 0;
  p.max_num_partitions = 1;
  p.write_partials = 0;
  dim3 grid(a->num_kv_heads * p.gtiles, a->num_seqs, 1);
  return dispatch(a, p, tile, grid, stream);
}

B200_API int b200_paged_attention_v2(const b200_paged_attn_args_t* a, b200_stream_t stream_) {
  int st = check_common(a);
  if (st != B200_OK || a->num_seqs == 0) return st;
  if (!a->exp_sums || !a->max_logits || !a->tmp_out || a->max_num_partitions <= 0) return B200_ERR_ARG;
  if (kV2PartitionTokens % a->block_size != 0) return B200_ERR_BLOCK_SIZE;
  DeviceGuard guard;
  B200_CUDA_TRY(guard.set(a->device));
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AttnParams p = base_params(a);
  const TilePlan tile = plan_tile(a);
  const int gt = tile.gt;
  p.gtiles = (p.group + gt - 1) / gt;
  p.exp_sums = a->exp_sums;
  p.max_logits = a->max_logits;
  p.tmp_out = a->tmp_out;
  p.partition_blocks = kV2PartitionTokens / a->block_size;
  p.max_num_partitions = a->max_num_partitions;
  p.write_partials = 1;
  // partitions actually needed for max_context_len (never more than the caller's scratch holds)
  int np = (a->max_context_len + kV2PartitionTokens - 1) / kV2PartitionTokens;
  np = std::max(1, std::min(np, a->max_num_partitions));
  dim3 grid(a->num_kv_heads * p.gtiles, a->num_seqs, np);
  st = dispatch(a, p, tile, grid, stream);
  if (st != B200_OK) return st;
  return launch_reduce(a, a->out, a->exp_sums, a->max_logits, a->tmp_out, a->max_num_partitions,
                       kV2PartitionTokens, stream);
}

}  // extern "C"
