// One decode step of the hot path for one attention layer, as three programmatically chained launches:
//   decode metadata (b200_decode_metadata_step)  <- ModelRunner._prepare_decode, vllm/worker/model_runner.py:274-399
//   RoPE + cache write                           <- rotary_emb + reshape_and_cache, vllm/model_executor/models/llama.py:156-158,
//                                                   .../attention/ops/paged_attn.py:21-36
//   paged attention v1 / v2                      <- PagedAttentionImpl.forward_decode, paged_attn.py:38-111
// The reference captures exactly this sequence (per layer) into its decode CUDA graphs (model_runner.py:826-873);
// here it is also ONE host call, and the 2nd and 3rd kernels are launched with programmatic stream serialization:
// each kernel releases its successor as soon as it has started (griddepcontrol.launch_dependents) and every kernel
// orders itself after its predecessor's memory with griddepcontrol.wait, so launch latency and kernel prologues
// (barrier init, shared-memory carve-up) overlap the predecessor instead of adding to an ~100 us step.
#include "common.cuh"

using namespace b200;

extern "C" B200_API int b200_decode_step(const b200_decode_step_args_t* a, b200_stream_t stream) {
  if (!a) return B200_ERR_ARG;
  const b200_paged_attn_args_t& at = a->attn;
  if (at.num_seqs == 0) return B200_OK;
  if (at.num_seqs < 0 || !a->key || !a->value || !a->slot_mapping || !at.query || !at.key_cache || !at.value_cache)
    return B200_ERR_ARG;
  const int cs = dtype_size(at.cache_dtype);
  if (cs == 0) return B200_ERR_DTYPE;
  DeviceGuard guard;
  B200_CUDA_TRY(guard.set(at.device));
  int st = B200_OK;
  if (a->state_block_tables != nullptr) {
    if (!a->positions) return B200_ERR_ARG;
    st = b200_decode_metadata_step(a->state_block_tables, a->state_seq_lens, a->batch_rows, at.num_seqs,
                                   a->state_max_blocks_per_seq, at.block_size, a->sliding_window, a->positions,
                                   a->slot_mapping, const_cast<int32_t*>(at.context_lens),
                                   const_cast<int32_t*>(at.block_tables), at.max_num_blocks_per_seq, stream);
    if (st != B200_OK) return st;
  }
  {
    PdlScope pdl(true);
    if (a->cos_sin_cache != nullptr) {
      if (!a->positions) return B200_ERR_ARG;
      st = b200_rotary_embedding_reshape_and_cache(
          a->positions, const_cast<void*>(at.query), a->key, a->value, a->cos_sin_cache, at.num_seqs, at.num_heads,
          at.num_kv_heads, at.head_size, a->rot_dim, at.q_stride, a->key_stride, a->value_stride, a->is_neox,
          const_cast<void*>(at.key_cache), const_cast<void*>(at.value_cache), a->slot_mapping, at.block_size, 16 / cs,
          a->num_blocks, at.query_dtype, at.cache_dtype, stream);
    } else {
      st = b200_reshape_and_cache(a->key, a->value, const_cast<void*>(at.key_cache), const_cast<void*>(at.value_cache),
                                  a->slot_mapping, at.num_seqs, at.num_kv_heads, at.head_size, at.block_size, 16 / cs,
                                  a->key_stride, a->value_stride, a->num_blocks, at.query_dtype, at.cache_dtype, stream);
    }
  }
  if (st != B200_OK) return st;
  // the attention launch is always a programmatic dependent (launch_attn)
  return at.exp_sums != nullptr ? b200_paged_attention_v2(&at, stream) : b200_paged_attention_v1(&at, stream);
}
