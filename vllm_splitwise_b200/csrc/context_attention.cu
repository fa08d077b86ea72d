// Host side of b200_context_attention_fwd: argument checks and dispatch.
// Replaces the host function context_attention_fwd of the reference
// (vllm/model_executor/layers/attention/ops/prefix_prefill.py:621-745).
#include <stdlib.h>
#include <string.h>

#include "context_attention.cuh"

extern "C" int b200_context_attention_fwd(const b200_context_attn_args_t* a, b200_stream_t stream) {
  using namespace b200;
  if (!a) return B200_ERR_ARG;
  if (a->batch == 0 || a->max_input_len <= 0) return B200_OK;
  if (!a->out || !a->query || !a->key || !a->value || !a->key_cache || !a->value_cache || !a->block_tables ||
      !a->start_loc || !a->seq_lens || !a->ctx_lens)
    return B200_ERR_ARG;
  if (a->batch < 0 || a->num_heads <= 0 || a->num_kv_heads <= 0 || a->num_heads % a->num_kv_heads != 0)
    return B200_ERR_ARG;
  if (a->dtype != B200_F16 && a->dtype != B200_BF16) return B200_ERR_DTYPE;
  // 16-byte granules everywhere: every row / chunk the kernel copies must start on a 16-byte boundary
  const int64_t strides[] = {a->q_stride_t, a->q_stride_h, a->k_stride_t, a->k_stride_h, a->v_stride_t, a->v_stride_h,
                             a->o_stride_t, a->o_stride_h, a->kc_block_stride, a->kc_head_stride,
                             a->vc_block_stride, a->vc_head_stride};
  for (int64_t s : strides)
    if (s % 8) return B200_ERR_ALIGN;
  const void* ptrs[] = {a->out, a->query, a->key, a->value, a->key_cache, a->value_cache};
  for (const void* q : ptrs)
    if (reinterpret_cast<uintptr_t>(q) % 16) return B200_ERR_ALIGN;

  CtxAttnParams p{};
  p.q = a->query; p.k = a->key; p.v = a->value; p.o = a->out;
  p.k_cache = a->key_cache; p.v_cache = a->value_cache;
  p.b_loc = a->block_tables; p.b_start_loc = a->start_loc; p.b_seq_len = a->seq_lens; p.b_ctx_len = a->ctx_lens;
  p.alibi_slopes = a->alibi_slopes;
  p.scale_log2 = a->scale * 1.4426950408889634f;
  p.q_ts = a->q_stride_t; p.q_hs = a->q_stride_h; p.k_ts = a->k_stride_t; p.k_hs = a->k_stride_h;
  p.v_ts = a->v_stride_t; p.v_hs = a->v_stride_h; p.o_ts = a->o_stride_t; p.o_hs = a->o_stride_h;
  p.kc_bs = a->kc_block_stride; p.kc_hs = a->kc_head_stride; p.vc_bs = a->vc_block_stride; p.vc_hs = a->vc_head_stride;
  p.b_loc_stride = a->block_table_stride;
  p.num_heads = a->num_heads;
  p.num_kv_heads = a->num_kv_heads;
  p.row_limit = (a->max_input_len + 127) / 128 * 128;
  static const int pair_group = getenv("B200_CTX_PAIR_GROUP") ? atoi(getenv("B200_CTX_PAIR_GROUP")) : 32;  // tuning knob (measured: profiles/r05_prefix_prefill_lpt.jsonl)
  p.pair_group = pair_group;
#ifdef B200_CTX_TRACE
  p.trace = getenv("B200_CTX_TRACE_PTR") ? reinterpret_cast<unsigned long long*>(strtoull(getenv("B200_CTX_TRACE_PTR"), nullptr, 16)) : nullptr;
#endif

  DeviceGuard guard;
  B200_CUDA_TRY(guard.set(a->device));
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  // development knob: B200_CTX_ATTN=mma forces the warp-level MMA kernel for head size 128 too
  static const bool tc = [] {
    const char* e = getenv("B200_CTX_ATTN");
    return !(e && strcmp(e, "mma") == 0);
  }();
  if (a->dtype == B200_F16)
    return launch_ctx_attn_dtype<__half>(p, a->head_size, a->block_size, a->batch, a->max_input_len, tc, st);
  return launch_ctx_attn_dtype<__nv_bfloat16>(p, a->head_size, a->block_size, a->batch, a->max_input_len, tc, st);
}
