/*
 * b200_kv.h -- C ABI of libb200kv.so: the B200 (sm_100a) paged-KV-cache hot path.
 *
 * Every entry point takes plain device/host pointers, sizes, strides and enums; there are no
 * torch (or any C++) types in any signature.  Each function replaces one op of the reference's
 * `vllm._C` extension (or of its MSCCL++ KV-transfer layer); the reference interface it replaces
 * is cited as file:line relative to the reference tree.  The reference-facing Python binding
 * (`vllm_splitwise_b200._C`) is a thin ctypes layer over exactly these symbols; see
 * INTEGRATION.md for the stub a maintainer of the reference would add.
 *
 * Conventions
 *   - all strides are in ELEMENTS of the tensor they describe (as torch's .stride()).
 *   - `stream` is a cudaStream_t passed as an opaque pointer (NULL = legacy default stream).
 *   - every function returns B200_OK (0) or a negative b200_status_t; nothing throws.
 *     No launch is followed by a device sync -- asynchronous faults surface at the caller's next
 *     sync, exactly as in the reference (csrc/attention/attention_kernels.cu has no post-launch
 *     check either).
 *   - functions never allocate device memory and never synchronise, so the attention and
 *     reshape_and_cache entry points are CUDA-graph capturable (reference: decode is captured,
 *     vllm/worker/model_runner.py:826-834).
 */
#ifndef B200_KV_H_
#define B200_KV_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_KV_ABI_VERSION 2

#if defined(__GNUC__)
#define B200_API __attribute__((visibility("default")))
#else
#define B200_API
#endif

typedef void* b200_stream_t; /* cudaStream_t */

typedef enum {
  B200_OK = 0,
  B200_ERR_HEAD_SIZE = -1,  /* reference: TORCH_CHECK "Unsupported head size"  attention_kernels.cu:674 */
  B200_ERR_BLOCK_SIZE = -2, /* reference: TORCH_CHECK "Unsupported block size" attention_kernels.cu:706 */
  B200_ERR_DTYPE = -3,      /* reference: TORCH_CHECK "Unsupported data type"  attention_kernels.cu:732,745 */
  B200_ERR_ARG = -4,        /* null pointer / negative size / inconsistent head counts */
  B200_ERR_ALIGN = -5,      /* cache chunk not 16-byte aligned (bulk-copy requirement) */
  B200_ERR_CUDA = -6,       /* a CUDA runtime call failed; see b200_kv_last_cuda_error() */
  B200_ERR_DEVICE = -7,     /* swap_blocks: src and dst on different GPUs (cache_kernels.cu:28-32) */
  B200_ERR_TIMEOUT = -8,    /* kv wait: bounded spin expired (kv_comm_kernels.cu:7) */
  B200_ERR_SCRATCH_LIMIT = -9 /* b200_paged_attention_v2_static_scratch: the call needs more than B200_V2_STATIC_SCRATCH_BYTES */
} b200_status_t;

typedef enum {
  B200_F32 = 0,
  B200_F16 = 1,
  B200_BF16 = 2,
  B200_U8_E5M2 = 3 /* uint8 storage holding fp8 e5m2, unscaled (kv_cache_dtype == "fp8_e5m2") */
} b200_dtype_t;

B200_API const char* b200_kv_strerror(int status);
B200_API const char* b200_kv_last_cuda_error(void);
B200_API int b200_kv_abi_version(void);
/* Number of kernels this library has launched in this process (all threads) since load; bench.py
 * reports the difference across its timed region as "gpu_launches". */
B200_API uint64_t b200_kv_launch_count(void);

/* ------------------------------------------------------------------------------------------
 * Paged attention decode.
 * Replaces: paged_attention_v1 (csrc/ops.h:5-17, csrc/attention/attention_kernels.cu:711-747)
 *           paged_attention_v2 (csrc/ops.h:19-34, csrc/attention/attention_kernels.cu:901-940)
 *
 *   out[s,h,:] = softmax_t( scale * q[s,h,:].K[t,:] + alibi[h]*(t - ctx_s + 1) ) . V[t,:]
 *   for t < context_lens[s], kv head = h / (num_heads / num_kv_heads).
 *
 * Layouts (reference vllm/worker/cache_engine.py:60-75):
 *   query        [num_seqs, num_heads, head_size]            T, row stride q_stride
 *   key_cache    [num_blocks, num_kv_heads, head_size/x, block_size, x]   x = 16/sizeof(cache elt)
 *   value_cache  [num_blocks, num_kv_heads, head_size, block_size]
 *   block_tables int32 [num_seqs, max_num_blocks_per_seq], context_lens int32 [num_seqs]
 *   out          [num_seqs, num_heads, head_size] T, dense
 * v2 only (caller-allocated scratch, reference paged_attn.py:83-93), P = max_num_partitions:
 *   tmp_out [num_seqs, num_heads, P, head_size] T; exp_sums, max_logits f32 [num_seqs, num_heads, P]
 *   partition p covers tokens [512p, 512p+512) exactly as in the reference.
 * Optional workspace (v1 only): when the (sequence, kv-head) items of a call cannot fill the persistent grid
 *   the library splits contexts into partitions and combines the partials inside the same kernel; it needs
 *   b200_paged_attention_workspace_bytes(args) bytes of caller-owned scratch for exactly these arguments
 *   (0 when the call is not split; a few MB at most; less is fine: the split is limited to what fits).
 *   The first 256 KiB hold reduce tickets: they must be ZERO before the first call that uses the buffer
 *   and every call leaves them zero again, so one cudaMemset at allocation is enough.
 *   Two calls that may EXECUTE concurrently (different streams; a graph replay against an eager call)
 *   must not share a workspace: keep one per stream and one per captured graph, as the torch front does.
 *   workspace == NULL disables splitting (never an error).
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  void* out;
  const void* query;
  const void* key_cache;
  const void* value_cache;
  const int32_t* block_tables;
  const int32_t* context_lens;
  const float* alibi_slopes; /* [num_heads] f32 or NULL */
  float scale;
  int32_t num_seqs;
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t head_size;  /* 64, 80, 96, 112, 128, 256 */
  int32_t block_size; /* 8, 16, 32 */
  int32_t max_context_len;
  int32_t max_num_blocks_per_seq;
  int64_t q_stride;
  int64_t kv_block_stride;
  int64_t kv_head_stride;
  int32_t query_dtype; /* b200_dtype_t: F32/F16/BF16 */
  int32_t cache_dtype; /* == query_dtype ("auto") or B200_U8_E5M2 */
  /* v2 scratch */
  float* exp_sums;
  float* max_logits;
  void* tmp_out;
  int32_t max_num_partitions;
  /* v1 optional scratch */
  void* workspace;
  int64_t workspace_bytes;
  int32_t device; /* CUDA device ordinal the pointers live on, or -1 = current device */
  int32_t reserved;
} b200_paged_attn_args_t;

B200_API int b200_paged_attention_v1(const b200_paged_attn_args_t* args, b200_stream_t stream);
B200_API int b200_paged_attention_v2(const b200_paged_attn_args_t* args, b200_stream_t stream);
B200_API int64_t b200_paged_attention_workspace_bytes(const b200_paged_attn_args_t* args);

/* paged_attention_v2 for callers that cannot allocate (code under CUDA graph capture, serving loops and harnesses
 * that forbid cudaMalloc).  No counterpart in the reference.
 * Same kernels and the same results, bit for bit, as b200_paged_attention_v2, but the v2 scratch (tmp_out, exp_sums,
 * max_logits) lives in a file-scope __device__ array of B200_V2_STATIC_SCRATCH_BYTES inside the library, resolved
 * once per device with cudaGetSymbolAddress; args->tmp_out / exp_sums / max_logits / max_num_partitions are ignored
 * (partitions = ceil(max_context_len / 512)).  The call allocates nothing (no cudaMalloc / cudaFree, no allocating
 * library routine) and never synchronises.
 *   limit   num_seqs * num_heads * partitions * (head_size * sizeof(query element) + 8) bytes, each of the three
 *           pieces rounded up to 256 -- b200_paged_attention_v2_static_scratch_bytes(args) -- must not exceed
 *           B200_V2_STATIC_SCRATCH_BYTES, else B200_ERR_SCRATCH_LIMIT and nothing is launched
 *           (8 MiB = e.g. 64 sequences x 32 heads x 4096 tokens of context at head size 128, f16/bf16).
 *   sharing every call on a device uses the SAME array (it keeps the last call's contents; nothing in it has to
 *           start at zero).  Calls that can overlap on the GPU -- different streams, or graphs replayed
 *           concurrently -- race on it: serialise them, or use b200_paged_attention_v2 with caller scratch.
 * Every GPU that loads the library holds the array until the process exits. */
#define B200_V2_STATIC_SCRATCH_BYTES (8 << 20)
B200_API int b200_paged_attention_v2_static_scratch(const b200_paged_attn_args_t* args, b200_stream_t stream);
B200_API int64_t b200_paged_attention_v2_static_scratch_bytes(const b200_paged_attn_args_t* args);

/* ------------------------------------------------------------------------------------------
 * Cache writes and block moves.
 * ------------------------------------------------------------------------------------------ */

/* Replaces reshape_and_cache (csrc/cache.h:18-24, csrc/cache_kernels.cu:154-268).
 * key/value [num_tokens, num_heads, head_size] with row strides key_stride/value_stride;
 * slot_mapping int64 [num_tokens], negative = padding (skipped). src_dtype F32/F16/BF16;
 * cache_dtype == src_dtype or U8_E5M2 (saturating RNE convert, quant_utils.cuh:164-192). */
B200_API int b200_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                           const int64_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                           int32_t head_size, int32_t block_size, int32_t x, int64_t key_stride,
                           int64_t value_stride, int64_t num_blocks, int32_t src_dtype,
                           int32_t cache_dtype, b200_stream_t stream);

/* Replaces copy_blocks (csrc/cache.h:13-16, csrc/cache_kernels.cu:65-150).
 * key_cache_ptrs/value_cache_ptrs: HOST arrays of num_layers device base addresses;
 * block_pairs: HOST array [num_pairs][2] = (src_block, dst_block); block_bytes = bytes of one
 * block of one layer's K (== V).  Tables travel as kernel parameters: no H2D copy, no sync
 * (the reference does a synchronous .to(device), cache_kernels.cu:129-134). */
B200_API int b200_copy_blocks(const int64_t* key_cache_ptrs, const int64_t* value_cache_ptrs,
                     int32_t num_layers, const int64_t* block_pairs, int32_t num_pairs,
                     int64_t block_bytes, b200_stream_t stream);

typedef enum { B200_SWAP_D2D = 0, B200_SWAP_D2H = 1, B200_SWAP_H2D = 2 } b200_swap_kind_t;

/* Replaces swap_blocks (csrc/cache.h:8-11, csrc/cache_kernels.cu:21-60).
 * block_pairs HOST [num_pairs][2] (src,dst) in the iteration order of the reference's std::map
 * (ascending src).  D2D: one gather kernel. D2H/H2D: consecutive (src,dst) runs are merged into one
 * cudaMemcpyAsync each.  src_device/dst_device: CUDA ordinals or -1 for host. */
B200_API int b200_swap_blocks(const void* src, void* dst, int64_t block_bytes, const int64_t* block_pairs,
                     int32_t num_pairs, int32_t kind, int32_t src_device, int32_t dst_device,
                     b200_stream_t stream);

/* Replaces convert_fp8_e5m2 (csrc/cache.h:26-29, csrc/cache_kernels.cu:272-320): elementwise
 * {F32,F16,BF16} -> U8_E5M2 or back, numel elements. */
B200_API int b200_convert_fp8_e5m2(const void* src, void* dst, int64_t numel, int32_t src_dtype,
                          int32_t dst_dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Rotary position embedding, alone and fused with the cache write (the op the model runs right
 * before reshape_and_cache: vllm/model_executor/models/llama.py:156-158).
 * ------------------------------------------------------------------------------------------ */

/* Replaces rotary_embedding (csrc/ops.h:48-54, csrc/pos_encoding_kernels.cu:77-95,126-165) and, with
 * cos_sin_cache_offsets != NULL, batched_rotary_embedding (csrc/ops.h:56-63, :97-123,167-209).
 * In place on query [num_tokens, num_heads, head_size] (row stride query_stride) and key
 * [num_tokens, num_kv_heads, head_size] (row stride key_stride); positions int64 [num_tokens];
 * cos_sin_cache [max_position, rot_dim] of the query dtype: rot_dim/2 cosines then rot_dim/2 sines.
 * is_neox: pairs (i, i + rot_dim/2); otherwise GPT-J pairs (2i, 2i+1).  dtype F32/F16/BF16.
 * Every product and sum is rounded to the element type, as the reference's scalar_t arithmetic. */
B200_API int b200_rotary_embedding(const int64_t* positions, void* query, void* key, const void* cos_sin_cache,
                          const int64_t* cos_sin_cache_offsets, int64_t num_tokens, int32_t num_heads,
                          int32_t num_kv_heads, int32_t head_size, int32_t rot_dim, int64_t query_stride,
                          int64_t key_stride, int32_t is_neox, int32_t dtype, b200_stream_t stream);

/* rotary_embedding(positions, query, key) followed by reshape_and_cache(key, value, ...) with the same
 * results bit for bit: query and key are rotated in place, the rotated key and the value go to the
 * paged cache (negative slots skip the cache write only).  Decode-sized calls are ONE launch (the K
 * row is read once); prefill-sized calls run the two kernels back to back.  Arguments as the two ops. */
B200_API int b200_rotary_embedding_reshape_and_cache(
    const int64_t* positions, void* query, void* key, const void* value, const void* cos_sin_cache,
    int64_t num_tokens, int32_t num_heads, int32_t num_kv_heads, int32_t head_size, int32_t rot_dim,
    int64_t query_stride, int64_t key_stride, int64_t value_stride, int32_t is_neox, void* key_cache,
    void* value_cache, const int64_t* slot_mapping, int32_t block_size, int32_t x, int64_t num_blocks,
    int32_t src_dtype, int32_t cache_dtype, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Prefix ("context") prefill attention over the paged cache.
 * Replaces context_attention_fwd, the Triton kernels _fwd_kernel / _fwd_kernel_alibi
 * (vllm/model_executor/layers/attention/ops/prefix_prefill.py:621-745, :10-199, :392-618); caller
 * PagedAttentionImpl.forward_prefix (.../ops/paged_attn.py:113-138).
 *
 *   for sequence b with ctx = ctx_lens[b] cached tokens and L = seq_lens[b] - ctx new tokens at rows
 *   start_loc[b] .. start_loc[b]+L-1 of query/key/value/out, and m < L:
 *     out[m,h,:] = softmax_n( scale*q[m,h].k[n] + alibi[h]*(pos_n - (ctx+m)) ) . v[n]
 *   over the cached tokens n < ctx (read through block_tables from key_cache / value_cache) and the new
 *   tokens n <= m (read from key / value); kv head = h / (num_heads / num_kv_heads).
 *
 * query/out [num_tokens, num_heads, head_size], key/value [num_tokens, num_kv_heads, head_size] with
 * token / head strides in elements (last dim dense); caches in the paged layouts of paged_attention
 * (x = 8: 2-byte elements only, as the reference hard-codes, prefix_prefill.py:719); block_tables int32
 * [batch, block_table_stride]; start_loc, seq_lens, ctx_lens int32 [batch]; dtype F16 / BF16;
 * head_size 16/32/64/128 (prefix_prefill.py:639), block_size 8/16/32.  Rows m >= max_input_len are not
 * computed (the reference's grid, :645).  Never allocates, never synchronises.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  void* out;
  const void* query;
  const void* key;
  const void* value;
  const void* key_cache;
  const void* value_cache;
  const int32_t* block_tables;
  const int32_t* start_loc;
  const int32_t* seq_lens;
  const int32_t* ctx_lens;
  const float* alibi_slopes; /* [num_heads] f32 or NULL */
  float scale;
  int32_t batch;
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t head_size;
  int32_t block_size;
  int32_t max_input_len;
  int32_t block_table_stride;
  int64_t q_stride_t, q_stride_h;
  int64_t k_stride_t, k_stride_h;
  int64_t v_stride_t, v_stride_h;
  int64_t o_stride_t, o_stride_h;
  int64_t kc_block_stride, kc_head_stride;
  int64_t vc_block_stride, vc_head_stride;
  int32_t dtype;  /* b200_dtype_t: F16 / BF16 */
  int32_t device; /* CUDA device ordinal, or -1 = current device */
} b200_context_attn_args_t;

B200_API int b200_context_attention_fwd(const b200_context_attn_args_t* args, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * Device-side decode metadata.
 * Replaces the per-step Python list building + H2D copies of ModelRunner._prepare_decode
 * (vllm/worker/model_runner.py:274-399).  Persistent device state, owned by the caller:
 *   state_block_tables int32 [max_seqs, max_blocks_per_seq]   physical block of every logical block
 *   state_seq_lens     int32 [max_seqs]                       tokens of the sequence so far
 * b200_block_tables_append: appends = int32 triples (state_row, block_index, physical_block) on the device --
 *   the blocks the block manager allocated since the last step (block_manager.py:268-309).
 * b200_decode_metadata_step: for batch entry b with row r = batch_rows[b] (r < 0 = CUDA-graph padding entry):
 *   seq_len = ++state_seq_lens[r]; positions[b] = seq_len-1; context_lens[b] = min(seq_len, sliding_window) (or
 *   seq_len when sliding_window <= 0); slot_mapping[b] = table[r][pos / bs]*bs + pos % bs;
 *   block_tables[b, :] = the sequence's blocks (the last sliding_window/bs of them with a window), zero padded to
 *   out_blocks.  Padding entries get position 0, slot -1 (_PAD_SLOT_ID), context 1, a zero row (:336-341).
 *   A row must appear at most once per batch.  Never allocates, never synchronises (CUDA-graph capturable).
 * ------------------------------------------------------------------------------------------ */
B200_API int b200_block_tables_append(int32_t* state_block_tables, const int32_t* appends, int32_t num_appends,
                                      int32_t max_blocks_per_seq, b200_stream_t stream);
B200_API int b200_decode_metadata_step(const int32_t* state_block_tables, int32_t* state_seq_lens,
                                       const int32_t* batch_rows, int32_t batch, int32_t max_blocks_per_seq,
                                       int32_t block_size, int32_t sliding_window, int64_t* positions,
                                       int64_t* slot_mapping, int32_t* context_lens, int32_t* block_tables,
                                       int32_t out_blocks, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * One decode step of one attention layer in a single call: [decode metadata ->] RoPE + cache write -> paged
 * attention, the sequence the reference runs per layer inside its captured decode graphs
 * (vllm/worker/model_runner.py:274-399 then :826-873; vllm/model_executor/models/llama.py:156-158;
 * .../attention/ops/paged_attn.py:21-36, 38-111).  Same results, bit for bit, as calling
 * b200_decode_metadata_step, b200_rotary_embedding_reshape_and_cache (or b200_reshape_and_cache when
 * cos_sin_cache == NULL) and b200_paged_attention_v1 (v2 when attn.exp_sums != NULL) one after the other;
 * the launches are chained with programmatic dependent launch so their prologues overlap.
 *   attn                   the attention call of the step.  attn.query is rotated in place first;
 *                          attn.block_tables / attn.context_lens are the tensors the metadata kernel WRITES
 *                          (int32 [num_seqs, attn.max_num_blocks_per_seq] / [num_seqs]) when state_block_tables
 *                          != NULL, plain inputs otherwise.
 *   state_*, batch_rows, sliding_window, positions, slot_mapping   as b200_decode_metadata_step;
 *                          state_block_tables == NULL skips the metadata kernel (layers 2..L of a model step
 *                          reuse the tensors the first layer's call produced).
 *   key, value             the step's new K / V rows [num_seqs, num_kv_heads, head_size], row strides in elements.
 * Never allocates, never synchronises: CUDA-graph capturable.
 * ------------------------------------------------------------------------------------------ */
typedef struct {
  b200_paged_attn_args_t attn;
  const int32_t* state_block_tables; /* NULL: no metadata kernel */
  int32_t* state_seq_lens;
  const int32_t* batch_rows;
  int32_t state_max_blocks_per_seq;
  int32_t sliding_window;
  int64_t* positions;    /* int64 [num_seqs]: written by the metadata kernel, read by RoPE */
  int64_t* slot_mapping; /* int64 [num_seqs]: written by the metadata kernel, read by the cache write */
  void* key;
  const void* value;
  const void* cos_sin_cache; /* [max_position, rot_dim] of the query dtype, or NULL = no rotary embedding */
  int32_t rot_dim;
  int32_t is_neox;
  int64_t key_stride;
  int64_t value_stride;
  int64_t num_blocks; /* blocks in key_cache / value_cache (slots beyond are never written) */
} b200_decode_step_args_t;

B200_API int b200_decode_step(const b200_decode_step_args_t* args, b200_stream_t stream);

/* ------------------------------------------------------------------------------------------
 * One-shot all-reduce (sum) over NVLink peer memory: the tensor-parallel reduction after the
 * attention output projection.  Replaces the reference's custom all-reduce, 1-stage kernel
 * (csrc/custom_all_reduce.cuh:215-262, :420-470; vllm/model_executor/parallel_utils/custom_all_reduce.py).
 * Every rank owns an input buffer and two uint64 flag arrays start[world], done[world] (zero at creation), mapped
 * into all peers with b200_kv_ipc_* / peer access.  Per call (the same `epoch` = 1, 2, 3, ... on all ranks):
 *   peer_inputs[r]      rank r's input buffer as seen from this rank (own buffer for r == rank)
 *   peer_start_slots[r] address of start[rank] inside rank r's start array;  peer_done_slots[r] likewise for done
 *   my_start / my_done  this rank's own arrays;  counter: local uint32, zero between calls
 *   out[i] = sum over r = 0..world-1 (fp32, in rank order -> identical bits on every rank) of input_r[i]
 * numel * sizeof(elt) must be a multiple of 16.  Spins are bounded (max_spins, default 5e7): on expiry *timeout_flag
 * is set to 1 and the kernel completes.  Never allocates, never synchronises the host.
 * ------------------------------------------------------------------------------------------ */
B200_API int b200_allreduce_oneshot(const int64_t* peer_inputs, const int64_t* peer_start_slots,
                                    const int64_t* peer_done_slots, const uint64_t* my_start, const uint64_t* my_done,
                                    uint32_t* counter, int32_t* timeout_flag, void* out, int64_t numel, int32_t dtype,
                                    int32_t rank, int32_t world, uint64_t epoch, int64_t max_spins, int32_t max_ctas,
                                    b200_stream_t stream);

/* Two-stage variant for large messages (reference cross_device_reduce_2stage, custom_all_reduce.cuh:264-330):
 * reduce-scatter into a peer-visible scratch buffer (peer_tmps[r] = rank r's scratch as seen from this rank, at least
 * as large as the message), a per-CTA flag barrier (mid array uint64 [160][world] per rank, zero at creation;
 * peer_mid_slots[r] = address of mid[0][rank] inside rank r's array), then an all-gather of the slices.  Same results,
 * bit for bit, as the one-shot kernel.  All ranks must call it with the same numel. */
B200_API int b200_allreduce_twostage(const int64_t* peer_inputs, const int64_t* peer_tmps,
                                     const int64_t* peer_start_slots, const int64_t* peer_done_slots,
                                     const int64_t* peer_mid_slots, const uint64_t* my_start, const uint64_t* my_done,
                                     const uint64_t* my_mid, uint32_t* counter, int32_t* timeout_flag, void* out,
                                     int64_t numel, int32_t dtype, int32_t rank, int32_t world, uint64_t epoch,
                                     int64_t max_spins, b200_stream_t stream);

/* Replaces cuda_utils (csrc/cuda_utils_kernels.cu:5-35). */
B200_API int b200_get_device_attribute(int32_t attribute, int32_t device_id);
B200_API int b200_get_max_shared_memory_per_block_device_attribute(int32_t device_id);

/* ------------------------------------------------------------------------------------------
 * Splitwise prefill -> decode KV-block transfer over NVLink peer memory.
 * Replaces the MSCCL++ proxy-channel kernels (csrc/kv_comm_kernels.cu:3-29) and the calls the
 * reference makes on them (vllm/worker/comm_utils.py:117-147): put / signal+flush / wait.
 * There is no host proxy: the prompt GPU's kernel stores straight into the token GPU's cache.
 * ------------------------------------------------------------------------------------------ */
#define B200_IPC_HANDLE_BYTES 64

/* Export dev_ptr's allocation for another process; offset = dev_ptr - allocation base. */
B200_API int b200_kv_ipc_export(const void* dev_ptr, unsigned char handle[B200_IPC_HANDLE_BYTES],
                       int64_t* offset, int64_t* alloc_bytes);
B200_API int b200_kv_ipc_open(const unsigned char handle[B200_IPC_HANDLE_BYTES], void** base);
B200_API int b200_kv_ipc_close(void* base);
/* Same-process multi-GPU: enable dev -> peer access (idempotent). */
B200_API int b200_kv_enable_peer_access(int32_t dev, int32_t peer_dev);

/* Copy block runs of `num_bufs` buffers to the same byte offsets in their peer-mapped twins.
 * src_ptrs/dst_ptrs: DEVICE int64 arrays of device addresses (dst = peer mapping), one per buffer
 * (layer x {K,V}), built once at setup; runs: HOST [num_runs][2] = (block_start, num_blocks) as produced by the
 * reference's coalesce_blocks (vllm/utils.py:403-426).  One launch, max_ctas CTAs (0 = default).
 * Replaces nw_cache_out_kernel's put (kv_comm_kernels.cu:11-20); offsets are 64-bit. */
/* OR this into a positive max_ctas: the caller guarantees that the source blocks of this put were complete when
 * it was enqueued by something other than stream order behind the previous put (e.g. an event the stream waits
 * for).  The put may then overlap the tail of the put before it (programmatic dependent launch); completion
 * stays in stream order.  Without the flag a put starts reading only after everything before it has finished. */
#define B200_KV_PUT_CHAINED (1 << 30)
B200_API int b200_kv_put_blocks(const int64_t* src_ptrs, const int64_t* dst_ptrs, int32_t num_bufs,
                       const int32_t* runs, int32_t num_runs, int64_t block_bytes,
                       int32_t max_ctas, b200_stream_t stream);

/* Process-wide put configuration.  variant: 0 = SIMT (every warp moves 4 KiB chunks with ld.global / st.global),
 * 1 = bulk (one thread per CTA drives cp.async.bulk global -> shared ring -> peer global), < 0 = keep;
 * default_ctas: CTAs a put launch uses when max_ctas <= 0 (the transfer overlaps prefill compute, so it is a
 * budget of SMs, not "all of them"), <= 0 = keep.  Initial values: B200_KV_PUT_VARIANT / B200_KV_PUT_CTAS or
 * the measured defaults (profiles/r05_kv_put_sweep*.jsonl).  No reference counterpart: MSCCL++'s proxy uses the
 * copy engine (comm_utils.py:173-197). */
B200_API int b200_kv_put_configure(int32_t variant, int32_t default_ctas);
/* The configuration in force (either pointer may be NULL). */
B200_API int b200_kv_put_config(int32_t* variant, int32_t* default_ctas);

/* Run lists longer than one launch's parameter table (several hundred runs: many fragmented prompts put in
 * one go).  b200_kv_put_table_build (host only, no CUDA call) turns runs [num_runs][2] into a table of
 * 3*num_runs+1 int32 -- starts, lengths, exclusive tile prefix -- that the caller copies to the device ONCE
 * and reuses for every layer (the backends put the same block list 2 x num_layers times,
 * flash_attn.py:81-86); b200_kv_put_blocks_table then moves all of it in one launch. */
B200_API int b200_kv_put_table_build(const int32_t* runs, int32_t num_runs, int64_t block_bytes, int32_t* table,
                            int32_t* tile_vecs_out, int64_t* tiles_per_buf_out);
B200_API int b200_kv_put_blocks_table(const int64_t* src_ptrs, const int64_t* dst_ptrs, int32_t num_bufs,
                             const int32_t* dev_table, int32_t num_runs, int32_t tile_vecs, int64_t tiles_per_buf,
                             int64_t block_bytes, int32_t max_ctas, b200_stream_t stream);

/* After all prior work on `stream`: system-scope release store of `value` to *remote_flag (a
 * peer-mapped uint64).  Replaces nw_cache_out_signal_kernel (kv_comm_kernels.cu:22-29). */
B200_API int b200_kv_signal(uint64_t* remote_flag, uint64_t value, b200_stream_t stream);

/* Enqueue a one-warp kernel that acquire-spins until *local_flag >= value or max_spins polls
 * passed; on expiry it sets *timeout_flag (device int32, may be NULL) to 1 and returns.
 * Replaces nw_cache_in_kernel (kv_comm_kernels.cu:3-9). */
B200_API int b200_kv_wait(const uint64_t* local_flag, uint64_t value, int64_t max_spins,
                 int32_t* timeout_flag, b200_stream_t stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_KV_H_ */
