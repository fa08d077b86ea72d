// Host side of b200_paged_attention_v1/v2: argument checks, grid/partition planning, dispatch.
// Replaces the reference launchers (csrc/attention/attention_kernels.cu:606-747, 784-940).
#include <algorithm>
#include <atomic>

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
  p.num_seqs = a->num_seqs;
  p.group = a->num_heads / a->num_kv_heads;
  p.max_num_blocks_per_seq = a->max_num_blocks_per_seq;
  p.q_stride = a->q_stride;
  p.kv_block_stride = a->kv_block_stride;
  p.kv_head_stride = a->kv_head_stride;
#ifdef B200_ATTN_TRACE
  p.trace = getenv("B200_ATTN_TRACE_PTR") ? reinterpret_cast<unsigned long long*>(strtoull(getenv("B200_ATTN_TRACE_PTR"), nullptr, 16)) : nullptr;
#endif
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

int dispatch_mma(const b200_paged_attn_args_t* a, const AttnParams& p, int gt, int budget, int sms, cudaStream_t stream) {
  const bool fp8 = a->cache_dtype == B200_U8_E5M2;
  if (a->query_dtype == B200_F16)
    return fp8 ? launch_mma<__half, uint8_t>(p, a->head_size, a->block_size, gt, budget, sms, stream)
               : launch_mma<__half, __half>(p, a->head_size, a->block_size, gt, budget, sms, stream);
  if (a->query_dtype == B200_BF16)
    return fp8 ? launch_mma<__nv_bfloat16, uint8_t>(p, a->head_size, a->block_size, gt, budget, sms, stream)
               : launch_mma<__nv_bfloat16, __nv_bfloat16>(p, a->head_size, a->block_size, gt, budget, sms, stream);
  return B200_ERR_DTYPE;
}

int dispatch_simt(const b200_paged_attn_args_t* a, const AttnParams& p, int gt, int budget, int sms, cudaStream_t stream) {
  const bool fp8 = a->cache_dtype == B200_U8_E5M2;
  const int hs = a->head_size, bs = a->block_size;
  switch (a->query_dtype) {
    case B200_F32:
      return fp8 ? launch_simt<float, uint8_t>(p, hs, bs, gt, budget, sms, stream)
                 : launch_simt<float, float>(p, hs, bs, gt, budget, sms, stream);
    case B200_F16:
      return fp8 ? launch_simt<__half, uint8_t>(p, hs, bs, gt, budget, sms, stream)
                 : launch_simt<__half, __half>(p, hs, bs, gt, budget, sms, stream);
    case B200_BF16:
      return fp8 ? launch_simt<__nv_bfloat16, uint8_t>(p, hs, bs, gt, budget, sms, stream)
                 : launch_simt<__nv_bfloat16, __nv_bfloat16>(p, hs, bs, gt, budget, sms, stream);
    default:
      return B200_ERR_DTYPE;
  }
}

// Merge of the v2 partials in a second kernel, launched with programmatic stream serialization so that its
// launch overlaps the attention kernel's tail (the kernel orders itself with griddepcontrol.wait).
template <typename T>
int launch_reduce_t(const b200_paged_attn_args_t* a, int max_parts, int partition_tokens, cudaStream_t stream) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(a->num_seqs, a->num_heads);
  cfg.blockDim = dim3(128);
  cfg.dynamicSmemBytes = (size_t)max_parts * sizeof(float);
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  B200_CUDA_TRY(cudaLaunchKernelEx(&cfg, paged_attn_reduce_kernel<T>, static_cast<T*>(a->out),
                                   static_cast<const float*>(a->exp_sums), static_cast<const float*>(a->max_logits),
                                   static_cast<const T*>(a->tmp_out), a->context_lens, (int)a->head_size, max_parts,
                                   partition_tokens));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int launch_reduce(const b200_paged_attn_args_t* a, int max_parts, int partition_tokens, cudaStream_t stream) {
  switch (a->query_dtype) {
    case B200_F32: return launch_reduce_t<float>(a, max_parts, partition_tokens, stream);
    case B200_F16: return launch_reduce_t<__half>(a, max_parts, partition_tokens, stream);
    case B200_BF16: return launch_reduce_t<__nv_bfloat16>(a, max_parts, partition_tokens, stream);
    default: return B200_ERR_DTYPE;
  }
}

// Shared-memory budget per CTA: aim for two resident CTAs per SM (2 x ~112 KiB of the 227 KiB), fall back
// to one when a ring of 4 stages does not fit in that (fp32, D=256, bs=32).
int pick_smem_budget(const b200_paged_attn_args_t* a, const TilePlan& t, int* ctas_per_sm) {
  const int stage = 2 * a->head_size * a->block_size * dtype_size(a->cache_dtype);
  const int fixed = t.mma ? 0 : t.gt * a->head_size * 4 + kConsumerWarps * t.gt * a->block_size * 4;
  const int tail = attn_smem_tail(fixed, t.gt, a->head_size, dtype_size(a->query_dtype));
  // two resident CTAs per SM when each still gets a ring of at least 4 stages, else one.  (Three CTAs with
  // 12-stage rings fit for 1-byte caches but measured 2 % slower on C4: profiles/README.md.)
  int want = 2;
  if (const char* e = getenv("B200_ATTN_CTAS_PER_SM")) want = atoi(e);  // tuning knob
  for (int c = 3; c >= 1; --c) {
    if (c > want) continue;
    const int budget = (227 * 1024) / c - 1024;  // 1 KiB per CTA is reserved by the system
    const int n = (budget - tail) / stage;
    if (c == 1 || n >= (c == 3 ? 12 : 4)) {
      *ctas_per_sm = c;
      return budget;
    }
  }
  *ctas_per_sm = 1;
  return 226 * 1024;
}

int dispatch(const b200_paged_attn_args_t* a, const AttnParams& p, const TilePlan& t, int budget, int sms,
             cudaStream_t stream) {
  return t.mma ? dispatch_mma(a, p, t.gt, budget, sms, stream) : dispatch_simt(a, p, t.gt, budget, sms, stream);
}

// Work counters of the persistent kernel: one per launch in flight.  A launch leaves its counter at zero, so
// a slot is reusable as soon as that launch has finished.
//   eager launches    round-robin over kEagerSlots: two launches can only collide if more than a thousand
//                     attention kernels are EXECUTING at once on one device.
//   captured launches (the reference captures decode, model_runner.py:826-834) bake the slot address into the
//                     graph, and a replay may overlap any later eager call or other graph -- so every captured
//                     launch gets a slot of its own, never handed out again.  When those run out the launch
//                     falls back to a static round-robin schedule (p.sched == nullptr), which needs no counter.
constexpr unsigned kEagerSlots = 1024;
constexpr unsigned kCaptureSlots = 31 * 1024;
__device__ unsigned int g_sched[kEagerSlots + kCaptureSlots];
std::atomic<unsigned> g_sched_seq{0};
std::atomic<unsigned> g_capture_seq[64];

int attach_sched(AttnParams& p, cudaStream_t stream) {
  static unsigned int* base[64] = {nullptr};
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  dev &= 63;
  if (!base[dev]) {
    void* sym = nullptr;
    B200_CUDA_TRY(cudaGetSymbolAddress(&sym, g_sched));
    base[dev] = static_cast<unsigned int*>(sym);
  }
  cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
  if (cudaStreamIsCapturing(stream, &cs) != cudaSuccess) {
    (void)cudaGetLastError();
    cs = cudaStreamCaptureStatusNone;
  }
  static const bool force_static = getenv("B200_ATTN_STATIC_SCHED") != nullptr;  // test knob
  if (force_static) {
    p.sched = nullptr;
  } else if (cs != cudaStreamCaptureStatusNone) {
    const unsigned k = g_capture_seq[dev].fetch_add(1, std::memory_order_relaxed);
    p.sched = k < kCaptureSlots ? base[dev] + kEagerSlots + k : nullptr;
  } else {
    p.sched = base[dev] + (g_sched_seq.fetch_add(1, std::memory_order_relaxed) % kEagerSlots);
  }
  return B200_OK;
}

constexpr int kTicketInts = 1 << 16;                       // head of the v1 workspace: in-kernel reduce tickets
constexpr int64_t kTicketBytes = (int64_t)kTicketInts * 4;
constexpr int kMaxSplit = 64;

// Internal split of a v1 call (needs caller workspace).  The persistent grid runs `slots` CTAs; CTA c takes
// items c, c + slots, ...  With equal contexts the call lasts ceil(items / slots) rounds of
// (partition_blocks + per-item overhead) block times, so choose the partition count that minimises that:
// it fills the last round without making items so small that the per-item epilogue shows.
int plan_v1_split(const b200_paged_attn_args_t* a, int tiles_per_seq, int slots, int64_t ws_avail, int* partition_blocks,
                  int64_t* need_out = nullptr) {
  *partition_blocks = 0;
  if (need_out) *need_out = 0;
  const long long base_items = (long long)a->num_seqs * tiles_per_seq;
  const int max_blocks = (a->max_context_len + a->block_size - 1) / a->block_size;
  if (ws_avail <= 0 || base_items > kTicketInts || max_blocks < 16) return 1;
  auto need_for = [&](int np_eff) {
    return kTicketBytes + (long long)a->num_seqs * a->num_heads * np_eff *
                              ((long long)a->head_size * dtype_size(a->query_dtype) + 8) + 64;
  };
  if (const char* e = getenv("B200_ATTN_SPLIT")) {  // tuning knob: force a partition count
    const int np = std::max(1, std::min(atoi(e), kMaxSplit));
    if (np == 1) return 1;
    const int pb = (max_blocks + np - 1) / np;
    const int np_eff = (max_blocks + pb - 1) / pb;
    if (need_for(np_eff) > ws_avail) return 1;
    *partition_blocks = pb;
    if (need_out) *need_out = need_for(np_eff);
    return np_eff;
  }
  // Measured (profiles/r02_split_sweep.txt): with dynamically scheduled items an unsplit call runs at the HBM
  // ceiling as soon as the items cover ~3/4 of the resident CTAs, and every extra partition costs a few
  // percent (partials + per-item epilogue).  So split only to fill the machine, never to balance it.
  if (base_items * 4 >= (long long)slots * 3) return 1;
  const int min_pb = 8;
  // The largest partition count whose items still run in ONE wave (items <= resident CTAs): one item more than
  // that and a second, nearly empty wave doubles the call (measured, profiles/r05_attn_quad.txt: the reference
  // benchmark shape, 64 base items on 296 CTAs, takes 35.7 us with 4 partitions and 48.3 us with 5).
  int np = (int)std::min<long long>(std::max<long long>(1, slots / base_items), kMaxSplit);
  np = std::min(np, std::max(1, max_blocks / min_pb));
  int best_np = 1, best_pb = 0;
  for (; np > 1; --np) {
    const int pb = (max_blocks + np - 1) / np;
    const int np_eff = (max_blocks + pb - 1) / pb;
    if (need_for(np_eff) <= ws_avail) {
      best_np = np_eff;
      best_pb = pb;
      if (need_out) *need_out = need_for(np_eff);
      break;
    }
  }
  *partition_blocks = best_pb;
  return best_np;
}

// Scratch of b200_paged_attention_v2_static_scratch: tmp_out | exp_sums | max_logits of ONE v2 call, held in the
// library's own image instead of caller tensors (every device that loads the library keeps these bytes until the
// process exits).  In host code `g_v2_static_scratch` is only a handle for cudaGetSymbolAddress -- never a device
// address -- so the host side resolves it once per device, with that device current.
static __device__ __align__(256) unsigned char g_v2_static_scratch[B200_V2_STATIC_SCRATCH_BYTES];

int static_scratch_base(unsigned char** out) {
  static std::atomic<unsigned char*> base[64];
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  if (dev < 0 || dev >= 64) return B200_ERR_ARG;
  unsigned char* b = base[dev].load(std::memory_order_acquire);
  if (!b) {
    void* sym = nullptr;
    B200_CUDA_TRY(cudaGetSymbolAddress(&sym, g_v2_static_scratch));
    b = static_cast<unsigned char*>(sym);
    base[dev].store(b, std::memory_order_release);  // racing threads resolve the same address
  }
  *out = b;
  return B200_OK;
}

constexpr int64_t align256(int64_t v) { return (v + 255) / 256 * 256; }

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

B200_API int64_t b200_paged_attention_workspace_bytes(const b200_paged_attn_args_t* a) {
  // Exactly what a v1 call with these arguments uses (0 when the planner does not split it): the split only
  // happens when the (sequence, head tile) items cannot fill the resident CTAs, so the partials stay small
  // (< 2 x resident CTAs x 16 rows), a few MB at most.
  if (!a || check_common(a) != B200_OK || a->num_seqs <= 0) return 0;
  const TilePlan tile = plan_tile(a);
  const int group = a->num_heads / a->num_kv_heads;
  const int tiles_per_seq = a->num_kv_heads * ((group + tile.gt - 1) / tile.gt);
  int ctas_per_sm = 2, pb = 0;
  (void)pick_smem_budget(a, tile, &ctas_per_sm);
  int64_t need = 0;
  (void)plan_v1_split(a, tiles_per_seq, ctas_per_sm * num_sms(a->device), INT64_MAX, &pb, &need);
  return need;
}

B200_API int b200_paged_attention_v1(const b200_paged_attn_args_t* a, b200_stream_t stream_) {
  int st = check_common(a);
  if (st != B200_OK || a->num_seqs == 0) return st;
  DeviceGuard guard;
  B200_CUDA_TRY(guard.set(a->device));
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  AttnParams p = base_params(a);
  const TilePlan tile = plan_tile(a);
  p.gtiles = (p.group + tile.gt - 1) / tile.gt;
  p.tiles_per_seq = a->num_kv_heads * p.gtiles;
  const int sms = num_sms(a->device);
  int ctas_per_sm = 2;
  const int budget = pick_smem_budget(a, tile, &ctas_per_sm);
  int pb = 0;
  const int np = plan_v1_split(a, p.tiles_per_seq, ctas_per_sm * sms, a->workspace ? a->workspace_bytes : 0, &pb);
  p.partition_blocks = pb;
  p.max_num_partitions = np;
  p.always_partials = 0;
  if (np > 1) {
    // carve the workspace: tickets | tmp_out | exp_sums | max_logits
    const long long rows = (long long)a->num_seqs * a->num_heads * np;
    unsigned char* ws = static_cast<unsigned char*>(a->workspace);
    p.tickets = reinterpret_cast<int*>(ws);
    p.tmp_out = ws + kTicketBytes;
    long long off = kTicketBytes + rows * a->head_size * dtype_size(a->query_dtype);
    off = (off + 15) / 16 * 16;
    p.exp_sums = reinterpret_cast<float*>(ws + off);
    p.max_logits = p.exp_sums + rows;
  }
  p.total_items = (long long)np * a->num_seqs * p.tiles_per_seq;
  st = attach_sched(p, stream);
  if (st != B200_OK) return st;
  return dispatch(a, p, tile, budget, sms, stream);
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
  p.gtiles = (p.group + tile.gt - 1) / tile.gt;
  p.tiles_per_seq = a->num_kv_heads * p.gtiles;
  p.exp_sums = a->exp_sums;
  p.max_logits = a->max_logits;
  p.tmp_out = a->tmp_out;
  p.partition_blocks = kV2PartitionTokens / a->block_size;
  p.max_num_partitions = a->max_num_partitions;
  p.always_partials = 1;
  // partitions actually needed for max_context_len (never more than the caller's scratch holds)
  int np = (a->max_context_len + kV2PartitionTokens - 1) / kV2PartitionTokens;
  np = std::max(1, std::min(np, a->max_num_partitions));
  p.total_items = (long long)np * a->num_seqs * p.tiles_per_seq;
  const int sms = num_sms(a->device);
  int ctas_per_sm = 2;
  const int budget = pick_smem_budget(a, tile, &ctas_per_sm);
  st = attach_sched(p, stream);
  if (st != B200_OK) return st;
  st = dispatch(a, p, tile, budget, sms, stream);
  if (st != B200_OK) return st;
  return launch_reduce(a, a->max_num_partitions, kV2PartitionTokens, stream);
}

B200_API int64_t b200_paged_attention_v2_static_scratch_bytes(const b200_paged_attn_args_t* a) {
  if (!a || a->num_seqs <= 0 || a->num_heads <= 0 || a->head_size <= 0 || a->max_context_len <= 0) return 0;
  const int64_t np = (a->max_context_len + kV2PartitionTokens - 1) / kV2PartitionTokens;
  const int64_t rows = (int64_t)a->num_seqs * a->num_heads * np;
  return align256(rows * a->head_size * dtype_size(a->query_dtype)) + 2 * align256(rows * 4);
}

B200_API int b200_paged_attention_v2_static_scratch(const b200_paged_attn_args_t* a, b200_stream_t stream) {
  int st = check_common(a);
  if (st != B200_OK || a->num_seqs == 0) return st;
  if (a->max_context_len <= 0) return B200_ERR_ARG;
  const int64_t need = b200_paged_attention_v2_static_scratch_bytes(a);
  if (need > (int64_t)B200_V2_STATIC_SCRATCH_BYTES) return B200_ERR_SCRATCH_LIMIT;
  unsigned char* base = nullptr;
  {
    DeviceGuard guard;  // the symbol's address is per device: resolve it with the call's device current
    B200_CUDA_TRY(guard.set(a->device));
    st = static_scratch_base(&base);
    if (st != B200_OK) return st;
  }
  // Nothing in the scratch has to start at zero: the attention kernel writes every partial the reduce kernel reads
  // (partitions past a sequence's context are never read), so the previous call's contents are simply overwritten.
  const int64_t np = (a->max_context_len + kV2PartitionTokens - 1) / kV2PartitionTokens;
  const int64_t rows = (int64_t)a->num_seqs * a->num_heads * np;
  b200_paged_attn_args_t v2 = *a;
  v2.tmp_out = base;
  v2.exp_sums = reinterpret_cast<float*>(base + align256(rows * a->head_size * dtype_size(a->query_dtype)));
  v2.max_logits = v2.exp_sums + align256(rows * 4) / 4;
  v2.max_num_partitions = (int32_t)np;
  return b200_paged_attention_v2(&v2, stream);
}

}  // extern "C"
