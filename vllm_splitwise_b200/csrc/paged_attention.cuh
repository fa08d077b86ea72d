// Paged-KV attention decode for sm_100a: persistent kernel skeleton + the SIMT consumer.
//
// Replaces the reference's paged_attention_kernel / v1 / v2 / v2_reduce
// (csrc/attention/attention_kernels.cu:83-575).  Same math (fp32 logits, fp32 softmax with the
// 1e-6 guard on the denominator, fp32 P.V accumulation, tail tokens of the last block masked on
// both the logit and the V side) -- different execution model:
//
//  * PERSISTENT grid: (resident CTAs per SM) x (SM count) CTAs loop over work items
//    (partition, sequence, kv-head tile).  All query heads that share a KV head are processed by ONE
//    item, so every KV block is read from HBM once per KV head (the reference reads it once per QUERY
//    head, grid.x = num_heads, attention_kernels.cu:131-134).  Items are small (a partition of the
//    context), so 148 SMs stay evenly loaded whatever the batch / head count is, and the copy pipeline
//    below never drains between items.
//  * a producer warp walks the block table (one coalesced load per group of ring slots) and streams each
//    (block, kv_head) chunk -- contiguous D*bs*elt bytes for K and for V -- into a shared-memory ring
//    with cp.async.bulk (TMA engine, SASS UBLKCP) completing on mbarriers; each lane owns one ring slot,
//    so address math and barrier polls run 16-32 wide and only the copy issue itself is serial.
//  * consumer warps own ring slots round-robin and run an online (flash-decoding) softmax, so
//    shared memory is O(stage), not O(context) as in the reference's v1 (attention_kernels.cu:640-649),
//    and the context is walked once, not twice.  Two consumers exist: SIMT (this file; MHA, fp32,
//    block_size 8) and warp-level tensor cores (paged_attention_mma.cuh; grouped queries).
//  * per-warp partial states are merged per item; with partitions the partial state goes out in the
//    reference's own tmp_out / exp_sums / max_logits format and is combined either by the last CTA to
//    finish a sequence's partitions (v1, in-kernel, self-resetting tickets) or by
//    paged_attn_reduce_kernel (v2 signature, the reference's formula :484-575).
#pragma once

#include <float.h>
#include <stdlib.h>

#include "common.cuh"

namespace b200 {

struct AttnParams {
  void* out;
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int* block_tables;
  const int* context_lens;
  const float* alibi_slopes;
  float* exp_sums;
  float* max_logits;
  void* tmp_out;
  int* tickets;          // [num_seqs * tiles_per_seq], zero between launches; non-null = in-kernel reduce
  unsigned int* sched;   // work counter of this launch (zero between launches, self-resetting)
  float scale;
  int num_seqs;
  int num_heads;
  int num_kv_heads;
  int group;             // num_heads / num_kv_heads
  int gtiles;            // ceil(group / GT)
  int tiles_per_seq;     // num_kv_heads * gtiles
  int max_num_blocks_per_seq;
  int partition_blocks;  // KV blocks per partition; 0 = whole context
  int max_num_partitions;
  int always_partials;   // 1: v2 signature (every item writes tmp_out/exp_sums/max_logits)
  int nstages;           // ring depth; always a multiple of nconsumers
  int nconsumers;        // consumer warps that take blocks (<= kConsumerWarps)
  int q_bulk;            // 1: the query rows of an item reach the consumers through shared memory (bulk copy)
  int quad;              // 1: v2 signature on the tensor-core consumer -- every consumer warp owns a whole item
  int parts_per_item;    // quad mode: partitions one warp walks with ONE set of query fragments (1 or 2; 0 = 1)
  long long total_items; // max_num_partitions * num_seqs * tiles_per_seq
  long long q_stride;
  long long kv_block_stride;
  long long kv_head_stride;
#ifdef B200_ATTN_TRACE
  unsigned long long* trace;  // development builds only (make TRACE=1): clock64 stamps of CTA 0, tools/attn_trace.py
#endif
};
#ifdef B200_ATTN_TRACE
// who: consumer warp 0-3; quad (k-th work unit of the CTA) < 32; event < 8
#define B200_ATTN_STAMP(p, who, k, ev)                                                        \
  do {                                                                                        \
    if ((p).trace && blockIdx.x == 0 && (k) < 32 && (threadIdx.x & 31) == 0)                  \
      (p).trace[(((who) * 32 + (k)) * 8) + (ev)] = (unsigned long long)clock64();             \
  } while (0)
#else
#define B200_ATTN_STAMP(p, who, k, ev) do {} while (0)
#endif

constexpr int kConsumerWarps = 4;
constexpr int kConsumerThreads = kConsumerWarps * 32;
constexpr int kAttnThreads = (kConsumerWarps + 1) * 32;
constexpr int kMaxStages = 32;
constexpr int kItemQueue = 4;  // work items the producer may run ahead of the consumers

// One unit of work of the persistent grid.
struct ItemInfo {
  int seq, part, kvh, tile, head0, ng;  // tile = kvh * gtiles + gt; first query head of the tile, valid heads in it
  int ctx, nblk_ctx;              // context length (tokens, blocks)
  int b0, nblk;                   // first block and block count of this partition (nblk == 0: nothing to read)
  int nparts;                     // partitions this sequence has (0 for an empty context)
};

// ctx_known >= 0: the context length was already fetched (the producer passes it through the item queue so
// that consumers do not pay a global-memory round trip per item).
template <int GT, int BS>
__device__ __forceinline__ ItemInfo decode_item(const AttnParams& p, long long item, int ctx_known = -1) {
  ItemInfo it;
  const int per_part = p.num_seqs * p.tiles_per_seq;
  // 64-bit division is ~100 instructions; item ids fit 32 bits for every realistic launch (uniform branch)
  if (p.total_items <= 0x7fffffffLL) it.part = (int)((unsigned)item / (unsigned)per_part);
  else it.part = (int)(item / per_part);
  const int rem = (int)(item - (long long)it.part * per_part);
  it.seq = rem / p.tiles_per_seq;
  it.tile = rem - it.seq * p.tiles_per_seq;
  it.kvh = it.tile / p.gtiles;
  const int gt = it.tile - it.kvh * p.gtiles;
  it.head0 = it.kvh * p.group + gt * GT;
  it.ng = min(GT, p.group - gt * GT);
  it.ctx = ctx_known >= 0 ? ctx_known : p.context_lens[it.seq];
  it.nblk_ctx = (it.ctx + BS - 1) / BS;
  if (p.partition_blocks > 0) {
    // parts_per_item = 2 (quad mode): the item covers partitions 2 * idx and 2 * idx + 1; it.part names the first
    const int ppi = p.parts_per_item > 1 ? p.parts_per_item : 1;
    const int item_blocks = p.partition_blocks * ppi;
    it.part *= ppi;
    it.b0 = it.part * p.partition_blocks;
    it.nblk = max(0, min(it.nblk_ctx, it.b0 + item_blocks) - it.b0);
    it.nparts = (it.nblk_ctx + p.partition_blocks - 1) / p.partition_blocks;
  } else {
    it.b0 = 0;
    it.nblk = it.nblk_ctx;
    it.nparts = it.nblk_ctx > 0 ? 1 : 0;
  }
  return it;
}

// bulk copy without an eviction hint (query rows: re-read by the sequence's other partitions)
__device__ __forceinline__ void bulk_g2s_plain(void* smem_dst, const void* gsrc, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
               ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

__device__ __forceinline__ void consumer_bar() {  // the 4 consumer warps only; the producer never joins
  asm volatile("bar.sync 1, %0;" ::"n"(kConsumerThreads) : "memory");
}

// ---- producer ---------------------------------------------------------------------------------
// Ring position of the producer, carried across items: global block index modulo the ring depth and
// the number of times the ring has wrapped.
struct RingPos {
  int slot0;
  int fill0;
};

// Streams blocks [b0, b0+nblk) of one item into the ring.  Lane j of the warp handles block n0 + j of each
// group of G = min(nstages, 32) blocks, i.e. its own ring slot: the block-table load, the 64-bit address
// math, the empty-barrier poll and the expect_tx arrive all run G lanes wide; only UBLKCP itself is issued
// lane by lane.  Lanes whose slot is still being consumed keep polling while the others issue.
template <typename CT>
__device__ __forceinline__ void attn_produce_item(const AttnParams& p, const ItemInfo& it, RingPos& pos, int lane,
                                                  unsigned char* stages, int stage_bytes, int kbytes,
                                                  uint64_t* full_bar, uint64_t* empty_bar, uint64_t pol) {
  const int nst = p.nstages;
  const int G = nst < 32 ? nst : 32;
  const int* table = p.block_tables + (long long)it.seq * p.max_num_blocks_per_seq + it.b0;
  const unsigned char* kbase = static_cast<const unsigned char*>(p.k_cache) + (long long)it.kvh * p.kv_head_stride * (long long)sizeof(CT);
  const unsigned char* vbase = static_cast<const unsigned char*>(p.v_cache) + (long long)it.kvh * p.kv_head_stride * (long long)sizeof(CT);
  const long long blk_bytes = p.kv_block_stride * (long long)sizeof(CT);
  const int nblk = it.nblk;
  int phys_next = (lane < G && lane < nblk) ? table[lane] : 0;
  for (int n0 = 0; n0 < nblk; n0 += G) {
    const int phys = phys_next;
    const int nn = n0 + G + lane;
    phys_next = (lane < G && nn < nblk) ? table[nn] : 0;  // in flight while this group is issued
    int slot = pos.slot0 + lane, fill = pos.fill0;
    if (slot >= nst) { slot -= nst; ++fill; }
    bool pending = lane < G && n0 + lane < nblk;
    const long long off = (long long)phys * blk_bytes;
    unsigned char* dst = stages + (size_t)slot * stage_bytes;
    while (__any_sync(0xffffffffu, pending)) {
      if (pending && (fill == 0 || mbar_try_wait(&empty_bar[slot], (fill - 1) & 1))) {
        mbar_arrive_expect_tx(&full_bar[slot], 2 * kbytes);
        bulk_g2s(dst, kbase + off, kbytes, &full_bar[slot], pol);
        bulk_g2s(dst + kbytes, vbase + off, kbytes, &full_bar[slot], pol);
        pending = false;
      }
    }
    const int cnt = min(G, nblk - n0);
    pos.slot0 += cnt;
    if (pos.slot0 >= nst) { pos.slot0 -= nst; ++pos.fill0; }
  }
}

// ---- per-item epilogue (the 128 consumer threads) -------------------------------------------------
// Merge the consumer warps' partial (max, sum, acc) states of one item and write either the final output
// rows or the partition partials in the reference's tmp_out / exp_sums / max_logits format.
// red_acc [NW][GT][D] f32, red_ml [NW][GT][2] f32 (shared memory).  Returns nothing; `partial` says which.
template <typename T, int D, int GT>
__device__ __forceinline__ void attn_merge_store(const AttnParams& p, const ItemInfo& it, const float* red_acc,
                                                 const float* red_ml, bool partial, int ctid) {
  constexpr int NW = kConsumerWarps;
  T* out = static_cast<T*>(p.out);
  for (int idx = ctid; idx < it.ng * D; idx += kConsumerThreads) {
    const int g = idx / D, d = idx - g * D;
    float M = -FLT_MAX;
#pragma unroll
    for (int w = 0; w < NW; ++w) M = fmaxf(M, red_ml[(w * GT + g) * 2]);
    float L = 0.f, A = 0.f;
#pragma unroll
    for (int w = 0; w < NW; ++w) {
      const float sc = __expf(red_ml[(w * GT + g) * 2] - M);
      L = fmaf(red_ml[(w * GT + g) * 2 + 1], sc, L);
      A = fmaf(red_acc[(w * GT + g) * D + d], sc, A);
    }
    const float o = A * __fdividef(1.f, L + 1e-6f);
    const long long sh = (long long)it.seq * p.num_heads + it.head0 + g;
    if (partial) {
      const long long pi = sh * p.max_num_partitions + it.part;
      static_cast<T*>(p.tmp_out)[pi * D + d] = from_f32<T>(o);
      if (d == 0) {
        p.exp_sums[pi] = L;
        p.max_logits[pi] = M;
      }
    } else {
      out[sh * D + d] = from_f32<T>(o);
    }
  }
}

// Combine the partitions of one (sequence, head tile): the reference's v2 reduce formula
// (attention_kernels.cu:484-575): out = sum_j tmp_j * exp_sums_j * e^(max_j - max) / (sum + 1e-6).
// Run by the 128 consumer threads of the CTA that finished the sequence's last partition; partials of
// the other partitions were written by other CTAs, so they are read past L1 (ld.global.cg).
template <typename T, int D, int GT>
__device__ __forceinline__ void attn_reduce_in_kernel(const AttnParams& p, const ItemInfo& it, float* scratch, int ctid) {
  // scratch: shared memory, >= 2 * GT * np floats (the merge buffers are free at this point).
  // Step 1: all (head, partition) statistics in one round of independent loads.
  T* out = static_cast<T*>(p.out);
  const T* tmp_out = static_cast<const T*>(p.tmp_out);
  const int np = it.nparts;
  float* ml_s = scratch;
  float* w_s = scratch + GT * np;
  const long long sh0 = (long long)it.seq * p.num_heads + it.head0;
  for (int i = ctid; i < it.ng * np; i += kConsumerThreads) {
    const int g = i / np, j = i - g * np;
    ml_s[i] = __ldcg(p.max_logits + (sh0 + g) * p.max_num_partitions + j);
    w_s[i] = __ldcg(p.exp_sums + (sh0 + g) * p.max_num_partitions + j);
  }
  consumer_bar();
  // Step 2: per output element, weights from shared memory and the partials four loads at a time.
  for (int idx = ctid; idx < it.ng * D; idx += kConsumerThreads) {
    const int g = idx / D, d = idx - g * D;
    const T* tmp = tmp_out + (sh0 + g) * p.max_num_partitions * D + d;
    float M = -FLT_MAX;
    for (int j = 0; j < np; ++j) M = fmaxf(M, ml_s[g * np + j]);
    float sum = 0.f, a = 0.f;
    int j = 0;
    for (; j + 4 <= np; j += 4) {
      float v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) v[u] = ldcg_f32<T>(tmp + (long long)(j + u) * D);
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const float w = w_s[g * np + j + u] * __expf(ml_s[g * np + j + u] - M);
        sum += w;
        a = fmaf(v[u], w, a);
      }
    }
    for (; j < np; ++j) {
      const float w = w_s[g * np + j] * __expf(ml_s[g * np + j] - M);
      sum += w;
      a = fmaf(ldcg_f32<T>(tmp + (long long)j * D), w, a);
    }
    out[(sh0 + g) * D + d] = from_f32<T>(a * __fdividef(1.f, sum + 1e-6f));
  }
}

// ---- quad mode (v2 signature, tensor-core consumer) ----------------------------------------------
// Under the v2 signature every item is one 512-token partition: 32 blocks shared by four warps is 8 blocks per
// warp, then a cross-warp merge through shared memory and two CTA-wide barriers -- measured, that per-item
// epilogue left C4 through the v2 signature at 0.67 of the copy peak against 0.95 for the unsplit v1 call
// (profiles/r05_attn_short_items.txt).  Here the unit of work is a QUAD of four consecutive items and every
// consumer warp owns one of them outright: its own query fragments, all of its partition's blocks, and the
// partial written straight from its registers (Impl::store_partial) -- no merge, no barrier between items.
// The producer interleaves the four block streams (virtual block v -> warp v % 4, round v / 4), so ring slot
// ownership is unchanged (global block index modulo 4); rounds a shorter item does not have are passed through the
// ring as empty slots.  Query rows are staged in the (otherwise unused) merge buffer, two quads deep.
constexpr int kQuadQueue = 2;

template <class Impl>
__device__ __forceinline__ void attn_quad_mode(const AttnParams& p, unsigned char* stages, unsigned char* q_stage,
                                               long long* quad_q, int* ctx8, uint64_t* full_bar, uint64_t* empty_bar,
                                               uint64_t* iq_full, uint64_t* iq_empty, int warp, int lane) {
  using T = typename Impl::T;
  using CT = typename Impl::CT;
  constexpr int D = Impl::D, BS = Impl::BS, GT = Impl::GT;
  constexpr int NW = kConsumerWarps;
  constexpr int QB = GT * D * (int)sizeof(T);
  static_assert(NW == 4 && sizeof(T) == 2, "quad mode: four consumer warps, query staging aliases the merge buffer");
  const int nst = p.nstages;
  const long long total_quads = (p.total_items + 3) >> 2;

  if (warp == NW) {
    // ================= producer =================
    const uint64_t pol = l2_policy_evict_first();
    const int G = nst < 32 ? nst : 32;  // a multiple of 4 (nstages is)
    const int w = lane & 3;             // the consumer warp this lane feeds: virtual block n0 + lane, n0 % 4 == 0
    const int kbytes = Impl::kKBytes;
    const long long blk_bytes = p.kv_block_stride * (long long)sizeof(CT);
    int slot0 = 0, fill0 = 0;
    long long quad = blockIdx.x;
    const int per_part = p.num_seqs * p.tiles_per_seq;
    // The next quad is drawn while the last two groups of the current one are being issued, and its context lengths
    // and first block numbers are fetched behind that: the three dependent global round trips (ticket, context
    // length, block table) would otherwise drain the ring at every quad boundary.  Drawing only ~2 ring depths early
    // keeps the work a CTA has claimed but not started small (a whole item ahead measured slower, see the item mode).
    int ctx_pref = -1, phys_pref = 0;
    bool have_pref = false;
    for (int k = 0;; ++k) {
      const int qs = k % kQuadQueue;
      const bool done = quad >= total_quads;
      const long long item = 4 * quad + w;
      const bool valid = !done && item < p.total_items;
      ItemInfo it;
      it.nblk = 0; it.ctx = -1; it.seq = 0; it.kvh = 0; it.b0 = 0; it.ng = 0; it.head0 = 0;
      if (valid) it = decode_item<GT, BS>(p, item, have_pref ? ctx_pref : -1);
      const int my_n = valid ? it.nblk : 0;
      int rounds = my_n;
      rounds = max(rounds, __shfl_xor_sync(0xffffffffu, rounds, 1));
      rounds = max(rounds, __shfl_xor_sync(0xffffffffu, rounds, 2));
      // ---- queue entry: quad id, the four context lengths, the four query tiles ----
      if (lane == 0 && k >= kQuadQueue) mbar_wait(&iq_empty[qs], ((k / kQuadQueue) - 1) & 1);
      __syncwarp();
      if (lane < 4) ctx8[qs * 4 + lane] = valid ? it.ctx : -1;
      if (lane == 0) quad_q[qs] = done ? -1 : (quad | ((long long)rounds << 40));  // quad ids stay below 2^40
      uint32_t qbytes = (lane < 4 && my_n > 0 && p.q_bulk) ? (uint32_t)(it.ng * D * (int)sizeof(T)) : 0u;
      uint32_t qtot = qbytes + __shfl_xor_sync(0xffffffffu, qbytes, 1);
      qtot += __shfl_xor_sync(0xffffffffu, qtot, 2);
      __syncwarp();
      if (lane == 0) {
        if (qtot) mbar_arrive_expect_tx(&iq_full[qs], qtot);
        else mbar_arrive(&iq_full[qs]);
      }
      __syncwarp();
      if (qbytes)
        bulk_g2s_plain(q_stage + (qs * 4 + lane) * QB,
                       static_cast<const T*>(p.q) + (long long)it.seq * p.q_stride + (long long)it.head0 * D, qbytes,
                       &iq_full[qs]);
      if (done) break;
      // ---- the four block streams, interleaved ----
      const int* table = p.block_tables + (long long)it.seq * p.max_num_blocks_per_seq + it.b0;
      const unsigned char* kbase = static_cast<const unsigned char*>(p.k_cache) + (long long)it.kvh * p.kv_head_stride * (long long)sizeof(CT);
      const unsigned char* vbase = static_cast<const unsigned char*>(p.v_cache) + (long long)it.kvh * p.kv_head_stride * (long long)sizeof(CT);
      const int nvirt = 4 * rounds;
      int phys_next = have_pref ? phys_pref : ((lane < G && (lane >> 2) < my_n) ? table[lane >> 2] : 0);
      // next quad: 0 = not drawn, 1 = ticket in flight, 2 = id known and prefetch issued
      int stage_next = 0;
      unsigned int raw = 0;
      long long next_quad = 0;
      int ctx_next = -1, phys_n = 0;
      auto draw = [&]() {
        if (p.sched != nullptr && lane == 0) {
          raw = atomicAdd(p.sched, 1u);
          if ((long long)raw == total_quads - 1) *p.sched = 0u;  // the last draw of the launch: leave the counter at zero
        }
        stage_next = 1;
      };
      auto prefetch = [&]() {
        next_quad = p.sched == nullptr ? quad + gridDim.x : (long long)__shfl_sync(0xffffffffu, raw, 0) + gridDim.x;
        const long long ni = 4 * next_quad + w;
        ctx_next = -1;
        phys_n = 0;
        if (next_quad < total_quads && ni < p.total_items) {
          const int part = (int)(ni / per_part);
          const int seq = (int)(ni - (long long)part * per_part) / p.tiles_per_seq;
          ctx_next = p.context_lens[seq];
          const int b = part * p.partition_blocks * (p.parts_per_item > 1 ? p.parts_per_item : 1) + (lane >> 2);
          if (lane < G && b < p.max_num_blocks_per_seq) phys_n = p.block_tables[(long long)seq * p.max_num_blocks_per_seq + b];
        }
        stage_next = 2;
      };
      for (int n0 = 0; n0 < nvirt; n0 += G) {
        if (stage_next == 1) prefetch();
        else if (stage_next == 0 && n0 + 2 * G >= nvirt) draw();
        const int r = (n0 + lane) >> 2;
        const bool virt = lane < G && n0 + lane < nvirt;
        const bool real = virt && r < my_n;
        const int phys = phys_next;
        const int rn = (n0 + G + lane) >> 2;
        phys_next = (lane < G && rn < my_n) ? table[rn] : 0;  // in flight while this group is issued
        int slot = slot0 + lane, fill = fill0;
        if (slot >= nst) { slot -= nst; ++fill; }
        bool pending = virt;
        const long long off = (long long)phys * blk_bytes;
        unsigned char* dst = stages + (size_t)slot * Impl::kStageBytes;
        while (__any_sync(0xffffffffu, pending)) {
          if (pending && (fill == 0 || mbar_try_wait(&empty_bar[slot], (fill - 1) & 1))) {
            if (real) {
              mbar_arrive_expect_tx(&full_bar[slot], 2 * kbytes);
              bulk_g2s(dst, kbase + off, kbytes, &full_bar[slot], pol);
              bulk_g2s(dst + kbytes, vbase + off, kbytes, &full_bar[slot], pol);
            } else {
              mbar_arrive(&full_bar[slot]);  // a round this warp's item does not have: an empty slot
            }
            pending = false;
          }
        }
        const int cnt = min(G, nvirt - n0);
        slot0 += cnt;
        if (slot0 >= nst) { slot0 -= nst; ++fill0; }
      }
      if (stage_next == 0) draw();
      if (stage_next == 1) prefetch();
      quad = next_quad;
      ctx_pref = ctx_next;
      phys_pref = phys_n;
      have_pref = ctx_next >= 0;
    }
    return;
  }

  // ================= consumers: warp w owns item 4 * quad + w =================
  int slot = warp;
  uint32_t parity = 0;
  typename Impl::State st;
  for (int k = 0;; ++k) {
    const int qs = k % kQuadQueue;
    mbar_wait(&iq_full[qs], (k / kQuadQueue) & 1);
    const long long entry = quad_q[qs];
    if (entry < 0) break;
    B200_ATTN_STAMP(p, warp, k, 0);
    const long long quad = entry & ((1LL << 40) - 1);
    const int rounds = (int)(entry >> 40);  // the longest of the four items, from the producer
    const long long item = 4 * quad + warp;
    const int ctx = ctx8[qs * 4 + warp];
    ItemInfo it;
    it.nblk = 0;
    if (ctx >= 0) it = decode_item<GT, BS>(p, item, ctx);
    const int my_n = ctx >= 0 ? it.nblk : 0;
    if (my_n > 0) {
      Impl::init(st);
      const T* qrows = p.q_bulk ? reinterpret_cast<const T*>(q_stage + (qs * 4 + warp) * QB)
                                : static_cast<const T*>(p.q) + (long long)it.seq * p.q_stride + (long long)it.head0 * D;
      Impl::begin_item(st, p, it, qrows, nullptr, warp, lane);
    }
    __syncwarp();
    if (lane == 0) mbar_arrive(&iq_empty[qs]);  // every warp arrives, with or without an item
    B200_ATTN_STAMP(p, warp, k, 1);
    const int b_first = it.b0;
    int part_end = p.partition_blocks;  // first round of the item's next partition
    for (int r = 0; r < rounds; ++r) {
      unsigned char* ks = stages + (size_t)slot * Impl::kStageBytes;
      mbar_wait(&full_bar[slot], parity);
      if (r == 0) B200_ATTN_STAMP(p, warp, k, 2);
      if (r == 1) B200_ATTN_STAMP(p, warp, k, 5);
      if (r < my_n) {
        if (r > 0 && r == part_end) {
          // two partitions per item: the first one is complete -- store it, keep the query fragments, start over
          Impl::store_partial(st, p, it, lane);
          Impl::init(st);
          ++it.part;
          part_end += p.partition_blocks;
        }
        Impl::block(st, p, it, ks, ks + Impl::kKBytes, b_first + r, nullptr, warp, lane);
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[slot]);
      slot += NW;
      if (slot >= nst) { slot -= nst; parity ^= 1u; }
    }
    B200_ATTN_STAMP(p, warp, k, 3);
    if (my_n > 0) Impl::store_partial(st, p, it, lane);
    B200_ATTN_STAMP(p, warp, k, 4);
  }
}

// ---- the persistent kernel ------------------------------------------------------------------
// Impl supplies the consumer: types T (query/out) and CT (cache), tile constants D, BS, GT,
// kKBytes/kStageBytes, kFixedBytes of private shared memory, a per-thread State and
//   init(State&)                                   once per item, registers only
//   begin_item(State&, p, it, fixed, warp, lane)   load the item's queries (may use consumer_bar)
//   block(State&, p, it, ks, vs, blk, fixed, warp, lane)   one KV block from ring memory
//   end_item(State&, it, red_acc, red_ml, warp, lane)      park the warp's (max, sum, acc) state
template <class Impl>
__global__ void __launch_bounds__(kAttnThreads) paged_attn_kernel(const __grid_constant__ AttnParams p) {
  using T = typename Impl::T;
  using CT = typename Impl::CT;
  constexpr int D = Impl::D, BS = Impl::BS, GT = Impl::GT;
  constexpr int NW = kConsumerWarps;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int nst = p.nstages;

  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* stages = smem;
  unsigned char* fixed = smem + (size_t)nst * Impl::kStageBytes;
  float* red_acc = reinterpret_cast<float*>(fixed + Impl::kFixedBytes);
  float* red_ml = red_acc + NW * GT * D;
  int* s_flag = reinterpret_cast<int*>(red_ml + NW * GT * 2);
  long long* item_q = reinterpret_cast<long long*>(s_flag + 4);  // [kItemQueue] items handed producer -> consumers
  int* ctx_q = reinterpret_cast<int*>(item_q + kItemQueue);      // [kItemQueue] their context lengths
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(ctx_q + kItemQueue);
  uint64_t* empty_bar = full_bar + kMaxStages;
  uint64_t* iq_full = empty_bar + kMaxStages;
  uint64_t* iq_empty = iq_full + kItemQueue;
  constexpr int QB = GT * D * (int)sizeof(T);  // query rows of one item
  unsigned char* q_stage = reinterpret_cast<unsigned char*>(iq_empty + kItemQueue);  // [kItemQueue][QB]

  if (tid == 0) {
    for (int s = 0; s < nst; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    for (int s = 0; s < kItemQueue; ++s) {
      mbar_init(&iq_full[s], 1);
      mbar_init(&iq_empty[s], NW);
    }
    mbar_fence_init();
  }
  __syncthreads();
  // Everything above overlaps the tail of the previous kernel of the stream when this launch is a programmatic
  // dependent (it always is: launch_attn); nothing below may run before that kernel's memory is visible.
  griddep_wait();

  if constexpr (Impl::kQuad) {
    if (p.quad) {  // query staging (2 quads x 4 warps) aliases the merge buffer; ids and context lengths the item queue
      attn_quad_mode<Impl>(p, stages, reinterpret_cast<unsigned char*>(red_acc), item_q,
                           reinterpret_cast<int*>(item_q + kQuadQueue), full_bar, empty_bar, iq_full, iq_empty, warp, lane);
      return;
    }
  }

  if (warp == NW) {
    // ================= producer =================
#ifdef B200_ATTN_NO_HINT
    const uint64_t pol = 0;
#else
    const uint64_t pol = l2_policy_evict_first();
#endif
    // Work items are drawn from a device-wide counter (the first one is blockIdx.x), so CTAs on slower SMs
    // simply take fewer; every item id is handed to the consumer warps through a small shared-memory queue.
    RingPos pos{0, 0};
    const T* qbase = static_cast<const T*>(p.q);
    // hand item `item` (context length ctx) to the consumers: id, context length and -- q_bulk -- its query rows
    auto enqueue = [&](int k, long long item, bool done, const ItemInfo& it) {
      const int qs = k % kItemQueue;
      if (lane == 0) {
        if (k >= kItemQueue) mbar_wait(&iq_empty[qs], ((k / kItemQueue) - 1) & 1);
        item_q[qs] = item;
        ctx_q[qs] = done ? 0 : it.ctx;
        if (!done && p.q_bulk && it.nblk > 0) {
          const uint32_t bytes = (uint32_t)(it.ng * D * (int)sizeof(T));
          mbar_arrive_expect_tx(&iq_full[qs], bytes);
          bulk_g2s_plain(q_stage + qs * QB, qbase + (long long)it.seq * p.q_stride + (long long)it.head0 * D, bytes,
                         &iq_full[qs]);
        } else {
          mbar_arrive(&iq_full[qs]);
        }
      }
    };
    long long item = blockIdx.x;
    for (int k = 0;; ++k) {
      const bool done = item >= p.total_items;
      ItemInfo it;
      if (!done) it = decode_item<GT, BS>(p, item);
      enqueue(k, item, done, it);
      if (done) break;
      if (it.nblk > 0)
        attn_produce_item<CT>(p, it, pos, lane, stages, Impl::kStageBytes, Impl::kKBytes, full_bar, empty_bar, pol);
      // (Measured, profiles/r05_attn_short_items.txt: drawing and prefetching two items ahead to hide the ticket /
      // context-length / block-table round trips between short items made every short-item shape SLOWER -- the
      // extra claimed item costs more at the tail than the bubbles it removes.)
      if (p.sched == nullptr) {  // no counter slot left for this launch (see attach_sched): static round-robin
        item += gridDim.x;
        continue;
      }
      unsigned int raw = 0;
      if (lane == 0) {
        raw = atomicAdd(p.sched, 1u);
        if ((long long)raw == p.total_items - 1) *p.sched = 0u;  // the last draw of the launch: leave the counter at zero
      }
      raw = __shfl_sync(0xffffffffu, raw, 0);
      item = (long long)raw + gridDim.x;
    }
    return;
  }

  // ================= consumers =================
  // Warp w owns ring slots w, w + nwa, ... (nstages % nwa == 0), i.e. global block indices congruent to w:
  // every fill of a slot is consumed by the same warp, in order, so a parity wait is unambiguous.
  const int nwa = p.nconsumers;
  const bool active = warp < nwa;
  int my_gi = warp;        // next global block index of this warp
  int slot = warp;         // = my_gi % nstages
  uint32_t parity = 0;     // = (my_gi / nstages) & 1
  int gbase = 0;           // global block index of the current item's first block
  typename Impl::State st;

  for (int k = 0;; ++k) {
    const int qs = k % kItemQueue;
    mbar_wait(&iq_full[qs], (k / kItemQueue) & 1);
    const long long item = item_q[qs];
    const int ctx_known = ctx_q[qs];
    __syncwarp();
    if (item >= p.total_items) break;
    const ItemInfo it = decode_item<GT, BS>(p, item, ctx_known);
    if (it.nblk == 0) {
      if (lane == 0) mbar_arrive(&iq_empty[qs]);
      // nothing to read.  An empty context is softmax over nothing -> zeros (reference: exp_sum = 0, acc = 0);
      // under the v2 signature the reduce kernel writes them (it sees zero partitions).
      if (it.nblk_ctx == 0 && it.part == 0 && !p.always_partials) {
        T* out = static_cast<T*>(p.out);
        for (int i = tid; i < it.ng * D; i += kConsumerThreads)
          out[((long long)it.seq * p.num_heads + it.head0) * D + i] = from_f32<T>(0.f);
      }
      continue;
    }
    Impl::init(st);
    // the item's query rows [ng][D]: staged by the producer next to the queue entry, or read from global memory
    const T* qrows = p.q_bulk ? reinterpret_cast<const T*>(q_stage + qs * QB)
                              : static_cast<const T*>(p.q) + (long long)it.seq * p.q_stride + (long long)it.head0 * D;
    Impl::begin_item(st, p, it, qrows, fixed, warp, lane);
    __syncwarp();
    if (lane == 0) mbar_arrive(&iq_empty[qs]);  // the queue entry (and its query rows) may be reused
    if (active) {
      const int gend = gbase + it.nblk;
      for (; my_gi < gend; my_gi += nwa) {
        unsigned char* ks = stages + (size_t)slot * Impl::kStageBytes;
        mbar_wait(&full_bar[slot], parity);
        Impl::block(st, p, it, ks, ks + Impl::kKBytes, it.b0 + (my_gi - gbase), fixed, warp, lane);
        __syncwarp();
        if (lane == 0) mbar_arrive(&empty_bar[slot]);
        slot += nwa;
        if (slot >= nst) { slot -= nst; parity ^= 1u; }
      }
    }
    gbase += it.nblk;
    Impl::end_item(st, it, red_acc, red_ml, warp, lane);
    consumer_bar();
    const bool partial = p.always_partials || it.nparts > 1;
    attn_merge_store<T, D, GT>(p, it, red_acc, red_ml, partial, tid);
    if (partial && p.tickets != nullptr) {
      // v1 with an internal split: the CTA that completes the sequence's last partition combines them.
      __threadfence();
      consumer_bar();
      if (tid == 0) {
        int* ticket = p.tickets + it.seq * p.tiles_per_seq + it.tile;
        const int prev = atomicAdd(ticket, 1);
        const int last = prev == it.nparts - 1;
        if (last) *ticket = 0;  // self-resetting: the array is all-zero again when the kernel ends
        *s_flag = last;
      }
      consumer_bar();
      if (*s_flag) {
        __threadfence();
        attn_reduce_in_kernel<T, D, GT>(p, it, red_acc, tid);  // red_acc: NW*GT*D floats >= 2*GT*64
      }
    }
    consumer_bar();  // red_acc / red_ml / the impl's query tile may be rewritten by the next item
  }
}

// ---- SIMT consumer ----------------------------------------------------------------------------
template <typename CT, int N>
__device__ __forceinline__ void load_unpack(const unsigned char* p, float* f) {
  // N cache elements starting at p (N * sizeof(CT) is 16, or 8 for fp8 with block_size 8)
  if constexpr (N * sizeof(CT) == 16) {
    Unpack16<CT>::run(*reinterpret_cast<const uint4*>(p), f);
  } else {
    static_assert(sizeof(CT) == 1 && N == 8, "only fp8 x 8 tokens is narrower than 16 bytes");
    const uint2 v = *reinterpret_cast<const uint2*>(p);
    float2 a = e5m2x2_lo_to_float2(v.x), b = e5m2x2_hi_to_float2(v.x);
    float2 c = e5m2x2_lo_to_float2(v.y), d = e5m2x2_hi_to_float2(v.y);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
  }
}

// One token per thread group for Q.K, 16-byte shared-memory vectors, fp32 everywhere.  GT_ query heads of
// one KV head per item (1 for MHA; 8 for grouped queries that cannot take the tensor-core consumer).
template <typename T_, typename CT_, int D_, int BS_, int GT_>
struct SimtImpl {
  using T = T_;
  using CT = CT_;
  static constexpr int D = D_, BS = BS_, GT = GT_;
  static constexpr int NW = kConsumerWarps;
  static constexpr int X = 16 / (int)sizeof(CT);
  static constexpr int kKBytes = D * BS * (int)sizeof(CT);
  static constexpr int kStageBytes = 2 * kKBytes;
  static constexpr int kFixedBytes = GT * D * 4 + NW * GT * BS * 4;  // q (fp32) | per-warp probabilities
  static constexpr int NVK = (D / X) * BS;           // 16-byte vectors in a K chunk
  static constexpr int KIT = (NVK + 31) / 32;
  static constexpr int VV = (X < BS) ? X : BS;       // tokens per V vector
  static constexpr int VBYTES = VV * (int)sizeof(CT);
  static constexpr int VPR = BS / VV;                // vectors per V row
  static constexpr int NVV = D * VPR;
  static constexpr int VIT = (NVV + 31) / 32;

  struct State {
    float m_run[GT], l_run[GT], alibi[GT];
    float acc[GT][VIT];
  };
  static constexpr bool kQuad = false;  // the SIMT consumer shares its query tile between the warps
  __device__ __forceinline__ static void store_partial(State&, const AttnParams&, const ItemInfo&, int) {}

  __device__ __forceinline__ static void init(State& st) {
#pragma unroll
    for (int g = 0; g < GT; ++g) {
      st.m_run[g] = -FLT_MAX;
      st.l_run[g] = 0.f;
#pragma unroll
      for (int i = 0; i < VIT; ++i) st.acc[g][i] = 0.f;
    }
  }

  __device__ __forceinline__ static void begin_item(State& st, const AttnParams& p, const ItemInfo& it, const T* q,
                                                    unsigned char* fixed, int warp, int lane) {
    float* q_s = reinterpret_cast<float*>(fixed);
    for (int i = warp * 32 + lane; i < GT * D; i += kConsumerThreads) q_s[i] = (i < it.ng * D) ? to_f32<T>(q[i]) : 0.f;
#pragma unroll
    for (int g = 0; g < GT; ++g)
      st.alibi[g] = (p.alibi_slopes != nullptr && g < it.ng) ? p.alibi_slopes[it.head0 + g] : 0.f;
    consumer_bar();
  }

  __device__ __forceinline__ static void block(State& st, const AttnParams& p, const ItemInfo& it,
                                               const unsigned char* ks, const unsigned char* vs, int blk,
                                               unsigned char* fixed, int warp, int lane) {
    const float* q_s = reinterpret_cast<const float*>(fixed);
    float* plog_w = reinterpret_cast<float*>(fixed) + GT * D + warp * GT * BS;
    const int tok = lane % BS;
    const int vpart = lane % VPR;
    const int ctx = it.ctx;

    // ---- logits: one token per thread group, 16-byte vectors, fp32 accumulate ----
    float s[GT];
#pragma unroll
    for (int g = 0; g < GT; ++g) s[g] = 0.f;
#pragma unroll
    for (int i = 0; i < KIT; ++i) {
      const int v = lane + 32 * i;
      if ((NVK % 32 == 0) || v < NVK) {
        const int chunk = v / BS;
        float kf[X];
        load_unpack<CT, X>(ks + v * 16, kf);
#pragma unroll
        for (int g = 0; g < GT; ++g) {
          const float4* qp = reinterpret_cast<const float4*>(q_s + g * D + chunk * X);
#pragma unroll
          for (int e = 0; e < X / 4; ++e) {
            const float4 qq = qp[e];
            s[g] = fmaf(qq.x, kf[4 * e + 0], s[g]);
            s[g] = fmaf(qq.y, kf[4 * e + 1], s[g]);
            s[g] = fmaf(qq.z, kf[4 * e + 2], s[g]);
            s[g] = fmaf(qq.w, kf[4 * e + 3], s[g]);
          }
        }
      }
    }
    const int t_idx = blk * BS + tok;
    const bool valid = t_idx < ctx;
    float corr[GT];
#pragma unroll
    for (int g = 0; g < GT; ++g) {
#pragma unroll
      for (int m = 16; m >= BS; m >>= 1) s[g] += __shfl_xor_sync(0xffffffffu, s[g], m);
      float sv = s[g] * p.scale;
      sv += (st.alibi[g] != 0.f) ? st.alibi[g] * (float)(t_idx - ctx + 1) : 0.f;
      sv = valid ? sv : -FLT_MAX;
      float bmax = sv;
#pragma unroll
      for (int m = BS / 2; m >= 1; m >>= 1) bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, m));
      const float m_new = fmaxf(st.m_run[g], bmax);
      corr[g] = __expf(st.m_run[g] - m_new);
      const float pe = valid ? __expf(sv - m_new) : 0.f;
      st.l_run[g] = st.l_run[g] * corr[g] + (lane < BS ? pe : 0.f);
      st.m_run[g] = m_new;
      if (lane < BS) plog_w[g * BS + tok] = pe;
    }
    __syncwarp();

    // ---- P.V: each lane owns VV consecutive tokens of some rows of the [D, BS] tile ----
    float pv[GT][VV];
#pragma unroll
    for (int g = 0; g < GT; ++g) {
#pragma unroll
      for (int j = 0; j < VV; j += 4) {
        const float4 t4 = *reinterpret_cast<const float4*>(plog_w + g * BS + vpart * VV + j);
        pv[g][j] = t4.x; pv[g][j + 1] = t4.y; pv[g][j + 2] = t4.z; pv[g][j + 3] = t4.w;
      }
    }
    const bool last = (blk == it.nblk_ctx - 1);
#pragma unroll
    for (int i = 0; i < VIT; ++i) {
      const int v = lane + 32 * i;
      if ((NVV % 32 == 0) || v < NVV) {
        float vf[VV];
        load_unpack<CT, VV>(vs + v * VBYTES, vf);
        if (last) {
          // stale slots past the context may hold NaN/Inf bit patterns (reference :342-351)
#pragma unroll
          for (int j = 0; j < VV; ++j)
            if (blk * BS + vpart * VV + j >= ctx) vf[j] = 0.f;
        }
#pragma unroll
        for (int g = 0; g < GT; ++g) {
          float a = st.acc[g][i] * corr[g];
#pragma unroll
          for (int j = 0; j < VV; ++j) a = fmaf(pv[g][j], vf[j], a);
          st.acc[g][i] = a;
        }
      }
    }
  }

  __device__ __forceinline__ static void end_item(State& st, const ItemInfo& it, float* red_acc, float* red_ml,
                                                  int warp, int lane) {
    const int vpart = lane % VPR;
#pragma unroll
    for (int g = 0; g < GT; ++g) {
      float l = st.l_run[g];
#pragma unroll
      for (int m = 16; m >= 1; m >>= 1) l += __shfl_xor_sync(0xffffffffu, l, m);
      if (lane == 0) {
        red_ml[(warp * GT + g) * 2 + 0] = st.m_run[g];
        red_ml[(warp * GT + g) * 2 + 1] = l;
      }
#pragma unroll
      for (int i = 0; i < VIT; ++i) {
        float a = st.acc[g][i];
#pragma unroll
        for (int m = VPR / 2; m >= 1; m >>= 1) a += __shfl_xor_sync(0xffffffffu, a, m);
        const int v = lane + 32 * i;
        if (vpart == 0 && ((NVV % 32 == 0) || v < NVV)) red_acc[(warp * GT + g) * D + v / VPR] = a;
      }
    }
  }
};

// ---- v2 signature: separate merge of the partition partials -------------------------------------
// Same formula as the reference's paged_attention_v2_reduce_kernel (attention_kernels.cu:484-575).
// Launched with programmatic stream serialization: its CTAs may be resident before the attention
// kernel has finished; griddepcontrol.wait orders it after that kernel's memory.
template <typename T>
__global__ void __launch_bounds__(128) paged_attn_reduce_kernel(
    T* __restrict__ out, const float* __restrict__ exp_sums, const float* __restrict__ max_logits,
    const T* __restrict__ tmp_out, const int* __restrict__ context_lens, int head_size,
    int max_num_partitions, int partition_tokens) {
  const int head = blockIdx.y, seq = blockIdx.x, num_heads = gridDim.y;  // sequences on x: no 65 535 limit
  const int ctx = context_lens[seq];
  const int np = (ctx + partition_tokens - 1) / partition_tokens;
  const long long sh = (long long)seq * num_heads + head;
  const T* tmp = tmp_out + sh * max_num_partitions * head_size;
  T* o = out + sh * head_size;
  asm volatile("griddepcontrol.wait;" ::: "memory");
  if (np <= 1) {
    for (int d = threadIdx.x; d < head_size; d += blockDim.x) o[d] = np == 1 ? tmp[d] : from_f32<T>(0.f);
    return;
  }
  extern __shared__ float wsm[];  // [np] rescaled exp sums
  __shared__ float red[8];
  const float* ml = max_logits + sh * max_num_partitions;
  const float* es = exp_sums + sh * max_num_partitions;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  float m = -FLT_MAX;
  for (int j = threadIdx.x; j < np; j += blockDim.x) m = fmaxf(m, ml[j]);
#pragma unroll
  for (int k = 16; k >= 1; k >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, k));
  if (lane == 0) red[warp] = m;
  __syncthreads();
  m = fmaxf(fmaxf(red[0], red[1]), fmaxf(red[2], red[3]));
  float sum = 0.f;
  for (int j = threadIdx.x; j < np; j += blockDim.x) {
    const float w = es[j] * expf(ml[j] - m);
    wsm[j] = w;
    sum += w;
  }
#pragma unroll
  for (int k = 16; k >= 1; k >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, k);
  if (lane == 0) red[4 + warp] = sum;
  __syncthreads();
  sum = red[4] + red[5] + red[6] + red[7];
  const float inv = __fdividef(1.f, sum + 1e-6f);
  for (int d = threadIdx.x; d < head_size; d += blockDim.x) {
    float a = 0.f;
    for (int j = 0; j < np; ++j) a += to_f32<T>(tmp[(long long)j * head_size + d]) * wsm[j] * inv;
    o[d] = from_f32<T>(a);
  }
}

// ---- host-side launch plumbing shared by both consumers ---------------------------------------------
// Ring depth and shared-memory size for a consumer under a per-CTA budget.
inline int attn_max_stages(int stage_bytes) {
  // (a 24-stage ring for 4 KiB stages measured no faster than 16 on any shape: profiles/r05_attn_short_items.txt)
  static const int env = getenv("B200_ATTN_MAX_STAGES") ? atoi(getenv("B200_ATTN_MAX_STAGES")) : 0;  // tuning knob
  (void)stage_bytes;
  if (env >= 2 && env <= kMaxStages) return env;
  return 16;
}
inline int attn_smem_tail(int fixed_bytes, int gt, int head_size, int q_elt_bytes) {
  return fixed_bytes + kConsumerWarps * gt * head_size * 4 + kConsumerWarps * gt * 2 * 4 + 16 + kItemQueue * 16 +
         2 * kMaxStages * 8 + 2 * kItemQueue * 8 + kItemQueue * gt * head_size * q_elt_bytes;
}
inline bool attn_smem_plan(int stage_bytes, int fixed_bytes, int gt, int head_size, int q_elt_bytes, int smem_budget,
                           int* nstages, int* nconsumers, int* smem_bytes) {
  const int tail = attn_smem_tail(fixed_bytes, gt, head_size, q_elt_bytes);
  int n = (smem_budget - tail) / stage_bytes;
  if (n > attn_max_stages(stage_bytes)) n = attn_max_stages(stage_bytes);
  // ring depth must be a multiple of the number of consuming warps (same warp per slot, see the kernel)
  const int nwa = n < kConsumerWarps ? n : kConsumerWarps;
  if (nwa > 0) n = n / nwa * nwa;
  *nstages = n;
  *nconsumers = nwa;
  *smem_bytes = n * stage_bytes + tail;
  return n >= 2;
}

// Launch the persistent kernel for one consumer: grid = resident CTAs per SM x SMs, capped by the item count.
template <class Impl>
int launch_attn(const AttnParams& p_in, int smem_budget, int sms, cudaStream_t stream) {
  AttnParams p = p_in;
  int nst = 0, nwa = 0, smem = 0;
  if (!attn_smem_plan(Impl::kStageBytes, Impl::kFixedBytes, Impl::GT, Impl::D, (int)sizeof(typename Impl::T), smem_budget,
                      &nst, &nwa, &smem))
    return B200_ERR_ARG;
  p.nstages = nst;
  p.nconsumers = nwa;
  // query rows travel through shared memory when a bulk copy can fetch them (16-byte aligned rows)
  static const bool no_q_bulk = getenv("B200_ATTN_NO_QBULK") != nullptr;      // A/B knob
  p.q_bulk = !no_q_bulk && (reinterpret_cast<uintptr_t>(p.q) % 16 == 0) &&
             ((p.q_stride * (long long)sizeof(typename Impl::T)) % 16 == 0) &&
             ((Impl::D * (int)sizeof(typename Impl::T)) % 16 == 0);

  auto kern = paged_attn_kernel<Impl>;
  // per instantiation and per device (function attributes are per device): smem size -> resident CTAs per SM
  static int occ_smem[64] = {0}, occ_val[64] = {0};
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  dev &= 63;
  if (occ_smem[dev] != smem) {
    B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    int occ = 0;
    B200_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&occ, kern, kAttnThreads, smem));
    if (occ < 1) return B200_ERR_ARG;
    occ_val[dev] = occ;
    occ_smem[dev] = smem;
  }
  static const bool no_quad = getenv("B200_ATTN_NO_QUAD") != nullptr;  // A/B knob
  long long grid = (long long)occ_val[dev] * sms;
  // quad mode needs enough quads to keep every resident CTA busy (measured, profiles/r05_attn_quad.txt: with 128
  // quads for 296 CTAs the reference benchmark shape loses 7 %, with 512 quads C3 tp=8 gains 13 %)
  p.quad = Impl::kQuad && !no_quad && p.always_partials && nwa == kConsumerWarps && p.partition_blocks > 0 &&
           (p.total_items + 3) / 4 >= grid * 3 / 2;
  p.parts_per_item = 1;
  if (p.quad) {
    // Two partitions per warp-item: the per-item prologue (item decode, query fragments: ~2 100 cycles per 32-block
    // partition, profiles/r05_attn_trace_c4_v2.txt) is paid once for both, the partials are still written per
    // partition.  Only with at least six of the doubled quads per resident CTA -- with fewer, the tail of whole
    // quads costs more than the prologues (measured: C4, 6.9 per CTA, 0.843 -> 0.904; ctx 32 768 at batch 16, 3.5 per
    // CTA, 1.03 -> 1.00).
    static const bool no_pair = getenv("B200_ATTN_NO_PAIR") != nullptr;  // A/B knob
    static const long long pair_min =  // test knob: doubled quads per resident CTA from which pairs are used
        getenv("B200_ATTN_PAIR_MIN_QUADS") ? atoll(getenv("B200_ATTN_PAIR_MIN_QUADS")) : 6;
    const long long per_part = (long long)p.num_seqs * p.tiles_per_seq;
    const long long np = p.total_items / per_part;
    const long long pair_items = ((np + 1) / 2) * per_part;
    if (!no_pair && np >= 2 && (pair_items + 3) / 4 >= grid * pair_min) {
      p.parts_per_item = 2;
      p.total_items = pair_items;
    }
  }
  const long long units = p.quad ? (p.total_items + 3) / 4 : p.total_items;
  if (grid > units) grid = units;
  B200_CUDA_TRY(launch_kernel(kern, dim3((unsigned)grid), dim3(kAttnThreads), (size_t)smem, stream, /*pdl=*/true, p));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// One instantiation family per (T, CT); defined in attn_inst_*.cu / attn_mma_inst_*.cu so they compile in
// parallel.  gt: query heads per item (SIMT: 1 or 8; MMA: 8 or 16).
template <typename T, typename CT>
int launch_simt(const AttnParams& p, int head_size, int block_size, int gt, int smem_budget, int sms, cudaStream_t stream);
template <typename T, typename CT>
int launch_mma(const AttnParams& p, int head_size, int block_size, int gt, int smem_budget, int sms, cudaStream_t stream);

}  // namespace b200
