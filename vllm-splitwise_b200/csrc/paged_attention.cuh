// Paged-KV attention decode for sm_100a: kernel templates.
//
// Replaces the reference's paged_attention_kernel / v1 / v2 / v2_reduce
// (csrc/attention/attention_kernels.cu:83-575).  Same math (fp32 logits, fp32 softmax with the
// 1e-6 guard on the denominator, fp32 P.V accumulation, tail tokens of the last block masked on
// both the logit and the V side) -- different execution model:
//
//  * grid = (kv_head x q-group tile, seq, partition): the query heads that share a KV head are
//    processed by ONE CTA, so every KV block is read from HBM once per KV head (the reference
//    reads it once per QUERY head, grid.x = num_heads, attention_kernels.cu:131-134).
//  * a producer warp walks the block table (32 entries per coalesced load, prefetched one group
//    ahead) and streams each (block, kv_head) chunk -- contiguous D*bs*elt bytes for K and for V --
//    into a shared-memory ring with cp.async.bulk (TMA engine, SASS UBLKCP) completing on
//    mbarriers.  Tens of KiB are in flight per SM without costing a register.
//  * consumer warps take ring stages round-robin and run an online (flash-decoding) softmax, so
//    shared memory is O(stage), not O(context) as in the reference's v1 (attention_kernels.cu:640-649),
//    and the context is walked once, not twice.
//  * per-warp partial states are merged once at the end; with partitions (v2, or the internal
//    split used to fill the GPU) the partial state goes out in the reference's own
//    tmp_out / exp_sums / max_logits format.
#pragma once

#include <float.h>

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
  float scale;
  int num_heads;
  int num_kv_heads;
  int group;             // num_heads / num_kv_heads
  int gtiles;            // ceil(group / GT)
  int max_num_blocks_per_seq;
  int partition_blocks;  // KV blocks per partition; 0 = whole context
  int max_num_partitions;
  int write_partials;    // 1: tmp_out/exp_sums/max_logits, 0: final out
  int nstages;           // ring depth; always a multiple of nconsumers (see simt_smem_plan)
  int nconsumers;        // consumer warps that take blocks (<= kConsumerWarps)
  long long q_stride;
  long long kv_block_stride;
  long long kv_head_stride;
};

constexpr int kConsumerWarps = 4;
constexpr int kAttnThreads = (kConsumerWarps + 1) * 32;

template <typename CT, int D, int BS, int GT>
struct SimtSmem {
  static constexpr int kKBytes = D * BS * (int)sizeof(CT);
  static constexpr int kStageBytes = 2 * kKBytes;
  static constexpr int kFixedBytes =
      GT * D * 4 +                           // q (fp32)
      kConsumerWarps * GT * BS * 4 +         // per-warp probabilities
      kConsumerWarps * GT * D * 4 +          // per-warp accumulators for the final merge
      kConsumerWarps * GT * 2 * 4 +          // per-warp (max, sum)
      2 * 32 * 8;                            // mbarriers (up to 32 stages)
  static constexpr int kMaxStages = 32;
};

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

// Merge the consumer warps' partial (max, sum, acc) states and write either the final output row or
// the partition partials in the reference's tmp_out / exp_sums / max_logits format.
// red_acc [NW][GT][D] f32, red_ml [NW][GT][2] f32 (shared memory).
template <typename T, int D, int GT>
__device__ __forceinline__ void attn_merge_store(const AttnParams& p, const float* red_acc, const float* red_ml,
                                                 int ng, int seq, int head0, int part, int tid) {
  constexpr int NW = kConsumerWarps;
  T* out = static_cast<T*>(p.out);
  for (int idx = tid; idx < ng * D; idx += kAttnThreads) {
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
    const long long sh = (long long)seq * p.num_heads + head0 + g;
    if (p.write_partials) {
      const long long pi = sh * p.max_num_partitions + part;
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

// Producer warp (shared by the SIMT and the tensor-core kernels): walks the block table of `seq`
// from block b0 (32 entries per coalesced load, the next group prefetched while this one is issued)
// and streams each (block, kv_head) chunk -- kbytes contiguous bytes of K, then of V -- into ring
// stage n % nstages with two bulk copies completing on full_bar[stage].
template <typename CT>
__device__ __forceinline__ void attn_producer(const AttnParams& p, int seq, int kvh, int b0, int nblk, int lane,
                                              unsigned char* stages, int stage_bytes, int kbytes,
                                              uint64_t* full_bar, uint64_t* empty_bar) {
  const int nst = p.nstages;
  const int* table = p.block_tables + (long long)seq * p.max_num_blocks_per_seq + b0;
  const unsigned char* kbase = static_cast<const unsigned char*>(p.k_cache) + (long long)kvh * p.kv_head_stride * (long long)sizeof(CT);
  const unsigned char* vbase = static_cast<const unsigned char*>(p.v_cache) + (long long)kvh * p.kv_head_stride * (long long)sizeof(CT);
  const long long blk_bytes = p.kv_block_stride * (long long)sizeof(CT);
#ifdef B200_ATTN_NO_HINT
  const uint64_t pol = 0;
#else
  const uint64_t pol = l2_policy_evict_first();
#endif
  int phys_next = lane < nblk ? table[lane] : 0;
  int stage = 0, fill = 0;  // fill = how many times the ring has wrapped
  for (int base = 0; base < nblk; base += 32) {
    const int phys = phys_next;
    const int nxt = base + 32 + lane;
    phys_next = nxt < nblk ? table[nxt] : 0;  // in flight while this group is issued
    const int cnt = min(32, nblk - base);
    for (int j = 0; j < cnt; ++j) {
      const long long pb = __shfl_sync(0xffffffffu, phys, j);
      if (lane == 0) {
        if (fill > 0) mbar_wait(&empty_bar[stage], (fill - 1) & 1);
        unsigned char* dst = stages + (size_t)stage * stage_bytes;
        mbar_arrive_expect_tx(&full_bar[stage], 2 * kbytes);
        bulk_g2s(dst, kbase + pb * blk_bytes, kbytes, &full_bar[stage], pol);
        bulk_g2s(dst + kbytes, vbase + pb * blk_bytes, kbytes, &full_bar[stage], pol);
      }
      if (++stage == nst) { stage = 0; ++fill; }
    }
  }
}

template <typename T, typename CT, int D, int BS, int GT>
__global__ void __launch_bounds__(kAttnThreads) paged_attn_simt_kernel(const __grid_constant__ AttnParams p) {
  using SM = SimtSmem<CT, D, BS, GT>;
  constexpr int X = 16 / (int)sizeof(CT);
  constexpr int KB = SM::kKBytes;
  constexpr int NW = kConsumerWarps;

  const int seq = blockIdx.y;
  const int part = blockIdx.z;
  const int kvh = blockIdx.x / p.gtiles;
  const int gt = blockIdx.x - kvh * p.gtiles;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int head0 = kvh * p.group + gt * GT;           // first query head of this CTA
  const int ng = min(GT, p.group - gt * GT);            // valid query heads in this tile

  const int ctx = p.context_lens[seq];
  const int nblk_ctx = (ctx + BS - 1) / BS;
  int b0 = 0, b1 = nblk_ctx;
  if (p.partition_blocks > 0) {
    b0 = part * p.partition_blocks;
    b1 = min(nblk_ctx, b0 + p.partition_blocks);
  }
  T* out = static_cast<T*>(p.out);
  if (b0 >= b1) {
    // empty context: softmax over nothing -> zeros (reference: exp_sum = 0, acc = 0)
    if (!p.write_partials && part == 0)
      for (int i = tid; i < ng * D; i += kAttnThreads)
        out[((long long)seq * p.num_heads + head0) * D + i] = from_f32<T>(0.f);
    return;
  }
  const int nblk = b1 - b0;
  const int nst = p.nstages;

  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* stages = smem;
  float* q_s = reinterpret_cast<float*>(smem + (size_t)nst * SM::kStageBytes);
  float* plog = q_s + GT * D;
  float* red_acc = plog + NW * GT * BS;
  float* red_ml = red_acc + NW * GT * D;
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(red_ml + NW * GT * 2);
  uint64_t* empty_bar = full_bar + SM::kMaxStages;

  if (tid == 0) {
    for (int s = 0; s < nst; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_fence_init();
  }
  {
    const T* q = static_cast<const T*>(p.q) + (long long)seq * p.q_stride + (long long)head0 * D;
    for (int i = tid; i < GT * D; i += kAttnThreads) q_s[i] = (i < ng * D) ? to_f32<T>(q[i]) : 0.f;
  }
  __syncthreads();

  if (warp == NW) {
    attn_producer<CT>(p, seq, kvh, b0, nblk, lane, stages, SM::kStageBytes, KB, full_bar, empty_bar);
  } else {
    // ================= consumers: online softmax over ring stages =================
    constexpr int NVK = (D / X) * BS;           // 16-byte vectors in a K chunk
    constexpr int KIT = (NVK + 31) / 32;
    constexpr int VV = (X < BS) ? X : BS;       // tokens per V vector
    constexpr int VBYTES = VV * (int)sizeof(CT);
    constexpr int VPR = BS / VV;                // vectors per V row
    constexpr int NVV = D * VPR;
    constexpr int VIT = (NVV + 31) / 32;
    const int tok = lane % BS;
    const int vpart = lane % VPR;
    float* plog_w = plog + warp * GT * BS;

    float m_run[GT], l_run[GT], alibi[GT];
    float acc[GT][VIT];
#pragma unroll
    for (int g = 0; g < GT; ++g) {
      m_run[g] = -FLT_MAX;
      l_run[g] = 0.f;
      alibi[g] = (p.alibi_slopes != nullptr && g < ng) ? p.alibi_slopes[head0 + g] : 0.f;
#pragma unroll
      for (int i = 0; i < VIT; ++i) acc[g][i] = 0.f;
    }

    // Blocks go to the active consumer warps round-robin.  nst % nwa == 0, so every fill of a given
    // stage is consumed by the SAME warp, in program order: a parity wait can only tell adjacent
    // phases apart, and bulk copies land out of order, so a different warp waiting for fill f+1 of a
    // stage could otherwise slip through while fill f is still in flight.
    const int nwa = p.nconsumers;
    for (int n = warp; n < nblk && warp < nwa; n += nwa) {
      const int stage = n % nst;
      const uint32_t parity = (n / nst) & 1;
      const unsigned char* ks = stages + (size_t)stage * SM::kStageBytes;
      const unsigned char* vs = ks + KB;
      const int blk = b0 + n;
      mbar_wait(&full_bar[stage], parity);

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
        sv += (alibi[g] != 0.f) ? alibi[g] * (float)(t_idx - ctx + 1) : 0.f;
        sv = valid ? sv : -FLT_MAX;
        float bmax = sv;
#pragma unroll
        for (int m = BS / 2; m >= 1; m >>= 1) bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, m));
        const float m_new = fmaxf(m_run[g], bmax);
        corr[g] = __expf(m_run[g] - m_new);
        const float pe = valid ? __expf(sv - m_new) : 0.f;
        l_run[g] = l_run[g] * corr[g] + (lane < BS ? pe : 0.f);
        m_run[g] = m_new;
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
      const bool last = (blk == nblk_ctx - 1);
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
            float a = acc[g][i] * corr[g];
#pragma unroll
            for (int j = 0; j < VV; ++j) a = fmaf(pv[g][j], vf[j], a);
            acc[g][i] = a;
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[stage]);
    }

    // ---- per-warp totals -> shared memory ----
#pragma unroll
    for (int g = 0; g < GT; ++g) {
      float l = l_run[g];
#pragma unroll
      for (int m = 16; m >= 1; m >>= 1) l += __shfl_xor_sync(0xffffffffu, l, m);
      if (lane == 0) {
        red_ml[(warp * GT + g) * 2 + 0] = m_run[g];
        red_ml[(warp * GT + g) * 2 + 1] = l;
      }
#pragma unroll
      for (int i = 0; i < VIT; ++i) {
        float a = acc[g][i];
#pragma unroll
        for (int m = VPR / 2; m >= 1; m >>= 1) a += __shfl_xor_sync(0xffffffffu, a, m);
        const int v = lane + 32 * i;
        if (vpart == 0 && ((NVV % 32 == 0) || v < NVV)) red_acc[(warp * GT + g) * D + v / VPR] = a;
      }
    }
  }
  __syncthreads();

  attn_merge_store<T, D, GT>(p, red_acc, red_ml, ng, seq, head0, part, tid);
}

// Merge of partition partials; same formula as the reference's paged_attention_v2_reduce_kernel
// (attention_kernels.cu:484-575): out = sum_j tmp_j * exp_sums_j * e^(max_j - max) / (sum + 1e-6).
template <typename T>
__global__ void __launch_bounds__(128) paged_attn_reduce_kernel(
    T* __restrict__ out, const float* __restrict__ exp_sums, const float* __restrict__ max_logits,
    const T* __restrict__ tmp_out, const int* __restrict__ context_lens, int head_size,
    int max_num_partitions, int partition_tokens) {
  const int head = blockIdx.x, seq = blockIdx.y, num_heads = gridDim.x;
  const int ctx = context_lens[seq];
  const int np = (ctx + partition_tokens - 1) / partition_tokens;
  const long long sh = (long long)seq * num_heads + head;
  const T* tmp = tmp_out + sh * max_num_partitions * head_size;
  T* o = out + sh * head_size;
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

// One instantiation family per (T, CT); defined in attn_inst_*.cu so they compile in parallel.
template <typename T, typename CT>
int launch_simt(const AttnParams& p, int head_size, int block_size, int gt, dim3 grid,
                int smem_budget, cudaStream_t stream);

template <typename CT>
inline bool simt_smem_plan(int head_size, int block_size, int gt, int smem_budget, int* nstages, int* nconsumers,
                           int* smem_bytes) {
  const int kbytes = head_size * block_size * (int)sizeof(CT);
  const int stage = 2 * kbytes;
  const int fixed = gt * head_size * 4 + kConsumerWarps * gt * block_size * 4 +
                    kConsumerWarps * gt * head_size * 4 + kConsumerWarps * gt * 2 * 4 + 2 * 32 * 8;
  int n = (smem_budget - fixed) / stage;
  if (n > 16) n = 16;
  // ring depth must be a multiple of the number of consuming warps (same warp per stage, see kernel)
  const int nwa = n < kConsumerWarps ? n : kConsumerWarps;
  if (nwa > 0) n = n / nwa * nwa;
  *nstages = n;
  *nconsumers = nwa;
  *smem_bytes = n * stage + fixed;
  return n >= 2;
}

}  // namespace b200
