// Paged-KV attention decode for sm_100a, grouped-query variant on the warp-level tensor-core path.
//
// When several query heads share a KV head (GQA / MQA) the SIMT kernel does G multiply-adds per KV
// element and becomes issue-bound before HBM does (ncu: profiles/).  Here the
// G query heads of one KV head are the M dimension of mma.sync.m16n8k16 tiles, so the per-element work
// moves to the tensor pipe and the kernel goes back to being an HBM streamer.  tcgen05 is not used on
// purpose: its minimum tile is M=64 with a TMEM round trip, and a decode step has 1 query row per
// head -- the legacy warp-level MMA has an order of magnitude more throughput than this path needs.
//
// Same ring / producer / partition / epilogue contract as paged_attn_simt_kernel; only the consumer
// warps differ.  The paged layouts are already fragment-shaped:
//   K [D/x][bs][x]: row (chunk j, token t) is 16 contiguous bytes -> one ldmatrix row; 8 tokens of a
//                   chunk form one 8x8 matrix = the B fragment of S^T = Q.K^T   (n = token, k = d)
//   V [D][bs]:      row d holds the block's tokens contiguously -> ldmatrix rows = d give the B
//                   fragment of O = P.V                                          (n = d, k = token)
// and the accumulator layout of S^T is the A-fragment layout of P, so probabilities never leave
// registers.  fp8 (e5m2) caches are widened to f16 in registers with a byte permute (exact); a fixed
// permutation of the k index (d for Q.K^T, token for P.V) absorbs the 4-elements-per-word layout that
// ldmatrix gives for 1-byte data.
#pragma once

#include "paged_attention.cuh"

namespace b200 {

// ---- fragment helpers ---------------------------------------------------------------------
template <typename MT>
__device__ __forceinline__ void mma16816(float* d, const uint32_t* a, uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma16816<__half>(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
template <>
__device__ __forceinline__ void mma16816<__nv_bfloat16>(float* d, const uint32_t* a, uint32_t b0, uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
      : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}

__device__ __forceinline__ void ldsm_x4(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}
__device__ __forceinline__ void ldsm_x2(uint32_t addr, uint32_t& r0, uint32_t& r1) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x2.shared.b16 {%0,%1}, [%2];" : "=r"(r0), "=r"(r1) : "r"(addr));
}

template <typename MT>
__device__ __forceinline__ uint32_t pack2(float lo, float hi);
template <>
__device__ __forceinline__ uint32_t pack2<__half>(float lo, float hi) {
  __half2 h = __floats2half2_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}
template <>
__device__ __forceinline__ uint32_t pack2<__nv_bfloat16>(float lo, float hi) {
  __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<uint32_t*>(&h);
}

// two e5m2 bytes -> packed f16x2 (e5m2 is the high byte of an IEEE half: exact)
__device__ __forceinline__ uint32_t e5m2x2_lo_to_h2(uint32_t w) { return __byte_perm(w, 0u, 0x1404); }
__device__ __forceinline__ uint32_t e5m2x2_hi_to_h2(uint32_t w) { return __byte_perm(w, 0u, 0x3424); }

// MMA element type: the cache type, or f16 for fp8 caches (the query is converted to match)
template <typename T, typename CT> struct MmaElt { using type = CT; };
template <typename T> struct MmaElt<T, uint8_t> { using type = __half; };

constexpr int kMmaHeads = 16;  // query heads per CTA tile (the M of m16n8k16)

template <typename CT, int D, int BS>
struct MmaSmem {
  static constexpr int kKBytes = D * BS * (int)sizeof(CT);
  static constexpr int kStageBytes = 2 * kKBytes;
  static constexpr int kMaxStages = 32;
  static constexpr int kMergeBytes = kConsumerWarps * kMmaHeads * D * 4;  // aliased onto the ring after the main loop
  static constexpr int kFixedBytes = kConsumerWarps * kMmaHeads * 2 * 4 + 2 * kMaxStages * 8;
};

// mma with only rows 0..7 of the M=16 tile live (<= 8 query heads per KV head): a1 = a3 = 0, the
// upper accumulator half is dead, so its registers are never allocated.
template <typename MT>
__device__ __forceinline__ void mma16816_lo(float& d0, float& d1, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1);
template <>
__device__ __forceinline__ void mma16816_lo<__half>(float& d0, float& d1, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  float x2, x3;  // dead upper half of the accumulator tile
  (void)x2; (void)x3;
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%5}, {%7,%8}, {%0,%1,%9,%9};"
      : "+f"(d0), "+f"(d1), "=f"(x2), "=f"(x3)
      : "r"(a0), "r"(0u), "r"(a2), "r"(b0), "r"(b1), "f"(0.f));
}
template <>
__device__ __forceinline__ void mma16816_lo<__nv_bfloat16>(float& d0, float& d1, uint32_t a0, uint32_t a2, uint32_t b0, uint32_t b1) {
  float x2, x3;  // dead upper half of the accumulator tile
  (void)x2; (void)x3;
  asm volatile(
      "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%5}, {%7,%8}, {%0,%1,%9,%9};"
      : "+f"(d0), "+f"(d1), "=f"(x2), "=f"(x3)
      : "r"(a0), "r"(0u), "r"(a2), "r"(b0), "r"(b1), "f"(0.f));
}

// HI = rows 8..15 of the tile carry query heads (more than 8 query heads per KV head).
template <typename T, typename CT, int D, int BS, bool HI>
__global__ void __launch_bounds__(kAttnThreads) paged_attn_mma_kernel(const __grid_constant__ AttnParams p) {
  using MT = typename MmaElt<T, CT>::type;
  using SM = MmaSmem<CT, D, BS>;
  constexpr bool FP8 = sizeof(CT) == 1;
  constexpr int KB = SM::kKBytes;
  constexpr int NW = kConsumerWarps;
  constexpr int GT = HI ? 16 : 8;   // query heads per CTA
  constexpr int R = HI ? 2 : 1;     // row sets per thread (head g, head g + 8)
  constexpr int KS = D / 16;        // k-steps of Q.K^T
  constexpr int NT = D / 8;         // n-tiles of P.V
  constexpr int SB = BS / 16;       // 16-token sub-blocks per KV block
  static_assert(D % 16 == 0 && BS % 16 == 0, "tile shapes");

  const int seq = blockIdx.y;
  const int part = blockIdx.z;
  const int kvh = blockIdx.x / p.gtiles;
  const int gt = blockIdx.x - kvh * p.gtiles;
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int head0 = kvh * p.group + gt * GT;
  const int ng = min(GT, p.group - gt * GT);

  const int ctx = p.context_lens[seq];
  const int nblk_ctx = (ctx + BS - 1) / BS;
  int b0 = 0, b1 = nblk_ctx;
  if (p.partition_blocks > 0) {
    b0 = part * p.partition_blocks;
    b1 = min(nblk_ctx, b0 + p.partition_blocks);
  }
  if (b0 >= b1) {
    if (!p.write_partials && part == 0) {
      T* out = static_cast<T*>(p.out);
      for (int i = tid; i < ng * D; i += kAttnThreads)
        out[((long long)seq * p.num_heads + head0) * D + i] = from_f32<T>(0.f);
    }
    return;
  }
  const int nblk = b1 - b0;
  const int nst = p.nstages;

  extern __shared__ __align__(128) unsigned char smem[];
  unsigned char* stages = smem;
  float* red_ml = reinterpret_cast<float*>(smem + (size_t)nst * SM::kStageBytes);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(red_ml + NW * kMmaHeads * 2);
  uint64_t* empty_bar = full_bar + SM::kMaxStages;
  float* red_acc = reinterpret_cast<float*>(stages);  // aliased onto the ring once it has drained

  if (tid == 0) {
    for (int s = 0; s < nst; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_fence_init();
  }
  __syncthreads();

  const int g = lane >> 2, t = lane & 3;
  float m_run[R], l_run[R];
  float o[NT][2 * R];
#pragma unroll
  for (int r = 0; r < R; ++r) { m_run[r] = -FLT_MAX; l_run[r] = 0.f; }
#pragma unroll
  for (int i = 0; i < NT; ++i)
#pragma unroll
    for (int j = 0; j < 2 * R; ++j) o[i][j] = 0.f;

  if (warp == NW) {
    attn_producer<CT>(p, seq, kvh, b0, nblk, lane, stages, SM::kStageBytes, KB, full_bar, empty_bar);
  } else if (warp < p.nconsumers) {
    // ---- Q fragments (A operand of S^T = Q.K^T), in registers for the whole kernel ----
    // k-slot -> d mapping inside a 16-wide k-step: natural for 2-byte caches; for fp8 the thread's
    // ldmatrix word holds d = 4t..4t+3, so slots (2t,2t+1) <- d (4t,4t+1), (2t+8,2t+9) <- d (4t+2,4t+3).
    uint32_t qa[KS][2 * R];
    {
      const T* qrow = static_cast<const T*>(p.q) + (long long)seq * p.q_stride + (long long)head0 * D;
#pragma unroll
      for (int kk = 0; kk < KS; ++kk) {
        const int d_lo = FP8 ? 16 * kk + 4 * t : 16 * kk + 2 * t;
        const int d_hi = FP8 ? d_lo + 2 : d_lo + 8;
#pragma unroll
        for (int r = 0; r < R; ++r) {
          const int h = g + 8 * r;
          float x0 = 0.f, x1 = 0.f, y0 = 0.f, y1 = 0.f;
          if (h < ng) {
            const T* qh = qrow + h * D;
            x0 = to_f32<T>(qh[d_lo]); x1 = to_f32<T>(qh[d_lo + 1]);
            y0 = to_f32<T>(qh[d_hi]); y1 = to_f32<T>(qh[d_hi + 1]);
          }
          qa[kk][r] = pack2<MT>(x0, x1);       // a0 (r=0) / a1 (r=1)
          qa[kk][R + r] = pack2<MT>(y0, y1);   // a2 / a3
        }
      }
    }
    float alibi[R];
#pragma unroll
    for (int r = 0; r < R; ++r)
      alibi[r] = (p.alibi_slopes != nullptr && g + 8 * r < ng) ? p.alibi_slopes[head0 + g + 8 * r] : 0.f;

    // per-lane ldmatrix row offsets (bytes) inside a K chunk / V chunk, excluding the k-step term
    const int mi = lane >> 3, rr = lane & 7;
    int k_lane_off, v_lane_off;
    if constexpr (FP8) {
      // K x2: matrix (mi & 1) = token tile; row rr -> token (rr>>1)*4 + (rr&1) + 2*tile
      k_lane_off = ((rr >> 1) * 4 + (rr & 1) + 2 * (mi & 1)) * 16;
      // V x2: matrix (mi & 1) = d tile; row rr -> d = 8*(mi&1) + rr, 16 tokens = 16 bytes
      v_lane_off = ((mi & 1) * 8 + rr) * BS;
    } else {
      // K x4: matrices [(chunk 2kk, tok 0-7), (2kk+1, 0-7), (2kk, 8-15), (2kk+1, 8-15)]
      k_lane_off = ((mi & 1) * BS + (mi >> 1) * 8 + rr) * 16;
      // V x4: matrices [(d tile, tok 0-7), (d tile, tok 8-15), (d tile+1, 0-7), (d tile+1, 8-15)]
      v_lane_off = (((mi >> 1) * 8 + rr) * BS + (mi & 1) * 8) * 2;
    }

    const int nwa = p.nconsumers;
    for (int n = warp; n < nblk; n += nwa) {
      const int stage = n % nst;
      const uint32_t parity = (n / nst) & 1;
      unsigned char* ks = stages + (size_t)stage * SM::kStageBytes;
      unsigned char* vs = ks + KB;
      const int blk = b0 + n;
      mbar_wait(&full_bar[stage], parity);

      if (blk == nblk_ctx - 1 && ctx < nblk_ctx * BS) {
        // stale slots past the context may hold NaN/Inf bit patterns (reference :342-351): 0 * NaN would
        // poison the tensor-core product, so clear those V columns in shared memory first.
        const int rem = ctx - blk * BS;
        CT* vt = reinterpret_cast<CT*>(vs);
        for (int i = lane; i < D * BS; i += 32)
          if ((i % BS) >= rem) vt[i] = CT(0);
        __syncwarp();
      }
      const uint32_t ks_u = smem_u32(ks), vs_u = smem_u32(vs);

#pragma unroll
      for (int sb = 0; sb < SB; ++sb) {
        // ---- S^T[head][token] = Q.K^T : two independent accumulator chains per token tile ----
        float s[2][2][2 * R];
#pragma unroll
        for (int c = 0; c < 2; ++c)
#pragma unroll
          for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int j = 0; j < 2 * R; ++j) s[c][nt][j] = 0.f;
#pragma unroll
        for (int kk = 0; kk < KS; ++kk) {
          uint32_t b[2][2];  // [token tile][b0, b1]
          if constexpr (FP8) {
            uint32_t w0, w1;
            ldsm_x2(ks_u + (kk * BS + sb * 16) * 16 + k_lane_off, w0, w1);
            b[0][0] = e5m2x2_lo_to_h2(w0); b[0][1] = e5m2x2_hi_to_h2(w0);
            b[1][0] = e5m2x2_lo_to_h2(w1); b[1][1] = e5m2x2_hi_to_h2(w1);
          } else {
            ldsm_x4(ks_u + (2 * kk * BS + sb * 16) * 16 + k_lane_off, b[0][0], b[0][1], b[1][0], b[1][1]);
          }
#pragma unroll
          for (int nt = 0; nt < 2; ++nt) {
            if constexpr (HI) mma16816<MT>(s[kk & 1][nt], qa[kk], b[nt][0], b[nt][1]);
            else mma16816_lo<MT>(s[kk & 1][nt][0], s[kk & 1][nt][1], qa[kk][0], qa[kk][1], b[nt][0], b[nt][1]);
          }
        }
        // ---- online softmax; thread (g, t) holds heads g (+8) x tokens {tile nt, slots 2t, 2t+1} ----
        uint32_t pa[4] = {0u, 0u, 0u, 0u};
#pragma unroll
        for (int r = 0; r < R; ++r) {
          float sv[2][2];
          bool valid[2][2];
          float bmax = -FLT_MAX;
#pragma unroll
          for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              const int tok_in = FP8 ? (4 * t + 2 * nt + e) : (8 * nt + 2 * t + e);
              const int t_idx = blk * BS + sb * 16 + tok_in;
              valid[nt][e] = t_idx < ctx;
              float x = (s[0][nt][2 * r + e] + s[1][nt][2 * r + e]) * p.scale;
              x += (alibi[r] != 0.f) ? alibi[r] * (float)(t_idx - ctx + 1) : 0.f;
              x = valid[nt][e] ? x : -FLT_MAX;
              sv[nt][e] = x;
              bmax = fmaxf(bmax, x);
            }
          bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, 1));
          bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, 2));
          const float m_new = fmaxf(m_run[r], bmax);
          const float corr = __expf(m_run[r] - m_new);
          float pe[2][2], psum = 0.f;
#pragma unroll
          for (int nt = 0; nt < 2; ++nt)
#pragma unroll
            for (int e = 0; e < 2; ++e) {
              pe[nt][e] = valid[nt][e] ? __expf(sv[nt][e] - m_new) : 0.f;
              psum += pe[nt][e];
            }
          l_run[r] = l_run[r] * corr + psum;  // per-thread partial; the quad is summed once at the end
          m_run[r] = m_new;
          pa[r] = pack2<MT>(pe[0][0], pe[0][1]);      // a0 / a1: k slots 2t, 2t+1
          pa[2 + r] = pack2<MT>(pe[1][0], pe[1][1]);  // a2 / a3: k slots 2t+8, 2t+9
#pragma unroll
          for (int i = 0; i < NT; ++i) {
            o[i][2 * r] *= corr;
            o[i][2 * r + 1] *= corr;
          }
        }
        // ---- O[head][d] += P.V ----
#pragma unroll
        for (int dn = 0; dn < NT; dn += 2) {
          uint32_t b[2][2];  // [d tile][b0, b1]
          if constexpr (FP8) {
            uint32_t w0, w1;
            ldsm_x2(vs_u + dn * 8 * BS + sb * 16 + v_lane_off, w0, w1);
            b[0][0] = e5m2x2_lo_to_h2(w0); b[0][1] = e5m2x2_hi_to_h2(w0);
            b[1][0] = e5m2x2_lo_to_h2(w1); b[1][1] = e5m2x2_hi_to_h2(w1);
          } else {
            ldsm_x4(vs_u + (dn * 8 * BS + sb * 16) * 2 + v_lane_off, b[0][0], b[0][1], b[1][0], b[1][1]);
          }
#pragma unroll
          for (int i = 0; i < 2; ++i) {
            if constexpr (HI) mma16816<MT>(o[dn + i], pa, b[i][0], b[i][1]);
            else mma16816_lo<MT>(o[dn + i][0], o[dn + i][1], pa[0], pa[2], b[i][0], b[i][1]);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty_bar[stage]);
    }
  }
  // every issued stage has been consumed: the ring is idle and can hold the merge buffers
  __syncthreads();
  if (warp < NW) {
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float l = l_run[r];
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      const int h = g + 8 * r;
      if (h < ng) {
        if (t == 0) {
          red_ml[(warp * GT + h) * 2 + 0] = m_run[r];
          red_ml[(warp * GT + h) * 2 + 1] = l;
        }
#pragma unroll
        for (int i = 0; i < NT; ++i)
          *reinterpret_cast<float2*>(&red_acc[(warp * GT + h) * D + i * 8 + 2 * t]) = make_float2(o[i][2 * r], o[i][2 * r + 1]);
      }
    }
  }
  __syncthreads();
  attn_merge_store<T, D, GT>(p, red_acc, red_ml, ng, seq, head0, part, tid);
}

// One instantiation family per (T, CT); defined in attn_mma_inst_*.cu.  hi = more than 8 heads per tile.
template <typename T, typename CT>
int launch_mma(const AttnParams& p, int head_size, int block_size, bool hi, dim3 grid, int smem_budget,
               cudaStream_t stream);

template <typename CT>
inline bool mma_smem_plan(int head_size, int block_size, int smem_budget, int* nstages, int* nconsumers, int* smem_bytes) {
  const int stage = 2 * head_size * block_size * (int)sizeof(CT);
  const int fixed = kConsumerWarps * kMmaHeads * 2 * 4 + 2 * 32 * 8;
  const int merge = kConsumerWarps * kMmaHeads * head_size * 4;
  int n = (smem_budget - fixed) / stage;
  if (n > 16) n = 16;
  const int nwa = n < kConsumerWarps ? n : kConsumerWarps;
  if (nwa > 0) n = n / nwa * nwa;
  *nstages = n;
  *nconsumers = nwa;
  *smem_bytes = n * stage + fixed;
  return n >= 2 && n * stage >= merge;
}

}  // namespace b200
