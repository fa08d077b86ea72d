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

#include <type_traits>

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


__device__ __forceinline__ float ex2_approx(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}

// Tensor-core consumer.  GT_ = 8: rows 0..7 of the M=16 tile carry the query heads of one KV head (the upper
// half of every fragment is dead and costs no registers); GT_ = 16: all rows live (more than 8 heads per KV head).
template <typename T_, typename CT_, int D_, int BS_, int GT_>
struct MmaImpl {
  using T = T_;
  using CT = CT_;
  using MT = typename MmaElt<T, CT>::type;
  static constexpr int D = D_, BS = BS_, GT = GT_;
  static constexpr bool HI = GT_ == 16;
  static constexpr bool FP8 = sizeof(CT) == 1;
  // bf16 queries over an fp8 cache run the f16 tensor-core path (e5m2 widens to f16 exactly, by a byte permute).
  // bf16 has 8 exponent bits, f16 five: each query row is first scaled by a power of two that brings its
  // largest magnitude into [1, 2) (exact), and the logit scale carries the inverse, so no query overflows
  // or flushes where the reference (fp8 -> bf16 widening, fp32 dot; dtype_fp8_e5m2.cuh, attention_utils.cuh:31-37)
  // would not.  Elements more than 2^24 below their row's maximum round to zero: below fp32 dot-product noise.
  static constexpr bool NORM = FP8 && sizeof(T) == 2 && !std::is_same<T, __half>::value;
  static constexpr int R = HI ? 2 : 1;     // row sets per thread (head g, head g + 8)
  static constexpr int KS = D / 16;        // k-steps of Q.K^T
  static constexpr int NT = D / 8;         // n-tiles of P.V
  static constexpr int SB = BS / 16;       // 16-token sub-blocks per KV block
  static constexpr int kKBytes = D * BS * (int)sizeof(CT);
  static constexpr int kStageBytes = 2 * kKBytes;
  static constexpr int kFixedBytes = 0;    // queries and probabilities live in registers
  static_assert(GT_ == 8 || GT_ == 16, "tile rows");
  static_assert(D % 16 == 0 && BS % 16 == 0, "tile shapes");
  static constexpr float kLog2e = 1.4426950408889634f;
  static constexpr float kLn2 = 0.6931471805599453f;

  struct State {
    float m_run[R], l_run[R];   // running max (log2 units) and per-thread partial of the running sum
    float o[NT][2 * R];         // O[head g (+8)][d = 8*tile + 2t, +1]
    uint32_t qa[KS][2 * R];     // A fragments of Q: [k-step][a0, (a1), a2, (a3)]
    float alibi2[R];            // alibi slope * log2(e)
    float scale2[R];            // softmax scale * log2(e) (* the row's power-of-two query normaliser, NORM)
    int k_lane_off, v_lane_off; // per-lane ldmatrix row offsets (bytes) inside a K / V chunk
  };

  __device__ __forceinline__ static void init(State& st) {
#pragma unroll
    for (int r = 0; r < R; ++r) { st.m_run[r] = -FLT_MAX; st.l_run[r] = 0.f; }
#pragma unroll
    for (int i = 0; i < NT; ++i)
#pragma unroll
      for (int j = 0; j < 2 * R; ++j) st.o[i][j] = 0.f;
  }

  __device__ __forceinline__ static void begin_item(State& st, const AttnParams& p, const ItemInfo& it, const T* qrow,
                                                    unsigned char*, int, int lane) {
    const int g = lane >> 2, t = lane & 3;
    // Q fragments (A operand of S^T = Q.K^T).  k-slot -> d mapping inside a 16-wide k-step: natural for
    // 2-byte caches; for fp8 the thread's ldmatrix word holds d = 4t..4t+3, so slots (2t,2t+1) <- d (4t,4t+1)
    // and (2t+8,2t+9) <- d (4t+2,4t+3).
    float qmul[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      qmul[r] = 1.f;
      st.scale2[r] = p.scale * kLog2e;
    }
    // The four query elements a thread needs per k-step: d_lo, d_lo+1 (a0/a1) and d_hi, d_hi+1 (a2/a3).  fp8 caches:
    // d_lo = 16kk + 4t, d_hi = d_lo + 2 -- one contiguous 8-byte word; 2-byte caches: d_lo = 16kk + 2t, d_hi = d_lo + 8
    // -- two 4-byte words.  Rows staged in shared memory (q_bulk) are 16-byte aligned, so the words are loaded as such;
    // rows read straight from global memory may not be, and go element by element.
    float qv[R][KS][4];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int h = g + 8 * r;
      const T* qh = qrow + h * D;
#pragma unroll
      for (int kk = 0; kk < KS; ++kk) {
        const int d_lo = FP8 ? 16 * kk + 4 * t : 16 * kk + 2 * t;
        const int d_hi = FP8 ? d_lo + 2 : d_lo + 8;
        if (h >= it.ng) {
          qv[r][kk][0] = qv[r][kk][1] = qv[r][kk][2] = qv[r][kk][3] = 0.f;
        } else if (p.q_bulk) {
          alignas(8) T w[4];
          if constexpr (FP8) {
            *reinterpret_cast<uint2*>(w) = *reinterpret_cast<const uint2*>(qh + d_lo);
          } else {
            *reinterpret_cast<uint32_t*>(w) = *reinterpret_cast<const uint32_t*>(qh + d_lo);
            *reinterpret_cast<uint32_t*>(w + 2) = *reinterpret_cast<const uint32_t*>(qh + d_hi);
          }
#pragma unroll
          for (int e = 0; e < 4; ++e) qv[r][kk][e] = to_f32<T>(w[e]);
        } else {
          qv[r][kk][0] = to_f32<T>(qh[d_lo]); qv[r][kk][1] = to_f32<T>(qh[d_lo + 1]);
          qv[r][kk][2] = to_f32<T>(qh[d_hi]); qv[r][kk][3] = to_f32<T>(qh[d_hi + 1]);
        }
      }
    }
    if constexpr (NORM) {
      // the quad (t = 0..3) of a row holds all D elements of head g (+8) between them
#pragma unroll
      for (int r = 0; r < R; ++r) {
        float mx = 0.f;
#pragma unroll
        for (int kk = 0; kk < KS; ++kk)
#pragma unroll
          for (int e = 0; e < 4; ++e) mx = fmaxf(mx, fabsf(qv[r][kk][e]));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
        mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
        int ex = (int)((__float_as_uint(mx) >> 23) & 0xffu) - 127;  // floor(log2(mx)) for normal mx
        if (!(mx > 0.f) || ex > 127) ex = 0;                        // all-zero row, Inf/NaN: leave as is
        ex = max(-100, min(100, ex));
        qmul[r] = __uint_as_float((uint32_t)(127 - ex) << 23);      // 2^-ex
        st.scale2[r] *= __uint_as_float((uint32_t)(127 + ex) << 23);
      }
    }
#pragma unroll
    for (int kk = 0; kk < KS; ++kk) {
#pragma unroll
      for (int r = 0; r < R; ++r) {
        st.qa[kk][r] = pack2<MT>(qv[r][kk][0] * qmul[r], qv[r][kk][1] * qmul[r]);       // a0 (r=0) / a1 (r=1)
        st.qa[kk][R + r] = pack2<MT>(qv[r][kk][2] * qmul[r], qv[r][kk][3] * qmul[r]);   // a2 / a3
      }
    }
#pragma unroll
    for (int r = 0; r < R; ++r)
      st.alibi2[r] = (p.alibi_slopes != nullptr && g + 8 * r < it.ng) ? p.alibi_slopes[it.head0 + g + 8 * r] * kLog2e : 0.f;
    const int mi = lane >> 3, rr = lane & 7;
    if constexpr (FP8) {
      // K x2: matrix (mi & 1) = token tile; row rr -> token (rr>>1)*4 + (rr&1) + 2*tile
      st.k_lane_off = ((rr >> 1) * 4 + (rr & 1) + 2 * (mi & 1)) * 16;
      // V x2: matrix (mi & 1) = d tile; row rr -> d = 8*(mi&1) + rr, one byte per token
      st.v_lane_off = ((mi & 1) * 8 + rr) * BS;
    } else {
      // K x4: matrices [(chunk 2kk, tok 0-7), (2kk+1, 0-7), (2kk, 8-15), (2kk+1, 8-15)]
      st.k_lane_off = ((mi & 1) * BS + (mi >> 1) * 8 + rr) * 16;
      // V x4: matrices [(d tile, tok 0-7), (d tile, tok 8-15), (d tile+1, 0-7), (d tile+1, 8-15)]
      st.v_lane_off = (((mi >> 1) * 8 + rr) * BS + (mi & 1) * 8) * 2;
    }
  }

  __device__ __forceinline__ static void block(State& st, const AttnParams& p, const ItemInfo& it,
                                               unsigned char* ks, unsigned char* vs, int blk, unsigned char*,
                                               int, int lane) {
    const int t = lane & 3;
    const int ctx = it.ctx;
    const bool last = blk == it.nblk_ctx - 1;
    if (last && ctx < it.nblk_ctx * BS) {
      // stale slots past the context may hold NaN/Inf bit patterns (reference :342-351): 0 * NaN would
      // poison the tensor-core product, so clear those V columns in shared memory first.
      const int rem = ctx - blk * BS;
      CT* vt = reinterpret_cast<CT*>(vs);
      for (int i = lane; i < D * BS; i += 32)
        if ((i % BS) >= rem) vt[i] = CT(0);
      // These are generic-proxy writes into a stage the producer will refill through the async proxy (TMA):
      // order them before that refill (the empty-barrier arrive that follows releases the stage).
      asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
      __syncwarp();
    }
    const uint32_t ks_u = smem_u32(ks), vs_u = smem_u32(vs);
    const bool has_alibi = p.alibi_slopes != nullptr;

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
          ldsm_x2(ks_u + (kk * BS + sb * 16) * 16 + st.k_lane_off, w0, w1);
          b[0][0] = e5m2x2_lo_to_h2(w0); b[0][1] = e5m2x2_hi_to_h2(w0);
          b[1][0] = e5m2x2_lo_to_h2(w1); b[1][1] = e5m2x2_hi_to_h2(w1);
        } else {
          ldsm_x4(ks_u + (2 * kk * BS + sb * 16) * 16 + st.k_lane_off, b[0][0], b[0][1], b[1][0], b[1][1]);
        }
#pragma unroll
        for (int nt = 0; nt < 2; ++nt) {
          if constexpr (HI) mma16816<MT>(s[kk & 1][nt], st.qa[kk], b[nt][0], b[nt][1]);
          else mma16816_lo<MT>(s[kk & 1][nt][0], s[kk & 1][nt][1], st.qa[kk][0], st.qa[kk][1], b[nt][0], b[nt][1]);
        }
      }
      // ---- online softmax in log2 units; thread (g, t) holds heads g (+8) x tokens {tile nt, slots 2t, 2t+1} ----
      uint32_t pa[4] = {0u, 0u, 0u, 0u};
      float corr[R];
#pragma unroll
      for (int r = 0; r < R; ++r) {
        float sv[2][2];
        float bmax = -FLT_MAX;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            const int tok_in = FP8 ? (4 * t + 2 * nt + e) : (8 * nt + 2 * t + e);
            const int t_idx = blk * BS + sb * 16 + tok_in;
            float x = (s[0][nt][2 * r + e] + s[1][nt][2 * r + e]) * st.scale2[r];
            if (has_alibi) x = fmaf(st.alibi2[r], (float)(t_idx - ctx + 1), x);
            if (last) x = t_idx < ctx ? x : -FLT_MAX;
            sv[nt][e] = x;
            bmax = fmaxf(bmax, x);
          }
        bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, 1));
        bmax = fmaxf(bmax, __shfl_xor_sync(0xffffffffu, bmax, 2));
        const float m_new = fmaxf(st.m_run[r], bmax);
        corr[r] = ex2_approx(st.m_run[r] - m_new);
        float pe[2][2], psum = 0.f;
#pragma unroll
        for (int nt = 0; nt < 2; ++nt)
#pragma unroll
          for (int e = 0; e < 2; ++e) {
            // masked tokens: -FLT_MAX - m_new -> ex2 = 0 (m_new is finite: the block has a valid token or
            // m_run already is), so no select is needed
            pe[nt][e] = ex2_approx(sv[nt][e] - m_new);
            if (last) pe[nt][e] = sv[nt][e] == -FLT_MAX ? 0.f : pe[nt][e];
            psum += pe[nt][e];
          }
        st.l_run[r] = st.l_run[r] * corr[r] + psum;  // per-thread partial; the quad is summed once per item
        st.m_run[r] = m_new;
        pa[r] = pack2<MT>(pe[0][0], pe[0][1]);      // a0 / a1: k slots 2t, 2t+1
        pa[2 + r] = pack2<MT>(pe[1][0], pe[1][1]);  // a2 / a3: k slots 2t+8, 2t+9
      }
      // rescale O only when some row's running max moved (it rarely does after the first blocks)
      bool moved = corr[0] != 1.f;
      if constexpr (HI) moved = moved || corr[1] != 1.f;
      if (__any_sync(0xffffffffu, moved)) {
#pragma unroll
        for (int r = 0; r < R; ++r)
#pragma unroll
          for (int i = 0; i < NT; ++i) {
            st.o[i][2 * r] *= corr[r];
            st.o[i][2 * r + 1] *= corr[r];
          }
      }
      // ---- O[head][d] += P.V ----
#pragma unroll
      for (int dn = 0; dn < NT; dn += 2) {
        uint32_t b[2][2];  // [d tile][b0, b1]
        if constexpr (FP8) {
          uint32_t w0, w1;
          ldsm_x2(vs_u + dn * 8 * BS + sb * 16 + st.v_lane_off, w0, w1);
          b[0][0] = e5m2x2_lo_to_h2(w0); b[0][1] = e5m2x2_hi_to_h2(w0);
          b[1][0] = e5m2x2_lo_to_h2(w1); b[1][1] = e5m2x2_hi_to_h2(w1);
        } else {
          ldsm_x4(vs_u + (dn * 8 * BS + sb * 16) * 2 + st.v_lane_off, b[0][0], b[0][1], b[1][0], b[1][1]);
        }
#pragma unroll
        for (int i = 0; i < 2; ++i) {
          if constexpr (HI) mma16816<MT>(st.o[dn + i], pa, b[i][0], b[i][1]);
          else mma16816_lo<MT>(st.o[dn + i][0], st.o[dn + i][1], pa[0], pa[2], b[i][0], b[i][1]);
        }
      }
    }
  }

  // One warp processed a whole partition by itself (quad mode): normalise its state and write the partition's
  // partial straight from registers in the reference's tmp_out / exp_sums / max_logits format -- the same formula as
  // attn_merge_store with a single contributor.
  static constexpr bool kQuad = true;
  __device__ __forceinline__ static void store_partial(State& st, const AttnParams& p, const ItemInfo& it, int lane) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float l = st.l_run[r];
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      const int h = g + 8 * r;
      if (h < it.ng) {
        const float inv = __fdividef(1.f, l + 1e-6f);
        const long long pi = ((long long)it.seq * p.num_heads + it.head0 + h) * p.max_num_partitions + it.part;
        T* row = static_cast<T*>(p.tmp_out) + pi * D;
#pragma unroll
        for (int i = 0; i < NT; ++i)
          *reinterpret_cast<uint32_t*>(row + i * 8 + 2 * t) = pack2<T>(st.o[i][2 * r] * inv, st.o[i][2 * r + 1] * inv);
        if (t == 0) {
          p.exp_sums[pi] = l;
          p.max_logits[pi] = st.m_run[r] * kLn2;
        }
      }
    }
  }

  __device__ __forceinline__ static void end_item(State& st, const ItemInfo& it, float* red_acc, float* red_ml,
                                                  int warp, int lane) {
    const int g = lane >> 2, t = lane & 3;
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float l = st.l_run[r];
      l += __shfl_xor_sync(0xffffffffu, l, 1);
      l += __shfl_xor_sync(0xffffffffu, l, 2);
      const int h = g + 8 * r;
      if (h < it.ng) {
        if (t == 0) {
          red_ml[(warp * GT + h) * 2 + 0] = st.m_run[r] * kLn2;  // back to natural-log units for the merge
          red_ml[(warp * GT + h) * 2 + 1] = l;
        }
#pragma unroll
        for (int i = 0; i < NT; ++i)
          *reinterpret_cast<float2*>(&red_acc[(warp * GT + h) * D + i * 8 + 2 * t]) =
              make_float2(st.o[i][2 * r], st.o[i][2 * r + 1]);
      }
    }
  }
};

}  // namespace b200
