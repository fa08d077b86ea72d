// Prefix ("context") prefill attention over the paged KV cache for sm_100a  (SURVEY 8f row N1).
//
// Replaces the reference's Triton kernels _fwd_kernel / _fwd_kernel_alibi
// (vllm/model_executor/layers/attention/ops/prefix_prefill.py:10-199, 392-618): for every new token m of a
// sequence whose first ctx_len tokens already sit in the paged cache,
//     out[m] = softmax_n( scale * q[m].k[n] + slope * (pos_n - pos_m) ) . v[n]
// over n in { cached tokens 0..ctx_len-1 } U { new tokens 0..m } (causal inside the new tokens).
//
// One CTA = 128 query rows of one (sequence, query head): 8 warps x 16 rows, mma.sync.m16n8k16 with fp32
// accumulators, online softmax in the log2 domain, probabilities never leave registers (the accumulator
// layout of S is the A-fragment layout of P).  KV tiles of 64 tokens travel global -> shared with 16-byte
// cp.async through a 3-stage ring.  The paged layouts are fragment-shaped already:
//   K cache [D/8][bs][8]: (d-octet j, token t) is one 16-byte granule = one ldmatrix row of the B
//                         fragment of S = Q.K^T (n = token, k = d);
//   V cache [D][bs]:      (d, 8 tokens) is one granule = one ldmatrix row of the B fragment of O = P.V
//                         (n = d, k = token); granules are regrouped as [token-octet][d] on the way in so
//                         that the 8 rows of an ldmatrix tile are one 128-byte line (no bank conflicts);
//   new K [token][D]:     stored in the SAME shared layout as cached K (XOR-swizzled [j][token]), so the
//                         Q.K^T loop is shared by both phases;
//   new V [token][D]:     XOR-swizzled rows, read with ldmatrix.trans.
// Stale slots of the last cached block and rows past the sequence end are zeroed on the V side and
// masked on the logit side, so uninitialised cache memory (NaN bit patterns included) cannot leak.
//
// The legacy warp-level MMA is the first correct path for this row; a tcgen05/TMEM version is the
// follow-up.
#pragma once

#include "paged_attention_mma.cuh"

namespace b200 {

struct CtxAttnParams {
  const void* q;
  const void* k;
  const void* v;
  void* o;
  const void* k_cache;
  const void* v_cache;
  const int32_t* b_loc;        // [batch, b_loc_stride] block table
  const int32_t* b_start_loc;  // [batch] first row of the sequence's new tokens in q/k/v/o
  const int32_t* b_seq_len;    // [batch] ctx_len + number of new tokens
  const int32_t* b_ctx_len;    // [batch] tokens already in the cache
  const float* alibi_slopes;   // [num_heads] or null
  float scale_log2;            // sm_scale * log2(e)
  int64_t q_ts, q_hs, k_ts, k_hs, v_ts, v_hs, o_ts, o_hs;  // token / head strides in elements
  int64_t kc_bs, kc_hs, vc_bs, vc_hs;                      // cache block / head strides in elements
  int32_t b_loc_stride;
  int32_t num_heads;
  int32_t num_kv_heads;
  int32_t row_limit;  // new-token rows >= row_limit are not computed: ceil(max_input_len / 128) * 128 (the reference's grid)
  int32_t pair_group; // (head, sequence) pairs whose row blocks are issued longest-first together (ctx_cta_coords); <= 1: one
#ifdef B200_CTX_TRACE
  unsigned long long* trace;  // development builds only (make TRACE=1): clock64 stamps of CTA (0,0,0), tools/ctx_trace.py
#endif
};
#ifdef B200_CTX_TRACE
// who: 0-7 = softmax warps (0-3 block A, 4-7 block B), 8 = MMA warp, 9 = V loader (warp 12), 10 = K loader (warp 9); tile < 64; event < 8
#define B200_CTX_STAMP(p, who, tile, ev)                                                              \
  do {                                                                                                \
    if ((p).trace && blockIdx.x == 0 && blockIdx.y == 0 && blockIdx.z == 0 && (tile) < 64)            \
      (p).trace[(((who) * 64 + (tile)) * 8) + (ev)] = (unsigned long long)clock64();                 \
  } while (0)
#else
#define B200_CTX_STAMP(p, who, tile, ev) do {} while (0)
#endif

// CTA -> (row block, head, sequence) for the tcgen05 kernels.  CTAs are issued with blockIdx.x fastest.  Causal work grows
// with the row block, so a plain (row block, head, sequence) grid ends on the heavy-to-light sawtooth of its last pairs and
// the last heavy CTA sets the tail.  Here the grid is walked in groups of `pair_group` (head, sequence) pairs: inside a
// group the heaviest row block of EVERY pair goes first, then the next heaviest, ... (longest-first across the group), so
// the grid ends on a whole group's lightest CTAs.  Consecutive heads stay in one group, so the query heads of a KV head
// still run together and share its K/V in L2; groups keep the K/V live at any time to a few pairs' worth.
__device__ __forceinline__ void ctx_cta_coords(int pair_group, int& mblk, int& head, int& seq) {
  const int M = gridDim.x, H = gridDim.y, P = H * gridDim.z;
  const int G = max(1, min(pair_group, P));
  const int L = blockIdx.x + M * (blockIdx.y + H * blockIdx.z);
  const int g = L / (M * G);
  const int r = L - g * M * G;
  const int Gp = min(G, P - g * G);  // the last group may be short
  const int mr = r / Gp;
  const int pair = g * G + (r - mr * Gp);
  mblk = M - 1 - mr;
  seq = pair / H;
  head = pair - seq * H;
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void ldsm_x4_trans(uint32_t addr, uint32_t& r0, uint32_t& r1, uint32_t& r2, uint32_t& r3) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3)
               : "r"(addr));
}

constexpr int kCtxBM = 128;      // query rows per CTA
constexpr int kCtxBN = 64;       // KV tokens per tile
constexpr int kCtxStages = 3;
constexpr int kCtxThreads = 256;

template <int D>
constexpr int ctx_attn_smem_bytes() {
  return kCtxStages * 2 * kCtxBN * D * 2;
}

template <typename MT, int D, int BS>
__global__ void __launch_bounds__(kCtxThreads, 1) context_attn_kernel(const CtxAttnParams p) {
  constexpr int BM = kCtxBM, BN = kCtxBN, NST = kCtxStages;
  constexpr int G = D / 8;           // 16-byte granules per token row
  constexpr int KS = D / 16;         // k-steps of Q.K^T
  constexpr int NT = D / 8;          // n-tiles of P.V
  constexpr int TILE_UNITS = BN * G; // granules of one K (or V) tile
  constexpr int HG = BS / 8;         // token octets per cache block
  static_assert(D % 16 == 0 && BN % BS == 0 && BS % 8 == 0, "tile shapes");
  // XOR swizzle of the new-token V rows: spreads the 8 rows of an ldmatrix.trans tile over 8 bank groups
  constexpr int SWZ_DIV = G >= 8 ? 1 : 8 / G;
  constexpr int SWZ_MASK = G >= 8 ? 7 : G - 1;

  extern __shared__ __align__(128) uint4 ctx_smem[];
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31, g = lane >> 2, t = lane & 3;
  const int mblk = gridDim.x - 1 - blockIdx.x;  // heaviest (most causal columns) row blocks first
  const int head = blockIdx.y, seq = blockIdx.z;
  const int ctx_len = p.b_ctx_len[seq];
  const int q_len = p.b_seq_len[seq] - ctx_len;
  const int m0 = mblk * BM;
  if (m0 >= q_len) return;
  const int64_t start = p.b_start_loc[seq];
  const int kvh = head / (p.num_heads / p.num_kv_heads);
  const int n_ctx_tiles = (ctx_len + BN - 1) / BN;
  const int new_end = min(q_len, m0 + BM);
  const int n_tiles = n_ctx_tiles + (new_end + BN - 1) / BN;

  const MT* kc = static_cast<const MT*>(p.k_cache) + (int64_t)kvh * p.kc_hs;
  const MT* vc = static_cast<const MT*>(p.v_cache) + (int64_t)kvh * p.vc_hs;
  const MT* kn = static_cast<const MT*>(p.k) + start * p.k_ts + (int64_t)kvh * p.k_hs;
  const MT* vn = static_cast<const MT*>(p.v) + start * p.v_ts + (int64_t)kvh * p.v_hs;
  const int32_t* bt = p.b_loc + (int64_t)seq * p.b_loc_stride;

  auto load_tile = [&](int ti, int stage) {
    uint4* Ks = ctx_smem + stage * 2 * TILE_UNITS;
    uint4* Vs = Ks + TILE_UNITS;
    if (ti < n_ctx_tiles) {
      const int n0 = ti * BN;
#pragma unroll
      for (int e = tid; e < TILE_UNITS; e += kCtxThreads) {
        const int tt = e % BS, j = (e / BS) % G, c = e / (BS * G);
        const int tok = c * BS + tt;
        if (n0 + tok < ctx_len) {
          const int64_t bn = bt[(n0 + tok) / BS];
          cp_async16(&Ks[j * BN + (tok ^ (j & 7))], kc + bn * p.kc_bs + (j * BS + tt) * 8);
        }
      }
#pragma unroll
      for (int e = tid; e < TILE_UNITS; e += kCtxThreads) {
        const int hh = e % HG, d = (e / HG) % D, c = e / (HG * D);
        const int tg = c * HG + hh;
        uint4* dst = &Vs[tg * D + d];
        if (n0 + tg * 8 < ctx_len) {
          const int64_t bn = bt[(n0 + tg * 8) / BS];
          cp_async16(dst, vc + bn * p.vc_bs + d * BS + hh * 8);
        } else {
          *dst = make_uint4(0u, 0u, 0u, 0u);
        }
      }
    } else {
      const int n0 = (ti - n_ctx_tiles) * BN;
#pragma unroll
      for (int e = tid; e < TILE_UNITS; e += kCtxThreads) {
        const int j = e % G, tok = e / G;
        uint4* vdst = &Vs[tok * G + (j ^ ((tok / SWZ_DIV) & SWZ_MASK))];
        if (n0 + tok < q_len) {
          cp_async16(&Ks[j * BN + (tok ^ (j & 7))], kn + (int64_t)(n0 + tok) * p.k_ts + j * 8);
          cp_async16(vdst, vn + (int64_t)(n0 + tok) * p.v_ts + j * 8);
        } else {
          *vdst = make_uint4(0u, 0u, 0u, 0u);
        }
      }
    }
  };

  // prologue: first tiles in flight while the query fragments are fetched
#pragma unroll
  for (int s = 0; s < NST - 1; ++s) {
    if (s < n_tiles) load_tile(s, s);
    cp_async_commit();
  }

  // ---- Q fragments (A operand), rows past the sequence end read as zero --------------------
  const int row0 = m0 + warp * 16 + g;  // this thread's rows: row0 and row0 + 8 (new-token index)
  uint32_t qa[KS][4];
  {
    const MT* qb = static_cast<const MT*>(p.q) + start * p.q_ts + (int64_t)head * p.q_hs;
    const bool ok0 = row0 < q_len, ok1 = row0 + 8 < q_len;
    const MT* q0 = qb + (int64_t)row0 * p.q_ts;
    const MT* q1 = qb + (int64_t)(row0 + 8) * p.q_ts;
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
      const int d = 16 * ks + 2 * t;
      qa[ks][0] = ok0 ? *reinterpret_cast<const uint32_t*>(q0 + d) : 0u;
      qa[ks][1] = ok1 ? *reinterpret_cast<const uint32_t*>(q1 + d) : 0u;
      qa[ks][2] = ok0 ? *reinterpret_cast<const uint32_t*>(q0 + d + 8) : 0u;
      qa[ks][3] = ok1 ? *reinterpret_cast<const uint32_t*>(q1 + d + 8) : 0u;
    }
  }

  float o[NT][4];
#pragma unroll
  for (int i = 0; i < NT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.f;
  float m_run[2] = {-INFINITY, -INFINITY}, l_run[2] = {0.f, 0.f};
  const float slope2 = p.alibi_slopes ? p.alibi_slopes[head] * 1.4426950408889634f : 0.f;
  const bool has_alibi = p.alibi_slopes != nullptr;
  const int warp_row_max = m0 + warp * 16 + 15;

  for (int ti = 0; ti < n_tiles; ++ti) {
    cp_async_wait<NST - 2>();
    __syncthreads();  // tile ti has landed for every thread; everyone is done with tile ti-1's stage
    {
      const int nx = ti + NST - 1;
      if (nx < n_tiles) load_tile(nx, nx % NST);
      cp_async_commit();
    }
    const bool is_ctx = ti < n_ctx_tiles;
    const int n0 = (is_ctx ? ti : ti - n_ctx_tiles) * BN;
    uint4* Ks = ctx_smem + (ti % NST) * 2 * TILE_UNITS;
    uint4* Vs = Ks + TILE_UNITS;
    if (is_ctx && n0 + BN > ctx_len && ((ctx_len - n0) & 7)) {
      // the octet that straddles ctx_len: slots past the end are stale cache memory
      const int tgb = (ctx_len - n0) >> 3, rem = (ctx_len - n0) & 7;
      for (int d = tid; d < D; d += kCtxThreads) {
        MT* gp = reinterpret_cast<MT*>(&Vs[tgb * D + d]);
        for (int i = rem; i < 8; ++i) gp[i] = MT(0.f);
      }
      __syncthreads();
    }
    if (!is_ctx && n0 > warp_row_max) continue;  // tile entirely above this warp's diagonal

    // ---- S = Q.K^T -------------------------------------------------------------------------
    float s[BN / 8][4];
    const uint32_t ks_u = smem_u32(Ks);
#pragma unroll
    for (int nt = 0; nt < BN / 8; ++nt) {
      s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.f;
#pragma unroll
      for (int k2 = 0; k2 < KS / 2; ++k2) {
        const int j = 4 * k2 + (lane >> 3);
        uint32_t b0, b1, b2, b3;
        ldsm_x4(ks_u + (j * BN + nt * 8 + ((lane & 7) ^ (j & 7))) * 16, b0, b1, b2, b3);
        mma16816<MT>(s[nt], qa[2 * k2], b0, b1);
        mma16816<MT>(s[nt], qa[2 * k2 + 1], b2, b3);
      }
      if constexpr (KS & 1) {
        const int j = 2 * (KS - 1) + ((lane >> 3) & 1);
        uint32_t b0, b1;
        ldsm_x2(ks_u + (j * BN + nt * 8 + ((lane & 7) ^ (j & 7))) * 16, b0, b1);
        mma16816<MT>(s[nt], qa[KS - 1], b0, b1);
      }
    }

    // ---- scale, bias, mask, online softmax (log2 domain) ----------------------------------
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < BN / 8; ++nt) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const int col = n0 + nt * 8 + 2 * t + (i & 1);
        const int row = row0 + (i >> 1) * 8;
        float v = s[nt][i] * p.scale_log2;
        // key position minus query position: cached token `col` vs new token `row` at ctx_len + row
        if (has_alibi) v += slope2 * (float)(is_ctx ? col - ctx_len - row : col - row);
        const bool valid = is_ctx ? (col < ctx_len) : (col <= row);
        v = valid ? v : -INFINITY;
        s[nt][i] = v;
        mx[i >> 1] = fmaxf(mx[i >> 1], v);
      }
    }
    float alpha[2];
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
      mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
      const float m_new = fmaxf(m_run[r], mx[r]);
      alpha[r] = ex2_approx(m_run[r] - m_new);
      m_run[r] = m_new;
    }
    float sum[2] = {0.f, 0.f};
#pragma unroll
    for (int nt = 0; nt < BN / 8; ++nt) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const float e = ex2_approx(s[nt][i] - m_run[i >> 1]);
        s[nt][i] = e;
        sum[i >> 1] += e;
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) l_run[r] = l_run[r] * alpha[r] + sum[r];
#pragma unroll
    for (int dn = 0; dn < NT; ++dn) {
      o[dn][0] *= alpha[0]; o[dn][1] *= alpha[0];
      o[dn][2] *= alpha[1]; o[dn][3] *= alpha[1];
    }

    // ---- O += P.V ----------------------------------------------------------------------------
    const uint32_t vs_u = smem_u32(Vs);
    const int mat = lane >> 3;
#pragma unroll
    for (int kt = 0; kt < BN / 16; ++kt) {
      uint32_t a[4];
      a[0] = pack2<MT>(s[2 * kt][0], s[2 * kt][1]);
      a[1] = pack2<MT>(s[2 * kt][2], s[2 * kt][3]);
      a[2] = pack2<MT>(s[2 * kt + 1][0], s[2 * kt + 1][1]);
      a[3] = pack2<MT>(s[2 * kt + 1][2], s[2 * kt + 1][3]);
#pragma unroll
      for (int d2 = 0; d2 < NT / 2; ++d2) {
        uint32_t b0, b1, b2, b3;
        if (is_ctx) {
          const int d = 8 * (2 * d2 + (mat >> 1)) + (lane & 7);
          ldsm_x4(vs_u + ((kt * 2 + (mat & 1)) * D + d) * 16, b0, b1, b2, b3);
        } else {
          const int tok = 16 * kt + 8 * (mat & 1) + (lane & 7);
          const int jd = 2 * d2 + (mat >> 1);
          ldsm_x4_trans(vs_u + (tok * G + (jd ^ ((tok / SWZ_DIV) & SWZ_MASK))) * 16, b0, b1, b2, b3);
        }
        mma16816<MT>(o[2 * d2], a, b0, b1);
        mma16816<MT>(o[2 * d2 + 1], a, b2, b3);
      }
    }
  }
  cp_async_wait<0>();

  // ---- epilogue: normalise and store -----------------------------------------------------------
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 1);
    l_run[r] += __shfl_xor_sync(0xffffffffu, l_run[r], 2);
  }
  MT* ob = static_cast<MT*>(p.o) + start * p.o_ts + (int64_t)head * p.o_hs;
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    const int row = row0 + 8 * r;
    if (row < q_len) {
      const float inv = 1.f / l_run[r];
      MT* orow = ob + (int64_t)row * p.o_ts;
#pragma unroll
      for (int dn = 0; dn < NT; ++dn)
        *reinterpret_cast<uint32_t*>(orow + 8 * dn + 2 * t) = pack2<MT>(o[dn][2 * r] * inv, o[dn][2 * r + 1] * inv);
    }
  }
}

template <typename MT, int D, int BS>
int launch_ctx_attn(const CtxAttnParams& p, int batch, int max_input_len, cudaStream_t stream) {
  constexpr int smem = ctx_attn_smem_bytes<D>();
  static bool attr_set[64] = {false};  // per instantiation and per device (function attributes are per device)
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  dev &= 63;
  if (!attr_set[dev]) {
    B200_CUDA_TRY(cudaFuncSetAttribute(context_attn_kernel<MT, D, BS>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    attr_set[dev] = true;
  }
  dim3 grid((max_input_len + kCtxBM - 1) / kCtxBM, p.num_heads, batch);
  context_attn_kernel<MT, D, BS><<<grid, kCtxThreads, smem, stream>>>(p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename MT>
int launch_ctx_attn_dtype(const CtxAttnParams& p, int head_size, int block_size, int batch, int max_input_len,
                          bool tensor_core_gen5, cudaStream_t stream);

}  // namespace b200
