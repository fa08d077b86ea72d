// Prefix prefill on tcgen05/TMEM, 256 query rows per CTA (two 128-row blocks A and B), head sizes 128 and 64.
//
// context_attn_tc_kernel (context_attention_tc.cuh) does 128 flop per byte it pulls out of L2 -- every 128-row CTA
// reads all of its K/V -- and ncu (profiles/r05_ctx_tc_stalls_after.txt) shows where that ends: the MMA thread waits
// for V tiles 28 % of its time, the softmax warps for S 35 % of theirs, L2 throughput at 48 %.  Here one CTA owns 256
// consecutive query rows of a (sequence, head): every K/V tile is loaded ONCE and used by both row blocks (half the
// L2 requests and loader work per flop).
//
// A first version with 128-token tiles (S_A | S_B | O_A | O_B fill the 512 TMEM columns, so a block's next Q.K^T can
// only be issued after its P.V) reached 647 TFLOP/s with the softmax warps waiting for S half of the time
// (profiles/r05_ctx_tc256_stalls.txt): softmax -> P.V -> Q.K^T -> softmax is a serial chain per block and the two
// blocks run it in phase.  With 64-token tiles an S buffer is 64 columns, every block has TWO of them
// (S_A0 S_A1 | S_B0 S_B1 | O_A | O_B), and Q.K^T runs a tile ahead of the softmax: the chain per block is the softmax
// alone.
//
// Warps (16): 0-3 softmax of block A, 4-7 of block B -- ONE thread per query row holding the row's 64 scores in
// registers (no cross-thread exchange), 8 = TMEM allocator + the MMA-issuing thread, 9-10 = K (+Q) loaders, 11 idle,
// 12-15 = V loaders.  Shared memory: Q_A, Q_B (32 KiB each), kStg K and kStgV V stages of 16 KiB.
// Layouts, descriptors, loaders and the lazy rescale are those of context_attention_tc.cuh.
#pragma once

#include "context_attention_tc.cuh"

namespace b200 {
namespace tc256 {

using namespace tc;

constexpr int kRows = 2 * kBM;   // query rows per CTA
constexpr int kThreads256 = 512;
constexpr int kKLoad = 64;  // K (+Q) loader threads; V loader threads: Cfg::kVLoad
constexpr int kMaxStg = 4;
// TN = KV tokens per tile.  TN = 128: one S buffer per block (Q.K^T of tile i+1 follows P.V of tile i), two 32 KiB
// stages each for K and V, 192 registers per softmax thread (128 scores).  TN = 64: two S buffers per block (Q.K^T a
// tile ahead), four 16 KiB stages, 160 registers.  Measured: see default_tn().
template <int TN, int D = kD>
struct Cfg {
  static constexpr int kNSB = 128 / TN;              // S buffers per row block
  static constexpr int kStg = 256 / TN;              // K and V ring depth
  static constexpr int kQTileBytes = kBM * D * 2;
  static constexpr int kKVTileBytes = TN * D * 2;
  static constexpr int kVLoad = D == 128 ? 128 : 64;  // V loader threads (a pass of the cached-V loader covers 32-128 d rows)
  static constexpr int kSmem = 2 * kQTileBytes + 2 * kStg * kKVTileBytes + 512;
  static_assert(D == 64 || D == 128, "head sizes of this kernel");
  static_assert(TN == 64 || TN == 128, "tile sizes");
  static_assert(kStg <= kMaxStg, "ring depth");
};

struct Barriers256 {
  uint64_t q_full;
  uint64_t k_full[kMaxStg], k_empty[kMaxStg], v_full[kMaxStg], v_empty[kMaxStg];
  uint64_t s_full[2][2];  // [row block][S buffer]
  uint64_t p_full[2][2];  // [row block][S buffer]: the softmax may finish tile i+1 before the MMA thread has looked at
                          // tile i's P (S_{i+1} is ready early), so a single barrier per block could run two phases
                          // ahead of its waiter -- per buffer it cannot (S_{i+2} needs P_i.V issued)
  uint64_t o_full[2][4];  // per row block, tile t on [t % 4] (observed lazily, see context_attention_tc.cuh)
  uint32_t tmem_base;
};
static_assert(sizeof(Barriers256) <= 512, "barrier block");

// One softmax thread = one query row of block X: all TN columns of every tile.
template <typename MT, int TN, int D>
__device__ __forceinline__ void softmax_rows(const CtxAttnParams& p, Barriers256* bar, uint32_t tmem, int X, int warp, int tid,
                                             int head, int m0, int q_len, int ctx_len, int n_ctx_tiles, int n_x, int64_t start) {
  const int row = tid & 127;
  const uint32_t lane_off = (uint32_t)(32 * (warp & 3)) << 16;
  const int qrow = m0 + kBM * X + row;
  const bool has_alibi = p.alibi_slopes != nullptr;
  const float slope2 = has_alibi ? p.alibi_slopes[head] * 1.4426950408889634f : 0.f;
  const uint32_t o_addr = tmem + lane_off + 256 + X * 128;
  float m_used = -INFINITY, l_run = 0.f;  // lazily updated reference point (see context_attention_tc.cuh)
  int o_next = 0;                         // first tile whose P.V completion this thread has not observed yet
  auto observe_o_upto = [&](int last) {
    for (; o_next <= last; ++o_next) mbar_wait(&bar->o_full[X][o_next & 3], (o_next >> 2) & 1);
  };

  constexpr int kTN = TN, NSB = Cfg<TN>::kNSB;
  for (int i = 0; i < n_x; ++i) {
    const bool is_ctx = i < n_ctx_tiles;
    const int n0 = (is_ctx ? i : i - n_ctx_tiles) * kTN;
    const int limit = is_ctx ? ctx_len - n0 : qrow - n0 + 1;  // columns [0, limit) are visible to this row
    const float bias0 = slope2 * (float)(is_ctx ? n0 - ctx_len - qrow : n0 - qrow);
    const uint32_t s_addr = tmem + lane_off + X * 128 + (i % NSB) * kTN;
    mbar_wait(&bar->s_full[X][i % NSB], (i / NSB) & 1);
    tc_fence_after();
    if ((tid & 31) == 0) B200_CTX_STAMP(p, warp, i, 0);
    float s[kTN];
#pragma unroll
    for (int c = 0; c < kTN / 32; ++c) tmem_ld32(s_addr + 32 * c, reinterpret_cast<uint32_t*>(&s[32 * c]));
    tmem_ld_wait();
    if ((tid & 31) == 0) B200_CTX_STAMP(p, warp, i, 1);
    // warp-uniform fast path: nothing masked, no bias -> raw maximum, scale folded into the exponent's FMA
    const bool plain = !has_alibi && __all_sync(0xffffffffu, limit >= kTN);
    float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
    float mx, mul;
    if (plain) {
#pragma unroll
      for (int k = 0; k < kTN; ++k) mx4[k & 3] = fmaxf(mx4[k & 3], s[k]);
      mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;  // scale > 0
      mul = p.scale_log2;
    } else {
#pragma unroll
      for (int k = 0; k < kTN; ++k) {
        float v = fmaf(s[k], p.scale_log2, has_alibi ? fmaf(slope2, (float)k, bias0) : 0.f);
        v = k < limit ? v : -INFINITY;
        s[k] = v;
        mx4[k & 3] = fmaxf(mx4[k & 3], v);
      }
      mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
      mul = 1.f;
    }
    if (__any_sync(0xffffffffu, mx > m_used + 8.f)) {
      const float m_new = fmaxf(m_used, mx);
      const float alpha = ex2_approx(m_used - m_new);
      if (i > 0) {
        observe_o_upto(i - 1);
        tc_fence_after();
#pragma unroll 1
        for (int c = 0; c < D / 32; ++c) {
          uint32_t r[32];
          tmem_ld32(o_addr + 32 * c, r);
          tmem_ld_wait();
#pragma unroll
          for (int k = 0; k < 32; ++k) r[k] = __float_as_uint(__uint_as_float(r[k]) * alpha);
          tmem_st16(o_addr + 32 * c, r);
          tmem_st16(o_addr + 32 * c + 16, r + 16);
        }
      }
      l_run *= alpha;
      m_used = m_new;
    }
    if ((tid & 31) == 0) B200_CTX_STAMP(p, warp, i, 2);
    // P = 2^(score - reference) over the S columns already consumed (two elements per 32-bit column)
    float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int c = 0; c < kTN / 32; ++c) {
      uint32_t pk[16];
#pragma unroll
      for (int k = 0; k < 32; k += 2) {
        const float e0 = ex2_approx(fmaf(s[32 * c + k], mul, -m_used));  // masked columns are -inf -> 0
        const float e1 = ex2_approx(fmaf(s[32 * c + k + 1], mul, -m_used));
        sum4[(k >> 1) & 3] += e0 + e1;
        pk[k >> 1] = pack2<MT>(e0, e1);
      }
      tmem_st16(s_addr + 16 * c, pk);
    }
    tmem_st_wait();
    if ((tid & 31) == 0) B200_CTX_STAMP(p, warp, i, 3);
    observe_o_upto(i - 4);
    tc_fence_before();
    mbar_arrive(&bar->p_full[X][i % NSB]);
    if ((tid & 31) == 0) B200_CTX_STAMP(p, warp, i, 4);
    l_run += (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
  }

  if (n_x > 0) {
    observe_o_upto(n_x - 1);
    tc_fence_after();
    const float inv = 1.f / l_run;
    MT* orow = static_cast<MT*>(p.o) + (start + qrow) * p.o_ts + (int64_t)head * p.o_hs;
#pragma unroll 1
    for (int c = 0; c < D / 32; ++c) {
      uint32_t r[32];
      tmem_ld32(o_addr + 32 * c, r);
      tmem_ld_wait();
      if (qrow < min(q_len, p.row_limit)) {
#pragma unroll
        for (int k = 0; k < 32; k += 8) {
          uint4 w;
          w.x = pack2<MT>(__uint_as_float(r[k]) * inv, __uint_as_float(r[k + 1]) * inv);
          w.y = pack2<MT>(__uint_as_float(r[k + 2]) * inv, __uint_as_float(r[k + 3]) * inv);
          w.z = pack2<MT>(__uint_as_float(r[k + 4]) * inv, __uint_as_float(r[k + 5]) * inv);
          w.w = pack2<MT>(__uint_as_float(r[k + 6]) * inv, __uint_as_float(r[k + 7]) * inv);
          *reinterpret_cast<uint4*>(orow + 32 * c + k) = w;
        }
      }
    }
  }
}

template <typename MT, int BS, int TN, int D>
__global__ void __launch_bounds__(kThreads256, 1) context_attn_tc256_kernel(const CtxAttnParams p) {
  using C = Cfg<TN, D>;
  constexpr int kTN = TN, NSB = C::kNSB, kStg = C::kStg, kStgV = C::kStg, kVLoad = C::kVLoad;
  constexpr int kKVTileBytes = C::kKVTileBytes, kQTileBytes = C::kQTileBytes;
  constexpr int Q_UNITS = kBM * (D / 8);    // 16-byte granules of a Q tile
  constexpr int KV_UNITS = kTN * (D / 8);   // ... of a K or V tile
  extern __shared__ __align__(1024) uint8_t tc256_smem[];
  uint4* q_s = reinterpret_cast<uint4*>(tc256_smem);  // Q_A, Q_B
  uint4* k_s = q_s + 2 * Q_UNITS;
  uint4* v_s = k_s + kStg * KV_UNITS;
  Barriers256* bar = reinterpret_cast<Barriers256*>(v_s + kStgV * KV_UNITS);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int mblk, head, seq;
  ctx_cta_coords(p.pair_group, mblk, head, seq);
  const int ctx_len = p.b_ctx_len[seq];
  const int q_len = p.b_seq_len[seq] - ctx_len;
  const int m0 = mblk * kRows;
  if (m0 >= q_len) return;
  const int64_t start = p.b_start_loc[seq];
  const int kvh = head / (p.num_heads / p.num_kv_heads);
  const int n_ctx_tiles = (ctx_len + kTN - 1) / kTN;
  // tiles each row block attends: the cached prefix, then the new tokens up to its own last row
  const int n_a = n_ctx_tiles + (min(q_len, m0 + kBM) + kTN - 1) / kTN;
  const int n_b = m0 + kBM < min(q_len, p.row_limit) ? n_ctx_tiles + (min(q_len, m0 + kRows) + kTN - 1) / kTN : 0;
  const int n_load = max(n_a, n_b);

  if (tid == 0) {
    mbar_init(&bar->q_full, kKLoad);
    for (int s = 0; s < kStg; ++s) {
      mbar_init(&bar->k_full[s], kKLoad);
      mbar_init(&bar->k_empty[s], 1);
    }
    for (int s = 0; s < kStgV; ++s) {
      mbar_init(&bar->v_full[s], kVLoad);
      mbar_init(&bar->v_empty[s], 1);
    }
    for (int x = 0; x < 2; ++x) {
      mbar_init(&bar->s_full[x][0], 1);
      mbar_init(&bar->s_full[x][1], 1);
      mbar_init(&bar->p_full[x][0], 128);
      mbar_init(&bar->p_full[x][1], 128);
      for (int s = 0; s < 4; ++s) mbar_init(&bar->o_full[x][s], 1);
    }
    mbar_fence_init();
  }
  if (warp == 8) tmem_alloc_512(&bar->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp < 8) {
    if constexpr (TN == 128) asm volatile("setmaxnreg.inc.sync.aligned.u32 192;");  // (8 x 192 + 8 x 64) x 32 = 64 Ki
    else asm volatile("setmaxnreg.inc.sync.aligned.u32 160;");                       // (8 x 160 + 8 x 96) x 32
    const int X = warp >> 2;
    softmax_rows<MT, TN, D>(p, bar, tmem, X, warp, tid, head, m0, q_len, ctx_len, n_ctx_tiles, X ? n_b : n_a, start);
  } else {
    if constexpr (TN == 128) asm volatile("setmaxnreg.dec.sync.aligned.u32 64;");
    else asm volatile("setmaxnreg.dec.sync.aligned.u32 96;");
    if (warp == 8) {
      // ================= the MMA warp (uniform: every lane runs the loop, elect.sync issues) =================
      {
        const uint32_t idesc_qk = instr_desc<MT>(kTN, false);  // S = Q.K^T: N = 64 tokens, K-major B
        const uint32_t idesc_pv = instr_desc<MT>(D, false);    // O += P.V over cached V: N = 128, K-major B
        const uint32_t idesc_mn = instr_desc<MT>(D, true);     // ... over new V: MN-major B
        // Shared-memory descriptors are built ONCE: the start-address field is the low 14 bits (address >> 4), so the
        // descriptor of "base + off" is the base descriptor plus off >> 4.  (ncu, r05: building two descriptors per
        // tcgen05.mma from scratch made the issuing thread the bottleneck -- 80 % busy, tensor pipe 26 % active.)
        const uint64_t qd[2] = {smem_desc(smem_u32(q_s), kBM * 16, 128), smem_desc(smem_u32(q_s) + kQTileBytes, kBM * 16, 128)};
        const uint64_t kd0 = smem_desc(smem_u32(k_s), kTN * 16, 128);      // K-major [d-octet][token]
        const uint64_t vd0_ctx = smem_desc(smem_u32(v_s), D * 16, 128);    // cached V: K-major [token-octet][d]
        const uint64_t vd0_new = smem_desc(smem_u32(v_s), 128, kTN * 16);  // new V: MN-major [d-octet][token]
        constexpr uint64_t kStageStep = kKVTileBytes >> 4;
        mbar_wait(&bar->q_full, 0);
        auto issue_qk = [&](int X, int i) {  // S_X[i & 1] = Q_X.K_i^T
          const int st = i % kStg;
          if (lane == 0 && i > 0) B200_CTX_STAMP(p, 8, i - 1, 5 + 2 * X);  // P.V of tile i - 1 issued, about to wait for K tile i
          mbar_wait(&bar->k_full[st], (i / kStg) & 1);
          tc_fence_after();
          const uint64_t kd = kd0 + (uint64_t)st * kStageStep;
          const uint64_t qx = qd[X];
          const uint32_t d_addr = tmem + X * 128 + (i % NSB) * kTN;
#pragma unroll
          for (int ks = 0; ks < D / 16; ++ks)
            mma_ss_u(d_addr, qx + (uint64_t)(ks * ((2 * kBM * 16) >> 4)), kd + (uint64_t)(ks * ((2 * kTN * 16) >> 4)), idesc_qk,
                   ks > 0);
          if (X == 1 || i >= n_b) tc_commit_u(&bar->k_empty[st]);  // last reader of K tile i
          tc_commit_u(&bar->s_full[X][i % NSB]);
        };
        auto issue_pv = [&](int X, int i) {  // O_X += P_X.V_i
          const int st = i % kStgV;
          mbar_wait(&bar->v_full[st], (i / kStgV) & 1);
          if (lane == 0) B200_CTX_STAMP(p, 8, i, 4 + 2 * X);
          mbar_wait(&bar->p_full[X][i % NSB], (i / NSB) & 1);
          tc_fence_after();
          if (lane == 0) B200_CTX_STAMP(p, 8, i, 2 * X);
          const bool is_ctx = i < n_ctx_tiles;
          // cached V: k-step = two token octets (2 * kD granules); new V: k-step = 16 tokens (16 granules)
          const uint64_t vd = (is_ctx ? vd0_ctx : vd0_new) + (uint64_t)st * kStageStep;
          const uint64_t step = is_ctx ? (uint64_t)((2 * D * 16) >> 4) : (uint64_t)((16 * 16) >> 4);
          const uint32_t idesc = is_ctx ? idesc_pv : idesc_mn;
          const uint32_t o_addr = tmem + 256 + X * 128, p_addr = tmem + X * 128 + (i % NSB) * kTN;
#pragma unroll
          for (int kt = 0; kt < kTN / 16; ++kt) mma_ts_u(o_addr, p_addr + kt * 8, vd + kt * step, idesc, (i | kt) > 0);
          if (X == 1 || i >= n_b) tc_commit_u(&bar->v_empty[st]);  // last reader of V tile i
          tc_commit_u(&bar->o_full[X][i & 3]);
        };
        // Q.K^T runs NSB - 1 tiles ahead of the softmax in each block: S_X[(i + NSB) % NSB] is free again, in pipe
        // order, as soon as P_X(i).V has been issued.  (Block B's Q.K_t^T is always issued before block A's Q.K_{t+1}^T:
        // the K loader announces tile t+1 only after it has seen tile t's stage released.)
        for (int t = 0; t < NSB; ++t) {
          if (t < n_a) issue_qk(0, t);
          if (t < n_b) issue_qk(1, t);
        }
        for (int i = 0; i < n_load; ++i) {
          if (i < n_a) {
            issue_pv(0, i);
            if (i + NSB < n_a) issue_qk(0, i + NSB);
            if (lane == 0) B200_CTX_STAMP(p, 8, i, 1);
          }
          if (i < n_b) {
            issue_pv(1, i);
            if (i + NSB < n_b) issue_qk(1, i + NSB);
            if (lane == 0) B200_CTX_STAMP(p, 8, i, 3);
          }
        }
      }
      __syncwarp();
    } else if (warp < 11) {
      // ================= K (and Q) loaders =================
      const int lt = tid - 9 * 32;
      const MT* kc = static_cast<const MT*>(p.k_cache) + (int64_t)kvh * p.kc_hs;
      const MT* kn = static_cast<const MT*>(p.k) + start * p.k_ts + (int64_t)kvh * p.k_hs;
      const int32_t* bt = p.b_loc + (int64_t)seq * p.b_loc_stride;
      // block numbers of a cached-prefix tile (lane c: the c-th block), fetched before the wait for the stage
      auto k_blocks = [&](int ti) -> int {
        return ti < n_ctx_tiles ? ctx_tile_blocks<BS, kTN>(bt + ti * kTN / BS, ctx_len - ti * kTN, lane) : 0;
      };
      auto issue_k = [&](int ti, int my_bn) {
        uint4* Ks = k_s + (ti % kStg) * KV_UNITS;
        if (ti < n_ctx_tiles) {
          load_k_ctx_bn<MT, BS, kKLoad, kTN, D>(Ks, kc, p.kc_bs, my_bn, ctx_len - ti * kTN, lt);
        } else {
          const int n0 = (ti - n_ctx_tiles) * kTN;
          load_rows<MT, kKLoad, false, kTN, D>(Ks, kn + (int64_t)n0 * p.k_ts, p.k_ts, q_len - n0, lt);
        }
      };
      const MT* qb = static_cast<const MT*>(p.q) + (start + m0) * p.q_ts + (int64_t)head * p.q_hs;
      load_rows<MT, kKLoad, true, kBM, D>(q_s, qb, p.q_ts, q_len - m0, lt);
      load_rows<MT, kKLoad, true, kBM, D>(q_s + Q_UNITS, qb + (int64_t)kBM * p.q_ts, p.q_ts, q_len - m0 - kBM, lt);
      cp_async_commit();
      const int pro = min(kStg, n_load);
      for (int t = 0; t < pro; ++t) {
        issue_k(t, k_blocks(t));
        cp_async_commit();
      }
      cp_async_wait_upto(pro);  // Q (the oldest group) has landed
      fence_proxy_async_smem();
      mbar_arrive(&bar->q_full);
      // A tile is announced as soon as it has landed, one tile AHEAD of the refill that follows it in program order.
      // (r05 trace, tools/ctx_trace.py: with "announce i, wait for stage i to drain, issue i + kStg, announce i + 1" the
      // announcement of tile i + 1 -- long landed -- sat behind the wait for the MMA warp AND the ~2-3 k cycles it takes
      // to issue the next tile's copies, and the MMA warp spent 60 % of every period waiting for exactly that.)
      auto announce_k = [&](int t) {
        fence_proxy_async_smem();
        mbar_arrive(&bar->k_full[t % kStg]);
        if (tid == 9 * 32) B200_CTX_STAMP(p, 10, t, 2);
      };
      if (n_load > 0) {
        cp_async_wait_upto(pro - 1);
        announce_k(0);
      }
      // When the stage of the next refill has ALREADY drained (the consumer is ahead: small head sizes, short tiles) the
      // refill goes first, so that two tiles stay in flight; otherwise the landed tile is announced before the loader
      // blocks on the stage.  The decision is per warp (one instruction reads the barrier for all its lanes).
      int issued = pro;  // tile groups committed so far
      for (int i = 0; i < n_load; ++i) {
        const bool more = i + kStg < n_load;
        const int bn = more ? k_blocks(i + kStg) : 0;
        auto refill = [&]() {
          if (tid == 9 * 32) B200_CTX_STAMP(p, 10, i + kStg, 0);
          issue_k(i + kStg, bn);
          cp_async_commit();
          ++issued;
          if (tid == 9 * 32) B200_CTX_STAMP(p, 10, i + kStg, 1);
        };
        bool refilled = false;
        if (more && __all_sync(0xffffffffu, mbar_test(&bar->k_empty[i % kStg], (i / kStg) & 1))) {
          refill();
          refilled = true;
        }
        if (i + 1 < n_load) {
          cp_async_wait_upto(issued - (i + 2));  // groups younger than tile i + 1
          announce_k(i + 1);
        }
        if (more && !refilled) {
          mbar_wait(&bar->k_empty[i % kStg], (i / kStg) & 1);
          refill();
        }
      }
    } else if (warp >= 12 && warp < 12 + kVLoad / 32) {
      // ================= V loaders =================
      const int lt = tid - 12 * 32;
      const MT* vc = static_cast<const MT*>(p.v_cache) + (int64_t)kvh * p.vc_hs;
      const MT* vn = static_cast<const MT*>(p.v) + start * p.v_ts + (int64_t)kvh * p.v_hs;
      const int32_t* bt = p.b_loc + (int64_t)seq * p.b_loc_stride;
      auto v_blocks = [&](int ti) -> int {
        return ti < n_ctx_tiles ? ctx_tile_blocks<BS, kTN>(bt + ti * kTN / BS, ctx_len - ti * kTN, lane) : 0;
      };
      auto issue_v = [&](int ti, int my_bn) {
        uint4* Vs = v_s + (ti % kStgV) * KV_UNITS;
        if (ti < n_ctx_tiles) {
          load_v_ctx_bn<MT, BS, kVLoad, kTN, D>(Vs, vc, p.vc_bs, my_bn, ctx_len - ti * kTN, lt);
        } else {
          const int n0 = (ti - n_ctx_tiles) * kTN;
          load_rows<MT, kVLoad, true, kTN, D>(Vs, vn + (int64_t)n0 * p.v_ts, p.v_ts, q_len - n0, lt);
        }
      };
      const int pro = min(kStgV, n_load);
      for (int t = 0; t < pro; ++t) {
        issue_v(t, v_blocks(t));
        cp_async_commit();
      }
      auto announce_v = [&](int t) {  // the copies of tile t have landed (this thread's)
        if (t < n_ctx_tiles && t * kTN + kTN > ctx_len && ((ctx_len - t * kTN) & 7)) {
          // the token octet that straddles ctx_len: slots past the end are stale cache memory (may be NaN / Inf)
          asm volatile("bar.sync 1, %0;" ::"n"(kVLoad) : "memory");  // every V loader's copies of tile t landed
          const int tgb = (ctx_len - t * kTN) >> 3, rem = (ctx_len - t * kTN) & 7;
          for (int d = lt; d < D; d += kVLoad) {
            MT* gp = reinterpret_cast<MT*>(&v_s[(t % kStgV) * KV_UNITS + tgb * D + d]);
            for (int x = rem; x < 8; ++x) gp[x] = MT(0.f);
          }
        }
        fence_proxy_async_smem();
        mbar_arrive(&bar->v_full[t % kStgV]);
        if (tid == 12 * 32) B200_CTX_STAMP(p, 9, t, 2);
      };
      if (n_load > 0) {
        cp_async_wait_upto(pro - 1);
        announce_v(0);
      }
      int issued = pro;
      for (int i = 0; i < n_load; ++i) {  // the K loader's order: refill first if the stage is free, else announce first
        const bool more = i + kStgV < n_load;
        const int bn = more ? v_blocks(i + kStgV) : 0;
        auto refill = [&]() {
          if (tid == 12 * 32) B200_CTX_STAMP(p, 9, i + kStgV, 0);
          issue_v(i + kStgV, bn);
          cp_async_commit();
          ++issued;
          if (tid == 12 * 32) B200_CTX_STAMP(p, 9, i + kStgV, 1);
        };
        bool refilled = false;
        if (more && __all_sync(0xffffffffu, mbar_test(&bar->v_empty[i % kStgV], (i / kStgV) & 1))) {
          refill();
          refilled = true;
        }
        if (i + 1 < n_load) {
          cp_async_wait_upto(issued - (i + 2));
          announce_v(i + 1);
        }
        if (more && !refilled) {
          mbar_wait(&bar->v_empty[i % kStgV], (i / kStgV) & 1);
          refill();
        }
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == 8) tmem_dealloc_512(tmem);
}

// Tile size (B200_CTX_TN=64|128 overrides, A/B knob).  While the loaders announced a landed tile only after the next
// refill, 128-token tiles were faster (647 vs 598 TFLOP/s); with the loaders off the critical path
// (profiles/r05_prefix_prefill_loaders.jsonl) the look-ahead Q.K^T of the 64-token variant pays: head size 128 --
// 783 vs 736 (2k + 2k), 597 vs 581 (4k, no prefix), 789 vs 742, 875 vs 841, 484 vs 451; head size 64 -- 481 vs 502.
template <int D>
constexpr int default_tn() { return D == 128 ? 64 : 128; }

template <typename MT, int BS, int TN, int D>
int launch_ctx_attn_tc256_tn(const CtxAttnParams& p, int batch, int max_input_len, cudaStream_t stream) {
  static bool attr_set[64] = {false};
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  dev &= 63;
  if (!attr_set[dev]) {
    B200_CUDA_TRY(cudaFuncSetAttribute(context_attn_tc256_kernel<MT, BS, TN, D>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                       Cfg<TN, D>::kSmem));
    attr_set[dev] = true;
  }
  dim3 grid((max_input_len + kRows - 1) / kRows, p.num_heads, batch);
  context_attn_tc256_kernel<MT, BS, TN, D><<<grid, kThreads256, Cfg<TN, D>::kSmem, stream>>>(p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename MT, int BS, int D = kD>
int launch_ctx_attn_tc256(const CtxAttnParams& p, int batch, int max_input_len, cudaStream_t stream) {
  static const int tn = getenv("B200_CTX_TN") ? atoi(getenv("B200_CTX_TN")) : default_tn<D>();
  if (tn == 64) return launch_ctx_attn_tc256_tn<MT, BS, 64, D>(p, batch, max_input_len, stream);
  return launch_ctx_attn_tc256_tn<MT, BS, 128, D>(p, batch, max_input_len, stream);
}

}  // namespace tc256
}  // namespace b200
