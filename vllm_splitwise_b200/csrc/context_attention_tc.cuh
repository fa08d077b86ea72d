// Prefix ("context") prefill attention on the 5th-generation tensor cores (tcgen05 + TMEM), head size 128.
//
// Same contract as context_attn_kernel (context_attention.cuh); this is the GEMM-shaped version:
//   S = Q.K^T   : tcgen05.mma M=128 (query rows) x N=128 (KV tokens) x K=16, both operands from shared memory,
//                 fp32 accumulator in TMEM (three S buffers: Q.K^T runs two tiles ahead of the softmax)
//   P           : the softmax warps read S from TMEM (two threads per query row, 64 columns each; the row maximum
//                 is exchanged through shared memory once per tile) and write P in the element type back to TMEM over
//                 the S columns they have consumed
//   O += P.V    : tcgen05.mma with A = P from TMEM, B = V from shared memory, one fp32 accumulator in TMEM for the whole
//                 row block.  The softmax reference point follows the running maximum lazily (it moves only when the
//                 maximum grew by more than 2^8), so the accumulator is rescaled in TMEM rarely.
// Warp roles (12 warps): 0-7 softmax/epilogue (warp w reaches TMEM lanes 32*(w%4)..+31; warps 0-3 take tile columns
// 0-63, warps 4-7 columns 64-127), 8 TMEM allocator + the single MMA-issuing thread, 9 K loader (+ the Q tile),
// 10-11 V loaders.  Loaders move 16-byte granules with cp.async straight into the no-swizzle canonical UMMA layouts
// -- the paged layouts are already made of those granules:
//   Q, K (cached or new): [d-octet j][row]   K-major   core matrix = 8 rows x 16 B contiguous, SBO 128 B, LBO rows*16 B
//   V cached  [D][bs]  -> [token-octet][d]   K-major B operand of P.V (n = d, k = token)
//   V new   [token][D] -> [d-octet][token]   MN-major B operand
// Pipelines: k_full/k_empty and v_full/v_empty (2 stages each, loaders <-> MMA via tcgen05.commit), s_full, p_full,
// o_full (MMA <-> softmax).  tcgen05.mma's execute in issue order, which is what frees an S buffer (P.V of tile i
// is issued before Q.K^T of tile i+3).
#pragma once

#include <type_traits>

#include "context_attention.cuh"

namespace b200 {
namespace tc {

constexpr int kD = 128;        // head size of this kernel
constexpr int kBM = 128;       // query rows per CTA = UMMA M
constexpr int kBN = 128;       // KV tokens per tile = UMMA N of Q.K^T
// K and V ring depth.  ncu (profiles/r05_ctx_tc_stalls.txt): with two stages the V loaders spend 80 % of a tile period
// between issuing a tile and seeing it land (L2 latency under load ~2 us), i.e. the kernel was bound by the bytes it
// could keep in flight (one K + one V tile per SM); three stages is what 227 KiB of shared memory holds next to Q.
constexpr int kStages = 3;
constexpr int kTileBytes = kBN * kD * 2;  // 32 KiB: one Q, K or V tile
constexpr int kSoftmaxThreads = 256, kKLoaderThreads = 64, kVLoaderThreads = 128;
constexpr int kMmaWarp = kSoftmaxThreads / 32;
// 16 warps = 4 warpgroups: 0-1 softmax, 2 = {MMA warp, 2 K-loader warps, 1 spare}, 3 = 4 V-loader warps.  The kernel is
// launched with 128 registers per thread; setmaxnreg moves registers from the loader groups to the softmax groups.
constexpr int kThreads = 512;
constexpr int kRegsSoftmax = 168, kRegsOther = 88;  // (2*168 + 2*88) * 128 threads = 64 Ki registers
constexpr int kSmemBytes = (1 + 2 * kStages) * kTileBytes + 256 + 2 * 2 * 128 * 4;

// ---- tcgen05 wrappers --------------------------------------------------------------------------
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void fence_proxy_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }

__device__ __forceinline__ void tmem_alloc_512(uint32_t* smem_slot) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], 512;" ::"r"(smem_u32(smem_slot)) : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc_512(uint32_t taddr) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, 512;" ::"r"(taddr) : "memory");
}
// arrives on the mbarrier once every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void tc_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
// D[tmem] (+)= A[smem] . B[smem]
__device__ __forceinline__ void mma_ss(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// D[tmem] (+)= A[tmem] . B[smem]
__device__ __forceinline__ void mma_ts(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// Warp-uniform issue: every lane of the (converged) MMA warp executes the call, elect.sync picks the one lane that issues.
// Issued from inside an `if (lane == 0)` region the operands count as per-thread values and ptxas wraps every
// tcgen05.mma in R2UR moves and an elect loop (~18 instructions per MMA, the issuing thread became the bottleneck of
// the 256-row kernel); from uniform code the descriptors stay in uniform registers.
__device__ __forceinline__ void tc_commit_u(uint64_t* bar) {
  asm volatile(
      "{\n\t.reg .pred q;\n\telect.sync _|q, 0xffffffff;\n\t"
      "@q tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];\n\t}"
      ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mma_ss_u(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\telect.sync _|q, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void mma_ts_u(uint32_t d_tmem, uint32_t a_tmem, uint64_t b_desc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p, q;\n\telect.sync _|q, 0xffffffff;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "@q tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}"
      ::"r"(d_tmem), "r"(a_tmem), "l"(b_desc), "r"(idesc), "r"(accumulate)
      : "memory");
}

// shared-memory matrix descriptor, no swizzle (layout type 0), sm_100 descriptor version 1
// (bit layout: cute/arch/mma_sm100_desc.hpp SmemDescriptor): start>>4 [0,14), LBO>>4 [16,30), SBO>>4 [32,46)
__device__ __forceinline__ uint64_t smem_desc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  return (uint64_t)((saddr >> 4) & 0x3fffu) | ((uint64_t)((lbo_bytes >> 4) & 0x3fffu) << 16) |
         ((uint64_t)((sbo_bytes >> 4) & 0x3fffu) << 32) | (1ull << 46);
}
// instruction descriptor for kind::f16 (InstrDescriptor in the same header): fp32 accumulate, M=128
template <typename MT>
__device__ __forceinline__ constexpr uint32_t instr_desc(int n, bool b_mn_major) {
  const uint32_t fmt = sizeof(MT) == 2 && std::is_same<MT, __nv_bfloat16>::value ? 1u : 0u;  // 0 = f16, 1 = bf16
  return (1u << 4) | (fmt << 7) | (fmt << 10) | ((b_mn_major ? 1u : 0u) << 16) | ((uint32_t)(n >> 3) << 17) |
         ((uint32_t)(kBM >> 4) << 24);
}

// 32 consecutive fp32 columns of this thread's TMEM lane
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
}
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t* r) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};"
      ::"r"(taddr), "r"(r[0]), "r"(r[1]), "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7]), "r"(r[8]),
      "r"(r[9]), "r"(r[10]), "r"(r[11]), "r"(r[12]), "r"(r[13]), "r"(r[14]), "r"(r[15])
      : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

// cp.async.wait_group with a run-time bound (0 .. kStages - 1 groups may stay in flight)
__device__ __forceinline__ void cp_async_wait_upto(int n) {
  if (n <= 0) cp_async_wait<0>();
  else if (n == 1) cp_async_wait<1>();
  else if (n == 2) cp_async_wait<2>();
  else cp_async_wait<3>();
}
static_assert(kStages <= 4, "cp_async_wait_upto covers up to three groups in flight");

struct Barriers {
  uint64_t q_full;
  uint64_t k_full[kStages], k_empty[kStages], v_full[kStages], v_empty[kStages];
  uint64_t s_full[3], p_full[3], o_full[4];
  uint32_t tmem_base;
};
struct Exchange {        // between the two threads of a query row
  float m[2][2][128];    // [tile parity][column half][row] tile maximum; after the last tile the free parity slot carries
};                       // the partial denominators

// granule index -> (row, d-octet) for sources laid out [row][D]: a warp covers 8 rows x 4 octets, i.e. 64-byte
// global segments and four conflict-free 128-byte shared-memory lines
__device__ __forceinline__ void row_major_granule(int e, int& row, int& j) {
  row = ((e >> 5) & 15) * 8 + (e & 7);
  j = (e >> 9) * 4 + ((e >> 3) & 3);
}

// ---- loaders: 16-byte granules, global -> canonical UMMA layouts in shared memory ------------------------------
// Everything that depends on the lane is hoisted; per granule the loops issue one cp.async plus a pointer add.

// [row][D] source (new K, new V, Q) -> [d-octet][row] granules.  A warp instruction covers 8 rows x 4 octets:
// 64-byte global segments and four conflict-free 128-byte shared-memory lines.
template <typename MT, int NT, bool ZERO_FILL, int ROWS = kBN, int D = kD>  // ROWS = rows of the tile, D = head size
__device__ __forceinline__ void load_rows(uint4* tile, const MT* src, int64_t row_stride, int rows_valid, int lt) {
  constexpr int JW = NT / 8;  // d-octets covered per pass
  static_assert(JW <= D / 8 && (D / 8) % JW == 0, "loader threads vs head size");
  const int row_lo = lt & 7, j_lo = lt >> 3;
  const MT* s0 = src + (int64_t)row_lo * row_stride + j_lo * 8;
  uint4* d0 = tile + j_lo * ROWS + row_lo;
  const int64_t step = 8 * row_stride;
  if (rows_valid >= ROWS) {
#pragma unroll
    for (int jh = 0; jh < (D / 8) / JW; ++jh) {
      const MT* sp = s0 + jh * JW * 8;
#pragma unroll
      for (int rh = 0; rh < ROWS / 8; ++rh, sp += step) cp_async16(d0 + jh * JW * ROWS + rh * 8, sp);
    }
  } else {
#pragma unroll
    for (int jh = 0; jh < (D / 8) / JW; ++jh) {
      const MT* sp = s0 + jh * JW * 8;
#pragma unroll 4
      for (int rh = 0; rh < ROWS / 8; ++rh, sp += step) {
        if (rh * 8 + row_lo < rows_valid) cp_async16(d0 + jh * JW * ROWS + rh * 8, sp);
        else if (ZERO_FILL) d0[jh * JW * ROWS + rh * 8] = make_uint4(0u, 0u, 0u, 0u);
      }
    }
  }
}

// The physical number of a cached-prefix tile's c-th block, held by lane c (0 in the other lanes and past the context).
// Loaded separately from the copies so that a loader can fetch it BEFORE it waits for its stage to drain.
template <int BS, int TN>
__device__ __forceinline__ int ctx_tile_blocks(const int32_t* bt, int toks_valid, int lane) {
  return (lane < TN / BS && lane * BS < toks_valid) ? bt[lane] : 0;
}

// (Tried, r05: cached K by bulk copies -- one cp.async.bulk of BS * 16 bytes per (block, d-octet), completing k_full by
// itself.  128 copies of 256 bytes per tile took the two loader warps ~4 000 cycles to ISSUE, 1 000 more than the 2 048
// 16-byte cp.async they replace, and the 2k+2k shape went from 712 to 665 TFLOP/s.  Removed.)

// cached K: chunks [D/8][BS][8] of the blocks bt[0..] -> [d-octet][token] granules
template <typename MT, int BS, int NT, int TN, int D>
__device__ __forceinline__ void load_k_ctx_bn(uint4* tile, const MT* kc, int64_t block_stride, int my_bn, int toks_valid,
                                              int lt);
template <typename MT, int BS, int NT, int TN = kBN, int D = kD>  // TN = tokens of the tile
__device__ __forceinline__ void load_k_ctx(uint4* tile, const MT* kc, int64_t block_stride, const int32_t* bt,
                                           int toks_valid, int lt, int lane) {
  constexpr int JW = NT / BS;  // d-octets covered per pass
  static_assert(JW <= D / 8 && (D / 8) % JW == 0, "loader threads vs head size");
  // lane c holds the physical number of the tile's c-th block; the loop gets it by shuffle, so no load sits in
  // front of a cp.async
  load_k_ctx_bn<MT, BS, NT, TN, D>(tile, kc, block_stride, ctx_tile_blocks<BS, TN>(bt, toks_valid, lane), toks_valid, lt);
}
template <typename MT, int BS, int NT, int TN, int D>
__device__ __forceinline__ void load_k_ctx_bn(uint4* tile, const MT* kc, int64_t block_stride, int my_bn, int toks_valid,
                                              int lt) {
  constexpr int JW = NT / BS;  // d-octets covered per pass
  const int tt = lt % BS, jl = lt / BS;
#pragma unroll
  for (int c = 0; c < TN / BS; ++c) {
    const int64_t bn = __shfl_sync(0xffffffffu, my_bn, c);
    if (c * BS + tt < toks_valid) {
      const MT* sp = kc + bn * block_stride + (jl * BS + tt) * 8;
      uint4* dp = tile + jl * TN + c * BS + tt;
#pragma unroll
      for (int q = 0; q < (D / 8) / JW; ++q) cp_async16(dp + q * JW * TN, sp + q * JW * BS * 8);
    }
  }
}

// cached V: chunks [D][BS] -> [token-octet][d] granules; octets past the context end are zero-filled
template <typename MT, int BS, int NT, int TN, int D>
__device__ __forceinline__ void load_v_ctx_bn(uint4* tile, const MT* vc, int64_t block_stride, int my_bn, int toks_valid,
                                              int lt);
template <typename MT, int BS, int NT, int TN = kBN, int D = kD>
__device__ __forceinline__ void load_v_ctx(uint4* tile, const MT* vc, int64_t block_stride, const int32_t* bt,
                                           int toks_valid, int lt, int lane) {
  constexpr int HG = BS / 8;   // token octets per block
  constexpr int DW = NT / HG;  // d rows covered per pass
  static_assert(DW <= D && D % DW == 0, "loader threads vs head size");
  load_v_ctx_bn<MT, BS, NT, TN, D>(tile, vc, block_stride, ctx_tile_blocks<BS, TN>(bt, toks_valid, lane), toks_valid, lt);
}
template <typename MT, int BS, int NT, int TN, int D>
__device__ __forceinline__ void load_v_ctx_bn(uint4* tile, const MT* vc, int64_t block_stride, int my_bn, int toks_valid,
                                              int lt) {
  constexpr int HG = BS / 8;   // token octets per block
  constexpr int DW = NT / HG;  // d rows covered per pass
  const int hh = lt % HG, dl = lt / HG;
#pragma unroll
  for (int c = 0; c < TN / BS; ++c) {
    const int64_t bn = __shfl_sync(0xffffffffu, my_bn, c);
    const int tg = c * HG + hh;
    uint4* dp = tile + tg * D + dl;
    if (tg * 8 < toks_valid) {
      const MT* sp = vc + bn * block_stride + dl * BS + hh * 8;
#pragma unroll
      for (int q = 0; q < D / DW; ++q) cp_async16(dp + q * DW, sp + q * DW * BS);
    } else {
#pragma unroll
      for (int q = 0; q < D / DW; ++q) dp[q * DW] = make_uint4(0u, 0u, 0u, 0u);
    }
  }
}

template <typename MT, int BS>
__global__ void __launch_bounds__(kThreads, 1) context_attn_tc_kernel(const CtxAttnParams p) {
  constexpr int G = kD / 8;    // 16-byte granules per row
  constexpr int HG = BS / 8;   // token octets per cache block
  constexpr int TILE_UNITS = kBN * G;
  extern __shared__ __align__(1024) uint8_t tc_smem[];
  uint4* q_s = reinterpret_cast<uint4*>(tc_smem);
  uint4* k_s = q_s + TILE_UNITS;
  uint4* v_s = k_s + kStages * TILE_UNITS;
  Barriers* bar = reinterpret_cast<Barriers*>(v_s + kStages * TILE_UNITS);
  Exchange* xch = reinterpret_cast<Exchange*>(reinterpret_cast<uint8_t*>(bar) + 256);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  int mblk, head, seq;
  ctx_cta_coords(p.pair_group, mblk, head, seq);
  const int ctx_len = p.b_ctx_len[seq];
  const int q_len = p.b_seq_len[seq] - ctx_len;
  const int m0 = mblk * kBM;
  if (m0 >= q_len) return;
  const int64_t start = p.b_start_loc[seq];
  const int kvh = head / (p.num_heads / p.num_kv_heads);
  const int n_ctx_tiles = (ctx_len + kBN - 1) / kBN;
  const int new_end = min(q_len, m0 + kBM);
  const int n_tiles = n_ctx_tiles + (new_end + kBN - 1) / kBN;

  if (tid == 0) {
    mbar_init(&bar->q_full, kKLoaderThreads);
    for (int s = 0; s < kStages; ++s) {
      mbar_init(&bar->k_full[s], kKLoaderThreads);
      mbar_init(&bar->v_full[s], kVLoaderThreads);
      mbar_init(&bar->k_empty[s], 1);
      mbar_init(&bar->v_empty[s], 1);
    }
    for (int s = 0; s < 3; ++s) {
      mbar_init(&bar->s_full[s], 1);
      mbar_init(&bar->p_full[s], kSoftmaxThreads);
    }
    for (int s = 0; s < 4; ++s) mbar_init(&bar->o_full[s], 1);
    mbar_fence_init();
  }
  if (warp == kMmaWarp) tmem_alloc_512(&bar->tmem_base);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem = bar->tmem_base;

  if (warp < 8) {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(kRegsSoftmax));
    // ================= softmax + epilogue: two threads per query row, 64 columns of every tile each =================
    // warps 0-3 take columns [0,64), warps 4-7 columns [64,128) of the same 128 rows (a warp reaches the TMEM lanes
    // 32*(warp%4)..+31).  The two halves of a row agree on the reference point through shared memory once per tile.
    const int half = warp >> 2, row = tid & 127;
    const uint32_t lane_off = (uint32_t)(32 * (warp & 3)) << 16;
    const int qrow = m0 + row;
    const bool has_alibi = p.alibi_slopes != nullptr;
    const float slope2 = has_alibi ? p.alibi_slopes[head] * 1.4426950408889634f : 0.f;
    const uint32_t o_addr = tmem + lane_off + 384 + 64 * half;
    // m_used: the reference point of P and O.  It follows the running maximum lazily: it moves only when the
    // maximum has grown by more than 2^8, so that the O accumulator in TMEM is rescaled rarely (exact either way:
    // numerator and denominator carry the same reference point).
    float m_used = -INFINITY, l_run = 0.f;
    // Completion of P_t.V_t is signalled on o_full[t % 4].  The softmax threads need it only when they rescale O
    // (rare) and at the very end, but an mbarrier phase must be observed before the one after it on the same barrier
    // can complete -- so tile t is observed, in order, at the latest before these threads release tile t + 4
    // (by then it completed long ago: no stall), or right away when a rescale needs O.
    int o_next = 0;  // first tile whose P.V completion these threads have not observed yet
    auto observe_o_upto = [&](int last) {  // tiles o_next .. last
      for (; o_next <= last; ++o_next) mbar_wait(&bar->o_full[o_next & 3], (o_next >> 2) & 1);
    };

    for (int i = 0; i < n_tiles; ++i) {
      const int sb = i % 3, st = i & 1;  // S/P buffer, exchange-slot parity
      const bool is_ctx = i < n_ctx_tiles;
      const int n0 = (is_ctx ? i : i - n_ctx_tiles) * kBN + 64 * half;  // first column of this thread's half
      // columns [0, limit) of this half are visible to this row
      const int limit = is_ctx ? ctx_len - n0 : qrow - n0 + 1;
      // key position minus query position of this half's column 0 (ALiBi)
      const float bias0 = slope2 * (float)(is_ctx ? n0 - ctx_len - qrow : n0 - qrow);
      mbar_wait(&bar->s_full[sb], (i / 3) & 1);
      tc_fence_after();
      const uint32_t s_addr = tmem + lane_off + sb * 128;
      float s[64];
      tmem_ld32(s_addr + 64 * half, reinterpret_cast<uint32_t*>(&s[0]));
      tmem_ld32(s_addr + 64 * half + 32, reinterpret_cast<uint32_t*>(&s[32]));
      tmem_ld_wait();
      float mx4[4] = {-INFINITY, -INFINITY, -INFINITY, -INFINITY};
      // fast path (warp-uniform): every column of this half is visible to every row of the warp and there is no
      // bias -> the maximum is taken on the raw scores and the scale is folded into the exponent's FMA
      const bool plain = !has_alibi && __all_sync(0xffffffffu, limit >= 64);
      float mx, mul;
      if (plain) {
#pragma unroll
        for (int col = 0; col < 64; ++col) mx4[col & 3] = fmaxf(mx4[col & 3], s[col]);
        mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3])) * p.scale_log2;  // scale > 0
        mul = p.scale_log2;
      } else {
#pragma unroll
        for (int col = 0; col < 64; ++col) {
          float v = fmaf(s[col], p.scale_log2, has_alibi ? fmaf(slope2, (float)col, bias0) : 0.f);
          v = col < limit ? v : -INFINITY;
          s[col] = v;
          mx4[col & 3] = fmaxf(mx4[col & 3], v);
        }
        mx = fmaxf(fmaxf(mx4[0], mx4[1]), fmaxf(mx4[2], mx4[3]));
        mul = 1.f;
      }
      // the row's other half: exchange maxima (double-buffered by tile parity; one barrier per tile).  After the
      // barrier both halves have S in registers, so P may overwrite the S columns.
      xch->m[st][half][row] = mx;
      asm volatile("bar.sync %0, 64;" ::"r"(2 + (warp & 3)) : "memory");  // the two warps that share these 32 rows
      mx = fmaxf(mx, xch->m[st][half ^ 1][row]);
      // O is touched by these threads only when the reference point moves: then P_{i-1}.V_{i-1} must have completed
      // (observe_o_upto); otherwise its completion is observed up to four tiles late (ncu, r05: waiting for it at the
      // end of every tile was 16 % of the softmax warps' time).
      if (__any_sync(0xffffffffu, mx > m_used + 8.f)) {
        const float m_new = fmaxf(m_used, mx);
        const float alpha = ex2_approx(m_used - m_new);
        if (i > 0) {
          observe_o_upto(i - 1);
          tc_fence_after();
#pragma unroll
          for (int c = 0; c < 2; ++c) {
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
      float sum4[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int c = 0; c < 2; ++c) {
        uint32_t pk[16];
#pragma unroll
        for (int k = 0; k < 32; k += 2) {
          const float e0 = ex2_approx(fmaf(s[32 * c + k], mul, -m_used));   // masked columns are -inf -> 0
          const float e1 = ex2_approx(fmaf(s[32 * c + k + 1], mul, -m_used));
          sum4[(k >> 1) & 3] += e0 + e1;
          pk[k >> 1] = pack2<MT>(e0, e1);
        }
        tmem_st16(s_addr + 32 * half + 16 * c, pk);  // P (2 elements per column) over the consumed S columns
      }
      tmem_st_wait();
      observe_o_upto(i - 4);
      tc_fence_before();
      mbar_arrive(&bar->p_full[sb]);
      l_run += (sum4[0] + sum4[1]) + (sum4[2] + sum4[3]);
    }

    xch->m[n_tiles & 1][half][row] = l_run;  // the slot the last tile did not use
    asm volatile("bar.sync %0, 64;" ::"r"(2 + (warp & 3)) : "memory");
    const float inv = 1.f / (l_run + xch->m[n_tiles & 1][half ^ 1][row]);
    observe_o_upto(n_tiles - 1);
    tc_fence_after();
    MT* orow = static_cast<MT*>(p.o) + (start + qrow) * p.o_ts + (int64_t)head * p.o_hs + 64 * half;
#pragma unroll
    for (int c = 0; c < 2; ++c) {
      uint32_t r[32];
      tmem_ld32(o_addr + 32 * c, r);
      tmem_ld_wait();
      if (qrow < q_len) {
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
  } else if (warp < 12) {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
   if (warp == kMmaWarp) {
    // ================= the MMA warp (uniform: every lane runs the loop, elect.sync issues) =================
    {
      const uint32_t idesc_k = instr_desc<MT>(kBN, false);   // Q.K^T and P.V over cached V: K-major B
      const uint32_t idesc_mn = instr_desc<MT>(kD, true);    // P.V over new V: MN-major B
      // descriptors are built once and advanced by integer adds on the start-address field (address >> 4)
      const uint64_t qd = smem_desc(smem_u32(q_s), kBM * 16, 128);
      const uint64_t kd0 = smem_desc(smem_u32(k_s), kBN * 16, 128);
      const uint64_t vd0_ctx = smem_desc(smem_u32(v_s), kD * 16, 128);   // cached V: K-major [token-octet][d]
      const uint64_t vd0_new = smem_desc(smem_u32(v_s), 128, kBN * 16);  // new V: MN-major [d-octet][token]
      constexpr uint64_t kStageStep = kTileBytes >> 4;
      mbar_wait(&bar->q_full, 0);
      auto issue_qk = [&](int i) {  // S_i = Q.K_i^T into S buffer i%3
        const int st = i % kStages;
        mbar_wait(&bar->k_full[st], (i / kStages) & 1);
        tc_fence_after();
        const uint64_t kd = kd0 + (uint64_t)st * kStageStep;
#pragma unroll
        for (int ks = 0; ks < kD / 16; ++ks)
          mma_ss_u(tmem + (i % 3) * 128, qd + (uint64_t)(ks * ((2 * kBM * 16) >> 4)),
                   kd + (uint64_t)(ks * ((2 * kBN * 16) >> 4)), idesc_k, ks > 0);
        tc_commit_u(&bar->k_empty[st]);
        tc_commit_u(&bar->s_full[i % 3]);
      };
      // Q.K^T runs three tiles ahead of P.V (three S buffers; a buffer is free again, in pipe order, as soon as the
      // P.V that reads it has been issued): S_{j+1} and S_{j+2} are ready when the softmax warps finish tile j
      issue_qk(0);
      if (n_tiles > 1) issue_qk(1);
      if (n_tiles > 2) issue_qk(2);
      for (int j = 0; j < n_tiles; ++j) {
        {  // O += P_j.V_j (first: it is on the softmax warps' critical path)
          const int st = j % kStages, sb = j % 3;
          mbar_wait(&bar->v_full[st], (j / kStages) & 1);
          mbar_wait(&bar->p_full[sb], (j / 3) & 1);
          tc_fence_after();
          const bool is_ctx = j < n_ctx_tiles;
          // cached V: [token-octet][d] granules (K-major, LBO = next token octet, SBO = next 8 d rows), k-step = two
          // token octets; new V: [d-octet][token] granules (MN-major, LBO = next 8 tokens, SBO = next d octet), k-step =
          // 16 tokens
          const uint64_t vd = (is_ctx ? vd0_ctx : vd0_new) + (uint64_t)st * kStageStep;
          const uint64_t step = is_ctx ? (uint64_t)((2 * kD * 16) >> 4) : (uint64_t)((16 * 16) >> 4);
          const uint32_t idesc = is_ctx ? idesc_k : idesc_mn;
#pragma unroll
          for (int kt = 0; kt < kBN / 16; ++kt)
            mma_ts_u(tmem + 384, tmem + sb * 128 + kt * 8, vd + kt * step, idesc, (j | kt) > 0);
          tc_commit_u(&bar->v_empty[st]);
          tc_commit_u(&bar->o_full[j & 3]);
        }
        // S buffer j % 3 is free again in pipe order (P_j.V_j was just issued and tcgen05.mma's execute in order)
        if (j + 3 < n_tiles) issue_qk(j + 3);
      }
    }
    __syncwarp();
   } else if (warp < kMmaWarp + 3) {
    // ================= K (and Q) loaders =================
    const int lt = tid - (kMmaWarp + 1) * 32;
    const MT* kc = static_cast<const MT*>(p.k_cache) + (int64_t)kvh * p.kc_hs;
    const MT* kn = static_cast<const MT*>(p.k) + start * p.k_ts + (int64_t)kvh * p.k_hs;
    const int32_t* bt = p.b_loc + (int64_t)seq * p.b_loc_stride;
    auto issue_k = [&](int ti, int st) {
      uint4* Ks = k_s + st * TILE_UNITS;
      if (ti < n_ctx_tiles) {
        const int n0 = ti * kBN;
        load_k_ctx<MT, BS, kKLoaderThreads>(Ks, kc, p.kc_bs, bt + n0 / BS, ctx_len - n0, lt, lane);
      } else {
        const int n0 = (ti - n_ctx_tiles) * kBN;
        load_rows<MT, kKLoaderThreads, false>(Ks, kn + (int64_t)n0 * p.k_ts, p.k_ts, q_len - n0, lt);
      }
    };
    load_rows<MT, kKLoaderThreads, true>(q_s, static_cast<const MT*>(p.q) + (start + m0) * p.q_ts + (int64_t)head * p.q_hs,
                                         p.q_ts, q_len - m0, lt);
    cp_async_commit();
    issue_k(0, 0);
    cp_async_commit();
    cp_async_wait<1>();
    fence_proxy_async_smem();
    mbar_arrive(&bar->q_full);
    for (int t = 1; t < kStages && t < n_tiles; ++t) {
      issue_k(t, t);
      cp_async_commit();
    }
    // A landed tile is announced AHEAD of the refill that follows it in program order -- or the refill goes first when
    // its stage has already drained -- see the loaders of context_attention_tc256.cuh (r05 trace: the announcement of a
    // tile that had long landed sat behind the wait for the previous stage plus the issue of the next tile's copies).
    int issued = min(kStages, n_tiles);  // tile groups committed so far
    auto announce_k = [&](int t) {
      fence_proxy_async_smem();
      mbar_arrive(&bar->k_full[t % kStages]);
    };
    cp_async_wait_upto(issued - 1);
    announce_k(0);
    for (int i = 0; i < n_tiles; ++i) {
      const bool more = i + kStages < n_tiles;
      bool refilled = false;
      if (more && __all_sync(0xffffffffu, mbar_test(&bar->k_empty[i % kStages], (i / kStages) & 1))) {
        issue_k(i + kStages, i % kStages);
        cp_async_commit();
        ++issued;
        refilled = true;
      }
      if (i + 1 < n_tiles) {
        cp_async_wait_upto(issued - (i + 2));  // groups younger than tile i + 1
        announce_k(i + 1);
      }
      if (more && !refilled) {
        mbar_wait(&bar->k_empty[i % kStages], (i / kStages) & 1);
        issue_k(i + kStages, i % kStages);
        cp_async_commit();
        ++issued;
      }
    }
   }
  } else {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(kRegsOther));
    // ================= V loaders =================
    const int lt = tid - 12 * 32;
    const MT* vc = static_cast<const MT*>(p.v_cache) + (int64_t)kvh * p.vc_hs;
    const MT* vn = static_cast<const MT*>(p.v) + start * p.v_ts + (int64_t)kvh * p.v_hs;
    const int32_t* bt = p.b_loc + (int64_t)seq * p.b_loc_stride;
    auto issue_v = [&](int ti, int st) {
      uint4* Vs = v_s + st * TILE_UNITS;
      if (ti < n_ctx_tiles) {
        const int n0 = ti * kBN;
        load_v_ctx<MT, BS, kVLoaderThreads>(Vs, vc, p.vc_bs, bt + n0 / BS, ctx_len - n0, lt, lane);
      } else {
        const int n0 = (ti - n_ctx_tiles) * kBN;
        load_rows<MT, kVLoaderThreads, true>(Vs, vn + (int64_t)n0 * p.v_ts, p.v_ts, q_len - n0, lt);
      }
    };
    for (int t = 0; t < kStages && t < n_tiles; ++t) {
      issue_v(t, t);
      cp_async_commit();
    }
    auto announce_v = [&](int t) {  // this thread's copies of tile t have landed
      if (t < n_ctx_tiles && t * kBN + kBN > ctx_len && ((ctx_len - t * kBN) & 7)) {
        // the token octet that straddles ctx_len: slots past the end are stale cache memory (may be NaN / Inf)
        asm volatile("bar.sync 1, %0;" ::"n"(kVLoaderThreads) : "memory");  // every V loader's copies of tile t landed
        const int tgb = (ctx_len - t * kBN) >> 3, rem = (ctx_len - t * kBN) & 7;
        for (int d = lt; d < kD; d += kVLoaderThreads) {
          MT* gp = reinterpret_cast<MT*>(&v_s[(t % kStages) * TILE_UNITS + tgb * kD + d]);
          for (int x = rem; x < 8; ++x) gp[x] = MT(0.f);
        }
      }
      fence_proxy_async_smem();
      mbar_arrive(&bar->v_full[t % kStages]);
    };
    int issued = min(kStages, n_tiles);
    if (n_tiles > 0) {
      cp_async_wait_upto(issued - 1);
      announce_v(0);
    }
    for (int i = 0; i < n_tiles; ++i) {  // the K loader's order
      const bool more = i + kStages < n_tiles;
      bool refilled = false;
      if (more && __all_sync(0xffffffffu, mbar_test(&bar->v_empty[i % kStages], (i / kStages) & 1))) {
        issue_v(i + kStages, i % kStages);
        cp_async_commit();
        ++issued;
        refilled = true;
      }
      if (i + 1 < n_tiles) {
        cp_async_wait_upto(issued - (i + 2));
        announce_v(i + 1);
      }
      if (more && !refilled) {
        mbar_wait(&bar->v_empty[i % kStages], (i / kStages) & 1);
        issue_v(i + kStages, i % kStages);
        cp_async_commit();
        ++issued;
      }
    }
  }

  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  if (warp == kMmaWarp) tmem_dealloc_512(tmem);
}

template <typename MT, int BS>
int launch_ctx_attn_tc(const CtxAttnParams& p, int batch, int max_input_len, cudaStream_t stream) {
  static bool attr_set[64] = {false};
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  dev &= 63;
  if (!attr_set[dev]) {
    B200_CUDA_TRY(cudaFuncSetAttribute(context_attn_tc_kernel<MT, BS>, cudaFuncAttributeMaxDynamicSharedMemorySize, kSmemBytes));
    attr_set[dev] = true;
  }
  dim3 grid((max_input_len + kBM - 1) / kBM, p.num_heads, batch);
  context_attn_tc_kernel<MT, BS><<<grid, kThreads, kSmemBytes, stream>>>(p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace tc
}  // namespace b200
