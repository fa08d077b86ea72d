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
