// Cache writes and block moves for the paged KV cache (sm_100a).
//   b200_reshape_and_cache  <- reference csrc/cache_kernels.cu:154-268
//   b200_copy_blocks        <- reference csrc/cache_kernels.cu:65-150
//   b200_swap_blocks        <- reference csrc/cache_kernels.cu:21-60
//   b200_convert_fp8_e5m2   <- reference csrc/cache_kernels.cu:272-320
// All four are HBM-bound byte movers: the design rules are coalesced 16-byte accesses,
// tables carried as kernel parameters (no H2D copy, no sync) and grids sized from the work.
#include <algorithm>
#include <stdlib.h>
#include <atomic>
#include <cstring>
#include <mutex>
#include <string>
#include <type_traits>

#include "common.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------
// error plumbing (shared by all translation units)
// ------------------------------------------------------------------------------------------
static thread_local char g_last_cuda_error[512] = "";

void set_last_cuda_error(cudaError_t e, const char* where) {
  snprintf(g_last_cuda_error, sizeof(g_last_cuda_error), "%s: %s (%s)", where,
           cudaGetErrorName(e), cudaGetErrorString(e));
}

thread_local bool tl_pdl_launch = false;
bool pdl_disabled() {
  static const bool off = getenv("B200_NO_PDL") != nullptr;
  return off;
}

static std::atomic<unsigned long long> g_launches{0};
void count_launch() { g_launches.fetch_add(1, std::memory_order_relaxed); }

int num_sms(int device) {
  static int cached[64];
  static std::once_flag once;
  std::call_once(once, [] { std::fill(cached, cached + 64, 0); });
  int dev = device;
  if (dev < 0 && cudaGetDevice(&dev) != cudaSuccess) return 148;
  if (dev >= 64) dev = 63;
  if (cached[dev] == 0) {
    int v = 0;
    if (cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess || v <= 0)
      v = 148;
    cached[dev] = v;
  }
  return cached[dev];
}

// ------------------------------------------------------------------------------------------
// reshape_and_cache
// ------------------------------------------------------------------------------------------
// N source elements starting at p: 16-byte vector loads when the host proved alignment (VEC), else scalar.
template <typename T, int N, bool VEC>
__device__ __forceinline__ void load_elems(const T* __restrict__ p, T* out /* 16-byte aligned */) {
  if constexpr (VEC) {
    constexpr int PER = 16 / (int)sizeof(T);
    static_assert(N % PER == 0, "whole vectors");
#pragma unroll
    for (int i = 0; i < N / PER; ++i)
      *reinterpret_cast<uint4*>(out + i * PER) = ld_nc_v4(reinterpret_cast<const uint4*>(p) + i);
  } else {
#pragma unroll
    for (int i = 0; i < N; ++i) out[i] = p[i];
  }
}

template <typename T, typename CT, bool FP8>
__device__ __forceinline__ CT cache_cvt(T v) {
  if constexpr (FP8) return to_e5m2<T>(v);
  else return v;
}
// N (even) source elements -> N cache elements; fp8 goes two elements per conversion instruction
template <typename T, typename CT, bool FP8, int N>
__device__ __forceinline__ void cache_cvt_vec(const T* src, CT* dst /* 2-byte aligned */) {
  if constexpr (FP8) {
    static_assert(N % 2 == 0, "pairs");
#pragma unroll
    for (int e = 0; e < N; e += 2) *reinterpret_cast<uint16_t*>(dst + e) = to_e5m2_x2<T>(src[e], src[e + 1]);
  } else {
#pragma unroll
    for (int e = 0; e < N; ++e) dst[e] = src[e];
  }
}

// The same for source elements that sit PACKED in registers (16-byte vectors read from global or shared memory): a bf16
// pair widens with one shift and one mask, no per-element extraction.
template <typename T, typename CT, bool FP8, int N>
__device__ __forceinline__ void cache_cvt_vec_packed(const T* src /* 4-byte aligned */, CT* dst) {
  if constexpr (FP8 && std::is_same<T, __nv_bfloat16>::value) {
#pragma unroll
    for (int e = 0; e < N; e += 2) {
      const uint32_t w = *reinterpret_cast<const uint32_t*>(src + e);
      *reinterpret_cast<uint16_t*>(dst + e) = (uint16_t)__nv_cvt_float2_to_fp8x2(
          make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u)), __NV_SATFINITE, __NV_E5M2);
    }
  } else {
    cache_cvt_vec<T, CT, FP8, N>(src, dst);
  }
}

// Token-major kernel (decode-sized calls): one CTA per token.  K: every (head, d/x chunk) is x contiguous
// source elements and x contiguous destination elements (16 bytes in the cache), so it moves as one
// vector.  V: element (h, d) lands at [blk, h, d, off] -- a scatter with stride block_size that one token
// cannot avoid; the source side is read in 16-byte vectors.
template <typename T, typename CT, bool FP8, int X, bool VEC>
__global__ void __launch_bounds__(512) reshape_and_cache_token_kernel(
    const T* __restrict__ key, const T* __restrict__ value, CT* __restrict__ key_cache,
    CT* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int64_t key_stride,
    int64_t value_stride, int num_heads, int head_size, int block_size, int64_t num_blocks) {
  constexpr int PER = 16 / (int)sizeof(T);
  griddep_launch_dependents();  // the attention kernel that follows may set itself up now
  griddep_wait();               // (launched as a programmatic dependent inside b200_decode_step)
  const int64_t token = blockIdx.x;
  const int n = num_heads * head_size;
  const int n_kvec = n / X;    // 16-byte cache vectors of the K row
  const int n_vvec = n / PER;  // 16-byte source vectors of the V row
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  // The slot and the source vectors are independent loads: all of them are issued before anything is
  // consumed, so a decode-sized call costs one global-memory round trip, not three in a row.
  const int64_t slot = slot_mapping[token];
  for (int base = 0; base < n_vvec; base += blockDim.x) {  // n_vvec >= n_kvec (PER <= X)
    const int i = base + threadIdx.x;
    alignas(16) T ks[X];
    alignas(16) T vs[PER];
    const bool has_k = i < n_kvec, has_v = i < n_vvec;
    // (h * chunks + c) * X == i * X: heads are dense in the source row and in the cache block
    if (has_k) load_elems<T, X, VEC>(ksrc + (int64_t)i * X, ks);
    if (has_v) load_elems<T, PER, VEC>(vsrc + (int64_t)i * PER, vs);
    if (slot < 0) continue;  // padding (reference cache_kernels.cu:169-172)
    const int64_t blk = slot / block_size;
    const int off = static_cast<int>(slot - blk * block_size);
    if (blk >= num_blocks) continue;  // out-of-range slot: never write outside the cache
    CT* kdst = key_cache + blk * (int64_t)n * block_size;
    CT* vdst = value_cache + blk * (int64_t)n * block_size;
    if (has_k) {
      alignas(16) CT tmp[X];
#pragma unroll
      for (int e = 0; e < X; ++e) tmp[e] = cache_cvt<T, CT, FP8>(ks[e]);
      *reinterpret_cast<uint4*>(kdst + ((int64_t)i * block_size + off) * X) = *reinterpret_cast<const uint4*>(tmp);
    }
    if (has_v) {
#pragma unroll
      for (int e = 0; e < PER; ++e)
        vdst[(int64_t)(i * PER + e) * block_size + off] = cache_cvt<T, CT, FP8>(vs[e]);
    }
  }
}

// Tile kernel (prefill-sized calls): one CTA per (16 consecutive tokens, head).
// K needs no staging: thread (chunk c, token t), token-fastest, reads its x source elements (16-64 bytes of
// one row; the lanes of a warp cover whole 32-byte sectors between them) and writes one 16-byte cache vector;
// consecutive slots give 256 contiguous bytes per chunk.  Every thread looks its own slot up, so the K path
// has no barrier and all of its global loads are in flight from the first instruction.
// V is the transposing side: the [16, D] source tile is read with 16-byte vectors into shared memory (rows
// padded by 16 bytes); cache row d of a block holds that block's tokens contiguously, so when a group of
// G = min(block_size, 16) tile tokens occupies an aligned run of G slots of one block (the normal prefill
// layout) each thread gathers its row's G values from the shared tile and stores them as 8-byte words --
// a warp writes 32 adjacent rows, i.e. one contiguous span -- instead of the reference's isolated 1-2 byte
// stores block_size elements apart (cache_kernels.cu:188-191,204).  Any other mapping (padding, shuffled or
// unaligned slots) takes the element-wise path of the same kernel.

template <typename T, typename CT, bool FP8, int X, bool VEC, bool KSMEM, int TOK>
__global__ void __launch_bounds__(256) reshape_and_cache_tile_kernel(
    const T* __restrict__ key, const T* __restrict__ value, CT* __restrict__ key_cache,
    CT* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int64_t key_stride,
    int64_t value_stride, int num_tokens, int num_heads, int head_size, int block_size,
    int64_t num_blocks, int head_major) {
  static_assert(TOK == 16 || TOK == 32, "the slot decode runs on one warp");
  constexpr int PER = 16 / (int)sizeof(T);
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int pitch = head_size + PER;  // elements; rows stay 16-byte aligned
  T* vt = reinterpret_cast<T*>(smem_raw);
  T* kt = vt + TOK * pitch;  // KSMEM only
  __shared__ int64_t s_base[TOK];  // element offset of (blk, h, 0, 0) or -1 = skip
  __shared__ int s_off[TOK];
  __shared__ int s_fast;

  // head_major: consecutive CTAs take the heads of one token tile, so CTAs that run together read one
  // contiguous span of each source row (all heads) instead of 256-byte pieces a whole row apart.
  const int tile = head_major ? blockIdx.y : blockIdx.x;
  const int h = head_major ? blockIdx.x : blockIdx.y;
  const int t0 = tile * TOK;
  const int nt = min(TOK, num_tokens - t0);
  const int G = min(block_size, TOK);
  const int64_t head_elems = (int64_t)head_size * block_size;

  // ---- source tiles -> shared memory (issued first: the longest dependency chain) ----
  const T* vsrc = value + (int64_t)t0 * value_stride + h * head_size;
  const T* ksrc = key + (int64_t)t0 * key_stride + h * head_size;
  if constexpr (VEC) {
    const int vpr = head_size / PER;  // 16-byte vectors per row
    for (int v = threadIdx.x; v < nt * vpr; v += blockDim.x) {
      const int t = v / vpr, c = v - t * vpr;
      uint4 kv;
      if constexpr (KSMEM) kv = ld_nc_v4(reinterpret_cast<const uint4*>(ksrc + (int64_t)t * key_stride) + c);
      *reinterpret_cast<uint4*>(vt + t * pitch + c * PER) =
          ld_nc_v4(reinterpret_cast<const uint4*>(vsrc + (int64_t)t * value_stride) + c);
      if constexpr (KSMEM) *reinterpret_cast<uint4*>(kt + t * pitch + c * PER) = kv;
    }
  } else {
    for (int i = threadIdx.x; i < nt * head_size; i += blockDim.x) {
      const int t = i / head_size, d = i - t * head_size;
      vt[t * pitch + d] = vsrc[(int64_t)t * value_stride + d];
      if constexpr (KSMEM) kt[t * pitch + d] = ksrc[(int64_t)t * key_stride + d];
    }
  }

  // ---- slots of the tile (first warp) ----
  if (threadIdx.x < 32) {
    const int t = threadIdx.x;
    int64_t base = -1;
    int off = 0;
    if (t < nt) {
      const int64_t slot = slot_mapping[t0 + t];
      if (slot >= 0) {
        const int64_t blk = slot / block_size;
        off = static_cast<int>(slot - blk * block_size);
        if (blk < num_blocks) base = (blk * num_heads + h) * head_elems;
      }
    }
    if (t < TOK) {
      s_base[t] = base;
      s_off[t] = off;
    }
    // group leader = first token of this token's aligned group of G
    const int lead = t - (t % G);
    const int64_t lbase = __shfl_sync(0xffffffffu, base, lead & 31);
    const int loff = __shfl_sync(0xffffffffu, off, lead & 31);
    const bool ok = t >= TOK || (nt == TOK && base >= 0 && base == lbase && off == loff + (t - lead) && (loff % G) == 0);
    const bool all_ok = __all_sync(0xffffffffu, ok);
    // whole groups only: a block size that does not divide the tile (e.g. 24 with 32-token tiles) would leave
    // the tile's last tokens without a group, so it takes the element-wise path
    if (t == 0) s_fast = VEC && all_ok && (TOK % G == 0);
  }

  // ---- K: one 16-byte cache vector per (token, chunk), token-fastest ----
  const int chunks = head_size / X;
  if constexpr (KSMEM) __syncthreads();
  for (int v = threadIdx.x; v < nt * chunks; v += blockDim.x) {
    const int c = v / nt, t = v - c * nt;
    const int64_t slot = slot_mapping[t0 + t];
    alignas(16) T src[X];
    if constexpr (KSMEM) {
#pragma unroll
      for (int i = 0; i < X / PER; ++i)
        *reinterpret_cast<uint4*>(src + i * PER) = *reinterpret_cast<const uint4*>(kt + t * pitch + c * X + i * PER);
    } else {
      load_elems<T, X, VEC>(ksrc + (int64_t)t * key_stride + c * X, src);
    }
    if (slot < 0) continue;
    const int64_t blk = slot / block_size;
    if (blk >= num_blocks) continue;
    const int off = static_cast<int>(slot - blk * block_size);
    alignas(16) CT tmp[X];
    cache_cvt_vec<T, CT, FP8, X>(src, tmp);
    *reinterpret_cast<uint4*>(key_cache + (blk * num_heads + h) * head_elems + ((int64_t)c * block_size + off) * X) =
        *reinterpret_cast<const uint4*>(tmp);
  }
  if constexpr (!KSMEM) __syncthreads();

  // ---- V ----
  if (s_fast) {
    const int ngroups = TOK / G;
    constexpr int WPE = 8 / (int)sizeof(CT);  // cache elements per 8-byte word
    constexpr int VPE = 16 / (int)sizeof(CT); // cache elements per 16-byte vector
    if ((G % VPE) == 0 && (block_size % VPE) == 0) {
      // a group's row is whole 16-byte vectors (block_size 16/32; fp32 also 8): one st.v4 per 16 bytes, consecutive
      // lanes write consecutive rows (one contiguous span per warp); fp8 conversions go two elements per instruction.
      // (Tried, r05: RD = 4 adjacent rows per thread from one 8-byte shared-memory read per token -- fewer
      // instructions, but the lanes' stores end up 64 bytes apart and the kernel lost 8-20 %.)
      for (int i = threadIdx.x; i < head_size * ngroups; i += blockDim.x) {
        const int g = i / head_size, d = i - g * head_size;
        const int tl = g * G;
        CT* row = value_cache + s_base[tl] + (int64_t)d * block_size + s_off[tl];
        const T* col = vt + tl * pitch + d;
#pragma unroll
        for (int w = 0; w < TOK / VPE; ++w) {
          if (w * VPE < G) {
            alignas(16) T line[VPE];
#pragma unroll
            for (int e = 0; e < VPE; ++e) line[e] = col[(w * VPE + e) * pitch];
            alignas(16) CT tmp[VPE];
            cache_cvt_vec<T, CT, FP8, VPE>(line, tmp);
            *reinterpret_cast<uint4*>(row + w * VPE) = *reinterpret_cast<const uint4*>(tmp);
          }
        }
      }
    } else
    for (int i = threadIdx.x; i < head_size * ngroups; i += blockDim.x) {
      const int g = i / head_size, d = i - g * head_size;
      const int tl = g * G;
      CT* row = value_cache + s_base[tl] + (int64_t)d * block_size + s_off[tl];
      const T* col = vt + tl * pitch + d;
#pragma unroll
      for (int w = 0; w < TOK / WPE; ++w) {
        if (w * WPE < G) {
          alignas(8) CT tmp[WPE];
#pragma unroll
          for (int e = 0; e < WPE; ++e) tmp[e] = cache_cvt<T, CT, FP8>(col[(w * WPE + e) * pitch]);
          *reinterpret_cast<uint2*>(row + w * WPE) = *reinterpret_cast<const uint2*>(tmp);
        }
      }
    }
  } else {
    for (int i = threadIdx.x; i < nt * head_size; i += blockDim.x) {
      const int d = i / nt, t = i - d * nt;
      if (s_base[t] < 0) continue;
      value_cache[s_base[t] + (int64_t)d * block_size + s_off[t]] = cache_cvt<T, CT, FP8>(vt[t * pitch + d]);
    }
  }
}

// Prefill-sized calls with 16-byte aligned rows, streamed: the tile kernel above is one (token tile, head) per CTA --
// load, barrier, store, exit -- and ncu shows what that costs (r05: 69 % of the warps resident, every CTA's loads
// exposed once).  Here a resident grid walks the (token tile, head) items with a ring of STAGES shared-memory stages
// filled by cp.async: the loads of items i+1 .. i+STAGES-1 are in flight while item i is converted and stored, one
// __syncthreads per item.  Same stores, same bits as the tile kernel's vector path.

__device__ __forceinline__ void rc_cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc)
               : "memory");
}
__device__ __forceinline__ void rc_cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void rc_cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

template <typename T, typename CT, bool FP8, int X, int TOK, int STAGES>
__global__ void __launch_bounds__(256) reshape_and_cache_stream_kernel(
    const T* __restrict__ key, const T* __restrict__ value, CT* __restrict__ key_cache,
    CT* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int64_t key_stride,
    int64_t value_stride, int num_tokens, int num_heads, int head_size, int block_size,
    int64_t num_blocks, int num_tiles, int wide) {
  static_assert(TOK == 16 || TOK == 32, "the slot decode runs on one warp");
  constexpr int PER = 16 / (int)sizeof(T);
  constexpr int VPE = 16 / (int)sizeof(CT);  // cache elements per 16-byte vector
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int pitch = head_size + PER;  // elements; rows stay 16-byte aligned
  const int stage_elems = 2 * TOK * pitch;  // V tile, then K tile
  static_assert(STAGES == 2 || STAGES == 3, "ring depth");
  __shared__ int64_t s_base[STAGES][TOK];  // element offset of (blk, h, 0, 0) or -1 = skip
  __shared__ int s_off[STAGES][TOK];
  __shared__ int s_fast[STAGES];

  // items = (token tile, head), heads fastest: CTAs that run together read one contiguous span of each source row
  const int vpr = head_size / PER;  // 16-byte vectors per source row
  const int G = min(block_size, TOK);
  const int64_t head_elems = (int64_t)head_size * block_size;
  const bool v_vec = (G % VPE) == 0 && (block_size % VPE) == 0 && (TOK % G) == 0;
  const bool v_wide = wide && (G % (2 * VPE)) == 0 && (block_size % (2 * VPE)) == 0;  // rows of whole 32-byte sectors

  // cp.async of one item's V and K tiles.  256 threads cover 256 / vpr rows per pass when vpr divides 256.
  const bool even = (256 % vpr) == 0;
  const int my_t = even ? threadIdx.x / vpr : 0, my_c = even ? threadIdx.x - my_t * vpr : 0, rows_per_pass = even ? 256 / vpr : 0;
  auto issue = [&](int tile, int h, int stage) {
    const int t0 = tile * TOK;
    const int nt = min(TOK, num_tokens - t0);
    T* vt = reinterpret_cast<T*>(smem_raw) + (size_t)stage * stage_elems;
    T* kt = vt + TOK * pitch;
    const T* vsrc = value + (int64_t)t0 * value_stride + h * head_size;
    const T* ksrc = key + (int64_t)t0 * key_stride + h * head_size;
    if (even) {
      for (int t = my_t; t < nt; t += rows_per_pass) {
        rc_cp_async16(vt + t * pitch + my_c * PER, vsrc + (int64_t)t * value_stride + my_c * PER);
        rc_cp_async16(kt + t * pitch + my_c * PER, ksrc + (int64_t)t * key_stride + my_c * PER);
      }
    } else {
      for (int v = threadIdx.x; v < nt * vpr; v += blockDim.x) {
        const int t = v / vpr, c = v - t * vpr;
        rc_cp_async16(vt + t * pitch + c * PER, vsrc + (int64_t)t * value_stride + c * PER);
        rc_cp_async16(kt + t * pitch + c * PER, ksrc + (int64_t)t * key_stride + c * PER);
      }
    }
  };
  // Slots (first warp).  The slot of lane t is LOADED one iteration before it is decoded, so the global-memory round
  // trip hides behind an item's stores instead of holding the first warp -- and with it the CTA's next barrier -- up.
  auto load_slot = [&](int tile) -> int64_t {
    const int t0 = tile * TOK;
    return (int)threadIdx.x < min(TOK, num_tokens - t0) ? slot_mapping[t0 + threadIdx.x] : -1;
  };
  auto decode_slot = [&](int tile, int h, int stage, int64_t slot) {
    const int nt = min(TOK, num_tokens - tile * TOK);
    const int t = threadIdx.x;
    int64_t base = -1;
    int off = 0;
    if (slot >= 0) {
      int64_t blk;
      if ((slot >> 31) == 0) blk = (uint32_t)slot / (uint32_t)block_size;  // the usual case without a 64-bit division
      else blk = slot / block_size;
      off = static_cast<int>(slot - blk * block_size);
      if (blk < num_blocks) base = (blk * num_heads + h) * head_elems;
    }
    if (t < TOK) {
      s_base[stage][t] = base;
      s_off[stage][t] = off;
    }
    const int lead = t & ~(G - 1);  // first token of this token's aligned group of G (G is a power of two here)
    const int64_t lbase = __shfl_sync(0xffffffffu, base, lead & 31);
    const int loff = __shfl_sync(0xffffffffu, off, lead & 31);
    const bool ok = t >= TOK || (nt == TOK && base >= 0 && base == lbase && off == loff + (t - lead) && (loff & (G - 1)) == 0);
    const bool all_ok = __all_sync(0xffffffffu, ok);
    if (t == 0) s_fast[stage] = all_ok && v_vec;
  };

  // Items are walked with (tile, head) cursors advanced by the grid stride -- no division per item.
  struct Cursor {
    int tile, h;
  };
  const int step_t = (int)(gridDim.x / (unsigned)num_heads), step_h = (int)(gridDim.x % (unsigned)num_heads);
  auto advance = [&](Cursor& c) {
    c.tile += step_t;
    c.h += step_h;
    if (c.h >= num_heads) {
      c.h -= num_heads;
      ++c.tile;
    }
  };
  const bool slot_warp = threadIdx.x < 32;
  const int vg0 = threadIdx.x / head_size, vd0 = threadIdx.x - vg0 * head_size;
  Cursor cur{(int)(blockIdx.x / (unsigned)num_heads), (int)(blockIdx.x % (unsigned)num_heads)};
  Cursor nxt = cur;  // the item whose copies are issued next
#pragma unroll
  for (int s = 0; s < STAGES - 1; ++s) {
    if (nxt.tile < num_tiles) {
      issue(nxt.tile, nxt.h, s);
      if (slot_warp) decode_slot(nxt.tile, nxt.h, s, load_slot(nxt.tile));
    }
    rc_cp_async_commit();  // one group per stage, empty or not: the wait below counts groups
    advance(nxt);
  }
  Cursor pf = nxt;       // the item whose slots are loaded next: one ahead of nxt
  int64_t slot_pf = -1;  // slots of item nxt
  if (slot_warp && nxt.tile < num_tiles) slot_pf = load_slot(nxt.tile);
  advance(pf);
  int stage = 0;
  for (; cur.tile < num_tiles; advance(cur)) {
    rc_cp_async_wait<STAGES - 2>();  // this thread's copies of the current item have landed
    __syncthreads();                 // everybody's have, and everybody is done with the stage refilled next
    {
      const int ns = stage == 0 ? STAGES - 1 : stage - 1;
      if (nxt.tile < num_tiles) {
        issue(nxt.tile, nxt.h, ns);
        if (slot_warp) decode_slot(nxt.tile, nxt.h, ns, slot_pf);
      }
      rc_cp_async_commit();
      if (slot_warp) slot_pf = pf.tile < num_tiles ? load_slot(pf.tile) : -1;
      advance(nxt);
      advance(pf);
    }
    const int tile = cur.tile;
    const int nt = min(TOK, num_tokens - tile * TOK);
    const T* vt = reinterpret_cast<const T*>(smem_raw) + (size_t)stage * stage_elems;
    const T* kt = vt + TOK * pitch;
    // ---- K: one 16-byte cache vector per (token, chunk), token-fastest ----
    const int chunks = head_size / X;
    for (int v = threadIdx.x; v < nt * chunks; v += blockDim.x) {
      const int c = nt == TOK ? v / TOK : v / nt, t = v - c * nt;
      const int64_t base = s_base[stage][t];
      if (base < 0) continue;
      alignas(16) T src[X];
#pragma unroll
      for (int i = 0; i < X / PER; ++i)
        *reinterpret_cast<uint4*>(src + i * PER) = *reinterpret_cast<const uint4*>(kt + t * pitch + c * X + i * PER);
      alignas(16) CT tmp[X];
      cache_cvt_vec_packed<T, CT, FP8, X>(src, tmp);
      *reinterpret_cast<uint4*>(key_cache + base + ((int64_t)c * block_size + s_off[stage][t]) * X) =
          *reinterpret_cast<const uint4*>(tmp);
    }
    // ---- V: a group's row is whole 16-byte vectors; consecutive lanes write consecutive rows ----
    if (s_fast[stage]) {
      const int ngroups = TOK / G;
      for (int i = threadIdx.x, pass = 0; i < head_size * ngroups; i += blockDim.x, ++pass) {
        int g = vg0, d = vd0;  // first pass: computed once per kernel
        if (pass != 0) {
          g = i / head_size;
          d = i - g * head_size;
        }
        const int tl = g * G;
        CT* row = value_cache + s_base[stage][tl] + (int64_t)d * block_size + s_off[stage][tl];
        const T* col = vt + tl * pitch + d;
        if (v_wide) {
          // a group's row is 32 bytes (16 two-byte tokens): one 256-bit store per lane, whole sectors only
#pragma unroll
          for (int w = 0; w < TOK / (2 * VPE); ++w) {
            if (w * 2 * VPE < G) {
              alignas(16) T line[2 * VPE];
#pragma unroll
              for (int e = 0; e < 2 * VPE; ++e) line[e] = col[(w * 2 * VPE + e) * pitch];
              alignas(16) CT tmp[2 * VPE];
              cache_cvt_vec<T, CT, FP8, 2 * VPE>(line, tmp);
              st_v8(row + w * 2 * VPE, *reinterpret_cast<const uint4*>(tmp), *reinterpret_cast<const uint4*>(tmp + VPE));
            }
          }
          continue;
        }
#pragma unroll
        for (int w = 0; w < TOK / VPE; ++w) {
          if (w * VPE < G) {
            alignas(16) T line[VPE];
#pragma unroll
            for (int e = 0; e < VPE; ++e) line[e] = col[(w * VPE + e) * pitch];
            alignas(16) CT tmp[VPE];
            cache_cvt_vec<T, CT, FP8, VPE>(line, tmp);
            *reinterpret_cast<uint4*>(row + w * VPE) = *reinterpret_cast<const uint4*>(tmp);
          }
        }
      }
    } else {
      for (int i = threadIdx.x; i < nt * head_size; i += blockDim.x) {
        const int d = i / nt, t = i - d * nt;
        if (s_base[stage][t] < 0) continue;
        value_cache[s_base[stage][t] + (int64_t)d * block_size + s_off[stage][t]] = cache_cvt<T, CT, FP8>(vt[t * pitch + d]);
      }
    }
    stage = stage + 1 == STAGES ? 0 : stage + 1;
  }
  rc_cp_async_wait<0>();
}

template <typename T, typename CT, bool FP8>
static int launch_reshape_and_cache(const void* key, const void* value, void* key_cache,
                                    void* value_cache, const int64_t* slot_mapping, int num_tokens,
                                    int num_heads, int head_size, int block_size, int64_t key_stride,
                                    int64_t value_stride, int64_t num_blocks, cudaStream_t stream) {
  constexpr int X = 16 / sizeof(CT);
  constexpr int PER = 16 / (int)sizeof(T);
  if (head_size % X != 0) return B200_ERR_HEAD_SIZE;
  // 16-byte vector path: every source row and the caches are 16-byte aligned and rows are whole vectors
  const bool vec = (reinterpret_cast<uintptr_t>(key) % 16 == 0) && (reinterpret_cast<uintptr_t>(value) % 16 == 0) &&
                   ((key_stride * (int64_t)sizeof(T)) % 16 == 0) && ((value_stride * (int64_t)sizeof(T)) % 16 == 0) &&
                   (head_size % PER == 0) && (reinterpret_cast<uintptr_t>(value_cache) % 16 == 0) && (block_size % 8 == 0);
#define B200_RC_LAUNCH(KERNEL, GRID, THREADS, SMEM, ...)                                                  \
  do {                                                                                                    \
    if (vec) B200_CUDA_TRY(launch_kernel(KERNEL<T, CT, FP8, X, true>, dim3(GRID), dim3(THREADS), SMEM, stream,   \
                                         tl_pdl_launch, __VA_ARGS__));                                    \
    else B200_CUDA_TRY(launch_kernel(KERNEL<T, CT, FP8, X, false>, dim3(GRID), dim3(THREADS), SMEM, stream,      \
                                     tl_pdl_launch, __VA_ARGS__));                                        \
  } while (0)
#define B200_RC_LAUNCH_TILE(KS, TK, GRID, SMEM, ...)                                                      \
  do {                                                                                                    \
    if (vec) reshape_and_cache_tile_kernel<T, CT, FP8, X, true, KS, TK><<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);   \
    else reshape_and_cache_tile_kernel<T, CT, FP8, X, false, KS, TK><<<GRID, 256, SMEM, stream>>>(__VA_ARGS__);      \
  } while (0)
  // Prefill-sized calls take the tile kernel; decode-sized calls (one token per sequence, every token in a
  // different block: up to a few hundred tokens) take the token kernel.  The slots cannot be inspected on
  // the host, so the split is by size: measured (profiles/), a 128-token decode batch costs 2x more through
  // the tile kernel (32 CTAs, element-wise path) than through the token kernel.
  if (num_tokens >= 512) {
    static const int variant = getenv("B200_RC_VARIANT") ? atoi(getenv("B200_RC_VARIANT")) : 1;  // tuning knobs
    static const int tok_env = getenv("B200_RC_TOK") ? atoi(getenv("B200_RC_TOK")) : 32;
    const bool ksmem = variant & 1;
    // 32-token tiles while they still fit the default 48 KiB of dynamic shared memory, else 16
    const size_t row = (size_t)(head_size + PER) * sizeof(T);
    // and while 32-token tiles still give every SM a few waves of CTAs (measured: C4 prefill +20 %, C5 -6 %)
    const bool many = (long long)((num_tokens + 31) / 32) * num_heads >= 4ll * 8 * num_sms(-1);
    const int tok = (tok_env == 32 && many && (ksmem ? 2 : 1) * 32 * row <= 48 * 1024) ? 32 : 16;
    // streamed variant: vector path only, and only where V rows of a group are whole 16-byte vectors
    static const bool no_stream = getenv("B200_RC_NO_STREAM") != nullptr;  // A/B knob
    constexpr int VPE = 16 / (int)sizeof(CT);
    if (vec && !no_stream && block_size % VPE == 0) {
      static const int stok_env = getenv("B200_RC_STREAM_TOK") ? atoi(getenv("B200_RC_STREAM_TOK")) : 32;  // tuning knobs
      static const int stages = getenv("B200_RC_STREAM_STAGES") ? atoi(getenv("B200_RC_STREAM_STAGES")) : 3;
      const int nst = stages == 2 ? 2 : 3;
      const int stok = (stok_env == 32 && 3 * 2 * 32 * row <= 56 * 1024) ? 32 : 16;
      const size_t ssmem = (size_t)nst * 2 * stok * row;
      if (ssmem <= 100 * 1024 && stok % std::min(block_size, stok) == 0) {
        const int stiles = (num_tokens + stok - 1) / stok;
        const long long items = (long long)stiles * num_heads;
        const int per_sm = (int)std::min<size_t>(8, (227 * 1024) / (ssmem + 2048));
        const long long resident = (long long)per_sm * num_sms(-1);
        // the ring only pays when every resident CTA walks several items; smaller calls keep one CTA per tile
        if (items >= 4 * resident) {
        const int grid = (int)resident;
        auto kern = stok == 32 ? (nst == 3 ? reshape_and_cache_stream_kernel<T, CT, FP8, X, 32, 3>
                                           : reshape_and_cache_stream_kernel<T, CT, FP8, X, 32, 2>)
                               : (nst == 3 ? reshape_and_cache_stream_kernel<T, CT, FP8, X, 16, 3>
                                           : reshape_and_cache_stream_kernel<T, CT, FP8, X, 16, 2>);
        static const bool no_wide = getenv("B200_NO_ST256") != nullptr;  // A/B knob
        const int swide = !no_wide && reinterpret_cast<uintptr_t>(value_cache) % 32 == 0;
        B200_CUDA_TRY(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)ssmem));
        kern<<<grid, 256, ssmem, stream>>>((const T*)key, (const T*)value, (CT*)key_cache, (CT*)value_cache, slot_mapping,
                                           key_stride, value_stride, num_tokens, num_heads, head_size, block_size,
                                           num_blocks, stiles, swide);
        B200_LAUNCH_CHECK();
        return B200_OK;
        }
      }
    }
    const int tiles = (num_tokens + tok - 1) / tok;
    const int head_major = (variant & 2) && tiles <= 65535;
    dim3 grid = head_major ? dim3(num_heads, tiles) : dim3(tiles, num_heads);
    const size_t smem = (size_t)(ksmem ? 2 : 1) * tok * row;
#define B200_RC_TILE_ARGS (const T*)key, (const T*)value, (CT*)key_cache, (CT*)value_cache, slot_mapping, key_stride, \
                          value_stride, num_tokens, num_heads, head_size, block_size, num_blocks, head_major
    if (ksmem && tok == 32) B200_RC_LAUNCH_TILE(true, 32, grid, smem, B200_RC_TILE_ARGS);
    else if (ksmem) B200_RC_LAUNCH_TILE(true, 16, grid, smem, B200_RC_TILE_ARGS);
    else if (tok == 32) B200_RC_LAUNCH_TILE(false, 32, grid, smem, B200_RC_TILE_ARGS);
    else B200_RC_LAUNCH_TILE(false, 16, grid, smem, B200_RC_TILE_ARGS);
#undef B200_RC_TILE_ARGS
  } else {
    const int work = num_heads * head_size / PER;  // 16-byte source vectors per token
    const int threads = std::min(512, std::max(32, (work + 31) / 32 * 32));
    B200_RC_LAUNCH(reshape_and_cache_token_kernel, num_tokens, threads, 0, (const T*)key, (const T*)value,
                   (CT*)key_cache, (CT*)value_cache, slot_mapping, key_stride, value_stride, num_heads, head_size,
                   block_size, num_blocks);
  }
#undef B200_RC_LAUNCH
#undef B200_RC_LAUNCH_TILE
  B200_LAUNCH_CHECK();
  return B200_OK;
}

// ------------------------------------------------------------------------------------------
// copy_blocks / swap_blocks (device-to-device): table-in-parameters gather copy
// ------------------------------------------------------------------------------------------
constexpr int kMaxCopyBufs = 64;    // (layer, K|V) base pointers per launch
constexpr int kMaxCopyPairs = 320;  // (src, dst) block pairs per launch

struct CopyTable {
  unsigned long long src_base[kMaxCopyBufs];
  unsigned long long dst_base[kMaxCopyBufs];
  int src_blk[kMaxCopyPairs];
  int dst_blk[kMaxCopyPairs];
};  // 64*8*2 + 320*4*2 = 3584 bytes: fits the classic 4 KiB kernel-parameter space with room

// grid = (splits_per_block, num_pairs, num_bufs); each CTA copies one slice of one block.
template <typename V>
__global__ void __launch_bounds__(256) copy_blocks_kernel(const __grid_constant__ CopyTable tab,
                                                          long long block_bytes) {
  const int pair = blockIdx.y, buf = blockIdx.z;
  const char* src = reinterpret_cast<const char*>(tab.src_base[buf]) + (long long)tab.src_blk[pair] * block_bytes;
  char* dst = reinterpret_cast<char*>(tab.dst_base[buf]) + (long long)tab.dst_blk[pair] * block_bytes;
  const long long n = block_bytes / (long long)sizeof(V);
  const long long per = (n + gridDim.x - 1) / gridDim.x;
  const long long lo = per * blockIdx.x, hi = min(n, lo + per);
  const V* s = reinterpret_cast<const V*>(src);
  V* d = reinterpret_cast<V*>(dst);
  long long i = lo + threadIdx.x;
  // 4 independent 16-byte loads in flight per thread before the first store.
  for (; i + 3 * (long long)blockDim.x < hi; i += 4 * (long long)blockDim.x) {
    V a = s[i], b = s[i + blockDim.x], c = s[i + 2 * blockDim.x], e = s[i + 3 * blockDim.x];
    d[i] = a; d[i + blockDim.x] = b; d[i + 2 * blockDim.x] = c; d[i + 3 * blockDim.x] = e;
  }
  for (; i < hi; i += blockDim.x) d[i] = s[i];
}

static int launch_copy(const int64_t* src_bases, const int64_t* dst_bases, int num_bufs,
                       const int64_t* pairs, int num_pairs, int64_t block_bytes,
                       cudaStream_t stream) {
  if (num_bufs == 0 || num_pairs == 0 || block_bytes == 0) return B200_OK;
  // vector width from the common alignment of every address involved
  uint64_t align_or = (uint64_t)block_bytes;
  for (int b = 0; b < num_bufs; ++b) align_or |= (uint64_t)src_bases[b] | (uint64_t)dst_bases[b];
  const int vec = (align_or % 16 == 0) ? 16 : (align_or % 4 == 0) ? 4 : 1;
  const int64_t nvec = block_bytes / vec;
  int splits = (int)std::min<int64_t>(std::max<int64_t>(1, nvec / (256 * 8)), 64);
  static_assert(sizeof(CopyTable) + 8 <= 4096, "kernel parameter space");
  for (int b0 = 0; b0 < num_bufs; b0 += kMaxCopyBufs) {
    const int nb = std::min(kMaxCopyBufs, num_bufs - b0);
    for (int p0 = 0; p0 < num_pairs; p0 += kMaxCopyPairs) {
      const int np = std::min(kMaxCopyPairs, num_pairs - p0);
      CopyTable tab;
      for (int b = 0; b < nb; ++b) {
        tab.src_base[b] = (unsigned long long)src_bases[b0 + b];
        tab.dst_base[b] = (unsigned long long)dst_bases[b0 + b];
      }
      for (int p = 0; p < np; ++p) {
        const int64_t s = pairs[2 * (p0 + p)], d = pairs[2 * (p0 + p) + 1];
        if (s < 0 || d < 0 || s > INT32_MAX || d > INT32_MAX) return B200_ERR_ARG;
        tab.src_blk[p] = (int)s;
        tab.dst_blk[p] = (int)d;
      }
      dim3 grid(splits, np, nb);
      if (vec == 16) copy_blocks_kernel<uint4><<<grid, 256, 0, stream>>>(tab, block_bytes);
      else if (vec == 4) copy_blocks_kernel<uint32_t><<<grid, 256, 0, stream>>>(tab, block_bytes);
      else copy_blocks_kernel<uint8_t><<<grid, 256, 0, stream>>>(tab, block_bytes);
      B200_LAUNCH_CHECK();
    }
  }
  return B200_OK;
}

// ------------------------------------------------------------------------------------------
// convert_fp8_e5m2
// ------------------------------------------------------------------------------------------
// 16 elements per thread per step: 16-byte vectors on both sides when the host proved alignment; the
// tail (numel % 16) and unaligned tensors take the scalar loop of the same kernel.
// U groups per thread per step: all of a step's loads are issued before the first conversion, so a thread keeps
// U x 32 (to fp8) / U x 16 (from fp8) bytes in flight -- one group per step left the kernel latency-bound
// (profiles/r05_cache_ops_ncu.txt: from_fp8 at 0.70 of the copy peak with 57 % of the warps resident).
constexpr int kCvtUnroll = 4;
template <typename T, bool VEC>
__global__ void __launch_bounds__(256) to_fp8_kernel(const T* __restrict__ src, uint8_t* __restrict__ dst, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t done = 0;
  if constexpr (VEC) {
    const int64_t groups = n / 16;
    for (int64_t g0 = tid; g0 < groups; g0 += kCvtUnroll * stride) {
      alignas(16) T in[kCvtUnroll][16];
#pragma unroll
      for (int u = 0; u < kCvtUnroll; ++u)
        if (g0 + u * stride < groups) load_elems<T, 16, true>(src + (g0 + u * stride) * 16, in[u]);
#pragma unroll
      for (int u = 0; u < kCvtUnroll; ++u) {
        if (g0 + u * stride < groups) {
          alignas(16) uint8_t out[16];
#pragma unroll
          for (int e = 0; e < 16; e += 2) *reinterpret_cast<uint16_t*>(out + e) = to_e5m2_x2<T>(in[u][e], in[u][e + 1]);
          st_na_v4(dst + (g0 + u * stride) * 16, *reinterpret_cast<const uint4*>(out));
        }
      }
    }
    done = groups * 16;
  }
  for (int64_t i = done + tid; i < n; i += stride) dst[i] = to_e5m2<T>(src[i]);
}
template <typename T, bool VEC>
__global__ void __launch_bounds__(256) from_fp8_kernel(const uint8_t* __restrict__ src, T* __restrict__ dst, int64_t n,
                                                       int wide) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  int64_t done = 0;
  if constexpr (VEC) {
    constexpr int PER = 16 / (int)sizeof(T);
    const int64_t groups = n / 16;
    for (int64_t g0 = tid; g0 < groups; g0 += kCvtUnroll * stride) {
      uint4 raw[kCvtUnroll];
#pragma unroll
      for (int u = 0; u < kCvtUnroll; ++u)
        if (g0 + u * stride < groups) raw[u] = ld_nc_v4(src + (g0 + u * stride) * 16);
#pragma unroll
      for (int u = 0; u < kCvtUnroll; ++u) {
        if (g0 + u * stride < groups) {
          const uint8_t* in = reinterpret_cast<const uint8_t*>(&raw[u]);
          alignas(16) T out[16];
#pragma unroll
          for (int e = 0; e < 16; ++e) out[e] = from_e5m2<T>(in[e]);
          if (sizeof(T) == 2 && wide) {
            // 16 two-byte elements = one 32-byte sector per lane, one 256-bit store: a warp writes 1 KiB in one piece
            // (two 16-byte stores per lane left every store instruction with half-filled sectors: 0.71 of the HBM peak)
            st_v8(dst + (g0 + u * stride) * 16, *reinterpret_cast<const uint4*>(out), *reinterpret_cast<const uint4*>(out + 8));
          } else {
#pragma unroll
            for (int v = 0; v < 16 / PER; ++v)
              st_na_v4(dst + (g0 + u * stride) * 16 + v * PER, *reinterpret_cast<const uint4*>(out + v * PER));
          }
        }
      }
    }
    done = groups * 16;
  }
  for (int64_t i = done + tid; i < n; i += stride) dst[i] = from_e5m2<T>(src[i]);
}

template <typename T>
static void launch_to_fp8(const void* src, void* dst, int64_t n, int blocks, bool vec, cudaStream_t stream) {
  if (vec) to_fp8_kernel<T, true><<<blocks, 256, 0, stream>>>((const T*)src, (uint8_t*)dst, n);
  else to_fp8_kernel<T, false><<<blocks, 256, 0, stream>>>((const T*)src, (uint8_t*)dst, n);
}
template <typename T>
static void launch_from_fp8(const void* src, void* dst, int64_t n, int blocks, bool vec, cudaStream_t stream) {
  static const bool no_wide = getenv("B200_NO_ST256") != nullptr;  // A/B knob
  const int wide = !no_wide && reinterpret_cast<uintptr_t>(dst) % 32 == 0;
  if (vec) from_fp8_kernel<T, true><<<blocks, 256, 0, stream>>>((const uint8_t*)src, (T*)dst, n, wide);
  else from_fp8_kernel<T, false><<<blocks, 256, 0, stream>>>((const uint8_t*)src, (T*)dst, n, wide);
}

}  // namespace b200

using namespace b200;

extern "C" {

B200_API int b200_kv_abi_version(void) { return B200_KV_ABI_VERSION; }

B200_API const char* b200_kv_last_cuda_error(void) { return g_last_cuda_error; }

B200_API uint64_t b200_kv_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

B200_API const char* b200_kv_strerror(int status) {
  switch (status) {
    case B200_OK: return "ok";
    case B200_ERR_HEAD_SIZE: return "Unsupported head size";
    case B200_ERR_BLOCK_SIZE: return "Unsupported block size";
    case B200_ERR_DTYPE: return "Unsupported data type";
    case B200_ERR_ARG: return "Invalid argument";
    case B200_ERR_ALIGN: return "KV cache chunk is not 16-byte aligned";
    case B200_ERR_CUDA: return "CUDA runtime error";
    case B200_ERR_DEVICE: return "src and dst must be on the same GPU";
    case B200_ERR_TIMEOUT: return "KV wait timed out";
    case B200_ERR_SCRATCH_LIMIT: return "call needs more scratch than the library's static array holds (B200_V2_STATIC_SCRATCH_BYTES)";
    default: return "unknown status";
  }
}

B200_API int b200_reshape_and_cache(const void* key, const void* value, void* key_cache, void* value_cache,
                           const int64_t* slot_mapping, int32_t num_tokens, int32_t num_heads,
                           int32_t head_size, int32_t block_size, int32_t x, int64_t key_stride,
                           int64_t value_stride, int64_t num_blocks, int32_t src_dtype,
                           int32_t cache_dtype, b200_stream_t stream_) {
  if (num_tokens == 0) return B200_OK;
  if (!key || !value || !key_cache || !value_cache || !slot_mapping || num_tokens < 0 ||
      num_heads <= 0 || head_size <= 0 || block_size <= 0)
    return B200_ERR_ARG;
  const int cs = dtype_size(cache_dtype);
  if (cs == 0 || dtype_size(src_dtype) == 0 || src_dtype == B200_U8_E5M2) return B200_ERR_DTYPE;
  if (cache_dtype != B200_U8_E5M2 && cache_dtype != src_dtype) return B200_ERR_DTYPE;
  if (x != 16 / cs) return B200_ERR_ARG;  // layout contract: x = 16 / sizeof(cache element)
  if ((reinterpret_cast<uintptr_t>(key_cache) % 16) != 0) return B200_ERR_ALIGN;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
#define RC(T, CT, FP8)                                                                              \
  return launch_reshape_and_cache<T, CT, FP8>(key, value, key_cache, value_cache, slot_mapping,    \
                                              num_tokens, num_heads, head_size, block_size,         \
                                              key_stride, value_stride, num_blocks, stream)
  if (cache_dtype == B200_U8_E5M2) {
    if (src_dtype == B200_F32) RC(float, uint8_t, true);
    if (src_dtype == B200_F16) RC(__half, uint8_t, true);
    if (src_dtype == B200_BF16) RC(__nv_bfloat16, uint8_t, true);
  } else {
    // "auto": a pure byte move; 2-byte types share one instantiation
    if (src_dtype == B200_F32) RC(uint32_t, uint32_t, false);
    RC(uint16_t, uint16_t, false);
  }
#undef RC
  return B200_ERR_DTYPE;
}

B200_API int b200_copy_blocks(const int64_t* key_cache_ptrs, const int64_t* value_cache_ptrs,
                     int32_t num_layers, const int64_t* block_pairs, int32_t num_pairs,
                     int64_t block_bytes, b200_stream_t stream_) {
  if (num_layers == 0 || num_pairs == 0) return B200_OK;
  if (!key_cache_ptrs || !value_cache_ptrs || !block_pairs || num_layers < 0 || num_pairs < 0 ||
      block_bytes <= 0)
    return B200_ERR_ARG;
  // one buffer list: K of every layer, then V of every layer (in-place copy: src base == dst base)
  std::string scratch((size_t)num_layers * 2 * sizeof(int64_t), '\0');
  int64_t* bases = reinterpret_cast<int64_t*>(&scratch[0]);
  for (int l = 0; l < num_layers; ++l) {
    bases[l] = key_cache_ptrs[l];
    bases[num_layers + l] = value_cache_ptrs[l];
  }
  return launch_copy(bases, bases, 2 * num_layers, block_pairs, num_pairs, block_bytes,
                     static_cast<cudaStream_t>(stream_));
}

B200_API int b200_swap_blocks(const void* src, void* dst, int64_t block_bytes, const int64_t* block_pairs,
                     int32_t num_pairs, int32_t kind, int32_t src_device, int32_t dst_device,
                     b200_stream_t stream_) {
  if (num_pairs == 0) return B200_OK;
  if (!src || !dst || !block_pairs || num_pairs < 0 || block_bytes <= 0) return B200_ERR_ARG;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (kind == B200_SWAP_D2D) {
    if (src_device != dst_device) return B200_ERR_DEVICE;
    int64_t s = reinterpret_cast<int64_t>(src), d = reinterpret_cast<int64_t>(dst);
    return launch_copy(&s, &d, 1, block_pairs, num_pairs, block_bytes, stream);
  }
  if (kind != B200_SWAP_D2H && kind != B200_SWAP_H2D) return B200_ERR_ARG;
  const cudaMemcpyKind mk = kind == B200_SWAP_D2H ? cudaMemcpyDeviceToHost : cudaMemcpyHostToDevice;
  const char* s = static_cast<const char*>(src);
  char* d = static_cast<char*>(dst);
  // runs where src and dst both advance by one block
  int runs = 0;
  for (int i = 0; i < num_pairs; ++i)
    if (i == 0 || block_pairs[2 * i] != block_pairs[2 * (i - 1)] + 1 ||
        block_pairs[2 * i + 1] != block_pairs[2 * (i - 1) + 1] + 1)
      ++runs;
  // Fragmented block lists (the usual case: a swapped sequence's blocks are scattered on both sides) against
  // PINNED host memory: one gather kernel that reads / writes the host side through its device mapping, so
  // the PCIe link sees one stream of 16-byte accesses instead of one DMA descriptor per 128 KiB block
  // (the reference issues one cudaMemcpyAsync per block, cache_kernels.cu:47-58).  Pageable host memory and
  // lists that merge into a few long runs keep the copy engine.
  static const bool no_zero_copy = getenv("B200_SWAP_NO_ZERO_COPY") != nullptr;  // A/B knob
  // (measured, profiles/: H2D through the mapping beats the copy engine even for one long run: 50 vs 38-45 GB/s)
  if ((runs > 2 || kind == B200_SWAP_H2D) && !no_zero_copy) {
    const void* host = kind == B200_SWAP_D2H ? (const void*)dst : src;
    cudaPointerAttributes attr;
    if (cudaPointerGetAttributes(&attr, host) == cudaSuccess && attr.type == cudaMemoryTypeHost &&
        attr.devicePointer != nullptr) {
      int64_t sp = kind == B200_SWAP_D2H ? reinterpret_cast<int64_t>(src) : reinterpret_cast<int64_t>(attr.devicePointer);
      int64_t dp = kind == B200_SWAP_D2H ? reinterpret_cast<int64_t>(attr.devicePointer) : reinterpret_cast<int64_t>(dst);
      return launch_copy(&sp, &dp, 1, block_pairs, num_pairs, block_bytes, stream);
    }
    (void)cudaGetLastError();
  }
  int i = 0;
  while (i < num_pairs) {
    int j = i + 1;
    while (j < num_pairs && block_pairs[2 * j] == block_pairs[2 * (j - 1)] + 1 &&
           block_pairs[2 * j + 1] == block_pairs[2 * (j - 1) + 1] + 1)
      ++j;
    B200_CUDA_TRY(cudaMemcpyAsync(d + block_pairs[2 * i + 1] * block_bytes,
                                  s + block_pairs[2 * i] * block_bytes,
                                  (size_t)(j - i) * block_bytes, mk, stream));
    i = j;
  }
  return B200_OK;
}

B200_API int b200_convert_fp8_e5m2(const void* src, void* dst, int64_t numel, int32_t src_dtype,
                          int32_t dst_dtype, b200_stream_t stream_) {
  if (numel == 0) return B200_OK;
  if (!src || !dst || numel < 0) return B200_ERR_ARG;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const bool vec = (reinterpret_cast<uintptr_t>(src) % 16 == 0) && (reinterpret_cast<uintptr_t>(dst) % 16 == 0);
  const int64_t work = vec ? (numel + 15) / 16 : numel;
  const int blocks = (int)std::min<int64_t>((work + 255) / 256, (int64_t)num_sms(-1) * 16);
  if (dst_dtype == B200_U8_E5M2) {
    if (src_dtype == B200_F32) launch_to_fp8<float>(src, dst, numel, blocks, vec, stream);
    else if (src_dtype == B200_F16) launch_to_fp8<__half>(src, dst, numel, blocks, vec, stream);
    else if (src_dtype == B200_BF16) launch_to_fp8<__nv_bfloat16>(src, dst, numel, blocks, vec, stream);
    else return B200_ERR_DTYPE;
  } else if (src_dtype == B200_U8_E5M2) {
    if (dst_dtype == B200_F32) launch_from_fp8<float>(src, dst, numel, blocks, vec, stream);
    else if (dst_dtype == B200_F16) launch_from_fp8<__half>(src, dst, numel, blocks, vec, stream);
    else if (dst_dtype == B200_BF16) launch_from_fp8<__nv_bfloat16>(src, dst, numel, blocks, vec, stream);
    else return B200_ERR_DTYPE;
  } else {
    return B200_ERR_DTYPE;
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_get_device_attribute(int32_t attribute, int32_t device_id) {
  int device = device_id, value = 0;
  if (device_id < 0 && cudaGetDevice(&device) != cudaSuccess) return -1;
  if (cudaDeviceGetAttribute(&value, static_cast<cudaDeviceAttr>(attribute), device) != cudaSuccess) {
    (void)cudaGetLastError();
    return -1;
  }
  return value;
}

B200_API int b200_get_max_shared_memory_per_block_device_attribute(int32_t device_id) {
  return b200_get_device_attribute((int)cudaDevAttrMaxSharedMemoryPerBlockOptin, device_id);
}

}  // extern "C"
