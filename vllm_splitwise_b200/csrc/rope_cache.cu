// Rotary position embedding, stand-alone and fused with the KV-cache write.
//   b200_rotary_embedding                   <- reference csrc/pos_encoding_kernels.cu:9-95, 126-165
//   b200_rotary_embedding_reshape_and_cache <- that op followed by reshape_and_cache
//                                              (csrc/cache_kernels.cu:154-268), as the model calls them back
//                                              to back (vllm/model_executor/models/llama.py:156-158).
// In a decode step the two reference ops are two launches that both touch the step's K rows; fused, the K row
// is read once, rotated in shared memory, and leaves to BOTH its in-place home (the caller may still read it)
// and the paged cache, and V rides along -- one launch instead of two per layer.
//
// Arithmetic is the reference's, operation by operation: x' = x*cos - y*sin, y' = y*cos + x*sin with every
// product and the sum rounded to the element type (the reference computes in scalar_t, c10::Half / BFloat16
// operators: float op, round), so f16/bf16 results are bit-identical to rotary_embedding + reshape_and_cache.
#include <algorithm>
#include <stdlib.h>

#include "common.cuh"

namespace b200 {
namespace {

template <typename T> __device__ __forceinline__ T rnd_mul(T a, T b);
template <typename T> __device__ __forceinline__ T rnd_sub(T a, T b);
template <typename T> __device__ __forceinline__ T rnd_add(T a, T b);
template <> __device__ __forceinline__ float rnd_mul<float>(float a, float b) { return a * b; }
template <> __device__ __forceinline__ float rnd_sub<float>(float a, float b) { return a - b; }
template <> __device__ __forceinline__ float rnd_add<float>(float a, float b) { return a + b; }
#define B200_RND_OPS(T, TO_F, FROM_F)                                                                              \
  template <> __device__ __forceinline__ T rnd_mul<T>(T a, T b) { return FROM_F(__fmul_rn(TO_F(a), TO_F(b))); }    \
  template <> __device__ __forceinline__ T rnd_sub<T>(T a, T b) { return FROM_F(__fsub_rn(TO_F(a), TO_F(b))); }    \
  template <> __device__ __forceinline__ T rnd_add<T>(T a, T b) { return FROM_F(__fadd_rn(TO_F(a), TO_F(b))); }
B200_RND_OPS(__half, __half2float, __float2half_rn)
B200_RND_OPS(__nv_bfloat16, __bfloat162float, __float2bfloat16_rn)
#undef B200_RND_OPS

// One (x, y) pair of one head (reference apply_token_rotary_embedding, pos_encoding_kernels.cu:9-38).
// For float the expression is written exactly as the reference writes it, so the compiler contracts it the
// same way (x*cos - y*sin -> fma); for 2-byte types every step is rounded explicitly.
template <typename T, bool NEOX>
__device__ __forceinline__ void rotate_pair(T* arr, const T* __restrict__ cos_ptr, const T* __restrict__ sin_ptr,
                                            int rot_offset, int embed_dim) {
  int xi, yi;
  T c, s;
  if (NEOX) {
    xi = rot_offset;
    yi = embed_dim + rot_offset;
    c = cos_ptr[xi];
    s = sin_ptr[xi];
  } else {
    xi = 2 * rot_offset;
    yi = 2 * rot_offset + 1;
    c = cos_ptr[rot_offset];
    s = sin_ptr[rot_offset];
  }
  const T x = arr[xi], y = arr[yi];
  if constexpr (sizeof(T) == 4) {
    arr[xi] = x * c - y * s;
    arr[yi] = y * c + x * s;
  } else {
    arr[xi] = rnd_sub<T>(rnd_mul<T>(x, c), rnd_mul<T>(y, s));
    arr[yi] = rnd_add<T>(rnd_mul<T>(y, c), rnd_mul<T>(x, s));
  }
}

// Stand-alone op: one CTA per token, query heads then key heads, in place.
template <typename T, bool NEOX>
__global__ void __launch_bounds__(512) rotary_embedding_kernel(
    const int64_t* __restrict__ positions, T* __restrict__ query, T* __restrict__ key,
    const T* __restrict__ cos_sin_cache, const int64_t* __restrict__ cache_offsets, int rot_dim,
    int64_t query_stride, int64_t key_stride, int num_heads, int num_kv_heads, int head_size) {
  const int64_t token = blockIdx.x;
  int64_t pos = positions[token];
  if (cache_offsets != nullptr) pos += cache_offsets[token];  // batched_rotary_embedding (:97-123)
  const int embed = rot_dim / 2;
  const T* cos_ptr = cos_sin_cache + pos * rot_dim;
  const T* sin_ptr = cos_ptr + embed;
  const int nq = num_heads * embed;
  for (int i = threadIdx.x; i < nq; i += blockDim.x) {
    const int hd = i / embed;
    rotate_pair<T, NEOX>(query + token * query_stride + (int64_t)hd * head_size, cos_ptr, sin_ptr, i - hd * embed, embed);
  }
  const int nk = num_kv_heads * embed;
  for (int i = threadIdx.x; i < nk; i += blockDim.x) {
    const int hd = i / embed;
    rotate_pair<T, NEOX>(key + token * key_stride + (int64_t)hd * head_size, cos_ptr, sin_ptr, i - hd * embed, embed);
  }
}

template <typename T, typename CT, bool FP8>
__device__ __forceinline__ CT cache_cvt(T v) {
  if constexpr (FP8) return to_e5m2<T>(v);
  else return v;
}

// Fused op, token-major (decode-sized calls): one CTA per token.
//   q: rotated in place (global).
//   k: row -> shared memory, rotated there, then written to its in-place home and to the paged K cache
//      (one 16-byte vector per (head, d/x chunk)).
//   v: straight to the paged V cache (stride-block_size scatter, as in reshape_and_cache).
// A padding slot (< 0) still rotates q and k (rotary_embedding has no notion of padding) and skips the cache.
template <typename T, typename CT, bool FP8, int X, bool NEOX>
__global__ void __launch_bounds__(256) rope_reshape_and_cache_token_kernel(
    const int64_t* __restrict__ positions, T* __restrict__ query, T* key, const T* __restrict__ value,
    const T* __restrict__ cos_sin_cache, int rot_dim, int64_t query_stride, int64_t key_stride, int64_t value_stride,
    int num_heads, int num_kv_heads, int head_size, CT* __restrict__ key_cache, CT* __restrict__ value_cache,
    const int64_t* __restrict__ slot_mapping, int block_size, int64_t num_blocks) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  T* ks = reinterpret_cast<T*>(smem_raw);  // [num_kv_heads * head_size]
  griddep_launch_dependents();
  griddep_wait();  // positions / slot_mapping may come from the decode-metadata kernel right before (b200_decode_step)
  const int64_t token = blockIdx.x;
  const int64_t pos = positions[token];
  const int64_t slot = slot_mapping[token];
  const int embed = rot_dim / 2;
  const T* cos_ptr = cos_sin_cache + pos * rot_dim;
  const T* sin_ptr = cos_ptr + embed;
  const int nkv = num_kv_heads * head_size;
  T* krow = key + token * key_stride;
  for (int i = threadIdx.x; i < nkv; i += blockDim.x) ks[i] = krow[i];
  // q while the K row lands
  const int nq = num_heads * embed;
  for (int i = threadIdx.x; i < nq; i += blockDim.x) {
    const int hd = i / embed;
    rotate_pair<T, NEOX>(query + token * query_stride + (int64_t)hd * head_size, cos_ptr, sin_ptr, i - hd * embed, embed);
  }
  __syncthreads();
  const int nk = num_kv_heads * embed;
  for (int i = threadIdx.x; i < nk; i += blockDim.x) {
    const int hd = i / embed;
    rotate_pair<T, NEOX>(ks + hd * head_size, cos_ptr, sin_ptr, i - hd * embed, embed);
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nkv; i += blockDim.x) krow[i] = ks[i];
  if (slot < 0) return;
  const int64_t blk = slot / block_size;
  const int off = static_cast<int>(slot - blk * block_size);
  if (blk >= num_blocks) return;
  CT* kdst = key_cache + blk * (int64_t)nkv * block_size;
  CT* vdst = value_cache + blk * (int64_t)nkv * block_size;
  for (int v = threadIdx.x; v < nkv / X; v += blockDim.x) {
    alignas(16) CT tmp[X];
#pragma unroll
    for (int e = 0; e < X; ++e) tmp[e] = cache_cvt<T, CT, FP8>(ks[v * X + e]);
    *reinterpret_cast<uint4*>(kdst + ((int64_t)v * block_size + off) * X) = *reinterpret_cast<const uint4*>(tmp);
  }
  const T* vrow = value + token * value_stride;
  for (int i = threadIdx.x; i < nkv; i += blockDim.x)
    vdst[(int64_t)i * block_size + off] = cache_cvt<T, CT, FP8>(vrow[i]);
}

// ---- vectorised path ---------------------------------------------------------------------------
// The scalar kernels above walk (head, pair) items with 2-byte loads in a loop whose iterations the compiler
// must keep in order (in-place updates may alias); with many heads that is several dependent global round
// trips per token.  Here one work item is a 16-byte vector (q) or one x-element cache chunk (k) and a token's
// items are spread over up to 512 threads, so every global load of the token is in flight at once:
//   q item  NeoX: vectors j and j + embed of a head (PER pairs);   GPT-J: one vector (PER/2 adjacent pairs)
//   k item  NeoX: chunks c and c + embed/X of a head (X pairs);    GPT-J: one chunk (X/2 pairs);
//           rotated in registers, stored in place AND (CACHE) as 16-byte vectors of the paged K cache;
//           chunks past rot_dim are pass-through (cache write only)
//   v item  (CACHE) one 16-byte source vector scattered to the paged V cache.
template <typename T, int N>
__device__ __forceinline__ void ld_row(const T* p, T* out) {  // N * sizeof(T) is 8 or a multiple of 16
  constexpr int BYTES = N * (int)sizeof(T);
  if constexpr (BYTES % 16 == 0) {
#pragma unroll
    for (int i = 0; i < BYTES / 16; ++i) reinterpret_cast<uint4*>(out)[i] = reinterpret_cast<const uint4*>(p)[i];
  } else {
    static_assert(BYTES == 8, "8-byte tail only");
    *reinterpret_cast<uint2*>(out) = *reinterpret_cast<const uint2*>(p);
  }
}
template <typename T, int N>
__device__ __forceinline__ void st_row(T* p, const T* in) {
  constexpr int BYTES = N * (int)sizeof(T);
  static_assert(BYTES % 16 == 0, "whole vectors");
#pragma unroll
  for (int i = 0; i < BYTES / 16; ++i) reinterpret_cast<uint4*>(p)[i] = reinterpret_cast<const uint4*>(in)[i];
}

template <typename T>
__device__ __forceinline__ void rotate_xy(T& x, T& y, T c, T s) {
  const T x0 = x, y0 = y;
  if constexpr (sizeof(T) == 4) {
    x = x0 * c - y0 * s;
    y = y0 * c + x0 * s;
  } else {
    x = rnd_sub<T>(rnd_mul<T>(x0, c), rnd_mul<T>(y0, s));
    y = rnd_add<T>(rnd_mul<T>(y0, c), rnd_mul<T>(x0, s));
  }
}

// N elements at xp (and, NeoX, N more at xp + embed) rotated in registers; results left in xs / ys.
template <typename T, int N, bool NEOX>
__device__ __forceinline__ void rotate_item(const T* xp, int embed, const T* __restrict__ cos_p,
                                            const T* __restrict__ sin_p, T* xs, T* ys) {
  if constexpr (NEOX) {
    alignas(16) T cs[N], sn[N];
    ld_row<T, N>(xp, xs);
    ld_row<T, N>(xp + embed, ys);
    ld_row<T, N>(cos_p, cs);
    ld_row<T, N>(sin_p, sn);
#pragma unroll
    for (int e = 0; e < N; ++e) rotate_xy<T>(xs[e], ys[e], cs[e], sn[e]);
  } else {
    alignas(16) T cs[N / 2], sn[N / 2];
    ld_row<T, N>(xp, xs);
    ld_row<T, N / 2>(cos_p, cs);
    ld_row<T, N / 2>(sin_p, sn);
#pragma unroll
    for (int e = 0; e < N / 2; ++e) rotate_xy<T>(xs[2 * e], xs[2 * e + 1], cs[e], sn[e]);
  }
}

template <typename T, typename CT, bool FP8, int X, bool NEOX, bool CACHE>
__global__ void __launch_bounds__(512) rope_vec_kernel(
    const int64_t* __restrict__ positions, const int64_t* __restrict__ cache_offsets, T* query, T* key,
    const T* __restrict__ value, const T* __restrict__ cos_sin_cache, int rot_dim, int64_t query_stride,
    int64_t key_stride, int64_t value_stride, int num_heads, int num_kv_heads, int head_size,
    CT* __restrict__ key_cache, CT* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int block_size,
    int64_t num_blocks) {
  constexpr int PER = 16 / (int)sizeof(T);
  griddep_launch_dependents();
  griddep_wait();  // positions / slot_mapping may come from the decode-metadata kernel right before (b200_decode_step)
  const int64_t token = blockIdx.x;
  int64_t pos = positions[token];
  if (cache_offsets != nullptr) pos += cache_offsets[token];
  const int embed = rot_dim / 2;
  const T* cos_ptr = cos_sin_cache + pos * rot_dim;
  const T* sin_ptr = cos_ptr + embed;
  T* qrow = query + token * query_stride;
  T* krow = key + token * key_stride;

  // cache destination of this token (CACHE only); a negative slot is padding: rotate, skip the cache
  bool cache_ok = false;
  int off = 0;
  CT* kdst = nullptr;
  CT* vdst = nullptr;
  const int nkv = num_kv_heads * head_size;
  if constexpr (CACHE) {
    const int64_t slot = slot_mapping[token];
    if (slot >= 0) {
      const int64_t blk = slot / block_size;
      off = static_cast<int>(slot - blk * block_size);
      cache_ok = blk < num_blocks;
      kdst = key_cache + blk * (int64_t)nkv * block_size;
      vdst = value_cache + blk * (int64_t)nkv * block_size;
    }
  }

  const int q_per_head = (NEOX ? embed : rot_dim) / PER;
  const int nq = num_heads * q_per_head;
  const int chunks = head_size / X;                         // cache chunks per head
  const int k_rot = (NEOX ? embed : rot_dim) / X;           // rotating k items per head
  const int k_pass = CACHE ? (head_size - rot_dim) / X : 0; // pass-through chunks per head
  const int k_per_head = k_rot + k_pass;
  const int nk = num_kv_heads * k_per_head;
  const int nv = CACHE ? nkv / PER : 0;
  const int total = nq + nk + nv;

  for (int it = threadIdx.x; it < total; it += blockDim.x) {
    if (it < nq) {
      const int h = it / q_per_head, j = it - h * q_per_head;
      T* xp = qrow + (int64_t)h * head_size + j * PER;
      alignas(16) T xs[PER], ys[PER];
      rotate_item<T, PER, NEOX>(xp, embed, cos_ptr + (NEOX ? j * PER : j * PER / 2), sin_ptr + (NEOX ? j * PER : j * PER / 2),
                                xs, ys);
      st_row<T, PER>(xp, xs);
      if constexpr (NEOX) st_row<T, PER>(xp + embed, ys);
    } else if (it < nq + nk) {
      const int i = it - nq;
      const int h = i / k_per_head, c = i - h * k_per_head;
      T* hrow = krow + (int64_t)h * head_size;
      alignas(16) T xs[X], ys[X];
      alignas(16) CT tmp[X];
      if (c < k_rot) {
        T* xp = hrow + c * X;
        rotate_item<T, X, NEOX>(xp, embed, cos_ptr + (NEOX ? c * X : c * X / 2), sin_ptr + (NEOX ? c * X : c * X / 2), xs, ys);
        st_row<T, X>(xp, xs);
        if constexpr (NEOX) st_row<T, X>(xp + embed, ys);
        if (CACHE && cache_ok) {
#pragma unroll
          for (int e = 0; e < X; ++e) tmp[e] = cache_cvt<T, CT, FP8>(xs[e]);
          *reinterpret_cast<uint4*>(kdst + ((int64_t)(h * chunks + c) * block_size + off) * X) =
              *reinterpret_cast<const uint4*>(tmp);
          if constexpr (NEOX) {
#pragma unroll
            for (int e = 0; e < X; ++e) tmp[e] = cache_cvt<T, CT, FP8>(ys[e]);
            *reinterpret_cast<uint4*>(kdst + ((int64_t)(h * chunks + c + embed / X) * block_size + off) * X) =
                *reinterpret_cast<const uint4*>(tmp);
          }
        }
      } else if (CACHE && cache_ok) {  // pass-through chunk: d in [rot_dim, head_size)
        const int cc = rot_dim / X + (c - k_rot);
        ld_row<T, X>(hrow + cc * X, xs);
#pragma unroll
        for (int e = 0; e < X; ++e) tmp[e] = cache_cvt<T, CT, FP8>(xs[e]);
        *reinterpret_cast<uint4*>(kdst + ((int64_t)(h * chunks + cc) * block_size + off) * X) =
            *reinterpret_cast<const uint4*>(tmp);
      }
    } else if (CACHE) {
      const int i = it - nq - nk;
      alignas(16) T vs[PER];
      *reinterpret_cast<uint4*>(vs) = ld_nc_v4(value + token * value_stride + (int64_t)i * PER);
      if (cache_ok) {
#pragma unroll
        for (int e = 0; e < PER; ++e)
          vdst[(int64_t)(i * PER + e) * block_size + off] = cache_cvt<T, CT, FP8>(vs[e]);
      }
    }
  }
}

// Host-side proof that the vectorised kernel's 16-byte (and 8-byte) accesses are aligned.
inline bool rope_vec_ok(const void* query, const void* key, const void* value, const void* cos_sin_cache, int elt,
                        int head_size, int rot_dim, int64_t qs, int64_t ks, int64_t vs, bool neox, int x) {
  const int embed = rot_dim / 2;
  auto a16 = [](const void* p) { return reinterpret_cast<uintptr_t>(p) % 16 == 0; };
  if (!a16(query) || !a16(key) || (value && !a16(value)) || !a16(cos_sin_cache)) return false;
  if ((qs * elt) % 16 || (ks * elt) % 16 || (vs * elt) % 16 || (head_size * elt) % 16) return false;
  if ((rot_dim * elt) % 16 || (embed * elt) % 16) return false;
  if (head_size % x || rot_dim % x) return false;
  return neox ? embed % x == 0 : true;
}

template <typename T>
int launch_rotary(const int64_t* positions, void* query, void* key, const void* cos_sin_cache, const int64_t* offsets,
                  int64_t num_tokens, int num_heads, int num_kv_heads, int head_size, int rot_dim, int64_t query_stride,
                  int64_t key_stride, bool neox, cudaStream_t stream) {
  constexpr int PER = 16 / (int)sizeof(T);
  static const bool no_vec = getenv("B200_ROPE_NO_VEC") != nullptr;  // A/B knob
  if (!no_vec && rope_vec_ok(query, key, nullptr, cos_sin_cache, (int)sizeof(T), head_size, rot_dim, query_stride,
                             key_stride, 0, neox, PER)) {
    const int items = (num_heads + num_kv_heads) * (neox ? rot_dim / 2 : rot_dim) / PER;
    const int threads = std::max(32, std::min((items + 31) / 32 * 32, 512));
    if (neox)
      rope_vec_kernel<T, T, false, PER, true, false><<<(unsigned)num_tokens, threads, 0, stream>>>(
          positions, offsets, (T*)query, (T*)key, nullptr, (const T*)cos_sin_cache, rot_dim, query_stride, key_stride, 0,
          num_heads, num_kv_heads, head_size, nullptr, nullptr, nullptr, 0, 0);
    else
      rope_vec_kernel<T, T, false, PER, false, false><<<(unsigned)num_tokens, threads, 0, stream>>>(
          positions, offsets, (T*)query, (T*)key, nullptr, (const T*)cos_sin_cache, rot_dim, query_stride, key_stride, 0,
          num_heads, num_kv_heads, head_size, nullptr, nullptr, nullptr, 0, 0);
    B200_LAUNCH_CHECK();
    return B200_OK;
  }
  const int threads = std::max(32, std::min(num_heads * rot_dim / 2, 512));  // reference :139-140
  if (neox)
    rotary_embedding_kernel<T, true><<<(unsigned)num_tokens, threads, 0, stream>>>(
        positions, (T*)query, (T*)key, (const T*)cos_sin_cache, offsets, rot_dim, query_stride, key_stride, num_heads,
        num_kv_heads, head_size);
  else
    rotary_embedding_kernel<T, false><<<(unsigned)num_tokens, threads, 0, stream>>>(
        positions, (T*)query, (T*)key, (const T*)cos_sin_cache, offsets, rot_dim, query_stride, key_stride, num_heads,
        num_kv_heads, head_size);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <typename T, typename CT, bool FP8>
int launch_fused_token(const int64_t* positions, void* query, void* key, const void* value, const void* cos_sin_cache,
                       int64_t num_tokens, int num_heads, int num_kv_heads, int head_size, int rot_dim, int64_t qs,
                       int64_t ks, int64_t vs, bool neox, void* key_cache, void* value_cache, const int64_t* slots,
                       int block_size, int64_t num_blocks, cudaStream_t stream) {
  constexpr int X = 16 / (int)sizeof(CT);
  constexpr int PER = 16 / (int)sizeof(T);
  if (head_size % X != 0) return B200_ERR_HEAD_SIZE;
  static const bool no_vec = getenv("B200_ROPE_NO_VEC") != nullptr;  // A/B knob
  if (!no_vec &&
      rope_vec_ok(query, key, value, cos_sin_cache, (int)sizeof(T), head_size, rot_dim, qs, ks, vs, neox, X)) {
    const int half = neox ? rot_dim / 2 : rot_dim;
    const int items = num_heads * half / PER + num_kv_heads * (half / X + (head_size - rot_dim) / X) +
                      num_kv_heads * head_size / PER;
    const int threads = std::max(32, std::min((items + 31) / 32 * 32, 512));
#define B200_FUSED_VEC(NEOX)                                                                                      \
  B200_CUDA_TRY(launch_kernel(rope_vec_kernel<T, CT, FP8, X, NEOX, true>, dim3((unsigned)num_tokens), dim3(threads), 0, \
      stream, tl_pdl_launch,                                                                                      \
      positions, nullptr, (T*)query, (T*)key, (const T*)value, (const T*)cos_sin_cache, rot_dim, qs, ks, vs,      \
      num_heads, num_kv_heads, head_size, (CT*)key_cache, (CT*)value_cache, slots, block_size, num_blocks))
    if (neox) B200_FUSED_VEC(true);
    else B200_FUSED_VEC(false);
#undef B200_FUSED_VEC
    B200_LAUNCH_CHECK();
    return B200_OK;
  }
  const size_t smem = (size_t)num_kv_heads * head_size * sizeof(T);
  if (smem > 48 * 1024) return B200_ERR_ARG;
#define B200_FUSED(NEOX)                                                                                          \
  B200_CUDA_TRY(launch_kernel(rope_reshape_and_cache_token_kernel<T, CT, FP8, X, NEOX>, dim3((unsigned)num_tokens),  \
      dim3(256), smem, stream, tl_pdl_launch,                                                                     \
      positions, (T*)query, (T*)key, (const T*)value, (const T*)cos_sin_cache, rot_dim, qs, ks, vs, num_heads,    \
      num_kv_heads, head_size, (CT*)key_cache, (CT*)value_cache, slots, block_size, num_blocks))
  if (neox) B200_FUSED(true);
  else B200_FUSED(false);
#undef B200_FUSED
  B200_LAUNCH_CHECK();
  return B200_OK;
}

int check_rotary(const void* positions, const void* query, const void* key, const void* cache, int64_t num_tokens,
                 int num_heads, int num_kv_heads, int head_size, int rot_dim, int dtype) {
  if (num_tokens < 0 || num_heads <= 0 || num_kv_heads <= 0 || head_size <= 0) return B200_ERR_ARG;
  if (num_tokens > 0 && (!positions || !query || !key || !cache)) return B200_ERR_ARG;
  if (rot_dim <= 0 || rot_dim % 2 != 0 || rot_dim > head_size) return B200_ERR_ARG;
  if (dtype != B200_F32 && dtype != B200_F16 && dtype != B200_BF16) return B200_ERR_DTYPE;
  return B200_OK;
}

}  // namespace
}  // namespace b200

using namespace b200;

extern "C" {

B200_API int b200_rotary_embedding(const int64_t* positions, void* query, void* key, const void* cos_sin_cache,
                                   const int64_t* cos_sin_cache_offsets, int64_t num_tokens, int32_t num_heads,
                                   int32_t num_kv_heads, int32_t head_size, int32_t rot_dim, int64_t query_stride,
                                   int64_t key_stride, int32_t is_neox, int32_t dtype, b200_stream_t stream_) {
  int st = check_rotary(positions, query, key, cos_sin_cache, num_tokens, num_heads, num_kv_heads, head_size, rot_dim, dtype);
  if (st != B200_OK || num_tokens == 0) return st;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
#define B200_ROT(T)                                                                                               \
  return launch_rotary<T>(positions, query, key, cos_sin_cache, cos_sin_cache_offsets, num_tokens, num_heads,    \
                          num_kv_heads, head_size, rot_dim, query_stride, key_stride, is_neox != 0, stream)
  if (dtype == B200_F32) B200_ROT(float);
  if (dtype == B200_F16) B200_ROT(__half);
  B200_ROT(__nv_bfloat16);
#undef B200_ROT
}

B200_API int b200_rotary_embedding_reshape_and_cache(
    const int64_t* positions, void* query, void* key, const void* value, const void* cos_sin_cache, int64_t num_tokens,
    int32_t num_heads, int32_t num_kv_heads, int32_t head_size, int32_t rot_dim, int64_t query_stride, int64_t key_stride,
    int64_t value_stride, int32_t is_neox, void* key_cache, void* value_cache, const int64_t* slot_mapping,
    int32_t block_size, int32_t x, int64_t num_blocks, int32_t src_dtype, int32_t cache_dtype, b200_stream_t stream_) {
  int st = check_rotary(positions, query, key, cos_sin_cache, num_tokens, num_heads, num_kv_heads, head_size, rot_dim, src_dtype);
  if (st != B200_OK || num_tokens == 0) return st;
  if (!value || !key_cache || !value_cache || !slot_mapping || block_size <= 0) return B200_ERR_ARG;
  const int cs = dtype_size(cache_dtype);
  if (cs == 0) return B200_ERR_DTYPE;
  if (cache_dtype != B200_U8_E5M2 && cache_dtype != src_dtype) return B200_ERR_DTYPE;
  if (x != 16 / cs) return B200_ERR_ARG;
  if ((reinterpret_cast<uintptr_t>(key_cache) % 16) != 0) return B200_ERR_ALIGN;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  // Prefill-sized calls: rotate in place, then the tiled cache write (two large launches; the K rows the second
  // one reads are still in L2).  Decode-sized calls: the single fused launch.
  const size_t smem = (size_t)num_kv_heads * head_size * dtype_size(src_dtype);
  if (num_tokens >= 512 || smem > 48 * 1024) {
    st = b200_rotary_embedding(positions, query, key, cos_sin_cache, nullptr, num_tokens, num_heads, num_kv_heads,
                               head_size, rot_dim, query_stride, key_stride, is_neox, src_dtype, stream_);
    if (st != B200_OK) return st;
    return b200_reshape_and_cache(key, value, key_cache, value_cache, slot_mapping, (int32_t)num_tokens, num_kv_heads,
                                  head_size, block_size, x, key_stride, value_stride, num_blocks, src_dtype, cache_dtype,
                                  stream_);
  }
#define B200_FT(T, CT, FP8)                                                                                        \
  return launch_fused_token<T, CT, FP8>(positions, query, key, value, cos_sin_cache, num_tokens, num_heads,       \
                                        num_kv_heads, head_size, rot_dim, query_stride, key_stride, value_stride, \
                                        is_neox != 0, key_cache, value_cache, slot_mapping, block_size, num_blocks, stream)
  if (cache_dtype == B200_U8_E5M2) {
    if (src_dtype == B200_F32) B200_FT(float, uint8_t, true);
    if (src_dtype == B200_F16) B200_FT(__half, uint8_t, true);
    B200_FT(__nv_bfloat16, uint8_t, true);
  }
  if (src_dtype == B200_F32) B200_FT(float, float, false);
  if (src_dtype == B200_F16) B200_FT(__half, __half, false);
  B200_FT(__nv_bfloat16, __nv_bfloat16, false);
#undef B200_FT
}

}  // extern "C"
