// Cache writes and block moves for the paged KV cache (sm_100a).
//   b200_reshape_and_cache  <- reference csrc/cache_kernels.cu:154-268
//   b200_copy_blocks        <- reference csrc/cache_kernels.cu:65-150
//   b200_swap_blocks        <- reference csrc/cache_kernels.cu:21-60
//   b200_convert_fp8_e5m2   <- reference csrc/cache_kernels.cu:272-320
// All four are HBM-bound byte movers: the design rules are coalesced 16-byte accesses,
// tables carried as kernel parameters (no H2D copy, no sync) and grids sized from the work.
#include <algorithm>
#include <atomic>
#include <cstring>
#include <mutex>
#include <string>

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
// Token-major kernel: one CTA per token.  K: every (head, d/x chunk) is x contiguous source
// elements and x contiguous destination elements (16 bytes in the cache), so it moves as one
// vector.  V: element (h, d) lands at [blk, h, d, off] -- a scatter with stride block_size; the
// tile kernel below handles the prefill case where consecutive tokens fill consecutive offsets.
template <typename T, typename CT, bool FP8, int X>
__global__ void __launch_bounds__(256) reshape_and_cache_token_kernel(
    const T* __restrict__ key, const T* __restrict__ value, CT* __restrict__ key_cache,
    CT* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int64_t key_stride,
    int64_t value_stride, int num_heads, int head_size, int block_size, int64_t num_blocks) {
  const int64_t token = blockIdx.x;
  const int64_t slot = slot_mapping[token];
  if (slot < 0) return;  // padding (reference cache_kernels.cu:169-172)
  const int64_t blk = slot / block_size;
  const int off = static_cast<int>(slot % block_size);
  if (blk >= num_blocks) return;  // out-of-range slot: never write outside the cache
  const int chunks = head_size / X;
  const int n_vec = num_heads * chunks;
  const T* ksrc = key + token * key_stride;
  const T* vsrc = value + token * value_stride;
  CT* kdst = key_cache + blk * (int64_t)num_heads * head_size * block_size;
  CT* vdst = value_cache + blk * (int64_t)num_heads * head_size * block_size;

  for (int v = threadIdx.x; v < n_vec; v += blockDim.x) {
    const int h = v / chunks, c = v - h * chunks;
    const T* s = ksrc + h * head_size + c * X;
    CT* d = kdst + ((int64_t)(h * chunks + c) * block_size + off) * X;
    alignas(16) CT tmp[X];
#pragma unroll
    for (int e = 0; e < X; ++e) {
      if constexpr (FP8) tmp[e] = to_e5m2<T>(s[e]);
      else tmp[e] = s[e];
    }
    *reinterpret_cast<uint4*>(d) = *reinterpret_cast<const uint4*>(tmp);
  }
  const int n = num_heads * head_size;
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    CT o;
    if constexpr (FP8) o = to_e5m2<T>(vsrc[i]);
    else o = vsrc[i];
    vdst[(int64_t)i * block_size + off] = o;
  }
}

// Tile kernel (prefill): one CTA per (group of TOK consecutive tokens, head).  K moves as one
// 16-byte vector per (token, d/x chunk).  V is read coalesced along d, transposed through shared
// memory and written token-fastest, so when consecutive tokens occupy consecutive slots (the
// normal prefill layout) every row d of the [D, block_size] tile receives one contiguous run
// instead of `n` isolated 1-2 byte stores block_size elements apart (the reference's pattern,
// cache_kernels.cu:188-191,204).  Slots are looked up per token, so any mapping is correct.
template <typename T, typename CT, bool FP8, int X, int TOK>
__global__ void __launch_bounds__(256) reshape_and_cache_tile_kernel(
    const T* __restrict__ key, const T* __restrict__ value, CT* __restrict__ key_cache,
    CT* __restrict__ value_cache, const int64_t* __restrict__ slot_mapping, int64_t key_stride,
    int64_t value_stride, int num_tokens, int num_heads, int head_size, int block_size,
    int64_t num_blocks) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  CT* vt = reinterpret_cast<CT*>(smem_raw);  // [head_size][TOK + PAD]
  constexpr int PAD = 16 / sizeof(CT);
  constexpr int LD = TOK + PAD;
  __shared__ int64_t s_base[TOK];  // element offset of (blk, h, 0, 0) or -1 = skip
  __shared__ int s_off[TOK];

  const int t0 = blockIdx.x * TOK;
  const int h = blockIdx.y;
  const int nt = min(TOK, num_tokens - t0);
  if (threadIdx.x < TOK) {
    int64_t base = -1;
    int off = 0;
    if (threadIdx.x < nt) {
      const int64_t slot = slot_mapping[t0 + threadIdx.x];
      if (slot >= 0) {
        const int64_t blk = slot / block_size;
        off = static_cast<int>(slot - blk * block_size);
        if (blk < num_blocks) base = (blk * num_heads + h) * (int64_t)head_size * block_size;
      }
    }
    s_base[threadIdx.x] = base;
    s_off[threadIdx.x] = off;
  }
  __syncthreads();
  const int chunks = head_size / X;

  for (int v = threadIdx.x; v < nt * chunks; v += blockDim.x) {
    const int c = v / nt, t = v - c * nt;  // t fastest: neighbouring threads hit neighbouring slots
    if (s_base[t] < 0) continue;
    const T* s = key + (int64_t)(t0 + t) * key_stride + h * head_size + c * X;
    alignas(16) CT tmp[X];
#pragma unroll
    for (int e = 0; e < X; ++e) {
      if constexpr (FP8) tmp[e] = to_e5m2<T>(s[e]);
      else tmp[e] = s[e];
    }
    *reinterpret_cast<uint4*>(key_cache + s_base[t] + ((int64_t)c * block_size + s_off[t]) * X) =
        *reinterpret_cast<const uint4*>(tmp);
  }
  for (int i = threadIdx.x; i < nt * head_size; i += blockDim.x) {
    const int t = i / head_size, d = i - t * head_size;
    const T sv = value[(int64_t)(t0 + t) * value_stride + h * head_size + d];
    CT o;
    if constexpr (FP8) o = to_e5m2<T>(sv);
    else o = sv;
    vt[d * LD + t] = o;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < nt * head_size; i += blockDim.x) {
    const int d = i / nt, t = i - d * nt;
    if (s_base[t] < 0) continue;
    value_cache[s_base[t] + (int64_t)d * block_size + s_off[t]] = vt[d * LD + t];
  }
}

template <typename T, typename CT, bool FP8>
static int launch_reshape_and_cache(const void* key, const void* value, void* key_cache,
                                    void* value_cache, const int64_t* slot_mapping, int num_tokens,
                                    int num_heads, int head_size, int block_size, int64_t key_stride,
                                    int64_t value_stride, int64_t num_blocks, cudaStream_t stream) {
  constexpr int X = 16 / sizeof(CT);
  if (head_size % X != 0) return B200_ERR_HEAD_SIZE;
  constexpr int TOK = 16;
  // Prefill-sized calls take the tile kernel; decode-sized calls (one token per sequence, every
  // token in a different block) take the token kernel.
  if (num_tokens >= 4 * TOK) {
    dim3 grid((num_tokens + TOK - 1) / TOK, num_heads);
    size_t smem = (size_t)head_size * (TOK + 16 / sizeof(CT)) * sizeof(CT);
    reshape_and_cache_tile_kernel<T, CT, FP8, X, TOK><<<grid, 256, smem, stream>>>(
        (const T*)key, (const T*)value, (CT*)key_cache, (CT*)value_cache, slot_mapping, key_stride,
        value_stride, num_tokens, num_heads, head_size, block_size, num_blocks);
  } else {
    int threads = std::min(256, std::max(32, (num_heads * head_size + 31) / 32 * 32));
    reshape_and_cache_token_kernel<T, CT, FP8, X><<<num_tokens, threads, 0, stream>>>(
        (const T*)key, (const T*)value, (CT*)key_cache, (CT*)value_cache, slot_mapping, key_stride,
        value_stride, num_heads, head_size, block_size, num_blocks);
  }
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
template <typename T>
__global__ void __launch_bounds__(256) to_fp8_kernel(const T* __restrict__ src, uint8_t* __restrict__ dst, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = to_e5m2<T>(src[i]);
}
template <typename T>
__global__ void __launch_bounds__(256) from_fp8_kernel(const uint8_t* __restrict__ src, T* __restrict__ dst, int64_t n) {
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) dst[i] = from_e5m2<T>(src[i]);
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
  // merge runs where src and dst both advance by one block: one copy per run
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
  const int blocks = (int)std::min<int64_t>((numel + 255) / 256, (int64_t)num_sms(-1) * 16);
  if (dst_dtype == B200_U8_E5M2) {
    if (src_dtype == B200_F32) to_fp8_kernel<float><<<blocks, 256, 0, stream>>>((const float*)src, (uint8_t*)dst, numel);
    else if (src_dtype == B200_F16) to_fp8_kernel<__half><<<blocks, 256, 0, stream>>>((const __half*)src, (uint8_t*)dst, numel);
    else if (src_dtype == B200_BF16) to_fp8_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>((const __nv_bfloat16*)src, (uint8_t*)dst, numel);
    else return B200_ERR_DTYPE;
  } else if (src_dtype == B200_U8_E5M2) {
    if (dst_dtype == B200_F32) from_fp8_kernel<float><<<blocks, 256, 0, stream>>>((const uint8_t*)src, (float*)dst, numel);
    else if (dst_dtype == B200_F16) from_fp8_kernel<__half><<<blocks, 256, 0, stream>>>((const uint8_t*)src, (__half*)dst, numel);
    else if (dst_dtype == B200_BF16) from_fp8_kernel<__nv_bfloat16><<<blocks, 256, 0, stream>>>((const uint8_t*)src, (__nv_bfloat16*)dst, numel);
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
