// Splitwise prefill -> decode KV-block transfer over NVLink peer memory (sm_100a).
//
// Replaces the reference's MSCCL++ proxy-channel path: three one-thread kernels that push a
// trigger into a host FIFO, a host proxy thread that turns each trigger into a cudaMemcpyAsync
// over CudaIpc, and a semaphore (csrc/kv_comm_kernels.cu:3-29, vllm/worker/comm_utils.py:117-147).
// Here the prompt GPU's SMs store the blocks straight into the token GPU's cache through a peer
// mapping -- no host hop, one launch for any number of (layer, K|V) buffers and block runs --
// and a system-scope release/acquire flag replaces the semaphore.
#include <algorithm>
#include <cuda.h>

#include "common.cuh"

namespace b200 {
namespace {

constexpr int kMaxRuns = 440;        // (block_start, num_blocks) runs per launch (kernel params)
constexpr int kPutThreads = 256;
constexpr int kTileVecs = 2048;      // 16-byte vectors per tile = 32 KiB

struct PutTable {
  int start[kMaxRuns];
  int len[kMaxRuns];
  int num_runs;
  int num_bufs;
  int tile_vecs;              // 16-byte vectors per tile (host: divides a block evenly when it can)
  long long block_bytes;
  const long long* src_ptrs;  // device arrays [num_bufs]
  const long long* dst_ptrs;
};
static_assert(sizeof(PutTable) <= 4000, "kernel parameter space");

// CTAs take tiles round-robin; a tile lies inside one (buffer, run).  Each thread keeps up to 8 independent
// 16-byte loads from local HBM in flight and then issues the stores to the peer mapping.
__device__ __forceinline__ void put_tiles(const int* start, const int* len, const int* run_tile0, int num_runs,
                                          int num_bufs, int tv, long long block_bytes, const long long* src_ptrs,
                                          const long long* dst_ptrs) {
  const long long tpb = run_tile0[num_runs];
  const long long total = tpb * num_bufs;
  for (long long tile = blockIdx.x; tile < total; tile += gridDim.x) {
    const int buf = (int)(tile / tpb);
    const int rt = (int)(tile - (long long)buf * tpb);
    int lo = 0, hi = num_runs - 1;  // last run with run_tile0[r] <= rt
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (run_tile0[mid] <= rt) lo = mid; else hi = mid - 1;
    }
    const long long run_vecs = (long long)len[lo] * block_bytes / 16;
    const long long v0 = (long long)(rt - run_tile0[lo]) * tv;
    const long long v1 = min(run_vecs, v0 + tv);
    const long long byte0 = (long long)start[lo] * block_bytes;
    const uint4* src = reinterpret_cast<const uint4*>(src_ptrs[buf] + byte0);
    uint4* dst = reinterpret_cast<uint4*>(dst_ptrs[buf] + byte0);
    for (long long i = v0 + threadIdx.x; i < v1; i += 8 * kPutThreads) {
      uint4 r[8];
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (i + k * kPutThreads < v1) r[k] = ld_nc_v4(src + i + k * kPutThreads);
#pragma unroll
      for (int k = 0; k < 8; ++k)
        if (i + k * kPutThreads < v1) st_na_v4(dst + i + k * kPutThreads, r[k]);
    }
  }
}

// Work = num_bufs x (sum of run lengths) blocks, cut into tiles of `tile_vecs` 16-byte vectors
// (<= 32 KiB).  CTAs take tiles round-robin; a tile lies inside one (buffer, run).  Each thread keeps up
// to 8 independent 16-byte loads from local HBM in flight and then issues the stores to the peer mapping.
__global__ void __launch_bounds__(kPutThreads) kv_put_blocks_kernel(const __grid_constant__ PutTable t) {
  __shared__ int run_tile0[kMaxRuns + 1];  // exclusive prefix of tiles per run
  __shared__ int warp_sum[kPutThreads / 32];
  const int tv = t.tile_vecs;
  {
    // block-wide exclusive scan of tiles per run: thread i owns runs 2i and 2i+1 (kMaxRuns <= 2 * threads)
    static_assert(kMaxRuns <= 2 * kPutThreads, "two runs per thread");
    const int r0 = 2 * threadIdx.x, r1 = r0 + 1;
    const int c0 = r0 < t.num_runs ? (int)(((long long)t.len[r0] * t.block_bytes / 16 + tv - 1) / tv) : 0;
    const int c1 = r1 < t.num_runs ? (int)(((long long)t.len[r1] * t.block_bytes / 16 + tv - 1) / tv) : 0;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    int incl = c0 + c1;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += n;
    }
    if (lane == 31) warp_sum[warp] = incl;
    __syncthreads();
    int base = 0;
    for (int w = 0; w < warp; ++w) base += warp_sum[w];
    const int excl = base + incl - (c0 + c1);
    if (r0 < t.num_runs) run_tile0[r0] = excl;
    if (r1 < t.num_runs) run_tile0[r1] = excl + c0;
    if (threadIdx.x == kPutThreads - 1) run_tile0[t.num_runs] = base + incl;
    __syncthreads();
  }
  put_tiles(t.start, t.len, run_tile0, t.num_runs, t.num_bufs, tv, t.block_bytes, t.src_ptrs, t.dst_ptrs);
}

// Same copy for run lists that do not fit the kernel-parameter table: starts, lengths and the exclusive tile
// prefix live in DEVICE memory (built once per run list by b200_kv_put_table_build and reused for every
// layer -- the backends put the same block list 2 x num_layers times); each CTA stages them in shared memory.
// tab = [start[n] | len[n] | tile0[n + 1]] int32.
__global__ void __launch_bounds__(kPutThreads) kv_put_blocks_table_kernel(
    const int* __restrict__ tab, int num_runs, int num_bufs, int tile_vecs, long long block_bytes,
    const long long* __restrict__ src_ptrs, const long long* __restrict__ dst_ptrs) {
  extern __shared__ int s_tab[];
  for (int i = threadIdx.x; i < 3 * num_runs + 1; i += blockDim.x) s_tab[i] = tab[i];
  __syncthreads();
  put_tiles(s_tab, s_tab + num_runs, s_tab + 2 * num_runs, num_runs, num_bufs, tile_vecs, block_bytes, src_ptrs, dst_ptrs);
}

__global__ void kv_signal_kernel(unsigned long long* flag, unsigned long long value) {
  // everything this stream did before is ordered before the flag becomes visible to the peer
  __threadfence_system();
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(value) : "memory");
}

__global__ void kv_wait_kernel(const unsigned long long* flag, unsigned long long value, long long max_spins,
                               int* timeout_flag) {
  if (threadIdx.x != 0) return;
  unsigned long long seen = 0;
  long long spins = 0;
  unsigned ns = 32;
  for (;;) {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(flag) : "memory");
    if (seen >= value) break;
    if (++spins >= max_spins) {  // bounded like the reference's wait(100000000), kv_comm_kernels.cu:7
      if (timeout_flag) *timeout_flag = 1;
      break;
    }
    __nanosleep(ns);
    if (ns < 1024) ns <<= 1;
  }
  __threadfence_system();
}

}  // namespace
}  // namespace b200

using namespace b200;

// Tile = the largest even split of one block that stays <= 32 KiB (a 40 KiB block -> two 20 KiB tiles, not
// 32 + 8), so single-block runs of a fragmented block list load the CTAs evenly.
static int put_tile_vecs(int64_t block_bytes) {
  const long long block_vecs = block_bytes / 16;
  const long long pieces = (block_vecs + kTileVecs - 1) / kTileVecs;
  return (int)((block_vecs + pieces - 1) / pieces);
}

extern "C" {

B200_API int b200_kv_ipc_export(const void* dev_ptr, unsigned char handle[B200_IPC_HANDLE_BYTES], int64_t* offset,
                       int64_t* alloc_bytes) {
  if (!dev_ptr || !handle) return B200_ERR_ARG;
  static_assert(sizeof(cudaIpcMemHandle_t) == B200_IPC_HANDLE_BYTES, "ipc handle size");
  CUdeviceptr base = 0;
  size_t size = 0;
  B200_CUDA_TRY(cudaFree(0));  // make sure the primary context exists
  // libcuda is resolved at run time (the library must load on driverless build boxes)
  typedef CUresult (*get_range_fn)(CUdeviceptr*, size_t*, CUdeviceptr);
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult qres;
  B200_CUDA_TRY(cudaGetDriverEntryPoint("cuMemGetAddressRange", &fn, cudaEnableDefault, &qres));
  if (!fn || qres != cudaDriverEntryPointSuccess ||
      reinterpret_cast<get_range_fn>(fn)(&base, &size, (CUdeviceptr)dev_ptr) != CUDA_SUCCESS) {
    set_last_cuda_error(cudaErrorInvalidValue, "cuMemGetAddressRange");
    return B200_ERR_CUDA;
  }
  cudaIpcMemHandle_t h;
  B200_CUDA_TRY(cudaIpcGetMemHandle(&h, (void*)base));
  memcpy(handle, &h, sizeof(h));
  if (offset) *offset = (int64_t)((CUdeviceptr)dev_ptr - base);
  if (alloc_bytes) *alloc_bytes = (int64_t)size;
  return B200_OK;
}

B200_API int b200_kv_ipc_open(const unsigned char handle[B200_IPC_HANDLE_BYTES], void** base) {
  if (!handle || !base) return B200_ERR_ARG;
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof(h));
  B200_CUDA_TRY(cudaIpcOpenMemHandle(base, h, cudaIpcMemLazyEnablePeerAccess));
  return B200_OK;
}

B200_API int b200_kv_ipc_close(void* base) {
  if (!base) return B200_OK;
  B200_CUDA_TRY(cudaIpcCloseMemHandle(base));
  return B200_OK;
}

B200_API int b200_kv_enable_peer_access(int32_t dev, int32_t peer_dev) {
  if (dev == peer_dev) return B200_OK;
  int can = 0;
  B200_CUDA_TRY(cudaDeviceCanAccessPeer(&can, dev, peer_dev));
  if (!can) return B200_ERR_DEVICE;
  DeviceGuard guard;
  B200_CUDA_TRY(guard.set(dev));
  cudaError_t e = cudaDeviceEnablePeerAccess(peer_dev, 0);
  if (e == cudaErrorPeerAccessAlreadyEnabled) {
    (void)cudaGetLastError();
    return B200_OK;
  }
  B200_CUDA_TRY(e);
  return B200_OK;
}

B200_API int b200_kv_put_blocks(const int64_t* src_ptrs, const int64_t* dst_ptrs, int32_t num_bufs, const int32_t* runs,
                       int32_t num_runs, int64_t block_bytes, int32_t max_ctas, b200_stream_t stream_) {
  if (num_bufs == 0 || num_runs == 0) return B200_OK;
  if (!src_ptrs || !dst_ptrs || !runs || num_bufs < 0 || num_runs < 0 || block_bytes <= 0) return B200_ERR_ARG;
  if (block_bytes % 16 != 0) return B200_ERR_ALIGN;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (max_ctas <= 0) max_ctas = 2 * num_sms(-1);
  // Tile = the largest even split of one block that stays <= 32 KiB (a 40 KiB block -> two 20 KiB tiles, not
  // 32 + 8), so single-block runs of a fragmented block list load the CTAs evenly.
  const long long block_vecs = block_bytes / 16;
  const int tile_vecs = put_tile_vecs(block_bytes);
  for (int r0 = 0; r0 < num_runs; r0 += kMaxRuns) {
    PutTable t;
    t.num_runs = std::min(kMaxRuns, num_runs - r0);
    long long tiles = 0;
    for (int r = 0; r < t.num_runs; ++r) {
      t.start[r] = runs[2 * (r0 + r)];
      t.len[r] = runs[2 * (r0 + r) + 1];
      if (t.start[r] < 0 || t.len[r] < 0) return B200_ERR_ARG;
      tiles += ((long long)t.len[r] * block_vecs + tile_vecs - 1) / tile_vecs;
    }
    if (tiles * num_bufs > INT32_MAX) return B200_ERR_ARG;
    t.tile_vecs = tile_vecs;
    t.num_bufs = num_bufs;
    t.block_bytes = block_bytes;
    t.src_ptrs = reinterpret_cast<const long long*>(src_ptrs);
    t.dst_ptrs = reinterpret_cast<const long long*>(dst_ptrs);
    tiles *= num_bufs;
    if (tiles == 0) continue;
    const int grid = (int)std::min<long long>(tiles, max_ctas);
    kv_put_blocks_kernel<<<grid, kPutThreads, 0, stream>>>(t);
    B200_LAUNCH_CHECK();
  }
  return B200_OK;
}

B200_API int b200_kv_put_table_build(const int32_t* runs, int32_t num_runs, int64_t block_bytes, int32_t* table,
                            int32_t* tile_vecs_out, int64_t* tiles_per_buf_out) {
  if (!runs || !table || !tile_vecs_out || !tiles_per_buf_out || num_runs <= 0 || block_bytes <= 0) return B200_ERR_ARG;
  if (block_bytes % 16 != 0) return B200_ERR_ALIGN;
  const int tv = put_tile_vecs(block_bytes);
  long long acc = 0;
  for (int r = 0; r < num_runs; ++r) {
    const int start = runs[2 * r], len = runs[2 * r + 1];
    if (start < 0 || len < 0) return B200_ERR_ARG;
    table[r] = start;
    table[num_runs + r] = len;
    table[2 * num_runs + r] = (int32_t)acc;
    acc += ((long long)len * (block_bytes / 16) + tv - 1) / tv;
    if (acc > INT32_MAX) return B200_ERR_ARG;
  }
  table[3 * num_runs] = (int32_t)acc;
  *tile_vecs_out = tv;
  *tiles_per_buf_out = acc;
  return B200_OK;
}

B200_API int b200_kv_put_blocks_table(const int64_t* src_ptrs, const int64_t* dst_ptrs, int32_t num_bufs,
                             const int32_t* dev_table, int32_t num_runs, int32_t tile_vecs, int64_t tiles_per_buf,
                             int64_t block_bytes, int32_t max_ctas, b200_stream_t stream_) {
  if (num_bufs == 0 || num_runs == 0 || tiles_per_buf == 0) return B200_OK;
  if (!src_ptrs || !dst_ptrs || !dev_table || num_bufs < 0 || num_runs < 0 || tile_vecs <= 0 || tiles_per_buf < 0 ||
      block_bytes <= 0)
    return B200_ERR_ARG;
  if (block_bytes % 16 != 0) return B200_ERR_ALIGN;
  if (tiles_per_buf * num_bufs > INT32_MAX) return B200_ERR_ARG;
  const size_t smem = ((size_t)3 * num_runs + 1) * sizeof(int);
  if (smem > 200 * 1024) return B200_ERR_ARG;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  if (max_ctas <= 0) max_ctas = 2 * num_sms(-1);
  static size_t attr_set[64] = {0};
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  dev &= 63;
  if (smem > 48 * 1024 && attr_set[dev] < smem) {
    B200_CUDA_TRY(cudaFuncSetAttribute(kv_put_blocks_table_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
    attr_set[dev] = 200 * 1024;
  }
  const int grid = (int)std::min<long long>(tiles_per_buf * num_bufs, max_ctas);
  kv_put_blocks_table_kernel<<<grid, kPutThreads, smem, stream>>>(
      dev_table, num_runs, num_bufs, tile_vecs, block_bytes, reinterpret_cast<const long long*>(src_ptrs),
      reinterpret_cast<const long long*>(dst_ptrs));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_kv_signal(uint64_t* remote_flag, uint64_t value, b200_stream_t stream_) {
  if (!remote_flag) return B200_ERR_ARG;
  kv_signal_kernel<<<1, 1, 0, static_cast<cudaStream_t>(stream_)>>>(
      reinterpret_cast<unsigned long long*>(remote_flag), (unsigned long long)value);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

B200_API int b200_kv_wait(const uint64_t* local_flag, uint64_t value, int64_t max_spins, int32_t* timeout_flag,
                 b200_stream_t stream_) {
  if (!local_flag) return B200_ERR_ARG;
  if (max_spins <= 0) max_spins = 100000000LL;
  kv_wait_kernel<<<1, 32, 0, static_cast<cudaStream_t>(stream_)>>>(
      reinterpret_cast<const unsigned long long*>(local_flag), (unsigned long long)value, (long long)max_spins,
      timeout_flag);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // extern "C"
