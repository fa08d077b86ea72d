// Splitwise prefill -> decode KV-block transfer over NVLink peer memory (sm_100a).
//
// Replaces the reference's MSCCL++ proxy-channel path: three one-thread kernels that push a
// trigger into a host FIFO, a host proxy thread that turns each trigger into a cudaMemcpyAsync
// over CudaIpc, and a semaphore (csrc/kv_comm_kernels.cu:3-29, vllm/worker/comm_utils.py:117-147).
// Here the prompt GPU's SMs store the blocks straight into the token GPU's cache through a peer
// mapping -- no host hop, one launch for any number of (layer, K|V) buffers and block runs --
// and a system-scope release/acquire flag replaces the semaphore.
#include <algorithm>
#include <atomic>
#include <stdlib.h>
#include <cuda.h>

#include "common.cuh"

namespace b200 {
namespace {

constexpr int kMaxRuns = 440;        // (block_start, num_blocks) runs per launch (kernel params)
constexpr int kPutThreads = 512;     // SIMT variant: 16 warps per CTA, each moving its own 4 KiB chunks
constexpr int kChunkVecs = 256;      // SIMT variant: 16-byte vectors per warp chunk (32 lanes x 8 loads in flight)
constexpr int kTileVecs = 2048;      // bulk variant: 16-byte vectors per tile = 32 KiB (one ring stage)
constexpr int kBulkStages = 6;       // bulk variant: 6 x 32 KiB ring per CTA, up to 5 loads in flight
constexpr int kBulkThreads = 32;
constexpr int kBulkMaxRuns = 1360;   // bulk variant: run table that still fits beside the ring

struct PutTable {
  int start[kMaxRuns];
  int len[kMaxRuns];
  int num_runs;
  int num_bufs;
  int tile_vecs;              // bulk variant: 16-byte vectors per tile (host: divides a block evenly when it can)
  long long block_bytes;
  const long long* src_ptrs;  // device arrays [num_bufs]
  const long long* dst_ptrs;
  int chained;                // see put_begin / put_end
};
static_assert(sizeof(PutTable) <= 4000, "kernel parameter space");

// Stage the run list in shared memory and build the exclusive prefix of work units per run (unit = `unit_vecs`
// 16-byte vectors; the last unit of a run may be short).  s_tab = [start[n] | len[n] | unit0[n + 1]].
__device__ __forceinline__ void stage_runs(int* s_tab, const int* g_start, const int* g_len, int n, long long block_vecs,
                                           int unit_vecs) {
  __shared__ int s_part[32];
  int* s_start = s_tab;
  int* s_len = s_tab + n;
  int* s_unit0 = s_tab + 2 * n;
  const int nthr = blockDim.x, per = (n + nthr - 1) / nthr;
  const int lo = min(n, (int)threadIdx.x * per), hi = min(n, lo + per);
  int sum = 0;
  for (int i = lo; i < hi; ++i) {
    const int st = g_start[i], ln = g_len[i];
    s_start[i] = st;
    s_len[i] = ln;
    sum += (int)(((long long)ln * block_vecs + unit_vecs - 1) / unit_vecs);
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  int incl = sum;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    const int v = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += v;
  }
  if (lane == 31) s_part[warp] = incl;
  __syncthreads();
  int base = 0;
  for (int w = 0; w < warp; ++w) base += s_part[w];
  int run = base + incl - sum;
  for (int i = lo; i < hi; ++i) {
    s_unit0[i] = run;
    run += (int)(((long long)s_len[i] * block_vecs + unit_vecs - 1) / unit_vecs);
  }
  if ((int)threadIdx.x == nthr - 1) s_unit0[n] = base + incl;
  __syncthreads();
}

// unit `u` of one buffer -> (run, first vector, end vector) ; unit0 is the exclusive prefix built by stage_runs
__device__ __forceinline__ int find_run(const int* unit0, int n, int u) {
  int lo = 0, hi = n - 1;  // last run with unit0[r] <= u
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (unit0[mid] <= u) lo = mid; else hi = mid - 1;
  }
  return lo;
}

// Programmatic dependent launch between the puts of one communicator.  Every put lets its successor start at once
// (launch_dependents).  A put launched as `chained` -- the caller guarantees that its source blocks were complete
// when it was enqueued (side stream ordered by an event, comm_utils.py) -- streams immediately, overlapping the tail
// of its predecessor and hiding its own ramp, and waits for the predecessor only before it EXITS: completion stays
// transitive, so the signal kernel (an ordinary launch) still fires after every byte of every earlier put.  An
// unchained put waits before it reads anything, like any dependent kernel.
__device__ __forceinline__ void put_begin(int chained) {
  griddep_launch_dependents();
  if (!chained) griddep_wait();
}
__device__ __forceinline__ void put_end(int chained) {
  if (chained) griddep_wait();
}

// ---- SIMT variant -------------------------------------------------------------------------------------------
// Work = num_bufs x (sum of run lengths) blocks cut into 4 KiB chunks.  Every WARP takes chunks on its own
// (neighbouring warps take neighbouring chunks): 8 independent 16-byte loads per lane from local HBM, then the 8
// stores into the peer mapping, no CTA-wide barrier in the loop -- so a handful of CTAs keeps megabytes in flight.
__device__ __forceinline__ void put_chunks(const int* s_tab, int n, int num_bufs, long long block_bytes,
                                           const long long* src_ptrs, const long long* dst_ptrs) {
  const int* start = s_tab;
  const int* len = s_tab + n;
  const int* unit0 = s_tab + 2 * n;
  const long long cpb = unit0[n];
  const long long total = cpb * num_bufs;
  const int warps = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (long long c = (long long)blockIdx.x * warps + warp; c < total; c += (long long)gridDim.x * warps) {
    const int buf = (int)(c / cpb);
    const int rc = (int)(c - (long long)buf * cpb);
    const int r = find_run(unit0, n, rc);
    const long long run_vecs = (long long)len[r] * block_bytes / 16;
    const long long v0 = (long long)(rc - unit0[r]) * kChunkVecs;
    const long long v1 = min(run_vecs, v0 + kChunkVecs);
    const long long byte0 = (long long)start[r] * block_bytes;
    const uint4* src = reinterpret_cast<const uint4*>(src_ptrs[buf] + byte0);
    uint4* dst = reinterpret_cast<uint4*>(dst_ptrs[buf] + byte0);
    uint4 v[8];
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (v0 + lane + 32 * k < v1) v[k] = ld_nc_v4(src + v0 + lane + 32 * k);
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (v0 + lane + 32 * k < v1) st_na_v4(dst + v0 + lane + 32 * k, v[k]);
  }
}

__global__ void __launch_bounds__(kPutThreads) kv_put_blocks_kernel(const __grid_constant__ PutTable t) {
  extern __shared__ __align__(16) int s_tab[];
  put_begin(t.chained);
  stage_runs(s_tab, t.start, t.len, t.num_runs, t.block_bytes / 16, kChunkVecs);
  put_chunks(s_tab, t.num_runs, t.num_bufs, t.block_bytes, t.src_ptrs, t.dst_ptrs);
  put_end(t.chained);
}

// Same copy for run lists that do not fit the kernel-parameter table: starts and lengths live in DEVICE memory
// (built once per run list by b200_kv_put_table_build and reused for every layer -- the backends put the same
// block list 2 x num_layers times).  tab = [start[n] | len[n] | host-side tile prefix (unused here)] int32.
__global__ void __launch_bounds__(kPutThreads) kv_put_blocks_table_kernel(
    const int* __restrict__ tab, int num_runs, int num_bufs, long long block_bytes,
    const long long* __restrict__ src_ptrs, const long long* __restrict__ dst_ptrs, int chained) {
  extern __shared__ __align__(16) int s_tab[];
  put_begin(chained);
  stage_runs(s_tab, tab, tab + num_runs, num_runs, block_bytes / 16, kChunkVecs);
  put_chunks(s_tab, num_runs, num_bufs, block_bytes, src_ptrs, dst_ptrs);
  put_end(chained);
}

// ---- bulk-copy variant ----------------------------------------------------------------------------------------
// One thread per CTA drives the TMA engine: tiles (<= 32 KiB, inside one (buffer, run)) travel global -> shared
// memory ring (cp.async.bulk + mbarrier) -> peer global memory (cp.async.bulk.global.shared::cta), up to
// kBulkStages - 1 loads in flight per CTA.  No registers hold data and no warp spins on loads, so an SM that runs
// one of these CTAs keeps its issue slots for the prefill kernels the transfer overlaps.
__device__ __forceinline__ void bulk_s2g(void* gdst, const void* smem_src, uint32_t bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_read() { asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory"); }
template <int N>
__device__ __forceinline__ void bulk_wait_all() { asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory"); }

struct BulkShared {
  uint64_t full[kBulkStages];
  unsigned long long dst[kBulkStages];
  uint32_t bytes[kBulkStages];
};

__device__ __forceinline__ void put_tiles_bulk(uint8_t* ring, BulkShared* sh, const int* s_tab, int n, int num_bufs, int tv,
                                               long long block_bytes, const long long* src_ptrs,
                                               const long long* dst_ptrs) {
  const int* start = s_tab;
  const int* len = s_tab + n;
  const int* unit0 = s_tab + 2 * n;
  const long long tpb = unit0[n];
  const long long total = tpb * num_bufs;
  if (threadIdx.x != 0 || blockIdx.x >= total) return;
  const long long n_my = (total - blockIdx.x + gridDim.x - 1) / gridDim.x;
  const uint64_t policy = l2_policy_evict_first();
  auto issue_load = [&](long long k) {  // k-th tile of this CTA into stage k % kBulkStages
    const long long tile = blockIdx.x + k * gridDim.x;
    const int buf = (int)(tile / tpb);
    const int rt = (int)(tile - (long long)buf * tpb);
    const int r = find_run(unit0, n, rt);
    const long long run_vecs = (long long)len[r] * block_bytes / 16;
    const long long v0 = (long long)(rt - unit0[r]) * tv;
    const long long v1 = min(run_vecs, v0 + tv);
    const long long byte0 = (long long)start[r] * block_bytes + v0 * 16;
    const uint32_t bytes = (uint32_t)((v1 - v0) * 16);
    const int s = (int)(k % kBulkStages);
    sh->dst[s] = (unsigned long long)(dst_ptrs[buf] + byte0);
    sh->bytes[s] = bytes;
    mbar_arrive_expect_tx(&sh->full[s], bytes);
    bulk_g2s(ring + (size_t)s * kTileVecs * 16, reinterpret_cast<const void*>(src_ptrs[buf] + byte0), bytes, &sh->full[s],
             policy);
  };
  const long long pro = n_my < kBulkStages ? n_my : kBulkStages;
  for (long long k = 0; k < pro; ++k) issue_load(k);
  for (long long k = 0; k < n_my; ++k) {
    const int s = (int)(k % kBulkStages);
    mbar_wait(&sh->full[s], (uint32_t)((k / kBulkStages) & 1));
    bulk_s2g(reinterpret_cast<void*>(sh->dst[s]), ring + (size_t)s * kTileVecs * 16, sh->bytes[s]);
    bulk_commit();
    // the stage of tile k-1 is free once its store has read shared memory (all groups but the newest are done reading)
    if (k >= 1 && k - 1 + kBulkStages < n_my) {
      bulk_wait_read<1>();
      issue_load(k - 1 + kBulkStages);
    }
  }
  bulk_wait_all<0>();  // the writes are performed before the grid ends (the signal kernel follows in stream order)
}

__device__ __forceinline__ BulkShared* bulk_setup(uint8_t* smem) {
  BulkShared* sh = reinterpret_cast<BulkShared*>(smem + (size_t)kBulkStages * kTileVecs * 16);
  if (threadIdx.x == 0) {
    for (int s = 0; s < kBulkStages; ++s) mbar_init(&sh->full[s], 1);
    mbar_fence_init();
  }
  return sh;
}

__global__ void __launch_bounds__(kBulkThreads) kv_put_blocks_bulk_kernel(const __grid_constant__ PutTable t) {
  extern __shared__ __align__(128) uint8_t bulk_smem[];
  BulkShared* sh = bulk_setup(bulk_smem);
  int* s_tab = reinterpret_cast<int*>(sh + 1);
  put_begin(t.chained);
  stage_runs(s_tab, t.start, t.len, t.num_runs, t.block_bytes / 16, t.tile_vecs);
  put_tiles_bulk(bulk_smem, sh, s_tab, t.num_runs, t.num_bufs, t.tile_vecs, t.block_bytes, t.src_ptrs, t.dst_ptrs);
  put_end(t.chained);
}

__global__ void __launch_bounds__(kBulkThreads) kv_put_blocks_table_bulk_kernel(
    const int* __restrict__ tab, int num_runs, int num_bufs, int tile_vecs, long long block_bytes,
    const long long* __restrict__ src_ptrs, const long long* __restrict__ dst_ptrs, int chained) {
  extern __shared__ __align__(128) uint8_t bulk_smem[];
  BulkShared* sh = bulk_setup(bulk_smem);
  int* s_tab = reinterpret_cast<int*>(sh + 1);
  put_begin(chained);
  stage_runs(s_tab, tab, tab + num_runs, num_runs, block_bytes / 16, tile_vecs);
  put_tiles_bulk(bulk_smem, sh, s_tab, num_runs, num_bufs, tile_vecs, block_bytes, src_ptrs, dst_ptrs);
  put_end(chained);
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

// Bulk-variant tile = the largest even split of one block that stays <= 32 KiB (a 40 KiB block -> two 20 KiB tiles,
// not 32 + 8), so single-block runs of a fragmented block list load the CTAs evenly.
static int put_tile_vecs(int64_t block_bytes) {
  const long long block_vecs = block_bytes / 16;
  const long long pieces = (block_vecs + kTileVecs - 1) / kTileVecs;
  return (int)((block_vecs + pieces - 1) / pieces);
}

// Process-wide put configuration (b200_kv_put_configure; initial values from B200_KV_PUT_VARIANT / B200_KV_PUT_CTAS).
// variant 0 = SIMT (ld.global / st.global per warp chunk), 1 = bulk (TMA through a shared-memory ring).
// Defaults from the CTA sweep on 2 x B200 (profiles/r05_kv_put_sweep.jsonl).
static std::atomic<int> g_put_variant{-1}, g_put_ctas{-1};
static int put_variant() {
  int v = g_put_variant.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("B200_KV_PUT_VARIANT");
    v = e ? atoi(e) : 0;
    g_put_variant.store(v, std::memory_order_relaxed);
  }
  return v;
}
static int put_default_ctas() {
  int v = g_put_ctas.load(std::memory_order_relaxed);
  if (v < 0) {
    const char* e = getenv("B200_KV_PUT_CTAS");
    v = e ? atoi(e) : 32;
    if (v <= 0) v = 32;
    g_put_ctas.store(v, std::memory_order_relaxed);
  }
  return v;
}
static size_t bulk_smem_bytes(int num_runs) {
  return (size_t)kBulkStages * kTileVecs * 16 + sizeof(BulkShared) + ((size_t)3 * num_runs + 1) * sizeof(int);
}
template <typename K>
static cudaError_t allow_smem(K kernel, size_t bytes, size_t* granted /* [64] per device */) {
  if (bytes <= 48 * 1024) return cudaSuccess;
  int dev = 0;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) return e;
  dev &= 63;
  if (granted[dev] >= bytes) return cudaSuccess;
  const size_t want = 225 * 1024;
  e = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)want);
  if (e == cudaSuccess) granted[dev] = want;
  return e;
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

B200_API int b200_kv_put_configure(int32_t variant, int32_t default_ctas) {
  if (variant >= 0) {
    if (variant > 1) return B200_ERR_ARG;
    g_put_variant.store(variant, std::memory_order_relaxed);
  }
  if (default_ctas > 0) g_put_ctas.store(default_ctas, std::memory_order_relaxed);
  return B200_OK;
}

B200_API int b200_kv_put_config(int32_t* variant, int32_t* default_ctas) {
  if (variant) *variant = put_variant();
  if (default_ctas) *default_ctas = put_default_ctas();
  return B200_OK;
}

B200_API int b200_kv_put_blocks(const int64_t* src_ptrs, const int64_t* dst_ptrs, int32_t num_bufs, const int32_t* runs,
                       int32_t num_runs, int64_t block_bytes, int32_t max_ctas, b200_stream_t stream_) {
  if (num_bufs == 0 || num_runs == 0) return B200_OK;
  if (!src_ptrs || !dst_ptrs || !runs || num_bufs < 0 || num_runs < 0 || block_bytes <= 0) return B200_ERR_ARG;
  if (block_bytes % 16 != 0) return B200_ERR_ALIGN;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int chained = (max_ctas > 0 && (max_ctas & B200_KV_PUT_CHAINED)) ? 1 : 0;
  if (max_ctas > 0) max_ctas &= ~B200_KV_PUT_CHAINED;
  if (max_ctas <= 0) max_ctas = put_default_ctas();
  const bool bulk = put_variant() == 1;
  const long long block_vecs = block_bytes / 16;
  const int unit_vecs = bulk ? put_tile_vecs(block_bytes) : kChunkVecs;
  static size_t granted[64] = {0};
  for (int r0 = 0; r0 < num_runs; r0 += kMaxRuns) {
    PutTable t;
    t.num_runs = std::min(kMaxRuns, num_runs - r0);
    long long units = 0;
    for (int r = 0; r < t.num_runs; ++r) {
      t.start[r] = runs[2 * (r0 + r)];
      t.len[r] = runs[2 * (r0 + r) + 1];
      if (t.start[r] < 0 || t.len[r] < 0) return B200_ERR_ARG;
      units += ((long long)t.len[r] * block_vecs + unit_vecs - 1) / unit_vecs;
    }
    if (units * num_bufs > INT32_MAX) return B200_ERR_ARG;
    t.tile_vecs = unit_vecs;
    t.num_bufs = num_bufs;
    t.block_bytes = block_bytes;
    t.src_ptrs = reinterpret_cast<const long long*>(src_ptrs);
    t.dst_ptrs = reinterpret_cast<const long long*>(dst_ptrs);
    t.chained = chained;
    units *= num_bufs;
    if (units == 0) continue;
    if (bulk) {
      const size_t smem = bulk_smem_bytes(t.num_runs);
      B200_CUDA_TRY(allow_smem(kv_put_blocks_bulk_kernel, smem, granted));
      const int grid = (int)std::min<long long>(units, max_ctas);
      B200_CUDA_TRY(launch_kernel(kv_put_blocks_bulk_kernel, dim3(grid), dim3(kBulkThreads), smem, stream, true, t));
    } else {
      const size_t smem = ((size_t)3 * t.num_runs + 1) * sizeof(int);
      const long long want = (units + kPutThreads / 32 - 1) / (kPutThreads / 32);
      const int grid = (int)std::min<long long>(want, max_ctas);
      B200_CUDA_TRY(launch_kernel(kv_put_blocks_kernel, dim3(grid), dim3(kPutThreads), smem, stream, true, t));
    }
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
  if (tile_vecs != put_tile_vecs(block_bytes)) return B200_ERR_ARG;  // the table was built for another block size
  if (tiles_per_buf * num_bufs > INT32_MAX) return B200_ERR_ARG;
  const size_t tab_bytes = ((size_t)3 * num_runs + 1) * sizeof(int);
  if (tab_bytes > 200 * 1024) return B200_ERR_ARG;
  cudaStream_t stream = static_cast<cudaStream_t>(stream_);
  const int chained = (max_ctas > 0 && (max_ctas & B200_KV_PUT_CHAINED)) ? 1 : 0;
  if (max_ctas > 0) max_ctas &= ~B200_KV_PUT_CHAINED;
  if (max_ctas <= 0) max_ctas = put_default_ctas();
  const long long* sp = reinterpret_cast<const long long*>(src_ptrs);
  const long long* dp = reinterpret_cast<const long long*>(dst_ptrs);
  static size_t granted_bulk[64] = {0}, granted_simt[64] = {0};
  if (put_variant() == 1 && num_runs <= kBulkMaxRuns) {
    const size_t smem = bulk_smem_bytes(num_runs);
    B200_CUDA_TRY(allow_smem(kv_put_blocks_table_bulk_kernel, smem, granted_bulk));
    const int grid = (int)std::min<long long>(tiles_per_buf * num_bufs, max_ctas);
    B200_CUDA_TRY(launch_kernel(kv_put_blocks_table_bulk_kernel, dim3(grid), dim3(kBulkThreads), smem, stream, true,
                                dev_table, num_runs, num_bufs, tile_vecs, (long long)block_bytes, sp, dp, chained));
  } else {
    B200_CUDA_TRY(allow_smem(kv_put_blocks_table_kernel, tab_bytes, granted_simt));
    // chunks per buffer <= (tile_vecs / kChunkVecs + 1) per tile: size the grid from the tile count
    const long long chunks = tiles_per_buf * ((tile_vecs + kChunkVecs - 1) / kChunkVecs) * num_bufs;
    const long long want = (chunks + kPutThreads / 32 - 1) / (kPutThreads / 32);
    const int grid = (int)std::min<long long>(want, max_ctas);
    B200_CUDA_TRY(launch_kernel(kv_put_blocks_table_kernel, dim3(grid), dim3(kPutThreads), tab_bytes, stream, true, dev_table,
                                num_runs, num_bufs, (long long)block_bytes, sp, dp, chained));
  }
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
