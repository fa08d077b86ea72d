// One-shot all-reduce over NVLink peer memory: the tensor-parallel sum that follows the attention
// output projection in every layer.  Replaces the reference's custom all-reduce, 1-stage variant
// (csrc/custom_all_reduce.cuh:215-262 cross_device_reduce_1stage, launcher :420-470; Python side
// vllm/model_executor/parallel_utils/custom_all_reduce.py) for the decode-sized messages it is meant for.
//
// Every rank owns a registered input buffer and two flag arrays, all mapped into its peers (CUDA IPC or peer
// access).  One kernel per call:
//   start  CTA 0 publishes "my input for call #epoch is complete" into every peer's start[rank] (system-scope
//          release after a system fence; the input was written by earlier work on the same stream);
//          every CTA waits until its own start[p] >= epoch for all peers p (acquire, bounded spin);
//   sum    16-byte vectors are read from all ranks' buffers (L1 bypassed: peer lines must not be served stale),
//          accumulated in fp32 in rank order 0..world-1 -- the same order on every rank, so all ranks produce
//          identical bits -- and written to the local output;
//   end    the last CTA to finish publishes done[rank] = epoch to every peer and waits for theirs, so a rank's
//          input buffer is not reused while a peer may still be reading it.
// Epochs are monotonic per communicator; flags never need resetting.  Bound: NVLink latency for decode-sized
// messages, NVLink bandwidth (world-1 remote reads per element) for large ones.
#include "common.cuh"

namespace b200 {
namespace {

constexpr int kMaxRanks = 8;

struct ARParams {
  const void* in[kMaxRanks];                   // rank r's input buffer as seen from this rank
  unsigned long long* peer_start[kMaxRanks];   // &start_of_rank_p[my rank]
  unsigned long long* peer_done[kMaxRanks];    // &done_of_rank_p[my rank]
  const unsigned long long* my_start;          // [world] local
  const unsigned long long* my_done;           // [world] local
  unsigned int* counter;                       // local, zero between calls
  int* timeout;                                // local, set to 1 if a bounded spin expired (or null)
  void* out;
  long long n16;                               // 16-byte vectors
  unsigned long long epoch;
  long long max_spins;
  int rank, world;
};

__device__ __forceinline__ void st_release_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ bool spin_until(const unsigned long long* p, unsigned long long v, long long max_spins) {
  unsigned long long seen;
  unsigned ns = 32;
  for (long long s = 0; s < max_spins; ++s) {
    asm volatile("ld.acquire.sys.global.u64 %0, [%1];" : "=l"(seen) : "l"(p) : "memory");
    if (seen >= v) return true;
    if (s > 2000) {  // the peers are normally within microseconds of each other: poll hard first, back off later
      __nanosleep(ns);
      if (ns < 256) ns <<= 1;
    }
  }
  return false;
}

template <typename T>
struct Vec16;
template <>
struct Vec16<float> {
  static constexpr int N = 4;
  __device__ static void add(float* a, const uint4& v) {
    a[0] += __uint_as_float(v.x); a[1] += __uint_as_float(v.y); a[2] += __uint_as_float(v.z); a[3] += __uint_as_float(v.w);
  }
  __device__ static uint4 pack(const float* a) {
    return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]), __float_as_uint(a[3]));
  }
};
template <typename T>
struct Vec16 {  // __half / __nv_bfloat16
  static constexpr int N = 8;
  __device__ static void add(float* a, const uint4& v) {
    float f[8];
    Unpack16<T>::run(v, f);
#pragma unroll
    for (int i = 0; i < 8; ++i) a[i] += f[i];
  }
  __device__ static uint4 pack(const float* a) {
    uint4 w;
    T* h = reinterpret_cast<T*>(&w);
#pragma unroll
    for (int i = 0; i < 8; ++i) h[i] = from_f32<T>(a[i]);
    return w;
  }
};

template <typename T>
__global__ void __launch_bounds__(512) allreduce_oneshot_kernel(const ARParams p) {
  __shared__ int last_cta;
  const int t = threadIdx.x;
  if (blockIdx.x == 0 && t < p.world) {
    __threadfence_system();
    st_release_sys(p.peer_start[t], p.epoch);
  }
  if (t < p.world && !spin_until(p.my_start + t, p.epoch, p.max_spins) && p.timeout) *p.timeout = 1;
  __syncthreads();

  const long long stride = (long long)gridDim.x * blockDim.x;
  for (long long i = (long long)blockIdx.x * blockDim.x + t; i < p.n16; i += stride) {
    float acc[Vec16<T>::N];
#pragma unroll
    for (int k = 0; k < Vec16<T>::N; ++k) acc[k] = 0.f;
    uint4 v[kMaxRanks];
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      if (r < p.world) v[r] = __ldcg(reinterpret_cast<const uint4*>(p.in[r]) + i);  // all loads in flight at once
#pragma unroll
    for (int r = 0; r < kMaxRanks; ++r)
      if (r < p.world) Vec16<T>::add(acc, v[r]);
    reinterpret_cast<uint4*>(p.out)[i] = Vec16<T>::pack(acc);
  }

  __syncthreads();
  if (t == 0) {
    __threadfence();
    last_cta = atomicAdd(p.counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last_cta) {
    if (t < p.world) {
      st_release_sys(p.peer_done[t], p.epoch);
      if (!spin_until(p.my_done + t, p.epoch, p.max_spins) && p.timeout) *p.timeout = 1;
    }
    if (t == 0) *p.counter = 0;
  }
}

// Two-stage variant for large messages (reference: cross_device_reduce_2stage, custom_all_reduce.cuh:264-330): CTA b
// of every rank owns the same element range; rank r reduces the r-th slice of it into its peer-visible scratch
// (reduce-scatter), the CTAs b of all ranks meet at a per-CTA flag barrier, then every rank gathers the world slices.
// Remote reads per rank: 2*(world-1)/world of the message instead of (world-1) times the message.
constexpr int kMaxCtas2 = 160;

struct AR2Params {
  ARParams a;
  const void* tmp[kMaxRanks];                 // rank r's scratch as seen from this rank
  unsigned long long* peer_mid[kMaxRanks];    // &mid_of_rank_p[CTA 0][my rank]; CTA b is b*world slots further
  const unsigned long long* my_mid;           // [kMaxCtas2][world] local
};

template <typename T>
__global__ void __launch_bounds__(512) allreduce_twostage_kernel(const AR2Params q) {
  const ARParams& p = q.a;
  __shared__ int last_cta;
  const int t = threadIdx.x, b = blockIdx.x;
  if (b == 0 && t < p.world) {
    __threadfence_system();
    st_release_sys(p.peer_start[t], p.epoch);
  }
  if (t < p.world && !spin_until(p.my_start + t, p.epoch, p.max_spins) && p.timeout) *p.timeout = 1;
  __syncthreads();

  const long long per_cta = (p.n16 + gridDim.x - 1) / gridDim.x;
  const long long lo = min((long long)b * per_cta, p.n16), hi = min(lo + per_cta, p.n16);
  const long long slice = (hi - lo + p.world - 1) / p.world;
  {  // stage 1: my slice of this CTA's range, summed over the ranks in rank order
    const long long s_lo = min(lo + p.rank * slice, hi), s_hi = min(s_lo + slice, hi);
    uint4* mine = reinterpret_cast<uint4*>(const_cast<void*>(q.tmp[p.rank]));
    for (long long i = s_lo + t; i < s_hi; i += blockDim.x) {
      float acc[Vec16<T>::N];
#pragma unroll
      for (int k = 0; k < Vec16<T>::N; ++k) acc[k] = 0.f;
      uint4 v[kMaxRanks];
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r)
        if (r < p.world) v[r] = __ldcg(reinterpret_cast<const uint4*>(p.in[r]) + i);
#pragma unroll
      for (int r = 0; r < kMaxRanks; ++r)
        if (r < p.world) Vec16<T>::add(acc, v[r]);
      mine[i] = Vec16<T>::pack(acc);
    }
  }
  __threadfence_system();
  __syncthreads();
  if (t < p.world) {  // CTA b of every rank has written its slice
    st_release_sys(q.peer_mid[t] + (long long)b * p.world, p.epoch);
    if (!spin_until(q.my_mid + (long long)b * p.world + t, p.epoch, p.max_spins) && p.timeout) *p.timeout = 1;
  }
  __syncthreads();
  // stage 2: gather the slices
  for (long long i = lo + t; i < hi; i += blockDim.x) {
    const int r = slice > 0 ? (int)((i - lo) / slice) : 0;
    reinterpret_cast<uint4*>(p.out)[i] = __ldcg(reinterpret_cast<const uint4*>(q.tmp[r]) + i);
  }

  __syncthreads();
  if (t == 0) {
    __threadfence();
    last_cta = atomicAdd(p.counter, 1u) == gridDim.x - 1;
  }
  __syncthreads();
  if (last_cta) {
    if (t < p.world) {
      st_release_sys(p.peer_done[t], p.epoch);
      if (!spin_until(p.my_done + t, p.epoch, p.max_spins) && p.timeout) *p.timeout = 1;
    }
    if (t == 0) *p.counter = 0;
  }
}

}  // namespace
}  // namespace b200

extern "C" int b200_allreduce_twostage(const int64_t* peer_inputs, const int64_t* peer_tmps,
                                       const int64_t* peer_start_slots, const int64_t* peer_done_slots,
                                       const int64_t* peer_mid_slots, const uint64_t* my_start, const uint64_t* my_done,
                                       const uint64_t* my_mid, uint32_t* counter, int32_t* timeout_flag, void* out,
                                       int64_t numel, int32_t dtype, int32_t rank, int32_t world, uint64_t epoch,
                                       int64_t max_spins, b200_stream_t stream) {
  using namespace b200;
  if (numel == 0) return B200_OK;
  if (!peer_inputs || !peer_tmps || !peer_start_slots || !peer_done_slots || !peer_mid_slots || !my_start || !my_done ||
      !my_mid || !counter || !out || world < 1 || world > kMaxRanks || rank < 0 || rank >= world || numel < 0 || epoch == 0)
    return B200_ERR_ARG;
  const int es = dtype_size(dtype);
  if (dtype != B200_F32 && dtype != B200_F16 && dtype != B200_BF16) return B200_ERR_DTYPE;
  if ((numel * es) % 16 || reinterpret_cast<uintptr_t>(out) % 16) return B200_ERR_ALIGN;
  AR2Params q{};
  ARParams& p = q.a;
  for (int r = 0; r < world; ++r) {
    if (peer_inputs[r] % 16 || peer_tmps[r] % 16) return B200_ERR_ALIGN;
    p.in[r] = reinterpret_cast<const void*>(peer_inputs[r]);
    q.tmp[r] = reinterpret_cast<const void*>(peer_tmps[r]);
    p.peer_start[r] = reinterpret_cast<unsigned long long*>(peer_start_slots[r]);
    p.peer_done[r] = reinterpret_cast<unsigned long long*>(peer_done_slots[r]);
    q.peer_mid[r] = reinterpret_cast<unsigned long long*>(peer_mid_slots[r]);
  }
  p.my_start = reinterpret_cast<const unsigned long long*>(my_start);
  p.my_done = reinterpret_cast<const unsigned long long*>(my_done);
  q.my_mid = reinterpret_cast<const unsigned long long*>(my_mid);
  p.counter = counter;
  p.timeout = timeout_flag;
  p.out = out;
  p.n16 = numel * es / 16;
  p.epoch = epoch;
  p.max_spins = max_spins > 0 ? max_spins : 50000000LL;
  p.rank = rank;
  p.world = world;
  const int threads = 512;
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  // the same grid on every rank (CTA b of all ranks meet at a barrier): a function of the message size only,
  // capped by the flag array and by what is resident at once
  int ctas = (int)((p.n16 + 2 * threads - 1) / (2 * threads));
  const int cap = num_sms(dev) < 128 ? num_sms(dev) : 128;
  ctas = ctas < 1 ? 1 : (ctas > cap ? cap : ctas);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200_F32: allreduce_twostage_kernel<float><<<ctas, threads, 0, st>>>(q); break;
    case B200_F16: allreduce_twostage_kernel<__half><<<ctas, threads, 0, st>>>(q); break;
    default: allreduce_twostage_kernel<__nv_bfloat16><<<ctas, threads, 0, st>>>(q); break;
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_allreduce_oneshot(const int64_t* peer_inputs, const int64_t* peer_start_slots,
                                      const int64_t* peer_done_slots, const uint64_t* my_start, const uint64_t* my_done,
                                      uint32_t* counter, int32_t* timeout_flag, void* out, int64_t numel, int32_t dtype,
                                      int32_t rank, int32_t world, uint64_t epoch, int64_t max_spins, int32_t max_ctas,
                                      b200_stream_t stream) {
  using namespace b200;
  if (numel == 0) return B200_OK;
  if (!peer_inputs || !peer_start_slots || !peer_done_slots || !my_start || !my_done || !counter || !out ||
      world < 1 || world > kMaxRanks || rank < 0 || rank >= world || numel < 0 || epoch == 0)
    return B200_ERR_ARG;
  const int es = dtype_size(dtype);
  if (dtype != B200_F32 && dtype != B200_F16 && dtype != B200_BF16) return B200_ERR_DTYPE;
  if ((numel * es) % 16 || reinterpret_cast<uintptr_t>(out) % 16) return B200_ERR_ALIGN;
  ARParams p{};
  for (int r = 0; r < world; ++r) {
    if (peer_inputs[r] % 16) return B200_ERR_ALIGN;
    p.in[r] = reinterpret_cast<const void*>(peer_inputs[r]);
    p.peer_start[r] = reinterpret_cast<unsigned long long*>(peer_start_slots[r]);
    p.peer_done[r] = reinterpret_cast<unsigned long long*>(peer_done_slots[r]);
  }
  p.my_start = reinterpret_cast<const unsigned long long*>(my_start);
  p.my_done = reinterpret_cast<const unsigned long long*>(my_done);
  p.counter = counter;
  p.timeout = timeout_flag;
  p.out = out;
  p.n16 = numel * es / 16;
  p.epoch = epoch;
  p.max_spins = max_spins > 0 ? max_spins : 50000000LL;
  p.rank = rank;
  p.world = world;
  const int threads = 512;
  // every CTA must be resident at once (the start wait sits in front of the work): at most one CTA per SM
  int dev = 0;
  B200_CUDA_TRY(cudaGetDevice(&dev));
  // (the reference caps its grid at 36 blocks, custom_all_reduce.cuh kMaxBlocks; here every thread keeps `world`
  // 16-byte remote loads in flight, and a full machine of them is what it takes to cover the NVLink latency)
  const int cap = max_ctas > 0 ? max_ctas : num_sms(dev);
  int ctas = (int)((p.n16 + 2 * threads - 1) / (2 * threads));
  ctas = ctas < 1 ? 1 : (ctas > cap ? cap : ctas);
  if (ctas > num_sms(dev)) ctas = num_sms(dev);
  cudaStream_t st = static_cast<cudaStream_t>(stream);
  switch (dtype) {
    case B200_F32: allreduce_oneshot_kernel<float><<<ctas, threads, 0, st>>>(p); break;
    case B200_F16: allreduce_oneshot_kernel<__half><<<ctas, threads, 0, st>>>(p); break;
    default: allreduce_oneshot_kernel<__nv_bfloat16><<<ctas, threads, 0, st>>>(p); break;
  }
  B200_LAUNCH_CHECK();
  return B200_OK;
}
