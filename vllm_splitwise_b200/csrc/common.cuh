// Shared device/host helpers for libb200kv (sm_100a only).
#pragma once

#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include "../../include/b200_kv.h"

#if defined(__CUDA_ARCH__) && (__CUDA_ARCH__ < 1000)
#error "libb200kv is written for sm_100a (B200) only"
#endif

namespace b200 {

// ---- host-side error plumbing -------------------------------------------------------------
void set_last_cuda_error(cudaError_t e, const char* where);
void count_launch();

#define B200_CUDA_TRY(expr)                            \
  do {                                                 \
    cudaError_t e__ = (expr);                          \
    if (e__ != cudaSuccess) {                          \
      ::b200::set_last_cuda_error(e__, #expr);         \
      (void)cudaGetLastError(); /* not sticky for the next call's launch check */ \
      return B200_ERR_CUDA;                            \
    }                                                  \
  } while (0)

#define B200_LAUNCH_CHECK()                            \
  do {                                                 \
    ::b200::count_launch();                            \
    cudaError_t e__ = cudaPeekAtLastError();           \
    if (e__ != cudaSuccess) {                          \
      ::b200::set_last_cuda_error(e__, "kernel launch"); \
      (void)cudaGetLastError();                        \
      return B200_ERR_CUDA;                            \
    }                                                  \
  } while (0)

inline int dtype_size(int dt) {
  switch (dt) {
    case B200_F32: return 4;
    case B200_F16: return 2;
    case B200_BF16: return 2;
    case B200_U8_E5M2: return 1;
    default: return 0;
  }
}

// RAII device switch for entry points that carry a device ordinal.
struct DeviceGuard {
  int prev = -1;
  bool active = false;
  cudaError_t set(int dev) {
    if (dev < 0) return cudaSuccess;
    cudaError_t e = cudaGetDevice(&prev);
    if (e != cudaSuccess) return e;
    if (prev != dev) {
      e = cudaSetDevice(dev);
      active = (e == cudaSuccess);
    }
    return e;
  }
  ~DeviceGuard() {
    if (active) (void)cudaSetDevice(prev);
  }
};

int num_sms(int device);

// ---- programmatic dependent launch (PDL) ----------------------------------------------------
// The kernels of one decode step (metadata -> RoPE + cache write -> attention) run back to back on one stream.
// Every one of them executes griddepcontrol.wait before it touches memory a predecessor may have written and
// griddepcontrol.launch_dependents as soon as it has started, so a successor launched with the programmatic
// stream-serialization attribute gets its CTAs resident (barrier init, shared-memory carve-up) while the
// predecessor is still running.  Both instructions are no-ops for a kernel launched the ordinary way.
// tl_pdl_launch: set by b200_decode_step around its 2nd and 3rd launch; the attention kernel always uses it.
extern thread_local bool tl_pdl_launch;
bool pdl_disabled();  // B200_NO_PDL=1: every launch is an ordinary one (A/B knob)
struct PdlScope {
  bool prev;
  explicit PdlScope(bool on) : prev(tl_pdl_launch) { tl_pdl_launch = on; }
  ~PdlScope() { tl_pdl_launch = prev; }
};

#ifdef __CUDACC__
// Launch `kern` on `stream`; with pdl the launch carries cudaLaunchAttributeProgrammaticStreamSerialization.
template <typename... KArgs, typename... Args>
inline cudaError_t launch_kernel(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t stream,
                                 bool pdl, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  attr[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = attr;
  cfg.numAttrs = (pdl && !pdl_disabled()) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}
#endif

// ---- device-side PTX helpers --------------------------------------------------------------
#ifdef __CUDACC__

// PDL: order this grid after the memory of the grid it programmatically depends on / let the next grid start.
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
               "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
// Non-blocking: has the phase with this parity completed?  (try_wait may suspend the thread for a while; this does not.)
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(smem_u32(bar)), "r"(parity)
      : "memory");
  return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  while (!mbar_try_wait(bar, parity)) {
  }
}

// L2 eviction policy for streamed-once data (KV blocks are read exactly once per decode step).
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t pol;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}

// 1-D bulk async copy global -> shared (TMA engine, no tensor map), completion on an mbarrier.
// SASS: UBLKCP.  Requires 16-byte aligned src/dst and size % 16 == 0.
__device__ __forceinline__ void bulk_g2s(void* smem_dst, const void* gsrc, uint32_t bytes,
                                         uint64_t* bar, uint64_t policy) {
#ifdef B200_ATTN_NO_HINT
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
      ::"r"(smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar))
      : "memory");
  (void)policy;
  return;
#endif
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint "
      "[%0], [%1], %2, [%3], %4;"
      ::"r"(smem_u32(smem_dst)),
      "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_na_v4(void* p, const uint4& v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x),
               "r"(v.y), "r"(v.z), "r"(v.w)
               : "memory");
}

// 256-bit store (sm_100: STG.E.256): one lane writes a whole 32-byte sector.  p must be 32-byte aligned.
__device__ __forceinline__ void st_v8(void* p, const uint4& lo, const uint4& hi) {
  asm volatile("st.global.v8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"l"(p), "r"(lo.x), "r"(lo.y), "r"(lo.z), "r"(lo.w),
               "r"(hi.x), "r"(hi.y), "r"(hi.z), "r"(hi.w)
               : "memory");
}
// ---- element conversions ------------------------------------------------------------------
// fp8 e5m2 is the high byte of an IEEE half, so widening is a byte permute (exact).
__device__ __forceinline__ float2 e5m2x2_lo_to_float2(uint32_t w) {  // bytes 0,1
  uint32_t h2 = __byte_perm(w, 0u, 0x1404);
  return __half22float2(*reinterpret_cast<const __half2*>(&h2));
}
__device__ __forceinline__ float2 e5m2x2_hi_to_float2(uint32_t w) {  // bytes 2,3
  uint32_t h2 = __byte_perm(w, 0u, 0x3424);
  return __half22float2(*reinterpret_cast<const __half2*>(&h2));
}
__device__ __forceinline__ float2 half2_bits_to_float2(uint32_t w) {
  return __half22float2(*reinterpret_cast<const __half2*>(&w));
}
__device__ __forceinline__ float2 bf162_bits_to_float2(uint32_t w) {
  return make_float2(__uint_as_float(w << 16), __uint_as_float(w & 0xffff0000u));
}

// Unpack one 16-byte vector of cache elements into floats.  N = 16 / sizeof(CT).
template <typename CT>
struct Unpack16;
template <>
struct Unpack16<float> {
  static constexpr int N = 4;
  __device__ __forceinline__ static void run(const uint4& v, float* f) {
    f[0] = __uint_as_float(v.x); f[1] = __uint_as_float(v.y);
    f[2] = __uint_as_float(v.z); f[3] = __uint_as_float(v.w);
  }
};
template <>
struct Unpack16<__half> {
  static constexpr int N = 8;
  __device__ __forceinline__ static void run(const uint4& v, float* f) {
    float2 a = half2_bits_to_float2(v.x), b = half2_bits_to_float2(v.y);
    float2 c = half2_bits_to_float2(v.z), d = half2_bits_to_float2(v.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
  }
};
template <>
struct Unpack16<__nv_bfloat16> {
  static constexpr int N = 8;
  __device__ __forceinline__ static void run(const uint4& v, float* f) {
    float2 a = bf162_bits_to_float2(v.x), b = bf162_bits_to_float2(v.y);
    float2 c = bf162_bits_to_float2(v.z), d = bf162_bits_to_float2(v.w);
    f[0] = a.x; f[1] = a.y; f[2] = b.x; f[3] = b.y; f[4] = c.x; f[5] = c.y; f[6] = d.x; f[7] = d.y;
  }
};
template <>
struct Unpack16<uint8_t> {  // fp8 e5m2
  static constexpr int N = 16;
  __device__ __forceinline__ static void run(const uint4& v, float* f) {
    const uint32_t w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      float2 lo = e5m2x2_lo_to_float2(w[i]), hi = e5m2x2_hi_to_float2(w[i]);
      f[4 * i + 0] = lo.x; f[4 * i + 1] = lo.y; f[4 * i + 2] = hi.x; f[4 * i + 3] = hi.y;
    }
  }
};

template <typename T> __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

// load past L1 (data written by another CTA of the same launch)
template <typename T> __device__ __forceinline__ float ldcg_f32(const T* p);
template <> __device__ __forceinline__ float ldcg_f32<float>(const float* p) { return __ldcg(p); }
template <> __device__ __forceinline__ float ldcg_f32<__half>(const __half* p) {
  return __half2float(__ushort_as_half(__ldcg(reinterpret_cast<const unsigned short*>(p))));
}
template <> __device__ __forceinline__ float ldcg_f32<__nv_bfloat16>(const __nv_bfloat16* p) {
  return __bfloat162float(__ushort_as_bfloat16(__ldcg(reinterpret_cast<const unsigned short*>(p))));
}

template <typename T> __device__ __forceinline__ T from_f32(float v);
template <> __device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ __half from_f32<__half>(float v) { return __float2half_rn(v); }
template <> __device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16_rn(v); }

// Saturating RNE narrow to e5m2 with the SAME intrinsics the reference uses
// (csrc/quantization/fp8_e5m2_kvcache/quant_utils.cuh:164-192) so results are bit-identical.
template <typename T> __device__ __forceinline__ uint8_t to_e5m2(T v);
template <> __device__ __forceinline__ uint8_t to_e5m2<float>(float v) {
  return (uint8_t)__nv_cvt_float_to_fp8(v, __NV_SATFINITE, __NV_E5M2);
}
template <> __device__ __forceinline__ uint8_t to_e5m2<__half>(__half v) {
  return (uint8_t)__nv_cvt_halfraw_to_fp8(static_cast<__half_raw>(v), __NV_SATFINITE, __NV_E5M2);
}
template <> __device__ __forceinline__ uint8_t to_e5m2<__nv_bfloat16>(__nv_bfloat16 v) {
  return (uint8_t)__nv_cvt_bfloat16raw_to_fp8(static_cast<__nv_bfloat16_raw>(v), __NV_SATFINITE, __NV_E5M2);
}
// Two elements per conversion instruction (cvt.rn.satfinite.e5m2x2.*): the same rounding as the single-element
// intrinsics above (those issue the very same instruction with a zero partner), half the instructions.
// Result: low byte = a, high byte = b.
template <typename T> __device__ __forceinline__ uint16_t to_e5m2_x2(T a, T b);
template <> __device__ __forceinline__ uint16_t to_e5m2_x2<float>(float a, float b) {
  return (uint16_t)__nv_cvt_float2_to_fp8x2(make_float2(a, b), __NV_SATFINITE, __NV_E5M2);
}
template <> __device__ __forceinline__ uint16_t to_e5m2_x2<__half>(__half a, __half b) {
  __half2_raw r;
  r.x = static_cast<__half_raw>(a).x;
  r.y = static_cast<__half_raw>(b).x;
  return (uint16_t)__nv_cvt_halfraw2_to_fp8x2(r, __NV_SATFINITE, __NV_E5M2);
}
template <> __device__ __forceinline__ uint16_t to_e5m2_x2<__nv_bfloat16>(__nv_bfloat16 a, __nv_bfloat16 b) {
  // bf16 widens to fp32 exactly, and the fp32 pair converts in ONE instruction (F2FP...E5M2.F32.PACK_AB).  The packed
  // bf16 intrinsic (__nv_cvt_bfloat16raw2_to_fp8x2) converts the halves one at a time and merges the bytes with integer
  // ops: ~9 instructions per pair, which made the bf16 -> fp8 cache write issue-bound (r05 ncu: issue slots 72 % busy).
  return (uint16_t)__nv_cvt_float2_to_fp8x2(make_float2(__bfloat162float(a), __bfloat162float(b)), __NV_SATFINITE, __NV_E5M2);
}
// e5m2 -> T following the reference's widening chain (quant_utils.cuh:23-81):
// fp8 -> half (intrinsic); bf16 goes half -> float -> bf16; float goes half -> float.
template <typename T> __device__ __forceinline__ T from_e5m2(uint8_t v);
template <> __device__ __forceinline__ __half from_e5m2<__half>(uint8_t v) {
  return __half(__nv_cvt_fp8_to_halfraw(v, __NV_E5M2));
}
template <> __device__ __forceinline__ float from_e5m2<float>(uint8_t v) {
  return __half2float(__half(__nv_cvt_fp8_to_halfraw(v, __NV_E5M2)));
}
template <> __device__ __forceinline__ __nv_bfloat16 from_e5m2<__nv_bfloat16>(uint8_t v) {
  return __float2bfloat16(__half2float(__half(__nv_cvt_fp8_to_halfraw(v, __NV_E5M2))));
}

#endif  // __CUDACC__
}  // namespace b200
