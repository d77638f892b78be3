// b200ops -- shared device/host helpers for the sm_100a decode hot path.
// No torch types here: the library is a plain C-ABI shared object (include/b200ops.h).
#pragma once

#include <cuda.h>
#include <cuda_bf16.h>
#include <cuda_fp16.h>
#include <cuda_fp8.h>
#include <cuda_runtime.h>
#include <stdint.h>

#include <cstdarg>
#include <cstdio>

#include "../../include/b200ops.h"

namespace b200 {

// ---------------------------------------------------------------- error plumbing
void set_last_error(const char* fmt, ...);
int cuda_fail(cudaError_t e, const char* what);

#define B200_CHECK(cond, code, ...)        \
  do {                                     \
    if (!(cond)) {                         \
      ::b200::set_last_error(__VA_ARGS__); \
      return (code);                       \
    }                                      \
  } while (0)

#define B200_CUDA(call)                                        \
  do {                                                         \
    cudaError_t _e = (call);                                   \
    if (_e != cudaSuccess) return ::b200::cuda_fail(_e, #call); \
  } while (0)

#define B200_LAUNCH_CHECK() B200_CUDA(cudaPeekAtLastError())

int sm_count();                 // SMs of the current device (cached per device)
int max_smem_optin();           // max dynamic smem per CTA (opt-in)
// programmatic dependent launch between consecutive kernels of a stream (b200_set_pdl / env B200_PDL):
// 0 = off, 1 = the next kernel's prologue overlaps the previous kernel's tail (default), 2 = the GEMMs also
// request their first ring pass of WEIGHT tiles before griddepcontrol.wait.
// The promise level 2 asks of the caller is wider than "the kernel just before": every b200 kernel signals
// launch_dependents at ENTRY, before its own wait, so a chain of programmatic launches can be resident several
// kernels deep and a GEMM's pre-wait prefetch may run beside ANY earlier kernel of that chain. The weight operand
// (and seq_lens for the attention tile plan) must therefore not be written by any kernel launched on the stream
// since the last NON-programmatic launch (a plain <<<>>> kernel, a memcpy / memset node, an event wait, or a
// collective at B200_PDL_AR=2 all end the chain) -- true for model weights, which are only written at load time.
int pdl_level();
bool pdl_enabled();

// Launch with the programmatic-stream-serialization attribute: the kernel may become resident while its
// predecessor on the stream is still draining; everything that touches data produced by (or buffers read by) an
// earlier kernel sits behind pdl_wait() inside the kernel. Captured into CUDA graphs as a programmatic edge.
// Every kernel launched through this helper MUST execute pdl_wait() in at least one thread of every CTA that
// reads or writes global memory, so that completion stays transitive along the stream.
// NVLink collectives and programmatic launch (env B200_PDL_AR): 2 (default) = collectives are launched and behave like
// plain kernels; 0 = they are launched programmatically and let their dependents launch at kernel entry; 1 = only after
// their last barrier. Measured at 2 ranks (profiles/r3e_*.json): with 0 / 1 the step is 1-2 % SLOWER than without any
// programmatic launch, with 2 it is 1.8 % faster (the other kernels still overlap their prologues).
int pdl_collective_mode();

template <typename... KArgs, typename... Args>
static inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                     cudaStream_t stream, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

static inline int dtype_size(int dt) {
  switch (dt) {
    case B200_F32: return 4;
    case B200_F16: return 2;
    case B200_BF16: return 2;
    case B200_FP8_E4M3: return 1;
    case B200_U8: return 1;
    case B200_I32: return 4;
    case B200_I64: return 8;
    default: return 0;
  }
}

// ---------------------------------------------------------------- device helpers
#ifdef __CUDACC__

template <typename T> struct Cvt;
template <> struct Cvt<float> {
  static __device__ __forceinline__ float to_f32(float x) { return x; }
  static __device__ __forceinline__ float from_f32(float x) { return x; }
};
template <> struct Cvt<__half> {
  static __device__ __forceinline__ float to_f32(__half x) { return __half2float(x); }
  static __device__ __forceinline__ __half from_f32(float x) { return __float2half_rn(x); }
};
template <> struct Cvt<__nv_bfloat16> {
  static __device__ __forceinline__ float to_f32(__nv_bfloat16 x) { return __bfloat162float(x); }
  static __device__ __forceinline__ __nv_bfloat16 from_f32(float x) { return __float2bfloat16_rn(x); }
};

// float -> e4m3 (RNE, saturate-to-finite), the conversion the reference uses everywhere
// (__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3); csrc/quantization/w8a8/fp8/nvidia/quant_utils.cuh:492-503)
__device__ __forceinline__ uint8_t f32_to_e4m3(float x) {
  return (uint8_t)__nv_cvt_float_to_fp8(x, __NV_SATFINITE, __NV_E4M3);
}
// two floats -> packed e4m3x2 (lo = a, hi = b)
__device__ __forceinline__ uint16_t f32x2_to_e4m3x2(float a, float b) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(b), "f"(a));
  return r;
}
__device__ __forceinline__ float e4m3_to_f32(uint8_t v) {
  __half_raw h = __nv_cvt_fp8_to_halfraw(v, __NV_E4M3);
  return __half2float(__half(h));
}
// packed e4m3x2 (16 bit) -> f16x2 (32 bit), exact
__device__ __forceinline__ uint32_t e4m3x2_to_f16x2(uint16_t v) {
  uint32_t r;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(r) : "h"(v));
  return r;
}

__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// sum over the CTA (<= 1024 threads); `red` = 32 floats of shared memory. The reduction tree depends
// only on blockDim.x, so two kernels with the same block size produce the same bits.
__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31, nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float r = l < nw ? red[l] : 0.f;
  return warp_sum(r);
}

// one rotary pair: (x, y) <- (x*c - y*s, y*c + x*s), fp32, each result rounded once
// (pos_encoding_kernels.cu:8-40 apply_token_rotary_embedding). Shared by rotary_kernel and the fused
// rope + KV-write kernel so both produce the same bits.
template <typename T>
__device__ __forceinline__ void rope_rotate(T& xr, T& yr, float c, float sn) {
  const float x = Cvt<T>::to_f32(xr), y = Cvt<T>::to_f32(yr);
  xr = Cvt<T>::from_f32(x * c - y * sn);
  yr = Cvt<T>::from_f32(y * c + x * sn);
}

__device__ __forceinline__ uint32_t smem_u32(const void* p) {
  return (uint32_t)__cvta_generic_to_shared(p);
}

// 128-bit streaming loads / stores (read-once data: do not allocate in L1)
__device__ __forceinline__ uint4 ld_nc_v4(const void* p) {
  uint4 r;
  asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p));
  return r;
}
__device__ __forceinline__ void st_na_v4(void* p, uint4 v) {
  asm volatile("st.global.L1::no_allocate.v4.u32 [%0], {%1,%2,%3,%4};" ::"l"(p), "r"(v.x), "r"(v.y),
               "r"(v.z), "r"(v.w));
}

// ---- programmatic dependent launch (griddepcontrol; no-ops when the kernel was launched without the attribute)
// launch_dependents: the next kernel on the stream may start its prologue as soon as SM resources free up;
// wait: blocks until the preceding kernel has completed and its memory operations are visible.
__device__ __forceinline__ void pdl_launch_dependents() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
}
__device__ __forceinline__ void pdl_wait() {
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ---- mbarrier / TMA (sm_90+; used on sm_100a)
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_fence_init() {
  asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void mbar_arrive(uint32_t bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(bar) : "memory");
}
__device__ __forceinline__ void mbar_arrive_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes)
               : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint32_t bar, uint32_t parity) {
  uint32_t ok;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(ok)
      : "r"(bar), "r"(parity)
      : "memory");
  return ok != 0;
}
// Bounded: a barrier that never completes (a lost TMA, a dead peer CTA) traps after ~2^24 polls (seconds)
// instead of hanging the GPU; the first poll is the fast path. Roles that would otherwise poll while the
// PREVIOUS kernel is still running call pdl_wait() first, so the bound only ever covers this kernel's own work.
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  if (mbar_try_wait(bar, parity)) return;
  uint32_t polls = 0;
  while (!mbar_try_wait(bar, parity)) {
    if (++polls > (1u << 24)) __trap();
  }
}
// One lane of a CONVERGED warp. Single-thread roles (TMA producer, tcgen05.mma issuer) must be written as
// warp-uniform loops with `if (elect_one())` around the issue instructions: under a divergent `if (lane == 0)`
// ptxas cannot prove the operands of the uniform-datapath instructions (UTMALDG / UTCQMMA / UTCBAR) uniform and
// wraps EVERY one of them in an ELECT ... BRA.U.ANY loop plus R2UR moves - measured on B200 as ~100 cycles per
// TMA issue and ~65 cycles per MMA issue, which bounded a k-block at ~480 cycles whatever the tile width (that
// trace was lost with its container; with elect.sync the per-k-block trace profiles/r2q_trace.txt shows 52-59 cycles
// from "slot free" to both TMA issues done and 58 cycles for the four MMAs + commit).
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void fence_proxy_async() {
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void prefetch_tmap(const void* tmap) {
  asm volatile("prefetch.tensormap [%0];" ::"l"(tmap) : "memory");
}
__device__ __forceinline__ uint64_t l2_policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_normal() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ uint64_t l2_policy_evict_last() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void tma_load_2d(uint32_t dst, const void* tmap, uint32_t bar, int c0,
                                            int c1, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tma_load_5d(uint32_t dst, const void* tmap, uint32_t bar, int c0,
                                            int c1, int c2, int c3, int c4, uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.5d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2], %8;" ::"r"(dst),
      "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4), "l"(policy)
      : "memory");
}
#endif  // __CUDACC__

// ---------------------------------------------------------------- host TMA descriptor helper
// cuTensorMapEncodeTiled is fetched through cudaGetDriverEntryPoint so the library has no link-time
// dependency on libcuda (it must dlopen on a CPU-only box for the ABI tests).
int encode_tensor_map(CUtensorMap* out, CUtensorMapDataType dt, int rank, const void* base,
                      const uint64_t* dims, const uint64_t* strides_bytes /* rank-1 */,
                      const uint32_t* box, CUtensorMapSwizzle swz, CUtensorMapL2promotion l2);

}  // namespace b200
