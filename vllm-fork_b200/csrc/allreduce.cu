// b200ops -- tensor-parallel all-reduce over NVLink peer memory (one process per GPU, CUDA IPC).
//
// Contract: torch.ops._C_custom_ar.all_reduce and its helpers (csrc/custom_all_reduce.cuh:7-120
// cross_device_reduce_1stage / 2stage, csrc/custom_collective_common.cuh barriers,
// vllm/distributed/device_communicators/custom_all_reduce.py:56-560): out = sum over ranks of inp,
// fp32 accumulation in rank order 0..N-1 on every rank (bit-identical results on all ranks),
// out-of-place, CUDA-graph capturable, no host synchronisation.
//
// B200 design. Every rank owns one cudaMalloc'ed SYMMETRIC buffer [signals | data heap], opened on all
// peers through CUDA IPC; NVSwitch gives each GPU full bandwidth to every peer, so the kernels simply
// issue 16-byte peer loads.
//   * inputs that live inside the symmetric heap (the decoder allocates its GEMM outputs there) are
//     read in place by the peers: zero copies; other inputs are first staged into the heap by the
//     same kernel;
//   * one-shot (small messages): barrier, every rank reads all N copies and reduces, barrier;
//   * two-shot (large): reduce-scatter of a 1/N slice into the heap, barrier (release/acquire),
//     all-gather of the reduced slices;
//   * fused TP epilogue (csrc/allreduce_fused.cuh): reduce-scatter + residual add + RMSNorm +
//     static FP8 quant + all-gather of the FP8 rows in ONE kernel - the reference's
//     fuse_allreduce_rms target (docs/design/fusions.md:23) done sequence-parallel style, moving
//     (N-1)/N * (2 + 1) bytes per element instead of (N-1)/N * (2 + 2).
// Barriers are flag arrays in peer memory with a per-CTA epoch kept on the device, so a captured
// graph replays without any host-side state.
#include "common.cuh"

#include <cstdlib>
#include <cstring>

namespace b200 {
namespace ar {

constexpr int MAX_RANKS = 8;
constexpr int MAX_BLOCKS = 160;
constexpr int MAX_REGIONS = 16;
constexpr int NTHREADS = 512;

// lives at the start of every rank's symmetric buffer
struct Signal {
  uint32_t arrive[2][MAX_BLOCKS][MAX_RANKS];  // [epoch parity][cta][source rank]
  uint32_t epoch[MAX_BLOCKS];                 // this rank's per-CTA barrier count
};
constexpr size_t SIGNAL_BYTES = (sizeof(Signal) + 1023) / 1024 * 1024;

struct Region {
  uint8_t* base[MAX_RANKS];  // the same buffer as seen from this process, per rank
  size_t bytes;
};

struct Context {
  int rank, world;
  int device;
  Region regions[MAX_REGIONS];
  int num_regions;
  size_t heap_bytes;      // data part of region 0 (after the signals)
  size_t staging_off;     // byte offset (from region 0 base) of the staging area = SIGNAL_BYTES
  size_t staging_bytes;   // the staging area covers the first half of the heap by convention
  uint8_t* mc_base;       // NVSwitch multicast (NVLS) address of region 0, or null (b200_allreduce_set_multicast)
};

struct Peers {
  Signal* sig[MAX_RANKS];
  const uint8_t* in[MAX_RANKS];   // input of every rank (already offset)
  uint8_t* tmp[MAX_RANKS];        // scratch of every rank (two-shot result slices)
  const uint8_t* mc_in;           // multicast address of the input (all ranks' copies at once), or null
  int pdl_mode;                   // pdl_collective_mode(): when this kernel lets its dependents launch
  unsigned long long spin_ns;     // barrier wait limit in ns (spin_timeout_ns()); 0 = wait for ever, like the reference
};

// ---- NVSwitch in-network reduction (NVLS): one load from a multicast address returns the SUM over every rank's
// copy, reduced inside the switch with fp32 accumulation; one store to it lands in every rank's copy. What the
// reference reaches through symm_mem multimem_all_reduce_ (device_communicators/symm_mem.py:127-150).
template <typename T> struct McRed;
template <> struct McRed<float> {
  static __device__ __forceinline__ uint4 ld(const void* p) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.v4.f32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
  }
};
template <> struct McRed<__half> {
  static __device__ __forceinline__ uint4 ld(const void* p) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.f16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
  }
};
template <> struct McRed<__nv_bfloat16> {
  static __device__ __forceinline__ uint4 ld(const void* p) {
    uint4 r;
    asm volatile("multimem.ld_reduce.relaxed.sys.global.add.acc::f32.v4.bf16x2 {%0,%1,%2,%3}, [%4];"
                 : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
  }
};
__device__ __forceinline__ void mc_st_v4(void* p, const uint4& v) {
  asm volatile("multimem.st.relaxed.sys.global.v4.f32 [%0], {%1,%2,%3,%4};" ::"l"(p), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)), "f"(__uint_as_float(v.z)), "f"(__uint_as_float(v.w)) : "memory");
}
__device__ __forceinline__ void mc_st_v2(void* p, const uint2& v) {
  asm volatile("multimem.st.relaxed.sys.global.v2.f32 [%0], {%1,%2};" ::"l"(p), "f"(__uint_as_float(v.x)),
               "f"(__uint_as_float(v.y)) : "memory");
}

__device__ __forceinline__ void st_release_sys(uint32_t* p, uint32_t v) {
  asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_peer_v4(const void* p) {
  uint4 r;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "l"(p)
               : "memory");
  return r;
}

// A dead or wedged peer should not hang the GPU for ever (SURVEY section 5 "never hang"): a flag wait is bounded by
// wall time (%globaltimer, checked every 1024 polls) and traps with a message, which turns a hang into a CUDA error
// on the stream. The limit has to outlast every LEGAL skew between ranks - a peer still compiling, loading weights,
// capturing its graph, stopped in a debugger or in a GC pause - because the trap is a sticky error that takes the
// whole TP group down: default 30 min, env B200_AR_TIMEOUT_S = seconds, 0 = never trap (the reference's behaviour:
// its barriers spin for ever, custom_all_reduce.cuh:236-305). Read once per process, carried in Peers::spin_ns.
inline unsigned long long spin_timeout_ns() {
  static const unsigned long long ns = [] {
    const char* e = getenv("B200_AR_TIMEOUT_S");
    double sec = 1800.0;
    if (e && *e) {
      char* end = nullptr;
      const double v = strtod(e, &end);
      if (end != e && v >= 0.0) sec = v;
    }
    return (unsigned long long)(sec * 1e9);
  }();
  return ns;
}

__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

__device__ __noinline__ void spin_timeout(int rank, int peer, uint32_t want, uint32_t got, unsigned long long ns) {
  printf("b200ops all-reduce: rank %d CTA %d waited > %llu s (B200_AR_TIMEOUT_S) for rank %d (flag %u, want %u): peer "
         "dead or collective call order differs between ranks\n", rank, (int)blockIdx.x, ns / 1000000000ull, peer, got,
         want);
  __trap();
}

// All CTAs with the same blockIdx.x on all ranks meet. `epoch` is the CTA's running barrier count
// (register copy of Signal::epoch[cta]); consecutive barriers alternate between two flag arrays, and
// a rank can be at most one barrier ahead of a peer, so a flag is never overwritten before it is read.
// ORDER = true : release/acquire at system scope - data written by this kernel before the barrier is
//                visible to the peers after it (staged inputs, two-shot slices, pushed rows).
// ORDER = false: pure rendezvous (volatile flag traffic, no fence) - enough when the data was written by
//                an EARLIER kernel (complete at the kernel boundary) or when the barrier only protects
//                reads (end of the all-reduce). Same split as barrier_at_start / barrier_at_end of the
//                reference (custom_collective_common.cuh). A fence costs ~1 us per barrier here.
template <int N, bool ORDER>
__device__ __forceinline__ void barrier(const Peers& pe, int rank, uint32_t& epoch) {
  __syncthreads();  // all of this CTA's prior global accesses are issued (loads have returned)
  epoch++;
  if (threadIdx.x < N) {
    uint32_t* theirs = &pe.sig[threadIdx.x]->arrive[epoch & 1][blockIdx.x][rank];
    const uint32_t* mine = &pe.sig[rank]->arrive[epoch & 1][blockIdx.x][threadIdx.x];
    uint32_t v;
    if constexpr (ORDER) {
      __threadfence_system();
      st_release_sys(theirs, epoch);
    } else {
      asm volatile("st.volatile.global.u32 [%0], %1;" ::"l"(theirs), "r"(epoch) : "memory");
    }
    unsigned long long t0 = 0;
    for (uint32_t polls = 1;; polls++) {
      if constexpr (ORDER) v = ld_acquire_sys(mine);
      else asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(v) : "l"(mine) : "memory");
      if (v == epoch) break;
      if ((polls & 1023u) == 0) {
        const unsigned long long now = global_timer_ns();
        if (t0 == 0) t0 = now;
        else if (pe.spin_ns != 0 && now - t0 > pe.spin_ns) spin_timeout(rank, (int)threadIdx.x, epoch, v, pe.spin_ns);
      }
    }
  }
  __syncthreads();
}

template <typename T> struct Pack;  // 16 bytes of T <-> fp32
template <> struct Pack<float> {
  static constexpr int N = 4;
  static __device__ __forceinline__ void add(float (&a)[8], const uint4& v) {
    const float* f = reinterpret_cast<const float*>(&v);
#pragma unroll
    for (int i = 0; i < 4; i++) a[i] += f[i];
  }
  static __device__ __forceinline__ uint4 pack(const float (&a)[8]) {
    return make_uint4(__float_as_uint(a[0]), __float_as_uint(a[1]), __float_as_uint(a[2]), __float_as_uint(a[3]));
  }
};
template <> struct Pack<__half> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void add(float (&a)[8], const uint4& v) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float2 t = __half22float2(h[i]);
      a[2 * i] += t.x; a[2 * i + 1] += t.y;
    }
  }
  static __device__ __forceinline__ uint4 pack(const float (&a)[8]) {
    uint4 r;
    __half2* h = reinterpret_cast<__half2*>(&r);
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = __floats2half2_rn(a[2 * i], a[2 * i + 1]);
    return r;
  }
};
template <> struct Pack<__nv_bfloat16> {
  static constexpr int N = 8;
  static __device__ __forceinline__ void add(float (&a)[8], const uint4& v) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      const float2 t = __bfloat1622float2(h[i]);
      a[2 * i] += t.x; a[2 * i + 1] += t.y;
    }
  }
  static __device__ __forceinline__ uint4 pack(const float (&a)[8]) {
    uint4 r;
    __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&r);
#pragma unroll
    for (int i = 0; i < 4; i++) h[i] = __floats2bfloat162_rn(a[2 * i], a[2 * i + 1]);
    return r;
  }
};

// sum of pack `idx` over all ranks, in rank order 0..N-1 (same order on every rank)
template <typename T, int N>
__device__ __forceinline__ uint4 reduce_pack(const Peers& pe, int64_t idx) {
  uint4 v[N];
#pragma unroll
  for (int r = 0; r < N; r++) v[r] = ld_peer_v4(pe.in[r] + idx * 16);
  float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int r = 0; r < N; r++) Pack<T>::add(a, v[r]);
  return Pack<T>::pack(a);
}

// If `stage_src` is set the input is outside the symmetric heap: copy it into this rank's staging
// area (pe.in[rank]) first; the start barrier then publishes it. The barrier only orders CTAs with the
// SAME blockIdx.x across ranks, so CTA b must stage exactly the packs the peers' CTA b will read: for
// every slice [s*part, (s+1)*part) the packs s*part + tid + k*stride - the index walk of reduce_range
// (one-shot: one slice = the whole message; two-shot: one slice per rank).
template <int N>
__device__ __forceinline__ void stage_input(const Peers& pe, int rank, const uint8_t* stage_src, int64_t packs,
                                            int64_t part) {
  if (stage_src) {
    uint4* dst = reinterpret_cast<uint4*>(const_cast<uint8_t*>(pe.in[rank]));
    const uint4* src = reinterpret_cast<const uint4*>(stage_src);
    const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    const int64_t stride = (int64_t)gridDim.x * blockDim.x;
    for (int64_t lo = 0; lo < packs; lo += part) {
      const int64_t hi = min(packs, lo + part);
      for (int64_t i = lo + tid; i < hi; i += stride) dst[i] = src[i];
    }
  }
}

// Row-wise variant for the fused kernel: its CTA b reads whole rows row_lo(s) + b + k*gridDim.x of slice s.
template <int N>
__device__ __forceinline__ void stage_rows(const Peers& pe, int rank, const uint8_t* stage_src, int M, int nv,
                                           int rows_per_rank) {
  if (stage_src) {
    uint4* dst = reinterpret_cast<uint4*>(const_cast<uint8_t*>(pe.in[rank]));
    const uint4* src = reinterpret_cast<const uint4*>(stage_src);
    for (int lo = 0; lo < M; lo += rows_per_rank) {
      const int hi = min(M, lo + rows_per_rank);
      for (int row = lo + blockIdx.x; row < hi; row += gridDim.x)
        for (int v = threadIdx.x; v < nv; v += blockDim.x) dst[(int64_t)row * nv + v] = src[(int64_t)row * nv + v];
    }
  }
}

// `U` packs per thread per trip: all U*N peer loads are issued before the first add, so a trip costs one
// NVLink round trip instead of U (the kernels are latency-bound: 2 MiB is 2.3 us of NVLink bandwidth).
template <typename T, int N, int U, typename F>
__device__ __forceinline__ void reduce_range(const Peers& pe, int64_t lo, int64_t hi, F&& store) {
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  for (int64_t i0 = lo + tid; i0 < hi; i0 += stride * U) {
    uint4 v[U][N];
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t i = i0 + u * stride;
      if (i < hi) {
#pragma unroll
        for (int r = 0; r < N; r++) v[u][r] = ld_peer_v4(pe.in[r] + i * 16);
      }
    }
#pragma unroll
    for (int u = 0; u < U; u++) {
      const int64_t i = i0 + u * stride;
      if (i < hi) {
        float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
        for (int r = 0; r < N; r++) Pack<T>::add(a, v[u][r]);
        store(i, Pack<T>::pack(a));
      }
    }
  }
}

template <typename T, int N>
__global__ void __launch_bounds__(NTHREADS, 1)
    allreduce_one_shot_kernel(Peers pe, uint8_t* __restrict__ out, const uint8_t* stage_src, int rank,
                              int64_t packs, int skip_end) {
  if (pe.pdl_mode == 0) pdl_launch_dependents();
  pdl_wait();   // the input (and this context's epoch words) come from earlier kernels on the stream
  uint32_t epoch = pe.sig[rank]->epoch[blockIdx.x];
  stage_input<N>(pe, rank, stage_src, packs, packs);
  if (stage_src) barrier<N, true>(pe, rank, epoch);
  else barrier<N, false>(pe, rank, epoch);
  reduce_range<T, N, (N <= 4 ? 4 : 2)>(pe, 0, packs, [&](int64_t i, uint4 v) { reinterpret_cast<uint4*>(out)[i] = v; });
  // peers may overwrite their inputs only after everyone has read them. skip_end: the caller promises
  // not to touch the input before the NEXT collective of this context (e.g. alternating buffers), whose
  // start barrier then plays this role. A STAGED input lives in the library-owned staging area, which the
  // next collective overwrites before its start barrier: the promise cannot cover it, so never skip then.
  if (!skip_end || stage_src) barrier<N, false>(pe, rank, epoch);
  if (threadIdx.x == 0) pe.sig[rank]->epoch[blockIdx.x] = epoch;
  if (pe.pdl_mode == 1) pdl_launch_dependents();
}

template <typename T, int N>
__global__ void __launch_bounds__(NTHREADS, 1)
    allreduce_two_shot_kernel(Peers pe, uint8_t* __restrict__ out, const uint8_t* stage_src, int rank,
                              int64_t packs, int skip_end) {
  if (pe.pdl_mode == 0) pdl_launch_dependents();
  pdl_wait();   // the input (and this context's epoch words) come from earlier kernels on the stream
  uint32_t epoch = pe.sig[rank]->epoch[blockIdx.x];
  const int64_t part = (packs + N - 1) / N;          // packs per rank slice
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  stage_input<N>(pe, rank, stage_src, packs, part);
  if (stage_src) barrier<N, true>(pe, rank, epoch);
  else barrier<N, false>(pe, rank, epoch);
  {  // stage 1: reduce my slice into my scratch
    const int64_t lo = rank * part, hi = min(packs, lo + part);
    uint4* mine = reinterpret_cast<uint4*>(pe.tmp[rank]);
    reduce_range<T, N, 2>(pe, lo, hi, [&](int64_t i, uint4 v) { mine[i - lo] = v; });
  }
  barrier<N, true>(pe, rank, epoch);
  // stage 2: gather every rank's reduced slice (the barrier above is release/acquire at system scope)
#pragma unroll
  for (int r = 0; r < N; r++) {
    const int src = (rank + r) % N;  // start with the local slice, spread the peers
    const int64_t lo = src * part, hi = min(packs, lo + part);
    const uint4* theirs = reinterpret_cast<const uint4*>(pe.tmp[src]);
    for (int64_t i = lo + tid; i < hi; i += stride) reinterpret_cast<uint4*>(out)[i] = ld_peer_v4(theirs + (i - lo));
  }
  // the scratch slices are re-written by the next two-shot call: always rendezvous here
  (void)skip_end;
  barrier<N, false>(pe, rank, epoch);
  if (threadIdx.x == 0) pe.sig[rank]->epoch[blockIdx.x] = epoch;
  if (pe.pdl_mode == 1) pdl_launch_dependents();
}

// NVLS all-reduce. mc_out != null (output inside the symmetric buffer): two-shot in the switch - rank r reduces
// slice r with multimem.ld_reduce and broadcasts it with multimem.st, so every rank ends with the SAME bits and
// receives the message once ((N-1)/N * bytes through its NVLink instead of (N-1) * bytes for the peer-pull
// one-shot). mc_out == null: every rank reduces the whole message into its private output (still one message of
// receive traffic). The switch's accumulation order is not the rank order of the peer-pull kernels: results agree
// within the rounding of one fp32 accumulation (bit-exact for exactly representable sums).
template <typename T, int N>
__global__ void __launch_bounds__(NTHREADS, 1)
    allreduce_multimem_kernel(Peers pe, uint8_t* __restrict__ out, uint8_t* mc_out, const uint8_t* stage_src, int rank,
                              int64_t packs) {
  if (pe.pdl_mode == 0) pdl_launch_dependents();
  pdl_wait();   // the input (and this context's epoch words) come from earlier kernels on the stream
  uint32_t epoch = pe.sig[rank]->epoch[blockIdx.x];
  const int64_t part = (packs + N - 1) / N;
  stage_input<N>(pe, rank, stage_src, packs, mc_out ? part : packs);
  if (stage_src) barrier<N, true>(pe, rank, epoch);
  else barrier<N, false>(pe, rank, epoch);
  const int64_t tid = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
  const int64_t stride = (int64_t)gridDim.x * blockDim.x;
  if (mc_out) {
    const int64_t lo = rank * part, hi = min(packs, lo + part);
    for (int64_t i = lo + tid; i < hi; i += stride) mc_st_v4(mc_out + i * 16, McRed<T>::ld(pe.mc_in + i * 16));
    barrier<N, true>(pe, rank, epoch);   // every rank's broadcast slice has landed everywhere
  } else {
    for (int64_t i = tid; i < packs; i += stride) reinterpret_cast<uint4*>(out)[i] = McRed<T>::ld(pe.mc_in + i * 16);
    barrier<N, false>(pe, rank, epoch);  // inputs may be overwritten only after everyone has reduced them
  }
  if (threadIdx.x == 0) pe.sig[rank]->epoch[blockIdx.x] = epoch;
  if (pe.pdl_mode == 1) pdl_launch_dependents();
}

}  // namespace ar
}  // namespace b200

#include "allreduce_fused.cuh"
#include "dcp_merge.cuh"
#include "allreduce_host.cuh"
