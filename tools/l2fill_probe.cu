// Development probe (not part of libb200ops.so): how fast can the SMs be FILLED by TMA, chip-wide, when
//   (a) every CTA streams its own (HBM-resident) weight rows,
//   (b) every CTA re-reads the SAME small (L2-resident) activation block,
//   (c) the shared block is multicast inside a thread-block cluster of 2 / 4 / 8 CTAs?
// This is the question behind the decode-batch FP8 GEMM tiling: at batch 256 the o_proj kernel was
// bound by L2 -> SM fill (99 MB delivered for 17 MB of unique bytes), not by HBM or the tensor pipe.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/l2fill_probe tools/l2fill_probe.cu
// Run:   tools/l2fill_probe      (prints one JSON line per configuration)
#include <cuda.h>
#include <cuda_runtime.h>

#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#define CK(x)                                                                      \
  do {                                                                             \
    cudaError_t e_ = (x);                                                          \
    if (e_ != cudaSuccess) {                                                       \
      fprintf(stderr, "%s:%d %s\n", __FILE__, __LINE__, cudaGetErrorString(e_));   \
      exit(1);                                                                     \
    }                                                                              \
  } while (0)

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void mbar_init(uint32_t bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(bar), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint32_t bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tWAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra DONE;\n\tbra WAIT;\n\tDONE:\n\t}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_2d(uint32_t dst, const void* tm, uint32_t bar, int c0, int c1) {
  asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
               ::"r"(dst), "l"(tm), "r"(bar), "r"(c0), "r"(c1)
               : "memory");
}
__device__ __forceinline__ void tma_2d_mc(uint32_t dst, const void* tm, uint32_t bar, int c0, int c1, uint16_t mask) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes.multicast::cluster"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(tm), "r"(bar), "r"(c0), "r"(c1), "h"(mask)
      : "memory");
}

struct P {
  int iters;         // k-blocks per CTA
  int w_rows;        // unique rows per CTA per k-block (128-byte rows), 0 = none
  int a_rows;        // shared rows per k-block (every CTA needs all of them), 0 = none
  int csz;           // cluster size (1 = unicast of the shared rows)
  int a_kblocks;     // the shared block has this many k-blocks (wraps)
  int w_kblocks;     // K / 128 of the weight matrix
  int stream_rows;   // > 0: pure HBM read test: every (CTA, iteration) reads its own contiguous tile of this many 128-byte rows
  int stages;
  int w_advance;     // 1: after K / 128 k-blocks the CTA moves on to fresh weight rows (no re-read: a true HBM stream of
                     //    [w_rows x 128 B] boxes at row pitch K, the access pattern of a row-major weight matrix)
  int w_halves;      // 2: the weight box is issued as two boxes of w_rows / 2 rows (same bytes, more TMA operations)
};


// stage layout: [w_rows x 128 B][a_rows x 128 B]; one thread issues, the same thread waits (no consumer: the
// question is delivery rate). With multicast each CTA issues a_rows / csz rows to every CTA of the cluster and
// expects the full a_rows on its own barrier.
__global__ void __launch_bounds__(128, 1) fill_kernel(const __grid_constant__ CUtensorMap tmW,
                                                     const __grid_constant__ CUtensorMap tmA,
                                                     const __grid_constant__ CUtensorMap tmAs, const P p) {
  extern __shared__ uint8_t raw[];
  const uint32_t base = (smem_u32(raw) + 1023u) & ~1023u;
  const int STAGES = p.stages;
  const uint32_t stage_bytes = (uint32_t)(p.w_rows + p.a_rows + p.stream_rows) * 128u;
  const uint32_t bars = base + STAGES * stage_bytes;
  if (threadIdx.x == 0) {
    for (int i = 0; i < STAGES; i++) mbar_init(bars + 8u * i, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  __syncthreads();
  if (p.csz > 1) cluster_sync();
  const uint32_t rank = p.csz > 1 ? cluster_ctarank() : 0;
  if (threadIdx.x == 0) {
    const int row0 = blockIdx.x * p.w_rows;
    const int slice = p.csz > 1 ? p.a_rows / p.csz : p.a_rows;
    for (int it = 0; it < p.iters + STAGES; it++) {
      const int st = it % STAGES;
      if (it >= STAGES) mbar_wait(bars + 8u * st, ((it / STAGES) - 1) & 1);
      if (it >= p.iters) continue;
      // with multicast a peer may still be waiting on this stage's previous phase: the cluster runs in lock step
      // through its own barrier waits (every CTA must have consumed phase n before anyone refills), enforced by
      // waiting before issuing above; peers' data for the NEXT phase can only be issued after THEIR wait, which
      // implies they were filled by us. Good enough for a delivery-rate probe.
      const uint32_t dst = base + st * stage_bytes;
      mbar_expect_tx(bars + 8u * st, stage_bytes);
      if (p.stream_rows) tma_2d(dst, &tmAs, bars + 8u * st, 0, (it * (int)gridDim.x + (int)blockIdx.x) * p.stream_rows);
      if (p.w_rows) {
        const int r0 = p.w_advance ? (blockIdx.x + (it / p.w_kblocks) * (int)gridDim.x) * p.w_rows : row0;
        if (p.w_halves == 2) {
          tma_2d(dst, &tmW, bars + 8u * st, (it % p.w_kblocks) * 128, r0);
          tma_2d(dst + (p.w_rows / 2) * 128u, &tmW, bars + 8u * st, (it % p.w_kblocks) * 128, r0 + p.w_rows / 2);
        } else {
          tma_2d(dst, &tmW, bars + 8u * st, (it % p.w_kblocks) * 128, r0);
        }
      }
      if (p.a_rows) {
        const int kb = it % p.a_kblocks;
        if (p.csz > 1)
          tma_2d_mc(dst + p.w_rows * 128u + rank * slice * 128u, &tmAs, bars + 8u * st, kb * 128, (int)rank * slice,
                    (uint16_t)((1u << p.csz) - 1));
        else
          tma_2d(dst + p.w_rows * 128u, &tmA, bars + 8u * st, kb * 128, 0);
      }
    }
  }
  __syncthreads();
  if (p.csz > 1) cluster_sync();
}

typedef CUresult (*PFN_enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                            const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                            CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_enc g_enc;

static CUtensorMap make_map(void* ptr, uint64_t k_bytes, uint64_t rows, uint32_t box_rows) {
  CUtensorMap tm;
  cuuint64_t dims[2] = {k_bytes, rows};
  cuuint64_t strides[1] = {k_bytes};
  cuuint32_t box[2] = {128, box_rows};
  cuuint32_t es[2] = {1, 1};
  CUresult r = g_enc(&tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, ptr, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                     CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    fprintf(stderr, "encode failed %d (rows %llu box %u)\n", (int)r, (unsigned long long)rows, box_rows);
    exit(1);
  }
  return tm;
}

int main() {
  void* fn = nullptr;
  cudaDriverEntryPointQueryResult q;
  CK(cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &fn, cudaEnableDefault, &q));
  g_enc = (PFN_enc)fn;
  cudaDeviceProp prop;
  CK(cudaGetDeviceProperties(&prop, 0));
  const int sms = prop.multiProcessorCount;
  int clk_khz = 0;
  CK(cudaDeviceGetAttribute(&clk_khz, cudaDevAttrClockRate, 0));
  const int K = 4096;                       // bytes per row of both matrices
  const int64_t w_rows_total = 148 * 128 * 64;   // 4.96 GB of "weights": far larger than L2
  uint8_t *w, *a;
  CK(cudaMalloc(&w, (size_t)w_rows_total * K));
  CK(cudaMalloc(&a, (size_t)256 * K));
  CK(cudaMemset(w, 1, (size_t)w_rows_total * K));
  CK(cudaMemset(a, 1, (size_t)256 * K));
  CK(cudaFuncSetAttribute(fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
  CK(cudaFuncSetAttribute(fill_kernel, cudaFuncAttributeNonPortableClusterSizeAllowed, 1));
  cudaEvent_t e0, e1;
  CK(cudaEventCreate(&e0));
  CK(cudaEventCreate(&e1));
  struct Cfg { int ctas, w_rows, a_rows, csz; const char* what; int stream_rows = 0; int stages = 4; int adv = 0; int halves = 1; int kbytes = 0; };
  std::vector<Cfg> cfgs = {
      {sms, 0, 0, 1, "pure HBM read, contiguous 32 KB tiles, 4 stages (128 KB in flight per SM)", 256, 4},
      {sms, 0, 0, 1, "pure HBM read, contiguous 32 KB tiles, 6 stages (192 KB in flight per SM)", 256, 6},
      {sms, 0, 0, 1, "pure HBM read, contiguous 16 KB tiles, 12 stages (192 KB in flight per SM)", 128, 12},
      {sms, 0, 0, 1, "pure HBM read, contiguous 8 KB tiles, 24 stages (192 KB in flight per SM)", 64, 24},
      {sms, 128, 0, 1, "weights only (8x re-read: mostly L2), all SMs"},
      {sms, 128, 0, 1, "ROW-MAJOR WEIGHT STREAM: 128-row x 128 B boxes at pitch 4096, fresh rows (HBM), 148 CTAs, 8 stages", 0, 8, 1},
      {128, 128, 0, 1, "row-major weight stream, pitch 4096, 128 CTAs, 8 stages", 0, 8, 1},
      {64, 128, 0, 1, "row-major weight stream, pitch 4096, 64 CTAs, 8 stages", 0, 8, 1},
      {32, 128, 0, 1, "row-major weight stream, pitch 4096, 32 CTAs, 8 stages", 0, 8, 1},
      {32, 128, 0, 1, "row-major weight stream, pitch 4096, 32 CTAs, 12 stages", 0, 12, 1},
      {sms, 64, 0, 1, "row-major weight stream, 64-row boxes, pitch 4096, 148 CTAs, 16 stages", 0, 16, 1},
      {sms, 128, 0, 1, "row-major weight stream, pitch 4096, 148 CTAs, 8 stages, two 64-row boxes per stage", 0, 8, 1, 2},
      {sms, 128, 0, 1, "row-major weight stream, pitch 14336, 148 CTAs, 8 stages", 0, 8, 1, 1, 14336},
      {32, 128, 0, 1, "row-major weight stream, pitch 14336, 32 CTAs, 8 stages", 0, 8, 1, 1, 14336},
      {sms, 128, 64, 1, "row-major weight stream (fresh rows) + 64 shared rows, pitch 4096, 148 CTAs, 8 stages (v2 BN=64 traffic)", 0, 8, 1},
      {128, 64, 64, 1, "row-major weight stream 64-row boxes + 64 shared rows, 128 CTAs, 12 stages (pair kernel traffic)", 0, 12, 1},
      {sms, 0, 256, 1, "shared 256-row block only, unicast, all SMs"},
      {16, 0, 256, 1, "shared block only, unicast, 16 SMs (per-SM ingest)"},
      {64, 0, 256, 1, "shared block only, unicast, 64 SMs"},
      {sms / 2 * 2, 0, 256, 2, "shared block only, multicast x2"},
      {sms / 4 * 4, 0, 256, 4, "shared block only, multicast x4"},
      {sms / 8 * 8, 0, 256, 8, "shared block only, multicast x8"},
      {sms, 128, 256, 1, "128 weight rows + 256 shared rows, unicast (v2 BN=256)"},
      {sms, 128, 128, 1, "128 weight rows + 128 shared rows, unicast (pair kernel per-SM mix)"},
      {sms, 128, 64, 1, "128 weight rows + 64 shared rows, unicast"},
      {sms / 2 * 2, 128, 256, 2, "128 weight rows + 256 shared rows, multicast x2"},
      {sms / 4 * 4, 128, 256, 4, "128 weight rows + 256 shared rows, multicast x4"},
      {sms / 8 * 8, 128, 256, 8, "128 weight rows + 256 shared rows, multicast x8"},
      {sms / 4 * 4, 128, 128, 4, "128 weight rows + 128 shared rows, multicast x4"},
  };
  for (const Cfg& c : cfgs) {
    P p;
    p.iters = 256;
    p.w_rows = c.w_rows;
    p.a_rows = c.a_rows;
    p.csz = c.csz;
    const int Kc = c.kbytes ? c.kbytes : K;
    p.a_kblocks = K / 128;
    p.w_kblocks = Kc / 128;
    p.w_advance = c.adv;
    p.w_halves = c.halves;
    p.stream_rows = c.stream_rows;
    p.stages = c.stages;
    if (c.stream_rows) p.iters = (int)((int64_t)w_rows_total * K / 128 / c.stream_rows / c.ctas);
    const uint64_t w_rows_c = (uint64_t)w_rows_total * K / Kc;
    CUtensorMap tmW = make_map(w, Kc, w_rows_c, c.w_rows ? c.w_rows / c.halves : 128);
    if (c.adv) p.iters = (4 * 148 / c.ctas) * p.w_kblocks;   // ~300 MB of fresh rows in total: beyond L2 on every repetition
    CUtensorMap tmA = make_map(a, K, 256, c.a_rows ? c.a_rows : 128);
    CUtensorMap tmAs = c.stream_rows ? make_map(w, 128, (uint64_t)w_rows_total * K / 128, c.stream_rows)
                                     : make_map(a, K, 256, c.a_rows ? c.a_rows / c.csz : 128);
    const size_t smem = (size_t)c.stages * (c.w_rows + c.a_rows + c.stream_rows) * 128 + 8 * c.stages + 64 + 1024;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(c.ctas);
    cfg.blockDim = dim3(128);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = c.csz;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    float best = 1e30f;
    for (int rep = 0; rep < 4; rep++) {
      CK(cudaEventRecord(e0));
      CK(cudaLaunchKernelEx(&cfg, fill_kernel, tmW, tmA, tmAs, p));
      CK(cudaEventRecord(e1));
      CK(cudaEventSynchronize(e1));
      float ms;
      CK(cudaEventElapsedTime(&ms, e0, e1));
      if (rep > 0 && ms < best) best = ms;
    }
    const double delivered = (double)c.ctas * p.iters * (c.w_rows + c.a_rows + c.stream_rows) * 128.0;
    const double unique_hbm = (double)c.ctas * p.iters * (c.w_rows + c.stream_rows) * 128.0;
    const double l2_reads = unique_hbm + (double)c.ctas * p.iters * c.a_rows * 128.0 / c.csz;
    printf("{\"what\": \"%s\", \"ctas\": %d, \"cluster\": %d, \"us\": %.2f, \"delivered_GBps\": %.0f, "
           "\"hbm_GBps\": %.0f, \"l2_request_GBps\": %.0f, \"delivered_B_per_clk_at_max_clock\": %.0f, "
           "\"per_sm_B_per_clk\": %.1f}\n",
           c.what, c.ctas, c.csz, best * 1e3, delivered / best / 1e6, unique_hbm / best / 1e6, l2_reads / best / 1e6,
           delivered / (best * 1e-3) / (clk_khz * 1e3), delivered / (best * 1e-3) / (clk_khz * 1e3) / c.ctas);
    fflush(stdout);
  }
  return 0;
}
