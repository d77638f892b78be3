// b200ops -- FP8 (E4M3) W8A8 GEMM for decode batches with few weight tiles: split-K across the CTA pairs of
// one thread-block cluster, partial sums exchanged through distributed shared memory (sm_100a).
//
// Contract: identical to gemm_fp8.cu (torch.ops._C.cutlass_scaled_mm,
// csrc/libtorch_stable/quantization/w8a8/cutlass/scaled_mm_entry.cu:197-270), M <= 256 only:
//   out[M,N] = a_scale (.) ( b_scale (.) (A[M,K] . B[K,N]) ) (+ bias[N]),  fp32 accumulate.
//
// Why (measured on B200): at batch 256 the qkv / o / down projections of an 8B model
// have 16-24 weight tiles of 256 rows, far fewer than the 74 CTA pairs of the chip, and K of 4096-14336.
// The one-CTA kernel answers with narrow batch tiles (every weight byte re-read from L2 by 4 CTAs, 24 KB of
// operands per 128 MMA cycles); stream-K through global memory (gemm_fp8_v3.cu) pays a 128 KB fp32 round trip
// per CTA through L2 plus flags and fences. Here the K range of one 256 x BN tile is cut over the SPLIT pairs
// of a cluster (2 * SPLIT CTAs); every pair runs tcgen05.mma.cta_group::2 over its k-blocks, then the pairs do
// a reduce-scatter over DSMEM: pair s owns the 16-column chunks c with c % SPLIT == s, the other pairs write
// their fp32 chunks straight into the owner's (now idle) pipeline shared memory with st.shared::cluster, one
// cluster barrier, and the owner adds the SPLIT partials in pair order (deterministic), scales, converts and
// writes its chunks with TMA stores. No global workspace, no flags, no atomics.
// Roles per CTA: warp 0 = TMA producer, warp 1 = MMA issuer (even CTA of the pair) + TMEM owner,
// warps 2-5 = epilogue.
#include "tcgen05.cuh"

#include <cmath>
#include <cstring>

namespace b200 {
namespace gemm8v4 {

constexpr int BM = 128;          // rows per CTA (UMMA M = 256 per pair)
constexpr int BK = 128;          // k-block bytes
constexpr int UMMA_K = 32;
constexpr int NTHREADS = 192;
constexpr int MAX_STAGES = 8;
constexpr int MAX_SPLIT = 4;
constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;  // clears the low CTA-rank bit of a shared::cluster address
constexpr uint32_t CHUNK_BYTES = BM * 16 * 4;   // one 16-column fp32 chunk of a CTA's 128 rows
constexpr uint32_t STG_CHUNK_BYTES = 16 * BM * 2;  // the same chunk as [16 m][128 r] 16-bit output

struct Params {
  const float* r_scale;   // weight scale (per-row vector or scalar)
  const float* c_scale;   // activation scale (per-token vector or scalar)
  const void* bias;       // [N] in out dtype, nullable
  int R, C, K;            // weight rows (N), batch (M), reduction
  int r_scale_vec, c_scale_vec;
  int out_bf16, stages, num_kb, split;
  int prefetch_w;         // PDL level 2: weight tiles of the first ring pass precede griddepcontrol.wait
};

using namespace b200::tc;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ void tma_load_2d_2sm(uint32_t dst, const void* tmap, uint32_t bar, int c0, int c1,
                                                uint64_t policy) {
  asm volatile(
      "cp.async.bulk.tensor.2d.cta_group::2.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint"
      " [%0], [%1, {%3, %4}], [%2], %5;" ::"r"(dst),
      "l"(tmap), "r"(bar & PEER_MASK), "r"(c0), "r"(c1), "l"(policy)
      : "memory");
}
__device__ __forceinline__ void tc_mma_f8_2sm(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc,
                                              uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::2.kind::f8f6f4 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
// arrive on the barrier at this smem offset in both CTAs of the pair (cluster ranks 2s, 2s + 1) when the
// prior MMAs complete
__device__ __forceinline__ void tc_commit_2sm(uint32_t bar, uint16_t mask) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"(mask)
      : "memory");
}
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}
__device__ __forceinline__ void st_cluster_v4(uint32_t addr, float a, float b, float c, float d) {
  asm volatile("st.shared::cluster.v4.f32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "f"(a), "f"(b), "f"(c), "f"(d)
               : "memory");
}

template <int BN>
__global__ void __launch_bounds__(NTHREADS, 1)
    gemm_fp8_v4_kernel(const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmC,
                       const __grid_constant__ CUtensorMap tmO, const Params p) {
  constexpr uint32_t R_BYTES = BM * BK;             // this CTA's 128 weight rows
  constexpr uint32_t C_BYTES = (BN / 2) * BK;       // this CTA's half of the batch operand
  constexpr uint32_t STAGE_BYTES = R_BYTES + C_BYTES;
  constexpr uint32_t TMEM_COLS = BN < 32 ? 32 : BN;
  constexpr int NCHUNK = BN / 16;
  // instruction descriptor: D = F32, A = B = E4M3, K-major, N = BN, M = 256 (pair)
  constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t base = (raw_base + 1023u) & ~1023u;   // pipeline stages; after barrier A: receive slots
  uint8_t* smem = smem_raw + (base - raw_base);
  const int stages = p.stages;
  const int split = p.split;
  const int jmax = (NCHUNK + split - 1) / split;      // chunks owned per pair
  const uint32_t staging = base + (uint32_t)stages * STAGE_BYTES;          // [jmax][16 m][128 r] 16-bit
  const uint32_t bar_base = staging + (uint32_t)jmax * STG_CHUNK_BYTES;
  const uint32_t bar_full = bar_base;                          // [stages]   (the even CTA's copy is used)
  const uint32_t bar_empty = bar_base + 8u * MAX_STAGES;       // [stages]   per CTA
  const uint32_t bar_tfull = bar_base + 16u * MAX_STAGES;      // per CTA
  volatile uint32_t* tmem_holder = reinterpret_cast<volatile uint32_t*>(
      smem + (size_t)stages * STAGE_BYTES + (size_t)jmax * STG_CHUNK_BYTES + 16 * MAX_STAGES + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t rank = cluster_ctarank();
  const int s = (int)(rank >> 1);          // pair inside the cluster = K slice
  const uint32_t h = rank & 1u;            // half of the 256-row tile
  const int tile = blockIdx.x / (2 * split);
  const int kb0 = p.num_kb * s / split, kb1 = p.num_kb * (s + 1) / split;
  const int row0 = tile * 2 * BM + (int)h * BM;   // this CTA's first weight row
  pdl_launch_dependents();

  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; i++) {
      mbar_init(bar_full + 8u * i, 1);
      mbar_init(bar_empty + 8u * i, 1);
    }
    mbar_init(bar_tfull, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    const uint32_t holder = smem_u32((const void*)tmem_holder);
    asm volatile("tcgen05.alloc.cta_group::2.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(holder), "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::2.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();   // every CTA's barriers are initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ============================== TMA producer (warp-uniform loop, one elected lane issues: common.cuh elect_one)
    prefetch_tmap(&tmR);
    prefetch_tmap(&tmC);
    const uint64_t pol_w = l2_policy_evict_first();   // weights stream once
    const uint64_t pol_a = l2_policy_evict_last();    // the batch is re-read by every cluster
    // Programmatic dependent launch: weight tiles of the first ring pass before griddepcontrol.wait (level 2),
    // the activation tiles (the previous kernel's output) after it
    const int nkb = kb1 - kb0;
    const int pre = nkb < stages ? nkb : stages;
    if (!p.prefetch_w) pdl_wait();
    if (elect_one()) {
      for (int i = 0; i < pre; i++) {
        if (h == 0) mbar_arrive_expect_tx(bar_full + 8u * i, 2 * STAGE_BYTES);
        tma_load_2d_2sm(base + (uint32_t)i * STAGE_BYTES, &tmR, bar_full + 8u * i, (kb0 + i) * BK, row0, pol_w);
      }
    }
    __syncwarp();
    if (p.prefetch_w) pdl_wait();
    if (elect_one()) {
      for (int i = 0; i < pre; i++)
        tma_load_2d_2sm(base + (uint32_t)i * STAGE_BYTES + R_BYTES, &tmC, bar_full + 8u * i, (kb0 + i) * BK,
                        (int)h * (BN / 2), pol_a);
    }
    __syncwarp();
    int stage = pre == stages ? 0 : pre;
    uint32_t phase = pre == stages ? 1u : 0u;
    for (int kb = kb0 + pre; kb < kb1; kb++) {
      mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
      if (elect_one()) {
        const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES;
        if (h == 0) mbar_arrive_expect_tx(bar_full + 8u * stage, 2 * STAGE_BYTES);
        tma_load_2d_2sm(dst, &tmR, bar_full + 8u * stage, kb * BK, row0, pol_w);
        tma_load_2d_2sm(dst + R_BYTES, &tmC, bar_full + 8u * stage, kb * BK, (int)h * (BN / 2), pol_a);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer: one elected lane of the even CTA drives both tensor cores
    pdl_wait();   // nothing to issue before the previous kernel's output arrives: block instead of polling
    if (h == 0) {
      const uint16_t pair_mask = (uint16_t)(3u << (2 * s));
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = kb0; kb < kb1; kb++) {
        mbar_wait(bar_full + 8u * stage, phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = base + (uint32_t)stage * STAGE_BYTES;
          const uint64_t da = make_smem_desc(sa);
          const uint64_t db = make_smem_desc(sa + R_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; k++)
            tc_mma_f8_2sm(tmem_base, da + (uint64_t)(k * UMMA_K >> 4), db + (uint64_t)(k * UMMA_K >> 4), IDESC,
                          (kb > kb0 || k > 0) ? 1u : 0u);
          tc_commit_2sm(bar_empty + 8u * stage, pair_mask);
          if (kb == kb1 - 1) tc_commit_2sm(bar_tfull, pair_mask);
        }
        __syncwarp();
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
    }
    __syncwarp();
  } else {
    pdl_wait();                // scales / bias / `out` are shared with the previous kernel on the stream
    mbar_wait(bar_tfull, 0);   // this pair's partial accumulator is complete, its pipeline smem idle
    tc_fence_after();
  }

  // ---- barrier A: every pair of the cluster is past its main loop, so pipeline smem can receive partials
  cluster_sync();

  const int wq = warp & 3;
  const int et = (warp - 2) * 32 + lane;    // 0..127 among the epilogue threads
  const int trow = wq * 32 + lane;          // row of this thread inside the CTA's 128
  const uint32_t tbase = tmem_base + ((uint32_t)(wq * 32) << 16);
  const int nchunks = (min(BN, p.C) + 15) >> 4;
  // receive slot of (source pair sp, owned chunk j, float4 plane pl, row): [sp][j][pl][row] x 16 bytes
  auto slot = [&](int sp, int j, int pl) -> uint32_t {
    return base + (uint32_t)((((sp * jmax + j) * 4 + pl) * BM + trow) * 16);
  };
  if (warp >= 2) {
    // ---- reduce-scatter, send side: chunk c belongs to pair c % split. Four TMEM loads in flight per wait.
    for (int c0 = 0; c0 < nchunks; c0 += 4) {
      uint32_t v[4][16];
#pragma unroll
      for (int i = 0; i < 4; i++)
        if (c0 + i < nchunks && (c0 + i) % split != s) tc_ld16(tbase + (uint32_t)((c0 + i) << 4), v[i]);
      tc_wait_ld();
#pragma unroll
      for (int i = 0; i < 4; i++) {
        const int c = c0 + i;
        if (c >= nchunks || c % split == s) continue;
        const uint32_t remote = mapa(slot(s, c / split, 0), (uint32_t)(2 * (c % split)) + h);
#pragma unroll
        for (int pl = 0; pl < 4; pl++)
          st_cluster_v4(remote + (uint32_t)(pl * BM * 16), __uint_as_float(v[i][4 * pl]),
                        __uint_as_float(v[i][4 * pl + 1]), __uint_as_float(v[i][4 * pl + 2]),
                        __uint_as_float(v[i][4 * pl + 3]));
      }
    }
  }

  // ---- barrier B: all partial chunks have landed (release / acquire at cluster scope)
  cluster_sync();

  if (warp >= 2) {
    const uint16_t* bias16 = reinterpret_cast<const uint16_t*>(p.bias);
    const int r = row0 + trow;
    const bool r_ok = r < p.R;
    const float rs = p.r_scale_vec ? (r_ok ? p.r_scale[r] : 0.f) : p.r_scale[0];
    const float cs0 = p.c_scale_vec ? 0.f : p.c_scale[0];
    float bias_r = 0.f;
    if (bias16 && r_ok)
      bias_r = p.out_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(bias16)[r])
                          : __half2float(reinterpret_cast<const __half*>(bias16)[r]);
    int nown = 0;
    uint32_t vnext[16];   // this pair's own partial of the NEXT owned chunk (TMEM load in flight during the adds)
    if (s < nchunks) tc_ld16(tbase + (uint32_t)(s << 4), vnext);
#pragma unroll 1
    for (int c = s; c < nchunks; c += split, nown++) {
      const int j = c / split;
      float own[16];
      tc_wait_ld();
#pragma unroll
      for (int e = 0; e < 16; e++) own[e] = __uint_as_float(vnext[e]);
      if (c + split < nchunks) tc_ld16(tbase + (uint32_t)((c + split) << 4), vnext);
      // sum the split partials in pair order (this pair's own one comes from TMEM)
      float a[16];
#pragma unroll
      for (int e = 0; e < 16; e++) a[e] = 0.f;
#pragma unroll 1
      for (int sp = 0; sp < split; sp++) {
        if (sp == s) {
#pragma unroll
          for (int e = 0; e < 16; e++) a[e] = (sp == 0) ? own[e] : a[e] + own[e];
        } else {
#pragma unroll
          for (int pl = 0; pl < 4; pl++) {
            float4 t;
            asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                         : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w)
                         : "r"(slot(sp, j, pl)));
            if (sp == 0) {
              a[4 * pl] = t.x; a[4 * pl + 1] = t.y; a[4 * pl + 2] = t.z; a[4 * pl + 3] = t.w;
            } else {
              a[4 * pl] += t.x; a[4 * pl + 1] += t.y; a[4 * pl + 2] += t.z; a[4 * pl + 3] += t.w;
            }
          }
        }
      }
      // out[m][n] = a_s[m] * (b_s[n] * acc) + bias[n], staged as [16 m][128 n] for one TMA store per chunk
      const int cbase = c << 4;
#pragma unroll
      for (int e = 0; e < 16; e += 2) {
        const float cs_a = p.c_scale_vec ? (cbase + e < p.C ? p.c_scale[cbase + e] : 0.f) : cs0;
        const float cs_b = p.c_scale_vec ? (cbase + e + 1 < p.C ? p.c_scale[cbase + e + 1] : 0.f) : cs0;
        const float oa = cs_a * (rs * a[e]) + bias_r;
        const float ob = cs_b * (rs * a[e + 1]) + bias_r;
        // 2x2 transpose between lane pairs: even lane keeps column e of rows (n, n+1), odd lane column e+1
        const float send = (lane & 1) ? oa : ob;
        const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
        const float lo_ = (lane & 1) ? recv : oa;
        const float hi_ = (lane & 1) ? ob : recv;
        uint32_t packed;
        if (p.out_bf16) {
          __nv_bfloat162 t = __floats2bfloat162_rn(lo_, hi_);
          packed = *reinterpret_cast<uint32_t*>(&t);
        } else {
          __half2 t = __floats2half2_rn(lo_, hi_);
          packed = *reinterpret_cast<uint32_t*>(&t);
        }
        const int m_local = e + (lane & 1);
        const uint32_t addr = staging + (uint32_t)j * STG_CHUNK_BYTES + (uint32_t)m_local * (BM * 2) +
                              (uint32_t)(wq * 32 + (lane & ~1)) * 2u;
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(packed) : "memory");
      }
    }
    fence_proxy_async();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (et == 0) {
      for (int j = 0; j < nown; j++) {
        const int c = s + j * split;
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tmO),
                     "r"(row0), "r"(c << 4), "r"(staging + (uint32_t)j * STG_CHUNK_BYTES)
                     : "memory");
      }
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
      asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();   // peers may still read this CTA's barriers / write its receive slots until here
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

static int make_tmap_2d(CUtensorMap* tm, const void* base, int rows, int K, int64_t ld_bytes, int box_rows) {
  uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld_bytes};
  uint32_t box[2] = {(uint32_t)BK, (uint32_t)box_rows};
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

template <int BN>
static int smem_plan(int split, int* stages_out) {
  constexpr int STAGE_BYTES = BM * BK + (BN / 2) * BK;
  const int jmax = (BN / 16 + split - 1) / split;
  const int fixed = 1024 + jmax * (int)STG_CHUNK_BYTES + 16 * MAX_STAGES + 64;
  int stages = (max_smem_optin() - fixed) / STAGE_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  // the receive slots of the reduce-scatter ([split][jmax] chunks) reuse the pipeline stages
  if (stages < 2 || (long long)stages * STAGE_BYTES < (long long)split * jmax * (int)CHUNK_BYTES) return -1;
  *stages_out = stages;
  return stages * STAGE_BYTES + fixed;
}

// largest number of co-resident clusters of 2 * split CTAs (cached per BN / split)
template <int BN>
static int max_clusters(int split, int smem) {
  static int cache[MAX_SPLIT + 1] = {0, 0, 0, 0, 0};
  if (cache[split] == 0) {
    cudaFuncSetAttribute(gemm_fp8_v4_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_smem_optin());
    cudaFuncSetAttribute(gemm_fp8_v4_kernel<BN>, cudaFuncAttributeNonPortableClusterSizeAllowed, 0);
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(2 * split * 64);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = smem;
    cudaLaunchAttribute at[1];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 2 * split;
    at[0].val.clusterDim.y = 1;
    at[0].val.clusterDim.z = 1;
    cfg.attrs = at;
    cfg.numAttrs = 1;
    int n = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&n, gemm_fp8_v4_kernel<BN>, &cfg);
    if (e != cudaSuccess || n <= 0) {
      cudaGetLastError();
      n = -1;
    }
    cache[split] = n;
  }
  return cache[split];
}

template <int BN>
static int launch(const CUtensorMap& tmR, const CUtensorMap& tmC, const CUtensorMap& tmO, Params& p, int tiles,
                  int split_req, cudaStream_t stream) {
  // split: as many K slices as keep every cluster co-resident in one wave (<= 4, <= k-blocks)
  int split = 0, smem = 0, stages = 0;
  for (int sp = MAX_SPLIT; sp >= 1; sp--) {
    if (split_req > 0 && sp != split_req) continue;
    if (sp > p.num_kb) continue;
    int st = 0;
    const int sm = smem_plan<BN>(sp, &st);
    if (sm < 0) continue;
    const int mc = max_clusters<BN>(sp, sm);
    if (mc <= 0) continue;
    if (split_req > 0 || tiles <= mc) { split = sp; smem = sm; stages = st; break; }
  }
  if (split == 0) {
    set_last_error("fp8_gemm_splitk: no split of %d weight tiles fits one wave of clusters", tiles);
    return B200_ERR_UNSUPPORTED;
  }
  p.split = split;
  p.stages = stages;
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(tiles * 2 * split);
  cfg.blockDim = dim3(NTHREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[2];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 2 * split;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = 1;
  at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[1].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = pdl_enabled() ? 2 : 1;
  p.prefetch_w = pdl_level() >= 2;
  B200_CUDA(cudaLaunchKernelEx(&cfg, gemm_fp8_v4_kernel<BN>, tmR, tmC, tmO, p));
  return B200_OK;
}

}  // namespace gemm8v4
}  // namespace b200

extern "C" int b200_fp8_gemm_splitk(void* out, const void* a, const void* b, const float* a_scale,
                                    const float* b_scale, const void* bias, int M, int N, int K, int64_t lda,
                                    int64_t ldb, int64_t ldo, int a_scale_per_row, int b_scale_per_col,
                                    int out_dtype, int split, b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::gemm8v4;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(M >= 0 && N > 0 && K > 0, B200_ERR_INVALID, "cutlass_scaled_mm: bad sizes (%d,%d,%d)", M, N, K);
  if (M == 0) return B200_OK;
  B200_CHECK(M <= 256, B200_ERR_UNSUPPORTED, "fp8_gemm_splitk: decode batches only (M <= 256), got %d", M);
  B200_CHECK(split >= 0 && split <= MAX_SPLIT, B200_ERR_INVALID, "fp8_gemm_splitk: split must be 0 (auto) .. %d", MAX_SPLIT);
  B200_CHECK(out && a && b && a_scale && b_scale, B200_ERR_INVALID, "cutlass_scaled_mm: null pointer");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "cutlass_scaled_mm: out dtype must be bf16 or f16");
  B200_CHECK(lda % 16 == 0 && ldb % 16 == 0 && K % 16 == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm: K and the row strides of a / b must be multiples of 16 (got K=%d lda=%lld ldb=%lld)",
             K, (long long)lda, (long long)ldb);
  B200_CHECK(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm: a / b must be 16-byte aligned");
  B200_CHECK(((uintptr_t)out & 15) == 0 && (ldo * 2) % 16 == 0, B200_ERR_UNSUPPORTED,
             "fp8_gemm_splitk: out must be 16-byte aligned with a 16-byte multiple row stride");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "cutlass_scaled_mm: no CUDA device (no CPU fallback)");
  Params p;
  memset(&p, 0, sizeof(p));
  p.bias = bias;
  p.K = K;
  p.out_bf16 = out_dtype == B200_BF16;
  p.num_kb = (K + BK - 1) / BK;
  p.R = N; p.C = M;
  p.r_scale = b_scale; p.r_scale_vec = b_scale_per_col;
  p.c_scale = a_scale; p.c_scale_vec = a_scale_per_row;
  const int bn = M <= 64 ? 64 : M <= 128 ? 128 : 256;
  CUtensorMap tmR, tmC, tmO;
  int rc = make_tmap_2d(&tmR, b, N, K, ldb, BM);
  if (rc) return rc;
  rc = make_tmap_2d(&tmC, a, M, K, lda, bn / 2);
  if (rc) return rc;
  {
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
    uint64_t strides[1] = {(uint64_t)ldo * 2};
    uint32_t box[2] = {(uint32_t)BM, 16u};
    rc = encode_tensor_map(&tmO, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, out, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE);
    if (rc) return rc;
  }
  const int tiles = (N + 2 * BM - 1) / (2 * BM);
  switch (bn) {
    case 64: return launch<64>(tmR, tmC, tmO, p, tiles, split, stream);
    case 128: return launch<128>(tmR, tmC, tmO, p, tiles, split, stream);
    default: return launch<256>(tmR, tmC, tmO, p, tiles, split, stream);
  }
}
