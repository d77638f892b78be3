// b200ops -- block-scaled FP8 (E4M3) GEMM, DeepSeek-style 1x128 activation / 128x128 weight scales, on
// tcgen05 / TMEM / TMA (sm_100a).
//
// Contract: the blockwise branch of torch.ops._C.cutlass_scaled_mm
// (csrc/libtorch_stable/quantization/w8a8/cutlass/c3x/scaled_mm_helper.hpp:40-60 -> blockwise_func,
//  c3x/scaled_mm_blockwise_sm100_fp8_dispatch.cuh; semantic oracle tests/kernels/quant_utils.py:91-155):
//   out[m, n] = sum_kb a_s[m, kb] * b_s[kb, n / 128] * (A[m, kb-th 128 columns] . B[same, n]),
//   a_s float [M, ceil(K/128)] (any strides: the reference passes it column-major), b_s float
//   [ceil(K/128), ceil(N/128)] (any strides), no bias, fp32 accumulate, out bf16 / f16.
//
// B200 design. The tensor core cannot apply a scale per k-block, so each 128-wide k-block is contracted
// into a PARTIAL accumulator in TMEM (4 x tcgen05.mma kind::f8f6f4 128 x BN x 32, accumulate reset per
// k-block) and promoted into fp32 registers by the epilogue warps: acc += partial * (a_s * b_s).
// Two TMEM slots alternate so the MMAs of k-block kb+1 overlap the promotion of kb.
// Orientation is always swap-AB: the 128 TMEM lanes are 128 weight rows (exactly one 128 x 128 weight
// scale block per k-block -> one scalar), the BN TMEM columns are tokens (each with its own a_s[m, kb],
// staged as a_s * b_s in shared memory once per k-block and read back as broadcast float4).
// Warp roles as csrc/gemm_fp8.cu: warp 0 TMA producer (SWIZZLE_128B boxes, mbarrier ring), warp 1 the
// single-thread MMA issuer, warps 2-5 promotion + epilogue (tile transposed through the idle pipeline
// smem and written by one TMA store).
// Bound: weight stream N*K bytes from HBM for decode batches (M <= 128: every weight byte is read once);
// for large M the promotion (BN FFMA + BN/4 LDS per thread per k-block) runs beside the MMAs.
#include "tcgen05.cuh"

#include <cstring>

namespace b200 {
namespace gemm8b {

constexpr int BM = 128;     // weight rows per tile = one scale block = UMMA M
constexpr int BK = 128;     // k-block = scale block = one 128-byte swizzle row
constexpr int UMMA_K = 32;
constexpr int NTHREADS = 192;
constexpr int MAX_STAGES = 8;

struct Params {
  void* out;
  const float* a_s;   // [M, KB] strides (as_m, as_k)
  const float* b_s;   // [KB, NB] strides (bs_k, bs_n)
  int64_t as_m, as_k, bs_k, bs_n;
  int M, N, K;
  int64_t ldo;
  int out_bf16;
  int stages;
  int tma_store;
  int split;   // K slices = CTAs per thread-block cluster (1, 1, split); 1 = no cluster
  // grouped (fused-MoE) mode: grid = (token blocks of BN sorted rows, weight slabs); block b uses expert
  // expert_ids[b] (weight slice e of the 3-D map, scale matrix b_s + e * bs_e) and is skipped when the id is
  // negative or the block starts at / after *num_tokens_post_pad. Null expert_ids = plain GEMM.
  const int32_t* expert_ids;
  const int32_t* num_tokens_post_pad;
  int64_t bs_e;
};

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t mapa(uint32_t addr, uint32_t rank) {
  uint32_t r;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(r) : "r"(addr), "r"(rank));
  return r;
}

using namespace b200::tc;

template <int BN>
__global__ void __launch_bounds__(NTHREADS, 1)
    gemm_fp8_block_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmA,
                          const __grid_constant__ CUtensorMap tmO, const Params p) {
  constexpr uint32_t W_BYTES = BM * BK;
  constexpr uint32_t A_BYTES = BN * BK;
  constexpr uint32_t STAGE_BYTES = W_BYTES + A_BYTES;
  constexpr uint32_t TMEM_COLS = 2 * BN <= 32 ? 32 : 2 * BN <= 64 ? 64 : 2 * BN <= 128 ? 128 : 256;
  constexpr uint32_t IDESC = (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(BN >> 3) << 17) |
                             ((uint32_t)(BM >> 4) << 24);
  constexpr int NCH = BN / 16;

  const bool grouped = p.expert_ids != nullptr;
  const int wt = grouped ? blockIdx.y : blockIdx.x;   // weight slab
  const int tt = grouped ? blockIdx.x : blockIdx.y;   // token tile
  int expert = 0;
  if (grouped) {
    expert = p.expert_ids[tt];
    if (expert < 0 || tt * BN >= *p.num_tokens_post_pad) return;   // uniform for the CTA
  }

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t base = (raw_base + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_base);
  const int stages = p.stages;
  // [stages x (W | A)] [bar_full x8] [bar_empty x8] [acc_full x2] [acc_empty x2] [tmem holder] [pad] [sc 2 x BN f32]
  const uint32_t bar_base = base + (uint32_t)stages * STAGE_BYTES;
  const uint32_t bar_full = bar_base;
  const uint32_t bar_empty = bar_base + 8u * MAX_STAGES;
  const uint32_t bar_acc_full = bar_base + 16u * MAX_STAGES;
  const uint32_t bar_acc_empty = bar_acc_full + 16u;
  uint8_t* tail = smem + (size_t)stages * STAGE_BYTES + 16 * MAX_STAGES + 32;
  volatile uint32_t* tmem_holder = reinterpret_cast<volatile uint32_t*>(tail);
  float* sc = reinterpret_cast<float*>(tail + 16);   // 16-byte aligned: base is 1024-aligned, offsets are x16

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = wt * BM;   // weight rows (output columns)
  const int col0 = tt * BN;   // tokens (output rows)
  // split-K: CTA `ks` of the cluster contracts k-blocks [kb0, kb0 + num_kb) of the tile
  const int total_kb = (p.K + BK - 1) / BK;
  const int ks = p.split > 1 ? (int)cluster_ctarank() : 0;
  const int kb0 = (int)((int64_t)ks * total_kb / p.split);
  const int num_kb = (int)((int64_t)(ks + 1) * total_kb / p.split) - kb0;

  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; i++) {
      mbar_init(bar_full + 8u * i, 1);
      mbar_init(bar_empty + 8u * i, 1);
    }
    for (int i = 0; i < 2; i++) {
      mbar_init(bar_acc_full + 8u * i, 1);
      mbar_init(bar_acc_empty + 8u * i, 128);
    }
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(smem_u32((const void*)tmem_holder), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_holder;

  if (warp == 0) {
    // ============================== TMA producer (warp-uniform loop, one elected lane issues: common.cuh elect_one)
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmA);
    const uint64_t pol_w = (grouped || gridDim.y > 1) ? l2_policy_evict_last() : l2_policy_evict_first();
    const uint64_t pol_a = l2_policy_evict_last();
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; kb++) {
      mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
      if (elect_one()) {
        const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES;
        mbar_arrive_expect_tx(bar_full + 8u * stage, STAGE_BYTES);
        tma_load_3d(dst, &tmW, bar_full + 8u * stage, (kb0 + kb) * BK, row0, expert, pol_w);
        tma_load_2d(dst + W_BYTES, &tmA, bar_full + 8u * stage, (kb0 + kb) * BK, col0, pol_a);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer (one elected lane): one partial accumulator per k-block
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; kb++) {
      const int slot = kb & 1;
      mbar_wait(bar_acc_empty + 8u * slot, (((uint32_t)kb >> 1) & 1u) ^ 1u);   // promotion of kb-2 done
      mbar_wait(bar_full + 8u * stage, phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sa = base + (uint32_t)stage * STAGE_BYTES;
        const uint64_t dw = make_smem_desc(sa);
        const uint64_t da = make_smem_desc(sa + W_BYTES);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; k++)
          tc_mma_f8(tmem_d + (uint32_t)(slot * BN), dw + (uint64_t)(k * UMMA_K >> 4),
                    da + (uint64_t)(k * UMMA_K >> 4), IDESC, k > 0 ? 1u : 0u);
        tc_commit(bar_empty + 8u * stage);        // smem slot free when the MMAs have read it
        tc_commit(bar_acc_full + 8u * slot);      // partial accumulator complete
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else {
    // ============================== promotion + epilogue (4 warps, one TMEM lane quarter each)
    const int wq = warp & 3;
    const int et = wq * 32 + lane;        // 0..127: TMEM lane = weight row of the tile
    const int r = row0 + et;              // output column n
    const bool r_ok = r < p.N;
    const float* bs_col = p.b_s + (int64_t)expert * p.bs_e + (int64_t)wt * p.bs_n;
    // this thread stages the combined scale of token (col0 + et) when et < BN
    const int my_m = col0 + et;
    const bool stage_it = et < BN;
    const float* as_row = p.a_s + (int64_t)(my_m < p.M ? my_m : 0) * p.as_m;
    float acc[BN];
#pragma unroll
    for (int i = 0; i < BN; i++) acc[i] = 0.f;
    const uint32_t tlane = tmem_d + ((uint32_t)(wq * 32) << 16);
    float s_next = 0.f;
    if (stage_it && num_kb > 0)
      s_next = (my_m < p.M ? as_row[(int64_t)kb0 * p.as_k] : 0.f) * bs_col[(int64_t)kb0 * p.bs_k];
    for (int kb = 0; kb < num_kb; kb++) {
      const int slot = kb & 1;
      float* scs = sc + slot * BN;
      if (stage_it) scs[et] = s_next;
      if (stage_it && kb + 1 < num_kb)
        s_next = (my_m < p.M ? as_row[(int64_t)(kb0 + kb + 1) * p.as_k] : 0.f) *
                 bs_col[(int64_t)(kb0 + kb + 1) * p.bs_k];
      asm volatile("bar.sync 1, 128;" ::: "memory");
      mbar_wait(bar_acc_full + 8u * slot, ((uint32_t)kb >> 1) & 1u);
      tc_fence_after();
      const uint32_t tb = tlane + (uint32_t)(slot * BN);
      uint32_t v0[16], v1[16];
      tc_ld16(tb, v0);
#pragma unroll
      for (int ch = 0; ch < NCH; ch += 2) {
        tc_wait_ld();
        if (ch + 1 < NCH) tc_ld16(tb + (uint32_t)((ch + 1) << 4), v1);
#pragma unroll
        for (int j = 0; j < 16; j += 4) {
          const float4 s4 = *reinterpret_cast<const float4*>(scs + (ch << 4) + j);
          acc[(ch << 4) + j + 0] = fmaf(__uint_as_float(v0[j + 0]), s4.x, acc[(ch << 4) + j + 0]);
          acc[(ch << 4) + j + 1] = fmaf(__uint_as_float(v0[j + 1]), s4.y, acc[(ch << 4) + j + 1]);
          acc[(ch << 4) + j + 2] = fmaf(__uint_as_float(v0[j + 2]), s4.z, acc[(ch << 4) + j + 2]);
          acc[(ch << 4) + j + 3] = fmaf(__uint_as_float(v0[j + 3]), s4.w, acc[(ch << 4) + j + 3]);
        }
        if (ch + 1 < NCH) {
          tc_wait_ld();
          if (ch + 2 < NCH) tc_ld16(tb + (uint32_t)((ch + 2) << 4), v0);
#pragma unroll
          for (int j = 0; j < 16; j += 4) {
            const float4 s4 = *reinterpret_cast<const float4*>(scs + ((ch + 1) << 4) + j);
            acc[((ch + 1) << 4) + j + 0] = fmaf(__uint_as_float(v1[j + 0]), s4.x, acc[((ch + 1) << 4) + j + 0]);
            acc[((ch + 1) << 4) + j + 1] = fmaf(__uint_as_float(v1[j + 1]), s4.y, acc[((ch + 1) << 4) + j + 1]);
            acc[((ch + 1) << 4) + j + 2] = fmaf(__uint_as_float(v1[j + 2]), s4.z, acc[((ch + 1) << 4) + j + 2]);
            acc[((ch + 1) << 4) + j + 3] = fmaf(__uint_as_float(v1[j + 3]), s4.w, acc[((ch + 1) << 4) + j + 3]);
          }
        }
      }
      tc_fence_before();
      mbar_arrive(bar_acc_empty + 8u * slot);   // TMEM slot may be overwritten
    }
    if (p.split > 1) {
      // ---- split-K reduction over distributed shared memory: slices 1.. push their fp32 tiles into slice 0's
      // idle pipeline smem ([slice-1][m][128 n] floats, lanes -> consecutive words), slice 0 adds them in slice
      // order (deterministic) and runs the epilogue alone.
      cluster_sync();   // A: every CTA of the cluster is past its main loop (its MMAs have all completed)
      if (ks > 0) {
        const uint32_t remote = mapa(base + (uint32_t)((ks - 1) * BN) * (BM * 4) + (uint32_t)et * 4u, 0);
#pragma unroll
        for (int j = 0; j < BN; j++)
          asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote + (uint32_t)j * (BM * 4)), "f"(acc[j])
                       : "memory");
      }
      cluster_sync();   // B: partial tiles have landed (release / acquire at cluster scope)
      for (int z = 1; ks == 0 && z < p.split; z++) {
        const uint32_t src = base + (uint32_t)((z - 1) * BN) * (BM * 4) + (uint32_t)et * 4u;
#pragma unroll
        for (int j = 0; j < BN; j++) {
          float v;
          asm volatile("ld.shared.f32 %0, [%1];" : "=f"(v) : "r"(src + (uint32_t)j * (BM * 4)) : "memory");
          acc[j] += v;
        }
      }
      asm volatile("bar.sync 1, 128;" ::: "memory");   // the staging below overwrites the partial tiles
    }
    if (ks == 0) {
    // ---- epilogue: out[m][n] = T(acc); lane = n, register index = m
    auto pack2 = [&](float lo, float hi) -> uint32_t {
      if (p.out_bf16) {
        __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&t);
      }
      __half2 t = __floats2half2_rn(lo, hi);
      return *reinterpret_cast<uint32_t*>(&t);
    };
    if (p.tma_store) {
      // stage the [BN m][128 n] tile in the idle pipeline smem (every MMA has completed: the last
      // acc_full commit covers all earlier tcgen05 work), one TMA store writes rows of 256 bytes
      const uint32_t stile = base;
#pragma unroll
      for (int j = 0; j < BN; j += 2) {
        // 2x2 transpose between lane pairs: even lane keeps token j of rows (n, n+1), odd lane token j+1
        const float oa = acc[j], ob = acc[j + 1];
        const float send = (lane & 1) ? oa : ob;
        const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
        const float lo = (lane & 1) ? recv : oa;
        const float hi = (lane & 1) ? ob : recv;
        const int m_local = j + (lane & 1);
        const uint32_t addr = stile + (uint32_t)m_local * (BM * 2) + (uint32_t)(wq * 32 + (lane & ~1)) * 2u;
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(pack2(lo, hi)) : "memory");
      }
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) tma_store_2d(&tmO, stile, row0, col0);
    } else if (r_ok) {
      uint16_t* out16 = reinterpret_cast<uint16_t*>(p.out);
#pragma unroll
      for (int j = 0; j < BN; j++) {
        const int m = col0 + j;
        if (m < p.M) out16[(int64_t)m * p.ldo + r] = (uint16_t)(pack2(acc[j], 0.f) & 0xffffu);
      }
    }
    }  // ks == 0
  }
  if (warp < 2 && p.split > 1) {   // producer / MMA warps take part in the two cluster barriers
    cluster_sync();
    cluster_sync();
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_d, TMEM_COLS);
  }
}

template <int BN>
static int launch(const CUtensorMap& tmW, const CUtensorMap& tmA, const CUtensorMap& tmO, Params& p, dim3 grid,
                  cudaStream_t stream) {
  constexpr uint32_t STAGE_BYTES = BM * BK + BN * BK;
  constexpr int TAIL = 16 * MAX_STAGES + 32 + 16 + 2 * BN * 4;
  const int smem_max = max_smem_optin();
  int stages = (smem_max - 1024 - TAIL - 64) / (int)STAGE_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) {
    set_last_error("cutlass_scaled_mm (blockwise): not enough shared memory");
    return B200_ERR_UNSUPPORTED;
  }
  p.stages = stages;
  const int smem = stages * (int)STAGE_BYTES + TAIL + 64 + 1024;
  static bool configured_dev[64] = {};   // the attribute is per device
  int cfg_dev_ = 0;
  cudaGetDevice(&cfg_dev_);
  bool& configured = configured_dev[cfg_dev_ & 63];
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(gemm_fp8_block_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   smem_max));
    configured = true;
  }
  // split-K (decode batches: one token tile, few weight tiles): K slices over the CTAs of a (1, 1, split)
  // cluster so that tiles x split approaches the SM count; every slice keeps >= 4 k-blocks and the
  // (split - 1) fp32 partial tiles must fit slice 0's pipeline smem.
  const int total_kb = (p.K + BK - 1) / BK;
  int split = 1;
  if (grid.y == 1 && !p.expert_ids) {
    const int want = p.split;   // 0 = auto
    for (int sp = 8; sp >= 2; sp >>= 1) {
      const bool fits = (sp - 1) * BN * BM * 4 <= stages * (int)STAGE_BYTES && sp <= total_kb;
      if (!fits) continue;
      if (want > 0 ? sp <= want : ((int)grid.x * sp <= sm_count() && total_kb / sp >= 4)) { split = sp; break; }
    }
  }
  p.split = split;
  if (split == 1) {
    gemm_fp8_block_kernel<BN><<<grid, NTHREADS, smem, stream>>>(tmW, tmA, tmO, p);
    B200_LAUNCH_CHECK();
    return B200_OK;
  }
  cudaLaunchConfig_t cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.gridDim = dim3(grid.x, 1, split);
  cfg.blockDim = dim3(NTHREADS);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeClusterDimension;
  at[0].val.clusterDim.x = 1;
  at[0].val.clusterDim.y = 1;
  at[0].val.clusterDim.z = split;
  cfg.attrs = at;
  cfg.numAttrs = 1;
  B200_CUDA(cudaLaunchKernelEx(&cfg, gemm_fp8_block_kernel<BN>, tmW, tmA, tmO, p));
  return B200_OK;
}

// weights: [K, N, E] (E = 1 for the plain GEMM)
static int make_tmap_w(CUtensorMap* tm, const void* base, int rows, int K, int64_t ld_bytes, int experts) {
  uint64_t dims[3] = {(uint64_t)K, (uint64_t)rows, (uint64_t)experts};
  uint64_t strides[2] = {(uint64_t)ld_bytes, (uint64_t)ld_bytes * (uint64_t)rows};
  uint32_t box[3] = {(uint32_t)BK, (uint32_t)BM, 1u};
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, base, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

static int make_tmap_2d(CUtensorMap* tm, const void* base, int rows, int K, int64_t ld_bytes, int box_rows) {
  uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld_bytes};
  uint32_t box[2] = {(uint32_t)BK, (uint32_t)box_rows};
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

}  // namespace gemm8b
}  // namespace b200

extern "C" int b200_fp8_gemm_blockscaled(void* out, const void* a, const void* b, const float* a_scales,
                                         const float* b_scales, int M, int N, int K, int64_t lda,
                                         int64_t ldb, int64_t ldo, int64_t as_stride_m, int64_t as_stride_k,
                                         int64_t bs_stride_k, int64_t bs_stride_n, int out_dtype, int split,
                                         b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::gemm8b;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(M >= 0 && N > 0 && K > 0, B200_ERR_INVALID, "cutlass_scaled_mm (blockwise): bad sizes (%d,%d,%d)",
             M, N, K);
  if (M == 0) return B200_OK;
  B200_CHECK(out && a && b && a_scales && b_scales, B200_ERR_INVALID,
             "cutlass_scaled_mm (blockwise): null pointer");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "cutlass_scaled_mm (blockwise): out dtype must be bf16 or f16");
  B200_CHECK(lda % 16 == 0 && ldb % 16 == 0 && K % 16 == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm (blockwise): K and the row strides of a / b must be multiples of 16 "
             "(got K=%d lda=%lld ldb=%lld)", K, (long long)lda, (long long)ldb);
  B200_CHECK(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm (blockwise): a / b must be 16-byte aligned");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "cutlass_scaled_mm (blockwise): no CUDA device (no CPU fallback)");
  Params p;
  memset(&p, 0, sizeof(p));
  p.out = out;
  p.a_s = a_scales; p.b_s = b_scales;
  p.as_m = as_stride_m; p.as_k = as_stride_k; p.bs_k = bs_stride_k; p.bs_n = bs_stride_n;
  p.M = M; p.N = N; p.K = K;
  p.ldo = ldo;
  p.out_bf16 = out_dtype == B200_BF16;
  B200_CHECK(split >= 0 && split <= 8, B200_ERR_INVALID, "cutlass_scaled_mm (blockwise): split must be 0 (auto) .. 8");
  p.split = split;
  const int bn = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128;
  CUtensorMap tmW, tmA, tmO;
  memset(&tmO, 0, sizeof(tmO));
  int rc = make_tmap_w(&tmW, b, N, K, ldb, 1);
  if (rc) return rc;
  rc = make_tmap_2d(&tmA, a, M, K, lda, bn);
  if (rc) return rc;
  if (((uintptr_t)out & 15) == 0 && (ldo * 2) % 16 == 0) {
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
    uint64_t strides[1] = {(uint64_t)ldo * 2};
    uint32_t box[2] = {(uint32_t)BM, (uint32_t)bn};
    rc = encode_tensor_map(&tmO, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, out, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE);
    if (rc) return rc;
    p.tma_store = 1;
  }
  dim3 grid((N + BM - 1) / BM, (M + bn - 1) / bn);
  B200_CHECK(grid.y <= 65535, B200_ERR_UNSUPPORTED, "cutlass_scaled_mm (blockwise): M too large (%d)", M);
  switch (bn) {
    case 16: return launch<16>(tmW, tmA, tmO, p, grid, stream);
    case 32: return launch<32>(tmW, tmA, tmO, p, grid, stream);
    case 64: return launch<64>(tmW, tmA, tmO, p, grid, stream);
    default: return launch<128>(tmW, tmA, tmO, p, grid, stream);
  }
}

// Grouped (fused-MoE) form: fused_experts with block_shape = [128, 128]
// (vllm/model_executor/layers/fused_moe/fused_moe.py:299-612 use_fp8_w8a8 + block_shape branch :478-560, w8a8 block
// scales a_scale [rows, K/128], b_scale [E, N/128, K/128]).
extern "C" int b200_moe_grouped_gemm_fp8_blockscaled(void* out, const void* a_sorted, const void* w,
                                                     const int32_t* expert_ids,
                                                     const int32_t* num_tokens_post_pad, const float* a_scales,
                                                     const float* w_scales, int64_t max_rows, int num_experts,
                                                     int N, int K, int block_size, int out_dtype,
                                                     b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::gemm8b;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(max_rows >= 0 && num_experts > 0 && N > 0 && K > 0, B200_ERR_INVALID,
             "moe_grouped_gemm_fp8_blockscaled: bad sizes");
  if (max_rows == 0) return B200_OK;
  B200_CHECK(out && a_sorted && w && expert_ids && num_tokens_post_pad && a_scales && w_scales, B200_ERR_INVALID,
             "moe_grouped_gemm_fp8_blockscaled: null pointer");
  B200_CHECK(block_size == 16 || block_size == 32 || block_size == 64 || block_size == 128, B200_ERR_UNSUPPORTED,
             "moe_grouped_gemm_fp8_blockscaled: block_size must be 16, 32, 64 or 128 (got %d)", block_size);
  B200_CHECK(max_rows % block_size == 0 && max_rows < (1LL << 31), B200_ERR_INVALID,
             "moe_grouped_gemm_fp8_blockscaled: the sorted buffer must hold a whole number of blocks");
  B200_CHECK(K % 16 == 0 && N % 8 == 0, B200_ERR_INVALID,
             "moe_grouped_gemm_fp8_blockscaled: K must be a multiple of 16 and N of 8");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "moe_grouped_gemm_fp8_blockscaled: out dtype must be bf16 or f16");
  B200_CHECK(((uintptr_t)a_sorted & 15) == 0 && ((uintptr_t)w & 15) == 0 && ((uintptr_t)out & 15) == 0,
             B200_ERR_INVALID, "moe_grouped_gemm_fp8_blockscaled: tensors must be 16-byte aligned");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "moe_grouped_gemm_fp8_blockscaled: no CUDA device (no CPU fallback)");
  const int KB = (K + BK - 1) / BK, NB = (N + BM - 1) / BM;
  Params p;
  memset(&p, 0, sizeof(p));
  p.out = out;
  p.a_s = a_scales; p.b_s = w_scales;
  p.as_m = KB; p.as_k = 1;          // a_scales [max_rows, KB] row-major
  p.bs_k = 1; p.bs_n = KB;          // w_scales [E, NB, KB]
  p.bs_e = (int64_t)NB * KB;
  p.M = (int)max_rows; p.N = N; p.K = K;
  p.ldo = N;
  p.out_bf16 = out_dtype == B200_BF16;
  p.split = 1;
  p.tma_store = 1;
  p.expert_ids = expert_ids;
  p.num_tokens_post_pad = num_tokens_post_pad;
  CUtensorMap tmW, tmA, tmO;
  int rc = make_tmap_w(&tmW, w, N, K, K, num_experts);
  if (rc) return rc;
  rc = make_tmap_2d(&tmA, a_sorted, (int)max_rows, K, K, block_size);
  if (rc) return rc;
  {
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)max_rows};
    uint64_t strides[1] = {(uint64_t)N * 2};
    uint32_t box[2] = {(uint32_t)BM, (uint32_t)block_size};
    rc = encode_tensor_map(&tmO, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, out, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE);
    if (rc) return rc;
  }
  dim3 grid((unsigned)(max_rows / block_size), NB);
  switch (block_size) {
    case 16: return launch<16>(tmW, tmA, tmO, p, grid, stream);
    case 32: return launch<32>(tmW, tmA, tmO, p, grid, stream);
    case 64: return launch<64>(tmW, tmA, tmO, p, grid, stream);
    default: return launch<128>(tmW, tmA, tmO, p, grid, stream);
  }
}
