// b200ops -- fused-MoE grouped FP8 GEMM on tcgen05 / TMEM / TMA (sm_100a).
//
// Contract: the GEMM inside fused_experts (vllm/model_executor/layers/fused_moe/fused_moe.py:299-612,
// launcher :763-905) / cutlass_moe_mm (csrc/libtorch_stable/quantization/w8a8/cutlass/moe/
// grouped_mm_c3x_sm100.cu:16-115): for every block b of `block_size` rows of the expert-sorted
// (permuted) activation buffer,
//     out[b*bs + i, :] = a_scale * w_scale[e] * (A8[b*bs + i, :] . W8[e]^T) (* row_weight[b*bs + i]),
//     e = expert_ids[b];  blocks with e < 0 or b*bs >= *num_tokens_post_pad are skipped.
//
// B200 design: same warp-specialised tcgen05 pipeline as gemm_fp8.cu in swap-AB orientation:
// rows operand = 128 output features of expert e (3-D TMA map [K, N, E]), cols operand = the block's
// tokens (UMMA N = block_size), accumulator in TMEM, epilogue staged in smem and written by one TMA
// store. Routing tables are read on the device: no host sync, CUDA-graph safe.
#include "tcgen05.cuh"

#include <cstdlib>
#include <cstring>

namespace b200 {
namespace moegemm {

using namespace b200::tc;

constexpr int BM = 128, BK = 128, UMMA_K = 32, NTHREADS = 192, MAX_STAGES = 8;

struct Params {
  const int32_t* expert_ids;
  const int32_t* num_tokens_post_pad;
  const float* a_scale;     // [1], or [P] one per sorted row (a_scale_per_row: per-token activation quant)
  const float* w_scale;     // [E] or [E, N]
  const float* row_weight;  // [P] nullable (topk weight of each sorted row)
  int N, K, w_scale_per_channel, a_scale_per_row, out_bf16, stages;
};

template <int BN>
__global__ void __launch_bounds__(NTHREADS, 1)
    moe_gemm_fp8_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmA,
                        const __grid_constant__ CUtensorMap tmO, const Params p) {
  constexpr uint32_t R_BYTES = BM * BK, C_BYTES = BN * BK, STAGE_BYTES = R_BYTES + C_BYTES;
  constexpr uint32_t TMEM_COLS = BN <= 32 ? 32 : BN <= 64 ? 64 : BN <= 128 ? 128 : 256;
  constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

  // grid = (token blocks, weight slabs): the blocks of one expert that need the same 128-row weight slab are
  // co-scheduled, so the slab is fetched from HBM once and the other blocks hit L2 (with slabs on x an expert
  // with two blocks streamed its whole weight twice: 3.96 TB/s algorithmic on the Mixtral shape, r1i profile)
  const int blk = blockIdx.x;
  const int e = p.expert_ids[blk];
  if (e < 0 || blk * BN >= *p.num_tokens_post_pad) return;  // uniform for the CTA

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t base = (raw_base + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_base);
  const int stages = p.stages;
  const uint32_t bar_full = base + (uint32_t)stages * STAGE_BYTES;
  const uint32_t bar_empty = bar_full + 8u * MAX_STAGES;
  const uint32_t bar_tmem = bar_full + 16u * MAX_STAGES;
  volatile uint32_t* tmem_holder =
      reinterpret_cast<volatile uint32_t*>(smem + (size_t)stages * STAGE_BYTES + 16 * MAX_STAGES + 8);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = blockIdx.y * BM;  // output feature
  const int col0 = blk * BN;         // sorted row
  const int num_kb = (p.K + BK - 1) / BK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; i++) {
      mbar_init(bar_full + 8u * i, 1);
      mbar_init(bar_empty + 8u * i, 1);
    }
    mbar_init(bar_tmem, 1);
    mbar_fence_init();
  }
  if (warp == 1) tmem_alloc(smem_u32((const void*)tmem_holder), TMEM_COLS);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_holder;

  if (warp == 0) {
    // TMA producer: warp-uniform loop, one elected lane issues (common.cuh elect_one)
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmA);
    const uint64_t pol_w = l2_policy_evict_first(), pol_a = l2_policy_evict_last();
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; kb++) {
      mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
      if (elect_one()) {
        const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES;
        mbar_arrive_expect_tx(bar_full + 8u * stage, STAGE_BYTES);
        tma_load_3d(dst, &tmW, bar_full + 8u * stage, kb * BK, row0, e, pol_w);
        tma_load_2d(dst + R_BYTES, &tmA, bar_full + 8u * stage, kb * BK, col0, pol_a);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // MMA issuer: warp-uniform loop, one elected lane issues
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; kb++) {
      mbar_wait(bar_full + 8u * stage, phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sa = base + (uint32_t)stage * STAGE_BYTES;
        const uint64_t da = make_smem_desc(sa), db = make_smem_desc(sa + R_BYTES);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; k++)
          tc_mma_f8(tmem_d, da + (uint64_t)(k * UMMA_K >> 4), db + (uint64_t)(k * UMMA_K >> 4), IDESC,
                    (kb > 0 || k > 0) ? 1u : 0u);
        tc_commit(bar_empty + 8u * stage);
        if (kb == num_kb - 1) tc_commit(bar_tmem);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else {
    const int wq = warp & 3;
    mbar_wait(bar_tmem, 0);
    tc_fence_after();
    const int n = row0 + wq * 32 + lane;
    const bool n_ok = n < p.N;
    const float ws = n_ok ? (p.w_scale_per_channel ? p.w_scale[(int64_t)e * p.N + n] : p.w_scale[e]) : 0.f;
    const float as = p.a_scale_per_row ? 0.f : p.a_scale[0];
    const uint32_t tbase = tmem_d + ((uint32_t)(wq * 32) << 16);
    const uint32_t stile = base;
    uint32_t v0[16], v1[16];
    tc_ld16(tbase, v0);
    auto stage_chunk = [&](const uint32_t (&vv)[16], int ch) {
#pragma unroll
      for (int j = 0; j < 16; j += 2) {
        const int c = col0 + (ch << 4) + j;
        const float rw_a = p.row_weight ? p.row_weight[c] : 1.f;
        const float rw_b = p.row_weight ? p.row_weight[c + 1] : 1.f;
        const float as_a = p.a_scale_per_row ? p.a_scale[c] : as;
        const float as_b = p.a_scale_per_row ? p.a_scale[c + 1] : as;
        const float oa = rw_a * (as_a * (ws * __uint_as_float(vv[j])));
        const float ob = rw_b * (as_b * (ws * __uint_as_float(vv[j + 1])));
        const float send = (lane & 1) ? oa : ob;
        const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
        const float lo = (lane & 1) ? recv : oa;
        const float hi = (lane & 1) ? ob : recv;
        uint32_t packed;
        if (p.out_bf16) {
          __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
          packed = *reinterpret_cast<uint32_t*>(&t);
        } else {
          __half2 t = __floats2half2_rn(lo, hi);
          packed = *reinterpret_cast<uint32_t*>(&t);
        }
        const int m_local = (ch << 4) + j + (lane & 1);
        const uint32_t addr = stile + (uint32_t)m_local * (BM * 2) + (uint32_t)(wq * 32 + (lane & ~1)) * 2u;
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(packed) : "memory");
      }
    };
    constexpr int nchunks = BN / 16;
    for (int ch = 0; ch < nchunks; ch += 2) {
      tc_wait_ld();
      if (ch + 1 < nchunks) tc_ld16(tbase + (uint32_t)((ch + 1) << 4), v1);
      stage_chunk(v0, ch);
      if (ch + 1 < nchunks) {
        tc_wait_ld();
        if (ch + 2 < nchunks) tc_ld16(tbase + (uint32_t)((ch + 2) << 4), v0);
        stage_chunk(v1, ch + 1);
      }
    }
    fence_proxy_async();
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (warp == 2 && lane == 0) tma_store_2d(&tmO, stile, row0, col0);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_d, TMEM_COLS);
  }
}

template <int BN>
static int launch(const CUtensorMap& tmW, const CUtensorMap& tmA, const CUtensorMap& tmO, Params& p,
                  dim3 grid, cudaStream_t stream) {
  constexpr uint32_t STAGE_BYTES = BM * BK + BN * BK;
  const int smem_max = max_smem_optin();
  int stages = (smem_max - 1024 - 256 - 64) / (int)STAGE_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  // many short CTAs (one token block x one weight slab each): a pipeline that fills the SM alone leaves its
  // prologue and epilogue uncovered (ncu r1k: DRAM traffic == algorithmic bytes but 4.4 TB/s). Half the ring lets
  // two CTAs share an SM so one streams while the other drains. B200_MOE_STAGES overrides (tuning).
  const int two_cta = ((228 * 1024) / 2 - 1024 - (16 * MAX_STAGES + 64 + 1024)) / (int)STAGE_BYTES;
  if ((int64_t)grid.x * grid.y >= 2 * (int64_t)sm_count() && two_cta >= 3 && two_cta < stages) stages = two_cta;
  static const int forced = [] { const char* e = getenv("B200_MOE_STAGES"); return e ? atoi(e) : 0; }();
  if (forced >= 2 && forced <= MAX_STAGES && forced * (int)STAGE_BYTES + 2048 <= smem_max) stages = forced;
  p.stages = stages;
  const int smem = stages * (int)STAGE_BYTES + 16 * MAX_STAGES + 64 + 1024;
  static bool configured_dev[64] = {};   // the attribute is per device
  int cfg_dev_ = 0;
  cudaGetDevice(&cfg_dev_);
  bool& configured = configured_dev[cfg_dev_ & 63];
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(moe_gemm_fp8_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    configured = true;
  }
  moe_gemm_fp8_kernel<BN><<<grid, NTHREADS, smem, stream>>>(tmW, tmA, tmO, p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace moegemm
}  // namespace b200

extern "C" int b200_moe_grouped_gemm_fp8(void* out, const void* a_sorted, const void* w,
                                         const int32_t* expert_ids,
                                         const int32_t* num_tokens_post_pad, const float* a_scale,
                                         const float* w_scale, const float* row_weight,
                                         int64_t max_rows, int num_experts, int N, int K,
                                         int block_size, int w_scale_per_channel, int a_scale_per_row,
                                         int out_dtype, b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::moegemm;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(max_rows >= 0 && num_experts > 0 && N > 0 && K > 0, B200_ERR_INVALID,
             "moe_grouped_gemm_fp8: bad sizes");
  if (max_rows == 0) return B200_OK;
  B200_CHECK(out && a_sorted && w && expert_ids && num_tokens_post_pad && a_scale && w_scale,
             B200_ERR_INVALID, "moe_grouped_gemm_fp8: null pointer");
  B200_CHECK(block_size == 16 || block_size == 32 || block_size == 64 || block_size == 128,
             B200_ERR_UNSUPPORTED, "moe_grouped_gemm_fp8: block_size must be 16, 32, 64 or 128 (got %d)",
             block_size);
  B200_CHECK(max_rows % block_size == 0, B200_ERR_INVALID,
             "moe_grouped_gemm_fp8: the sorted buffer must hold a whole number of blocks");
  B200_CHECK(K % 16 == 0 && N % 8 == 0, B200_ERR_INVALID,
             "moe_grouped_gemm_fp8: K must be a multiple of 16 and N of 8");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "moe_grouped_gemm_fp8: out dtype must be bf16 or f16");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "moe_grouped_gemm_fp8: no CUDA device (no CPU fallback)");
  CUtensorMap tmW, tmA, tmO;
  {
    uint64_t dims[3] = {(uint64_t)K, (uint64_t)N, (uint64_t)num_experts};
    uint64_t strides[2] = {(uint64_t)K, (uint64_t)K * (uint64_t)N};
    uint32_t box[3] = {(uint32_t)BK, (uint32_t)BM, 1u};
    int rc = encode_tensor_map(&tmW, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, w, dims, strides, box,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)K, (uint64_t)max_rows};
    uint64_t strides[1] = {(uint64_t)K};
    uint32_t box[2] = {(uint32_t)BK, (uint32_t)block_size};
    int rc = encode_tensor_map(&tmA, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, a_sorted, dims, strides, box,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)max_rows};
    uint64_t strides[1] = {(uint64_t)N * 2};
    uint32_t box[2] = {(uint32_t)BM, (uint32_t)block_size};
    int rc = encode_tensor_map(&tmO, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, out, dims, strides, box,
                               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE);
    if (rc) return rc;
  }
  Params p;
  p.expert_ids = expert_ids;
  p.num_tokens_post_pad = num_tokens_post_pad;
  p.a_scale = a_scale;
  p.w_scale = w_scale;
  p.row_weight = row_weight;
  p.N = N;
  p.K = K;
  p.w_scale_per_channel = w_scale_per_channel;
  p.a_scale_per_row = a_scale_per_row;
  p.out_bf16 = out_dtype == B200_BF16;
  p.stages = 0;
  dim3 grid((unsigned)(max_rows / block_size), (N + BM - 1) / BM);
  switch (block_size) {
    case 16: return launch<16>(tmW, tmA, tmO, p, grid, stream);
    case 32: return launch<32>(tmW, tmA, tmO, p, grid, stream);
    case 64: return launch<64>(tmW, tmA, tmO, p, grid, stream);
    default: return launch<128>(tmW, tmA, tmO, p, grid, stream);
  }
}
