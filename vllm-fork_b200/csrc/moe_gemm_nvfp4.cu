// b200ops -- NVFP4 (W4A4) grouped GEMM for packed MoE inputs on the block-scaled tcgen05 MMA (sm_100a).
//
// Contract: torch.ops._C.cutlass_fp4_group_mm (csrc/libtorch_stable/quantization/fp4/
// nvfp4_blockwise_moe_kernel.cu:676-746, schema csrc/libtorch_stable/torch_bindings.cpp:226-232,
// vllm/_custom_ops.py:1010-1046; caller run_cutlass_moe_fp4, fused_moe/experts/cutlass_moe.py:498-688):
// rows of `a` are sorted by expert (get_cutlass_moe_mm_data), expert e owns the m_e = problem_sizes[e][0]
// rows starting at expert_offsets[e]; for those rows
//     out[r, :] = alphas[e] * sum_k (a[r,k] * sfa_e[r - off_e, k/16]) * (w[e,n,k] * sfw[e,n,k/16]),
// a [m_topk, K/2] u8 and w [E, N, K/2] u8 hold two e2m1 codes per byte; a_blockscale is one swizzled
// 128x4-tile scale matrix per expert starting sf_offsets[e] (a multiple of 128) rows into the buffer
// (written by scaled_fp4_experts_quant); w_blockscale [E, round_up(N,128), round_up(K/16,4)] is one swizzled
// matrix per expert; alphas f32 [E]; out bf16 / f16 [m_topk, N]. Rows outside every expert are not written.
//
// B200 design: gemm_nvfp4.cu's warp-specialised pipeline in swap-AB orientation -- rows operand = 128 output
// features of expert e (3-D TMA map [K/2, N, E]), cols operand = BN packed tokens of that expert (UMMA N = BN,
// BN = 64 or 128 chosen on the host from the average expert load) -- driven by a device-side tile table: CTA
// (x, y) finds (expert, token tile) number x by a warp scan over problem_sizes (no host sync, CUDA-graph safe)
// and handles weight slab y, so the token tiles that share a weight slab are co-scheduled and the slab is
// streamed from HBM once. A 64-token tile at offset 64 of a 128-row scale tile reads its scales two TMEM columns
// in. The ragged last tile of an expert computes on whatever rows follow in the packed buffer (token columns
// are independent in the MMA) and the epilogue simply does not write them: the transposed tile is staged in
// the idle pipeline smem and written with row-guarded 16-byte stores (256 contiguous bytes per token row).
#include "tcgen05.cuh"

#include <cstdlib>
#include <cstring>

namespace b200 {
namespace moegemm4 {

using namespace b200::tc;

constexpr int BM = 128, BKB = 128, BK = 256, SF_TILE = 512, SF_KB = 4 * SF_TILE, NTHREADS = 192, MAX_STAGES = 8;

struct Params {
  void* out;
  const uint8_t* a_sf;
  const uint8_t* w_sf;
  const float* alphas;
  const int32_t* problem_sizes;   // [E, 3], [e][0] = rows of expert e
  const int32_t* expert_offsets;  // [E]
  const int32_t* sf_offsets;      // [E]
  int E, N, K, out_bf16, stages;
  int64_t max_rows;
};

// (expert, token tile) number `blk` in expert order; every lane returns the same answer
template <int BN>
__device__ __forceinline__ bool find_tile(const int32_t* __restrict__ ps, int E, int blk, int lane, int* e_out,
                                          int* t_out, int* m_out) {
  int base = 0;
  for (int c = 0; c < E; c += 32) {
    const int e = c + lane;
    const int m = e < E ? ps[e * 3] : 0;
    const int tiles = (m + BN - 1) / BN;
    int inc = tiles;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int n = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += n;
    }
    const int total = __shfl_sync(0xffffffffu, inc, 31);
    if (blk < base + total) {
      const unsigned hit = __ballot_sync(0xffffffffu, base + inc > blk);
      const int l = __ffs(hit) - 1;
      const int excl = __shfl_sync(0xffffffffu, inc - tiles, l);
      *e_out = c + l;
      *t_out = blk - base - excl;
      *m_out = __shfl_sync(0xffffffffu, m, l);
      return true;
    }
    base += total;
  }
  return false;
}

template <int BN>
__global__ void __launch_bounds__(NTHREADS, 1)
    moe_gemm_nvfp4_kernel(const __grid_constant__ CUtensorMap tmW, const __grid_constant__ CUtensorMap tmA,
                          const Params p) {
  constexpr uint32_t R_BYTES = BM * BKB, C_BYTES = BN * BKB;
  constexpr uint32_t STAGE_BYTES = R_BYTES + C_BYTES + 2 * SF_KB;
  // TMEM: [0, BN) accumulator, 16 columns of weight scales (4 per MMA), 16 columns of token scales
  constexpr uint32_t SFA_COL = BN, SFB_COL = BN + 16;
  constexpr uint32_t TMEM_COLS = BN + 32 <= 128 ? 128 : 256;
  constexpr uint32_t IDESC = (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);
  static_assert(BN == 64 || BN == 128, "token tile");

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  int e, t, m_e;
  if (!find_tile<BN>(p.problem_sizes, p.E, (int)blockIdx.x, lane, &e, &t, &m_e)) return;   // uniform for the CTA
  const int tok0 = t * BN;                               // first token of this tile inside the expert
  const int64_t row_base = (int64_t)p.expert_offsets[e] + tok0;   // ... inside the packed buffer
  const int ntok = min(BN, m_e - tok0);
  const int row0 = blockIdx.y * BM;                      // output feature

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
  const int num_kb = (p.K + BK - 1) / BK;
  const int k_tiles = (p.K + 63) / 64;

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
    // ============================== producer: codes by TMA, scale tiles by 1-D bulk copies
    prefetch_tmap(&tmW);
    prefetch_tmap(&tmA);
    const uint64_t pol_w = l2_policy_evict_first(), pol_a = l2_policy_evict_last();
    const int n_tiles = (p.N + 127) >> 7;
    const uint8_t* wsf = p.w_sf + ((int64_t)e * n_tiles + (row0 >> 7)) * k_tiles * SF_TILE;
    const uint8_t* asf = p.a_sf + ((int64_t)(p.sf_offsets[e] >> 7) + (tok0 >> 7)) * k_tiles * SF_TILE;
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; kb++) {
      mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
      if (elect_one()) {
        const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES;
        const int kt = min(4, k_tiles - kb * 4);
        const uint32_t sf_bytes = (uint32_t)kt * SF_TILE;
        mbar_arrive_expect_tx(bar_full + 8u * stage, R_BYTES + C_BYTES + 2 * sf_bytes);
        tma_load_3d(dst, &tmW, bar_full + 8u * stage, kb * BKB, row0, e, pol_w);
        tma_load_2d(dst + R_BYTES, &tmA, bar_full + 8u * stage, kb * BKB, (int)row_base, pol_a);
        bulk_load_1d(dst + R_BYTES + C_BYTES, wsf + (int64_t)kb * SF_KB, sf_bytes, bar_full + 8u * stage);
        bulk_load_1d(dst + R_BYTES + C_BYTES + SF_KB, asf + (int64_t)kb * SF_KB, sf_bytes, bar_full + 8u * stage);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer: scales smem -> TMEM, then the block-scaled MMAs
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t sfb_col_off = BN < 128 ? (uint32_t)((tok0 & 127) >> 5) : 0u;
    for (int kb = 0; kb < num_kb; kb++) {
      mbar_wait(bar_full + 8u * stage, phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sa = base + (uint32_t)stage * STAGE_BYTES;
        const uint32_t ssf_w = sa + R_BYTES + C_BYTES, ssf_a = ssf_w + SF_KB;
        const uint64_t da = make_smem_desc(sa), db = make_smem_desc(sa + R_BYTES);
        const int kt = min(4, k_tiles - kb * 4);
        for (int k = 0; k < kt; k++) {
          tc_cp_sf(tmem_d + SFA_COL + 4u * k, ssf_w + (uint32_t)k * SF_TILE);
          tc_cp_sf(tmem_d + SFB_COL + 4u * k, ssf_a + (uint32_t)k * SF_TILE);
        }
        for (int k = 0; k < kt; k++)
          tc_mma_nvf4(tmem_d, da + (uint64_t)(k * 32 >> 4), db + (uint64_t)(k * 32 >> 4), IDESC,
                      tmem_d + SFA_COL + 4u * k, tmem_d + SFB_COL + 4u * k + sfb_col_off,
                      (kb > 0 || k > 0) ? 1u : 0u);
        tc_commit(bar_empty + 8u * stage);
        if (kb == num_kb - 1) tc_commit(bar_tmem);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else {
    // ============================== epilogue: out[row_base + c][row0 + r] = alpha_e * acc[r][c]
    const int wq = warp & 3;
    const int et = (warp - 2) * 32 + lane;
    const float alpha = p.alphas[e];
    mbar_wait(bar_tmem, 0);
    tc_fence_after();
    auto pack2 = [&](float lo, float hi) -> uint32_t {
      if (p.out_bf16) {
        __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&v);
      }
      __half2 v = __floats2half2_rn(lo, hi);
      return *reinterpret_cast<uint32_t*>(&v);
    };
    const uint32_t tbase = tmem_d + ((uint32_t)(wq * 32) << 16);
    const uint32_t stile = base;   // [BN tokens][128 features] 16-bit, rows of 256 bytes
    const int nchunks = (ntok + 15) >> 4;
    uint32_t v0[16], v1[16];
    tc_ld16(tbase, v0);
    auto stage_chunk = [&](const uint32_t (&vv)[16], int ch) {
#pragma unroll
      for (int j = 0; j < 16; j += 2) {
        const float oa = alpha * __uint_as_float(vv[j]);
        const float ob = alpha * __uint_as_float(vv[j + 1]);
        const float send = (lane & 1) ? oa : ob;
        const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
        const float lo = (lane & 1) ? recv : oa;
        const float hi = (lane & 1) ? ob : recv;
        const int m_local = (ch << 4) + j + (lane & 1);
        const uint32_t addr = stile + (uint32_t)m_local * (BM * 2) + (uint32_t)(wq * 32 + (lane & ~1)) * 2u;
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(pack2(lo, hi)) : "memory");
      }
    };
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
    asm volatile("bar.sync 1, 128;" ::: "memory");
    // 16 threads x 16 bytes per token row, 8 rows per pass
    uint16_t* out16 = reinterpret_cast<uint16_t*>(p.out);
    const int seg = et & 15;
    const int feat = row0 + seg * 8;
    if (feat < p.N) {
      for (int c = et >> 4; c < ntok; c += 8) {
        uint4 v;
        asm volatile("ld.shared.v4.b32 {%0,%1,%2,%3}, [%4];"
                     : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w)
                     : "r"(stile + (uint32_t)c * (BM * 2) + (uint32_t)seg * 16u));
        *reinterpret_cast<uint4*>(out16 + (row_base + c) * p.N + feat) = v;
      }
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_d, TMEM_COLS);
  }
}

template <int BN>
static int launch(const CUtensorMap& tmW, const CUtensorMap& tmA, Params& p, cudaStream_t stream) {
  constexpr uint32_t STAGE_BYTES = BM * BKB + BN * BKB + 2 * SF_KB;
  const int smem_max = max_smem_optin();
  int stages = (smem_max - 1024 - 256 - 64) / (int)STAGE_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  // upper bound of sum_e ceil(m_e / BN) for any routing of max_rows rows over E experts
  int64_t tiles = p.max_rows / BN + p.E;
  if (tiles > p.max_rows) tiles = p.max_rows;
  dim3 grid((unsigned)tiles, (p.N + BM - 1) / BM);
  // many short CTAs: half the ring lets two CTAs share an SM so one streams while the other drains
  // (same reasoning and measurement as moe_gemm.cu)
  const int two_cta = ((228 * 1024) / 2 - 1024 - (16 * MAX_STAGES + 64 + 1024)) / (int)STAGE_BYTES;
  if ((int64_t)grid.x * grid.y >= 2 * (int64_t)sm_count() && two_cta >= 3 && two_cta < stages) stages = two_cta;
  if (stages * (int)STAGE_BYTES < BN * BM * 2) stages = (BN * BM * 2 + (int)STAGE_BYTES - 1) / (int)STAGE_BYTES;
  p.stages = stages;
  const int smem = stages * (int)STAGE_BYTES + 16 * MAX_STAGES + 64 + 1024;
  static bool configured[64] = {};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev >= 0 && dev < 64 && !configured[dev]) {
    B200_CUDA(cudaFuncSetAttribute(moe_gemm_nvfp4_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    configured[dev] = true;
  }
  moe_gemm_nvfp4_kernel<BN><<<grid, NTHREADS, smem, stream>>>(tmW, tmA, p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace moegemm4
}  // namespace b200

extern "C" int b200_nvfp4_moe_gemm(void* out, const void* a, const void* w, const void* a_blockscale,
                                   const void* w_blockscale, const float* alphas, const int32_t* problem_sizes,
                                   const int32_t* expert_offsets, const int32_t* sf_offsets, int64_t max_rows,
                                   int num_experts, int N, int K, int out_dtype, b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::moegemm4;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(max_rows >= 0 && num_experts > 0 && N > 0 && K > 0, B200_ERR_INVALID, "cutlass_fp4_group_mm: bad sizes");
  if (max_rows == 0) return B200_OK;
  B200_CHECK(out && a && w && a_blockscale && w_blockscale && alphas && problem_sizes && expert_offsets && sf_offsets,
             B200_ERR_INVALID, "cutlass_fp4_group_mm: null pointer");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "cutlass_fp4_group_mm: out dtype must be bf16 or f16");
  B200_CHECK(K % 32 == 0 && N % 8 == 0, B200_ERR_INVALID,
             "cutlass_fp4_group_mm: K must be a multiple of 32 and N of 8 (got N=%d K=%d)", N, K);
  B200_CHECK(((uintptr_t)a & 15) == 0 && ((uintptr_t)w & 15) == 0 && ((uintptr_t)a_blockscale & 15) == 0 &&
                 ((uintptr_t)w_blockscale & 15) == 0 && ((uintptr_t)out & 15) == 0,
             B200_ERR_INVALID, "cutlass_fp4_group_mm: tensors must be 16-byte aligned");
  B200_CHECK(max_rows < (1LL << 31), B200_ERR_UNSUPPORTED, "cutlass_fp4_group_mm: too many rows");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "cutlass_fp4_group_mm: no CUDA device (no CPU fallback)");
  const int bn = max_rows <= 64LL * num_experts ? 64 : 128;
  CUtensorMap tmW, tmA;
  {
    uint64_t dims[3] = {(uint64_t)K / 2, (uint64_t)N, (uint64_t)num_experts};
    uint64_t strides[2] = {(uint64_t)K / 2, (uint64_t)(K / 2) * (uint64_t)N};
    uint32_t box[3] = {(uint32_t)BKB, (uint32_t)BM, 1u};
    int rc = encode_tensor_map(&tmW, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, w, dims, strides, box,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
    if (rc) return rc;
  }
  {
    uint64_t dims[2] = {(uint64_t)K / 2, (uint64_t)max_rows};
    uint64_t strides[1] = {(uint64_t)K / 2};
    uint32_t box[2] = {(uint32_t)BKB, (uint32_t)bn};
    int rc = encode_tensor_map(&tmA, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, a, dims, strides, box,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
    if (rc) return rc;
  }
  Params p;
  memset(&p, 0, sizeof(p));
  p.out = out;
  p.a_sf = (const uint8_t*)a_blockscale;
  p.w_sf = (const uint8_t*)w_blockscale;
  p.alphas = alphas;
  p.problem_sizes = problem_sizes;
  p.expert_offsets = expert_offsets;
  p.sf_offsets = sf_offsets;
  p.E = num_experts;
  p.N = N;
  p.K = K;
  p.out_bf16 = out_dtype == B200_BF16;
  p.max_rows = max_rows;
  return bn == 64 ? launch<64>(tmW, tmA, p, stream) : launch<128>(tmW, tmA, p, stream);
}
