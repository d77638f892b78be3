// b200ops -- FP8 (E4M3) W8A8 GEMM for decode batches, CTA pairs sharing the batch operand (sm_100a).
//
// Contract: identical to gemm_fp8.cu (torch.ops._C.cutlass_scaled_mm,
// csrc/libtorch_stable/quantization/w8a8/cutlass/scaled_mm_entry.cu:197-270), M <= 256 only:
//   out[M,N] = a_scale (.) ( b_scale (.) (A[M,K] . B[K,N]) ) (+ bias[N]),  fp32 accumulate.
//
// Why (measured on B200, DESIGN.md section 3.4, profiles/r2s_l2fill.jsonl): an SM ingests at most ~65 bytes per
// clock from L2, whatever the number of active SMs, and TMA multicast does not raise that. A decode GEMM whose
// batch tile is re-read by every weight tile and whose weight tile is re-read by every batch tile is therefore
// bound by the bytes each SM has to pull in, not by HBM or the tensor pipe: the one-CTA kernel (gemm_fp8.cu) pulls
// 16 KB of weights + BN x 128 B of activations per k-block and SM. Here two SMs form one
// tcgen05.mma.cta_group::2 instruction of UMMA M = 128: each CTA stages only ITS 64 weight rows (8 KB) and HALF of
// the batch tile, and the tensor cores of the pair read each other's halves through the instruction itself - half
// the ingest per SM for the same tile, no split-K, no partial sums to exchange (the reference's CUTLASS kernel
// for this bucket is a 2-SM 128 x 128 tile too: c3x/scaled_mm_sm100_fp8_dispatch.cuh:118-134).
// Swap-AB as in gemm_fp8.cu: weight rows on UMMA M, the batch on UMMA N (BN in {32, 64, 128, 256}).
// TMEM layout of a 2-SM M = 128 accumulator in each CTA: lanes 0-63 = its 64 rows x columns [0, BN/2),
// lanes 64-127 = the same 64 rows x columns [BN/2, BN).
// Roles per CTA: warp 0 = TMA producer, warp 1 = MMA issuer (even CTA of the pair) + TMEM owner,
// warps 2-5 = epilogue (scale / bias -> 16-bit -> transposed staging tile -> one TMA store).
#include "tcgen05.cuh"

#include <cstdlib>
#include <cstring>

namespace b200 {
namespace gemm8v5 {

constexpr int ROWS = 64;         // weight rows per CTA (UMMA M = 128 per pair)
constexpr int BK = 128;          // k-block bytes
constexpr int UMMA_K = 32;
constexpr int NTHREADS = 192;
constexpr int MAX_STAGES = 12;
constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;  // clears the low CTA-rank bit of a shared::cluster address

struct Params {
  const float* r_scale;   // weight scale (per-row vector or scalar)
  const float* c_scale;   // activation scale (per-token vector or scalar)
  const void* bias;       // [N] in out dtype, nullable
  int R, C, K;            // weight rows (N), batch (M), reduction
  int r_scale_vec, c_scale_vec;
  int out_bf16, stages, num_kb;
  int prefetch_w;         // PDL level 2: weight tiles of the first ring pass precede griddepcontrol.wait
};

using namespace b200::tc;

__device__ __forceinline__ uint32_t cluster_ctarank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
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
// arrive on the barrier at this smem offset in both CTAs of the pair when the prior MMAs complete
__device__ __forceinline__ void tc_commit_2sm(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}

template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
    gemm_fp8_v5_kernel(const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmC,
                       const __grid_constant__ CUtensorMap tmO, const Params p) {
  constexpr uint32_t R_BYTES = ROWS * BK;           // this CTA's 64 weight rows
  constexpr uint32_t C_BYTES = (BN / 2) * BK;       // this CTA's half of the batch tile
  constexpr uint32_t STAGE_BYTES = R_BYTES + C_BYTES;
  constexpr uint32_t TMEM_COLS = BN / 2 < 32 ? 32 : BN / 2;
  constexpr int NCHUNK = BN / 32;                   // 16-column chunks per lane half
  // instruction descriptor: D = F32, A = B = E4M3, K-major, N = BN, M = 128 (pair)
  constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(128 >> 4) << 24);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t base = (raw_base + 1023u) & ~1023u;   // pipeline stages; after the main loop: staging tile
  uint8_t* smem = smem_raw + (base - raw_base);
  const int stages = p.stages;
  const uint32_t bar_base = base + (uint32_t)stages * STAGE_BYTES;
  const uint32_t bar_full = bar_base;                          // [stages]   (the even CTA's copy is used)
  const uint32_t bar_empty = bar_base + 8u * MAX_STAGES;       // [stages]   per CTA
  const uint32_t bar_tfull = bar_base + 16u * MAX_STAGES;      // per CTA
  volatile uint32_t* tmem_holder =
      reinterpret_cast<volatile uint32_t*>(smem + (size_t)stages * STAGE_BYTES + 16 * MAX_STAGES + 16);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t h = cluster_ctarank();    // half of the 128-row tile / of the batch tile
  const int tile = blockIdx.x >> 1;
  const int row0 = tile * 2 * ROWS + (int)h * ROWS;   // this CTA's first weight row
  const int col0 = blockIdx.y * BN;                   // first batch row of the pair's tile
  const int num_kb = p.num_kb;
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
  cluster_sync_all();   // both CTAs' barriers are initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  if (warp == 0) {
    // ============================== TMA producer (warp-uniform loop, one elected lane issues: common.cuh elect_one)
    prefetch_tmap(&tmR);
    prefetch_tmap(&tmC);
    const uint64_t pol_w = l2_policy_evict_first();   // weights stream once
    const uint64_t pol_a = l2_policy_evict_last();    // the batch is re-read by every pair
    const int acol = col0 + (int)h * (BN / 2);
    // Programmatic dependent launch: weight tiles of the first ring pass before griddepcontrol.wait (level 2),
    // the activation tiles (the previous kernel's output) after it
    const int pre = num_kb < stages ? num_kb : stages;
    if (!p.prefetch_w) pdl_wait();
    if (elect_one()) {
      for (int i = 0; i < pre; i++) {
        if (h == 0) mbar_arrive_expect_tx(bar_full + 8u * i, 2 * STAGE_BYTES);
        tma_load_2d_2sm(base + (uint32_t)i * STAGE_BYTES, &tmR, bar_full + 8u * i, i * BK, row0, pol_w);
        if (!p.prefetch_w)
          tma_load_2d_2sm(base + (uint32_t)i * STAGE_BYTES + R_BYTES, &tmC, bar_full + 8u * i, i * BK, acol, pol_a);
      }
    }
    __syncwarp();
    if (p.prefetch_w) {
      pdl_wait();
      if (elect_one()) {
        for (int i = 0; i < pre; i++)
          tma_load_2d_2sm(base + (uint32_t)i * STAGE_BYTES + R_BYTES, &tmC, bar_full + 8u * i, i * BK, acol, pol_a);
      }
      __syncwarp();
    }
    int stage = pre == stages ? 0 : pre;
    uint32_t phase = pre == stages ? 1u : 0u;
    for (int kb = pre; kb < num_kb; kb++) {
      mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
      if (elect_one()) {
        const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES;
        if (h == 0) mbar_arrive_expect_tx(bar_full + 8u * stage, 2 * STAGE_BYTES);
        tma_load_2d_2sm(dst, &tmR, bar_full + 8u * stage, kb * BK, row0, pol_w);
        tma_load_2d_2sm(dst + R_BYTES, &tmC, bar_full + 8u * stage, kb * BK, acol, pol_a);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer: one elected lane of the even CTA drives both tensor cores
    pdl_wait();   // nothing to issue before the previous kernel's output arrives: block instead of polling
    if (h == 0) {
      int stage = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < num_kb; kb++) {
        mbar_wait(bar_full + 8u * stage, phase);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t sa = base + (uint32_t)stage * STAGE_BYTES;
          const uint64_t da = make_smem_desc(sa);
          const uint64_t db = make_smem_desc(sa + R_BYTES);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; k++)
            tc_mma_f8_2sm(tmem_base, da + (uint64_t)(k * UMMA_K >> 4), db + (uint64_t)(k * UMMA_K >> 4), IDESC,
                          (kb > 0 || k > 0) ? 1u : 0u);
          tc_commit_2sm(bar_empty + 8u * stage);
          if (kb == num_kb - 1) tc_commit_2sm(bar_tfull);
        }
        __syncwarp();
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
    }
    __syncwarp();
  } else {
    // ============================== epilogue: 4 warps, one TMEM lane quarter each
    pdl_wait();                // scales / bias / `out` are shared with the previous kernel on the stream
    const int wq = warp & 3;
    const int row_l = (wq & 1) * 32 + lane;         // weight row inside this CTA's 64
    const int cbase = (wq >> 1) * (BN / 2);         // lanes 64-127 hold the second half of the batch columns
    const int r = row0 + row_l;
    const bool r_ok = r < p.R;
    const float rs = p.r_scale_vec ? (r_ok ? p.r_scale[r] : 0.f) : p.r_scale[0];
    const float cs0 = p.c_scale_vec ? 0.f : p.c_scale[0];
    const uint16_t* bias16 = reinterpret_cast<const uint16_t*>(p.bias);
    float bias_r = 0.f;
    if (bias16 && r_ok)
      bias_r = p.out_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(bias16)[r])
                          : __half2float(reinterpret_cast<const __half*>(bias16)[r]);
    mbar_wait(bar_tfull, 0);   // the pair's accumulator is complete, the pipeline smem of both CTAs idle
    tc_fence_after();
    const uint32_t tbase = tmem_base + ((uint32_t)(wq * 32) << 16);
    const uint32_t stile = base;   // [BN m][64 r] 16-bit = BN x 128 bytes, written by one TMA store
    uint32_t v0[16], v1[16];
    tc_ld16(tbase, v0);
    auto stage_chunk = [&](const uint32_t (&vv)[16], int ch) {
#pragma unroll
      for (int j = 0; j < 16; j += 2) {
        const int ml = cbase + (ch << 4) + j;       // batch column inside the tile
        const int c = col0 + ml;
        const float cs_a = p.c_scale_vec ? (c < p.C ? p.c_scale[c] : 0.f) : cs0;
        const float cs_b = p.c_scale_vec ? (c + 1 < p.C ? p.c_scale[c + 1] : 0.f) : cs0;
        // reference order a_s * (b_s * acc) + bias
        const float oa = cs_a * (rs * __uint_as_float(vv[j])) + bias_r;
        const float ob = cs_b * (rs * __uint_as_float(vv[j + 1])) + bias_r;
        // 2x2 transpose between lane pairs: even lane ends with column j of rows (n, n+1), odd lane with j+1
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
        const uint32_t addr = stile + (uint32_t)(ml + (lane & 1)) * (ROWS * 2) + (uint32_t)(row_l & ~1) * 2u;
        asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(packed) : "memory");
      }
    };
#pragma unroll 1
    for (int ch = 0; ch < NCHUNK; ch += 2) {
      tc_wait_ld();
      if (ch + 1 < NCHUNK) tc_ld16(tbase + (uint32_t)((ch + 1) << 4), v1);
      stage_chunk(v0, ch);
      if (ch + 1 < NCHUNK) {
        tc_wait_ld();
        if (ch + 2 < NCHUNK) tc_ld16(tbase + (uint32_t)((ch + 2) << 4), v0);
        stage_chunk(v1, ch + 1);
      }
    }
    fence_proxy_async();  // generic-proxy smem writes -> visible to the TMA (async proxy)
    asm volatile("bar.sync 1, 128;" ::: "memory");
    if (warp == 2 && lane == 0) tma_store_2d(&tmO, stile, row0, col0);
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync_all();   // the peer may still commit onto this CTA's barriers until here
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

template <int BN>
static int launch(const CUtensorMap& tmR, const CUtensorMap& tmC, const CUtensorMap& tmO, Params& p, dim3 grid,
                  cudaStream_t stream) {
  constexpr int STAGE_BYTES = ROWS * BK + (BN / 2) * BK;
  const int smem_max = max_smem_optin();
  const int fixed = 1024 + 16 * MAX_STAGES + 64;
  int stages = (smem_max - fixed) / STAGE_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  static const int forced_stages = [] { const char* e = getenv("B200_GEMM5_STAGES"); return e ? atoi(e) : 0; }();
  if (forced_stages >= 2 && forced_stages < stages) stages = forced_stages;
  // the staging tile of the epilogue ([BN][64] 16-bit) reuses the pipeline stages
  if (stages < 2 || stages * STAGE_BYTES < BN * ROWS * 2) {
    set_last_error("fp8_gemm_pair: not enough shared memory");
    return B200_ERR_UNSUPPORTED;
  }
  p.stages = stages;
  p.prefetch_w = pdl_level() >= 2;
  const int smem = stages * STAGE_BYTES + fixed;
  static bool configured_dev[64] = {};   // the attribute is per device
  int dev = 0;
  cudaGetDevice(&dev);
  if (!configured_dev[dev & 63]) {
    B200_CUDA(cudaFuncSetAttribute(gemm_fp8_v5_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    configured_dev[dev & 63] = true;
  }
  B200_CUDA(launch_pdl(gemm_fp8_v5_kernel<BN>, grid, dim3(NTHREADS), (size_t)smem, stream, tmR, tmC, tmO, p));
  return B200_OK;
}

static int make_tmap_2d(CUtensorMap* tm, const void* base, int rows, int K, int64_t ld_bytes, int box_rows) {
  uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld_bytes};
  uint32_t box[2] = {(uint32_t)BK, (uint32_t)box_rows};
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

// batch tile: the smallest BN whose grid (2 CTAs per 128-row weight tile x batch tiles) still fits one wave, i.e.
// as many SMs as possible pulling as few bytes each as possible; at least 32 (UMMA N of a 2-SM instruction)
static int pick_bn(int M, int N, int sms) {
  const int tiles = (N + 2 * ROWS - 1) / (2 * ROWS);
  const int cands[4] = {32, 64, 128, 256};
  for (int i = 0; i < 4; i++) {
    const int gy = (M + cands[i] - 1) / cands[i];
    if (2 * tiles * gy <= sms) return cands[i];
  }
  return 256;
}

}  // namespace gemm8v5
}  // namespace b200

extern "C" int b200_fp8_gemm_pair(void* out, const void* a, const void* b, const float* a_scale,
                                  const float* b_scale, const void* bias, int M, int N, int K, int64_t lda,
                                  int64_t ldb, int64_t ldo, int a_scale_per_row, int b_scale_per_col,
                                  int out_dtype, int batch_tile, b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::gemm8v5;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(M >= 0 && N > 0 && K > 0, B200_ERR_INVALID, "cutlass_scaled_mm: bad sizes (%d,%d,%d)", M, N, K);
  if (M == 0) return B200_OK;
  B200_CHECK(M <= 256, B200_ERR_UNSUPPORTED, "fp8_gemm_pair: decode batches only (M <= 256), got %d", M);
  B200_CHECK(batch_tile == 0 || batch_tile == 32 || batch_tile == 64 || batch_tile == 128 || batch_tile == 256,
             B200_ERR_INVALID, "fp8_gemm_pair: batch_tile must be 0 (auto), 32, 64, 128 or 256");
  B200_CHECK(out && a && b && a_scale && b_scale, B200_ERR_INVALID, "cutlass_scaled_mm: null pointer");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "cutlass_scaled_mm: out dtype must be bf16 or f16");
  B200_CHECK(lda % 16 == 0 && ldb % 16 == 0 && K % 16 == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm: K and the row strides of a / b must be multiples of 16 (got K=%d lda=%lld ldb=%lld)",
             K, (long long)lda, (long long)ldb);
  B200_CHECK(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm: a / b must be 16-byte aligned");
  B200_CHECK(((uintptr_t)out & 15) == 0 && (ldo * 2) % 16 == 0, B200_ERR_UNSUPPORTED,
             "fp8_gemm_pair: out must be 16-byte aligned with a 16-byte multiple row stride");
  const int sms = sm_count();
  B200_CHECK(sms > 0, B200_ERR_CUDA, "cutlass_scaled_mm: no CUDA device (no CPU fallback)");
  Params p;
  memset(&p, 0, sizeof(p));
  p.bias = bias;
  p.K = K;
  p.out_bf16 = out_dtype == B200_BF16;
  p.num_kb = (K + BK - 1) / BK;
  p.R = N; p.C = M;
  p.r_scale = b_scale; p.r_scale_vec = b_scale_per_col;
  p.c_scale = a_scale; p.c_scale_vec = a_scale_per_row;
  static const int forced_bn = [] { const char* e = getenv("B200_GEMM5_BN"); return e ? atoi(e) : 0; }();
  int bn = batch_tile ? batch_tile : pick_bn(M, N, sms);
  if (!batch_tile && (forced_bn == 32 || forced_bn == 64 || forced_bn == 128 || forced_bn == 256)) bn = forced_bn;
  CUtensorMap tmR, tmC, tmO;
  int rc = make_tmap_2d(&tmR, b, N, K, ldb, ROWS);
  if (rc) return rc;
  rc = make_tmap_2d(&tmC, a, M, K, lda, bn / 2);
  if (rc) return rc;
  {
    uint64_t dims[2] = {(uint64_t)N, (uint64_t)M};
    uint64_t strides[1] = {(uint64_t)ldo * 2};
    uint32_t box[2] = {(uint32_t)ROWS, (uint32_t)bn};
    rc = encode_tensor_map(&tmO, CU_TENSOR_MAP_DATA_TYPE_UINT16, 2, out, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE);
    if (rc) return rc;
  }
  const int tiles = (N + 2 * ROWS - 1) / (2 * ROWS);
  dim3 grid(2 * tiles, (M + bn - 1) / bn);
  switch (bn) {
    case 32: return launch<32>(tmR, tmC, tmO, p, grid, stream);
    case 64: return launch<64>(tmR, tmC, tmO, p, grid, stream);
    case 128: return launch<128>(tmR, tmC, tmO, p, grid, stream);
    default: return launch<256>(tmR, tmC, tmO, p, grid, stream);
  }
}
