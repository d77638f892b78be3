// b200ops -- NVFP4 (E2M1 values, E4M3 scale per 16 elements, fp32 global scale) W4A4 GEMM on
// tcgen05 block-scaled MMA / TMEM / TMA (sm_100a).
//
// Contract: torch.ops._C.cutlass_scaled_fp4_mm (csrc/libtorch_stable/quantization/fp4/
// nvfp4_scaled_mm_kernels.cu:76-140,279-348; nvfp4_scaled_mm_entry.cu; vllm/_custom_ops.py:733-745;
// reference test tests/kernels/quantization/test_nvfp4_scaled_mm.py:27-108):
//   out[M,N] = alpha * sum_k (a[m,k] * sfa[m,k/16]) * (b[n,k] * sfb[n,k/16]),  fp32 accumulate,
//   a [M,K/2] u8 and b [N,K/2] u8 hold two e2m1 codes per byte (low nibble = even k),
//   block_scale_a [round_up(M,128), round_up(K/16,4)] / block_scale_b [round_up(N,128), ...] are
//   e4m3 bytes in the 128x4 tile layout written by scaled_fp4_quant (nvfp4_utils.cuh:186-197),
//   alpha f32[1] on the device, out bf16 / f16.
//
// B200 design. Same skeleton as gemm_fp8.cu (TMA producer warp -> mbarrier ring -> one thread
// issuing tcgen05.mma -> 4 epilogue warps reading TMEM) with the block-scaled instruction
//   tcgen05.mma.cta_group::1.kind::mxf4nvf4.block_scale.block16  (128 x BN x 64 per instruction).
// A 128-byte swizzle row holds 256 e2m1 values = 4 MMAs. The scale factors of one 128-row tile and
// one 64-element k-slice are exactly one 512-byte tile of the reference's swizzled layout, which is
// also byte for byte the image `tcgen05.cp.32x128b.warpx4` wants (row r -> lane r%32, column r/32,
// byte = scale index): so the scales go global -> smem with one 1-D bulk copy per k-block (2 KB per
// 128 rows, contiguous) and smem -> TMEM with one tcgen05.cp per MMA, issued by the MMA thread right
// before the MMAs that consume them (tcgen05.cp and tcgen05.mma execute in issue order).
// Decode (M <= 256): swap-AB, weight rows on the MMA M axis, the padded batch on N, every weight
// byte read from HBM once (0.5625 B per parameter). Large M: activation rows on M, 256 weights on N.
#include "tcgen05.cuh"

#include <cmath>
#include <cstring>

namespace b200 {
namespace gemm4 {

constexpr int BM = 128;        // UMMA M
constexpr int BKB = 128;       // k-block bytes = one 128-byte swizzle row
constexpr int BK = 256;        // k-block elements
constexpr int UMMA_K = 64;     // e2m1 MMA K (32 bytes)
constexpr int SF_TILE = 512;   // bytes of scale factors for 128 rows x 64 k
constexpr int SF_KB = 4 * SF_TILE;  // per 128 rows per k-block
constexpr int NTHREADS = 192;
constexpr int MAX_STAGES = 8;

struct Params {
  void* out;
  const float* alpha;
  const uint8_t* r_sf;   // swizzled scale factors of the rows operand
  const uint8_t* c_sf;   // ... of the cols operand
  int R, C, K;           // rows-operand extent, cols-operand extent, reduction (elements)
  int64_t ldo;
  int swap;              // 1: rows operand = weight (out[c][r]); 0: rows operand = activation
  int out_bf16;
  int stages;
  int tma_store;
};

using namespace b200::tc;

template <int BN>
__global__ void __launch_bounds__(NTHREADS, 1)
    gemm_nvfp4_kernel(const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmC,
                      const __grid_constant__ CUtensorMap tmO, const Params p) {
  constexpr uint32_t R_BYTES = BM * BKB;
  constexpr uint32_t C_BYTES = BN * BKB;
  constexpr int C_TILES = (BN + 127) / 128;           // 128-row scale tiles of the cols operand
  constexpr uint32_t SFR_BYTES = SF_KB;
  constexpr uint32_t SFC_BYTES = C_TILES * SF_KB;
  constexpr uint32_t STAGE_BYTES = R_BYTES + C_BYTES + SFR_BYTES + SFC_BYTES;
  // TMEM: [0, BN) accumulator, then 16 columns of rows-operand scales (4 per MMA), then
  // 4 * 4 * C_TILES columns of cols-operand scales
  constexpr uint32_t ACC_COLS = BN < 32 ? 32 : BN;
  constexpr uint32_t SFA_COL = ACC_COLS;
  constexpr uint32_t SFB_COL = ACC_COLS + 16;
  constexpr uint32_t USED_COLS = SFB_COL + 16 * C_TILES;
  constexpr uint32_t TMEM_COLS = USED_COLS <= 64 ? 64 : USED_COLS <= 128 ? 128 : USED_COLS <= 256 ? 256 : 512;
  // block-scaled instruction descriptor (cute/arch/mma_sm100_desc.hpp InstrDescriptorBlockScaled):
  // a_format = b_format = E2M1 (1), scale_format = UE4M3 (0), K-major A and B, N >> 3, M >> 4, K64
  constexpr uint32_t IDESC = (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(BM >> 4) << 24);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t base = (raw_base + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_base);
  const int stages = p.stages;
  const uint32_t bar_base = base + (uint32_t)stages * STAGE_BYTES;
  const uint32_t bar_full = bar_base;
  const uint32_t bar_empty = bar_base + 8u * MAX_STAGES;
  const uint32_t bar_tmem = bar_base + 16u * MAX_STAGES;
  volatile uint32_t* tmem_holder =
      reinterpret_cast<volatile uint32_t*>(smem + (size_t)stages * STAGE_BYTES + 16 * MAX_STAGES + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = blockIdx.x * BM;
  const int col0 = blockIdx.y * BN;
  const int num_kb = (p.K + BK - 1) / BK;
  const int k_tiles = (p.K + 63) / 64;  // 512-byte scale tiles along K (= padded K/16 / 4)

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
    // ============================== producer: TMA tiles of codes + bulk copies of scale tiles
    // (warp-uniform loop, one elected lane issues: common.cuh elect_one)
    prefetch_tmap(&tmR);
    prefetch_tmap(&tmC);
    const uint64_t pol_w = l2_policy_evict_first();
    const uint64_t pol_a = l2_policy_evict_last();
    const uint64_t pol_r = p.swap ? pol_w : pol_a;
    const uint64_t pol_c = p.swap ? pol_a : pol_w;
    const int c_rows_pad = (p.C + 127) / 128 * 128;
    int c_tiles = 0;
    for (int t = 0; t < C_TILES; t++)
      if (col0 + t * 128 < c_rows_pad) c_tiles++;
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; kb++) {
      mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
      if (elect_one()) {
        const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES;
        const int kt = min(4, k_tiles - kb * 4);  // scale tiles (= MMAs) in this k-block
        const uint32_t sf_bytes = (uint32_t)kt * SF_TILE;
        mbar_arrive_expect_tx(bar_full + 8u * stage, R_BYTES + C_BYTES + sf_bytes * (1 + c_tiles));
        tma_load_2d(dst, &tmR, bar_full + 8u * stage, kb * BKB, row0, pol_r);
        tma_load_2d(dst + R_BYTES, &tmC, bar_full + 8u * stage, kb * BKB, col0, pol_c);
        bulk_load_1d(dst + R_BYTES + C_BYTES,
                     p.r_sf + ((int64_t)(row0 >> 7) * k_tiles + kb * 4) * SF_TILE, sf_bytes,
                     bar_full + 8u * stage);
        for (int t = 0; t < c_tiles; t++)
          bulk_load_1d(dst + R_BYTES + C_BYTES + SFR_BYTES + (uint32_t)t * SF_KB,
                       p.c_sf + ((int64_t)((col0 >> 7) + t) * k_tiles + kb * 4) * SF_TILE, sf_bytes,
                       bar_full + 8u * stage);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer (one elected lane): scales smem->TMEM, then the MMAs
    int stage = 0;
    uint32_t phase = 0;
    const uint32_t sfb_col_off = BN < 128 ? (uint32_t)((col0 & 127) >> 5) : 0u;
    for (int kb = 0; kb < num_kb; kb++) {
      mbar_wait(bar_full + 8u * stage, phase);
      tc_fence_after();
      if (elect_one()) {
        const uint32_t sa = base + (uint32_t)stage * STAGE_BYTES;
        const uint32_t ssf_r = sa + R_BYTES + C_BYTES;
        const uint32_t ssf_c = ssf_r + SFR_BYTES;
        const uint64_t da = make_smem_desc(sa);
        const uint64_t db = make_smem_desc(sa + R_BYTES);
        const int kt = min(4, k_tiles - kb * 4);
        for (int k = 0; k < kt; k++) {
          tc_cp_sf(tmem_d + SFA_COL + 4u * k, ssf_r + (uint32_t)k * SF_TILE);
#pragma unroll
          for (int t = 0; t < C_TILES; t++)
            tc_cp_sf(tmem_d + SFB_COL + 4u * C_TILES * k + 4u * t, ssf_c + (uint32_t)t * SF_KB + (uint32_t)k * SF_TILE);
        }
        for (int k = 0; k < kt; k++) {
          // 64 e2m1 = 32 bytes further inside the 128-byte swizzle row
          // a batch tile narrower than 128 starts (col0 % 128) / 32 columns into its 128-row scale tile
          tc_mma_nvf4(tmem_d, da + (uint64_t)(k * 32 >> 4), db + (uint64_t)(k * 32 >> 4), IDESC,
                      tmem_d + SFA_COL + 4u * k, tmem_d + SFB_COL + 4u * C_TILES * k + sfb_col_off,
                      (kb > 0 || k > 0) ? 1u : 0u);
        }
        tc_commit(bar_empty + 8u * stage);
        if (kb == num_kb - 1) tc_commit(bar_tmem);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else {
    // ============================== epilogue (4 warps, one TMEM lane quarter each): out = alpha * acc
    const int wq = warp & 3;
    mbar_wait(bar_tmem, 0);
    tc_fence_after();
    const int r = row0 + wq * 32 + lane;
    const bool r_ok = r < p.R;
    const float alpha = p.alpha[0];
    uint16_t* out16 = reinterpret_cast<uint16_t*>(p.out);
    auto pack2 = [&](float lo, float hi) -> uint32_t {
      if (p.out_bf16) {
        __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
        return *reinterpret_cast<uint32_t*>(&t);
      }
      __half2 t = __floats2half2_rn(lo, hi);
      return *reinterpret_cast<uint32_t*>(&t);
    };
    const int ncols = min(BN, p.C - col0);
    const int nchunks = (ncols + 15) >> 4;
    const uint32_t tbase = tmem_d + ((uint32_t)(wq * 32) << 16);
    uint32_t v0[16], v1[16];
    tc_ld16(tbase, v0);
    if (p.swap && p.tma_store) {
      // stage the transposed [batch][128 weights] tile in the idle pipeline smem, one TMA store
      const uint32_t stile = base;
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
      fence_proxy_async();
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) tma_store_2d(&tmO, stile, row0, col0);
    } else {
      auto store_chunk = [&](const uint32_t (&vv)[16], int ch) {
        const int c0 = ch << 4;
        if (!r_ok) return;
        if (p.swap) {
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const int c = col0 + c0 + j;
            if (c < p.C) out16[(int64_t)c * p.ldo + r] = (uint16_t)pack2(alpha * __uint_as_float(vv[j]), 0.f);
          }
        } else {
          uint32_t h[8];
#pragma unroll
          for (int j = 0; j < 16; j += 2)
            h[j >> 1] = pack2(alpha * __uint_as_float(vv[j]), alpha * __uint_as_float(vv[j + 1]));
          uint16_t* dst = out16 + (int64_t)r * p.ldo + col0 + c0;
          if (col0 + c0 + 16 <= p.C && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
            reinterpret_cast<uint4*>(dst)[0] = reinterpret_cast<const uint4*>(h)[0];
            reinterpret_cast<uint4*>(dst)[1] = reinterpret_cast<const uint4*>(h)[1];
          } else {
            const uint16_t* h16 = reinterpret_cast<const uint16_t*>(h);
            for (int j = 0; j < 16; j++)
              if (col0 + c0 + j < p.C) dst[j] = h16[j];
          }
        }
      };
      for (int ch = 0; ch < nchunks; ch += 2) {
        tc_wait_ld();
        if (ch + 1 < nchunks) tc_ld16(tbase + (uint32_t)((ch + 1) << 4), v1);
        store_chunk(v0, ch);
        if (ch + 1 < nchunks) {
          tc_wait_ld();
          if (ch + 2 < nchunks) tc_ld16(tbase + (uint32_t)((ch + 2) << 4), v0);
          store_chunk(v1, ch + 1);
        }
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
static int launch(const CUtensorMap& tmR, const CUtensorMap& tmC, const CUtensorMap& tmO, Params& p, dim3 grid,
                  cudaStream_t stream) {
  constexpr int C_TILES = (BN + 127) / 128;
  constexpr uint32_t STAGE_BYTES = BM * BKB + BN * BKB + SF_KB + C_TILES * SF_KB;
  const int smem_max = max_smem_optin();
  int stages = (smem_max - 1024 - 256 - 64) / (int)STAGE_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) {
    set_last_error("cutlass_scaled_fp4_mm: not enough shared memory");
    return B200_ERR_UNSUPPORTED;
  }
  p.stages = stages;
  const int smem = stages * (int)STAGE_BYTES + 16 * MAX_STAGES + 64 + 1024;
  static bool configured_dev[64] = {};   // the attribute is per device
  int cfg_dev_ = 0;
  cudaGetDevice(&cfg_dev_);
  bool& configured = configured_dev[cfg_dev_ & 63];
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(gemm_nvfp4_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   smem_max));
    configured = true;
  }
  gemm_nvfp4_kernel<BN><<<grid, NTHREADS, smem, stream>>>(tmR, tmC, tmO, p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

static int make_tmap_2d(CUtensorMap* tm, const void* base, int rows, int k_bytes, int64_t ld_bytes,
                        int box_rows) {
  uint64_t dims[2] = {(uint64_t)k_bytes, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld_bytes};
  uint32_t box[2] = {(uint32_t)BKB, (uint32_t)box_rows};
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

}  // namespace gemm4
}  // namespace b200

extern "C" int b200_nvfp4_gemm(void* out, const void* a, const void* b, const void* block_scale_a,
                               const void* block_scale_b, const float* alpha, int M, int N, int K,
                               int64_t lda_bytes, int64_t ldb_bytes, int64_t ldo, int out_dtype,
                               b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::gemm4;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(M >= 0 && N > 0 && K > 0, B200_ERR_INVALID, "cutlass_scaled_fp4_mm: bad sizes (%d,%d,%d)",
             M, N, K);
  if (M == 0) return B200_OK;
  B200_CHECK(out && a && b && block_scale_a && block_scale_b && alpha, B200_ERR_INVALID,
             "cutlass_scaled_fp4_mm: null pointer");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "cutlass_scaled_fp4_mm: out dtype must be bf16 or f16");
  B200_CHECK(K % 32 == 0, B200_ERR_INVALID,
             "cutlass_scaled_fp4_mm: Expected k to be divisible by 32, but got a shape (%d, %d)", M, K / 2);
  B200_CHECK(lda_bytes % 16 == 0 && ldb_bytes % 16 == 0, B200_ERR_INVALID,
             "cutlass_scaled_fp4_mm: row strides of a / b must be multiples of 16 bytes");
  B200_CHECK(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0 && ((uintptr_t)block_scale_a & 15) == 0 &&
                 ((uintptr_t)block_scale_b & 15) == 0,
             B200_ERR_INVALID, "cutlass_scaled_fp4_mm: a / b / block scales must be 16-byte aligned");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "cutlass_scaled_fp4_mm: no CUDA device (no CPU fallback)");
  Params p;
  p.out = out;
  p.alpha = alpha;
  p.K = K;
  p.ldo = ldo;
  p.out_bf16 = out_dtype == B200_BF16;
  p.stages = 0;
  const bool swap = M <= 256;
  p.swap = swap ? 1 : 0;
  p.tma_store = 0;
  CUtensorMap tmR, tmC, tmO;
  memset(&tmO, 0, sizeof(tmO));
  int rc, bn;
  if (swap) {
    // rows = weight [N,K], cols = activation [M,K]. Batch tile by the same small cost model as the FP8
    // kernel: per k-block a CTA needs max(MMA cycles, bytes / per-SM ingest); waves over the SMs; fixed
    // prologue + epilogue. Narrow tiles put more CTAs on a weight slab when N is small.
    {
      const int slabs = (N + BM - 1) / BM;
      const int num_kb = (K + BK - 1) / BK;
      const int sms = sm_count();
      const int cands[4] = {256, 128, 64, 32};
      double best = 1e30;
      bn = 32;
      for (int ci = 0; ci < 4; ci++) {
        const int c = cands[ci];
        if (c >= 2 * M && c > 32) continue;   // never pad the batch by more than 2x
        // a 32-wide tile at a non-zero offset would need its scale columns at an odd TMEM column, which
        // the block-scaled MMA rejects (misaligned address); 64-wide tiles use column offsets 0 / 2
        if (c == 32 && M > 32) continue;
        const int gy = (M + c - 1) / c;
        const double waves = ceil((double)slabs * gy / sms);
        const double mma = 4.0 * (128.0 * c / 256.0) / 2.0;           // fp4 MMAs run at twice the fp8 rate
        const double ingest = (BM * BKB + c * BKB + 2048.0 + (c > 128 ? 4096.0 : 2048.0)) / 70.0;
        const double t = waves * (num_kb * (mma > ingest ? mma : ingest) + 5000.0 + 12.0 * c);
        if (t < best) { best = t; bn = c; }
      }
    }
    p.R = N; p.C = M;
    p.r_sf = (const uint8_t*)block_scale_b;
    p.c_sf = (const uint8_t*)block_scale_a;
    rc = make_tmap_2d(&tmR, b, N, K / 2, ldb_bytes, BM);
    if (rc) return rc;
    rc = make_tmap_2d(&tmC, a, M, K / 2, lda_bytes, bn);
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
  } else {
    bn = 256;
    p.R = M; p.C = N;
    p.r_sf = (const uint8_t*)block_scale_a;
    p.c_sf = (const uint8_t*)block_scale_b;
    rc = make_tmap_2d(&tmR, a, M, K / 2, lda_bytes, BM);
    if (rc) return rc;
    rc = make_tmap_2d(&tmC, b, N, K / 2, ldb_bytes, bn);
    if (rc) return rc;
  }
  dim3 grid((p.R + BM - 1) / BM, (p.C + bn - 1) / bn);
  switch (bn) {
    case 32: return launch<32>(tmR, tmC, tmO, p, grid, stream);
    case 64: return launch<64>(tmR, tmC, tmO, p, grid, stream);
    case 128: return launch<128>(tmR, tmC, tmO, p, grid, stream);
    default: return launch<256>(tmR, tmC, tmO, p, grid, stream);
  }
}
