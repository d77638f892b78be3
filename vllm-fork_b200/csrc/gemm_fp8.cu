// b200ops -- FP8 (E4M3) W8A8 GEMM with scale epilogue on tcgen05 / TMEM / TMA (sm_100a).
//
// Contract: torch.ops._C.cutlass_scaled_mm (csrc/libtorch_stable/quantization/w8a8/cutlass/
// scaled_mm_entry.cu:197-270, sm100 dispatch c3x/scaled_mm_sm100_fp8_dispatch.cuh:18-295; epilogue
// cutlass_extensions/epilogue/scaled_mm_epilogues_c3x.hpp:151-236; vllm/_custom_ops.py:756-804):
//   out[M,N] = a_scale (.) ( b_scale (.) (A[M,K] . B[K,N]) ) (+ bias[N]),  fp32 accumulate,
//   A e4m3 row-major, B e4m3 column-major (= weight [N,K] row-major), scales per-tensor or
//   per-row(A) / per-column(B), out bf16 / f16.
//
// B200 design. Both operands are K-major, so one kernel serves two orientations:
//   "rows" operand -> UMMA M = 128 (TMEM lanes), "cols" operand -> UMMA N = BN (TMEM columns).
//   * decode (M <= 256): swap-AB like the reference's M<=64 buckets: rows = 128 weight rows,
//     cols = the whole (padded) batch, so every weight byte is streamed from HBM exactly once and
//     the accumulator [128 x M] lives in <= 256 TMEM columns;
//   * large M: rows = 128 activation rows, cols = 256 weight rows.
// Warp-specialised: warp 0 = TMA producer (SWIZZLE_128B boxes, mbarrier ring), warp 1 = single-thread
// tcgen05.mma issuer (kind::f8f6f4, 128xBNx32 per instruction) + TMEM owner, warps 2-5 = epilogue
// (tcgen05.ld 32x32b -> scale/bias -> bf16 -> global).
#include "tcgen05.cuh"

#include <cmath>
#include <cstdlib>
#include <cstring>

namespace b200 {
namespace gemm8 {

constexpr int BM = 128;          // UMMA M
constexpr int BK = 128;          // k-block = 128 fp8 = one 128-byte swizzle row
constexpr int UMMA_K = 32;       // fp8 MMA K
constexpr int NTHREADS = 192;
constexpr int MAX_STAGES = 8;

struct Params {
  void* out;
  const float* r_scale;  // scale of the rows operand   (per-row or scalar)
  const float* c_scale;  // scale of the cols operand   (per-col or scalar)
  const void* bias;      // [N] in out dtype, nullable (indexed by the weight dimension)
  int R, C, K;           // rows-operand extent, cols-operand extent, reduction
  int64_t ldo;           // output row stride (elements)
  int r_scale_vec, c_scale_vec;  // 1 = per-row / per-col vector, 0 = scalar
  int swap;              // 1: rows operand = weight (out[c][r]); 0: rows operand = activation (out[r][c])
  int out_bf16;
  int stages;
  int pol_mode;          // tuning knob (B200_GEMM_POL): 0 = weights evict-first / activations evict-last, 1 = both normal
  int prefetch_w;        // PDL level 2: weight tiles of the first ring pass are requested before griddepcontrol.wait
  int tma_store;         // swap mode: stage the tile in smem and write it with one TMA store
  // GATED kernel only (gate_up GEMM + silu_and_mul + static FP8 quant): rows r and gate_rows + r of
  // the weight are the gate / up rows of output column r
  int gate_rows;
  const float* quant_scale;
  // L2-friendly tile order for a 1-D grid (large M): consecutive CTAs walk `raster_group` row tiles of
  // one column tile, then the next column tile, so one wave of 148 CTAs touches ~16 row tiles x ~9 column
  // tiles (35 MB of operands at K = 8192) instead of all row tiles (ncu r1b: 649 MB DRAM reads for 134 MB
  // of operands with the row-major order). 0 = plain 2-D grid.
  int raster_group, row_tiles, col_tiles;
  // development probe (b200_debug_set_trace): 8 x u64 %globaltimer stamps per CTA, null in production
  unsigned long long* trace;
};

static unsigned long long* g_trace = nullptr;
// fine probe: CTA (0,0) only, clock64 per k-block (< 64) into trace[32768 + role * 64 + kb]
__device__ __forceinline__ void probe_kb(const Params& p, int role, int kb) {
  if (p.trace && blockIdx.x == 0 && blockIdx.y == 0 && kb < 64) p.trace[32768 + role * 64 + kb] = clock64();
}
__device__ __forceinline__ void probe(const Params& p, int slot) {
  if (p.trace) {
    unsigned long long t;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
    p.trace[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 8 + slot] = t;
  }
}

using namespace b200::tc;

// silu(x) * y with the roundings of the unfused pair (activation.cu silu_mul_rounded)
template <bool BF16>
__device__ __forceinline__ float gated_act(float g, float u) {
  if constexpr (BF16) {
    const float gf = __bfloat162float(__float2bfloat16_rn(g)), uf = __bfloat162float(__float2bfloat16_rn(u));
    const float s = __bfloat162float(__float2bfloat16_rn(__fdiv_rn(gf, 1.0f + expf(-gf))));
    return __bfloat162float(__float2bfloat16_rn(s * uf));
  } else {
    const float gf = __half2float(__float2half_rn(g)), uf = __half2float(__float2half_rn(u));
    const float s = __half2float(__float2half_rn(__fdiv_rn(gf, 1.0f + expf(-gf))));
    return __half2float(__float2half_rn(s * uf));
  }
}

// CB > 1 (decode, swap-AB): the CB batch tiles of one weight slab form a thread-block cluster (1, CB, 1). Each
// CTA loads 128 / CB rows of every weight k-block and MULTICASTS them into all CB CTAs, so a weight byte crosses
// the L2 -> SM fabric once per cluster instead of once per batch tile; a stage is recycled when the MMAs of all
// CB CTAs have read it (multicast tcgen05.commit on every CTA's `empty` barrier).
template <int BN, bool GATED = false, int CB = 1>
__global__ void __launch_bounds__(NTHREADS, 1)
    gemm_fp8_kernel(const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmC,
                    const __grid_constant__ CUtensorMap tmO, const Params p) {
  constexpr uint32_t R_BYTES = (GATED ? 2 : 1) * BM * BK;   // GATED: gate slab then up slab
  constexpr uint32_t C_BYTES = BN * BK;
  constexpr uint32_t STAGE_BYTES = R_BYTES + C_BYTES;
  constexpr uint32_t ACC_COLS = (GATED ? 2 : 1) * BN;
  constexpr uint32_t TMEM_COLS = ACC_COLS <= 32 ? 32 : ACC_COLS <= 64 ? 64 : ACC_COLS <= 128 ? 128 : ACC_COLS <= 256 ? 256 : 512;
  // instruction descriptor: D = F32, A = B = E4M3, both K-major, N = BN, M = 128
  constexpr uint32_t IDESC = (1u << 4) | (0u << 7) | (0u << 10) | ((uint32_t)(BN >> 3) << 17) |
                             ((uint32_t)(BM >> 4) << 24);

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
  if (threadIdx.x == 0) probe(p, 0);
  pdl_launch_dependents();
  int rt = blockIdx.x, ct = blockIdx.y;
  if (p.raster_group > 0) {
    const int per_group = p.raster_group * p.col_tiles;
    const int grp = blockIdx.x / per_group, within = blockIdx.x - grp * per_group;
    const int rows_here = min(p.raster_group, p.row_tiles - grp * p.raster_group);
    rt = grp * p.raster_group + within % rows_here;
    ct = within / rows_here;
  }
  const int row0 = rt * BM;
  const int col0 = ct * BN;
  const int num_kb = (p.K + BK - 1) / BK;

  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; i++) {
      mbar_init(bar_full + 8u * i, 1);
      mbar_init(bar_empty + 8u * i, CB);
    }
    mbar_init(bar_tmem, 1);
    mbar_fence_init();
  }
  if (warp == 1) {
    const uint32_t holder = smem_u32((const void*)tmem_holder);
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(holder),
                 "r"(TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (CB > 1) cluster_sync_all();   // every CTA's barriers exist before any multicast traffic
  tc_fence_after();
  const uint32_t tmem_d = *tmem_holder;
  if (threadIdx.x == 0) probe(p, 1);

  if (warp == 0) {
    // ============================== TMA producer (warp-uniform loop, one elected lane issues: see elect_one())
    prefetch_tmap(&tmR);
    prefetch_tmap(&tmC);
    // weights are streamed once; activations are re-read by every CTA
    const uint64_t pol_w = p.pol_mode ? l2_policy_evict_normal() : l2_policy_evict_first();
    const uint64_t pol_a = p.pol_mode ? l2_policy_evict_normal() : l2_policy_evict_last();
    const uint64_t pol_r = p.swap ? pol_w : pol_a;
    const uint64_t pol_c = p.swap ? pol_a : pol_w;
    // One k-block = the rows tile (+ the up slab when GATED) and the cols tile. No run-time branch sits between
    // elect_one() and the TMA instructions (ptxas would wrap each UTMALDG in an emulation loop, common.cuh).
    auto load_rows = [&](int stage, int kb) {
      const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES;
      if constexpr (CB > 1) {
        // this CTA's slice of the weight k-block, delivered to every batch tile of the cluster
        constexpr int SL = BM / CB;
        const int rk = blockIdx.y;   // cluster (1, CB, 1), gridDim.y == CB: rank in the cluster
        tma_load_2d_mc(dst + (uint32_t)(rk * SL * BK), &tmR, bar_full + 8u * stage, kb * BK, row0 + rk * SL,
                       (uint16_t)((1u << CB) - 1u), pol_r);
      } else {
        tma_load_2d(dst, &tmR, bar_full + 8u * stage, kb * BK, row0, pol_r);
        if constexpr (GATED) tma_load_2d(dst + BM * BK, &tmR, bar_full + 8u * stage, kb * BK, p.gate_rows + row0, pol_r);
      }
    };
    auto load_cols = [&](int stage, int kb) {
      tma_load_2d(base + (uint32_t)stage * STAGE_BYTES + R_BYTES, &tmC, bar_full + 8u * stage, kb * BK, col0, pol_c);
    };
    int stage = 0, kb_next = 0;
    uint32_t phase = 0;
    if (p.prefetch_w) {
      // Programmatic dependent launch, level 2 (decode, swap-AB: rows = weights): the weight tiles of the first
      // ring pass do not depend on the previous kernel; they are requested BEFORE griddepcontrol.wait and stream
      // from HBM while the predecessor drains. The activation tiles (its output) follow after the wait.
      const int pre = num_kb < stages ? num_kb : stages;
      if (elect_one()) {
        for (int kb = 0; kb < pre; kb++) {
          mbar_arrive_expect_tx(bar_full + 8u * kb, STAGE_BYTES);
          load_rows(kb, kb);
        }
      }
      __syncwarp();
      pdl_wait();
      if (elect_one()) {
        for (int kb = 0; kb < pre; kb++) load_cols(kb, kb);
        probe(p, 2);
      }
      __syncwarp();
      kb_next = pre;
      if (pre == stages) phase = 1u; else stage = pre;
    } else {
      pdl_wait();
    }
    for (int kb = kb_next; kb < num_kb; kb++) {
      mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
      if (elect_one()) {
        probe_kb(p, 0, kb);
        mbar_arrive_expect_tx(bar_full + 8u * stage, STAGE_BYTES);
        load_rows(stage, kb);
        load_cols(stage, kb);
        if (kb == 0) probe(p, 2);
        probe_kb(p, 1, kb);
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer (warp-uniform loop, one elected lane issues)
    pdl_wait();   // nothing to issue before the previous kernel's output arrives: block instead of polling
    int stage = 0;
    uint32_t phase = 0;
    for (int kb = 0; kb < num_kb; kb++) {
      mbar_wait(bar_full + 8u * stage, phase);
      tc_fence_after();
      if (elect_one()) {
        if (kb == 0) probe(p, 3);
        probe_kb(p, 2, kb);
        const uint32_t sa = base + (uint32_t)stage * STAGE_BYTES;
        const uint64_t da = make_smem_desc(sa);
        const uint64_t db = make_smem_desc(sa + R_BYTES);
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; k++) {
          // advance the start address by k * 32 bytes inside the 128-byte swizzle row
          tc_mma_f8(tmem_d, da + (uint64_t)(k * UMMA_K >> 4), db + (uint64_t)(k * UMMA_K >> 4), IDESC,
                    (kb > 0 || k > 0) ? 1u : 0u);
        }
        if constexpr (GATED) {   // the up slab against the same activation tile, second accumulator
          const uint64_t du = make_smem_desc(sa + BM * BK);
#pragma unroll
          for (int k = 0; k < BK / UMMA_K; k++)
            tc_mma_f8(tmem_d + BN, du + (uint64_t)(k * UMMA_K >> 4), db + (uint64_t)(k * UMMA_K >> 4), IDESC,
                      (kb > 0 || k > 0) ? 1u : 0u);
        }
        // frees the smem slot when the MMAs have read it (in every CTA that multicasts into it)
        if constexpr (CB > 1) tc_commit_mc(bar_empty + 8u * stage, (uint16_t)((1u << CB) - 1u));
        else tc_commit(bar_empty + 8u * stage);
        probe_kb(p, 3, kb);
        if (kb == num_kb - 1) {
          tc_commit(bar_tmem);  // accumulator complete
          probe(p, 4);
        }
      }
      __syncwarp();
      if (++stage == stages) { stage = 0; phase ^= 1u; }
    }
  } else {
    // ============================== epilogue (4 warps, one TMEM lane quarter each)
    const int wq = warp & 3;
    pdl_wait();   // scales / bias may be the previous kernel's output; `out` may still be read by it
    mbar_wait(bar_tmem, 0);
    tc_fence_after();
    if (warp == 2 && lane == 0) probe(p, 5);
    const int r = row0 + wq * 32 + lane;  // rows-operand index of this thread
    const bool r_ok = r < p.R;
    const float rs = p.r_scale_vec ? (r_ok ? p.r_scale[r] : 0.f) : p.r_scale[0];
    const float cs0 = p.c_scale_vec ? 0.f : p.c_scale[0];
    uint16_t* out16 = reinterpret_cast<uint16_t*>(p.out);
    const uint16_t* bias16 = reinterpret_cast<const uint16_t*>(p.bias);
    auto to_f32 = [&](uint16_t h) -> float {
      if (p.out_bf16) return __bfloat162float(*reinterpret_cast<const __nv_bfloat16*>(&h));
      return __half2float(*reinterpret_cast<const __half*>(&h));
    };
    auto from_f32 = [&](float f) -> uint16_t {
      if (p.out_bf16) {
        __nv_bfloat16 b = __float2bfloat16_rn(f);
        return *reinterpret_cast<uint16_t*>(&b);
      }
      __half h = __float2half_rn(f);
      return *reinterpret_cast<uint16_t*>(&h);
    };
    float bias_r = 0.f;
    if (p.swap && bias16 && r_ok) bias_r = to_f32(bias16[r]);
    const int ncols = min(BN, p.C - col0);
    const int nchunks = (ncols + 15) >> 4;
    const uint32_t tbase = tmem_d + ((uint32_t)(wq * 32) << 16);
    if constexpr (GATED) {
      // out8[m][r] = e4m3(clamp(act(gate, up) * (1 / quant_scale))) where gate / up are the GEMM outputs
      // rounded to the 16-bit activation dtype exactly like the unfused cutlass_scaled_mm ->
      // silu_and_mul_quant pair. rows = intermediate index r (TMEM lane), cols = batch index m.
      const float rs_u = p.r_scale_vec ? (r_ok ? p.r_scale[p.gate_rows + r] : 0.f) : rs;
      float bias_u = 0.f;
      if (bias16 && r_ok) bias_u = to_f32(bias16[p.gate_rows + r]);
      const float inv = __fdiv_rn(1.0f, *p.quant_scale);
      const uint32_t stile = base;   // [BN m][128 r] bytes, written by one TMA store
      uint8_t* out8 = reinterpret_cast<uint8_t*>(p.out);
      for (int ch = 0; ch < nchunks; ch++) {
        uint32_t g[16], u[16];
        tc_ld16(tbase + (uint32_t)(ch << 4), g);
        tc_ld16(tbase + (uint32_t)(BN + (ch << 4)), u);
        tc_wait_ld();
#pragma unroll
        for (int j = 0; j < 16; j++) {
          const int c = col0 + (ch << 4) + j;
          const float cs = p.c_scale_vec ? (c < p.C ? p.c_scale[c] : 0.f) : cs0;
          const float gv = cs * (rs * __uint_as_float(g[j])) + bias_r;
          const float uv = cs * (rs_u * __uint_as_float(u[j])) + bias_u;
          const float a = p.out_bf16 ? gated_act<true>(gv, uv) : gated_act<false>(gv, uv);
          const uint8_t q = f32_to_e4m3(fmaxf(-448.f, fminf(__fmul_rn(a, inv), 448.f)));
          if (p.tma_store) {
            const uint32_t addr = stile + (uint32_t)((ch << 4) + j) * BM + (uint32_t)(wq * 32 + lane);
            asm volatile("st.shared.u8 [%0], %1;" ::"r"(addr), "r"((uint32_t)q) : "memory");
          } else if (r_ok && c < p.C) {
            out8[(int64_t)c * p.ldo + r] = q;
          }
        }
      }
      if (p.tma_store) {
        fence_proxy_async();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (warp == 2 && lane == 0) tma_store_2d(&tmO, stile, row0, col0);
      }
    } else {
    uint32_t v0[16], v1[16];
    tc_ld16(tbase, v0);
    if (p.swap && p.tma_store) {
      // rows = weight index n (TMEM lane), cols = batch index m. Stage the [m][128 n] tile in the
      // (now idle) pipeline smem and let one TMA store write it: rows of 256 contiguous bytes.
      const uint32_t stile = base;  // 1024-aligned, >= BN * 256 bytes
      auto stage_chunk = [&](const uint32_t (&vv)[16], int ch) {
#pragma unroll
        for (int j = 0; j < 16; j += 2) {
          const int c = col0 + (ch << 4) + j;
          const float cs_a = p.c_scale_vec ? (c < p.C ? p.c_scale[c] : 0.f) : cs0;
          const float cs_b = p.c_scale_vec ? (c + 1 < p.C ? p.c_scale[c + 1] : 0.f) : cs0;
          const float oa = cs_a * (rs * __uint_as_float(vv[j])) + bias_r;
          const float ob = cs_b * (rs * __uint_as_float(vv[j + 1])) + bias_r;
          // 2x2 transpose between lane pairs: even lane ends with column j of rows (n, n+1),
          // odd lane with column j+1 of rows (n-1, n)
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
      fence_proxy_async();  // generic-proxy smem writes -> visible to the TMA (async proxy)
      asm volatile("bar.sync 1, 128;" ::: "memory");
      if (warp == 2 && lane == 0) {
        probe(p, 6);
        asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tmO),
                     "r"(row0), "r"(col0), "r"(stile)
                     : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
      }
    } else {
      auto store_chunk = [&](const uint32_t (&vv)[16], int ch) {
        const int c0 = ch << 4;
        if (!r_ok) return;
        if (p.swap) {
          // rows = weight index n, cols = batch index m : out[m][n]; reference order a_s * (b_s * acc)
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const int c = col0 + c0 + j;
            if (c < p.C) {
              const float cs = p.c_scale_vec ? p.c_scale[c] : cs0;
              const float o = cs * (rs * __uint_as_float(vv[j])) + bias_r;
              out16[(int64_t)c * p.ldo + r] = from_f32(o);
            }
          }
        } else {
          // rows = batch index m, cols = weight index n : out[m][n]
          uint16_t h[16];
#pragma unroll
          for (int j = 0; j < 16; j++) {
            const int c = col0 + c0 + j;
            float o = 0.f;
            if (c < p.C) {
              const float cs = p.c_scale_vec ? p.c_scale[c] : cs0;
              o = rs * (cs * __uint_as_float(vv[j]));
              if (bias16) o += to_f32(bias16[c]);
            }
            h[j] = from_f32(o);
          }
          uint16_t* dst = out16 + (int64_t)r * p.ldo + col0 + c0;
          if (col0 + c0 + 16 <= p.C && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
            reinterpret_cast<uint4*>(dst)[0] = reinterpret_cast<const uint4*>(h)[0];
            reinterpret_cast<uint4*>(dst)[1] = reinterpret_cast<const uint4*>(h)[1];
          } else {
            for (int j = 0; j < 16; j++)
              if (col0 + c0 + j < p.C) dst[j] = h[j];
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
    }  // !GATED
  }
  tc_fence_before();
  __syncthreads();
  if constexpr (CB > 1) cluster_sync_all();   // peers may still commit onto this CTA's barriers
  if (threadIdx.x == 0) probe(p, 7);
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_d), "r"(TMEM_COLS)
                 : "memory");
  }
}

template <int BN, bool GATED = false, int CB = 1>
static int launch(const CUtensorMap& tmR, const CUtensorMap& tmC, const CUtensorMap& tmO, Params& p, dim3 grid,
                  cudaStream_t stream) {
  constexpr uint32_t STAGE_BYTES = (GATED ? 2 : 1) * BM * BK + BN * BK;
  const int smem_max = max_smem_optin();
  int stages = (smem_max - 1024 - 256 - 64) / (int)STAGE_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) {
    set_last_error("fp8_gemm: not enough shared memory");
    return B200_ERR_UNSUPPORTED;
  }
  static const int forced_stages = [] { const char* e = getenv("B200_GEMM_STAGES"); return e ? atoi(e) : 0; }();
  if (forced_stages >= 2 && forced_stages < stages) stages = forced_stages;
  p.stages = stages;
  static const int pol_mode = [] { const char* e = getenv("B200_GEMM_POL"); return e ? atoi(e) : 0; }();
  p.pol_mode = pol_mode;
  p.prefetch_w = (p.swap && pdl_level() >= 2) ? 1 : 0;   // swap-AB: the rows operand is the weight
  p.trace = g_trace;
  const int smem = stages * (int)STAGE_BYTES + 16 * MAX_STAGES + 64 + 1024;
  static bool configured_dev[64] = {};   // the attribute is per device
  int cfg_dev_ = 0;
  cudaGetDevice(&cfg_dev_);
  bool& configured = configured_dev[cfg_dev_ & 63];
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(gemm_fp8_kernel<BN, GATED, CB>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                   smem_max));
    configured = true;
  }
  if constexpr (CB > 1) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid;
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = (size_t)smem;
    cfg.stream = stream;
    cudaLaunchAttribute at[2];
    at[0].id = cudaLaunchAttributeClusterDimension;
    at[0].val.clusterDim.x = 1;
    at[0].val.clusterDim.y = CB;
    at[0].val.clusterDim.z = 1;
    at[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    at[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = at;
    cfg.numAttrs = pdl_enabled() ? 2 : 1;
    B200_CUDA(cudaLaunchKernelEx(&cfg, gemm_fp8_kernel<BN, GATED, CB>, tmR, tmC, tmO, p));
    return B200_OK;
  }
  B200_CUDA(launch_pdl(gemm_fp8_kernel<BN, GATED>, grid, dim3(NTHREADS), (size_t)smem, stream, tmR, tmC, tmO, p));
  return B200_OK;
}

static int make_tmap_2d(CUtensorMap* tm, const void* base, int rows, int K, int64_t ld_bytes,
                        int box_rows) {
  uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld_bytes};
  uint32_t box[2] = {(uint32_t)BK, (uint32_t)box_rows};
  // tuning knob B200_GEMM_L2PROMO: 0 none, 1 64 B, 2 128 B, 3 256 B (default)
  static const int promo = [] { const char* e = getenv("B200_GEMM_L2PROMO"); return e ? atoi(e) : 3; }();
  const CUtensorMapL2promotion l2 = promo == 0 ? CU_TENSOR_MAP_L2_PROMOTION_NONE
                                    : promo == 1 ? CU_TENSOR_MAP_L2_PROMOTION_L2_64B
                                    : promo == 2 ? CU_TENSOR_MAP_L2_PROMOTION_L2_128B
                                                 : CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_128B, l2);
}

}  // namespace gemm8
}  // namespace b200

// development probe: device buffer of 8 x u64 per CTA receiving %globaltimer stamps of the v2 kernel (NULL = off)
extern "C" int b200_debug_set_trace(void* dev_ptr) {
  b200::gemm8::g_trace = reinterpret_cast<unsigned long long*>(dev_ptr);
  return B200_OK;
}

extern "C" int b200_fp8_gemm(void* out, const void* a, const void* b, const float* a_scale,
                             const float* b_scale, const void* bias, int M, int N, int K,
                             int64_t lda, int64_t ldb, int64_t ldo, int a_scale_per_row,
                             int b_scale_per_col, int out_dtype, b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::gemm8;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(M >= 0 && N > 0 && K > 0, B200_ERR_INVALID, "cutlass_scaled_mm: bad sizes (%d,%d,%d)", M,
             N, K);
  if (M == 0) return B200_OK;
  B200_CHECK(out && a && b && a_scale && b_scale, B200_ERR_INVALID,
             "cutlass_scaled_mm: null pointer");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "cutlass_scaled_mm: out dtype must be bf16 or f16");
  B200_CHECK(lda % 16 == 0 && ldb % 16 == 0 && K % 16 == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm: K and the row strides of a / b must be multiples of 16 (got K=%d lda=%lld ldb=%lld)",
             K, (long long)lda, (long long)ldb);
  B200_CHECK(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm: a / b must be 16-byte aligned");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "cutlass_scaled_mm: no CUDA device (no CPU fallback)");
  Params p;
  memset(&p, 0, sizeof(p));
  p.out = out;
  p.bias = bias;
  p.K = K;
  p.ldo = ldo;
  p.out_bf16 = out_dtype == B200_BF16;
  p.stages = 0;
  const bool swap = M <= 256;
  p.swap = swap ? 1 : 0;
  p.tma_store = 0;
  CUtensorMap tmR, tmC, tmO;
  memset(&tmO, 0, sizeof(tmO));
  int rc;
  int bn;
  int cb = 1;   // batch tiles per cluster (weight multicast), swap mode only
  const int sms = sm_count();
  if (swap) {
    // rows = weight [N,K], cols = activation [M,K]. Pick the batch tile BN (and with it the number
    // of CTAs per weight slab) with a small cost model: per k-block a CTA needs max(MMA cycles,
    // bytes / per-SM ingest rate); waves over the SMs; fixed prologue + epilogue.
    const int slabs = (N + BM - 1) / BM;
    const int num_kb = (K + BK - 1) / BK;
    const int cands[5] = {256, 128, 64, 32, 16};
    double best = 1e30;
    bn = 16;
    for (int ci = 0; ci < 5; ci++) {
      const int c = cands[ci];
      if (c >= 2 * M && c > 16) continue;  // never pad the batch by more than 2x
      const int gy = (M + c - 1) / c;
      const double ctas = (double)slabs * gy;
      const double waves = ceil(ctas / sms);
      const double mma = 4.0 * (128.0 * c / 256.0);
      const double ingest = (BM * BK + c * BK) / 70.0;
      const double t = waves * (num_kb * (mma > ingest ? mma : ingest) + 5000.0 + 12.0 * c);
      if (t < best) { best = t; bn = c; }
    }
    static const int forced_bn = [] { const char* e = getenv("B200_GEMM_BN"); return e ? atoi(e) : 0; }();
    if (forced_bn == 16 || forced_bn == 32 || forced_bn == 64 || forced_bn == 128 || forced_bn == 256) bn = forced_bn;
    p.R = N; p.C = M;
    p.r_scale = b_scale; p.r_scale_vec = b_scale_per_col;
    p.c_scale = a_scale; p.c_scale_vec = a_scale_per_row;
    // 2 or 4 batch tiles per weight slab: optionally one thread-block cluster per slab with the weight k-blocks
    // multicast (B200_GEMM_CLUSTER=1)
    // Off by default: measured on B200 (profiles/r2r_gemm_matrix.md) the multicast variant is never faster - the
    // k-block time of these kernels is set by the load latency x ring depth, not by L2 -> SM traffic.
    static const int use_cluster = [] { const char* e = getenv("B200_GEMM_CLUSTER"); return e ? atoi(e) : 0; }();
    const int gy = (M + bn - 1) / bn;
    // (only the <32|64|128, 2> and <32|64, 4> cluster kernels are built: four batch tiles of 128 stay unclustered,
    // the tensor maps below are encoded for the tile width of the kernel that is dispatched)
    if (use_cluster && ((gy == 2 && bn >= 32 && bn <= 128) || (gy == 4 && bn >= 32 && bn <= 64))) cb = gy;
    rc = make_tmap_2d(&tmR, b, N, K, ldb, BM / cb);
    if (rc) return rc;
    rc = make_tmap_2d(&tmC, a, M, K, lda, bn);
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
    p.r_scale = a_scale; p.r_scale_vec = a_scale_per_row;
    p.c_scale = b_scale; p.c_scale_vec = b_scale_per_col;
    rc = make_tmap_2d(&tmR, a, M, K, lda, BM);
    if (rc) return rc;
    rc = make_tmap_2d(&tmC, b, N, K, ldb, bn);
    if (rc) return rc;
  }
  dim3 grid((p.R + BM - 1) / BM, (p.C + bn - 1) / bn);
  if (!swap) {
    p.row_tiles = grid.x;
    p.col_tiles = grid.y;
    p.raster_group = 16;
    grid = dim3(grid.x * grid.y, 1);
  }
  if (cb == 2) {
    switch (bn) {
      case 32: return launch<32, false, 2>(tmR, tmC, tmO, p, grid, stream);
      case 64: return launch<64, false, 2>(tmR, tmC, tmO, p, grid, stream);
      default: return launch<128, false, 2>(tmR, tmC, tmO, p, grid, stream);
    }
  }
  if (cb == 4) {
    switch (bn) {
      case 32: return launch<32, false, 4>(tmR, tmC, tmO, p, grid, stream);
      default: return launch<64, false, 4>(tmR, tmC, tmO, p, grid, stream);
    }
  }
  switch (bn) {
    case 16: return launch<16>(tmR, tmC, tmO, p, grid, stream);
    case 32: return launch<32>(tmR, tmC, tmO, p, grid, stream);
    case 64: return launch<64>(tmR, tmC, tmO, p, grid, stream);
    case 128: return launch<128>(tmR, tmC, tmO, p, grid, stream);
    default: return launch<256>(tmR, tmC, tmO, p, grid, stream);
  }
}


// gate_up GEMM + silu_and_mul + static FP8 quant in one kernel (decode batches, M <= 256):
//   h8[m, i] = e4m3(clamp(T(T(silu(g)) * u) * (1 / *quant_scale))),  g = T(a_s * (b_s * A.Wg^T) + bias_g),
//   u likewise with the up rows; T = act_dtype (bf16 / f16). Bit-identical to cutlass_scaled_mm followed by
//   silu_and_mul_quant (csrc/libtorch_stable/quantization/activation_kernels.cu:33-104), without the
//   [M, 2I] 16-bit round trip through HBM.
extern "C" int b200_fp8_gemm_silu_mul_fp8(void* out8, const void* a, const void* w_gate_up, const float* a_scale,
                                          const float* w_scale, const void* bias, const float* quant_scale,
                                          int M, int inter, int K, int64_t lda, int64_t ldw, int64_t ldo,
                                          int a_scale_per_row, int w_scale_per_row, int act_dtype,
                                          b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::gemm8;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(M >= 0 && inter > 0 && K > 0, B200_ERR_INVALID, "fp8_gemm_silu_mul_fp8: bad sizes (%d,%d,%d)", M, inter, K);
  if (M == 0) return B200_OK;
  B200_CHECK(M <= 256, B200_ERR_UNSUPPORTED,
             "fp8_gemm_silu_mul_fp8: fused kernel covers decode batches (M <= 256); use the unfused pair above");
  B200_CHECK(inter % BM == 0, B200_ERR_UNSUPPORTED,
             "fp8_gemm_silu_mul_fp8: intermediate size %d must be a multiple of 128", inter);
  B200_CHECK(out8 && a && w_gate_up && a_scale && w_scale && quant_scale, B200_ERR_INVALID,
             "fp8_gemm_silu_mul_fp8: null pointer");
  B200_CHECK(act_dtype == B200_BF16 || act_dtype == B200_F16, B200_ERR_INVALID,
             "fp8_gemm_silu_mul_fp8: activation dtype must be bf16 or f16");
  B200_CHECK(lda % 16 == 0 && ldw % 16 == 0 && K % 16 == 0, B200_ERR_INVALID,
             "fp8_gemm_silu_mul_fp8: K and the row strides of a / w must be multiples of 16");
  B200_CHECK(((uintptr_t)a & 15) == 0 && ((uintptr_t)w_gate_up & 15) == 0, B200_ERR_INVALID,
             "fp8_gemm_silu_mul_fp8: a / w must be 16-byte aligned");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "fp8_gemm_silu_mul_fp8: no CUDA device (no CPU fallback)");
  Params p;
  memset(&p, 0, sizeof(p));
  p.out = out8;
  p.bias = bias;
  p.K = K;
  p.ldo = ldo;
  p.out_bf16 = act_dtype == B200_BF16;
  p.swap = 1;
  p.gate_rows = inter;
  p.quant_scale = quant_scale;
  p.R = inter; p.C = M;
  p.r_scale = w_scale; p.r_scale_vec = w_scale_per_row;
  p.c_scale = a_scale; p.c_scale_vec = a_scale_per_row;
  const int bn = M <= 32 ? 32 : M <= 64 ? 64 : M <= 128 ? 128 : 256;
  CUtensorMap tmR, tmC, tmO;
  memset(&tmO, 0, sizeof(tmO));
  int rc = make_tmap_2d(&tmR, w_gate_up, 2 * inter, K, ldw, BM);
  if (rc) return rc;
  rc = make_tmap_2d(&tmC, a, M, K, lda, bn);
  if (rc) return rc;
  if (((uintptr_t)out8 & 15) == 0 && ldo % 16 == 0) {
    uint64_t dims[2] = {(uint64_t)inter, (uint64_t)M};
    uint64_t strides[1] = {(uint64_t)ldo};
    uint32_t box[2] = {(uint32_t)BM, (uint32_t)bn};
    rc = encode_tensor_map(&tmO, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, out8, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE);
    if (rc) return rc;
    p.tma_store = 1;
  }
  dim3 grid(inter / BM, 1);
  switch (bn) {
    case 32: return launch<32, true>(tmR, tmC, tmO, p, grid, stream);
    case 64: return launch<64, true>(tmR, tmC, tmO, p, grid, stream);
    case 128: return launch<128, true>(tmR, tmC, tmO, p, grid, stream);
    default: return launch<256, true>(tmR, tmC, tmO, p, grid, stream);
  }
}
