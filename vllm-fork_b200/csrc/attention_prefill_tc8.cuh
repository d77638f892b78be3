// b200ops -- variable-length paged attention on tcgen05 with an FP8-E4M3 KV cache (included by attention_prefill.cu).
//
// Same pipeline as varlen_attn_tc_kernel (attention_prefill_tc.cuh); the tensor core cannot mix a 16-bit Q with 8-bit
// K / V in one instruction, so FP8 pages are expanded to f16 on the way (e4m3 -> f16 is exact, the per-tensor scales
// stay outside: k_scale folded into the softmax scale, v_scale applied to the output row):
//   TMA: raw FP8 pages (16 tokens x 128 B per box, SWIZZLE_128B) -> double-buffered raw stages (32 KB each)
//   4 converter warps: raw K -> f16 K tile in the K-major SWIZZLE_128B image, raw V -> f16 V tile in the same image
//     (= the canonical MN-major layout for P.V); rows past the sequence end are written as zeros
//   MMAs run in kind::f16 with F16 operands (a bf16 Q is converted to f16 when it is staged, P is written as f16).
// Warps 0-3 softmax, 4 TMA producer, 5 MMA issuer + TMEM owner, 6-9 converters.
#pragma once

#include "attention_prefill_tc.cuh"

namespace b200 {
namespace attnp {

constexpr int TC8_THREADS = 320;
constexpr uint32_t TC8_RAW = TC_BN * 128;   // raw FP8 K (or V) tile: 128 tokens x 128 B = 16 KB

template <bool QBF16, bool GENERAL>
__global__ void __launch_bounds__(TC8_THREADS, 1)
    varlen_attn_tc8_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                           const Params p) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kvh = blockIdx.y;
  const int BQ = p.block_q, g = p.g;

  const int bid = blockIdx.x;
  int s;
  {
    int lo = 0, hi = p.S;
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (p.cu_seqlens_q[mid] / BQ + mid <= bid) lo = mid; else hi = mid;
    }
    s = lo;
  }
  const int q_start = p.cu_seqlens_q[s];
  const int q_len = p.cu_seqlens_q[s + 1] - q_start;
  const int qb = bid - (q_start / BQ + s);
  if (qb * BQ >= q_len) return;
  const int L = p.seq_lens[s];
  const int ctx = L - q_len;
  const int first_q = qb * BQ;
  const int last_q = min(q_len, first_q + BQ) - 1;
  const int kv_hi = min(L, ctx + last_q + 1);
  const int kv_lo = p.window > 0 ? max(0, ctx + first_q - p.window + 1) : 0;
  const int j_lo = kv_lo / TC_BN, j_hi = (kv_hi + TC_BN - 1) / TC_BN;
  const int ntiles = j_hi - j_lo;

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t base = (raw_base + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_base);
  const uint32_t q_s = base;                       // Q tile (f16), K-major
  const uint32_t p_s = base + TC_TILE;             // P tile (f16), K-major
  const uint32_t k16_s = base + 2 * TC_TILE;       // expanded K tile
  const uint32_t v16_s = base + 3 * TC_TILE;       // expanded V tile
  const uint32_t raw_s = base + 4 * TC_TILE;       // [stage][raw K | raw V]
  const uint32_t bars = raw_s + 4 * TC8_RAW;
  const uint32_t bar_raw_full = bars, bar_raw_empty = bars + 16;                     // [2] each
  const uint32_t bar_k16_full = bars + 32, bar_k16_free = bars + 40;
  const uint32_t bar_v16_full = bars + 48, bar_v16_free = bars + 56;
  const uint32_t bar_s_full = bars + 64, bar_s_free = bars + 72, bar_p_full = bars + 80, bar_pv_done = bars + 88;
  volatile uint32_t* tmem_holder = reinterpret_cast<volatile uint32_t*>(smem + 4 * TC_TILE + 4 * TC8_RAW + 112);

  if (tid == 0) {
    for (int i = 0; i < 2; i++) {
      mbar_init(bar_raw_full + 8u * i, 1);
      mbar_init(bar_raw_empty + 8u * i, 128);
    }
    mbar_init(bar_k16_full, 128);
    mbar_init(bar_k16_free, 1);
    mbar_init(bar_v16_full, 128);
    mbar_init(bar_v16_free, 1);
    mbar_init(bar_s_full, 1);
    mbar_init(bar_s_free, 128);
    mbar_init(bar_p_full, 128);
    mbar_init(bar_pv_done, 1);
    mbar_fence_init();
  }
  if (warp == 5) {
    const uint32_t holder = smem_u32((const void*)tmem_holder);
    tmem_alloc(holder, 256);
  }

  const int row = tid;   // meaningful for tid < 128
  const int tl = row / g;
  const int qpos = first_q + tl;
  const int head = kvh * g + (row - tl * g);
  const bool valid = tid < 128 && tl < BQ && qpos < q_len;
  const int limit = ctx + qpos;
  if (tid < 128) {
    // Q row -> smem as f16, K-major SWIZZLE_128B (zero rows for padding)
    const uint16_t* qp = reinterpret_cast<const uint16_t*>(p.q) + (int64_t)(q_start + qpos) * p.q_stride_t +
                         (int64_t)head * p.q_stride_h;
#pragma unroll
    for (int c = 0; c < 16; c++) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (valid) v = *reinterpret_cast<const uint4*>(qp + 8 * c);
      if constexpr (QBF16) {
        v.x = bf16x2_to_f16x2(v.x); v.y = bf16x2_to_f16x2(v.y);
        v.z = bf16x2_to_f16x2(v.z); v.w = bf16x2_to_f16x2(v.w);
      }
      sts128(q_s + kmajor_off(row, c), v);
    }
    fence_proxy_async();
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;
  const uint32_t tm_S = tmem_base, tm_O = tmem_base + 128;

  if (warp == 4) {
    // ================================================================ TMA producer: raw FP8 pages
    if (lane == 0) {
      prefetch_tmap(&tmK);
      prefetch_tmap(&tmV);
    }
    const uint64_t pol = l2_policy_evict_last();
    const int32_t* bt = p.block_table + (int64_t)s * p.bt_stride;
    const int sub = lane & 7, is_v = (lane >> 3) & 1;
    for (int i = 0; i < ntiles; i++) {
      const int stage = i & 1;
      const int t0 = (j_lo + i) * TC_BN;
      const int nsub = min(8, (kv_hi - t0 + 15) >> 4);
      mbar_wait(bar_raw_empty + 8u * stage, ((uint32_t)(i >> 1) & 1u) ^ 1u);
      if (lane == 0) mbar_arrive_expect_tx(bar_raw_full + 8u * stage, (uint32_t)nsub * 2u * 2048u);
      __syncwarp();
      if (lane < 16 && sub < nsub) {
        const int token = t0 + 16 * sub;
        const int lb = token / p.bs;
        const int blk = bt[lb];
        const uint32_t dst = raw_s + (uint32_t)stage * 2u * TC8_RAW + (uint32_t)is_v * TC8_RAW + (uint32_t)sub * 2048u;
        tma_load_5d(dst, is_v ? &tmV : &tmK, bar_raw_full + 8u * stage, 0, token - lb * p.bs, 0, kvh, blk, pol);
      }
    }
  } else if (warp == 5) {
    // ================================================================ MMA issuer, kind::f16 / F16: warp-uniform
    // control flow, one elected lane issues (common.cuh elect_one)
    {
      constexpr uint32_t IDESC_S = (1u << 4) | ((uint32_t)(TC_BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
      constexpr uint32_t IDESC_O = (1u << 4) | (1u << 16) | ((uint32_t)(TC_D >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
      auto issue_s = [&](int i) {
        mbar_wait(bar_k16_full, (uint32_t)i & 1u);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < TC_D / 16; ks++) {
            const uint32_t off = (uint32_t)(ks >> 2) * TC_SEG + (uint32_t)(ks & 3) * 32u;
            tc_mma_f16(tm_S, desc_kmajor(q_s + off), desc_kmajor(k16_s + off), IDESC_S, ks > 0 ? 1u : 0u);
          }
          tc_commit(bar_k16_free);   // the expanded K tile may be overwritten
          tc_commit(bar_s_full);
        }
        __syncwarp();
      };
      issue_s(0);
      for (int i = 0; i < ntiles; i++) {
        if (i + 1 < ntiles) {
          mbar_wait(bar_s_free, (uint32_t)i & 1u);
          issue_s(i + 1);
        }
        mbar_wait(bar_p_full, (uint32_t)i & 1u);
        mbar_wait(bar_v16_full, (uint32_t)i & 1u);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < TC_BN / 16; ks++) {
            const uint32_t poff = (uint32_t)(ks >> 2) * TC_SEG + (uint32_t)(ks & 3) * 32u;
            tc_mma_f16(tm_O, desc_kmajor(p_s + poff), desc_mnmajor(v16_s + (uint32_t)ks * 2048u, TC_SEG, 1024u), IDESC_O,
                       (i > 0 || ks > 0) ? 1u : 0u);
          }
          tc_commit(bar_v16_free);
          tc_commit(bar_pv_done);
        }
        __syncwarp();
      }
    }
  } else if (warp >= 6) {
    // ================================================================ converters: raw e4m3 tile -> f16 tile
    const int crow = tid - 192;   // 0..127: one tile row (token) per thread
    auto expand = [&](uint32_t src, uint32_t dst, int rows_valid) {
      const bool live = crow < rows_valid;
#pragma unroll
      for (int c = 0; c < 8; c++) {
        uint4 o0 = make_uint4(0, 0, 0, 0), o1 = make_uint4(0, 0, 0, 0);
        if (live) {
          uint4 r;
          asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
                       : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
                       : "r"(src + (uint32_t)crow * 128u + (uint32_t)((c ^ (crow & 7)) << 4)));
          o0.x = e4m3x2_to_f16x2((uint16_t)(r.x & 0xffffu)); o0.y = e4m3x2_to_f16x2((uint16_t)(r.x >> 16));
          o0.z = e4m3x2_to_f16x2((uint16_t)(r.y & 0xffffu)); o0.w = e4m3x2_to_f16x2((uint16_t)(r.y >> 16));
          o1.x = e4m3x2_to_f16x2((uint16_t)(r.z & 0xffffu)); o1.y = e4m3x2_to_f16x2((uint16_t)(r.z >> 16));
          o1.z = e4m3x2_to_f16x2((uint16_t)(r.w & 0xffffu)); o1.w = e4m3x2_to_f16x2((uint16_t)(r.w >> 16));
        }
        sts128(dst + kmajor_off(crow, 2 * c), o0);
        sts128(dst + kmajor_off(crow, 2 * c + 1), o1);
      }
    };
    for (int i = 0; i < ntiles; i++) {
      const int stage = i & 1;
      const int t0 = (j_lo + i) * TC_BN;
      const int rows_valid = min(TC_BN, kv_hi - t0);   // rows past the sequence end become zeros (K and V)
      const uint32_t rk = raw_s + (uint32_t)stage * 2u * TC8_RAW, rv = rk + TC8_RAW;
      mbar_wait(bar_raw_full + 8u * stage, (uint32_t)(i >> 1) & 1u);
      if (i > 0) mbar_wait(bar_k16_free, (uint32_t)(i - 1) & 1u);   // S(i-1) has read the previous K tile
      expand(rk, k16_s, rows_valid);
      fence_proxy_async();
      mbar_arrive(bar_k16_full);
      if (i > 0) mbar_wait(bar_v16_free, (uint32_t)(i - 1) & 1u);   // P.V(i-1) has read the previous V tile
      expand(rv, v16_s, rows_valid);
      fence_proxy_async();
      mbar_arrive(bar_v16_full);
      mbar_arrive(bar_raw_empty + 8u * stage);
    }
  } else {
    // ================================================================ softmax / correction / epilogue
    SoftmaxCtx c;
    c.tm_S = tm_S; c.tm_O = tm_O; c.p_s = p_s;
    c.bar_s_full = bar_s_full; c.bar_s_free = bar_s_free; c.bar_p_full = bar_p_full; c.bar_pv_done = bar_pv_done;
    c.ntiles = ntiles; c.j_lo = j_lo; c.ctx = ctx; c.first_q = first_q; c.last_q = last_q; c.kv_hi = kv_hi;
    c.limit = limit; c.row = row; c.warp = warp; c.head = head; c.q_start = q_start; c.qpos = qpos;
    c.valid = valid;
    c.qk_scale = p.scale * (p.k_scale ? *p.k_scale : 1.f);
    c.v_scale = p.v_scale ? *p.v_scale : 1.f;
    softmax_warps<false, QBF16, GENERAL>(p, c);
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 5) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
}

}  // namespace attnp
}  // namespace b200
