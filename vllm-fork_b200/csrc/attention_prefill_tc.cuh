// b200ops -- variable-length paged attention on tcgen05 / TMEM / TMA (included by attention_prefill.cu).
//
// Same contract as varlen_attn_kernel (attention_prefill.cu); covers bf16 / f16 KV caches with head_size 128.
// One CTA = one q block of 128 rows ((128 / g query tokens) x the g query heads of one kv head), looping over
// KV tiles of 128 tokens:
//   S[128 x 128] = Q . K^T      tcgen05.mma kind::f16, Q and K K-major from smem (SWIZZLE_128B), S in TMEM
//   softmax                      4 warps, one thread per row: tcgen05.ld the row, exp2-domain online softmax with a
//                                LAZY running max (O is rescaled only when the max grows by > 2^8), P (bf16 / f16)
//                                written to smem in the K-major swizzled layout
//   O[128 x 128] += P . V        tcgen05.mma, P K-major from smem, V as an MN-major B operand: the paged
//                                [token][head_dim] image the TMA drops is exactly the canonical MN-major
//                                SWIZZLE_128B layout (64 head-dim elements contiguous, 8-token groups 1024 B apart)
// K/V pages are fetched by TMA, one 16-token x 128-byte box per (page, 128-byte segment), through the same 5-D
// tensor map geometry as the decode kernel; K and V tiles are double-buffered, S(j+1) is issued as soon as the
// softmax warps have pulled S(j) into registers, so the tensor pipe works under the exponentials.
// Warps 0-3 = softmax / correction / epilogue (TMEM lane quarter = warp), warp 4 = TMA producer,
// warp 5 = MMA issuer + TMEM owner.
#pragma once

#include "tcgen05.cuh"

namespace b200 {
namespace attnp {

constexpr int TC_BM = 128;        // rows per q block
constexpr int TC_BN = 128;        // kv tokens per tile
constexpr int TC_D = 128;
constexpr int TC_THREADS = 192;
constexpr uint32_t TC_SEG = TC_BN * 128;          // one 128-byte segment of a tile: 128 rows x 128 B = 16 KB
constexpr uint32_t TC_TILE = 2 * TC_SEG;          // K (or V, Q, P) tile: 2 segments = 32 KB
constexpr float TC_RESCALE = 8.0f;                // log2 units

using namespace b200::tc;

// K-major SWIZZLE_128B descriptor: rows 128 B apart, 8-row groups 1024 B apart (one 64-element k-atom)
__device__ __forceinline__ uint64_t desc_kmajor(uint32_t addr) { return make_smem_desc(addr); }
// MN-major SWIZZLE_128B descriptor (cute make_umma_desc<Major::MN>, mma_traits_sm100.hpp:238-269):
// ((8,n),(8,k)) : ((1,LBO),(8,SBO)) in 16-byte units: 64 MN elements contiguous, next 64 at LBO, k rows 128 B
// apart, 8-row groups at SBO
__device__ __forceinline__ uint64_t desc_mnmajor(uint32_t addr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((addr & 0x3FFFFu) >> 4);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFFu) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFFu) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
__device__ __forceinline__ void tc_mma_f16(uint32_t tmem_d, uint64_t da, uint64_t db, uint32_t idesc,
                                           uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(da), "l"(db), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tc_st16(uint32_t taddr, const uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16};" ::"r"(taddr),
      "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]), "r"(v[8]), "r"(v[9]),
      "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
      : "memory");
}
__device__ __forceinline__ void tc_wait_st() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ void sts128(uint32_t addr, uint4 v) {
  asm volatile("st.shared.v4.u32 [%0], {%1,%2,%3,%4};" ::"r"(addr), "r"(v.x), "r"(v.y), "r"(v.z), "r"(v.w) : "memory");
}
// byte offset of 16-byte chunk c (0..15) of row r inside a K-major SWIZZLE_128B [128 rows][128 elems x 2 B] tile
__device__ __forceinline__ uint32_t kmajor_off(int r, int c) {
  return (uint32_t)((c >> 3) * TC_SEG + r * 128 + (((c & 7) ^ (r & 7)) << 4));
}

// Everything the softmax warps need to know about their q block
struct SoftmaxCtx {
  uint32_t tm_S, tm_O, p_s, bar_s_full, bar_s_free, bar_p_full, bar_pv_done;
  int ntiles, j_lo, ctx, first_q, last_q, kv_hi, limit, row, warp, head, q_start, qpos;
  bool valid;
  float qk_scale, v_scale;
};

// Softmax / correction / epilogue warps (4 warps, one thread per q row). PBF16: P is written as bf16 (else f16);
// OBF16: output dtype; GENERAL: soft-cap and / or ALiBi.
template <bool PBF16, bool OBF16, bool GENERAL>
__device__ __forceinline__ void softmax_warps(const Params& p, const SoftmaxCtx& cx) {
  const float slope = (p.alibi && cx.valid) ? p.alibi[cx.head] : 0.f;
  const float qk_scale = cx.qk_scale;
  const float softcap = p.softcap;
  const uint32_t lane_base = (uint32_t)(cx.warp * 32) << 16;
  float m_ref = -INFINITY, l_sum = 0.f;
  for (int i = 0; i < cx.ntiles; i++) {
    const int t0 = (cx.j_lo + i) * TC_BN;
    mbar_wait(cx.bar_s_full, (uint32_t)i & 1u);
    tc_fence_after();
    float sv[TC_BN];
    {
      uint32_t v[8][16];
#pragma unroll
      for (int c = 0; c < 8; c++) tc_ld16(cx.tm_S + lane_base + (uint32_t)(c << 4), v[c]);
      tc_wait_ld();
#pragma unroll
      for (int c = 0; c < 8; c++)
#pragma unroll
        for (int e = 0; e < 16; e++) sv[c * 16 + e] = __uint_as_float(v[c][e]);
    }
    tc_fence_before();
    mbar_arrive(cx.bar_s_free);
    const bool need_mask = (t0 + TC_BN - 1 > cx.ctx + cx.first_q) || (t0 + TC_BN > cx.kv_hi) ||
                           (p.window > 0 && t0 < cx.ctx + cx.last_q - p.window + 1);
    // visible columns of this cx.row inside the tile: [lo_e, hi_e] (causal cx.limit, sequence end, sliding window)
    const int hi_e = min(cx.limit, cx.kv_hi - 1) - t0;
    const int lo_e = (p.window > 0 ? cx.limit - p.window + 1 : 0) - t0;
    if constexpr (GENERAL) {
      const float a0 = slope * (float)(t0 - cx.ctx);
#pragma unroll
      for (int e = 0; e < TC_BN; e++) {
        float v = sv[e] * qk_scale;
        if (softcap > 0.f) v = softcap * tanhf(v / softcap);
        sv[e] = (v + fmaf(slope, (float)e, a0)) * LOG2E;
      }
    }
    float mx = -INFINITY;
    if (need_mask) {
#pragma unroll
      for (int e = 0; e < TC_BN; e++) {
        if (e > hi_e || e < lo_e) sv[e] = -INFINITY;
        mx = fmaxf(mx, sv[e]);
      }
    } else {
#pragma unroll
      for (int e = 0; e < TC_BN; e++) mx = fmaxf(mx, sv[e]);
    }
    const float cexp = GENERAL ? 1.f : qk_scale * LOG2E;   // exp2 argument = sv * cexp - m_ref
    mx *= cexp;                                            // cexp > 0: max commutes, -inf stays -inf
    // lazy running max: re-base only when the cx.row max grew by more than 2^8 (exp2 arguments stay <= 8)
    const bool grow = i > 0 && mx > m_ref + TC_RESCALE;
    const float f = grow ? ex2(m_ref - mx) : 1.f;    // m_ref = -inf -> 0
    if (i == 0 || grow) m_ref = mx;
    l_sum *= f;
    const float mb = m_ref == -INFINITY ? 0.f : m_ref;
    // exponentials first (they only need S): this overlaps P.V of the previous tile on the tensor pipe
    uint32_t pw[TC_BN / 2];
#pragma unroll
    for (int k = 0; k < TC_BN / 2; k++) {
      const float e0 = ex2(fmaf(sv[2 * k], cexp, -mb)), e1 = ex2(fmaf(sv[2 * k + 1], cexp, -mb));
      l_sum += e0 + e1;
      pw[k] = pack2<PBF16>(e0, e1);
    }
    if (i > 0) {
      mbar_wait(cx.bar_pv_done, (uint32_t)(i - 1) & 1u);   // P smem free again, O complete up to tile i - 1
      tc_fence_after();
      if (__any_sync(0xffffffffu, grow)) {
#pragma unroll 1
        for (int c = 0; c < TC_D / 16; c++) {
          uint32_t o[16];
          tc_ld16(cx.tm_O + lane_base + (uint32_t)(c << 4), o);
          tc_wait_ld();
#pragma unroll
          for (int e = 0; e < 16; e++) o[e] = __float_as_uint(__uint_as_float(o[e]) * f);
          tc_st16(cx.tm_O + lane_base + (uint32_t)(c << 4), o);
        }
        tc_wait_st();
      }
    }
#pragma unroll
    for (int c = 0; c < 16; c++)
      sts128(cx.p_s + kmajor_off(cx.row, c), make_uint4(pw[4 * c], pw[4 * c + 1], pw[4 * c + 2], pw[4 * c + 3]));
    fence_proxy_async();    // P (generic proxy) -> tensor core (async proxy)
    tc_fence_before();      // O re-base (tcgen05.st) ordered before the MMA that accumulates into it
    mbar_arrive(cx.bar_p_full);
  }
  // ---- epilogue: O / l -> out
  mbar_wait(cx.bar_pv_done, (uint32_t)(cx.ntiles - 1) & 1u);
  tc_fence_after();
  const bool fp8_out = p.out_scale != nullptr;
  const float qinv = fp8_out ? __fdiv_rn(1.0f, *p.out_scale) : 1.f;
  const float inv = l_sum > 0.f ? cx.v_scale / l_sum : 0.f;
  const int64_t tok = cx.q_start + cx.qpos;
  if (cx.valid && p.lse) p.lse[tok * p.Hq + cx.head] = l_sum > 0.f ? (m_ref + lg2(l_sum)) * LN2 : -INFINITY;
  const int64_t obase = tok * p.out_stride_t + (int64_t)cx.head * p.out_stride_h;
#pragma unroll 1
  for (int c = 0; c < TC_D / 16; c++) {
    uint32_t o[16];
    tc_ld16(cx.tm_O + lane_base + (uint32_t)(c << 4), o);
    tc_wait_ld();
    if (!cx.valid) continue;
    if (fp8_out) {
      uint32_t w[4];
#pragma unroll
      for (int k = 0; k < 4; k++) {
        uint32_t b4 = 0;
#pragma unroll
        for (int h2 = 0; h2 < 2; h2++) {
          const float a = fmaxf(-448.f, fminf(__uint_as_float(o[4 * k + 2 * h2]) * inv * qinv, 448.f));
          const float b = fmaxf(-448.f, fminf(__uint_as_float(o[4 * k + 2 * h2 + 1]) * inv * qinv, 448.f));
          b4 |= (uint32_t)f32x2_to_e4m3x2(a, b) << (16 * h2);
        }
        w[k] = b4;
      }
      *reinterpret_cast<uint4*>(reinterpret_cast<uint8_t*>(p.out) + obase + 16 * c) = make_uint4(w[0], w[1], w[2], w[3]);
    } else {
      uint32_t w[8];
#pragma unroll
      for (int k = 0; k < 8; k++)
        w[k] = pack2<OBF16>(__uint_as_float(o[2 * k]) * inv, __uint_as_float(o[2 * k + 1]) * inv);
      uint4* dst = reinterpret_cast<uint4*>(reinterpret_cast<uint16_t*>(p.out) + obase + 16 * c);
      dst[0] = make_uint4(w[0], w[1], w[2], w[3]);
      dst[1] = make_uint4(w[4], w[5], w[6], w[7]);
    }
  }
  }

// GENERAL = soft-cap and / or ALiBi present (scores are transformed element by element); the plain kernel folds
// scale * log2(e) into the exponent's FFMA. Both keep the per-tile softmax code small (< ~1000 SASS instructions):
// the first version spent 38 k cycles per tile on a 3000-instruction unrolled loop, instruction-fetch bound.
template <bool BF16, bool GENERAL>
__global__ void __launch_bounds__(TC_THREADS, 1)
    varlen_attn_tc_kernel(const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                          const Params p) {
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int kvh = blockIdx.y;
  const int BQ = p.block_q, g = p.g;

  // ---- which sequence and q block (resolve_seq_and_query_len, triton_unified_attention.py:311-318)
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
  if (qb * BQ >= q_len) return;            // uniform over the CTA: nothing allocated yet
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
  const uint32_t q_s = base;                       // Q tile, K-major
  const uint32_t p_s = base + TC_TILE;             // P tile, K-major
  const uint32_t kv_s = base + 2 * TC_TILE;        // [stage][K tile | V tile]
  const uint32_t bars = base + 6 * TC_TILE;
  const uint32_t bar_k_full = bars, bar_v_full = bars + 16, bar_kv_empty = bars + 32;
  const uint32_t bar_s_full = bars + 48, bar_s_free = bars + 56, bar_p_full = bars + 64, bar_pv_done = bars + 72;
  const uint32_t bar_v_raw = bars + 80;
  volatile uint32_t* tmem_holder = reinterpret_cast<volatile uint32_t*>(smem + 6 * TC_TILE + 96);

  if (tid == 0) {
    for (int i = 0; i < 2; i++) {
      mbar_init(bar_k_full + 8u * i, 1);
      mbar_init(bar_v_full + 8u * i, 1);
      mbar_init(bar_kv_empty + 8u * i, 1);
    }
    mbar_init(bar_s_full, 1);
    mbar_init(bar_s_free, 128);
    mbar_init(bar_p_full, 128);
    mbar_init(bar_pv_done, 1);
    mbar_init(bar_v_raw, 1);
    mbar_fence_init();
  }
  if (warp == 5) {
    const uint32_t holder = smem_u32((const void*)tmem_holder);
    tmem_alloc(holder, 256);
  }

  // ---- row of this softmax thread: row m = (query token m / g, query head m % g of the group)
  const int row = tid;   // meaningful for tid < 128
  const int tl = row / g;
  const int qpos = first_q + tl;
  const int head = kvh * g + (row - tl * g);
  const bool valid = tid < 128 && tl < BQ && qpos < q_len;
  const int limit = ctx + qpos;
  if (tid < 128) {
    // Q row -> smem, K-major SWIZZLE_128B (zero rows for padding)
    const uint16_t* qp = reinterpret_cast<const uint16_t*>(p.q) + (int64_t)(q_start + qpos) * p.q_stride_t +
                         (int64_t)head * p.q_stride_h;
#pragma unroll
    for (int c = 0; c < 16; c++) {
      uint4 v = make_uint4(0, 0, 0, 0);
      if (valid) v = *reinterpret_cast<const uint4*>(qp + 8 * c);
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
    // ================================================================ TMA producer
    if (lane == 0) {
      prefetch_tmap(&tmK);
      prefetch_tmap(&tmV);
    }
    const uint64_t pol = l2_policy_evict_last();   // other q blocks / heads of the group re-read these pages
    const int32_t* bt = p.block_table + (int64_t)s * p.bt_stride;
    const int sub = lane & 7, seg = (lane >> 3) & 1, is_v = lane >> 4;
    for (int i = 0; i < ntiles; i++) {
      const int stage = i & 1;
      const int t0 = (j_lo + i) * TC_BN;
      const int nsub = min(8, (kv_hi - t0 + 15) >> 4);
      const bool ragged = kv_hi < t0 + TC_BN;     // only the last tile can be
      mbar_wait(bar_kv_empty + 8u * stage, ((uint32_t)(i >> 1) & 1u) ^ 1u);
      const uint32_t st_base = kv_s + (uint32_t)stage * 2u * TC_TILE;
      const uint32_t vbar = ragged ? bar_v_raw : bar_v_full + 8u * stage;
      if (lane == 0) {
        mbar_arrive_expect_tx(bar_k_full + 8u * stage, (uint32_t)nsub * 2u * 2048u);
        mbar_arrive_expect_tx(vbar, (uint32_t)nsub * 2u * 2048u);
      }
      __syncwarp();
      if (sub < nsub) {
        const int token = t0 + 16 * sub;
        const int lb = token / p.bs;
        const int blk = bt[lb];
        const uint32_t dst = st_base + (uint32_t)is_v * TC_TILE + (uint32_t)seg * TC_SEG + (uint32_t)sub * 2048u;
        tma_load_5d(dst, is_v ? &tmV : &tmK, is_v ? vbar : bar_k_full + 8u * stage, 0, token - lb * p.bs, seg, kvh,
                    blk, pol);
      }
      if (ragged) {
        // V rows past the end of the sequence would meet P = 0 as 0 x garbage (NaN patterns in unwritten
        // cache slots): zero them once the pages have landed, then publish the tile
        mbar_wait(bar_v_raw, 0);
        const int r0 = kv_hi - t0;
        const uint32_t vb = st_base + TC_TILE;
        for (int idx = lane; idx < (TC_BN - r0) * 16; idx += 32) {
          const int rr = r0 + (idx >> 4), rem = idx & 15;
          sts128(vb + (uint32_t)(rem >> 3) * TC_SEG + (uint32_t)rr * 128u + (uint32_t)(rem & 7) * 16u,
                 make_uint4(0, 0, 0, 0));
        }
        fence_proxy_async();
        __syncwarp();
        if (lane == 0) mbar_arrive(bar_v_full + 8u * stage);
      }
    }
  } else if (warp == 5) {
    // ================================================================ MMA issuer: warp-uniform control flow, one
    // elected lane issues (common.cuh elect_one: under `if (lane == 0)` ptxas wraps each of the 16 UTCHMMA of a
    // tile in an ELECT loop, ~65 cycles apiece - as much as the tensor work of the tile)
    {
      constexpr uint32_t FMT = BF16 ? 1u : 0u;
      constexpr uint32_t IDESC_S = (1u << 4) | (FMT << 7) | (FMT << 10) | ((uint32_t)(TC_BN >> 3) << 17) |
                                   ((uint32_t)(TC_BM >> 4) << 24);
      constexpr uint32_t IDESC_O = (1u << 4) | (FMT << 7) | (FMT << 10) | (1u << 16) | ((uint32_t)(TC_D >> 3) << 17) |
                                   ((uint32_t)(TC_BM >> 4) << 24);
      auto issue_s = [&](int i) {
        const uint32_t kb = kv_s + (uint32_t)(i & 1) * 2u * TC_TILE;
        mbar_wait(bar_k_full + 8u * (i & 1), (uint32_t)(i >> 1) & 1u);
        tc_fence_after();
        if (elect_one()) {
#pragma unroll
          for (int ks = 0; ks < TC_D / 16; ks++) {
            const uint32_t off = (uint32_t)(ks >> 2) * TC_SEG + (uint32_t)(ks & 3) * 32u;
            tc_mma_f16(tm_S, desc_kmajor(q_s + off), desc_kmajor(kb + off), IDESC_S, ks > 0 ? 1u : 0u);
          }
          tc_commit(bar_s_full);
        }
        __syncwarp();
      };
      issue_s(0);
      for (int i = 0; i < ntiles; i++) {
        const int stage = i & 1;
        if (i + 1 < ntiles) {
          mbar_wait(bar_s_free, (uint32_t)i & 1u);   // the softmax warps hold S(i) in registers
          issue_s(i + 1);
        }
        mbar_wait(bar_p_full, (uint32_t)i & 1u);
        mbar_wait(bar_v_full + 8u * stage, (uint32_t)(i >> 1) & 1u);
        tc_fence_after();
        if (elect_one()) {
          const uint32_t vb = kv_s + (uint32_t)stage * 2u * TC_TILE + TC_TILE;
#pragma unroll
          for (int ks = 0; ks < TC_BN / 16; ks++) {
            const uint32_t poff = (uint32_t)(ks >> 2) * TC_SEG + (uint32_t)(ks & 3) * 32u;
            tc_mma_f16(tm_O, desc_kmajor(p_s + poff), desc_mnmajor(vb + (uint32_t)ks * 2048u, TC_SEG, 1024u), IDESC_O,
                       (i > 0 || ks > 0) ? 1u : 0u);
          }
          tc_commit(bar_kv_empty + 8u * stage);
          tc_commit(bar_pv_done);
        }
        __syncwarp();
      }
    }
  } else {
    // ================================================================ softmax / correction / epilogue
    SoftmaxCtx c;
    c.tm_S = tm_S; c.tm_O = tm_O; c.p_s = p_s;
    c.bar_s_full = bar_s_full; c.bar_s_free = bar_s_free; c.bar_p_full = bar_p_full; c.bar_pv_done = bar_pv_done;
    c.ntiles = ntiles; c.j_lo = j_lo; c.ctx = ctx; c.first_q = first_q; c.last_q = last_q; c.kv_hi = kv_hi;
    c.limit = limit; c.row = row; c.warp = warp; c.head = head; c.q_start = q_start; c.qpos = qpos;
    c.valid = valid; c.qk_scale = p.scale; c.v_scale = 1.f;
    softmax_warps<BF16, BF16, GENERAL>(p, c);
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
