// b200ops -- FP8 (E4M3) W8A8 GEMM, third generation: persistent CTA PAIRS (cta_group::2), stream-K
// work split, double-buffered TMEM accumulators (sm_100a).
//
// Contract: identical to gemm_fp8.cu (torch.ops._C.cutlass_scaled_mm,
// csrc/libtorch_stable/quantization/w8a8/cutlass/scaled_mm_entry.cu:197-270):
//   out[M,N] = a_scale (.) ( b_scale (.) (A[M,K] . B[K,N]) ) (+ bias[N]),  fp32 accumulate.
//
// Why a third kernel (measured on B200, profiles/r1b_ncu_full.md): at decode M = 256 the one-CTA kernel is bound
// by what one SM can pull in per k-block (16 KB of weights + 32 KB of activations at ~100 GB/s per SM),
// loses ~25 % to wave quantisation (224 weight slabs on 148 SMs) and pays prologue + epilogue once per
// tile. Here:
//   * a CTA pair issues ONE tcgen05.mma.cta_group::2 of 256 x BN x 32: each CTA stages its own 128
//     "rows" and only HALF of the "cols" operand, so per-SM ingest per k-block drops from 48 to 32 KB;
//   * the grid is persistent (one pair per 2 SMs) and the flattened (tile, k-block) space is cut into
//     equal contiguous ranges (stream-K): every pair gets the same number of k-blocks whatever N is;
//     a tile cut by a range boundary is finished by the pair that holds its FIRST k-blocks (it gets
//     there last), the other pairs publish fp32 partials (L2-resident workspace + flag) first thing;
//   * two TMEM accumulator buffers (2 x BN columns) let the epilogue of piece i overlap the MMAs of
//     piece i+1.
// Roles per CTA: warp 0 = TMA producer, warp 1 = MMA issuer (leader CTA only) + TMEM owner,
// warps 2-5 = epilogue.
#include "tcgen05.cuh"

#include <cmath>
#include <cstring>
#include <type_traits>

namespace b200 {
namespace gemm8v3 {

#ifdef B200_V3_TRACE
// development-only timeline: [pair][role 0 = MMA thread, 1 = epilogue thread 0 of the leader CTA][piece < 8][4]
static unsigned long long* g_trace = nullptr;
#define V3_TRACE(role, piece, slot)                                                            \
  if (p.trace && (piece) < 8) p.trace[(((size_t)pair * 2 + (role)) * 8 + (piece)) * 4 + (slot)] = clock64()
#else
#define V3_TRACE(role, piece, slot)
#endif

constexpr int BM = 128;          // rows per CTA (UMMA M = 256 per pair)
constexpr int BK = 128;          // k-block bytes
constexpr int UMMA_K = 32;
constexpr int NTHREADS = 192;
constexpr int MAX_STAGES = 8;
constexpr uint32_t PEER_MASK = 0xFEFFFFFFu;  // clears the CTA-rank bit of a shared::cluster address

struct Params {
  void* out;
  const float* r_scale;
  const float* c_scale;
  const void* bias;
  float* ws_partials;     // [pairs][2 CTAs][BN/16][128][16] fp32
  uint32_t* ws_flags;     // [pairs][2]: arrive count, consumer-done count
  int R, C, K;
  int64_t ldo;
  int r_scale_vec, c_scale_vec;
  int swap, out_bf16, stages, tma_store;
  int prefetch_w;         // PDL level 2 (decode, swap): weight tiles of the first ring pass precede griddepcontrol.wait
  int row_tiles, col_tiles, num_kb;   // tiles of 256 rows x BN cols, k-blocks per tile
  int pairs;
  // Work split. interleave = 0: stream-K, pair i takes the i-th equal contiguous range of the flattened
  // (tile, k-block) space (decode: fewer tiles than pairs). interleave = 1: whole tiles, pair i takes tiles
  // i, i + pairs, ... of a grouped raster (large M): the `pairs` tiles in flight at any moment form a compact
  // raster_group x (pairs / raster_group) block, so their operands stay L2-resident (contiguous ranges put
  // concurrent pairs ~14 tiles apart = every column tile of B live at once).
  int interleave, raster_group;
  int staging_bytes;      // transposed-output staging tile (swap + TMA store only)
  unsigned long long* trace;   // B200_V3_TRACE builds only
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
// arrive on the barrier at this smem offset in BOTH CTAs of the pair when the prior MMAs complete
__device__ __forceinline__ void tc_commit_2sm(uint32_t bar) {
  asm volatile(
      "tcgen05.commit.cta_group::2.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 [%0], %1;" ::"r"(bar),
      "h"((uint16_t)3)
      : "memory");
}
// arrive on the LEADER CTA's copy of a barrier
__device__ __forceinline__ void mbar_arrive_leader(uint32_t bar) {
  asm volatile(
      "{\n\t.reg .b32 ra;\n\t"
      "mapa.shared::cluster.u32 ra, %0, 0;\n\t"
      "mbarrier.arrive.release.cluster.shared::cluster.b64 _, [ra];\n\t}" ::"r"(bar)
      : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_gpu(const uint32_t* p) {
  uint32_t v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}

// One contiguous piece of one tile handled by this pair.
struct Piece {
  int tile, kb0, kb1;   // k-blocks [kb0, kb1) of `tile`
};
// The unit range of pair p in the flattened (tile, k-block) space.
__device__ __forceinline__ void pair_range(const Params& p, int pair, int64_t* lo, int64_t* hi) {
  const int64_t units = (int64_t)p.row_tiles * p.col_tiles * p.num_kb;
  *lo = units * pair / p.pairs;
  *hi = units * (pair + 1) / p.pairs;
}
struct Sched {
  int64_t cur, hi;
};
__device__ __forceinline__ void sched_init(const Params& p, int pair, Sched* s) {
  if (p.interleave) {
    s->cur = pair;
    s->hi = (int64_t)p.row_tiles * p.col_tiles;
  } else {
    pair_range(p, pair, &s->cur, &s->hi);
  }
}
// next piece in processing order; returns false past the end
__device__ __forceinline__ bool next_piece(const Params& p, Sched& s, Piece* pc) {
  if (s.cur >= s.hi) return false;
  if (p.interleave) {
    pc->tile = (int)s.cur;
    pc->kb0 = 0;
    pc->kb1 = p.num_kb;
    s.cur += p.pairs;
    return true;
  }
  const int tile = (int)(s.cur / p.num_kb);
  const int kb0 = (int)(s.cur - (int64_t)tile * p.num_kb);
  const int64_t tile_end = (int64_t)(tile + 1) * p.num_kb;
  const int64_t end = tile_end < s.hi ? tile_end : s.hi;
  pc->tile = tile;
  pc->kb0 = kb0;
  pc->kb1 = kb0 + (int)(end - s.cur);
  s.cur = end;
  return true;
}
__device__ __forceinline__ void tile_coords(const Params& p, int tile, int* rt, int* ct) {
  if (p.raster_group > 0) {
    const int per_group = p.raster_group * p.col_tiles;
    const int grp = tile / per_group, within = tile - grp * per_group;
    const int rows_here = min(p.raster_group, p.row_tiles - grp * p.raster_group);
    *rt = grp * p.raster_group + within % rows_here;
    *ct = within / rows_here;
  } else {
    *rt = tile % p.row_tiles;
    *ct = tile / p.row_tiles;
  }
}
// 1-D bulk copy global -> this CTA's shared memory, completion on an mbarrier
__device__ __forceinline__ void bulk_g2s(uint32_t dst, const void* src, uint32_t bytes, uint32_t bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(dst),
               "l"(src), "r"(bytes), "r"(bar)
               : "memory");
}

template <int BN>
__global__ void __cluster_dims__(2, 1, 1) __launch_bounds__(NTHREADS, 1)
    gemm_fp8_v3_kernel(const __grid_constant__ CUtensorMap tmR, const __grid_constant__ CUtensorMap tmC,
                       const __grid_constant__ CUtensorMap tmO, const Params p) {
  constexpr uint32_t R_BYTES = BM * BK;             // this CTA's 128 rows
  constexpr uint32_t C_BYTES = (BN / 2) * BK;       // this CTA's half of the cols operand
  constexpr uint32_t STAGE_BYTES = R_BYTES + C_BYTES;
  constexpr uint32_t CHUNK_BYTES = BM * 16 * 4;      // one 16-column fp32 chunk of this CTA's 128 rows
  constexpr uint32_t FIX_BYTES = 2 * CHUNK_BYTES;    // fix-up buffer: 2 chunks (one drain iteration) of one contributor
  constexpr uint32_t TMEM_COLS = 2 * BN < 32 ? 32 : 2 * BN;   // two accumulator buffers
  // instruction descriptor: D = F32, A = B = E4M3, K-major, N = BN, M = 256 (pair)
  constexpr uint32_t IDESC = (1u << 4) | ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(256 >> 4) << 24);

  extern __shared__ uint8_t smem_raw[];
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t base = (raw_base + 1023u) & ~1023u;
  uint8_t* smem = smem_raw + (base - raw_base);
  const int stages = p.stages;
  const uint32_t staging = base + (uint32_t)stages * STAGE_BYTES;   // [BN m][128 r] 16-bit (swap + TMA store)
  const uint32_t bar_base = staging + (uint32_t)p.staging_bytes;
  const uint32_t bar_full = bar_base;                          // [stages]   (leader's copy is used)
  const uint32_t bar_empty = bar_base + 8u * MAX_STAGES;       // [stages]   per CTA
  const uint32_t bar_tfull = bar_base + 16u * MAX_STAGES;      // [2]        per CTA
  const uint32_t bar_tempty = bar_tfull + 16u;                 // [2]        (leader's copy is used)
  const uint32_t bar_fix = bar_tempty + 16u;                   // [MAX_STAGES] fix-up bulk copies, per CTA
  volatile uint32_t* tmem_holder = reinterpret_cast<volatile uint32_t*>(
      smem + (size_t)stages * STAGE_BYTES + p.staging_bytes + 24 * MAX_STAGES + 32 + 8);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const uint32_t cta = cluster_ctarank();
  const int pair = blockIdx.x >> 1;
  pdl_launch_dependents();

  if (threadIdx.x == 0) {
    for (int i = 0; i < stages; i++) {
      mbar_init(bar_full + 8u * i, 1);
      mbar_init(bar_empty + 8u * i, 1);
    }
    for (int b = 0; b < 2; b++) {
      mbar_init(bar_tfull + 8u * b, 1);
      mbar_init(bar_tempty + 8u * b, 8);   // 4 epilogue warps x 2 CTAs
    }
    for (int i = 0; i < MAX_STAGES; i++) mbar_init(bar_fix + 8u * i, 1);
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
  cluster_sync();   // both CTAs' barriers are initialised before any remote arrive / multicast commit
  tc_fence_after();
  const uint32_t tmem_base = *tmem_holder;

  Sched sched;
  sched_init(p, pair, &sched);

  if (warp == 0) {
    // ============================== TMA producer (warp-uniform loops, one elected lane issues: common.cuh elect_one)
    prefetch_tmap(&tmR);
    prefetch_tmap(&tmC);
    const uint64_t pol_w = l2_policy_evict_first();
    const uint64_t pol_a = l2_policy_evict_last();
    // decode: weights stream once (evict first), the batch is re-read by every pair (evict last);
    // large M: both operands are re-read by other tiles of the raster block
    const uint64_t pol_r = p.swap ? pol_w : pol_a;
    const uint64_t pol_c = pol_a;
    int stage = 0;
    uint32_t phase = 0;
    Piece pc;
    // Programmatic dependent launch, level 2 (decode): the weight tiles (rows operand) of the first ring pass do
    // not depend on the previous kernel; they are requested before griddepcontrol.wait from a copy of the
    // scheduler state, the activation tiles of the same k-blocks follow after the wait.
    int pre = 0;
    if (p.prefetch_w && p.swap) {
      Sched s1 = sched;
      while (pre < stages && next_piece(p, s1, &pc)) {
        int rt, ct;
        tile_coords(p, pc.tile, &rt, &ct);
        const int row0 = rt * 2 * BM + (int)cta * BM;
        for (int kb = pc.kb0; kb < pc.kb1 && pre < stages; kb++, pre++) {
          if (elect_one()) {
            if (cta == 0) mbar_arrive_expect_tx(bar_full + 8u * pre, 2 * STAGE_BYTES);
            tma_load_2d_2sm(base + (uint32_t)pre * STAGE_BYTES, &tmR, bar_full + 8u * pre, kb * BK, row0, pol_r);
          }
          __syncwarp();
        }
      }
    }
    pdl_wait();
    int n = 0;
    while (next_piece(p, sched, &pc)) {
      int rt, ct;
      tile_coords(p, pc.tile, &rt, &ct);
      const int row0 = rt * 2 * BM + (int)cta * BM;
      const int col0 = ct * BN + (int)cta * (BN / 2);
      for (int kb = pc.kb0; kb < pc.kb1; kb++, n++) {
        const bool w_in_flight = n < pre;   // first ring pass: slot free, weights already requested
        if (!w_in_flight) mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
        if (elect_one()) {
          const uint32_t dst = base + (uint32_t)stage * STAGE_BYTES;
          // the leader's barrier counts the bytes of BOTH CTAs (the peer's complete_tx may land first:
          // the tx-count simply goes negative until the leader's expect_tx)
          if (!w_in_flight) {
            if (cta == 0) mbar_arrive_expect_tx(bar_full + 8u * stage, 2 * STAGE_BYTES);
            tma_load_2d_2sm(dst, &tmR, bar_full + 8u * stage, kb * BK, row0, pol_r);
          }
          tma_load_2d_2sm(dst + R_BYTES, &tmC, bar_full + 8u * stage, kb * BK, col0, pol_c);
        }
        __syncwarp();
        if (++stage == stages) { stage = 0; phase ^= 1u; }
      }
    }
  } else if (warp == 1) {
    // ============================== MMA issuer: one elected lane of the LEADER CTA drives both tensor cores
    pdl_wait();   // nothing to issue before the previous kernel's output arrives: block instead of polling
    if (cta == 0) {
      int stage = 0;
      uint32_t phase = 0;
      uint32_t tphase[2] = {0, 0};
      Piece pc;
      int n = 0;
      while (next_piece(p, sched, &pc)) {
        const int buf = n & 1;
        if (lane == 0) V3_TRACE(0, n, 0);
        mbar_wait(bar_tempty + 8u * buf, tphase[buf] ^ 1u);   // epilogues of both CTAs drained this buffer
        if (lane == 0) V3_TRACE(0, n, 1);
        tphase[buf] ^= 1u;
        tc_fence_after();
        const uint32_t tmem_d = tmem_base + (uint32_t)(buf * BN);
        for (int kb = pc.kb0; kb < pc.kb1; kb++) {
          mbar_wait(bar_full + 8u * stage, phase);
          tc_fence_after();
          if (elect_one()) {
            const uint32_t sa = base + (uint32_t)stage * STAGE_BYTES;
            const uint64_t da = make_smem_desc(sa);
            const uint64_t db = make_smem_desc(sa + R_BYTES);
#pragma unroll
            for (int k = 0; k < BK / UMMA_K; k++)
              tc_mma_f8_2sm(tmem_d, da + (uint64_t)(k * UMMA_K >> 4), db + (uint64_t)(k * UMMA_K >> 4), IDESC,
                            (kb > pc.kb0 || k > 0) ? 1u : 0u);
            tc_commit_2sm(bar_empty + 8u * stage);
            if (kb == pc.kb1 - 1) tc_commit_2sm(bar_tfull + 8u * buf);
          }
          __syncwarp();
          if (++stage == stages) { stage = 0; phase ^= 1u; }
        }
        if (lane == 0) V3_TRACE(0, n, 2);
        n++;
      }
    }
  } else {
    // ============================== epilogue warps (4 per CTA, one TMEM lane quarter each)
    const int wq = warp & 3;
    const int et = (warp - 2) * 32 + lane;   // 0..127 among the epilogue threads
    uint32_t tphase[2] = {0, 0};
    Piece pc;
    int n = 0;
    pdl_wait();   // scales / workspace flags / `out` are shared with the previous kernel on the stream
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
    const float cs0 = p.c_scale_vec ? 0.f : p.c_scale[0];
    const int trow = wq * 32 + lane;        // row of this thread inside the CTA's 128
    // fix-up buffers live in the pipeline smem, idle once the LAST piece's accumulator is complete
    const int nbuf = min(MAX_STAGES, (int)((uint32_t)stages * STAGE_BYTES / FIX_BYTES));
    bool store_pending = false;
    while (next_piece(p, sched, &pc)) {
      const int buf = n & 1;
      int rt, ct;
      tile_coords(p, pc.tile, &rt, &ct);
      const int row0 = rt * 2 * BM + (int)cta * BM;   // this CTA's first row
      const int col0 = ct * BN;
      if (cta == 0 && et == 0) { V3_TRACE(1, n, 0); }
      mbar_wait(bar_tfull + 8u * buf, tphase[buf]);
      if (cta == 0 && et == 0) { V3_TRACE(1, n, 1); }
      tphase[buf] ^= 1u;
      tc_fence_after();
      const uint32_t tbase = tmem_base + (uint32_t)(buf * BN) + ((uint32_t)(wq * 32) << 16);
      const int ncols = min(BN, p.C - col0);
      const int nchunks = (ncols + 15) >> 4;
      const int nbatch = (nchunks + 3) >> 2;
      if (pc.kb0 > 0) {
        // ---------- not the owner of this tile: publish the fp32 partial, then raise the flag.
        // Layout per CTA slot: [chunk][j = 0..3][row 0..127][4 floats]: a chunk is 8 KB contiguous (one bulk
        // copy on the owner's side) and row-consecutive lanes are 16 bytes apart on both sides.
        float4* slot = reinterpret_cast<float4*>(p.ws_partials + ((size_t)(pair * 2 + cta) * (BN / 16) * BM) * 16);
        for (int b = 0; b < nbatch; b++) {
          uint32_t v[4][16];
#pragma unroll
          for (int c = 0; c < 4; c++)
            if (4 * b + c < nchunks) tc_ld16(tbase + (uint32_t)((4 * b + c) << 4), v[c]);
          tc_wait_ld();
#pragma unroll
          for (int c = 0; c < 4; c++) {
            if (4 * b + c < nchunks) {
              float4* dst = slot + (size_t)(4 * b + c) * 4 * BM + trow;
#pragma unroll
              for (int j = 0; j < 4; j++)
                dst[j * BM] = make_float4(__uint_as_float(v[c][4 * j]), __uint_as_float(v[c][4 * j + 1]),
                                          __uint_as_float(v[c][4 * j + 2]), __uint_as_float(v[c][4 * j + 3]));
            }
          }
        }
        tc_fence_before();
        __syncwarp();
        if (lane == 0) mbar_arrive_leader(bar_tempty + 8u * buf);   // TMEM buffer free for the next piece
        if (cta == 0 && et == 0) { V3_TRACE(1, n, 2); }
        __threadfence();
        asm volatile("bar.sync 1, 128;" ::: "memory");
        if (et == 0) atomicAdd(p.ws_flags + 2 * pair, 1u);
      } else {
        // ---------- owner: accumulator (+ partials of the pairs that hold the rest of this tile)
        const int r = row0 + trow;
        const bool r_ok = r < p.R;
        const float rs = p.r_scale_vec ? (r_ok ? p.r_scale[r] : 0.f) : p.r_scale[0];
        float bias_r = 0.f;
        if (p.swap && bias16 && r_ok) bias_r = to_f32(bias16[r]);
        // pairs q > pair whose range starts inside this tile hold its remaining k-blocks
        int q_end = pair + 1;
        if (pc.kb1 < p.num_kb) {
          const int64_t tile_end = (int64_t)(pc.tile + 1) * p.num_kb;
          while (q_end < p.pairs) {
            int64_t qlo, qhi;
            pair_range(p, q_end, &qlo, &qhi);
            if (qlo >= tile_end) break;
            q_end++;
          }
        }
        const int qn = q_end - pair - 1;        // contributors (their partials are added in pair order)
        const int total = ((nchunks + 1) >> 1) * qn;   // fix-up item i = (chunk pair i / qn, contributor i % qn)
        auto issue = [&](int i) {               // et == 0 only
          const int b = i / qn, q = pair + 1 + i % qn;
          const int nch = min(2, nchunks - 2 * b);
          const uint32_t bytes = (uint32_t)nch * CHUNK_BYTES;
          const float* src = p.ws_partials + ((size_t)(q * 2 + cta) * (BN / 16) + 2 * b) * (BM * 16);
          const uint32_t bar = bar_fix + 8u * (uint32_t)(i % nbuf);
          mbar_arrive_expect_tx(bar, bytes);
          bulk_g2s(base + (uint32_t)(i % nbuf) * FIX_BYTES, src, bytes, bar);
        };
        if (qn > 0 && et == 0) {
          // the contributors published first thing in their ranges; this pair got here last
          for (int q = pair + 1; q < q_end; q++) {
            const uint32_t* f = p.ws_flags + 2 * q;
            while (ld_acquire_gpu(f) < 2u) {
            }
          }
          asm volatile("fence.proxy.async;" ::: "memory");   // acquired generic writes -> async-proxy reads
          for (int i = 0; i < total && i < nbuf; i++) issue(i);
        }
        if (p.swap && p.tma_store && store_pending) {
          // the previous tile's TMA store must have read the staging tile
          if (et == 0) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
          asm volatile("bar.sync 1, 128;" ::: "memory");
        }
        // One chunk = 16 accumulator columns. Two chunks per iteration in a ROLLED loop, one store flavour per
        // instantiation: the unrolled 4-chunk / 3-flavour version was ~8000 SASS instructions of straight-line
        // code and ran at instruction-fetch speed (30k cycles per tile vs 16k of MMAs; round-1 timeline).
        auto pack2 = [&](float lo, float hi) -> uint32_t {
          if (p.out_bf16) {
            __nv_bfloat162 t = __floats2bfloat162_rn(lo, hi);
            return *reinterpret_cast<uint32_t*>(&t);
          }
          __half2 t = __floats2half2_rn(lo, hi);
          return *reinterpret_cast<uint32_t*>(&t);
        };
        auto emit_body = [&](auto mode_tag, auto full_tag, float (&a)[16], int ch) {
          constexpr int MODE = decltype(mode_tag)::value;   // 0 = out[r][c] direct, 1 = swap via staging, 2 = swap direct
          constexpr bool full = decltype(full_tag)::value;  // all 16 columns of the chunk exist: no per-element guards
          const int c0 = ch << 4;
          const int cbase = col0 + c0;
          // column factors of this chunk (per-column scale vector or the scalar)
          float cs[16];
          if (p.c_scale_vec) {
#pragma unroll
            for (int j = 0; j < 16; j++) cs[j] = (full || cbase + j < p.C) ? __ldg(p.c_scale + cbase + j) : 0.f;
          } else {
#pragma unroll
            for (int j = 0; j < 16; j++) cs[j] = cs0;
          }
          if constexpr (MODE == 0) {
            // out[r][c] = a_s[r] * (b_s[c] * acc) (+ bias[c])
#pragma unroll
            for (int j = 0; j < 16; j++) a[j] = rs * (cs[j] * a[j]);
            if (bias16) {
#pragma unroll
              for (int j = 0; j < 16; j++)
                if (full || cbase + j < p.C) a[j] += to_f32(bias16[cbase + j]);
            }
            if (r_ok) {
              uint32_t h[8];
#pragma unroll
              for (int j = 0; j < 8; j++) h[j] = pack2(a[2 * j], a[2 * j + 1]);
              uint16_t* dst = out16 + (int64_t)r * p.ldo + cbase;
              if (full && ((reinterpret_cast<uintptr_t>(dst) & 15) == 0)) {
                reinterpret_cast<uint4*>(dst)[0] = make_uint4(h[0], h[1], h[2], h[3]);
                reinterpret_cast<uint4*>(dst)[1] = make_uint4(h[4], h[5], h[6], h[7]);
              } else {
#pragma unroll
                for (int j = 0; j < 16; j++)
                  if (cbase + j < p.C) dst[j] = (uint16_t)(h[j >> 1] >> ((j & 1) * 16));
              }
            }
          } else {
            // out[c][r] = a_s[c] * (b_s[r] * acc) + bias[r]
#pragma unroll
            for (int j = 0; j < 16; j++) a[j] = cs[j] * (rs * a[j]) + bias_r;
            if constexpr (MODE == 1) {
#pragma unroll
              for (int j = 0; j < 16; j += 2) {
                const float oa = a[j], ob = a[j + 1];
                const float send = (lane & 1) ? oa : ob;
                const float recv = __shfl_xor_sync(0xffffffffu, send, 1);
                const uint32_t packed = pack2((lane & 1) ? recv : oa, (lane & 1) ? ob : recv);
                const int m_local = c0 + j + (lane & 1);
                const uint32_t addr = staging + (uint32_t)m_local * (BM * 2) + (uint32_t)(wq * 32 + (lane & ~1)) * 2u;
                asm volatile("st.shared.b32 [%0], %1;" ::"r"(addr), "r"(packed) : "memory");
              }
            } else if (r_ok) {
#pragma unroll
              for (int j = 0; j < 16; j++)
                if (full || cbase + j < p.C) out16[(int64_t)(cbase + j) * p.ldo + r] = from_f32(a[j]);
            }
          }
        };
        auto emit = [&](auto mode_tag, float (&a)[16], int ch) {
          if (col0 + (ch << 4) + 16 <= p.C) emit_body(mode_tag, std::true_type{}, a, ch);
          else emit_body(mode_tag, std::false_type{}, a, ch);
        };
        auto drain = [&](auto mode_tag) {
#pragma unroll 1
          for (int ch = 0; ch < nchunks; ch += 2) {
            const bool two = ch + 1 < nchunks;
            float a0[16], a1[16];
            {
              uint32_t v0[16], v1[16];
              tc_ld16(tbase + (uint32_t)(ch << 4), v0);
              if (two) tc_ld16(tbase + (uint32_t)((ch + 1) << 4), v1);
              tc_wait_ld();
#pragma unroll
              for (int j = 0; j < 16; j++) {
                a0[j] = __uint_as_float(v0[j]);
                a1[j] = two ? __uint_as_float(v1[j]) : 0.f;
              }
            }
            if (ch + 2 >= nchunks) {   // all TMEM reads of this buffer are done
              tc_fence_before();
              __syncwarp();
              if (lane == 0) mbar_arrive_leader(bar_tempty + 8u * buf);
              if (cta == 0 && et == 0) { V3_TRACE(1, n, 2); }
            }
            if (qn > 0) {
#pragma unroll 1
              for (int qi = 0; qi < qn; qi++) {
                const int i = (ch >> 1) * qn + qi;
                mbar_wait(bar_fix + 8u * (uint32_t)(i % nbuf), (uint32_t)(i / nbuf) & 1u);
                const uint32_t fb = base + (uint32_t)(i % nbuf) * FIX_BYTES + (uint32_t)trow * 16u;
#pragma unroll
                for (int j = 0; j < 4; j++) {
                  float4 t;
                  asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                               : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w)
                               : "r"(fb + (uint32_t)(j * BM) * 16u));
                  a0[4 * j] += t.x; a0[4 * j + 1] += t.y; a0[4 * j + 2] += t.z; a0[4 * j + 3] += t.w;
                }
                if (two) {
#pragma unroll
                  for (int j = 0; j < 4; j++) {
                    float4 t;
                    asm volatile("ld.shared.v4.f32 {%0,%1,%2,%3}, [%4];"
                                 : "=f"(t.x), "=f"(t.y), "=f"(t.z), "=f"(t.w)
                                 : "r"(fb + (uint32_t)((4 + j) * BM) * 16u));
                    a1[4 * j] += t.x; a1[4 * j + 1] += t.y; a1[4 * j + 2] += t.z; a1[4 * j + 3] += t.w;
                  }
                }
                if (i + nbuf < total) {   // recycle this buffer
                  asm volatile("bar.sync 1, 128;" ::: "memory");
                  if (et == 0) issue(i + nbuf);
                }
              }
            }
            emit(mode_tag, a0, ch);
            if (two) emit(mode_tag, a1, ch + 1);
          }
        };
        if (!p.swap) drain(std::integral_constant<int, 0>{});
        else if (p.tma_store) drain(std::integral_constant<int, 1>{});
        else drain(std::integral_constant<int, 2>{});
        if (p.swap && p.tma_store) {
          fence_proxy_async();
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (et == 0) {
            asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.bulk_group [%0, {%1, %2}], [%3];" ::"l"(&tmO),
                         "r"(row0), "r"(col0), "r"(staging)
                         : "memory");
            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
          }
          store_pending = true;
        }
        // release the partial slots: the second of the pair's two CTAs to finish resets the flags
        if (qn > 0) {
          asm volatile("bar.sync 1, 128;" ::: "memory");
          if (et == 0) {
            for (int q = pair + 1; q < q_end; q++) {
              if (atomicAdd(p.ws_flags + 2 * q + 1, 1u) == 1u) {
                p.ws_flags[2 * q + 1] = 0u;
                __threadfence();
                p.ws_flags[2 * q] = 0u;
              }
            }
          }
        }
      }
      if (cta == 0 && et == 0) { V3_TRACE(1, n, 3); }
      n++;
    }
    if (store_pending && et == 0) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  cluster_sync();   // the peer may still multicast-commit into / arrive on this CTA's barriers
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::2.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"(TMEM_COLS) : "memory");
  }
}

constexpr size_t WS_FLAG_BYTES = 4096;

template <int BN>
static int launch(const CUtensorMap& tmR, const CUtensorMap& tmC, const CUtensorMap& tmO, Params& p,
                  cudaStream_t stream) {
  constexpr uint32_t STAGE_BYTES = BM * BK + (BN / 2) * BK;
  const int staging = (p.swap && p.tma_store) ? BN * BM * 2 : 0;
  p.staging_bytes = staging;
  const int smem_max = max_smem_optin();
  int stages = (smem_max - 1024 - 256 - 128 - staging) / (int)STAGE_BYTES;
  if (stages > MAX_STAGES) stages = MAX_STAGES;
  if (stages < 2) {
    set_last_error("fp8_gemm_v3: not enough shared memory");
    return B200_ERR_UNSUPPORTED;
  }
  p.stages = stages;
  const int smem = stages * (int)STAGE_BYTES + staging + 24 * MAX_STAGES + 32 + 64 + 1024;
  static int max_pairs = 0;
  if (max_pairs == 0) {
    B200_CUDA(cudaFuncSetAttribute(gemm_fp8_v3_kernel<BN>, cudaFuncAttributeMaxDynamicSharedMemorySize, smem_max));
    // every pair must be co-resident (the tile owner spins on the other pairs' partials)
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(sm_count() / 2 * 2);
    cfg.blockDim = dim3(NTHREADS);
    cfg.dynamicSmemBytes = smem_max;
    int clusters = 0;
    cudaError_t e = cudaOccupancyMaxActiveClusters(&clusters, gemm_fp8_v3_kernel<BN>, &cfg);
    if (e != cudaSuccess || clusters <= 0) {
      cudaGetLastError();
      clusters = sm_count() / 2;
    }
    max_pairs = clusters < sm_count() / 2 ? clusters : sm_count() / 2;
  }
  p.pairs = max_pairs;
  const int64_t tiles = (int64_t)p.row_tiles * p.col_tiles;
  const int64_t units = tiles * p.num_kb;
  if (units < p.pairs) p.pairs = (int)units;
  // many tiles per pair: whole tiles in a grouped raster (L2 locality; the tail wave costs < 1 / (tiles / pairs));
  // few: stream-K so every pair streams the same number of k-blocks
  // (decode with more weight tiles than pairs, e.g. gate_up: whole tiles too -- the fp32 partial round trip of
  // a split tile costs more than the idle half wave: v3 stream-K 30.7-44.2 us vs 14-30 us for the others at the
  // qkv / o / down decode shapes, profiles/r2q_gemm_impls.jsonl)
  p.interleave = ((!p.swap && tiles >= 4 * (int64_t)p.pairs) || (p.swap && tiles >= (int64_t)p.pairs)) ? 1 : 0;
  p.raster_group = (p.interleave && !p.swap) ? 8 : 0;
  p.prefetch_w = pdl_level() >= 2;
  B200_CUDA(launch_pdl(gemm_fp8_v3_kernel<BN>, dim3(p.pairs * 2), dim3(NTHREADS), (size_t)smem, stream, tmR, tmC,
                       tmO, p));
  return B200_OK;
}

static int make_tmap_2d(CUtensorMap* tm, const void* base, int rows, int K, int64_t ld_bytes, int box_rows) {
  uint64_t dims[2] = {(uint64_t)K, (uint64_t)rows};
  uint64_t strides[1] = {(uint64_t)ld_bytes};
  uint32_t box[2] = {(uint32_t)BK, (uint32_t)box_rows};
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, base, dims, strides, box,
                           CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

}  // namespace gemm8v3
}  // namespace b200

#ifdef B200_V3_TRACE
extern "C" __attribute__((visibility("default"))) void b200_fp8_gemm_v3_set_trace(void* dev_ptr) {
  b200::gemm8v3::g_trace = reinterpret_cast<unsigned long long*>(dev_ptr);
}
#endif

extern "C" size_t b200_fp8_gemm_v3_workspace_bytes(void) {
  // flags + one fp32 partial tile (2 CTAs x 128 rows x 256 cols) per CTA pair
  const int pairs = b200::sm_count() > 0 ? b200::sm_count() / 2 : 74;
  return b200::gemm8v3::WS_FLAG_BYTES + (size_t)pairs * 2 * 128 * 256 * sizeof(float);
}

extern "C" int b200_fp8_gemm_v3(void* out, const void* a, const void* b, const float* a_scale,
                                const float* b_scale, const void* bias, int M, int N, int K, int64_t lda,
                                int64_t ldb, int64_t ldo, int a_scale_per_row, int b_scale_per_col,
                                int out_dtype, void* workspace, size_t workspace_bytes,
                                b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::gemm8v3;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(M >= 0 && N > 0 && K > 0, B200_ERR_INVALID, "cutlass_scaled_mm: bad sizes (%d,%d,%d)", M, N, K);
  if (M == 0) return B200_OK;
  B200_CHECK(out && a && b && a_scale && b_scale, B200_ERR_INVALID, "cutlass_scaled_mm: null pointer");
  B200_CHECK(out_dtype == B200_BF16 || out_dtype == B200_F16, B200_ERR_INVALID,
             "cutlass_scaled_mm: out dtype must be bf16 or f16");
  B200_CHECK(lda % 16 == 0 && ldb % 16 == 0 && K % 16 == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm: K and the row strides of a / b must be multiples of 16 (got K=%d lda=%lld ldb=%lld)",
             K, (long long)lda, (long long)ldb);
  B200_CHECK(((uintptr_t)a & 15) == 0 && ((uintptr_t)b & 15) == 0, B200_ERR_INVALID,
             "cutlass_scaled_mm: a / b must be 16-byte aligned");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "cutlass_scaled_mm: no CUDA device (no CPU fallback)");
  B200_CHECK(workspace && ((uintptr_t)workspace & 255) == 0 && workspace_bytes >= b200_fp8_gemm_v3_workspace_bytes(),
             B200_ERR_WORKSPACE, "cutlass_scaled_mm: workspace of %zu bytes (256-byte aligned, zero-initialised) required",
             b200_fp8_gemm_v3_workspace_bytes());
  Params p;
  memset(&p, 0, sizeof(p));
  p.out = out;
  p.bias = bias;
  p.K = K;
  p.ldo = ldo;
  p.out_bf16 = out_dtype == B200_BF16;
#ifdef B200_V3_TRACE
  p.trace = g_trace;
#endif
  p.ws_flags = reinterpret_cast<uint32_t*>(workspace);
  p.ws_partials = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(workspace) + WS_FLAG_BYTES);
  p.num_kb = (K + BK - 1) / BK;
  const bool swap = M <= 256;
  p.swap = swap ? 1 : 0;
  CUtensorMap tmR, tmC, tmO;
  memset(&tmO, 0, sizeof(tmO));
  int rc, bn;
  if (swap) {
    bn = M <= 64 ? 64 : M <= 128 ? 128 : 256;
    p.R = N; p.C = M;
    p.r_scale = b_scale; p.r_scale_vec = b_scale_per_col;
    p.c_scale = a_scale; p.c_scale_vec = a_scale_per_row;
    rc = make_tmap_2d(&tmR, b, N, K, ldb, BM);
    if (rc) return rc;
    rc = make_tmap_2d(&tmC, a, M, K, lda, bn / 2);
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
    rc = make_tmap_2d(&tmC, b, N, K, ldb, bn / 2);
    if (rc) return rc;
  }
  p.row_tiles = (p.R + 2 * BM - 1) / (2 * BM);
  p.col_tiles = (p.C + bn - 1) / bn;
  switch (bn) {
    case 64: return launch<64>(tmR, tmC, tmO, p, stream);
    case 128: return launch<128>(tmR, tmC, tmO, p, stream);
    default: return launch<256>(tmR, tmC, tmO, p, stream);
  }
}
