// b200ops -- variable-length (prefill / chunked-prefill / mixed) paged attention for sm_100a (B200).
//
// Contract: the query_len >= 1 branch of AttentionImpl.forward as implemented by the reference's Triton
// unified attention (vllm/v1/attention/ops/triton_unified_attention.py:179-683, wrapper :802-1140;
// test oracle tests/kernels/attention/test_triton_unified_attention.py:33-88): packed queries
// q[T, Hq, D] with cu_seqlens_q[S+1], paged KV [nb, bs, Hkv, D] through block_table[S, max_blocks],
// seq_lens[S] = context + query length, causal mask `kv <= context + q_pos`, optional sliding window
// / soft-cap / ALiBi, bf16 / f16 or FP8-E4M3 KV with per-tensor scales, fp32 softmax, optional fused
// static FP8 output and natural-log LSE.
//
// B200 design (first version: mma.sync tensor cores; a tcgen05 FA4-style version is the follow-up):
//  * grid = (q blocks, kv heads). A q block is 64 rows = (BLOCK_Q = 64 / g query tokens) x (the g query
//    heads of one kv head): the GQA group shares every K/V byte, as in the reference kernel's BLOCK_M.
//    q blocks are enumerated like the reference (`q.shape[0] // BLOCK_Q + num_seqs` upper bound, binary
//    search over cu_seqlens_q on the device: no host sync, CUDA-graph safe).
//  * 4 warps x 16 rows. K/V tiles of 64 tokens are gathered page by page with 16-byte cp.async into a
//    double-buffered, XOR-swizzled shared-memory image (rows past the sequence end are zero-filled so
//    stale cache bytes can never reach P.V).
//  * S = Q.K^T and O += P.V with mma.sync.m16n8k16 (bf16, or f16 when the KV cache is FP8: e4m3 -> f16 is
//    exact), fp32 accumulation. The head-dim contraction of Q.K^T is permuted so each lane reads K as
//    16-byte chunks; V^T fragments come from ldmatrix.trans (FP8: byte pairs split with PRMT, which
//    permutes the output head-dim inside 16-wide groups -- undone when the row is written).
//  * online softmax in the exp2 domain, quad-shuffle row reductions.
#include "common.cuh"

namespace b200 {
namespace attnp {

constexpr int BM = 64;   // query rows per CTA
constexpr int BN = 64;   // kv tokens per tile
constexpr int NTHREADS = 128;
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

enum Mode { BF16_BF16 = 0, F16_F16 = 1, BF16_FP8 = 2, F16_FP8 = 3 };

struct Params {
  void* out;
  float* lse;               // [T, Hq] natural log, nullable
  const void* q;
  const uint8_t* k_cache;
  const uint8_t* v_cache;
  const int32_t* cu_seqlens_q;
  const int32_t* seq_lens;
  const int32_t* block_table;
  const float* k_scale;
  const float* v_scale;
  const float* out_scale;
  const float* alibi;
  int64_t q_stride_t, q_stride_h, out_stride_t, out_stride_h;
  int64_t kv_block_stride, kv_token_stride, kv_head_stride;   // bytes
  int64_t bt_stride;
  int S, T, Hq, Hkv, g, bs, block_q;
  int window;               // 0 = none; else kv visible iff (context + q_pos - kv) < window
  float scale, softcap;
};

__device__ __forceinline__ float ex2(float x) {
  float y;
  asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ float lg2(float x) {
  float y;
  asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
  return y;
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
template <bool BF16>
__device__ __forceinline__ void mma16816(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
  if constexpr (BF16) {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  } else {
    asm volatile(
        "mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
        "{%0,%1,%2,%3};"
        : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
        : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
  }
}
template <bool BF16>
__device__ __forceinline__ uint32_t pack2(float lo, float hi) {
  if constexpr (BF16) {
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  } else {
    __half2 v = __floats2half2_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
  }
}
__device__ __forceinline__ uint32_t bf16x2_to_f16x2(uint32_t v) {
  __nv_bfloat162 b = *reinterpret_cast<__nv_bfloat162*>(&v);
  float2 f = __bfloat1622float2(b);
  __half2 h = __floats2half2_rn(f.x, f.y);
  return *reinterpret_cast<uint32_t*>(&h);
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 r;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];"
               : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w)
               : "r"(addr));
  return r;
}
__device__ __forceinline__ void ldmatrix_x4_trans(uint32_t addr, uint32_t (&r)[4]) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3])
               : "r"(addr));
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}

// byte offset of 16-byte chunk `c` of row `row` inside a [BN][RB] tile. The XOR keeps (a) the 8 rows of an
// ldmatrix phase and (b) the 2 rows x 4 chunks of a K-fragment quarter-warp on distinct bank groups.
template <int RB>
__device__ __forceinline__ uint32_t tile_off(int row, int c) {
  constexpr int CPR = RB / 16;
  if constexpr (CPR >= 8) {
    const int f = ((row & 1) << 2) | ((row >> 1) & 3);
    return (uint32_t)(row * RB + (((c & ~7) | ((c & 7) ^ f)) << 4));
  } else {
    return (uint32_t)(row * RB + ((c ^ ((row >> 1) & 3)) << 4));
  }
}

template <int D, int MODE>
__global__ void __launch_bounds__(NTHREADS) varlen_attn_kernel(const Params p) {
  constexpr bool FP8 = (MODE == BF16_FP8 || MODE == F16_FP8);
  constexpr bool QBF16 = (MODE == BF16_BF16 || MODE == BF16_FP8);
  constexpr bool MMA_BF16 = (MODE == BF16_BF16);   // FP8 KV is expanded to f16 (exact)
  constexpr int ES = FP8 ? 1 : 2;
  constexpr int RB = D * ES;          // bytes of one K / V row
  constexpr int CPR = RB / 16;        // 16-byte chunks per row
  constexpr int KS = D / 16;          // Q.K^T k-steps
  constexpr int NTO = D / 8;          // output n-tiles
  constexpr uint32_t TILE = BN * RB;

  extern __shared__ __align__(128) uint8_t smem_raw[];
  const uint32_t sbase = (smem_u32(smem_raw) + 127u) & ~127u;   // [stage][K tile | V tile]

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int r = lane >> 2, q4 = lane & 3;
  const int kvh = blockIdx.y;
  const int BQ = p.block_q, g = p.g;

  // ---- which sequence and q block (resolve_seq_and_query_len, triton_unified_attention.py:311-318)
  const int bid = blockIdx.x;
  int s;
  {
    int lo = 0, hi = p.S;   // invariant: start_block(lo) <= bid
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
  const int j_lo = kv_lo / BN, j_hi = (kv_hi + BN - 1) / BN;

  // ---- the two rows of this lane: row m = (query token m / g, query head m % g of the group)
  int qpos[2], head[2], limit[2];
  bool valid[2];
  float slope[2];
#pragma unroll
  for (int ri = 0; ri < 2; ri++) {
    const int m = warp * 16 + r + 8 * ri;
    const int tl = m / g;
    qpos[ri] = first_q + tl;
    head[ri] = kvh * g + (m - tl * g);
    valid[ri] = tl < BQ && qpos[ri] < q_len;
    limit[ri] = ctx + qpos[ri];   // last visible kv index
    slope[ri] = (p.alibi && valid[ri]) ? p.alibi[head[ri]] : 0.f;
  }

  // ---- Q fragments (A operand), head-dim permuted like the K reads below
  uint32_t qf[KS][4];
#pragma unroll
  for (int ri = 0; ri < 2; ri++) {
    const uint16_t* qp = reinterpret_cast<const uint16_t*>(p.q) + (int64_t)(q_start + qpos[ri]) * p.q_stride_t +
                         (int64_t)head[ri] * p.q_stride_h;
    if constexpr (!FP8) {
#pragma unroll
      for (int i = 0; i < D / 32; i++) {
        uint4 v = make_uint4(0, 0, 0, 0);
        if (valid[ri]) v = *reinterpret_cast<const uint4*>(qp + 32 * i + 8 * q4);
        qf[2 * i][ri] = v.x; qf[2 * i][2 + ri] = v.y;
        qf[2 * i + 1][ri] = v.z; qf[2 * i + 1][2 + ri] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < D / 64; i++) {
        uint4 v0 = make_uint4(0, 0, 0, 0), v1 = make_uint4(0, 0, 0, 0);
        if (valid[ri]) {
          v0 = *reinterpret_cast<const uint4*>(qp + 64 * i + 16 * q4);
          v1 = *reinterpret_cast<const uint4*>(qp + 64 * i + 16 * q4 + 8);
        }
        const uint32_t w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
        for (int k = 0; k < 4; k++) {
          qf[4 * i + k][ri] = QBF16 ? bf16x2_to_f16x2(w[2 * k]) : w[2 * k];
          qf[4 * i + k][2 + ri] = QBF16 ? bf16x2_to_f16x2(w[2 * k + 1]) : w[2 * k + 1];
        }
      }
    }
  }

  const float k_scale = (FP8 && p.k_scale) ? *p.k_scale : 1.f;
  const float v_scale = (FP8 && p.v_scale) ? *p.v_scale : 1.f;
  const float qk_scale = p.scale * k_scale;
  const float softcap = p.softcap;

  float acc[NTO][4];
#pragma unroll
  for (int nt = 0; nt < NTO; nt++)
#pragma unroll
    for (int c = 0; c < 4; c++) acc[nt][c] = 0.f;
  float mrow[2] = {-INFINITY, -INFINITY}, lrow[2] = {0.f, 0.f};

  // ---- gather of one 64-token K/V tile: 16-byte cp.async per (row, chunk), page looked up per row
  const int32_t* bt = p.block_table + (int64_t)s * p.bt_stride;
  auto load_tile = [&](int j, int stage) {
    const uint32_t kdst = sbase + (uint32_t)stage * 2u * TILE;
    const uint32_t vdst = kdst + TILE;
    const int t0 = j * BN;
#pragma unroll
    for (int it = 0; it < BN * CPR / NTHREADS; it++) {
      const int idx = tid + it * NTHREADS;
      const int row = idx / CPR, c = idx - row * CPR;
      const int t = t0 + row;
      const bool ok = t < kv_hi;
      int64_t off = 0;
      if (ok) {
        const int lb = t / p.bs;
        off = (int64_t)bt[lb] * p.kv_block_stride + (int64_t)(t - lb * p.bs) * p.kv_token_stride +
              (int64_t)kvh * p.kv_head_stride + c * 16;
      }
      const uint32_t so = tile_off<RB>(row, c);
      cp_async16(kdst + so, p.k_cache + off, ok ? 16u : 0u);
      cp_async16(vdst + so, p.v_cache + off, ok ? 16u : 0u);
    }
  };

  if (j_lo < j_hi) load_tile(j_lo, 0);
  cp_async_commit();

  for (int j = j_lo; j < j_hi; j++) {
    const int stage = (j - j_lo) & 1;
    cp_async_wait<0>();
    __syncthreads();                       // tile j visible to all warps; tile j-1 fully consumed
    if (j + 1 < j_hi) load_tile(j + 1, stage ^ 1);
    cp_async_commit();

    const uint32_t kbuf = sbase + (uint32_t)stage * 2u * TILE;
    const uint32_t vbuf = kbuf + TILE;
    const int t0 = j * BN;

    // ------------------------------------------------ S = Q . K^T   (16 rows x 64 tokens per warp)
    float sc[BN / 8][4];
#pragma unroll
    for (int nt = 0; nt < BN / 8; nt++) {
#pragma unroll
      for (int c = 0; c < 4; c++) sc[nt][c] = 0.f;
      const int row = nt * 8 + r;
      if constexpr (!FP8) {
#pragma unroll
        for (int i = 0; i < D / 32; i++) {
          const uint4 k = lds128(kbuf + tile_off<RB>(row, 4 * i + q4));
          mma16816<MMA_BF16>(sc[nt], qf[2 * i], k.x, k.y);
          mma16816<MMA_BF16>(sc[nt], qf[2 * i + 1], k.z, k.w);
        }
      } else {
#pragma unroll
        for (int i = 0; i < D / 64; i++) {
          const uint4 k = lds128(kbuf + tile_off<RB>(row, 4 * i + q4));
          const uint32_t kw[4] = {k.x, k.y, k.z, k.w};
#pragma unroll
          for (int w = 0; w < 4; w++)
            mma16816<MMA_BF16>(sc[nt], qf[4 * i + w], e4m3x2_to_f16x2((uint16_t)(kw[w] & 0xffffu)),
                               e4m3x2_to_f16x2((uint16_t)(kw[w] >> 16)));
        }
      }
    }

    // ------------------------------------------------ scale, soft-cap, ALiBi, masks (exp2 domain)
    const bool need_mask = (t0 + BN - 1 > ctx + first_q) || (t0 + BN > kv_hi) ||
                           (p.window > 0 && t0 < ctx + last_q - p.window + 1) || !(valid[0] && valid[1]);
    float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
    for (int nt = 0; nt < BN / 8; nt++) {
#pragma unroll
      for (int c = 0; c < 4; c++) {
        const int ri = c >> 1;
        const int t = t0 + nt * 8 + 2 * q4 + (c & 1);
        float v = sc[nt][c] * qk_scale;
        if (softcap > 0.f) v = softcap * tanhf(v / softcap);
        v += slope[ri] * (float)(t - ctx);
        v *= LOG2E;
        if (need_mask) {
          const bool vis = valid[ri] && t <= limit[ri] && t < kv_hi &&
                           (p.window <= 0 || limit[ri] - t < p.window);
          if (!vis) v = -INFINITY;
        }
        sc[nt][c] = v;
        mx[ri] = fmaxf(mx[ri], v);
      }
    }
    float alpha[2], mbase[2];
#pragma unroll
    for (int ri = 0; ri < 2; ri++) {
      mx[ri] = fmaxf(mx[ri], __shfl_xor_sync(0xffffffffu, mx[ri], 1));
      mx[ri] = fmaxf(mx[ri], __shfl_xor_sync(0xffffffffu, mx[ri], 2));
      const float mnew = fmaxf(mrow[ri], mx[ri]);
      alpha[ri] = (mnew == -INFINITY) ? 1.f : ex2(mrow[ri] - mnew);   // mrow = -inf -> 0
      mrow[ri] = mnew;
      mbase[ri] = (mnew == -INFINITY) ? 0.f : mnew;
      lrow[ri] *= alpha[ri];
    }
#pragma unroll
    for (int nt = 0; nt < NTO; nt++) {
      acc[nt][0] *= alpha[0]; acc[nt][1] *= alpha[0];
      acc[nt][2] *= alpha[1]; acc[nt][3] *= alpha[1];
    }
    // P as A fragments: k-step kt covers tokens 16 kt .. 16 kt + 15 = score n-tiles 2 kt, 2 kt + 1
    uint32_t pa[BN / 16][4];
#pragma unroll
    for (int kt = 0; kt < BN / 16; kt++) {
      float e[2][4];
#pragma unroll
      for (int h = 0; h < 2; h++)
#pragma unroll
        for (int c = 0; c < 4; c++) {
          e[h][c] = ex2(sc[2 * kt + h][c] - mbase[c >> 1]);
          lrow[c >> 1] += e[h][c];
        }
      pa[kt][0] = pack2<MMA_BF16>(e[0][0], e[0][1]);
      pa[kt][1] = pack2<MMA_BF16>(e[0][2], e[0][3]);
      pa[kt][2] = pack2<MMA_BF16>(e[1][0], e[1][1]);
      pa[kt][3] = pack2<MMA_BF16>(e[1][2], e[1][3]);
    }

    // ------------------------------------------------ O += P . V
    const int mi = lane >> 3, rin = lane & 7;
#pragma unroll
    for (int kt = 0; kt < BN / 16; kt++) {
      const int vrow = kt * 16 + (mi & 1) * 8 + rin;
#pragma unroll
      for (int jp = 0; jp < CPR / 2; jp++) {
        uint32_t R[4];
        ldmatrix_x4_trans(vbuf + tile_off<RB>(vrow, 2 * jp + (mi >> 1)), R);
        if constexpr (!FP8) {
          mma16816<MMA_BF16>(acc[2 * jp], pa[kt], R[0], R[1]);
          mma16816<MMA_BF16>(acc[2 * jp + 1], pa[kt], R[2], R[3]);
        } else {
          // R bytes: (V[t][2n], V[t][2n+1], V[t+1][2n], V[t+1][2n+1]) -> token pairs of d = 2n and d = 2n + 1
          uint32_t x[4];
#pragma unroll
          for (int k = 0; k < 4; k++) x[k] = prmt(R[k], 0u, 0x3120);
#pragma unroll
          for (int h = 0; h < 2; h++) {   // chunk 2 jp + h = 16 head-dim values
            const int c = 2 * jp + h;
            mma16816<MMA_BF16>(acc[2 * c], pa[kt], e4m3x2_to_f16x2((uint16_t)(x[2 * h] & 0xffffu)),
                               e4m3x2_to_f16x2((uint16_t)(x[2 * h + 1] & 0xffffu)));
            mma16816<MMA_BF16>(acc[2 * c + 1], pa[kt], e4m3x2_to_f16x2((uint16_t)(x[2 * h] >> 16)),
                               e4m3x2_to_f16x2((uint16_t)(x[2 * h + 1] >> 16)));
          }
        }
      }
    }
  }
  cp_async_wait<0>();

  // ---- finish: row sums over the quad, normalise, write
#pragma unroll
  for (int ri = 0; ri < 2; ri++) {
    lrow[ri] += __shfl_xor_sync(0xffffffffu, lrow[ri], 1);
    lrow[ri] += __shfl_xor_sync(0xffffffffu, lrow[ri], 2);
  }
  const bool fp8_out = p.out_scale != nullptr;
  const float qinv = fp8_out ? __fdiv_rn(1.0f, *p.out_scale) : 1.f;
#pragma unroll
  for (int ri = 0; ri < 2; ri++) {
    if (!valid[ri]) continue;
    const float inv = lrow[ri] > 0.f ? v_scale / lrow[ri] : 0.f;
    const int64_t tok = q_start + qpos[ri];
    if (p.lse && q4 == 0)
      p.lse[tok * p.Hq + head[ri]] = lrow[ri] > 0.f ? (mrow[ri] + lg2(lrow[ri])) * LN2 : -INFINITY;
    const int64_t obase = tok * p.out_stride_t + (int64_t)head[ri] * p.out_stride_h;
    if constexpr (!FP8) {
#pragma unroll
      for (int nt = 0; nt < NTO; nt++) {
        const float a = acc[nt][2 * ri] * inv, b = acc[nt][2 * ri + 1] * inv;
        const int d = nt * 8 + 2 * q4;
        if (fp8_out) {
          *reinterpret_cast<uint16_t*>(reinterpret_cast<uint8_t*>(p.out) + obase + d) =
              f32x2_to_e4m3x2(fmaxf(-448.f, fminf(a * qinv, 448.f)), fmaxf(-448.f, fminf(b * qinv, 448.f)));
        } else {
          *reinterpret_cast<uint32_t*>(reinterpret_cast<uint16_t*>(p.out) + obase + d) = pack2<QBF16>(a, b);
        }
      }
    } else {
#pragma unroll
      for (int c = 0; c < D / 16; c++) {
        // lane holds d = 16 c + 4 q4 + {0, 1, 2, 3} = (even, col 0), (odd, col 0), (even, col 1), (odd, col 1)
        const float o0 = acc[2 * c][2 * ri] * inv, o1 = acc[2 * c + 1][2 * ri] * inv;
        const float o2 = acc[2 * c][2 * ri + 1] * inv, o3 = acc[2 * c + 1][2 * ri + 1] * inv;
        const int d = 16 * c + 4 * q4;
        if (fp8_out) {
          const uint32_t lo8 = f32x2_to_e4m3x2(fmaxf(-448.f, fminf(o0 * qinv, 448.f)), fmaxf(-448.f, fminf(o1 * qinv, 448.f)));
          const uint32_t hi8 = f32x2_to_e4m3x2(fmaxf(-448.f, fminf(o2 * qinv, 448.f)), fmaxf(-448.f, fminf(o3 * qinv, 448.f)));
          *reinterpret_cast<uint32_t*>(reinterpret_cast<uint8_t*>(p.out) + obase + d) = lo8 | (hi8 << 16);
        } else {
          *reinterpret_cast<uint2*>(reinterpret_cast<uint16_t*>(p.out) + obase + d) =
              make_uint2(pack2<QBF16>(o0, o1), pack2<QBF16>(o2, o3));
        }
      }
    }
  }
}

}  // namespace attnp
}  // namespace b200

#include "attention_prefill_tc.cuh"
#include "attention_prefill_tc8.cuh"

#include <cstdlib>
#include <cstring>

namespace b200 {
namespace attnp {

// 5-D view of a paged K (or V) cache for the tcgen05 kernel: [64 elements = 128 B][token in block][128-B segment]
// [head][block]; one box = 16 tokens x 128 B of one head
static int make_kv_tmap(CUtensorMap* tm, const void* base, int64_t nb, int64_t sb, int64_t st, int64_t sh, int bs,
                        int hkv) {
  uint64_t dims[5] = {64, (uint64_t)bs, 2, (uint64_t)hkv, (uint64_t)nb};
  uint64_t strides[4] = {(uint64_t)(st * 2), 128, (uint64_t)(sh * 2), (uint64_t)(sb * 2)};
  uint32_t box[5] = {64, 16, 1, 1, 1};
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT16, 5, base, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

// FP8 cache: [128 bytes = the whole head][token in block][1][head][block]; one box = 16 tokens x 128 B
static int make_kv8_tmap(CUtensorMap* tm, const void* base, int64_t nb, int64_t sb, int64_t st, int64_t sh, int bs,
                         int hkv) {
  uint64_t dims[5] = {128, (uint64_t)bs, 1, (uint64_t)hkv, (uint64_t)nb};
  uint64_t strides[4] = {(uint64_t)st, 128, (uint64_t)sh, (uint64_t)sb};
  uint32_t box[5] = {128, 16, 1, 1, 1};
  return encode_tensor_map(tm, CU_TENSOR_MAP_DATA_TYPE_UINT8, 5, base, dims, strides, box, CU_TENSOR_MAP_SWIZZLE_128B,
                           CU_TENSOR_MAP_L2_PROMOTION_L2_256B);
}

template <bool QBF16, bool GENERAL>
static int launch_tc8(const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, cudaStream_t stream) {
  constexpr int smem = 4 * (int)TC_TILE + 4 * (int)TC8_RAW + 1024 + 256;
  static bool configured_dev[64] = {};   // the attribute is per device
  int cfg_dev_ = 0;
  cudaGetDevice(&cfg_dev_);
  bool& configured = configured_dev[cfg_dev_ & 63];
  auto kern = varlen_attn_tc8_kernel<QBF16, GENERAL>;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(p.T / p.block_q + p.S, p.Hkv);
  kern<<<grid, TC8_THREADS, smem, stream>>>(tmK, tmV, p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <bool BF16, bool GENERAL>
static int launch_tc(const CUtensorMap& tmK, const CUtensorMap& tmV, const Params& p, cudaStream_t stream) {
  constexpr int smem = 6 * (int)TC_TILE + 1024 + 256;
  static bool configured_dev[64] = {};   // the attribute is per device
  int cfg_dev_ = 0;
  cudaGetDevice(&cfg_dev_);
  bool& configured = configured_dev[cfg_dev_ & 63];
  auto kern = varlen_attn_tc_kernel<BF16, GENERAL>;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  dim3 grid(p.T / p.block_q + p.S, p.Hkv);
  kern<<<grid, TC_THREADS, smem, stream>>>(tmK, tmV, p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

template <int D, int MODE>
static int launch(const Params& p, cudaStream_t stream) {
  constexpr int ES = (MODE == BF16_FP8 || MODE == F16_FP8) ? 1 : 2;
  constexpr int smem = 2 * 2 * BN * D * ES + 128;
  static bool configured_dev[64] = {};   // the attribute is per device
  int cfg_dev_ = 0;
  cudaGetDevice(&cfg_dev_);
  bool& configured = configured_dev[cfg_dev_ & 63];
  auto kern = varlen_attn_kernel<D, MODE>;
  if (!configured) {
    B200_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, smem));
    configured = true;
  }
  // upper bound on the number of q blocks, as the reference: q.shape[0] // BLOCK_Q + num_seqs
  dim3 grid(p.T / p.block_q + p.S, p.Hkv);
  kern<<<grid, NTHREADS, smem, stream>>>(p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace attnp
}  // namespace b200

extern "C" int b200_paged_attention_varlen(const b200_varlen_attn_args* a, b200_stream_t stream_) {
  using namespace b200;
  using namespace b200::attnp;
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(a != nullptr, B200_ERR_INVALID, "unified_attention: null args");
  const int S = a->num_seqs, T = a->num_tokens, Hq = a->num_heads, Hkv = a->num_kv_heads, D = a->head_size;
  B200_CHECK(S >= 0 && T >= 0 && Hq > 0 && Hkv > 0 && D > 0, B200_ERR_INVALID,
             "unified_attention: bad sizes (seqs %d tokens %d heads %d kv_heads %d head_size %d)", S, T, Hq, Hkv, D);
  if (S == 0 || T == 0) return B200_OK;
  B200_CHECK(Hq % Hkv == 0, B200_ERR_INVALID,
             "unified_attention: num_heads (%d) must be divisible by num_kv_heads (%d)", Hq, Hkv);
  B200_CHECK(a->causal != 0, B200_ERR_UNSUPPORTED, "unified_attention: only causal attention is supported");
  B200_CHECK(a->out && a->q && a->k_cache && a->v_cache && a->block_table && a->seq_lens && a->cu_seqlens_q,
             B200_ERR_INVALID, "unified_attention: null pointer");
  B200_CHECK(a->block_size > 0 && a->block_size % 16 == 0, B200_ERR_INVALID,
             "unified_attention: block_size (%d) must be a multiple of 16", a->block_size);
  B200_CHECK(a->q_dtype == B200_BF16 || a->q_dtype == B200_F16, B200_ERR_INVALID,
             "unified_attention: query dtype must be bf16 or f16");
  const bool fp8 = a->kv_dtype == B200_FP8_E4M3;
  B200_CHECK(fp8 || a->kv_dtype == a->q_dtype, B200_ERR_INVALID,
             "unified_attention: kv dtype must equal query dtype (\"auto\") or be fp8_e4m3");
  const int es = fp8 ? 1 : 2;
  B200_CHECK(D == 64 || D == 128, B200_ERR_UNSUPPORTED, "unified_attention: head_size %d not supported (64, 128)", D);
  const int g = Hq / Hkv;
  B200_CHECK(g <= BM, B200_ERR_UNSUPPORTED, "unified_attention: more than %d query heads per kv head (%d)", BM, g);
  auto al16 = [](const void* ptr) { return (reinterpret_cast<uintptr_t>(ptr) & 15) == 0; };
  const int oes = a->output_scale ? 1 : 2;
  B200_CHECK(al16(a->q) && al16(a->out) && al16(a->k_cache) && al16(a->v_cache) && (a->q_stride_t * 2) % 16 == 0 &&
                 (a->q_stride_h * 2) % 16 == 0 && (a->out_stride_t * oes) % 8 == 0 && (a->out_stride_h * oes) % 8 == 0,
             B200_ERR_INVALID, "unified_attention: q/out/cache pointers and strides must be 16-byte aligned");
  B200_CHECK((a->kv_block_stride * es) % 16 == 0 && (a->kv_token_stride * es) % 16 == 0 &&
                 (a->kv_head_stride * es) % 16 == 0 && a->kv_block_stride > 0 && a->kv_token_stride > 0 &&
                 a->kv_head_stride > 0,
             B200_ERR_INVALID, "unified_attention: kv cache strides must be positive multiples of 16 bytes");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "unified_attention: no CUDA device (b200ops has no CPU fallback)");

  Params p;
  p.out = a->out;
  p.lse = a->lse;
  p.q = a->q;
  p.k_cache = reinterpret_cast<const uint8_t*>(a->k_cache);
  p.v_cache = reinterpret_cast<const uint8_t*>(a->v_cache);
  p.cu_seqlens_q = a->cu_seqlens_q;
  p.seq_lens = a->seq_lens;
  p.block_table = a->block_table;
  p.k_scale = fp8 ? a->k_scale : nullptr;
  p.v_scale = fp8 ? a->v_scale : nullptr;
  p.out_scale = a->output_scale;
  p.alibi = a->alibi_slopes;
  p.q_stride_t = a->q_stride_t; p.q_stride_h = a->q_stride_h;
  p.out_stride_t = a->out_stride_t; p.out_stride_h = a->out_stride_h;
  p.kv_block_stride = a->kv_block_stride * es;
  p.kv_token_stride = a->kv_token_stride * es;
  p.kv_head_stride = a->kv_head_stride * es;
  p.bt_stride = a->bt_stride;
  p.S = S; p.T = T; p.Hq = Hq; p.Hkv = Hkv; p.g = g; p.bs = a->block_size;
  p.block_q = BM / g;
  p.window = a->sliding_window > 0 ? a->sliding_window : 0;
  p.scale = a->scale;
  p.softcap = a->softcap > 0.f ? a->softcap : 0.f;

  // tcgen05 kernel: 16-bit caches, head_size 128 (B200_PREFILL_IMPL=mma forces the mma.sync kernel)
  const char* impl = std::getenv("B200_PREFILL_IMPL");
  const bool want_mma = impl && std::strcmp(impl, "mma") == 0;
  const bool tc_ok = D == TC_D && g <= TC_BM && a->num_blocks > 0 && !want_mma && (a->out_stride_t * oes) % 16 == 0 &&
                     (a->out_stride_h * oes) % 16 == 0;
  if (fp8 && tc_ok) {
    p.block_q = TC_BM / g;
    CUtensorMap tmK, tmV;
    int rc = make_kv8_tmap(&tmK, a->k_cache, a->num_blocks, a->kv_block_stride, a->kv_token_stride, a->kv_head_stride,
                           a->block_size, Hkv);
    if (rc) return rc;
    rc = make_kv8_tmap(&tmV, a->v_cache, a->num_blocks, a->kv_block_stride, a->kv_token_stride, a->kv_head_stride,
                       a->block_size, Hkv);
    if (rc) return rc;
    const bool general = p.softcap > 0.f || p.alibi != nullptr;
    if (a->q_dtype == B200_BF16)
      return general ? launch_tc8<true, true>(tmK, tmV, p, stream) : launch_tc8<true, false>(tmK, tmV, p, stream);
    return general ? launch_tc8<false, true>(tmK, tmV, p, stream) : launch_tc8<false, false>(tmK, tmV, p, stream);
  }
  if (!fp8 && tc_ok) {
    p.block_q = TC_BM / g;
    CUtensorMap tmK, tmV;
    int rc = make_kv_tmap(&tmK, a->k_cache, a->num_blocks, a->kv_block_stride, a->kv_token_stride, a->kv_head_stride,
                          a->block_size, Hkv);
    if (rc) return rc;
    rc = make_kv_tmap(&tmV, a->v_cache, a->num_blocks, a->kv_block_stride, a->kv_token_stride, a->kv_head_stride,
                      a->block_size, Hkv);
    if (rc) return rc;
    const bool general = p.softcap > 0.f || p.alibi != nullptr;
    if (a->q_dtype == B200_BF16)
      return general ? launch_tc<true, true>(tmK, tmV, p, stream) : launch_tc<true, false>(tmK, tmV, p, stream);
    return general ? launch_tc<false, true>(tmK, tmV, p, stream) : launch_tc<false, false>(tmK, tmV, p, stream);
  }
  const int mode = fp8 ? (a->q_dtype == B200_BF16 ? BF16_FP8 : F16_FP8)
                       : (a->q_dtype == B200_BF16 ? BF16_BF16 : F16_F16);
#define B200_VARLEN_CASE(DD, MM) \
  if (D == DD && mode == MM) return launch<DD, MM>(p, stream);
  B200_VARLEN_CASE(64, BF16_BF16) B200_VARLEN_CASE(64, F16_F16) B200_VARLEN_CASE(64, BF16_FP8) B200_VARLEN_CASE(64, F16_FP8)
  B200_VARLEN_CASE(128, BF16_BF16) B200_VARLEN_CASE(128, F16_F16) B200_VARLEN_CASE(128, BF16_FP8) B200_VARLEN_CASE(128, F16_FP8)
#undef B200_VARLEN_CASE
  set_last_error("unified_attention: no kernel for head_size %d, mode %d", D, mode);
  return B200_ERR_UNSUPPORTED;
}
