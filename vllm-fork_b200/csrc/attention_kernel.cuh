// b200ops -- paged-attention decode kernel body (included by attention.cu).
#pragma once

namespace b200 {
namespace attn {

__device__ __forceinline__ int seq_tiles(const int32_t* __restrict__ seq_lens, int s, int window,
                                         int& lo_tile) {
  const int L = seq_lens[s];
  lo_tile = 0;
  if (L <= 0) return 0;
  if (window > 0 && L > window) lo_tile = (L - window) >> 4;
  return ((L + 15) >> 4) - lo_tile;
}

// Consumer-side state of one (sequence segment, kv head) handled by one warp.
template <int D, int NT>
struct WarpState {
  float acc[NT][D / 16][4];  // O^T accumulators (mma C layout)
  float m[NT][2];            // running max (log2 domain) for q-heads 2q, 2q+1 of each n-tile
  float l[NT][2];            // thread-partial running sum
};

template <int D, int NT, int MODE>
__global__ void __launch_bounds__(NTHREADS, 1)
    paged_attn_decode_kernel(const __grid_constant__ CUtensorMap tmK,
                             const __grid_constant__ CUtensorMap tmV, const Params p) {
  constexpr bool FP8 = (MODE == BF16_FP8 || MODE == F16_FP8);
  constexpr bool QBF16 = (MODE == BF16_BF16 || MODE == BF16_FP8);
  constexpr bool MMA_BF16 = (MODE == BF16_BF16);  // fp8 KV is expanded to f16 (exact)
  constexpr int ES = FP8 ? 1 : 2;
  using G = Geo<D, ES>;
  constexpr int IB = G::IB, NSEG = G::NSEG, EPC = G::EPC, NCK = G::NCK, KS = G::KS, MT = G::MT;
  constexpr int HEAD_SLICE = NSEG * 16 * IB;  // bytes of one head inside a tile

  extern __shared__ uint8_t smem_raw[];
  __shared__ int s_wsum[NW + 1];
  // programmatic dependent launch: the reduce kernel (and whatever follows) may become resident now. q and the
  // layer's KV rows are the previous kernels' output: everything that touches them sits behind pdl_wait(). At level 2
  // (p.early_plan) the tile plan below, which reads only seq_lens, runs BEFORE the wait (the caller promises that
  // seq_lens is not written by the kernel launched immediately before this one: it is a step input).
  pdl_launch_dependents();
  if (!p.early_plan) pdl_wait();
  const uint32_t raw_base = smem_u32(smem_raw);
  const uint32_t base = (raw_base + 1023u) & ~1023u;  // swizzle atoms need 1024-B alignment
  uint8_t* smem = smem_raw + (base - raw_base);
  int* cum = reinterpret_cast<int*>(smem + p.off_cum);
  const uint32_t bar_full = base + p.off_bar;
  const uint32_t bar_empty = bar_full + 8u * MAX_STAGES;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int S = p.S;

  // ---------------------------------------------------------------- plan: prefix sum of tiles
  {
    const int per = (S + NTHREADS - 1) / NTHREADS;
    const int s0 = tid * per, s1 = min(S, s0 + per);
    int local = 0;
    for (int s = s0; s < s1; s++) {
      int lo;
      local += seq_tiles(p.seq_lens, s, p.window, lo);
    }
    int inc = local;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      int n = __shfl_up_sync(0xffffffffu, inc, o);
      if (lane >= o) inc += n;
    }
    if (lane == 31) s_wsum[warp] = inc;
    __syncthreads();
    int woff = 0;
    for (int w = 0; w < warp; w++) woff += s_wsum[w];
    int run = woff + inc - local;
    for (int s = s0; s < s1; s++) {
      int lo;
      cum[s] = run;
      run += seq_tiles(p.seq_lens, s, p.window, lo);
    }
    if (tid == NTHREADS - 1) {
      int tot = 0;
      for (int w = 0; w <= NW; w++) tot += s_wsum[w];
      cum[S] = tot;
    }
    if (tid == 0) {
      for (int i = 0; i < p.stages; i++) {
        mbar_init(bar_full + 8u * i, 1);
        mbar_init(bar_empty + 8u * i, NW);
      }
      mbar_fence_init();
    }
    __syncthreads();
  }
  if (p.early_plan) pdl_wait();   // from here on: global writes (workspace) and reads of q / KV
  const int T = cum[S];
  const int tpc = max((T + (int)gridDim.x - 1) / (int)gridDim.x, p.min_tpc);
  if (blockIdx.x == 0 && blockIdx.y == 0) {
    for (int i = tid; i <= S; i += NTHREADS) p.cum[i] = cum[i];
    if (tid == 0) p.cum[S + 1] = tpc;
  }
  const int a = blockIdx.x * tpc;
  const int b = min(T, a + tpc);
  if (a >= b) return;

  // first sequence of this CTA: largest s with cum[s] <= a (skipping empty sequences)
  int sfirst;
  {
    int lo = 0, hi = S;  // invariant: cum[lo] <= a < cum[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (cum[mid] <= a) lo = mid; else hi = mid;
    }
    sfirst = lo;
  }
  const int h0 = blockIdx.y * p.hc;  // first kv head of this CTA
  const int tg = p.tg;
  const uint32_t TB = p.tile_bytes;
  const uint32_t stage_bytes = (uint32_t)tg * 2u * TB;

  if (warp == 0) {
    // ================================================================ producer (TMA)
    if (lane == 0) {
      prefetch_tmap(&tmK);
      prefetch_tmap(&tmV);
    }
    const uint64_t pol = l2_policy_evict_first();  // KV bytes are read exactly once
    int stage = 0;
    uint32_t phase = 0;
    int pos = a, s = sfirst;
    while (pos < b) {
      const int c0 = cum[s], c1 = cum[s + 1];
      if (c1 <= pos) { s++; continue; }
      const int seg_end = min(b, c1);
      int lo_tile;
      seq_tiles(p.seq_lens, s, p.window, lo_tile);
      const int ta = pos - c0 + lo_tile, tb = seg_end - c0 + lo_tile;
      const int32_t* bt = p.block_table + (int64_t)s * p.bt_stride;
      for (int t0 = ta; t0 < tb; t0 += tg) {
        const int nt_ = min(tg, tb - t0);
        int blk = 0, tok_in = 0;
        if (lane < nt_) {
          const int token = (t0 + lane) << 4;
          const int lb = token / p.bs;
          tok_in = token - lb * p.bs;
          blk = bt[lb];
        }
        mbar_wait(bar_empty + 8u * stage, phase ^ 1u);
        if (lane == 0) mbar_arrive_expect_tx(bar_full + 8u * stage, (uint32_t)nt_ * 2u * TB);
        __syncwarp();
        if (lane < nt_) {
          const uint32_t dst = base + (uint32_t)stage * stage_bytes + (uint32_t)lane * 2u * TB;
          tma_load_5d(dst, &tmK, bar_full + 8u * stage, 0, tok_in, 0, h0, blk, pol);   // kv_pair: K and V in one box
          if (!p.kv_pair) tma_load_5d(dst + TB, &tmV, bar_full + 8u * stage, 0, tok_in, 0, h0, blk, pol);
        }
        if (++stage == p.stages) { stage = 0; phase ^= 1u; }
      }
      pos = seg_end;
      s++;
    }
    return;
  }

  // ================================================================ consumers
  const int cw = warp - 1;
  const int hl = cw % p.hc;           // kv head inside the CTA's head chunk
  const int tgi = cw / p.hc;          // token group
  const bool active = cw < p.hc * tg;
  const int r = lane >> 2, q4 = lane & 3;
  const float k_scale = (FP8 && p.k_scale) ? *p.k_scale : 1.f;
  const float v_scale = (FP8 && p.v_scale) ? *p.v_scale : 1.f;
  const float qk_scale = p.scale * k_scale;
  const float softcap = p.softcap;
  const int g = p.g;
  const int hq_chunk = p.hc * g;

  int stage = 0;
  uint32_t phase = 0;
  int pos = a, s = sfirst;
  while (pos < b) {
    const int c0 = cum[s], c1 = cum[s + 1];
    if (c1 <= pos) { s++; continue; }
    const int seg_end = min(b, c1);
    int lo_tile;
    seq_tiles(p.seq_lens, s, p.window, lo_tile);
    const int L = p.seq_lens[s];
    const int lo_tok = (p.window > 0 && L > p.window) ? L - p.window : 0;
    const int ta = pos - c0 + lo_tile, tb = seg_end - c0 + lo_tile;
    const bool whole = (pos == c0) && (seg_end == c1);
    const int slot = (pos == a) ? 0 : 1;

    // ---- Q fragments (B operand of S^T = K.Q^T), contraction index permuted like the K reads
    uint32_t qf[NT][KS][2];
    float slope[NT][2];
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
      const int gq = nt * 8 + r;  // q-head (within the GQA group) this lane supplies as column n = r
      const bool valid = active && gq < g;
      const int head = (h0 + hl) * g + gq;
      const uint16_t* qp = reinterpret_cast<const uint16_t*>(p.q) + (int64_t)s * p.q_stride_s +
                           (int64_t)head * p.q_stride_h;
#pragma unroll
      for (int i = 0; i < NCK; i++) {
        const int gc = NCK >= 2 ? 8 * (i >> 1) + 2 * q4 + (i & 1) : q4;
        const int d0 = gc * EPC;
        if constexpr (!FP8) {
          uint4 v = make_uint4(0, 0, 0, 0);
          if (valid) v = *reinterpret_cast<const uint4*>(qp + d0);
          qf[nt][2 * i][0] = v.x; qf[nt][2 * i][1] = v.y;
          qf[nt][2 * i + 1][0] = v.z; qf[nt][2 * i + 1][1] = v.w;
        } else {
          uint4 v0 = make_uint4(0, 0, 0, 0), v1 = make_uint4(0, 0, 0, 0);
          if (valid) {
            v0 = *reinterpret_cast<const uint4*>(qp + d0);
            v1 = *reinterpret_cast<const uint4*>(qp + d0 + 8);
          }
          uint32_t w[8] = {v0.x, v0.y, v0.z, v0.w, v1.x, v1.y, v1.z, v1.w};
#pragma unroll
          for (int pp = 0; pp < 8; pp++)
            qf[nt][4 * i + (pp >> 1)][pp & 1] = QBF16 ? bf16x2_to_f16x2(w[pp]) : w[pp];
        }
      }
#pragma unroll
      for (int a2 = 0; a2 < 2; a2++) {
        const int gq2 = nt * 8 + 2 * q4 + a2;  // q-heads whose scores / outputs this lane holds
        slope[nt][a2] = (p.alibi && gq2 < g) ? p.alibi[(h0 + hl) * g + gq2] : 0.f;
      }
    }

    WarpState<D, NT> st;
#pragma unroll
    for (int nt = 0; nt < NT; nt++) {
#pragma unroll
      for (int mt = 0; mt < MT; mt++)
#pragma unroll
        for (int c = 0; c < 4; c++) st.acc[nt][mt][c] = 0.f;
      st.m[nt][0] = st.m[nt][1] = -INFINITY;
      st.l[nt][0] = st.l[nt][1] = 0.f;
    }

    for (int t0 = ta; t0 < tb; t0 += tg) {
      const int tile = t0 + tgi;
      mbar_wait(bar_full + 8u * stage, phase);
      if (active && tile < tb) {
        const uint32_t tbase = base + (uint32_t)stage * stage_bytes + (uint32_t)tgi * 2u * TB +
                               (uint32_t)hl * HEAD_SLICE;
        const int tok0 = tile << 4;
        // ------------------------------------------------ S^T = K . Q^T
        float sc[NT][4];
#pragma unroll
        for (int nt = 0; nt < NT; nt++)
#pragma unroll
          for (int c = 0; c < 4; c++) sc[nt][c] = 0.f;
#pragma unroll
        for (int i = 0; i < NCK; i++) {
          const int sg = NCK >= 2 ? (i >> 1) : 0;
          const int ch = NCK >= 2 ? 2 * q4 + (i & 1) : q4;
          const uint32_t kb = tbase + (uint32_t)sg * 16 * IB;
          const uint4 k0 = lds128(kb + swz_off<IB>(r, ch));
          const uint4 k1 = lds128(kb + swz_off<IB>(r + 8, ch));
          const uint32_t w0[4] = {k0.x, k0.y, k0.z, k0.w};
          const uint32_t w1[4] = {k1.x, k1.y, k1.z, k1.w};
          if constexpr (FP8) {
#pragma unroll
            for (int w = 0; w < 4; w++) {
              uint32_t af[4];
              af[0] = e4m3x2_to_f16x2((uint16_t)(w0[w] & 0xffffu));
              af[1] = e4m3x2_to_f16x2((uint16_t)(w1[w] & 0xffffu));
              af[2] = e4m3x2_to_f16x2((uint16_t)(w0[w] >> 16));
              af[3] = e4m3x2_to_f16x2((uint16_t)(w1[w] >> 16));
#pragma unroll
              for (int nt = 0; nt < NT; nt++) mma16816<MMA_BF16>(sc[nt], af, qf[nt][4 * i + w]);
            }
          } else {
#pragma unroll
            for (int h = 0; h < 2; h++) {
              uint32_t af[4] = {w0[2 * h], w1[2 * h], w0[2 * h + 1], w1[2 * h + 1]};
#pragma unroll
              for (int nt = 0; nt < NT; nt++) mma16816<MMA_BF16>(sc[nt], af, qf[nt][2 * i + h]);
            }
          }
        }
        // ------------------------------------------------ online softmax (exp2 domain)
        const bool edge = (tok0 + 16 > L) || (tok0 < lo_tok);
        const int tA = tok0 + r, tB = tok0 + r + 8;
        bool need = false;
        float mnew[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
#pragma unroll
          for (int c = 0; c < 4; c++) {
            float v = sc[nt][c] * qk_scale;
            if (softcap > 0.f) v = softcap * tanhf(v / softcap);
            const int t = (c & 2) ? tB : tA;
            v += slope[nt][c & 1] * (float)(t - L + 1);
            v *= LOG2E;
            if (edge && (t >= L || t < lo_tok)) v = -INFINITY;
            sc[nt][c] = v;
          }
#pragma unroll
          for (int a2 = 0; a2 < 2; a2++) {
            float mx = fmaxf(sc[nt][a2], sc[nt][2 + a2]);
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 4));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 8));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 16));
            mnew[nt][a2] = fmaxf(st.m[nt][a2], mx);
            need |= mnew[nt][a2] > st.m[nt][a2] + RESCALE_THRESHOLD;
          }
        }
        if (__any_sync(0xffffffffu, need)) {
#pragma unroll
          for (int nt = 0; nt < NT; nt++) {
            // a fully masked column keeps m = -inf; exp2(-inf - -inf) would be NaN
            const float f0 = mnew[nt][0] == -INFINITY ? 1.f : ex2(st.m[nt][0] - mnew[nt][0]);
            const float f1 = mnew[nt][1] == -INFINITY ? 1.f : ex2(st.m[nt][1] - mnew[nt][1]);
            st.m[nt][0] = mnew[nt][0];
            st.m[nt][1] = mnew[nt][1];
            st.l[nt][0] *= f0;
            st.l[nt][1] *= f1;
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
              st.acc[nt][mt][0] *= f0; st.acc[nt][mt][1] *= f1;
              st.acc[nt][mt][2] *= f0; st.acc[nt][mt][3] *= f1;
            }
          }
        }
        uint32_t pb[NT][2];
#pragma unroll
        for (int nt = 0; nt < NT; nt++) {
          const float m0 = st.m[nt][0] == -INFINITY ? 0.f : st.m[nt][0];
          const float m1 = st.m[nt][1] == -INFINITY ? 0.f : st.m[nt][1];
          const float p0 = ex2(sc[nt][0] - m0), p1 = ex2(sc[nt][1] - m1);
          const float p2 = ex2(sc[nt][2] - m0), p3 = ex2(sc[nt][3] - m1);
          st.l[nt][0] += p0 + p2;
          st.l[nt][1] += p1 + p3;
          // P[tok][head] 8x8 tiles -> transposed = B operand (k = token, n = head)
          pb[nt][0] = movmatrix_trans(pack2<MMA_BF16>(p0, p1));
          pb[nt][1] = movmatrix_trans(pack2<MMA_BF16>(p2, p3));
        }
        // ------------------------------------------------ O^T += V^T . P^T
        const uint32_t vbase = tbase + TB;
        const int tk[4] = {2 * q4, 2 * q4 + 1, 2 * q4 + 8, 2 * q4 + 9};
        const bool vtail = tok0 + 16 > L;  // rows past the sequence end may hold NaN bit patterns
        if constexpr (G::VB >= 16) {
#pragma unroll
          for (int j = 0; j < G::NCV; j++) {
            uint4 v[4];
#pragma unroll
            for (int t = 0; t < 4; t++) {
              v[t] = lds128(vbase + (uint32_t)j * 16 * 128 + swz_off<128>(tk[t], r));
              if (vtail && tok0 + tk[t] >= L) v[t] = make_uint4(0, 0, 0, 0);
            }
#pragma unroll
            for (int w = 0; w < 4; w++) {
              const uint32_t x0 = (&v[0].x)[w], y0 = (&v[1].x)[w];
              const uint32_t x1 = (&v[2].x)[w], y1 = (&v[3].x)[w];
              if constexpr (FP8) {
                const uint32_t a01 = prmt(x0, y0, 0x5140), a23 = prmt(x0, y0, 0x7362);
                const uint32_t b01 = prmt(x1, y1, 0x5140), b23 = prmt(x1, y1, 0x7362);
                uint32_t f0[4], f1[4];
                f0[0] = e4m3x2_to_f16x2((uint16_t)(a01 & 0xffffu));
                f0[1] = e4m3x2_to_f16x2((uint16_t)(a01 >> 16));
                f0[2] = e4m3x2_to_f16x2((uint16_t)(b01 & 0xffffu));
                f0[3] = e4m3x2_to_f16x2((uint16_t)(b01 >> 16));
                f1[0] = e4m3x2_to_f16x2((uint16_t)(a23 & 0xffffu));
                f1[1] = e4m3x2_to_f16x2((uint16_t)(a23 >> 16));
                f1[2] = e4m3x2_to_f16x2((uint16_t)(b23 & 0xffffu));
                f1[3] = e4m3x2_to_f16x2((uint16_t)(b23 >> 16));
#pragma unroll
                for (int nt = 0; nt < NT; nt++) {
                  mma16816<MMA_BF16>(st.acc[nt][8 * j + 2 * w], f0, pb[nt]);
                  mma16816<MMA_BF16>(st.acc[nt][8 * j + 2 * w + 1], f1, pb[nt]);
                }
              } else {
                uint32_t f[4];
                f[0] = prmt(x0, y0, 0x5410);
                f[1] = prmt(x0, y0, 0x7632);
                f[2] = prmt(x1, y1, 0x5410);
                f[3] = prmt(x1, y1, 0x7632);
#pragma unroll
                for (int nt = 0; nt < NT; nt++) mma16816<MMA_BF16>(st.acc[nt][4 * j + w], f, pb[nt]);
              }
            }
          }
        } else {
          // VB == 8: 64-byte rows (IB == 64), each lane row owns 8 bytes
          uint2 v[4];
#pragma unroll
          for (int t = 0; t < 4; t++) {
            v[t] = lds64(vbase + swz_off<64>(tk[t], r >> 1) + (uint32_t)(r & 1) * 8u);
            if (vtail && tok0 + tk[t] >= L) v[t] = make_uint2(0, 0);
          }
#pragma unroll
          for (int w = 0; w < 2; w++) {
            const uint32_t x0 = (&v[0].x)[w], y0 = (&v[1].x)[w];
            const uint32_t x1 = (&v[2].x)[w], y1 = (&v[3].x)[w];
            if constexpr (FP8) {
              const uint32_t a01 = prmt(x0, y0, 0x5140), a23 = prmt(x0, y0, 0x7362);
              const uint32_t b01 = prmt(x1, y1, 0x5140), b23 = prmt(x1, y1, 0x7362);
              uint32_t f0[4], f1[4];
              f0[0] = e4m3x2_to_f16x2((uint16_t)(a01 & 0xffffu));
              f0[1] = e4m3x2_to_f16x2((uint16_t)(a01 >> 16));
              f0[2] = e4m3x2_to_f16x2((uint16_t)(b01 & 0xffffu));
              f0[3] = e4m3x2_to_f16x2((uint16_t)(b01 >> 16));
              f1[0] = e4m3x2_to_f16x2((uint16_t)(a23 & 0xffffu));
              f1[1] = e4m3x2_to_f16x2((uint16_t)(a23 >> 16));
              f1[2] = e4m3x2_to_f16x2((uint16_t)(b23 & 0xffffu));
              f1[3] = e4m3x2_to_f16x2((uint16_t)(b23 >> 16));
#pragma unroll
              for (int nt = 0; nt < NT; nt++) {
                mma16816<MMA_BF16>(st.acc[nt][2 * w], f0, pb[nt]);
                mma16816<MMA_BF16>(st.acc[nt][2 * w + 1], f1, pb[nt]);
              }
            } else {
              uint32_t f[4];
              f[0] = prmt(x0, y0, 0x5410);
              f[1] = prmt(x0, y0, 0x7632);
              f[2] = prmt(x1, y1, 0x5410);
              f[3] = prmt(x1, y1, 0x7632);
#pragma unroll
              for (int nt = 0; nt < NT; nt++) mma16816<MMA_BF16>(st.acc[nt][w], f, pb[nt]);
            }
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(bar_empty + 8u * stage);
      if (++stage == p.stages) { stage = 0; phase ^= 1u; }
    }

    // ---- row sums: lanes that share q4 hold the same q-heads
#pragma unroll
    for (int nt = 0; nt < NT; nt++)
#pragma unroll
      for (int a2 = 0; a2 < 2; a2++) {
        float v = st.l[nt][a2];
        v += __shfl_xor_sync(0xffffffffu, v, 4);
        v += __shfl_xor_sync(0xffffffffu, v, 8);
        v += __shfl_xor_sync(0xffffffffu, v, 16);
        st.l[nt][a2] = v;
      }

    // ---- merge the token groups of one head (only when tg > 1)
    if (tg > 1) {
      float* mbuf = reinterpret_cast<float*>(smem + p.off_merge);
      float* ml = mbuf;                                   // [(tg-1)*hc][NT*8][2]
      float4* ab = reinterpret_cast<float4*>(mbuf + (tg - 1) * p.hc * NT * 16);  // [..][NT*MT][32]
      if (active && tgi > 0) {
        const int wslot = (tgi - 1) * p.hc + hl;
        if (r == 0) {
#pragma unroll
          for (int nt = 0; nt < NT; nt++)
#pragma unroll
            for (int a2 = 0; a2 < 2; a2++) {
              ml[(wslot * NT * 8 + nt * 8 + 2 * q4 + a2) * 2 + 0] = st.m[nt][a2];
              ml[(wslot * NT * 8 + nt * 8 + 2 * q4 + a2) * 2 + 1] = st.l[nt][a2];
            }
        }
#pragma unroll
        for (int nt = 0; nt < NT; nt++)
#pragma unroll
          for (int mt = 0; mt < MT; mt++)
            ab[(wslot * NT * MT + nt * MT + mt) * 32 + lane] = make_float4(
                st.acc[nt][mt][0], st.acc[nt][mt][1], st.acc[nt][mt][2], st.acc[nt][mt][3]);
      }
      bar_sync_named(1, NW * 32);
      if (active && tgi == 0) {
        for (int t2 = 1; t2 < tg; t2++) {
          const int wslot = (t2 - 1) * p.hc + hl;
#pragma unroll
          for (int nt = 0; nt < NT; nt++) {
            float f1[2], f2[2];
#pragma unroll
            for (int a2 = 0; a2 < 2; a2++) {
              const float m2 = ml[(wslot * NT * 8 + nt * 8 + 2 * q4 + a2) * 2 + 0];
              const float l2 = ml[(wslot * NT * 8 + nt * 8 + 2 * q4 + a2) * 2 + 1];
              const float mn = fmaxf(st.m[nt][a2], m2);
              f1[a2] = mn == -INFINITY ? 1.f : ex2(st.m[nt][a2] - mn);
              f2[a2] = mn == -INFINITY ? 0.f : ex2(m2 - mn);
              st.m[nt][a2] = mn;
              st.l[nt][a2] = st.l[nt][a2] * f1[a2] + l2 * f2[a2];
            }
#pragma unroll
            for (int mt = 0; mt < MT; mt++) {
              const float4 o = ab[(wslot * NT * MT + nt * MT + mt) * 32 + lane];
              st.acc[nt][mt][0] = st.acc[nt][mt][0] * f1[0] + o.x * f2[0];
              st.acc[nt][mt][1] = st.acc[nt][mt][1] * f1[1] + o.y * f2[1];
              st.acc[nt][mt][2] = st.acc[nt][mt][2] * f1[0] + o.z * f2[0];
              st.acc[nt][mt][3] = st.acc[nt][mt][3] * f1[1] + o.w * f2[1];
            }
          }
        }
      }
      bar_sync_named(2, NW * 32);
    }

    // ---- write: final output (sequence wholly inside this CTA) or an fp32 partial
    if (active && tgi == 0) {
#pragma unroll
      for (int nt = 0; nt < NT; nt++)
#pragma unroll
        for (int a2 = 0; a2 < 2; a2++) {
          const int gq = nt * 8 + 2 * q4 + a2;
          if (gq >= g) continue;
          const int head = (h0 + hl) * g + gq;
          const float lsum = st.l[nt][a2];
          if (whole && p.out_scale) {
            // fused static FP8 output: e4m3(clamp((acc / l * v_scale) * (1 / out_scale)))
            const float inv = v_scale / lsum;
            const float qinv = __fdiv_rn(1.0f, *p.out_scale);
            uint8_t* op8 = reinterpret_cast<uint8_t*>(p.out) + (int64_t)s * p.out_stride_s +
                           (int64_t)head * p.out_stride_h;
            auto q2 = [&](float a, float b) -> uint16_t {
              return f32x2_to_e4m3x2(fmaxf(-448.f, fminf(a * inv * qinv, 448.f)),
                                     fmaxf(-448.f, fminf(b * inv * qinv, 448.f)));
            };
            if constexpr (G::VB >= 16) {
#pragma unroll
              for (int j = 0; j < G::NCV; j++) {
                uint16_t w8[EPC / 2];
#pragma unroll
                for (int e = 0; e < EPC; e += 2) {
                  const int i0 = j * EPC + e, i1 = i0 + 1;
                  w8[e / 2] = q2(st.acc[nt][i0 >> 1][(i0 & 1) * 2 + a2], st.acc[nt][i1 >> 1][(i1 & 1) * 2 + a2]);
                }
                uint8_t* dst = op8 + (r + 8 * j) * EPC;
#pragma unroll
                for (int c = 0; c < EPC / 4; c++)
                  *reinterpret_cast<uint32_t*>(dst + 4 * c) = (uint32_t)w8[2 * c] | ((uint32_t)w8[2 * c + 1] << 16);
              }
            } else {
              constexpr int NE = 8 / ES;
              uint8_t* dst = op8 + r * NE;
#pragma unroll
              for (int e = 0; e < NE; e += 2)
                *reinterpret_cast<uint16_t*>(dst + e) =
                    q2(st.acc[nt][e >> 1][(e & 1) * 2 + a2], st.acc[nt][(e + 1) >> 1][((e + 1) & 1) * 2 + a2]);
            }
            if (p.lse && r == 0)
              p.lse[(int64_t)s * p.Hq + head] = (st.m[nt][a2] + lg2(lsum)) * LN2;
          } else if (whole) {
            const float inv = v_scale / lsum;
            uint16_t* op = reinterpret_cast<uint16_t*>(p.out) + (int64_t)s * p.out_stride_s +
                           (int64_t)head * p.out_stride_h;
            if constexpr (G::VB >= 16) {
#pragma unroll
              for (int j = 0; j < G::NCV; j++) {
                uint32_t w[EPC / 2];
#pragma unroll
                for (int e = 0; e < EPC; e += 2) {
                  const int i0 = j * EPC + e, i1 = i0 + 1;
                  w[e / 2] = pack2<QBF16>(st.acc[nt][i0 >> 1][(i0 & 1) * 2 + a2] * inv,
                                          st.acc[nt][i1 >> 1][(i1 & 1) * 2 + a2] * inv);
                }
                uint16_t* dst = op + (r + 8 * j) * EPC;
#pragma unroll
                for (int c = 0; c < EPC / 8; c++)
                  *reinterpret_cast<uint4*>(dst + 8 * c) =
                      make_uint4(w[4 * c], w[4 * c + 1], w[4 * c + 2], w[4 * c + 3]);
              }
            } else {
              constexpr int NE = 8 / ES;  // elements per lane row
              uint32_t w[NE / 2];
#pragma unroll
              for (int e = 0; e < NE; e += 2)
                w[e / 2] = pack2<QBF16>(st.acc[nt][e >> 1][(e & 1) * 2 + a2] * inv,
                                        st.acc[nt][(e + 1) >> 1][((e + 1) & 1) * 2 + a2] * inv);
              uint16_t* dst = op + r * NE;
#pragma unroll
              for (int c = 0; c < NE / 2; c++) reinterpret_cast<uint32_t*>(dst)[c] = w[c];
            }
            if (p.lse && r == 0)
              p.lse[(int64_t)s * p.Hq + head] = (st.m[nt][a2] + lg2(lsum)) * LN2;
          } else {
            const int64_t pslot = ((int64_t)(blockIdx.y * gridDim.x + blockIdx.x)) * 2 + slot;
            const int hloc = hl * g + gq;
            float* pa = p.part_acc + (pslot * hq_chunk + hloc) * D;
            if constexpr (G::VB >= 16) {
#pragma unroll
              for (int j = 0; j < G::NCV; j++) {
                float* dst = pa + (r + 8 * j) * EPC;
#pragma unroll
                for (int e = 0; e < EPC; e += 4) {
                  float o[4];
#pragma unroll
                  for (int k = 0; k < 4; k++) {
                    const int i0 = j * EPC + e + k;
                    o[k] = st.acc[nt][i0 >> 1][(i0 & 1) * 2 + a2];
                  }
                  *reinterpret_cast<float4*>(dst + e) = make_float4(o[0], o[1], o[2], o[3]);
                }
              }
            } else {
              constexpr int NE = 8 / ES;
              float* dst = pa + r * NE;
#pragma unroll
              for (int e = 0; e < NE; e++) dst[e] = st.acc[nt][e >> 1][(e & 1) * 2 + a2];
            }
            if (r == 0) {
              float* pm = p.part_ml + (pslot * hq_chunk + hloc) * 2;
              pm[0] = st.m[nt][a2];
              pm[1] = lsum;
            }
          }
        }
    }
    pos = seg_end;
    s++;
  }
}

// Merge the partials of sequences that were cut by CTA boundaries; zero-fill empty sequences.
// grid (S, n_chunks); each thread owns 4 consecutive d of one q-head (float4 partial loads).
template <bool QBF16>
__global__ void __launch_bounds__(1024)
    paged_attn_reduce_kernel(const Params p, int D, int grid_x) {
  pdl_launch_dependents();
  pdl_wait();   // the partials and the tile prefix sums are the decode kernel's output
  const int s = blockIdx.x;
  const int y = blockIdx.y;
  const int g = p.g;
  const int hq_chunk = p.hc * g;
  const int c0 = p.cum[s], c1 = p.cum[s + 1];
  const int tpc = p.cum[p.S + 1];
  const float v_scale = (p.v_scale) ? *p.v_scale : 1.f;
  uint16_t* obase = reinterpret_cast<uint16_t*>(p.out) + (int64_t)s * p.out_stride_s;
  uint8_t* obase8 = reinterpret_cast<uint8_t*>(p.out) + (int64_t)s * p.out_stride_s;
  const bool fp8_out = p.out_scale != nullptr;
  const float qinv = fp8_out ? __fdiv_rn(1.0f, *p.out_scale) : 1.f;
  const int d4n = D >> 2;
  const int items = hq_chunk * d4n;
  if (c1 == c0) {
    // empty sequence: defined output (zeros), lse = -inf
    for (int i = threadIdx.x; i < items; i += blockDim.x) {
      const int h = i / d4n, d = (i - h * d4n) << 2;
      if (fp8_out)
        *reinterpret_cast<uint32_t*>(obase8 + (int64_t)(y * hq_chunk + h) * p.out_stride_h + d) = 0u;
      else
        *reinterpret_cast<uint2*>(obase + (int64_t)(y * hq_chunk + h) * p.out_stride_h + d) =
            make_uint2(0u, 0u);
    }
    if (p.lse)
      for (int h = threadIdx.x; h < hq_chunk; h += blockDim.x)
        p.lse[(int64_t)s * p.Hq + y * hq_chunk + h] = -INFINITY;
    return;
  }
  const int cta0 = c0 / tpc, cta1 = (c1 - 1) / tpc;
  if (cta0 == cta1) return;  // written directly by the attention kernel
  for (int i = threadIdx.x; i < items; i += blockDim.x) {
    const int h = i / d4n, d = (i - h * d4n) << 2;
    float mstar = -INFINITY;
    for (int c = cta0; c <= cta1; c++) {
      const int slot = (c * tpc >= c0) ? 0 : 1;
      const int64_t ps = ((int64_t)(y * grid_x + c)) * 2 + slot;
      mstar = fmaxf(mstar, __ldg(p.part_ml + (ps * hq_chunk + h) * 2));
    }
    float4 num = make_float4(0.f, 0.f, 0.f, 0.f);
    float den = 0.f;
    for (int c = cta0; c <= cta1; c++) {
      const int slot = (c * tpc >= c0) ? 0 : 1;
      const int64_t ps = ((int64_t)(y * grid_x + c)) * 2 + slot;
      const float2 ml = __ldg(reinterpret_cast<const float2*>(p.part_ml + (ps * hq_chunk + h) * 2));
      const float f = ex2(ml.x - mstar);
      const float4 a = __ldg(reinterpret_cast<const float4*>(p.part_acc + (ps * hq_chunk + h) * D + d));
      den += f * ml.y;
      num.x += f * a.x; num.y += f * a.y; num.z += f * a.z; num.w += f * a.w;
    }
    const float inv = v_scale / den;
    if (fp8_out) {
      const float a0 = fmaxf(-448.f, fminf(num.x * inv * qinv, 448.f)), a1 = fmaxf(-448.f, fminf(num.y * inv * qinv, 448.f));
      const float a2 = fmaxf(-448.f, fminf(num.z * inv * qinv, 448.f)), a3 = fmaxf(-448.f, fminf(num.w * inv * qinv, 448.f));
      *reinterpret_cast<uint32_t*>(obase8 + (int64_t)(y * hq_chunk + h) * p.out_stride_h + d) =
          (uint32_t)f32x2_to_e4m3x2(a0, a1) | ((uint32_t)f32x2_to_e4m3x2(a2, a3) << 16);
    } else {
      uint2 o;
      o.x = pack2<QBF16>(num.x * inv, num.y * inv);
      o.y = pack2<QBF16>(num.z * inv, num.w * inv);
      *reinterpret_cast<uint2*>(obase + (int64_t)(y * hq_chunk + h) * p.out_stride_h + d) = o;
    }
    if (p.lse && d == 0) p.lse[(int64_t)s * p.Hq + y * hq_chunk + h] = (mstar + lg2(den)) * LN2;
  }
}

}  // namespace attn
}  // namespace b200
