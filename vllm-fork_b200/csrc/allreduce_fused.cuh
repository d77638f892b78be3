// b200ops -- fused tensor-parallel epilogue (included by allreduce.cu):
//   reduce-scatter(x) -> z = T(T(sum_r x_r) + residual) -> residual = z -> RMSNorm(z) * w
//   -> static FP8 quant (or T output) -> all-gather, in ONE kernel over NVLink peer memory.
//
// Replaces the sequence  tensor_model_parallel_all_reduce (communication_op.py:12) ->
// fused_add_rms_norm_static_fp8_quant (layernorm_quant_kernels.cu:81-156)  that follows every
// row-parallel GEMM (linear.py:1766-1767); the reference fuses the same pair with its
// fuse_allreduce_rms pass (docs/design/fusions.md:23) through FlashInfer. Arithmetic is kept
// IDENTICAL to the unfused pair (the all-reduce result is rounded to T first, the norm uses the same
// 256-thread reduction tree as rms_norm_kernel in norm_rope.cu), so fused == unfused bit for bit.
//
// Sequence-parallel formulation: rank r owns rows [r*Mp, (r+1)*Mp) of the [M, hidden] activation and
// keeps ONLY those rows of the residual stream. It pulls its rows from all N ranks (reduce-scatter),
// normalises/quantises them, and PUSHES the finished rows into every rank's output buffer (posted
// NVLink stores), so the all-gather moves 1 byte per element instead of 2. Two barriers.
#pragma once

namespace b200 {
namespace ar {

struct FusedParams {
  uint8_t* mc_out;           // MC kernels: multicast address of the output buffer (one store reaches every rank)
  uint8_t* out[MAX_RANKS];   // output buffer of every rank ([M, hidden] e4m3 or T), symmetric heap
  void* residual;            // local [Mp, hidden] T: rows of this rank only (nullable: no add)
  const void* weight;        // [hidden] T
  const float* scale;        // static FP8 scale (QUANT) or unused
  float eps;
  int M, hidden, rows_per_rank;
};

// MC = true: the reduce leg is one multimem.ld_reduce per 16 bytes (sum over all ranks inside the NVSwitch, fp32
// accumulation, rounded to T - the same rounding point as the peer-pull sum) and the gather leg one multimem.st
// per finished 8 / 16 bytes instead of N posted stores. Per-rank NVLink traffic for [M, hidden]: M/N rows in
// (once, not N-1 times) and M/N FP8 rows out (once, not N-1 times).
template <typename T, int N, bool QUANT, int R, bool MC = false>
__global__ void __launch_bounds__(256, 1)
    allreduce_fused_norm_kernel(Peers pe, FusedParams fp, const uint8_t* stage_src, int rank) {
  constexpr int NP = 16 / sizeof(T);
  __shared__ float red[32];
  if (pe.pdl_mode == 0) pdl_launch_dependents();
  pdl_wait();   // the GEMM output being reduced comes from the previous kernel on the stream
  uint32_t epoch = pe.sig[rank]->epoch[blockIdx.x];
  const int nv = fp.hidden / NP;            // packs per row
  stage_rows<N>(pe, rank, stage_src, fp.M, nv, fp.rows_per_rank);
  if (stage_src) barrier<N, true>(pe, rank, epoch);
  else barrier<N, false>(pe, rank, epoch);
  const int row_lo = rank * fp.rows_per_rank;
  const int row_hi = min(fp.M, row_lo + fp.rows_per_rank);
  float inv = 1.f;
  if constexpr (QUANT) inv = __fdiv_rn(1.0f, *fp.scale);
  for (int row = row_lo + blockIdx.x; row < row_hi; row += gridDim.x) {
    T* res_row = fp.residual ? reinterpret_cast<T*>(fp.residual) + (int64_t)(row - row_lo) * fp.hidden : nullptr;
    uint4 regs[R];
    float ss = 0.f;
    // issue every peer load of the row before the first add when the registers allow it: one NVLink
    // round trip per row instead of R
    constexpr bool BATCH = !MC && R * N <= 16;
    uint4 pv[BATCH ? R : 1][MC ? 1 : N];
    uint4 mv[MC ? R : 1];
    if constexpr (MC) {   // all in-switch reductions of the row in flight before the first use
#pragma unroll
      for (int i = 0; i < R; i++) {
        const int v = threadIdx.x + i * 256;
        if (v < nv) mv[i] = McRed<T>::ld(pe.mc_in + ((int64_t)row * nv + v) * 16);
      }
    }
    if constexpr (BATCH) {
#pragma unroll
      for (int i = 0; i < R; i++) {
        const int v = threadIdx.x + i * 256;
        if (v < nv) {
#pragma unroll
          for (int r = 0; r < N; r++) pv[i][r] = ld_peer_v4(pe.in[r] + ((int64_t)row * nv + v) * 16);
        }
      }
    }
#pragma unroll
    for (int i = 0; i < R; i++) {
      const int v = threadIdx.x + i * 256;
      if (v < nv) {
        uint4 x;   // T(sum over ranks), rank order 0..N-1 (MC: the switch's order)
        if constexpr (MC) {
          x = mv[i];
        } else if constexpr (BATCH) {
          float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int r = 0; r < N; r++) Pack<T>::add(a, pv[i][r]);
          x = Pack<T>::pack(a);
        } else {
          x = reduce_pack<T, N>(pe, (int64_t)row * nv + v);
        }
        if (res_row) {
          const uint4 r = *reinterpret_cast<const uint4*>(res_row + v * NP);
          T* xe = reinterpret_cast<T*>(&x);
          const T* re = reinterpret_cast<const T*>(&r);
#pragma unroll
          for (int j = 0; j < NP; j++) xe[j] = Cvt<T>::from_f32(Cvt<T>::to_f32(xe[j]) + Cvt<T>::to_f32(re[j]));
          *reinterpret_cast<uint4*>(res_row + v * NP) = x;
        }
        const T* xe = reinterpret_cast<const T*>(&x);
#pragma unroll
        for (int j = 0; j < NP; j++) {
          const float f = Cvt<T>::to_f32(xe[j]);
          ss += f * f;
        }
        regs[i] = x;
      }
    }
    ss = block_sum(ss, red);
    const float s = rsqrtf(ss / (float)fp.hidden + fp.eps);
#pragma unroll
    for (int i = 0; i < R; i++) {
      const int v = threadIdx.x + i * 256;
      if (v < nv) {
        const uint4 w = *reinterpret_cast<const uint4*>(reinterpret_cast<const T*>(fp.weight) + v * NP);
        const T* xe = reinterpret_cast<const T*>(&regs[i]);
        const T* we = reinterpret_cast<const T*>(&w);
        if constexpr (QUANT) {
          uint16_t o[NP / 2];
#pragma unroll
          for (int j = 0; j < NP / 2; j++) {
            float a = Cvt<T>::to_f32(Cvt<T>::from_f32(Cvt<T>::to_f32(xe[2 * j]) * s * Cvt<T>::to_f32(we[2 * j])));
            float b = Cvt<T>::to_f32(Cvt<T>::from_f32(Cvt<T>::to_f32(xe[2 * j + 1]) * s * Cvt<T>::to_f32(we[2 * j + 1])));
            a = fmaxf(-448.f, fminf(__fmul_rn(a, inv), 448.f));
            b = fmaxf(-448.f, fminf(__fmul_rn(b, inv), 448.f));
            o[j] = f32x2_to_e4m3x2(a, b);
          }
          const int64_t off = (int64_t)row * fp.hidden + (int64_t)v * NP;  // bytes (1 B / element)
          if constexpr (MC) {
            if constexpr (NP == 8) {
              mc_st_v2(fp.mc_out + off, *reinterpret_cast<const uint2*>(o));
            }
          } else {
#pragma unroll
            for (int r = 0; r < N; r++) {
              uint8_t* dst = fp.out[(rank + r) % N] + off;
              *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(o);
            }
          }
        } else {
          uint4 rr;
          T* re = reinterpret_cast<T*>(&rr);
#pragma unroll
          for (int j = 0; j < NP; j++)
            re[j] = Cvt<T>::from_f32(Cvt<T>::to_f32(xe[j]) * s * Cvt<T>::to_f32(we[j]));
          const int64_t off = ((int64_t)row * fp.hidden + (int64_t)v * NP) * sizeof(T);
          if constexpr (MC) {
            mc_st_v4(fp.mc_out + off, rr);
          } else {
#pragma unroll
            for (int r = 0; r < N; r++) *reinterpret_cast<uint4*>(fp.out[(rank + r) % N] + off) = rr;
          }
        }
      }
    }
  }
  barrier<N, true>(pe, rank, epoch);  // every rank's pushes have landed; inputs may be overwritten
  if (threadIdx.x == 0) pe.sig[rank]->epoch[blockIdx.x] = epoch;
  if (pe.pdl_mode == 1) pdl_launch_dependents();
}

}  // namespace ar
}  // namespace b200
