// b200ops -- decode-context-parallel (DCP) attention merge across ranks over NVLink peer memory
// (included by allreduce.cu: shares Peers / barrier / Pack with the all-reduce kernels).
//
// Contract: cp_lse_ag_out_rs / cp_lse_ag_out_ar (vllm/v1/attention/ops/common.py:185-258) =
//   lses = all_gather(lse)                                     [N, B, H]
//   out, lse = correct_attn_out(out, lses, rank)               (:110-183, Triton kernel :13-93)
//       l_r   = lses[r] with NaN / +inf -> -inf ; m = max_r l_r (-inf -> 0)
//       lse   = log(sum_r exp(l_r - m)) + m
//       out_r = T(out_r * exp(lses[rank] - lse))               (NaN / +inf exponent -> factor 0 -> output 0)
//   out = reduce_scatter(out, dim = heads)   (rs)   |   out = all_reduce(out)   (ar)
// i.e. every rank holds the attention of ALL heads over ITS shard of the context; the merged result for a head is
// the LSE-weighted sum over ranks, and with `rs` rank r keeps heads [r*H/N, (r+1)*H/N).
//
// B200 design: the reference runs all-gather(lse) -> Triton correction kernel -> reduce-scatter(out): three
// launches and two NCCL collectives on a [B, H, D] tensor of a few hundred KB (pure latency). Here it is ONE
// kernel: rendezvous, every rank pulls the N partial (out, lse) pairs of the heads it owns straight from the peers'
// symmetric buffers (16-byte loads, all issued before the first use), applies the correction and sums in rank
// order with fp32 accumulation (bit-identical on every rank for `ar`), rendezvous. The intermediate rounding of
// the corrected partial to T is kept (the reference stores it in T before the collective).
#pragma once

namespace b200 {
namespace ar {

struct DcpParams {
  const float* lse[MAX_RANKS];  // [B, H] f32 of every rank (symmetric heap)
  void* out;                    // local [B, Hs, D] T
  float* lse_out;               // local [B, Hs] f32 or null
  int B, H, D, head_lo, Hs;     // this rank merges heads [head_lo, head_lo + Hs)
  int base_e;                   // 1: natural-log LSEs, 0: log2
};

__device__ __forceinline__ float dcp_exp(float x, int base_e) { return base_e ? expf(x) : exp2f(x); }
__device__ __forceinline__ float dcp_log(float x, int base_e) { return base_e ? logf(x) : log2f(x); }

template <typename T, int N>
__global__ void __launch_bounds__(256, 1) dcp_lse_merge_kernel(Peers pe, DcpParams p, int rank) {
  constexpr int NP = 16 / sizeof(T);
  uint32_t epoch = pe.sig[rank]->epoch[blockIdx.x];
  barrier<N, false>(pe, rank, epoch);   // the partials were written by an earlier kernel on every rank
  const int ppr = p.D / NP;             // packs per (token, head) row
  const int64_t total = (int64_t)p.B * p.Hs * ppr;
  const float ninf = -INFINITY;
  for (int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t)gridDim.x * blockDim.x) {
    const int v = (int)(i % ppr);
    const int64_t bh = i / ppr;
    const int hl = (int)(bh % p.Hs);
    const int64_t b = bh / p.Hs;
    const int64_t row = b * p.H + p.head_lo + hl;   // (token, global head)
    uint4 x[N];
    float l[N];
#pragma unroll
    for (int r = 0; r < N; r++) {
      x[r] = ld_peer_v4(pe.in[r] + (row * ppr + v) * 16);
      asm volatile("ld.volatile.global.f32 %0, [%1];" : "=f"(l[r]) : "l"(p.lse[r] + row) : "memory");
    }
    float m = ninf;
#pragma unroll
    for (int r = 0; r < N; r++) {
      const float t = (l[r] != l[r] || l[r] == INFINITY) ? ninf : l[r];
      m = fmaxf(m, t);
    }
    if (m == ninf) m = 0.f;
    float acc = 0.f;
#pragma unroll
    for (int r = 0; r < N; r++) {
      const float t = (l[r] != l[r] || l[r] == INFINITY) ? ninf : l[r];
      acc += dcp_exp(t - m, p.base_e);
    }
    const float lse = dcp_log(acc, p.base_e) + m;
    float a[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
#pragma unroll
    for (int r = 0; r < N; r++) {
      float d = l[r] - lse;
      if (d != d || d == INFINITY) d = ninf;
      const float f = dcp_exp(d, p.base_e);
      const T* xe = reinterpret_cast<const T*>(&x[r]);
#pragma unroll
      for (int j = 0; j < NP; j++) {
        // corrected partial rounded to T first (the reference stores it in T before the collective)
        const float c = f == 0.f ? 0.f : Cvt<T>::to_f32(Cvt<T>::from_f32(Cvt<T>::to_f32(xe[j]) * f));
        a[j] += c;
      }
    }
    uint4 o;
    T* oe = reinterpret_cast<T*>(&o);
#pragma unroll
    for (int j = 0; j < NP; j++) oe[j] = Cvt<T>::from_f32(a[j]);
    reinterpret_cast<uint4*>(p.out)[(b * p.Hs + hl) * ppr + v] = o;
    if (p.lse_out && v == 0) p.lse_out[b * p.Hs + hl] = lse;
  }
  barrier<N, false>(pe, rank, epoch);   // peers may overwrite their partials only after everyone has read them
  if (threadIdx.x == 0) pe.sig[rank]->epoch[blockIdx.x] = epoch;
}

}  // namespace ar
}  // namespace b200
