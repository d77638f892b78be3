// b200ops -- legacy-layout paged attention: torch.ops._C.paged_attention_v1 / paged_attention_v2.
//
// Contract (the op is absent from the reference tree and lives only in the wheel; semantics from
// tests/kernels/attention/test_attention.py:57-109 ref_single_query_cached_kv_attention,
// docs/design/paged_attention.md:45-139 and the python wrapper vllm/v1/attention/ops/paged_attn.py:15-51;
// schema SURVEY.md section 8b):
//   key_cache   [num_blocks, num_kv_heads, head_size/x, block_size, x]   x = 16 / sizeof(cache element)
//   value_cache [num_blocks, num_kv_heads, head_size, block_size]
//   out[s,h,:] = softmax_t( scale * q[s,h,:].K[t] + alibi[h]*(t - L_s + 1) ) . V,  t in [0, L_s)
//   FP8 KV: element = float(e4m3 byte) * k_scale / v_scale.
//   v2: 512-token partitions write max_logits / exp_sums / tmp_out [S, Hq, P(, D)] (natural-log
//   domain, tmp_out normalised by its partition's exp_sum + 1e-6), then a reduce pass combines them.
//
// B200 design. This layout is the OLD one (the flash layout + TMA kernel in attention.cu is the
// headline path); it cannot be described as one TMA box per tile, so this kernel uses plain 16-byte
// vector loads, arranged so that they coalesce: for K a warp reads, per 16-byte d-chunk, the
// `block_size` tokens of a block = one contiguous run of block_size*16 bytes; for V a lane owns whole
// head-dim rows (block_size contiguous tokens each), consecutive lanes = consecutive rows.
// One CTA per (partition, kv head x query-head chunk, sequence): all <= 8 query heads of a GQA group
// share every K/V byte (the legacy kernel re-read K/V once per query head). fp32 FMA math, exp2-domain
// online softmax per warp, 4 warps merged through shared memory.
#include "common.cuh"

namespace b200 {
namespace attn_legacy {

constexpr int NWARPS = 4;
constexpr int NTHREADS = NWARPS * 32;
constexpr int MAX_ROWS = 8;  // head-dim rows per lane: head_size <= 256
constexpr float LOG2E = 1.4426950408889634f;
constexpr float LN2 = 0.6931471805599453f;

struct Params {
  void* out;            // [S, Hq, D]
  float* exp_sums;      // [S, Hq, P]      (v2)
  float* max_logits;    // [S, Hq, P]      (v2)
  void* tmp_out;        // [S, Hq, P, D]   (v2)
  const void* q;
  const void* k_cache;
  const void* v_cache;
  const int32_t* block_tables;
  const int32_t* seq_lens;
  const float* alibi_slopes;
  const float* k_scale;
  const float* v_scale;
  int64_t q_stride, kv_block_stride, kv_head_stride, bt_stride;
  int num_heads, num_kv_heads, head_size, block_size;
  int partition_size;   // 0 = v1 (whole sequence in one CTA)
  int max_partitions;
  int q_chunks;         // ceil(group / GQ)
  int io_bf16;
  float scale;
};

template <typename T> struct KvLoad;
// 16-bit caches: 8 elements per 16 bytes
template <> struct KvLoad<__nv_bfloat16> {
  static constexpr int X = 8;
  static __device__ __forceinline__ void cvt(const uint4& v, float (&f)[16]) {
    const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&v);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float2 t = __bfloat1622float2(h[i]);
      f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
  }
};
template <> struct KvLoad<__half> {
  static constexpr int X = 8;
  static __device__ __forceinline__ void cvt(const uint4& v, float (&f)[16]) {
    const __half2* h = reinterpret_cast<const __half2*>(&v);
#pragma unroll
    for (int i = 0; i < 4; i++) {
      float2 t = __half22float2(h[i]);
      f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
  }
};
// fp8 e4m3 cache: 16 elements per 16 bytes
template <> struct KvLoad<uint8_t> {
  static constexpr int X = 16;
  static __device__ __forceinline__ void cvt(const uint4& v, float (&f)[16]) {
    const uint16_t* h = reinterpret_cast<const uint16_t*>(&v);
#pragma unroll
    for (int i = 0; i < 8; i++) {
      const uint32_t p = e4m3x2_to_f16x2(h[i]);
      float2 t = __half22float2(*reinterpret_cast<const __half2*>(&p));
      f[2 * i] = t.x; f[2 * i + 1] = t.y;
    }
  }
};

template <typename KV_T, int GQ, int ROWS>
__global__ void __launch_bounds__(NTHREADS)
    paged_attn_legacy_kernel(const Params p) {
  constexpr int X = KvLoad<KV_T>::X;
  extern __shared__ float smem_f[];
  const int D = p.head_size;
  float* q_s = smem_f;                         // [GQ][D]  (scale * log2e folded in)
  float* acc_s = q_s + GQ * D;                 // [NWARPS][GQ][D]
  float* ml_s = acc_s + NWARPS * GQ * D;       // [NWARPS][GQ][2]
  float* p_s = ml_s + NWARPS * GQ * 2;         // [NWARPS][GQ][32] probabilities of the warp's block

  const int part = blockIdx.x;
  const int kvh = blockIdx.y / p.q_chunks;
  const int qc = blockIdx.y % p.q_chunks;
  const int seq = blockIdx.z;
  const int group = p.num_heads / p.num_kv_heads;
  const int h0 = kvh * group + qc * GQ;              // first query head of this CTA
  const int nq = min(GQ, group - qc * GQ);           // valid query heads
  const int L = p.seq_lens[seq];
  const int bs = p.block_size;
  const int start_tok = p.partition_size ? part * p.partition_size : 0;
  if (p.partition_size && start_tok >= L) return;    // partition beyond the sequence: never read
  const int end_tok = p.partition_size ? min(L, start_tok + p.partition_size) : L;
  const int start_blk = start_tok / bs;
  const int end_blk = (end_tok + bs - 1) / bs;

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  // ---- q -> smem (fp32, pre-multiplied by scale*log2e; k_scale is applied to the scores)
  for (int i = threadIdx.x; i < GQ * D; i += NTHREADS) {
    const int g = i / D, d = i - g * D;
    float v = 0.f;
    if (g < nq) {
      const int64_t off = (int64_t)seq * p.q_stride + (int64_t)(h0 + g) * D + d;
      v = p.io_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.q)[off])
                    : __half2float(reinterpret_cast<const __half*>(p.q)[off]);
    }
    q_s[i] = v * p.scale * LOG2E;
  }
  __syncthreads();

  const float ks = p.k_scale ? p.k_scale[0] : 1.f;
  const int splits = 32 / bs;            // lanes per token along d-chunks (bs in {8,16,32})
  const int token = lane % bs;
  const int dsplit = lane / bs;
  const int n_chunks = D / X;
  float* pw = p_s + warp * GQ * 32;
  float alibi[GQ];
#pragma unroll
  for (int g = 0; g < GQ; g++)
    alibi[g] = (p.alibi_slopes && g < nq) ? p.alibi_slopes[h0 + g] * LOG2E : 0.f;

  float m[GQ], l[GQ], acc[GQ][ROWS];   // V rows owned by a lane: lane + 32*i
#pragma unroll
  for (int g = 0; g < GQ; g++) {
    m[g] = -INFINITY; l[g] = 0.f;
#pragma unroll
    for (int i = 0; i < ROWS; i++) acc[g][i] = 0.f;
  }
  const KV_T* kc = reinterpret_cast<const KV_T*>(p.k_cache);
  const KV_T* vc = reinterpret_cast<const KV_T*>(p.v_cache);

  for (int blk = start_blk + warp; blk < end_blk; blk += NWARPS) {
    const int64_t phys = p.block_tables[(int64_t)seq * p.bt_stride + blk];
    const KV_T* kb = kc + phys * p.kv_block_stride + (int64_t)kvh * p.kv_head_stride;
    const KV_T* vb = vc + phys * p.kv_block_stride + (int64_t)kvh * p.kv_head_stride;
    // ---------------- scores: s[g] = q[g] . k[token]
    float s[GQ];
#pragma unroll
    for (int g = 0; g < GQ; g++) s[g] = 0.f;
    for (int c = dsplit; c < n_chunks; c += splits) {
      const uint4 raw = ld_nc_v4(kb + ((int64_t)c * bs + token) * X);
      float kf[16];
      KvLoad<KV_T>::cvt(raw, kf);
#pragma unroll
      for (int g = 0; g < GQ; g++) {
        const float4* qv = reinterpret_cast<const float4*>(q_s + g * D + c * X);
#pragma unroll
        for (int j = 0; j < X / 4; j++) {
          const float4 t = qv[j];
          s[g] += t.x * kf[4 * j] + t.y * kf[4 * j + 1] + t.z * kf[4 * j + 2] + t.w * kf[4 * j + 3];
        }
      }
    }
    const int tok = blk * bs + token;
    const bool valid = tok < end_tok;
    __syncwarp();  // the previous block's reads of pw are done
#pragma unroll
    for (int g = 0; g < GQ; g++) {
      for (int o = bs; o < 32; o <<= 1) s[g] += __shfl_xor_sync(0xffffffffu, s[g], o);
      s[g] = valid ? s[g] * ks + alibi[g] * (float)(tok - L + 1) : -INFINITY;
      const float mb = warp_max(s[g]);
      const float mn = fmaxf(m[g], mb);       // finite: the block has >= 1 valid token
      const float corr = exp2f(m[g] - mn);    // m = -inf -> 0
      const float pr = valid ? exp2f(s[g] - mn) : 0.f;
      if (dsplit == 0) pw[g * 32 + token] = pr;
      const float psum = warp_sum(dsplit == 0 ? pr : 0.f);
      l[g] = l[g] * corr + psum;
      m[g] = mn;
#pragma unroll
      for (int i = 0; i < ROWS; i++) acc[g][i] *= corr;
    }
    __syncwarp();
    // ---------------- acc[g][row] += sum_t p[g][t] * v[row][t]
    for (int t0 = 0; t0 < bs; t0 += X) {
#pragma unroll
      for (int i = 0; i < ROWS; i++) {
        const int row = lane + 32 * i;
        if (row < D) {
          const uint4 raw = ld_nc_v4(vb + (int64_t)row * bs + t0);
          float vf[16];
          KvLoad<KV_T>::cvt(raw, vf);
#pragma unroll
          for (int g = 0; g < GQ; g++) {
            const float4* pv = reinterpret_cast<const float4*>(pw + g * 32 + t0);  // warp-wide broadcast
            float a = 0.f;
#pragma unroll
            for (int j = 0; j < X / 4; j++) {
              const float4 t = pv[j];
              a += t.x * vf[4 * j] + t.y * vf[4 * j + 1] + t.z * vf[4 * j + 2] + t.w * vf[4 * j + 3];
            }
            acc[g][i] += a;
          }
        }
      }
    }
  }
  // ---------------- merge the warps
#pragma unroll
  for (int g = 0; g < GQ; g++) {
    if (lane == 0) {
      ml_s[(warp * GQ + g) * 2] = m[g];
      ml_s[(warp * GQ + g) * 2 + 1] = l[g];
    }
#pragma unroll
    for (int i = 0; i < ROWS; i++) {
      const int row = lane + 32 * i;
      if (row < D) acc_s[(warp * GQ + g) * D + row] = acc[g][i];
    }
  }
  __syncthreads();
  const float vs = p.v_scale ? p.v_scale[0] : 1.f;
  for (int idx = threadIdx.x; idx < nq * D; idx += NTHREADS) {
    const int g = idx / D, d = idx - g * D;
    float M = -INFINITY;
#pragma unroll
    for (int w = 0; w < NWARPS; w++) M = fmaxf(M, ml_s[(w * GQ + g) * 2]);
    float lsum = 0.f, o = 0.f;
#pragma unroll
    for (int w = 0; w < NWARPS; w++) {
      const float mw = ml_s[(w * GQ + g) * 2];
      const float f = mw == -INFINITY ? 0.f : exp2f(mw - M);
      lsum += ml_s[(w * GQ + g) * 2 + 1] * f;
      o += acc_s[(w * GQ + g) * D + d] * f;
    }
    const float res = o * vs * __fdividef(1.f, lsum + 1e-6f);
    const int h = h0 + g;
    if (p.partition_size == 0) {
      const int64_t off = ((int64_t)seq * p.num_heads + h) * D + d;
      if (p.io_bf16) reinterpret_cast<__nv_bfloat16*>(p.out)[off] = __float2bfloat16_rn(res);
      else reinterpret_cast<__half*>(p.out)[off] = __float2half_rn(res);
    } else {
      const int64_t po = ((int64_t)seq * p.num_heads + h) * p.max_partitions + part;
      const int64_t off = po * D + d;
      if (p.io_bf16) reinterpret_cast<__nv_bfloat16*>(p.tmp_out)[off] = __float2bfloat16_rn(res);
      else reinterpret_cast<__half*>(p.tmp_out)[off] = __float2half_rn(res);
      if (d == 0) {
        p.max_logits[po] = M * LN2;   // natural-log domain, like the legacy kernel
        p.exp_sums[po] = lsum;
      }
    }
  }
}

// v2 second pass: out[s,h,:] = sum_p tmp_out[s,h,p,:] * w_p / (sum_p w_p + 1e-6),
// w_p = exp_sums[p] * exp(max_logits[p] - max_p' max_logits[p'])
__global__ void __launch_bounds__(128)
    paged_attn_legacy_reduce_kernel(const Params p) {
  const int h = blockIdx.x, seq = blockIdx.y;
  const int D = p.head_size;
  const int L = p.seq_lens[seq];
  const int np = (L + p.partition_size - 1) / p.partition_size;
  const int64_t po = ((int64_t)seq * p.num_heads + h) * p.max_partitions;
  float M = -INFINITY;
  for (int i = 0; i < np; i++) M = fmaxf(M, p.max_logits[po + i]);
  float wsum = 0.f;
  for (int i = 0; i < np; i++) wsum += p.exp_sums[po + i] * __expf(p.max_logits[po + i] - M);
  const float inv = __fdividef(1.f, wsum + 1e-6f);
  for (int d = threadIdx.x; d < D; d += blockDim.x) {
    float o = 0.f;
    for (int i = 0; i < np; i++) {
      const float w = p.exp_sums[po + i] * __expf(p.max_logits[po + i] - M);
      const int64_t off = (po + i) * D + d;
      const float t = p.io_bf16 ? __bfloat162float(reinterpret_cast<const __nv_bfloat16*>(p.tmp_out)[off])
                                : __half2float(reinterpret_cast<const __half*>(p.tmp_out)[off]);
      o += t * w;
    }
    const float res = np > 0 ? o * inv : 0.f;
    const int64_t off = ((int64_t)seq * p.num_heads + h) * D + d;
    if (p.io_bf16) reinterpret_cast<__nv_bfloat16*>(p.out)[off] = __float2bfloat16_rn(res);
    else reinterpret_cast<__half*>(p.out)[off] = __float2half_rn(res);
  }
}

template <typename KV_T, int GQ, int ROWS>
static int launch_rows(const Params& p, int num_seqs, int parts, cudaStream_t stream) {
  const size_t smem = ((size_t)GQ * p.head_size * (1 + NWARPS) + NWARPS * GQ * (2 + 32)) * sizeof(float);
  if (smem > 48 * 1024)
    B200_CUDA(cudaFuncSetAttribute(paged_attn_legacy_kernel<KV_T, GQ, ROWS>,
                                   cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  dim3 grid(parts, p.num_kv_heads * p.q_chunks, num_seqs);
  paged_attn_legacy_kernel<KV_T, GQ, ROWS><<<grid, NTHREADS, smem, stream>>>(p);
  B200_LAUNCH_CHECK();
  return B200_OK;
}
template <typename KV_T, int GQ>
static int launch(const Params& p, int num_seqs, int parts, cudaStream_t stream) {
  if (p.head_size <= 64) return launch_rows<KV_T, GQ, 2>(p, num_seqs, parts, stream);
  if (p.head_size <= 128) return launch_rows<KV_T, GQ, 4>(p, num_seqs, parts, stream);
  return launch_rows<KV_T, GQ, 8>(p, num_seqs, parts, stream);
}

template <typename KV_T>
static int dispatch_gq(Params& p, int num_seqs, int parts, cudaStream_t stream) {
  const int group = p.num_heads / p.num_kv_heads;
  const int gq = group >= 8 ? 8 : group >= 4 ? 4 : group >= 2 ? 2 : 1;
  p.q_chunks = (group + gq - 1) / gq;
  switch (gq) {
    case 8: return launch<KV_T, 8>(p, num_seqs, parts, stream);
    case 4: return launch<KV_T, 4>(p, num_seqs, parts, stream);
    case 2: return launch<KV_T, 2>(p, num_seqs, parts, stream);
    default: return launch<KV_T, 1>(p, num_seqs, parts, stream);
  }
}

static int run(void* out, float* exp_sums, float* max_logits, void* tmp_out, const void* query,
               const void* key_cache, const void* value_cache, int num_seqs, int num_heads,
               int num_kv_heads, int head_size, float scale, const int32_t* block_tables,
               const int32_t* seq_lens, int block_size, int max_seq_len, const float* alibi_slopes,
               int q_dtype, int kv_dtype, const float* k_scale, const float* v_scale, int64_t q_stride,
               int64_t kv_block_stride, int64_t kv_head_stride, int64_t bt_stride, int max_partitions,
               int blocksparse_vert_stride, cudaStream_t stream, const char* op) {
  B200_CHECK(num_seqs >= 0 && num_heads > 0 && num_kv_heads > 0 && head_size > 0, B200_ERR_INVALID,
             "%s: bad sizes", op);
  if (num_seqs == 0) return B200_OK;
  B200_CHECK(out && query && key_cache && value_cache && block_tables && seq_lens, B200_ERR_INVALID,
             "%s: null pointer", op);
  B200_CHECK(num_heads % num_kv_heads == 0, B200_ERR_INVALID,
             "%s: num_heads must be a multiple of num_kv_heads", op);
  B200_CHECK(q_dtype == B200_BF16 || q_dtype == B200_F16, B200_ERR_INVALID,
             "%s: query dtype must be bf16 or f16", op);
  B200_CHECK(kv_dtype == q_dtype || kv_dtype == B200_FP8_E4M3 || kv_dtype == B200_U8, B200_ERR_INVALID,
             "%s: kv_cache_dtype must be auto or fp8 (e4m3)", op);
  const bool fp8 = kv_dtype != q_dtype;
  B200_CHECK(!fp8 || (k_scale && v_scale), B200_ERR_INVALID, "%s: fp8 KV cache needs k_scale / v_scale", op);
  B200_CHECK(block_size == 8 || block_size == 16 || block_size == 32, B200_ERR_UNSUPPORTED,
             "%s: Unsupported block size: %d", op, block_size);
  const int x = fp8 ? 16 : 8;
  B200_CHECK(block_size % x == 0, B200_ERR_UNSUPPORTED,
             "%s: block size %d is smaller than one 16-byte vector of this cache dtype", op, block_size);
  B200_CHECK(head_size % x == 0 && head_size <= 32 * MAX_ROWS, B200_ERR_UNSUPPORTED,
             "%s: Unsupported head size: %d", op, head_size);
  B200_CHECK(blocksparse_vert_stride <= 1, B200_ERR_UNSUPPORTED,
             "%s: block-sparse attention is not supported", op);
  B200_CHECK(((uintptr_t)key_cache & 15) == 0 && ((uintptr_t)value_cache & 15) == 0 &&
                 (kv_block_stride * (fp8 ? 1 : 2)) % 16 == 0 && (kv_head_stride * (fp8 ? 1 : 2)) % 16 == 0,
             B200_ERR_INVALID, "%s: KV cache must be 16-byte aligned", op);
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "%s: no CUDA device (no CPU fallback)", op);
  Params p;
  p.out = out; p.exp_sums = exp_sums; p.max_logits = max_logits; p.tmp_out = tmp_out;
  p.q = query; p.k_cache = key_cache; p.v_cache = value_cache;
  p.block_tables = block_tables; p.seq_lens = seq_lens; p.alibi_slopes = alibi_slopes;
  p.k_scale = fp8 ? k_scale : nullptr; p.v_scale = fp8 ? v_scale : nullptr;
  p.q_stride = q_stride; p.kv_block_stride = kv_block_stride; p.kv_head_stride = kv_head_stride;
  p.bt_stride = bt_stride;
  p.num_heads = num_heads; p.num_kv_heads = num_kv_heads; p.head_size = head_size;
  p.block_size = block_size;
  p.partition_size = max_partitions > 0 ? 512 : 0;
  p.max_partitions = max_partitions;
  p.q_chunks = 1;
  p.io_bf16 = q_dtype == B200_BF16;
  p.scale = scale;
  (void)max_seq_len;
  const int parts = max_partitions > 0 ? max_partitions : 1;
  int rc;
  if (fp8) rc = dispatch_gq<uint8_t>(p, num_seqs, parts, stream);
  else if (q_dtype == B200_BF16) rc = dispatch_gq<__nv_bfloat16>(p, num_seqs, parts, stream);
  else rc = dispatch_gq<__half>(p, num_seqs, parts, stream);
  if (rc) return rc;
  if (max_partitions > 0) {
    paged_attn_legacy_reduce_kernel<<<dim3(num_heads, num_seqs), 128, 0, stream>>>(p);
    B200_LAUNCH_CHECK();
  }
  return B200_OK;
}

}  // namespace attn_legacy
}  // namespace b200

extern "C" int b200_paged_attention_v1(void* out, const void* query, const void* key_cache,
                                       const void* value_cache, int num_seqs, int num_heads,
                                       int num_kv_heads, int head_size, float scale,
                                       const int32_t* block_tables, const int32_t* seq_lens,
                                       int block_size, int max_seq_len, const float* alibi_slopes,
                                       int q_dtype, int kv_dtype, const float* k_scale,
                                       const float* v_scale, int64_t q_stride, int64_t kv_block_stride,
                                       int64_t kv_head_stride, int64_t bt_stride,
                                       int blocksparse_vert_stride, b200_stream_t stream) {
  return b200::attn_legacy::run(out, nullptr, nullptr, nullptr, query, key_cache, value_cache, num_seqs,
                                num_heads, num_kv_heads, head_size, scale, block_tables, seq_lens,
                                block_size, max_seq_len, alibi_slopes, q_dtype, kv_dtype, k_scale, v_scale,
                                q_stride, kv_block_stride, kv_head_stride, bt_stride, 0,
                                blocksparse_vert_stride, (cudaStream_t)stream, "paged_attention_v1");
}

extern "C" int b200_paged_attention_v2(void* out, float* exp_sums, float* max_logits, void* tmp_out,
                                       const void* query, const void* key_cache,
                                       const void* value_cache, int num_seqs, int num_heads,
                                       int num_kv_heads, int head_size, float scale,
                                       const int32_t* block_tables, const int32_t* seq_lens,
                                       int block_size, int max_seq_len, const float* alibi_slopes,
                                       int q_dtype, int kv_dtype, const float* k_scale,
                                       const float* v_scale, int64_t q_stride, int64_t kv_block_stride,
                                       int64_t kv_head_stride, int64_t bt_stride, int max_num_partitions,
                                       int blocksparse_vert_stride, b200_stream_t stream) {
  using namespace b200;
  B200_CHECK(exp_sums && max_logits && tmp_out, B200_ERR_INVALID, "paged_attention_v2: null pointer");
  B200_CHECK(max_num_partitions > 0 && (int64_t)max_num_partitions * 512 >= max_seq_len, B200_ERR_INVALID,
             "paged_attention_v2: exp_sums/max_logits/tmp_out need >= ceil(max_seq_len / 512) partitions");
  return b200::attn_legacy::run(out, exp_sums, max_logits, tmp_out, query, key_cache, value_cache, num_seqs,
                                num_heads, num_kv_heads, head_size, scale, block_tables, seq_lens,
                                block_size, max_seq_len, alibi_slopes, q_dtype, kv_dtype, k_scale, v_scale,
                                q_stride, kv_block_stride, kv_head_stride, bt_stride, max_num_partitions,
                                blocksparse_vert_stride, (cudaStream_t)stream, "paged_attention_v2");
}
