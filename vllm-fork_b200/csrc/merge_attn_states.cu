// b200ops -- merge two partial attention results by their log-sum-exp (prefix = KV-cache context,
// suffix = new tokens), optionally writing the merged LSE and/or quantising the result to FP8.
//
// Contract: torch.ops._C.merge_attn_states (csrc/libtorch_stable/attention/merge_attn_states.cu:13-200
// kernel, host below it; python vllm/v1/attention/ops/merge_attn_states.py:9-60):
//   token >= prefill_tokens_with_context : out = suffix_out,            out_lse = suffix_lse
//   max(p_lse, s_lse) infinite           : out = prefix_out,            out_lse = max
//   else  w_p = exp(p_lse - max) / (exp(p_lse - max) + exp(s_lse - max)),  w_s likewise
//         out = prefix_out * w_p + suffix_out * w_s,   out_lse = log(sum) + max
//   (+inf LSEs are treated as -inf, like the reference.) FP8 output: e4m3(clamp(out * (1/scale))).
//   outputs [T, H, D], LSEs [H, T] with explicit strides.
//
// B200 design: pure bandwidth (3 x T*H*D elements). One thread per 16-byte input pack; the two LSE
// scalars of a (token, head) are read once per pack through the read-only path (L1-resident).
#include "common.cuh"

namespace b200 {

struct MergeParams {
  void* out;
  float* out_lse;
  const void* prefix_out;
  const float* prefix_lse;
  const void* suffix_out;
  const float* suffix_lse;
  const float* out_scale;
  int64_t in_tok_stride, in_head_stride;    // shared by prefix / suffix (reference asserts equal head strides)
  int64_t sfx_tok_stride;
  int64_t out_tok_stride, out_head_stride;
  int64_t plse_h, plse_t, slse_h, slse_t, olse_h, olse_t;
  int64_t total_packs;
  int num_tokens, num_heads, head_size, prefix_tokens;
};

template <typename T, bool FP8_OUT>
__global__ void __launch_bounds__(256) merge_attn_states_kernel(const MergeParams p) {
  constexpr int N = 16 / sizeof(T);
  const int packs_per_head = p.head_size / N;
  const float inv_scale = FP8_OUT ? __fdiv_rn(1.0f, *p.out_scale) : 1.f;
  for (int64_t idx = (int64_t)blockIdx.x * blockDim.x + threadIdx.x; idx < p.total_packs;
       idx += (int64_t)gridDim.x * blockDim.x) {
    const int pack = (int)(idx % packs_per_head);
    const int64_t th = idx / packs_per_head;
    const int head = (int)(th % p.num_heads);
    const int64_t tok = th / p.num_heads;
    const T* pp = reinterpret_cast<const T*>(p.prefix_out) + tok * p.in_tok_stride + head * p.in_head_stride + pack * N;
    const T* sp = reinterpret_cast<const T*>(p.suffix_out) + tok * p.sfx_tok_stride + head * p.in_head_stride + pack * N;
    float wp, ws, lse;
    bool take_p = false, take_s = false;
    if (tok >= p.prefix_tokens) {
      take_s = true;
      lse = p.out_lse ? p.suffix_lse[head * p.slse_h + tok * p.slse_t] : 0.f;
      wp = 0.f; ws = 1.f;
    } else {
      float pl = p.prefix_lse[head * p.plse_h + tok * p.plse_t];
      float sl = p.suffix_lse[head * p.slse_h + tok * p.slse_t];
      pl = isinf(pl) ? -INFINITY : pl;
      sl = isinf(sl) ? -INFINITY : sl;
      const float mx = fmaxf(pl, sl);
      if (isinf(mx)) {
        take_p = true;
        lse = mx;
        wp = 1.f; ws = 0.f;
      } else {
        const float pe = expf(pl - mx), se = expf(sl - mx);
        const float sum = pe + se;
        wp = pe / sum;
        ws = se / sum;
        lse = logf(sum) + mx;
      }
    }
    uint4 pv = make_uint4(0, 0, 0, 0), sv = make_uint4(0, 0, 0, 0);
    if (!take_s) pv = *reinterpret_cast<const uint4*>(pp);
    if (!take_p) sv = *reinterpret_cast<const uint4*>(sp);
    const T* pe_ = reinterpret_cast<const T*>(&pv);
    const T* se_ = reinterpret_cast<const T*>(&sv);
    float o[N];
#pragma unroll
    for (int i = 0; i < N; i++) {
      if (take_s) o[i] = Cvt<T>::to_f32(se_[i]);
      else if (take_p) o[i] = Cvt<T>::to_f32(pe_[i]);
      else o[i] = Cvt<T>::to_f32(pe_[i]) * wp + (Cvt<T>::to_f32(se_[i]) * ws);
    }
    if constexpr (FP8_OUT) {
      uint8_t* dst = reinterpret_cast<uint8_t*>(p.out) + tok * p.out_tok_stride + head * p.out_head_stride + pack * N;
      uint8_t b[N];
#pragma unroll
      for (int i = 0; i < N; i++) b[i] = f32_to_e4m3(fmaxf(-448.f, fminf(o[i] * inv_scale, 448.f)));
      if constexpr (N == 8) *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(b);
      else *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(b);
    } else {
      T* dst = reinterpret_cast<T*>(p.out) + tok * p.out_tok_stride + head * p.out_head_stride + pack * N;
      uint4 r;
      T* re = reinterpret_cast<T*>(&r);
      if (take_s) r = sv;
      else if (take_p) r = pv;
      else {
#pragma unroll
        for (int i = 0; i < N; i++) re[i] = Cvt<T>::from_f32(o[i]);
      }
      *reinterpret_cast<uint4*>(dst) = r;
    }
    if (p.out_lse && pack == 0) p.out_lse[head * p.olse_h + tok * p.olse_t] = lse;
  }
}

}  // namespace b200

using namespace b200;

extern "C" int b200_merge_attn_states(void* output, float* output_lse, const void* prefix_output,
                                      const float* prefix_lse, const void* suffix_output,
                                      const float* suffix_lse, const float* output_scale,
                                      int64_t num_tokens, int num_heads, int head_size,
                                      int64_t prefix_tok_stride, int64_t suffix_tok_stride,
                                      int64_t in_head_stride, int64_t out_tok_stride,
                                      int64_t out_head_stride, int64_t prefix_lse_head_stride,
                                      int64_t prefix_lse_tok_stride, int64_t suffix_lse_head_stride,
                                      int64_t suffix_lse_tok_stride, int64_t out_lse_head_stride,
                                      int64_t out_lse_tok_stride, int64_t prefill_tokens_with_context,
                                      int in_dtype, int out_dtype, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(num_tokens >= 0 && num_heads > 0 && head_size > 0, B200_ERR_INVALID,
             "merge_attn_states: bad sizes");
  B200_CHECK(in_dtype == B200_F32 || in_dtype == B200_F16 || in_dtype == B200_BF16, B200_ERR_INVALID,
             "merge_attn_states: inputs must be float32 / float16 / bfloat16");
  const bool fp8_out = out_dtype == B200_FP8_E4M3;
  B200_CHECK(fp8_out || out_dtype == in_dtype, B200_ERR_INVALID,
             "merge_attn_states: output dtype must match the inputs or be FP8");
  B200_CHECK(!fp8_out || output_scale, B200_ERR_INVALID,
             "merge_attn_states: output_scale is required when output is FP8");
  B200_CHECK(fp8_out || !output_scale, B200_ERR_INVALID,
             "merge_attn_states: output must be FP8 when output_scale is provided");
  const int pack = 16 / dtype_size(in_dtype);
  B200_CHECK(head_size % pack == 0, B200_ERR_INVALID,
             "merge_attn_states: headsize must be multiple of pack_size: %d", pack);
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(output && prefix_output && prefix_lse && suffix_output && suffix_lse, B200_ERR_INVALID,
             "merge_attn_states: null pointer");
  const int esz = dtype_size(in_dtype);
  B200_CHECK(((uintptr_t)prefix_output & 15) == 0 && ((uintptr_t)suffix_output & 15) == 0 &&
                 (prefix_tok_stride * esz) % 16 == 0 && (suffix_tok_stride * esz) % 16 == 0 &&
                 (in_head_stride * esz) % 16 == 0,
             B200_ERR_INVALID, "merge_attn_states: inputs must be 16-byte aligned per head");
  const int osz = fp8_out ? 1 : esz;
  const int opack = pack * osz;
  B200_CHECK(((uintptr_t)output % opack) == 0 && (out_tok_stride * osz) % opack == 0 &&
                 (out_head_stride * osz) % opack == 0,
             B200_ERR_INVALID, "merge_attn_states: output must be aligned to its pack size");
  B200_CHECK(sm_count() > 0, B200_ERR_CUDA, "merge_attn_states: no CUDA device (no CPU fallback)");
  MergeParams p;
  p.out = output; p.out_lse = output_lse;
  p.prefix_out = prefix_output; p.prefix_lse = prefix_lse;
  p.suffix_out = suffix_output; p.suffix_lse = suffix_lse;
  p.out_scale = output_scale;
  p.in_tok_stride = prefix_tok_stride; p.sfx_tok_stride = suffix_tok_stride;
  p.in_head_stride = in_head_stride;
  p.out_tok_stride = out_tok_stride; p.out_head_stride = out_head_stride;
  p.plse_h = prefix_lse_head_stride; p.plse_t = prefix_lse_tok_stride;
  p.slse_h = suffix_lse_head_stride; p.slse_t = suffix_lse_tok_stride;
  p.olse_h = out_lse_head_stride; p.olse_t = out_lse_tok_stride;
  p.num_tokens = (int)num_tokens; p.num_heads = num_heads; p.head_size = head_size;
  const int64_t pt = prefill_tokens_with_context < 0 ? num_tokens : prefill_tokens_with_context;
  B200_CHECK(pt <= num_tokens, B200_ERR_INVALID,
             "merge_attn_states: prefill_tokens_with_context must be <= num_tokens");
  p.prefix_tokens = (int)pt;
  p.total_packs = num_tokens * num_heads * (head_size / pack);
  const int threads = 256;
  int64_t blocks = (p.total_packs + threads - 1) / threads;
  const int64_t cap = (int64_t)sm_count() * 16;
  if (blocks > cap) blocks = cap;
#define B200_MERGE_LAUNCH(T)                                                                    \
  do {                                                                                          \
    if (fp8_out) merge_attn_states_kernel<T, true><<<(int)blocks, threads, 0, stream>>>(p);     \
    else merge_attn_states_kernel<T, false><<<(int)blocks, threads, 0, stream>>>(p);            \
  } while (0)
  if (in_dtype == B200_F32) B200_MERGE_LAUNCH(float);
  else if (in_dtype == B200_F16) B200_MERGE_LAUNCH(__half);
  else B200_MERGE_LAUNCH(__nv_bfloat16);
#undef B200_MERGE_LAUNCH
  B200_LAUNCH_CHECK();
  return B200_OK;
}
