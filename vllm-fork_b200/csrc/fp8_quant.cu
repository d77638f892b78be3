// b200ops -- FP8-E4M3 activation quantisers (static / dynamic per-tensor / dynamic per-token).
//
// Behavioural contract (bit-exact): csrc/libtorch_stable/quantization/w8a8/fp8/common.cu:13-179 and
// csrc/quantization/w8a8/fp8/common.cuh:58-79 of the reference:
//   static / per-tensor : r = clamp(x * (1.0f / scale), -448, 448)   (reciprocal THEN multiply)
//   per-token           : r = clamp(x / scale, -448, 448)            (true division)
//   byte = cvt.rn.satfinite.e4m3(r); fminf/fmaxf drop NaN (NaN -> +448), as in the reference.
//
// B200 design: one CTA per token row, 16-byte vector loads, 8/16-byte vector stores. The per-token
// variant keeps the row in registers between the abs-max reduction and the quantisation, so the
// input is read from HBM once (the reference reads it twice).
#include "common.cuh"

namespace b200 {

static constexpr float kFp8Max = 448.0f;

template <bool INV>
__device__ __forceinline__ float scale_clamp(float x, float s) {
  float y = INV ? __fmul_rn(x, s) : __fdiv_rn(x, s);
  return fmaxf(-kFp8Max, fminf(y, kFp8Max));
}

template <typename in_t, bool INV>
__device__ __forceinline__ void quant_vec(const uint4& raw, uint8_t* dst, float s) {
  constexpr int N = 16 / sizeof(in_t);
  const in_t* e = reinterpret_cast<const in_t*>(&raw);
  uint16_t o[N / 2];
#pragma unroll
  for (int i = 0; i < N / 2; i++) {
    float a = scale_clamp<INV>(Cvt<in_t>::to_f32(e[2 * i]), s);
    float b = scale_clamp<INV>(Cvt<in_t>::to_f32(e[2 * i + 1]), s);
    o[i] = f32x2_to_e4m3x2(a, b);
  }
  if constexpr (N == 8)
    *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(o);
  else
    *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(o);
}

template <typename in_t>
__device__ __forceinline__ float absmax_vec(const uint4& raw, float m) {
  constexpr int N = 16 / sizeof(in_t);
  const in_t* e = reinterpret_cast<const in_t*>(&raw);
#pragma unroll
  for (int i = 0; i < N; i++) m = fmaxf(m, fabsf(Cvt<in_t>::to_f32(e[i])));
  return m;
}

__device__ __forceinline__ float block_max(float v, float* red) {
  v = warp_max(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  const int nw = (blockDim.x + 31) >> 5;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float r = l < nw ? red[l] : 0.f;
  return warp_max(r);
}

// ---- static scale(s): per-tensor, per-token, per-channel or 2-D groups
template <typename in_t, bool VEC>
__global__ void __launch_bounds__(256)
    fp8_quant_static_kernel(uint8_t* __restrict__ out, const in_t* __restrict__ in,
                            const float* __restrict__ scale, int hidden, int64_t in_row_stride,
                            int64_t out_row_stride, int group_m, int group_n,
                            int64_t scale_stride_i, int64_t scale_stride_j) {
  constexpr int N = 16 / sizeof(in_t);
  const int64_t row = blockIdx.x;
  const in_t* src = in + row * in_row_stride;
  uint8_t* dst = out + row * out_row_stride;
  const float* srow = scale + (scale_stride_i == 0 ? 0 : (row / group_m) * scale_stride_i);
  if constexpr (VEC) {
    // group_n % N == 0 or single scale per row guaranteed by the host
    int cached_g = -1;
    float inv = 0.f;
    for (int v = threadIdx.x; v < hidden / N; v += blockDim.x) {
      const int g = scale_stride_j == 0 ? 0 : (v * N) / group_n;
      if (g != cached_g) {
        inv = __fdiv_rn(1.0f, srow[g * scale_stride_j]);
        cached_g = g;
      }
      uint4 raw = *reinterpret_cast<const uint4*>(src + v * N);
      quant_vec<in_t, true>(raw, dst + v * N, inv);
    }
  } else {
    int cached_g = -1;
    float inv = 0.f;
    for (int n = threadIdx.x; n < hidden; n += blockDim.x) {
      const int g = scale_stride_j == 0 ? 0 : n / group_n;
      if (g != cached_g) {
        inv = __fdiv_rn(1.0f, srow[g * scale_stride_j]);
        cached_g = g;
      }
      dst[n] = f32_to_e4m3(scale_clamp<true>(Cvt<in_t>::to_f32(src[n]), inv));
    }
  }
}

// ---- dynamic per-tensor, pass 1: scale = max_rows(rowmax / 448) via float-as-int atomicMax (>= 0)
template <typename in_t, bool VEC>
__global__ void __launch_bounds__(256)
    fp8_absmax_kernel(float* __restrict__ scale, const in_t* __restrict__ in, int hidden,
                      int64_t in_row_stride) {
  constexpr int N = 16 / sizeof(in_t);
  __shared__ float red[32];
  const in_t* src = in + (int64_t)blockIdx.x * in_row_stride;
  float m = 0.f;
  if constexpr (VEC) {
    for (int v = threadIdx.x; v < hidden / N; v += blockDim.x)
      m = absmax_vec<in_t>(*reinterpret_cast<const uint4*>(src + v * N), m);
  } else {
    for (int n = threadIdx.x; n < hidden; n += blockDim.x)
      m = fmaxf(m, fabsf(Cvt<in_t>::to_f32(src[n])));
  }
  m = block_max(m, red);
  if (threadIdx.x == 0) atomicMax(reinterpret_cast<int*>(scale), __float_as_int(__fdiv_rn(m, kFp8Max)));
}

// ---- dynamic per-tensor, pass 2
template <typename in_t, bool VEC>
__global__ void __launch_bounds__(256)
    fp8_quant_tensor_kernel(uint8_t* __restrict__ out, const in_t* __restrict__ in,
                            const float* __restrict__ scale, int hidden, int64_t in_row_stride,
                            int64_t out_row_stride) {
  constexpr int N = 16 / sizeof(in_t);
  const in_t* src = in + (int64_t)blockIdx.x * in_row_stride;
  uint8_t* dst = out + (int64_t)blockIdx.x * out_row_stride;
  const float inv = __fdiv_rn(1.0f, *scale);
  if constexpr (VEC) {
    for (int v = threadIdx.x; v < hidden / N; v += blockDim.x)
      quant_vec<in_t, true>(*reinterpret_cast<const uint4*>(src + v * N), dst + v * N, inv);
  } else {
    for (int n = threadIdx.x; n < hidden; n += blockDim.x)
      dst[n] = f32_to_e4m3(scale_clamp<true>(Cvt<in_t>::to_f32(src[n]), inv));
  }
}

// ---- dynamic per-token: row kept in registers (R vectors / thread) when it fits
template <typename in_t, bool VEC, int R>
__global__ void __launch_bounds__(256)
    fp8_quant_token_kernel(uint8_t* __restrict__ out, float* __restrict__ scales,
                           const in_t* __restrict__ in, const float* __restrict__ scale_ub,
                           int hidden, int64_t in_row_stride, int64_t out_row_stride) {
  constexpr int N = 16 / sizeof(in_t);
  __shared__ float red[32];
  __shared__ float s_scale;
  const int64_t row = blockIdx.x;
  const in_t* src = in + row * in_row_stride;
  uint8_t* dst = out + row * out_row_stride;
  float m = 0.f;
  uint4 regs[R > 0 ? R : 1];
  if constexpr (VEC) {
    const int nv = hidden / N;
    if constexpr (R > 0) {
#pragma unroll
      for (int i = 0; i < R; i++) {
        const int v = threadIdx.x + i * blockDim.x;
        if (v < nv) {
          regs[i] = *reinterpret_cast<const uint4*>(src + v * N);
          m = absmax_vec<in_t>(regs[i], m);
        }
      }
    } else {
      for (int v = threadIdx.x; v < nv; v += blockDim.x)
        m = absmax_vec<in_t>(*reinterpret_cast<const uint4*>(src + v * N), m);
    }
  } else {
    for (int n = threadIdx.x; n < hidden; n += blockDim.x)
      m = fmaxf(m, fabsf(Cvt<in_t>::to_f32(src[n])));
  }
  m = block_max(m, red);
  if (threadIdx.x == 0) {
    float t = scale_ub ? fminf(m, *scale_ub) : m;
    t = fmaxf(__fdiv_rn(t, kFp8Max), 1.0f / (kFp8Max * 512.0f));
    scales[row] = t;
    s_scale = t;
  }
  __syncthreads();
  const float s = s_scale;
  if constexpr (VEC) {
    const int nv = hidden / N;
    if constexpr (R > 0) {
#pragma unroll
      for (int i = 0; i < R; i++) {
        const int v = threadIdx.x + i * blockDim.x;
        if (v < nv) quant_vec<in_t, false>(regs[i], dst + v * N, s);
      }
    } else {
      for (int v = threadIdx.x; v < nv; v += blockDim.x)
        quant_vec<in_t, false>(*reinterpret_cast<const uint4*>(src + v * N), dst + v * N, s);
    }
  } else {
    for (int n = threadIdx.x; n < hidden; n += blockDim.x)
      dst[n] = f32_to_e4m3(scale_clamp<false>(Cvt<in_t>::to_f32(src[n]), s));
  }
}

static bool vec_ok(const void* in, const void* out, int hidden, int64_t irs, int64_t ors, int isz) {
  const int N = 16 / isz;
  auto al = [](const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  return hidden % N == 0 && al(in, 16) && (irs * isz) % 16 == 0 && al(out, N) && ors % N == 0;
}

static int check_common(const char* op, const void* out, const void* in, int64_t num_tokens,
                        int hidden, int in_dtype) {
  B200_CHECK(num_tokens >= 0 && hidden > 0, B200_ERR_INVALID, "%s: bad sizes (%lld x %d)", op,
             (long long)num_tokens, hidden);
  B200_CHECK(num_tokens == 0 || (out && in), B200_ERR_INVALID, "%s: null pointer", op);
  B200_CHECK(in_dtype == B200_BF16 || in_dtype == B200_F16 || in_dtype == B200_F32,
             B200_ERR_INVALID, "%s: input must be float, half or bfloat16 (got dtype %d)", op,
             in_dtype);
  B200_CHECK(num_tokens <= 0x7fffffffLL, B200_ERR_UNSUPPORTED, "%s: more than 2^31-1 rows", op);
  return B200_OK;
}

}  // namespace b200

using namespace b200;

#define B200_DISPATCH_IN(dt, ...)                      \
  switch (dt) {                                        \
    case B200_BF16: { using in_t = __nv_bfloat16; __VA_ARGS__; break; } \
    case B200_F16: { using in_t = __half; __VA_ARGS__; break; }          \
    default: { using in_t = float; __VA_ARGS__; break; }                 \
  }

extern "C" int b200_fp8_quant_static(void* out, const void* input, const float* scale,
                                     int64_t num_tokens, int hidden, int64_t in_row_stride,
                                     int64_t out_row_stride, int group_m, int group_n,
                                     int64_t scale_stride_i, int64_t scale_stride_j, int in_dtype,
                                     b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc = check_common("static_scaled_fp8_quant", out, input, num_tokens, hidden, in_dtype);
  if (rc) return rc;
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(scale, B200_ERR_INVALID, "static_scaled_fp8_quant: null scale");
  B200_CHECK(group_m > 0 && group_n > 0, B200_ERR_INVALID,
             "static_scaled_fp8_quant: group shape must be positive (got %d x %d)", group_m, group_n);
  const int isz = dtype_size(in_dtype);
  const int N = 16 / isz;
  bool v = vec_ok(input, out, hidden, in_row_stride, out_row_stride, isz) &&
           (scale_stride_j == 0 || group_n % N == 0);
  B200_DISPATCH_IN(in_dtype, {
    if (v)
      fp8_quant_static_kernel<in_t, true><<<(unsigned)num_tokens, 256, 0, stream>>>(
          (uint8_t*)out, (const in_t*)input, scale, hidden, in_row_stride, out_row_stride, group_m,
          group_n, scale_stride_i, scale_stride_j);
    else
      fp8_quant_static_kernel<in_t, false><<<(unsigned)num_tokens, 256, 0, stream>>>(
          (uint8_t*)out, (const in_t*)input, scale, hidden, in_row_stride, out_row_stride, group_m,
          group_n, scale_stride_i, scale_stride_j);
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_fp8_quant_dynamic_tensor(void* out, const void* input, float* scale,
                                             int64_t num_tokens, int hidden, int64_t in_row_stride,
                                             int64_t out_row_stride, int in_dtype,
                                             b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc = check_common("dynamic_scaled_fp8_quant", out, input, num_tokens, hidden, in_dtype);
  if (rc) return rc;
  B200_CHECK(scale, B200_ERR_INVALID, "dynamic_scaled_fp8_quant: null scale");
  // scale must start at <= 0 before the atomic max (reference: cudaMemsetAsync, common.cu:352-354)
  B200_CUDA(cudaMemsetAsync(scale, 0, sizeof(float), stream));
  if (num_tokens == 0) return B200_OK;
  const bool v = vec_ok(input, out, hidden, in_row_stride, out_row_stride, dtype_size(in_dtype));
  B200_DISPATCH_IN(in_dtype, {
    if (v) {
      fp8_absmax_kernel<in_t, true><<<(unsigned)num_tokens, 256, 0, stream>>>(
          scale, (const in_t*)input, hidden, in_row_stride);
      fp8_quant_tensor_kernel<in_t, true><<<(unsigned)num_tokens, 256, 0, stream>>>(
          (uint8_t*)out, (const in_t*)input, scale, hidden, in_row_stride, out_row_stride);
    } else {
      fp8_absmax_kernel<in_t, false><<<(unsigned)num_tokens, 256, 0, stream>>>(
          scale, (const in_t*)input, hidden, in_row_stride);
      fp8_quant_tensor_kernel<in_t, false><<<(unsigned)num_tokens, 256, 0, stream>>>(
          (uint8_t*)out, (const in_t*)input, scale, hidden, in_row_stride, out_row_stride);
    }
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_fp8_quant_dynamic_token(void* out, const void* input, float* scales,
                                            const float* scale_ub, int64_t num_tokens, int hidden,
                                            int64_t in_row_stride, int64_t out_row_stride,
                                            int in_dtype, b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  int rc =
      check_common("dynamic_per_token_scaled_fp8_quant", out, input, num_tokens, hidden, in_dtype);
  if (rc) return rc;
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(scales, B200_ERR_INVALID, "dynamic_per_token_scaled_fp8_quant: null scales");
  const int isz = dtype_size(in_dtype);
  const bool v = vec_ok(input, out, hidden, in_row_stride, out_row_stride, isz);
  const int nv = hidden / (16 / isz);
  B200_DISPATCH_IN(in_dtype, {
    if (v && nv <= 256 * 2)
      fp8_quant_token_kernel<in_t, true, 2><<<(unsigned)num_tokens, 256, 0, stream>>>(
          (uint8_t*)out, scales, (const in_t*)input, scale_ub, hidden, in_row_stride,
          out_row_stride);
    else if (v && nv <= 256 * 8)
      fp8_quant_token_kernel<in_t, true, 8><<<(unsigned)num_tokens, 256, 0, stream>>>(
          (uint8_t*)out, scales, (const in_t*)input, scale_ub, hidden, in_row_stride,
          out_row_stride);
    else if (v)
      fp8_quant_token_kernel<in_t, true, 0><<<(unsigned)num_tokens, 256, 0, stream>>>(
          (uint8_t*)out, scales, (const in_t*)input, scale_ub, hidden, in_row_stride,
          out_row_stride);
    else
      fp8_quant_token_kernel<in_t, false, 0><<<(unsigned)num_tokens, 256, 0, stream>>>(
          (uint8_t*)out, scales, (const in_t*)input, scale_ub, hidden, in_row_stride,
          out_row_stride);
  });
  B200_LAUNCH_CHECK();
  return B200_OK;
}
