// b200ops -- the bandwidth-trivial neighbours of the hot path:
// RMSNorm (+ residual add) fused with the static FP8 quantisation of the next GEMM input, and RoPE.
//
// Contracts (reference, csrc/libtorch_stable):
//   rms_norm                              layernorm_kernels.cu:93      out = T(x * rsqrt(mean(x^2)+eps) * w)  (fp32, one rounding)
//   fused_add_rms_norm                    layernorm_kernels.cu         z = T(x + r); r = z; x = norm(z)
//   rms_norm_static_fp8_quant             layernorm_quant_kernels.cu:21-75
//        out8 = e4m3(clamp(float(T(x * s * w)) * (1/scale)))          (rounded through T first)
//   fused_add_rms_norm_static_fp8_quant   layernorm_quant_kernels.cu:81-156
//   rotary_embedding                      pos_encoding_kernels.cu:8-99  (neox / gpt-j, in place)
//
// B200 design: one CTA per token, the row stays in registers between the sum-of-squares pass and
// the normalise/quantise pass (one HBM read, one write), 16-byte vectors.
#include "common.cuh"

namespace b200 {

// MODE: 0 = norm -> T out, 1 = norm -> fp8 out
template <typename T, bool ADD, bool QUANT, int R>
__global__ void __launch_bounds__(256)
    rms_norm_kernel(void* __restrict__ out, T* __restrict__ input, T* __restrict__ residual,
                    const T* __restrict__ weight, const float* __restrict__ scale, float eps,
                    int hidden, int64_t in_stride, int64_t out_stride) {
  constexpr int N = 16 / sizeof(T);
  __shared__ float red[32];
  pdl_launch_dependents();
  pdl_wait();   // input / residual come from the previous kernel on the stream
  const int64_t row = blockIdx.x;
  T* in_row = input + row * in_stride;
  T* res_row = ADD ? residual + row * (int64_t)hidden : nullptr;
  const int nv = hidden / N;
  uint4 regs[R];
  float ss = 0.f;
#pragma unroll
  for (int i = 0; i < R; i++) {
    const int v = threadIdx.x + i * blockDim.x;
    if (v < nv) {
      uint4 x = *reinterpret_cast<const uint4*>(in_row + v * N);
      if constexpr (ADD) {
        const uint4 r = *reinterpret_cast<const uint4*>(res_row + v * N);
        T* xe = reinterpret_cast<T*>(&x);
        const T* re = reinterpret_cast<const T*>(&r);
#pragma unroll
        for (int j = 0; j < N; j++) xe[j] = Cvt<T>::from_f32(Cvt<T>::to_f32(xe[j]) + Cvt<T>::to_f32(re[j]));
        *reinterpret_cast<uint4*>(res_row + v * N) = x;
      }
      const T* xe = reinterpret_cast<const T*>(&x);
#pragma unroll
      for (int j = 0; j < N; j++) {
        const float f = Cvt<T>::to_f32(xe[j]);
        ss += f * f;
      }
      regs[i] = x;
    }
  }
  ss = block_sum(ss, red);
  const float s = rsqrtf(ss / (float)hidden + eps);
  float inv = 1.f;
  if constexpr (QUANT) inv = __fdiv_rn(1.0f, *scale);
#pragma unroll
  for (int i = 0; i < R; i++) {
    const int v = threadIdx.x + i * blockDim.x;
    if (v < nv) {
      const uint4 w = *reinterpret_cast<const uint4*>(weight + v * N);
      const T* xe = reinterpret_cast<const T*>(&regs[i]);
      const T* we = reinterpret_cast<const T*>(&w);
      if constexpr (QUANT) {
        uint16_t o[N / 2];
#pragma unroll
        for (int j = 0; j < N / 2; j++) {
          // x * s * w in fp32, rounded through T, then scaled (layernorm_quant_kernels.cu:66-71)
          float a = Cvt<T>::to_f32(Cvt<T>::from_f32(Cvt<T>::to_f32(xe[2 * j]) * s * Cvt<T>::to_f32(we[2 * j])));
          float b = Cvt<T>::to_f32(Cvt<T>::from_f32(Cvt<T>::to_f32(xe[2 * j + 1]) * s * Cvt<T>::to_f32(we[2 * j + 1])));
          a = fmaxf(-448.f, fminf(__fmul_rn(a, inv), 448.f));
          b = fmaxf(-448.f, fminf(__fmul_rn(b, inv), 448.f));
          o[j] = f32x2_to_e4m3x2(a, b);
        }
        uint8_t* dst = reinterpret_cast<uint8_t*>(out) + row * out_stride + v * N;
        if constexpr (N == 8) *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(o);
        else *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(o);
      } else {
        uint4 r;
        T* re = reinterpret_cast<T*>(&r);
#pragma unroll
        for (int j = 0; j < N; j++)
          re[j] = Cvt<T>::from_f32(Cvt<T>::to_f32(xe[j]) * s * Cvt<T>::to_f32(we[j]));  // one rounding (layernorm_kernels.cu:93,163)
        *reinterpret_cast<uint4*>(reinterpret_cast<T*>(out) + row * out_stride + v * N) = r;
      }
    }
  }
}

template <typename T, bool NEOX>
__global__ void __launch_bounds__(256)
    rotary_kernel(const int64_t* __restrict__ positions, T* __restrict__ query, T* __restrict__ key,
                  const T* __restrict__ cache, int rot_dim, int64_t q_stride, int64_t k_stride,
                  int64_t head_stride, int num_heads, int num_kv_heads) {
  const int tok = blockIdx.x;
  const int64_t pos = positions[tok];
  const int embed = rot_dim / 2;
  const T* cosp = cache + pos * rot_dim;
  const T* sinp = cosp + embed;
  const int nq = num_heads * embed, nk = key ? num_kv_heads * embed : 0;
  for (int i = threadIdx.x; i < nq + nk; i += blockDim.x) {
    const bool is_k = i >= nq;
    const int j = is_k ? i - nq : i;
    const int h = j / embed, ro = j - h * embed;
    T* arr = (is_k ? key + tok * k_stride : query + tok * q_stride) + h * head_stride;
    const int xi = NEOX ? ro : 2 * ro;
    const int yi = NEOX ? embed + ro : 2 * ro + 1;
    T x = arr[xi], y = arr[yi];
    rope_rotate<T>(x, y, Cvt<T>::to_f32(cosp[ro]), Cvt<T>::to_f32(sinp[ro]));
    arr[xi] = x;
    arr[yi] = y;
  }
}

template <typename T, bool ADD, bool QUANT>
static int launch_norm(void* out, void* input, void* residual, const void* weight, const float* scale,
                       float eps, int64_t num_tokens, int hidden, int64_t in_stride,
                       int64_t out_stride, cudaStream_t stream) {
  constexpr int N = 16 / sizeof(T);
  const int nv = hidden / N;
  dim3 grid((unsigned)num_tokens);
#define B200_NORM(RR)                                                                            \
  B200_CUDA(launch_pdl(rms_norm_kernel<T, ADD, QUANT, RR>, grid, dim3(256), 0, stream, out, (T*)input,   \
                       (T*)residual, (const T*)weight, scale, eps, hidden, in_stride, out_stride))
  if (nv <= 256) B200_NORM(1);
  else if (nv <= 512) B200_NORM(2);
  else if (nv <= 1024) B200_NORM(4);
  else B200_NORM(8);
#undef B200_NORM
  B200_LAUNCH_CHECK();
  return B200_OK;
}

static int norm_checks(const char* op, const void* out, const void* input, const void* weight,
                       int64_t num_tokens, int hidden, int64_t in_stride, int dtype) {
  B200_CHECK(num_tokens >= 0 && hidden > 0, B200_ERR_INVALID, "%s: bad sizes", op);
  B200_CHECK(num_tokens == 0 || (out && input && weight), B200_ERR_INVALID, "%s: null pointer", op);
  B200_CHECK(dtype == B200_BF16 || dtype == B200_F16, B200_ERR_UNSUPPORTED,
             "%s: only half / bfloat16 inputs are built", op);
  B200_CHECK(hidden % 8 == 0 && hidden <= 16384 && in_stride % 8 == 0, B200_ERR_UNSUPPORTED,
             "%s: hidden_size must be a multiple of 8 and <= 16384 (got %d)", op, hidden);
  B200_CHECK(((uintptr_t)input & 15) == 0 && ((uintptr_t)weight & 15) == 0 && ((uintptr_t)out & 7) == 0,
             B200_ERR_INVALID, "%s: pointers must be 16-byte aligned", op);
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_rms_norm(void* out, const void* input, const void* weight, float eps,
                             int64_t num_tokens, int hidden, int64_t in_row_stride, int dtype,
                             b200_stream_t stream) {
  int rc = norm_checks("rms_norm", out, input, weight, num_tokens, hidden, in_row_stride, dtype);
  if (rc || num_tokens == 0) return rc;
  if (dtype == B200_BF16)
    return launch_norm<__nv_bfloat16, false, false>(out, (void*)input, nullptr, weight, nullptr, eps, num_tokens, hidden, in_row_stride, hidden, (cudaStream_t)stream);
  return launch_norm<__half, false, false>(out, (void*)input, nullptr, weight, nullptr, eps, num_tokens, hidden, in_row_stride, hidden, (cudaStream_t)stream);
}

extern "C" int b200_fused_add_rms_norm(void* input, void* residual, const void* weight, float eps,
                                       int64_t num_tokens, int hidden, int64_t in_row_stride,
                                       int dtype, b200_stream_t stream) {
  int rc = norm_checks("fused_add_rms_norm", input, input, weight, num_tokens, hidden, in_row_stride, dtype);
  if (rc || num_tokens == 0) return rc;
  B200_CHECK(residual, B200_ERR_INVALID, "fused_add_rms_norm: null residual");
  if (dtype == B200_BF16)
    return launch_norm<__nv_bfloat16, true, false>(input, input, residual, weight, nullptr, eps, num_tokens, hidden, in_row_stride, in_row_stride, (cudaStream_t)stream);
  return launch_norm<__half, true, false>(input, input, residual, weight, nullptr, eps, num_tokens, hidden, in_row_stride, in_row_stride, (cudaStream_t)stream);
}

extern "C" int b200_rms_norm_static_fp8_quant(void* out, const void* input, const void* weight,
                                              const float* scale, float eps, int64_t num_tokens,
                                              int hidden, int64_t in_row_stride, int dtype,
                                              b200_stream_t stream) {
  int rc = norm_checks("rms_norm_static_fp8_quant", out, input, weight, num_tokens, hidden, in_row_stride, dtype);
  if (rc || num_tokens == 0) return rc;
  B200_CHECK(scale, B200_ERR_INVALID, "rms_norm_static_fp8_quant: null scale");
  if (dtype == B200_BF16)
    return launch_norm<__nv_bfloat16, false, true>(out, (void*)input, nullptr, weight, scale, eps, num_tokens, hidden, in_row_stride, hidden, (cudaStream_t)stream);
  return launch_norm<__half, false, true>(out, (void*)input, nullptr, weight, scale, eps, num_tokens, hidden, in_row_stride, hidden, (cudaStream_t)stream);
}

extern "C" int b200_fused_add_rms_norm_static_fp8_quant(void* out, void* input, void* residual,
                                                        const void* weight, const float* scale,
                                                        float eps, int64_t num_tokens, int hidden,
                                                        int64_t in_row_stride, int dtype,
                                                        b200_stream_t stream) {
  int rc = norm_checks("fused_add_rms_norm_static_fp8_quant", out, input, weight, num_tokens, hidden, in_row_stride, dtype);
  if (rc || num_tokens == 0) return rc;
  B200_CHECK(scale && residual, B200_ERR_INVALID, "fused_add_rms_norm_static_fp8_quant: null pointer");
  if (dtype == B200_BF16)
    return launch_norm<__nv_bfloat16, true, true>(out, input, residual, weight, scale, eps, num_tokens, hidden, in_row_stride, hidden, (cudaStream_t)stream);
  return launch_norm<__half, true, true>(out, input, residual, weight, scale, eps, num_tokens, hidden, in_row_stride, hidden, (cudaStream_t)stream);
}

extern "C" int b200_rotary_embedding(const int64_t* positions, void* query, void* key,
                                     const void* cos_sin_cache, int64_t num_tokens, int rot_dim,
                                     int64_t query_stride, int64_t key_stride, int64_t head_stride,
                                     int num_heads, int num_kv_heads, int is_neox, int dtype,
                                     b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(num_tokens >= 0 && rot_dim > 0 && rot_dim % 2 == 0 && num_heads > 0, B200_ERR_INVALID,
             "rotary_embedding: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(positions && query && cos_sin_cache, B200_ERR_INVALID, "rotary_embedding: null pointer");
  dim3 grid((unsigned)num_tokens);
#define B200_ROPE(T)                                                                               \
  if (is_neox)                                                                                     \
    rotary_kernel<T, true><<<grid, 256, 0, stream>>>(positions, (T*)query, (T*)key, (const T*)cos_sin_cache, \
                                                     rot_dim, query_stride, key_stride, head_stride, \
                                                     num_heads, num_kv_heads);                     \
  else                                                                                             \
    rotary_kernel<T, false><<<grid, 256, 0, stream>>>(positions, (T*)query, (T*)key, (const T*)cos_sin_cache, \
                                                      rot_dim, query_stride, key_stride, head_stride, \
                                                      num_heads, num_kv_heads)
  switch (dtype) {
    case B200_BF16: B200_ROPE(__nv_bfloat16); break;
    case B200_F16: B200_ROPE(__half); break;
    case B200_F32: B200_ROPE(float); break;
    default: set_last_error("rotary_embedding: unsupported dtype %d", dtype); return B200_ERR_INVALID;
  }
#undef B200_ROPE
  B200_LAUNCH_CHECK();
  return B200_OK;
}
