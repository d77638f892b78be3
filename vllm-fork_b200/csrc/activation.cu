// b200ops -- SiLU-and-mul between the two MLP / MoE GEMMs, optionally fused with the static FP8
// quantisation of the next GEMM's input.
//
// Contract: torch.ops._C.silu_and_mul (csrc/libtorch_stable/activation_kernels.cu:102-160,291-296)
//   out[t, i] = T( float(T(silu(x[t, i]))) * float(x[t, d + i]) ),  silu(x) = x / (1 + expf(-x))
// and torch.ops._C.silu_and_mul_quant (csrc/libtorch_stable/quantization/activation_kernels.cu:33-104)
//   out8[t, i] = e4m3(clamp(float(out[t, i]) * (1.0f / *scale), +-448)).
// Note both round silu() to the activation dtype before the multiply; we keep that.
//
// B200 design: bandwidth kernel; 16-byte vector loads of gate and up, 16/8-byte stores, grid =
// tokens x column-chunks so small decode batches still cover all SMs.
#include "common.cuh"

namespace b200 {

template <typename T>
__device__ __forceinline__ float silu_mul_rounded(T x, T y) {
  const float xf = Cvt<T>::to_f32(x);
  const T s = Cvt<T>::from_f32(__fdiv_rn(xf, 1.0f + expf(-xf)));
  return Cvt<T>::to_f32(Cvt<T>::from_f32(Cvt<T>::to_f32(s) * Cvt<T>::to_f32(y)));
}

template <typename T, bool QUANT, bool VEC>
__global__ void __launch_bounds__(256)
    silu_and_mul_kernel(void* __restrict__ out, const T* __restrict__ in, const float* __restrict__ scale,
                        int d, int64_t in_stride, int64_t out_stride) {
  constexpr int N = 16 / sizeof(T);
  pdl_launch_dependents();
  pdl_wait();   // `in` is the gate_up GEMM's output
  const int64_t tok = blockIdx.x;
  const T* xp = in + tok * in_stride;
  const T* yp = xp + d;
  float inv = 1.f;
  if constexpr (QUANT) inv = __fdiv_rn(1.0f, *scale);
  if constexpr (VEC) {
    const int nv = d / N;
    for (int v = blockIdx.y * blockDim.x + threadIdx.x; v < nv; v += gridDim.y * blockDim.x) {
      const uint4 xr = *reinterpret_cast<const uint4*>(xp + v * N);
      const uint4 yr = *reinterpret_cast<const uint4*>(yp + v * N);
      const T* xe = reinterpret_cast<const T*>(&xr);
      const T* ye = reinterpret_cast<const T*>(&yr);
      if constexpr (QUANT) {
        uint16_t o[N / 2];
#pragma unroll
        for (int i = 0; i < N / 2; i++) {
          float a = fmaxf(-448.f, fminf(__fmul_rn(silu_mul_rounded(xe[2 * i], ye[2 * i]), inv), 448.f));
          float b = fmaxf(-448.f, fminf(__fmul_rn(silu_mul_rounded(xe[2 * i + 1], ye[2 * i + 1]), inv), 448.f));
          o[i] = f32x2_to_e4m3x2(a, b);
        }
        uint8_t* dst = reinterpret_cast<uint8_t*>(out) + tok * out_stride + v * N;
        if constexpr (N == 8) *reinterpret_cast<uint2*>(dst) = *reinterpret_cast<const uint2*>(o);
        else *reinterpret_cast<uint32_t*>(dst) = *reinterpret_cast<const uint32_t*>(o);
      } else {
        uint4 r;
        T* re = reinterpret_cast<T*>(&r);
#pragma unroll
        for (int i = 0; i < N; i++) re[i] = Cvt<T>::from_f32(silu_mul_rounded(xe[i], ye[i]));
        *reinterpret_cast<uint4*>(reinterpret_cast<T*>(out) + tok * out_stride + v * N) = r;
      }
    }
  } else {
    for (int i = blockIdx.y * blockDim.x + threadIdx.x; i < d; i += gridDim.y * blockDim.x) {
      const float r = silu_mul_rounded(xp[i], yp[i]);
      if constexpr (QUANT)
        reinterpret_cast<uint8_t*>(out)[tok * out_stride + i] =
            f32_to_e4m3(fmaxf(-448.f, fminf(__fmul_rn(r, inv), 448.f)));
      else
        reinterpret_cast<T*>(out)[tok * out_stride + i] = Cvt<T>::from_f32(r);
    }
  }
}

template <typename T, bool QUANT>
static int launch_silu(void* out, const void* in, const float* scale, int64_t num_tokens, int d,
                       int64_t in_stride, int64_t out_stride, cudaStream_t stream) {
  constexpr int N = 16 / sizeof(T);
  const int osz = QUANT ? 1 : (int)sizeof(T);
  auto al = [](const void* p, int a) { return (reinterpret_cast<uintptr_t>(p) % a) == 0; };
  const bool vec = d % N == 0 && al(in, 16) && (in_stride * (int64_t)sizeof(T)) % 16 == 0 &&
                   al(out, N * osz) && (out_stride * osz) % (N * osz) == 0;
  const int work = vec ? d / N : d;
  int chunks = (work + 255) / 256;
  // enough CTAs for ~2 waves on 148 SMs at decode batch sizes, never more than needed
  int want = (int)((2 * 148 * 4 + num_tokens - 1) / num_tokens);
  if (chunks > want) chunks = want < 1 ? 1 : want;
  dim3 grid((unsigned)num_tokens, (unsigned)chunks);
  if (vec)
    B200_CUDA(launch_pdl(silu_and_mul_kernel<T, QUANT, true>, grid, dim3(256), 0, stream, out, (const T*)in, scale, d,
                         in_stride, out_stride));
  else
    B200_CUDA(launch_pdl(silu_and_mul_kernel<T, QUANT, false>, grid, dim3(256), 0, stream, out, (const T*)in, scale, d,
                         in_stride, out_stride));
  B200_LAUNCH_CHECK();
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_silu_and_mul(void* out, const void* input, int64_t num_tokens, int d,
                                 int64_t in_row_stride, int64_t out_row_stride, int dtype,
                                 b200_stream_t stream) {
  B200_CHECK(num_tokens >= 0 && d > 0, B200_ERR_INVALID, "silu_and_mul: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(out && input, B200_ERR_INVALID, "silu_and_mul: null pointer");
  switch (dtype) {
    case B200_BF16: return launch_silu<__nv_bfloat16, false>(out, input, nullptr, num_tokens, d, in_row_stride, out_row_stride, (cudaStream_t)stream);
    case B200_F16: return launch_silu<__half, false>(out, input, nullptr, num_tokens, d, in_row_stride, out_row_stride, (cudaStream_t)stream);
    case B200_F32: return launch_silu<float, false>(out, input, nullptr, num_tokens, d, in_row_stride, out_row_stride, (cudaStream_t)stream);
    default: set_last_error("silu_and_mul: unsupported dtype %d", dtype); return B200_ERR_INVALID;
  }
}

extern "C" int b200_silu_and_mul_fp8(void* out, const void* input, const float* scale,
                                     int64_t num_tokens, int d, int64_t in_row_stride,
                                     int64_t out_row_stride, int in_dtype, b200_stream_t stream) {
  B200_CHECK(num_tokens >= 0 && d > 0, B200_ERR_INVALID, "silu_and_mul_quant: bad sizes");
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(out && input && scale, B200_ERR_INVALID, "silu_and_mul_quant: null pointer");
  switch (in_dtype) {
    case B200_BF16: return launch_silu<__nv_bfloat16, true>(out, input, scale, num_tokens, d, in_row_stride, out_row_stride, (cudaStream_t)stream);
    case B200_F16: return launch_silu<__half, true>(out, input, scale, num_tokens, d, in_row_stride, out_row_stride, (cudaStream_t)stream);
    default: set_last_error("silu_and_mul_quant: input must be half or bfloat16"); return B200_ERR_INVALID;
  }
}
