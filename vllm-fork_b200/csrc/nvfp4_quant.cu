// b200ops -- NVFP4 activation quantiser (E2M1 values, one E4M3 scale per 16 elements, one fp32
// global scale), scale factors written in the tcgen05 128x4 swizzled layout or row-major.
//
// Contract (bit-exact instruction sequence): torch.ops._C.scaled_fp4_quant.out
//   csrc/libtorch_stable/quantization/fp4/nvfp4_quant_kernels.cu:39-112, 188-281 and
//   nvfp4_utils.cuh:158-302 (cvt_warp_fp16_to_fp4, cvt_quant_to_fp4_get_sf_out_offset):
//     SF      = e4m3( gs * (amax16 * rcp.approx.ftz(6)) )
//     outScale= SF != 0 ? rcp.approx.ftz( float(SF) * rcp.approx.ftz(gs) ) : 0
//     nibble  = cvt.rn.satfinite.e2m1x2.f32(x * outScale)      (low nibble = even element)
//   Swizzled SF offset of (row m, scale column k):
//     ((m/128 * numKTiles + k/4) << 9) | ((m%32) << 4) | (((m/32)%4) << 2) | (k%4)
//   Padded rows (up to a multiple of 128) and padded scale columns (up to a multiple of 4) are
//   written as zero.
//
// B200 design: bandwidth kernel (2.5625 B per element); one thread per 16-element block, two
// 16-byte loads, one 8-byte store, 2-D grid (row groups x column chunks) sized for >= 2 waves.
#include "common.cuh"

namespace b200 {

__device__ __forceinline__ float rcp_approx_ftz(float a) {
  float b;
  asm volatile("rcp.approx.ftz.f32 %0, %1;" : "=f"(b) : "f"(a));
  return b;
}
__device__ __forceinline__ uint32_t f32x8_to_e2m1(const float (&a)[8]) {
  uint32_t val;
  asm volatile(
      "{\n"
      ".reg .b8 b0, b1, b2, b3;\n"
      "cvt.rn.satfinite.e2m1x2.f32 b0, %2, %1;\n"
      "cvt.rn.satfinite.e2m1x2.f32 b1, %4, %3;\n"
      "cvt.rn.satfinite.e2m1x2.f32 b2, %6, %5;\n"
      "cvt.rn.satfinite.e2m1x2.f32 b3, %8, %7;\n"
      "mov.b32 %0, {b0, b1, b2, b3};\n"
      "}"
      : "=r"(val)
      : "f"(a[0]), "f"(a[1]), "f"(a[2]), "f"(a[3]), "f"(a[4]), "f"(a[5]), "f"(a[6]), "f"(a[7]));
  return val;
}

// silu(x) * y in fp32, rounded ONCE to the activation dtype: compute_silu_mul, nvfp4_utils.cuh:304-329
// (this differs from the stand-alone silu_and_mul op, which also rounds silu(x) first)
template <typename T>
__device__ __forceinline__ float silu_mul_f32_rounded(T x, T y) {
  const float xf = Cvt<T>::to_f32(x);
  const float s = __fdividef(xf, 1.f + __expf(-xf));
  return Cvt<T>::to_f32(Cvt<T>::from_f32(s * Cvt<T>::to_f32(y)));
}

// GATED: the input row is [gate(K) | up(K)] and the quantised tensor is silu(gate) * up
// (torch.ops._C.silu_and_mul_nvfp4_quant, activation_nvfp4_quant_fusion_kernels.cu:37-118).
template <typename T, bool GATED>
__global__ void __launch_bounds__(256)
    nvfp4_quant_kernel(uint8_t* __restrict__ out, uint8_t* __restrict__ sf_out,
                       const T* __restrict__ in, const float* __restrict__ global_scale, int M, int K,
                       int64_t in_stride, int swizzled) {
  const int sf_cols = K / 16;
  const int sf_cols_pad = swizzled ? (sf_cols + 3) / 4 * 4 : sf_cols;
  const int rows_pad = swizzled ? (M + 127) / 128 * 128 : M;
  const int num_k_tiles = (K + 63) / 64;
  const float gs = global_scale ? global_scale[0] : 1.0f;
  const int kb = blockIdx.y * blockDim.x + threadIdx.x;  // scale column of this thread
  if (kb >= sf_cols_pad) return;
  for (int row = blockIdx.x; row < rows_pad; row += gridDim.x) {
    const bool valid = row < M && kb < sf_cols;
    float x[16];
    if (valid) {
      const uint4 v0 = *reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + kb * 16);
      const uint4 v1 = *reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + kb * 16 + 8);
      const T* e0 = reinterpret_cast<const T*>(&v0);
      const T* e1 = reinterpret_cast<const T*>(&v1);
#pragma unroll
      for (int i = 0; i < 8; i++) {
        x[i] = Cvt<T>::to_f32(e0[i]);
        x[8 + i] = Cvt<T>::to_f32(e1[i]);
      }
      if constexpr (GATED) {
        const uint4 u0 = *reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + K + kb * 16);
        const uint4 u1 = *reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + K + kb * 16 + 8);
        const T* f0 = reinterpret_cast<const T*>(&u0);
        const T* f1 = reinterpret_cast<const T*>(&u1);
#pragma unroll
        for (int i = 0; i < 8; i++) {
          x[i] = silu_mul_f32_rounded(e0[i], f0[i]);
          x[8 + i] = silu_mul_f32_rounded(e1[i], f1[i]);
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; i++) x[i] = 0.f;
    }
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i++) amax = fmaxf(amax, fabsf(x[i]));
    float sfv = gs * (amax * rcp_approx_ftz(6.0f));
    const uint8_t sf8 = f32_to_e4m3(sfv);
    sfv = e4m3_to_f32(sf8);
    int64_t off;
    if (swizzled)
      off = (((int64_t)(row >> 7) * num_k_tiles + (kb >> 2)) << 9) | ((row & 31) << 4) |
            (((row >> 5) & 3) << 2) | (kb & 3);
    else
      off = (int64_t)row * sf_cols + kb;
    sf_out[off] = sf8;
    if (valid) {
      const float os = sfv != 0.0f ? rcp_approx_ftz(sfv * rcp_approx_ftz(gs)) : 0.0f;
      float y0[8], y1[8];
#pragma unroll
      for (int i = 0; i < 8; i++) {
        y0[i] = x[i] * os;
        y1[i] = x[8 + i] * os;
      }
      uint2 packed;
      packed.x = f32x8_to_e2m1(y0);
      packed.y = f32x8_to_e2m1(y1);
      *reinterpret_cast<uint2*>(out + (int64_t)row * (K / 2) + kb * 8) = packed;
    }
  }
}


// ------------------------------------------------------------------ packed-MoE variant
// torch.ops._C.scaled_fp4_experts_quant / silu_and_mul_scaled_fp4_experts_quant
// (csrc/libtorch_stable/quantization/fp4/nvfp4_experts_quant.cu:36-250, entry nvfp4_quant_entry.cu:120):
// rows are already sorted by expert, expert e owns rows [expert_offsets[e], expert_offsets[e+1]), uses its own
// global scale global_scale[e] and writes its scale factors as a separate swizzled matrix that starts
// blockscale_offsets[e] (a multiple of 128) rows into the scale buffer:
//   byte offset = blockscale_offsets[e] * numKTiles * 4 + swizzled(row - expert_offsets[e], scale column).
// Same arithmetic as nvfp4_quant_kernel. Rows of the scale buffer past an expert's last row are left
// untouched (the reference allocates the buffer with torch.empty and never reads them for a valid output row);
// padded scale COLUMNS of valid rows are zeroed so a k-tail MMA multiplies zero codes by zero scales.
template <typename T, bool GATED>
__global__ void __launch_bounds__(256)
    nvfp4_experts_quant_kernel(uint8_t* __restrict__ out, uint8_t* __restrict__ sf_out,
                               const T* __restrict__ in, const float* __restrict__ global_scale,
                               const int32_t* __restrict__ expert_offsets,
                               const int32_t* __restrict__ blockscale_offsets, int num_experts, int M, int K,
                               int64_t in_stride) {
  extern __shared__ int32_t s_off[];   // [num_experts + 1]
  for (int i = threadIdx.x; i <= num_experts; i += blockDim.x) s_off[i] = expert_offsets[i];
  __syncthreads();
  const int sf_cols = K / 16;
  const int sf_cols_pad = (sf_cols + 3) / 4 * 4;
  const int num_k_tiles = (K + 63) / 64;
  const int kb = blockIdx.y * blockDim.x + threadIdx.x;  // scale column of this thread
  if (kb >= sf_cols_pad) return;
  const int rows = min(M, s_off[num_experts]);
  for (int row = blockIdx.x; row < rows; row += gridDim.x) {
    // last expert whose first row is <= row (empty experts share their successor's offset)
    int lo = 0, hi = num_experts;   // invariant: s_off[lo] <= row < s_off[hi]
    while (hi - lo > 1) {
      const int mid = (lo + hi) >> 1;
      if (s_off[mid] <= row) lo = mid; else hi = mid;
    }
    const int e = lo;
    const int r = row - s_off[e];
    const bool valid = kb < sf_cols;
    const float gs = global_scale ? global_scale[e] : 1.0f;
    float x[16];
    if (valid) {
      const uint4 v0 = *reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + kb * 16);
      const uint4 v1 = *reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + kb * 16 + 8);
      const T* e0 = reinterpret_cast<const T*>(&v0);
      const T* e1 = reinterpret_cast<const T*>(&v1);
#pragma unroll
      for (int i = 0; i < 8; i++) {
        x[i] = Cvt<T>::to_f32(e0[i]);
        x[8 + i] = Cvt<T>::to_f32(e1[i]);
      }
      if constexpr (GATED) {
        const uint4 u0 = *reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + K + kb * 16);
        const uint4 u1 = *reinterpret_cast<const uint4*>(in + (int64_t)row * in_stride + K + kb * 16 + 8);
        const T* f0 = reinterpret_cast<const T*>(&u0);
        const T* f1 = reinterpret_cast<const T*>(&u1);
#pragma unroll
        for (int i = 0; i < 8; i++) {
          x[i] = silu_mul_f32_rounded(e0[i], f0[i]);
          x[8 + i] = silu_mul_f32_rounded(e1[i], f1[i]);
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < 16; i++) x[i] = 0.f;
    }
    float amax = 0.f;
#pragma unroll
    for (int i = 0; i < 16; i++) amax = fmaxf(amax, fabsf(x[i]));
    float sfv = gs * (amax * rcp_approx_ftz(6.0f));
    const uint8_t sf8 = f32_to_e4m3(sfv);
    sfv = e4m3_to_f32(sf8);
    const int64_t off = (int64_t)blockscale_offsets[e] * num_k_tiles * 4 +
                        ((((int64_t)(r >> 7) * num_k_tiles + (kb >> 2)) << 9) | ((r & 31) << 4) |
                         (((r >> 5) & 3) << 2) | (kb & 3));
    sf_out[off] = sf8;
    if (valid) {
      const float os = sfv != 0.0f ? rcp_approx_ftz(sfv * rcp_approx_ftz(gs)) : 0.0f;
      float y0[8], y1[8];
#pragma unroll
      for (int i = 0; i < 8; i++) {
        y0[i] = x[i] * os;
        y1[i] = x[8 + i] * os;
      }
      uint2 packed;
      packed.x = f32x8_to_e2m1(y0);
      packed.y = f32x8_to_e2m1(y1);
      *reinterpret_cast<uint2*>(out + (int64_t)row * (K / 2) + kb * 8) = packed;
    }
  }
}

}  // namespace b200

using namespace b200;

static int nvfp4_quant_launch(void* out, void* out_scale, const void* input, const float* global_scale,
                              int64_t num_tokens, int hidden, int64_t in_row_stride,
                              int is_sf_swizzled_layout, int in_dtype, bool gated, cudaStream_t stream,
                              const char* op) {
  B200_CHECK(num_tokens >= 0 && hidden > 0, B200_ERR_INVALID, "%s: bad sizes", op);
  B200_CHECK(hidden % 16 == 0, B200_ERR_INVALID, "%s: last dim has to be multiple of 16, but got %d", op,
             hidden);
  B200_CHECK(in_dtype == B200_BF16 || in_dtype == B200_F16, B200_ERR_INVALID,
             "%s: input.dtype needs to be fp16 or bf16", op);
  if (num_tokens == 0) return B200_OK;
  B200_CHECK(out && out_scale && input && global_scale, B200_ERR_INVALID, "%s: null pointer", op);
  B200_CHECK(((uintptr_t)input & 15) == 0 && (in_row_stride * 2) % 16 == 0 && ((uintptr_t)out & 7) == 0,
             B200_ERR_INVALID, "%s: input rows must be 16-byte aligned", op);
  B200_CHECK(num_tokens < (1LL << 31), B200_ERR_UNSUPPORTED, "%s: too many rows", op);
  const int M = (int)num_tokens;
  const int sf_cols_pad = is_sf_swizzled_layout ? (hidden / 16 + 3) / 4 * 4 : hidden / 16;
  const int rows_pad = is_sf_swizzled_layout ? (M + 127) / 128 * 128 : M;
  const int threads = sf_cols_pad < 256 ? (sf_cols_pad + 31) / 32 * 32 : 256;
  const int gy = (sf_cols_pad + threads - 1) / threads;
  int gx = rows_pad;
  const int sms = sm_count();
  B200_CHECK(sms > 0, B200_ERR_CUDA, "%s: no CUDA device (no CPU fallback)", op);
  const int cap = (sms * 8 + gy - 1) / gy;
  if (gx > cap) gx = cap;
  dim3 grid(gx, gy);
#define B200_NVFP4_LAUNCH(T, G)                                                                       \
  nvfp4_quant_kernel<T, G><<<grid, threads, 0, stream>>>((uint8_t*)out, (uint8_t*)out_scale,          \
                                                         (const T*)input, global_scale, M, hidden,    \
                                                         in_row_stride, is_sf_swizzled_layout)
  if (in_dtype == B200_BF16) {
    if (gated) B200_NVFP4_LAUNCH(__nv_bfloat16, true);
    else B200_NVFP4_LAUNCH(__nv_bfloat16, false);
  } else {
    if (gated) B200_NVFP4_LAUNCH(__half, true);
    else B200_NVFP4_LAUNCH(__half, false);
  }
#undef B200_NVFP4_LAUNCH
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_nvfp4_quant(void* out, void* out_scale, const void* input,
                                const float* global_scale, int64_t num_tokens, int hidden,
                                int64_t in_row_stride, int is_sf_swizzled_layout, int in_dtype,
                                b200_stream_t stream_) {
  return nvfp4_quant_launch(out, out_scale, input, global_scale, num_tokens, hidden, in_row_stride,
                            is_sf_swizzled_layout, in_dtype, false, (cudaStream_t)stream_,
                            "scaled_fp4_quant");
}

extern "C" int b200_silu_and_mul_nvfp4(void* out, void* out_scale, const void* input,
                                       const float* global_scale, int64_t num_tokens, int d,
                                       int64_t in_row_stride, int in_dtype, b200_stream_t stream_) {
  return nvfp4_quant_launch(out, out_scale, input, global_scale, num_tokens, d, in_row_stride, 1, in_dtype,
                            true, (cudaStream_t)stream_, "silu_and_mul_nvfp4_quant");
}

static int nvfp4_experts_quant_launch(void* out, void* out_scale, const void* input, const float* global_scale,
                                      const int32_t* expert_offsets, const int32_t* blockscale_offsets,
                                      int num_experts, int64_t m_topk, int k, int64_t in_row_stride, int in_dtype,
                                      bool gated, cudaStream_t stream, const char* op) {
  B200_CHECK(m_topk >= 0 && k > 0 && num_experts > 0, B200_ERR_INVALID, "%s: bad sizes", op);
  B200_CHECK(k % 16 == 0, B200_ERR_INVALID, "%s: k must be a multiple of 16, but got %d", op, k);
  B200_CHECK(in_dtype == B200_BF16 || in_dtype == B200_F16, B200_ERR_INVALID,
             "%s: input.dtype needs to be fp16 or bf16", op);
  B200_CHECK(num_experts <= 8192, B200_ERR_UNSUPPORTED, "%s: at most 8192 experts", op);
  if (m_topk == 0) return B200_OK;
  B200_CHECK(out && out_scale && input && expert_offsets && blockscale_offsets, B200_ERR_INVALID,
             "%s: null pointer", op);
  B200_CHECK(((uintptr_t)input & 15) == 0 && (in_row_stride * 2) % 16 == 0 && ((uintptr_t)out & 7) == 0,
             B200_ERR_INVALID, "%s: input rows must be 16-byte aligned", op);
  B200_CHECK(m_topk < (1LL << 31), B200_ERR_UNSUPPORTED, "%s: too many rows", op);
  const int M = (int)m_topk;
  const int sf_cols_pad = (k / 16 + 3) / 4 * 4;
  const int threads = sf_cols_pad < 256 ? (sf_cols_pad + 31) / 32 * 32 : 256;
  const int gy = (sf_cols_pad + threads - 1) / threads;
  int gx = M;
  const int sms = sm_count();
  B200_CHECK(sms > 0, B200_ERR_CUDA, "%s: no CUDA device (no CPU fallback)", op);
  const int cap = (sms * 8 + gy - 1) / gy;
  if (gx > cap) gx = cap;
  dim3 grid(gx, gy);
  const size_t smem = (size_t)(num_experts + 1) * sizeof(int32_t);
#define B200_NVFP4_LAUNCH(T, G)                                                                            \
  nvfp4_experts_quant_kernel<T, G><<<grid, threads, smem, stream>>>(                                       \
      (uint8_t*)out, (uint8_t*)out_scale, (const T*)input, global_scale, expert_offsets, blockscale_offsets, \
      num_experts, M, k, in_row_stride)
  if (in_dtype == B200_BF16) {
    if (gated) B200_NVFP4_LAUNCH(__nv_bfloat16, true);
    else B200_NVFP4_LAUNCH(__nv_bfloat16, false);
  } else {
    if (gated) B200_NVFP4_LAUNCH(__half, true);
    else B200_NVFP4_LAUNCH(__half, false);
  }
#undef B200_NVFP4_LAUNCH
  B200_LAUNCH_CHECK();
  return B200_OK;
}

extern "C" int b200_nvfp4_experts_quant(void* out, void* out_scale, const void* input,
                                        const float* global_scale, const int32_t* expert_offsets,
                                        const int32_t* blockscale_offsets, int num_experts, int64_t m_topk, int k,
                                        int64_t in_row_stride, int in_dtype, b200_stream_t stream_) {
  return nvfp4_experts_quant_launch(out, out_scale, input, global_scale, expert_offsets, blockscale_offsets,
                                    num_experts, m_topk, k, in_row_stride, in_dtype, false,
                                    (cudaStream_t)stream_, "scaled_fp4_experts_quant");
}

extern "C" int b200_silu_and_mul_nvfp4_experts_quant(void* out, void* out_scale, const void* input,
                                                     const float* global_scale, const int32_t* expert_offsets,
                                                     const int32_t* blockscale_offsets, int num_experts,
                                                     int64_t m_topk, int k, int64_t in_row_stride, int in_dtype,
                                                     b200_stream_t stream_) {
  return nvfp4_experts_quant_launch(out, out_scale, input, global_scale, expert_offsets, blockscale_offsets,
                                    num_experts, m_topk, k, in_row_stride, in_dtype, true,
                                    (cudaStream_t)stream_, "silu_and_mul_scaled_fp4_experts_quant");
}
