// b200ops -- per-token-group FP8 quantiser (the 1 x group_size activation scales of block-scaled FP8
// linears, DeepSeek-style).
//
// Behavioural contract (bit-exact): torch.ops._C.per_token_group_fp8_quant,
// csrc/libtorch_stable/quantization/w8a8/fp8/per_token_group_quant.cu:43-161 of the reference:
//   absmax = fmaxf-reduction of |x| over the group, starting from eps (NaN dropped)
//   y_s    = absmax / fp8_max                      (IEEE division)
//   ue8m0  : y_s = exp2f(ceilf(log2f(fmaxf(|y_s|, 1e-10f))))
//   q      = fminf(fmaxf(x / y_s, fp8_min), fp8_max) -> e4m3 (RNE)   (true division; NaN -> fp8_min)
//   scale (row r, group c) written at scales[r * stride_row + c * stride_col]  (row-major,
//   column-major and TMA-aligned column-major are all stride choices of the caller,
//   fp8_utils.py:575-597).
//
// B200 design: bandwidth kernel, algorithmic bytes M*K*(b_in + 1) + 4*M*K/g. A sub-warp of LANES lanes
// owns one group and keeps it in registers between the abs-max shuffle reduction and the conversion
// (one HBM read; the reference stages the group through shared memory). 16-byte loads, 8/4-byte stores.
// Grid = ceil(groups / groups-per-CTA): tens of thousands of CTAs at decode sizes, many waves per SM.
#include "common.cuh"

namespace b200 {

// x / y, correctly rounded, from the correctly rounded reciprocal r = RN(1 / y): q0 = RN(x r) is a faithful quotient,
// the remainder x - q0 y is exact in one FMA, and RN(q0 + rem r) is the correctly rounded quotient (Markstein's
// division theorem; no overflow / underflow can occur here: |x| <= 448 y up to the clamp, y >= eps / 448, and
// quotients small enough for the remainder to underflow round to e4m3 zero either way). 3 issue slots instead of
// the ~12 of __fdiv_rn, which made the kernel issue-bound (2.6 TB/s); bit-exactness vs the oracle's IEEE division
// and vs the wheel's kernel is what tests/test_gpu_fp8_quant.py and test_gpu_vs_wheel.py check.
__device__ __forceinline__ float div_by(float x, float y, float r) {
  const float q0 = __fmul_rn(x, r);
  const float rem = __fmaf_rn(-q0, y, x);
  const float q1 = __fmaf_rn(rem, r, q0);
  // inf / nan inputs: fall back to the plain quotient semantics (inf * r = inf, rem = nan)
  return (q0 - q0 == 0.f) ? q1 : __fdiv_rn(x, y);
}

template <typename in_t, int VPL /* 16-byte vectors per lane */, int LANES>
__global__ void __launch_bounds__(256)
    fp8_quant_token_group_kernel(uint8_t* __restrict__ out, float* __restrict__ scales,
                                 const in_t* __restrict__ in, int64_t num_groups, int groups_per_row,
                                 int group_size, int64_t in_row_stride, int64_t out_row_stride,
                                 int64_t s_stride_row, int64_t s_stride_col, float eps, float fp8_min,
                                 float fp8_max, int ue8m0) {
  constexpr int N = 16 / sizeof(in_t);
  const int sub = threadIdx.x / LANES, lane = threadIdx.x % LANES;
  const int64_t gid = (int64_t)blockIdx.x * (256 / LANES) + sub;
  const bool live = gid < num_groups;
  const int64_t row = live ? gid / groups_per_row : 0;
  const int gcol = live ? (int)(gid % groups_per_row) : 0;
  const in_t* src = in + row * in_row_stride + (int64_t)gcol * group_size;
  uint8_t* dst = out + row * out_row_stride + (int64_t)gcol * group_size;
  const int nvec = group_size / N;

  uint4 v[VPL];
  float m = eps;
#pragma unroll
  for (int i = 0; i < VPL; i++) {
    const int vi = lane + i * LANES;
    if (live && vi < nvec) {
      v[i] = ld_nc_v4(src + (int64_t)vi * N);
      const in_t* e = reinterpret_cast<const in_t*>(&v[i]);
#pragma unroll
      for (int j = 0; j < N; j++) m = fmaxf(m, fabsf(Cvt<in_t>::to_f32(e[j])));
    }
  }
#pragma unroll
  for (int o = LANES / 2; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
  float y_s = __fdiv_rn(m, fp8_max);
  if (ue8m0) y_s = exp2f(ceilf(log2f(fmaxf(fabsf(y_s), 1e-10f))));
  if (live && lane == 0) scales[row * s_stride_row + (int64_t)gcol * s_stride_col] = y_s;
  const float r_s = __frcp_rn(y_s);
#pragma unroll
  for (int i = 0; i < VPL; i++) {
    const int vi = lane + i * LANES;
    if (live && vi < nvec) {
      const in_t* e = reinterpret_cast<const in_t*>(&v[i]);
      uint16_t o[N / 2];
#pragma unroll
      for (int j = 0; j < N / 2; j++) {
        float a = fminf(fmaxf(div_by(Cvt<in_t>::to_f32(e[2 * j]), y_s, r_s), fp8_min), fp8_max);
        float b = fminf(fmaxf(div_by(Cvt<in_t>::to_f32(e[2 * j + 1]), y_s, r_s), fp8_min), fp8_max);
        o[j] = f32x2_to_e4m3x2(a, b);
      }
      if constexpr (N == 8)
        *reinterpret_cast<uint2*>(dst + (int64_t)vi * N) = *reinterpret_cast<const uint2*>(o);
      else
        *reinterpret_cast<uint32_t*>(dst + (int64_t)vi * N) = *reinterpret_cast<const uint32_t*>(o);
    }
  }
}

// Any group size / alignment: one warp per group, scalar accesses, the group is read twice.
template <typename in_t>
__global__ void __launch_bounds__(256)
    fp8_quant_token_group_scalar_kernel(uint8_t* __restrict__ out, float* __restrict__ scales,
                                        const in_t* __restrict__ in, int64_t num_groups,
                                        int groups_per_row, int group_size, int64_t in_row_stride,
                                        int64_t out_row_stride, int64_t s_stride_row,
                                        int64_t s_stride_col, float eps, float fp8_min, float fp8_max,
                                        int ue8m0) {
  const int lane = threadIdx.x & 31;
  const int64_t gid = (int64_t)blockIdx.x * 8 + (threadIdx.x >> 5);
  if (gid >= num_groups) return;  // warp-uniform
  const int64_t row = gid / groups_per_row;
  const int gcol = (int)(gid % groups_per_row);
  const in_t* src = in + row * in_row_stride + (int64_t)gcol * group_size;
  uint8_t* dst = out + row * out_row_stride + (int64_t)gcol * group_size;
  float m = eps;
  for (int i = lane; i < group_size; i += 32) m = fmaxf(m, fabsf(Cvt<in_t>::to_f32(src[i])));
  m = warp_max(m);
  float y_s = __fdiv_rn(m, fp8_max);
  if (ue8m0) y_s = exp2f(ceilf(log2f(fmaxf(fabsf(y_s), 1e-10f))));
  if (lane == 0) scales[row * s_stride_row + (int64_t)gcol * s_stride_col] = y_s;
  for (int i = lane; i < group_size; i += 32)
    dst[i] = f32_to_e4m3(fminf(fmaxf(__fdiv_rn(Cvt<in_t>::to_f32(src[i]), y_s), fp8_min), fp8_max));
}

}  // namespace b200

using namespace b200;

#define B200_DISPATCH_IN(dt, ...)                                         \
  switch (dt) {                                                           \
    case B200_BF16: { using in_t = __nv_bfloat16; __VA_ARGS__; break; }   \
    case B200_F16: { using in_t = __half; __VA_ARGS__; break; }           \
    default: { using in_t = float; __VA_ARGS__; break; }                  \
  }

extern "C" int b200_fp8_quant_token_group(void* out, float* scales, const void* input,
                                          int64_t num_tokens, int hidden, int group_size,
                                          int64_t in_row_stride, int64_t out_row_stride,
                                          int64_t scale_stride_row, int64_t scale_stride_col, float eps,
                                          float fp8_min, float fp8_max, int scale_ue8m0, int in_dtype,
                                          b200_stream_t stream_) {
  cudaStream_t stream = (cudaStream_t)stream_;
  B200_CHECK(in_dtype == B200_BF16 || in_dtype == B200_F16 || in_dtype == B200_F32, B200_ERR_INVALID,
             "per_token_group_fp8_quant: input must be bf16 / f16 / f32 (dtype %d)", in_dtype);
  B200_CHECK(num_tokens >= 0 && hidden >= 0, B200_ERR_INVALID,
             "per_token_group_fp8_quant: negative shape");
  B200_CHECK(group_size > 0 && hidden % group_size == 0, B200_ERR_INVALID,
             "per_token_group_fp8_quant: the last dimension %d must be divisible by group_size %d",
             hidden, group_size);
  if (num_tokens == 0 || hidden == 0) return B200_OK;
  B200_CHECK(out && scales && input, B200_ERR_INVALID, "per_token_group_fp8_quant: null pointer");
  B200_CHECK(in_row_stride >= hidden && out_row_stride >= hidden, B200_ERR_INVALID,
             "per_token_group_fp8_quant: row strides smaller than the row");
  B200_CHECK(fp8_max > 0.f && fp8_min < 0.f, B200_ERR_INVALID,
             "per_token_group_fp8_quant: bad fp8 range [%g, %g]", fp8_min, fp8_max);
  const int isz = dtype_size(in_dtype);
  const int N = 16 / isz;
  const int groups_per_row = hidden / group_size;
  const int64_t num_groups = num_tokens * groups_per_row;
  const bool vec = group_size % N == 0 && (uintptr_t)input % 16 == 0 && (uintptr_t)out % 8 == 0 &&
                   (in_row_stride * isz) % 16 == 0 && out_row_stride % 8 == 0;
  const int nvec = vec ? group_size / N : 0;
#define LAUNCH_G(VPL, LANES)                                                                          \
  fp8_quant_token_group_kernel<in_t, VPL, LANES>                                                      \
      <<<(unsigned)((num_groups + (256 / LANES) - 1) / (256 / LANES)), 256, 0, stream>>>(             \
          (uint8_t*)out, scales, (const in_t*)input, num_groups, groups_per_row, group_size,          \
          in_row_stride, out_row_stride, scale_stride_row, scale_stride_col, eps, fp8_min, fp8_max,   \
          scale_ue8m0)
  B200_DISPATCH_IN(in_dtype, {
    if (vec && nvec <= 8) LAUNCH_G(1, 8);
    else if (vec && nvec <= 16) LAUNCH_G(1, 16);
    else if (vec && nvec <= 32) LAUNCH_G(1, 32);
    else if (vec && nvec <= 64) LAUNCH_G(2, 32);
    else if (vec && nvec <= 128) LAUNCH_G(4, 32);
    else
      fp8_quant_token_group_scalar_kernel<in_t><<<(unsigned)((num_groups + 7) / 8), 256, 0, stream>>>(
          (uint8_t*)out, scales, (const in_t*)input, num_groups, groups_per_row, group_size,
          in_row_stride, out_row_stride, scale_stride_row, scale_stride_col, eps, fp8_min, fp8_max,
          scale_ue8m0);
  });
#undef LAUNCH_G
  B200_LAUNCH_CHECK();
  return B200_OK;
}
