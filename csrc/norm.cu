// K6 norm_residual: (x [+ x2] [+ residual]) -> new residual, LayerNorm / RMSNorm of it.
//
// Replaces the reference's separate ATen add + layer_norm launches (SURVEY.md G2, G13;
// gptj_modeling.py:296,310,496; gpt_bigcode_modeling.py:367,379,404,407) and provides the
// capability of the never-instantiated FastLayerNorm wrapper (layers.py:217-256, G18).
// One CTA per row, row held in registers (16-byte vector loads), fp32 statistics,
// two-pass variance.  Memory-bound: reads x(+res) once, writes normed(+res) once.
#include "common.cuh"

namespace {

constexpr int kThreads = 256;
constexpr int kMaxVec = 4;  // 256 thr * 4 vec * 8 elem = H <= 8192

__device__ __forceinline__ float block_sum(float v, float* red) {
  v = warp_sum(v);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  __syncthreads();
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = (l < kThreads / 32) ? red[l] : 0.f;
  return warp_sum(t);
}

template <typename T>
__global__ void __launch_bounds__(kThreads) norm_residual_kernel(
    const T* __restrict__ x, const T* __restrict__ x2, const T* residual_in, T* residual_out,
    const T* __restrict__ weight, const T* __restrict__ bias, T* __restrict__ out, int H, int64_t ldx,
    int64_t ldx2, int64_t ldr, int64_t ldo, float eps, int kind) {
  pdl_launch_dependents();
  pdl_wait();
  __shared__ float red[kThreads / 32];
  const int row = blockIdx.x;
  const int nvec = H >> 3;
  float v[kMaxVec][8];
  float sum = 0.f;
#pragma unroll
  for (int i = 0; i < kMaxVec; ++i) {
    const int c = threadIdx.x + i * kThreads;
    if (c < nvec) {
      uint4 raw = *reinterpret_cast<const uint4*>(x + row * ldx + c * 8);
      unpack8<T>(raw, v[i]);
      if (x2 != nullptr) {
        float t[8];
        raw = *reinterpret_cast<const uint4*>(x2 + row * ldx2 + c * 8);
        unpack8<T>(raw, t);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] += t[j];
      }
      if (residual_in != nullptr) {
        float t[8];
        raw = *reinterpret_cast<const uint4*>(residual_in + row * ldr + c * 8);
        unpack8<T>(raw, t);
#pragma unroll
        for (int j = 0; j < 8; ++j) v[i][j] += t[j];
      }
      if (residual_out != nullptr) {
        *reinterpret_cast<uint4*>(residual_out + row * ldr + c * 8) = pack8<T>(v[i]);
        // the normalisation sees the value as stored (same rounding as the unfused path)
        unpack8<T>(pack8<T>(v[i]), v[i]);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) sum += (kind == LLMSS_NORM_RMS) ? v[i][j] * v[i][j] : v[i][j];
    }
  }
  float mean = 0.f, rstd;
  if (kind == LLMSS_NORM_RMS) {
    float ss = block_sum(sum, red);
    rstd = rsqrtf(ss / (float)H + eps);
  } else {
    mean = block_sum(sum, red) / (float)H;
    float sq = 0.f;
#pragma unroll
    for (int i = 0; i < kMaxVec; ++i) {
      const int c = threadIdx.x + i * kThreads;
      if (c < nvec) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
          float d = v[i][j] - mean;
          sq += d * d;
        }
      }
    }
    float var = block_sum(sq, red) / (float)H;
    rstd = rsqrtf(var + eps);
  }
#pragma unroll
  for (int i = 0; i < kMaxVec; ++i) {
    const int c = threadIdx.x + i * kThreads;
    if (c < nvec) {
      float w[8], b[8], o[8];
      unpack8<T>(*reinterpret_cast<const uint4*>(weight + c * 8), w);
      if (bias != nullptr) unpack8<T>(*reinterpret_cast<const uint4*>(bias + c * 8), b);
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        float y = (v[i][j] - mean) * rstd * w[j];
        o[j] = (bias != nullptr) ? y + b[j] : y;
      }
      *reinterpret_cast<uint4*>(out + row * ldo + c * 8) = pack8<T>(o);
    }
  }
}

}  // namespace

LLMSS_API int llmss_norm_residual(const void* x, const void* x2, const void* residual_in, void* residual_out,
                                  const void* weight, const void* bias, void* out, int M, int H, int64_t ldx,
                                  int64_t ldx2, int64_t ldr, int64_t ldo, float eps, int kind, int dtype,
                                  cudaStream_t stream) {
  LLMSS_REQUIRE(H % 8 == 0 && H <= kThreads * kMaxVec * 8, "norm_residual: H=%d must be a multiple of 8 and <= %d", H,
                kThreads * kMaxVec * 8);
  LLMSS_REQUIRE(dtype == LLMSS_F16 || dtype == LLMSS_BF16, "norm_residual: dtype must be f16/bf16");
  LLMSS_REQUIRE(ldx % 8 == 0 && ldo % 8 == 0 && ldr % 8 == 0 && ldx2 % 8 == 0, "norm_residual: row strides must be multiples of 8");
  if (M == 0) return 0;
  if (dtype == LLMSS_F16)
    LLMSS_CHECK_CUDA(launch_pdl(norm_residual_kernel<__half>, dim3(M), dim3(kThreads), 0, stream, (const __half*)x, (const __half*)x2, (const __half*)residual_in, (__half*)residual_out, (const __half*)weight, (const __half*)bias, (__half*)out, H, ldx, ldx2, ldr, ldo, eps, kind));
  else
    LLMSS_CHECK_CUDA(launch_pdl(norm_residual_kernel<__nv_bfloat16>, dim3(M), dim3(kThreads), 0, stream, (const __nv_bfloat16*)x, (const __nv_bfloat16*)x2, (const __nv_bfloat16*)residual_in, (__nv_bfloat16*)residual_out, (const __nv_bfloat16*)weight, (const __nv_bfloat16*)bias, (__nv_bfloat16*)out, H, ldx, ldx2, ldr, ldo, eps, kind));
  LLMSS_CHECK_CUDA(cudaGetLastError());
  return 0;
}
