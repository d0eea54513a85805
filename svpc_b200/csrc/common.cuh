// Shared helpers for the svpc_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <math.h>

#define SVPC_OK 0
#define SVPC_ERR_ARG (-1)
#define SVPC_ERR_CUDA (-2)
#define SVPC_ERR_UNSUPPORTED (-3)

// activation codes shared by the GEMM epilogues (include/svpc_b200.h)
#define SVPC_ACT_NONE 0
#define SVPC_ACT_RELU 1
#define SVPC_ACT_GELU 2
#define SVPC_ACT_SIGMOID 3

void svpc_set_error(const char* fmt, ...);
int svpc_check_launch(const char* what);

#define SVPC_REQUIRE(cond, ...)                   \
  do {                                            \
    if (!(cond)) {                                \
      svpc_set_error(__VA_ARGS__);                \
      return SVPC_ERR_ARG;                        \
    }                                             \
  } while (0)

static inline int svpc_cdiv(long long a, long long b) { return (int)((a + b - 1) / b); }

__device__ __forceinline__ float svpc_gelu(float x) {
  // reference src/rtransformer/model.py:64  x * 0.5 * (1 + erf(x / sqrt(2)))
  return x * 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
}
__device__ __forceinline__ float svpc_gelu_grad(float x) {
  const float cdf = 0.5f * (1.0f + erff(x * 0.70710678118654752440f));
  const float pdf = 0.39894228040143267794f * expf(-0.5f * x * x);
  return cdf + x * pdf;
}
__device__ __forceinline__ float svpc_sigmoid(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float svpc_act(float v, int act) {
  switch (act) {
    case SVPC_ACT_RELU: return v > 0.f ? v : 0.f;
    case SVPC_ACT_GELU: return svpc_gelu(v);
    case SVPC_ACT_SIGMOID: return svpc_sigmoid(v);
    default: return v;
  }
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
  return v;
}

// Block-wide sum; `red` must hold >= 32 floats of shared memory. All threads get the result.
__device__ __forceinline__ float block_sum(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : 0.f;
  r = warp_sum(r);
  return r;
}
__device__ __forceinline__ float block_max(float v, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_max(v);
  __syncthreads();
  if (lane == 0) red[w] = v;
  __syncthreads();
  float r = (lane < nw) ? red[lane] : -INFINITY;
  r = warp_max(r);
  return r;
}

// Counter-based dropout mask: keep iff hash(seed, idx) >= p * 2^32.  Same function in fwd and bwd.
__device__ __forceinline__ uint32_t svpc_hash32(uint64_t seed, uint64_t idx) {
  uint64_t z = seed + idx * 0x9E3779B97F4A7C15ull;
  z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
  z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
  z = z ^ (z >> 31);
  return (uint32_t)(z >> 32);
}
__device__ __forceinline__ float svpc_dropout_scale(uint64_t seed, uint64_t idx, uint32_t thresh, float inv_keep) {
  return (svpc_hash32(seed, idx) >= thresh) ? inv_keep : 0.f;
}
