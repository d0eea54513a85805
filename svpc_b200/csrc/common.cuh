// Shared helpers for the svpc_b200 CUDA library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>
#include "../../include/svpc_b200.h"

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
static inline int svpc_imin(long long a, long long b) { return (int)(a < b ? a : b); }

#define SVPC_NUM_SMS 148

// ---- programmatic dependent launch (PDL).  The step is ~450 short kernels replayed from a CUDA graph; between two
// dependent kernels the GPU otherwise drains, then schedules and launches the next grid.  Every kernel of this library is
// launched with cudaLaunchAttributeProgrammaticStreamSerialization and starts with svpc_pdl_sync():
//   griddepcontrol.launch_dependents  the NEXT kernel's CTAs may be made resident as soon as all of this grid's CTAs are
//   griddepcontrol.wait               block until every prerequisite grid has completed and its memory is visible
// so nothing is read or written before the producer has finished, but launch latency and the tail of the previous grid
// overlap.  SVPC_PDL=0 in the environment falls back to plain stream order.
#if defined(__CUDACC__)
__device__ __forceinline__ void svpc_pdl_sync() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}
#endif
static inline int svpc_pdl_enabled() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("SVPC_PDL");
    on = e ? (atoi(e) != 0) : 1;
  }
  return on;
}
template <typename... KArgs, typename... Args>
static inline cudaError_t svpc_launch(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                                      Args... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = st;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = svpc_pdl_enabled() ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, static_cast<KArgs>(args)...);
}

// ---- typed loads / stores (activations are fp32 or bf16; math is always fp32)
typedef __nv_bfloat16 bf16;
__device__ __forceinline__ float ldf(const float* p) { return *p; }
__device__ __forceinline__ float ldf(const bf16* p) { return __bfloat162float(*p); }
__device__ __forceinline__ void stf(float* p, float v) { *p = v; }
__device__ __forceinline__ void stf(bf16* p, float v) { *p = __float2bfloat16_rn(v); }

// dtype dispatch: DT is SVPC_F32 / SVPC_BF16, binds type T
#define SVPC_DISPATCH(DT, T, ...)                                       \
  do {                                                                  \
    if ((DT) == SVPC_F32) { typedef float T; __VA_ARGS__; }             \
    else if ((DT) == SVPC_BF16) { typedef bf16 T; __VA_ARGS__; }        \
    else { svpc_set_error("bad dtype code %d", (int)(DT)); return SVPC_ERR_ARG; } \
  } while (0)

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

// ---- dropout (train mode): counter-based mask, recomputed wherever it is needed (forward, backward) instead of stored.
// keep(seed, site, idx) is a pure function of the step seed (read from device memory so that a captured CUDA graph sees
// a fresh seed every replay), the call site and the element index; kept elements are scaled by 1 / (1 - p).
struct SvpcDrop {
  const unsigned int* seed;   // device pointer, NULL = no dropout
  unsigned int site;
  float p;
};
__device__ __forceinline__ uint32_t svpc_mix32(uint32_t h) {
  h ^= h >> 16; h *= 0x85ebca6bu; h ^= h >> 13; h *= 0xc2b2ae35u; h ^= h >> 16;
  return h;
}
__device__ __forceinline__ uint32_t svpc_drop_key(const SvpcDrop& d) {
  return d.seed ? svpc_mix32(*d.seed * 0x9e3779b9u + d.site * 0x7f4a7c15u + 0x1234567u) : 0u;
}
// One 32-bit hash serves the element pair (idx & ~1, idx | 1): 16 bits each (p is resolved to 1/65536).
__device__ __forceinline__ uint32_t svpc_drop_hash(uint32_t key, unsigned long long pair) {
  return svpc_mix32((uint32_t)pair * 0x9e3779b1u + (uint32_t)(pair >> 32) * 0x85ebca6bu + key);
}
__device__ __forceinline__ uint32_t svpc_drop_thresh(float p) { return (uint32_t)(p * 65536.0f); }
// 0 for a dropped element, 1 / (1 - p) for a kept one
__device__ __forceinline__ float svpc_drop_scale(uint32_t key, float p, float inv_keep, unsigned long long idx) {
  const uint32_t h = svpc_drop_hash(key, idx >> 1);
  return ((h >> ((idx & 1) * 16)) & 0xffffu) >= svpc_drop_thresh(p) ? inv_keep : 0.f;
}
// both elements of an aligned pair (idx even) with one hash
__device__ __forceinline__ void svpc_drop_scale2(uint32_t key, float p, float inv_keep, unsigned long long idx_even,
                                                 float& s0, float& s1) {
  const uint32_t h = svpc_drop_hash(key, idx_even >> 1), t = svpc_drop_thresh(p);
  s0 = (h & 0xffffu) >= t ? inv_keep : 0.f;
  s1 = (h >> 16) >= t ? inv_keep : 0.f;
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
// two sums at once
__device__ __forceinline__ float2 block_sum2(float a, float b, float* red) {
  const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  a = warp_sum(a); b = warp_sum(b);
  __syncthreads();
  if (lane == 0) { red[w] = a; red[32 + w] = b; }
  __syncthreads();
  float ra = (lane < nw) ? red[lane] : 0.f;
  float rb = (lane < nw) ? red[32 + lane] : 0.f;
  return make_float2(warp_sum(ra), warp_sum(rb));
}
