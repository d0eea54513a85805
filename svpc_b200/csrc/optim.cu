// Fused training-iteration tail: global gradient clipping (src/train.py:141-142), BertAdam.step
// (src/rtransformer/optimization.py:284-331: per-parameter clip to max_grad_norm, Adam moments without bias correction,
// decoupled weight decay, scheduled learning rate) and the EMA shadow update (optimization.py:196-203) as three
// multi-tensor launches over ~180 separately allocated parameters instead of ~181 x ~12 library kernels:
//
//   optim_sqnorm   one pass over every gradient: per-chunk sum of squares (fixed order -> deterministic)
//   optim_coef     one CTA: per-tensor and total norms, combined clip coefficient per tensor
//   optim_apply    one pass over g, p, m, v, shadow: 20 B read + 16 B written per element, HBM-bound
//
// Tensors are addressed through device arrays of device pointers (they are torch allocations, not one flat buffer);
// a chunk table (tensor index, first element) gives every CTA a contiguous slice of one tensor.  Chunks of one tensor
// are consecutive in the table (tensor_first[t] .. tensor_first[t+1]).
#include "common.cuh"

#define OPT_THREADS 256

__device__ __forceinline__ float block_sum_256(float v, float* red) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  const int w = threadIdx.x >> 5, l = threadIdx.x & 31;
  if (l == 0) red[w] = v;
  __syncthreads();
  float t = 0.f;
  if (threadIdx.x < 32) {
    t = (threadIdx.x < OPT_THREADS / 32) ? red[threadIdx.x] : 0.f;
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) t += __shfl_xor_sync(0xffffffffu, t, o);
  }
  return t;   // valid in thread 0
}

__global__ void __launch_bounds__(OPT_THREADS) optim_sqnorm_kernel(const int* __restrict__ chunk_tensor,
                                                                   const long long* __restrict__ chunk_start, int chunk_elems,
                                                                   const float* const* __restrict__ grads,
                                                                   const long long* __restrict__ numel,
                                                                   float* __restrict__ chunk_sq) {
  svpc_pdl_sync();
  __shared__ float red[OPT_THREADS / 32];
  const int c = blockIdx.x, t = chunk_tensor[c];
  const long long s = chunk_start[c];
  const long long rem = numel[t] - s;
  const int n = (int)(rem < chunk_elems ? rem : chunk_elems);
  const float* g = grads[t] + s;
  float acc = 0.f;
  if (grads[t] == nullptr) {          // parameter without a gradient this step (BertAdam skips it, optimization.py:286-287)
  } else if ((((uintptr_t)g) & 15) == 0) {
    const int n4 = n >> 2;
    const float4* g4 = (const float4*)g;
    for (int i = threadIdx.x; i < n4; i += OPT_THREADS) {
      const float4 x = __ldg(g4 + i);
      acc += x.x * x.x + x.y * x.y + x.z * x.z + x.w * x.w;
    }
    for (int i = (n4 << 2) + threadIdx.x; i < n; i += OPT_THREADS) acc += g[i] * g[i];
  } else {
    for (int i = threadIdx.x; i < n; i += OPT_THREADS) acc += g[i] * g[i];
  }
  const float tot = block_sum_256(acc, red);
  if (threadIdx.x == 0) chunk_sq[c] = tot;
}

// coef[t] = c_global * c_param(t):  c_global = min(1, global_max / (||g||_all + 1e-6))  (torch clip_grad_norm_),
// c_param = min(1, param_max / (c_global * ||g_t|| + 1e-6)) (BertAdam's own clip, applied to the already scaled gradient).
// A non-positive max disables that clip.  norms[0] = total norm before clipping (what clip_grad_norm_ returns).
__global__ void __launch_bounds__(OPT_THREADS) optim_coef_kernel(int n_tensors, const int* __restrict__ tensor_first,
                                                                 const float* __restrict__ chunk_sq, float global_max,
                                                                 float param_max, float* __restrict__ coef,
                                                                 float* __restrict__ norms) {
  svpc_pdl_sync();
  __shared__ double red[OPT_THREADS];
  __shared__ float s_cg;
  double mine = 0.0;
  for (int t = threadIdx.x; t < n_tensors; t += OPT_THREADS) {
    double s = 0.0;
    for (int c = tensor_first[t]; c < tensor_first[t + 1]; ++c) s += (double)chunk_sq[c];
    coef[t] = (float)s;       // parked: squared norm of tensor t
    mine += s;
  }
  red[threadIdx.x] = mine;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
    for (int i = 0; i < OPT_THREADS; ++i) tot += red[i];
    const float total = (float)sqrt(tot);
    float cg = 1.f;
    if (global_max > 0.f) cg = fminf(global_max / (total + 1e-6f), 1.f);
    s_cg = cg;
    norms[0] = total;
    norms[1] = cg;
  }
  __syncthreads();
  const float cg = s_cg;
  for (int t = threadIdx.x; t < n_tensors; t += OPT_THREADS) {
    const float nt = cg * sqrtf(coef[t]);
    float cp = 1.f;
    if (param_max > 0.f) cp = fminf(param_max / (nt + 1e-6f), 1.f);
    coef[t] = cg * cp;
  }
}

struct OptimPtrs {
  float* const* grads;
  float* const* params;
  float* const* m;
  float* const* v;
  float* const* shadow;   // NULL table entry or NULL table: no EMA
};

// flags: bit0 = Adam update, bit1 = EMA update, bit2 = write the clipped gradient back (the reference clips in place)
template <int VEC>
__device__ __forceinline__ void optim_apply_elems(int n, float* __restrict__ g, float* __restrict__ p, float* __restrict__ m,
                                                  float* __restrict__ v, float* __restrict__ sh, float c, float lr, float wd,
                                                  float b1, float omb1, float b2, float omb2, float eps, float decay,
                                                  float omdecay, int flags) {
  const bool adam = flags & 1, ema = (flags & 2) && sh != nullptr, wb = flags & 4;
  const int nv = n / VEC;
  for (int i = threadIdx.x; i < nv; i += OPT_THREADS) {
    __align__(16) float gx[VEC], px[VEC], mx[VEC], vx[VEC], sx[VEC];
    if constexpr (VEC == 4) {
      *(float4*)px = ((const float4*)p)[i];
      if (adam) {
        *(float4*)gx = ((const float4*)g)[i];
        *(float4*)mx = ((const float4*)m)[i];
        *(float4*)vx = ((const float4*)v)[i];
      }
      if (ema) *(float4*)sx = ((const float4*)sh)[i];
    } else {
      px[0] = p[i];
      if (adam) { gx[0] = g[i]; mx[0] = m[i]; vx[0] = v[i]; }
      if (ema) sx[0] = sh[i];
    }
#pragma unroll
    for (int k = 0; k < VEC; ++k) {
      if (adam) {
        const float gg = gx[k] * c;
        gx[k] = gg;
        mx[k] = mx[k] * b1 + omb1 * gg;
        vx[k] = vx[k] * b2 + omb2 * gg * gg;
        float upd = mx[k] / (sqrtf(vx[k]) + eps);
        if (wd > 0.f) upd += wd * px[k];
        px[k] = px[k] - lr * upd;
      }
      if (ema) sx[k] = omdecay * px[k] + decay * sx[k];
    }
    if constexpr (VEC == 4) {
      if (adam) {
        ((float4*)p)[i] = *(float4*)px;
        ((float4*)m)[i] = *(float4*)mx;
        ((float4*)v)[i] = *(float4*)vx;
        if (wb) ((float4*)g)[i] = *(float4*)gx;
      }
      if (ema) ((float4*)sh)[i] = *(float4*)sx;
    } else {
      if (adam) { p[i] = px[0]; m[i] = mx[0]; v[i] = vx[0]; if (wb) g[i] = gx[0]; }
      if (ema) sh[i] = sx[0];
    }
  }
}

__global__ void __launch_bounds__(OPT_THREADS) optim_apply_kernel(const int* __restrict__ chunk_tensor,
                                                                  const long long* __restrict__ chunk_start, int chunk_elems,
                                                                  const OptimPtrs ptrs, const long long* __restrict__ numel,
                                                                  const float* __restrict__ coef,
                                                                  const float* __restrict__ tensor_lr,
                                                                  const float* __restrict__ tensor_wd, float b1, float omb1,
                                                                  float b2, float omb2, float eps,
                                                                  const float* __restrict__ ema_decay, int flags) {
  svpc_pdl_sync();
  const int c = blockIdx.x, t = chunk_tensor[c];
  const long long s = chunk_start[c];
  const long long rem = numel[t] - s;
  const int n = (int)(rem < chunk_elems ? rem : chunk_elems);
  const bool adam = (flags & 1) && ptrs.grads[t] != nullptr;      // no gradient: Adam skipped, EMA still applies
  if (!adam) flags &= ~1;
  float* g = adam ? ptrs.grads[t] + s : nullptr;
  float* p = ptrs.params[t] + s;
  float* m = adam ? ptrs.m[t] + s : nullptr;
  float* v = adam ? ptrs.v[t] + s : nullptr;
  float* sh = (ptrs.shadow != nullptr && ptrs.shadow[t] != nullptr) ? ptrs.shadow[t] + s : nullptr;
  const float cf = adam ? coef[t] : 1.f, lr = adam ? tensor_lr[t] : 0.f, wd = adam ? tensor_wd[t] : 0.f;
  const float decay = ema_decay != nullptr ? ema_decay[0] : 0.f, omdecay = ema_decay != nullptr ? ema_decay[1] : 1.f;
  const uintptr_t al = (uintptr_t)p | (uintptr_t)g | (uintptr_t)m | (uintptr_t)v | (uintptr_t)sh;
  if ((al & 15) == 0) {
    const int n4 = n & ~3;
    optim_apply_elems<4>(n4, g, p, m, v, sh, cf, lr, wd, b1, omb1, b2, omb2, eps, decay, omdecay, flags);
    if (n4 < n)
      optim_apply_elems<1>(n - n4, adam ? g + n4 : g, p + n4, adam ? m + n4 : m, adam ? v + n4 : v, sh ? sh + n4 : sh, cf, lr,
                           wd, b1, omb1, b2, omb2, eps, decay, omdecay, flags);
  } else {
    optim_apply_elems<1>(n, g, p, m, v, sh, cf, lr, wd, b1, omb1, b2, omb2, eps, decay, omdecay, flags);
  }
}

extern "C" int svpc_optim_sqnorm(int n_chunks, const int* chunk_tensor, const long long* chunk_start, int chunk_elems,
                                 const void* const* grads, const long long* numel, float* chunk_sq, svpc_stream_t stream) {
  if (n_chunks <= 0) return SVPC_OK;
  if (chunk_elems <= 0) {
    svpc_set_error("optim_sqnorm: chunk_elems must be positive");
    return SVPC_ERR_ARG;
  }
  svpc_launch(optim_sqnorm_kernel, n_chunks, OPT_THREADS, 0, (cudaStream_t)stream, chunk_tensor, chunk_start, chunk_elems,
              (const float* const*)grads, numel, chunk_sq);
  return svpc_check_launch("optim_sqnorm");
}

extern "C" int svpc_optim_coef(int n_tensors, const int* tensor_first, const float* chunk_sq, float global_max_norm,
                               float param_max_norm, float* coef, float* norms, svpc_stream_t stream) {
  if (n_tensors <= 0) return SVPC_OK;
  svpc_launch(optim_coef_kernel, 1, OPT_THREADS, 0, (cudaStream_t)stream, n_tensors, tensor_first, chunk_sq, global_max_norm,
              param_max_norm, coef, norms);
  return svpc_check_launch("optim_coef");
}

extern "C" int svpc_optim_apply(int n_chunks, const int* chunk_tensor, const long long* chunk_start, int chunk_elems,
                                void* const* grads, void* const* params, void* const* m, void* const* v, void* const* shadow,
                                const long long* numel, const float* coef, const float* tensor_lr, const float* tensor_wd,
                                double b1, double b2, double eps, const float* ema_decay, int flags, svpc_stream_t stream) {
  if (n_chunks <= 0) return SVPC_OK;
  if ((flags & 1) && (!grads || !m || !v || !coef || !tensor_lr || !tensor_wd)) {
    svpc_set_error("optim_apply: the Adam update needs grads, m, v, coef, lr and weight-decay tables");
    return SVPC_ERR_ARG;
  }
  if ((flags & 2) && (!shadow || !ema_decay)) {
    svpc_set_error("optim_apply: the EMA update needs the shadow table and the decay scalar");
    return SVPC_ERR_ARG;
  }
  OptimPtrs ptrs{(float* const*)grads, (float* const*)params, (float* const*)m, (float* const*)v, (float* const*)shadow};
  svpc_launch(optim_apply_kernel, n_chunks, OPT_THREADS, 0, (cudaStream_t)stream, chunk_tensor, chunk_start, chunk_elems, ptrs,
              numel, coef, tensor_lr, tensor_wd, (float)b1, (float)(1.0 - b1), (float)b2, (float)(1.0 - b2), (float)eps, ema_decay, flags);
  return svpc_check_launch("optim_apply");
}
