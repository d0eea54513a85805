// Bandwidth-bound row-wise kernels: LayerNorm (+residual, +position / token-type rows), activation backward + bias
// gradient, embedding gather + LN, video-feature LN staging, ingredient segment means.
// Layout rule: a CTA owns whole rows, threads stride the contiguous column dimension (coalesced), row statistics by
// warp shuffles; column reductions (dgamma / dbeta / dbias) are kept in registers across a grid-stride loop over rows
// and flushed with one red.global per column per CTA.
#include "common.cuh"
#include <mutex>

#define RW_THREADS 256
#define RW_MAXPT 16          // columns cached per thread (D <= RW_THREADS * RW_MAXPT = 4096)

static inline int rw_grid(long long rows) { return svpc_imin(rows, (long long)SVPC_NUM_SMS * 8); }

// ------------------------------------------------------------------------------------------------
// activation backward + column sum
// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void act_bwd_kernel(int rows, int cols, const T* __restrict__ dy, const T* __restrict__ ref, int act,
                               T* __restrict__ dz, float* __restrict__ dbias, const SvpcDrop drop) {
  svpc_pdl_sync();
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(drop);
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  // blockDim = (32, 8); a CTA owns a 32-column strip for a slice of rows
  const int c = blockIdx.x * 32 + threadIdx.x;
  __shared__ float red[8][33];
  float acc = 0.f;
  if (c < cols) {
    for (int r = blockIdx.y * 8 + threadIdx.y; r < rows; r += gridDim.y * 8) {
      const long long o = (long long)r * cols + c;
      float g = ldf(dy + o);
      if (dropping) g *= svpc_drop_scale(dkey, drop.p, dinv, (unsigned long long)o);
      if (act == SVPC_ACT_RELU) g = ldf(ref + o) > 0.f ? g : 0.f;
      else if (act == SVPC_ACT_GELU) g *= svpc_gelu_grad(ldf(ref + o));
      else if (act == SVPC_ACT_SIGMOID) { const float y = ldf(ref + o); g *= y * (1.f - y); }
      if (dz) stf(dz + o, g);
      acc += g;
    }
  }
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols && dbias) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i][threadIdx.x];
    atomicAdd(dbias + c, s);
  }
}

static int act_bwd_vec_launch(int rows, int cols, const void* dy, const void* ref, int dt, int act, void* dz, float* dbias,
                              SvpcDrop drop, cudaStream_t st);

extern "C" int svpc_act_bwd(int rows, int cols, const void* dy, const void* ref, int dt, int act, void* dz, float* dbias,
                            const unsigned int* drop_seed, unsigned int drop_site, float drop_p, svpc_stream_t stream) {
  if (rows <= 0 || cols <= 0) return SVPC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  if (dbias) cudaMemsetAsync(dbias, 0, sizeof(float) * cols, st);
  if (cols % 8 == 0 && (((uintptr_t)dy | (uintptr_t)ref | (uintptr_t)dz) & 15) == 0)
    return act_bwd_vec_launch(rows, cols, dy, ref, dt, act, dz, dbias, drop, st);
  dim3 grid(svpc_cdiv(cols, 32), svpc_imin(svpc_cdiv(rows, 8), 64));
  SVPC_DISPATCH(dt, T, (svpc_launch(act_bwd_kernel<T>, grid, dim3(32, 8), 0, st, rows, cols, (const T*)dy, (const T*)ref, act,
                                                                      (T*)dz, dbias, drop)));
  return svpc_check_launch("act_bwd");
}

extern "C" int svpc_colsum(int rows, int cols, const void* x, int dt, float* out, svpc_stream_t stream) {
  return svpc_act_bwd(rows, cols, x, nullptr, dt, SVPC_ACT_NONE, nullptr, out, nullptr, 0, 0.f, stream);
}

// out[c] += sum_r w[r] * x[r, c]
template <typename T>
__global__ void weighted_colsum_kernel(int rows, int cols, const T* __restrict__ x, const float* __restrict__ w,
                                       float* __restrict__ out) {
  svpc_pdl_sync();
  const int c = blockIdx.x * 32 + threadIdx.x;
  __shared__ float red[8][33];
  float acc = 0.f;
  if (c < cols)
    for (int r = blockIdx.y * 8 + threadIdx.y; r < rows; r += gridDim.y * 8) acc += w[r] * ldf(x + (long long)r * cols + c);
  red[threadIdx.y][threadIdx.x] = acc;
  __syncthreads();
  if (threadIdx.y == 0 && c < cols) {
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) s += red[i][threadIdx.x];
    atomicAdd(out + c, s);
  }
}

int svpc_weighted_colsum(int rows, int cols, const void* x, int dt, const float* w, float* out, cudaStream_t st) {
  dim3 grid(svpc_cdiv(cols, 32), svpc_imin(svpc_cdiv(rows, 8), 64));
  SVPC_DISPATCH(dt, T, (svpc_launch(weighted_colsum_kernel<T>, grid, dim3(32, 8), 0, st, rows, cols, (const T*)x, w, out)));
  return svpc_check_launch("weighted_colsum");
}

// ------------------------------------------------------------------------------------------------
// Warp-per-row LayerNorm for D = NCH * 256 (the model's hidden size 768 = 3 * 256): every lane owns 8 contiguous
// columns per 256-column chunk (16-byte loads / stores for bf16, 2 x 16 for fp32), row statistics by warp shuffles only,
// no block barrier.  Same two-pass statistics (mean, then centred second moment) as the generic kernel.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void ld8(const float* p, float (&v)[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void ld8(const bf16* p, float (&v)[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __low2float(h[i]); v[2 * i + 1] = __high2float(h[i]); }
}
// raw 8-element vectors: the load is issued without touching its result (a conversion right behind the load makes the thread
// wait there, before its other loads are even issued); cvt8 unpacks later
struct Raw8f { float4 a, b; };
struct Raw8h { uint4 u; };
__device__ __forceinline__ void ldraw8(const float* p, Raw8f& r) {
  r.a = *reinterpret_cast<const float4*>(p); r.b = *reinterpret_cast<const float4*>(p + 4);
}
__device__ __forceinline__ void ldraw8(const bf16* p, Raw8h& r) { r.u = *reinterpret_cast<const uint4*>(p); }
__device__ __forceinline__ void cvt8(const Raw8f& r, float (&v)[8]) {
  v[0] = r.a.x; v[1] = r.a.y; v[2] = r.a.z; v[3] = r.a.w; v[4] = r.b.x; v[5] = r.b.y; v[6] = r.b.z; v[7] = r.b.w;
}
__device__ __forceinline__ void cvt8(const Raw8h& r, float (&v)[8]) {
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&r.u);
#pragma unroll
  for (int i = 0; i < 4; ++i) { v[2 * i] = __low2float(h[i]); v[2 * i + 1] = __high2float(h[i]); }
}
template <typename T> struct Raw8;
template <> struct Raw8<float> { typedef Raw8f type; };
template <> struct Raw8<bf16> { typedef Raw8h type; };
__device__ __forceinline__ void st8(float* p, const float (&v)[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
__device__ __forceinline__ void st8(bf16* p, const float (&v)[8]) {
  uint4 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
#pragma unroll
  for (int i = 0; i < 4; ++i) h[i] = __floats2bfloat162_rn(v[2 * i], v[2 * i + 1]);
  *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void st4v(float* p, const float (&v)[4]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
}
__device__ __forceinline__ void st4v(bf16* p, const float (&v)[4]) {
  uint2 u;
  __nv_bfloat162* h = reinterpret_cast<__nv_bfloat162*>(&u);
  h[0] = __floats2bfloat162_rn(v[0], v[1]);
  h[1] = __floats2bfloat162_rn(v[2], v[3]);
  *reinterpret_cast<uint2*>(p) = u;
}
__device__ __forceinline__ float rnd_as(float v, const float*) { return v; }
__device__ __forceinline__ float rnd_as(float v, const bf16*) { return __bfloat162float(__float2bfloat16_rn(v)); }

template <typename TI, typename TO, int NCH>
__global__ void __launch_bounds__(256) ln_fwd_warp_kernel(int rows, const TI* __restrict__ x, const TI* __restrict__ res,
                                                          const float* __restrict__ gamma, const float* __restrict__ beta,
                                                          float eps, const float* __restrict__ pe, int pe_period,
                                                          const float* __restrict__ tt, const int* __restrict__ tt_ids,
                                                          TO* __restrict__ y, TI* __restrict__ pre,
                                                          float* __restrict__ mean, float* __restrict__ rstd) {
  svpc_pdl_sync();
  constexpr int D = NCH * 256;
  const int lane = threadIdx.x & 31;
  const int nw = gridDim.x * 8;
  for (int r = blockIdx.x * 8 + (threadIdx.x >> 5); r < rows; r += nw) {
    const long long base = (long long)r * D;
    float v[NCH][8];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = i * 256 + lane * 8;
      ld8(x + base + c, v[i]);
      if (res) {
        float t[8];
        ld8(res + base + c, t);
#pragma unroll
        for (int e = 0; e < 8; ++e) v[i][e] = rnd_as(v[i][e] + t[e], x);   // statistics on the stored (rounded) value
        st8(pre + base + c, v[i]);
      }
#pragma unroll
      for (int e = 0; e < 8; ++e) s += v[i][e];
    }
    const float mu = warp_sum(s) * (1.0f / D);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i)
#pragma unroll
      for (int e = 0; e < 8; ++e) { const float d = v[i][e] - mu; q = fmaf(d, d, q); }
    const float var = warp_sum(q) * (1.0f / D);
    const float rs = 1.0f / sqrtf(var + eps);
    if (lane == 0) { mean[r] = mu; rstd[r] = rs; }
    const float* per = pe ? pe + (long long)(r % pe_period) * D : nullptr;
    const float* ttr = tt ? tt + (long long)tt_ids[r] * D : nullptr;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = i * 256 + lane * 8;
      float g[8], b[8], o[8];
      ld8(gamma + c, g);
      ld8(beta + c, b);
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = (v[i][e] - mu) * rs * g[e] + b[e];
      if (per) { ld8(per + c, g);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += g[e]; }
      if (ttr) { ld8(ttr + c, g);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] += g[e]; }
      st8(y + base + c, o);
    }
  }
}

// FUSED: the LayerNorm input is y = drop(act(z)) (+ residual) with z the pre-activation of the Linear in front of it; the
// kernel then also writes dz = dpre * dropscale(r * D + c) * act'(ref) (what that Linear's dgrad / wgrad GEMMs consume) and
// accumulates dbias[c] = sum_r dz[r, c], so no separate activation-backward pass runs over the tensor.  ref: the Linear's
// pre-activation (GELU) or output (sigmoid); for ReLU without a residual the LN input itself is the activation output.
template <typename TG, typename TP, int NCH, bool FUSED>
__global__ void __launch_bounds__(256, 2) ln_bwd_warp_kernel(int rows, const TG* __restrict__ dy, const TP* __restrict__ pre,
                                                             const float* __restrict__ mean, const float* __restrict__ rstd,
                                                             const float* __restrict__ gamma, TP* __restrict__ dpre,
                                                             float* __restrict__ dgamma, float* __restrict__ dbeta,
                                                             TP* __restrict__ dz, int act, const TP* __restrict__ ref,
                                                             float* __restrict__ dbias, const SvpcDrop drop) {
  svpc_pdl_sync();
  constexpr int D = NCH * 256;
  extern __shared__ float ln_sm[];
  float (*part)[D] = reinterpret_cast<float (*)[D]>(ln_sm);              // [8][D] cross-warp reduction of the column sums
  // FUSED: the column sums of dy (dbeta) and dz (dbias) are accumulated in shared memory, lane-interleaved so that the
  // read-modify-writes are conflict-free (slot (i * 8 + e) * 32 + lane holds column i * 256 + lane * 8 + e): the register
  // file then holds one accumulator set (dgamma) instead of three and nothing spills
  float* bsum = ln_sm + 8 * D + (threadIdx.x >> 5) * D;                  // this warp's [D]
  float* zsum = ln_sm + 16 * D + (threadIdx.x >> 5) * D;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int nw = gridDim.x * 8;
  const bool dropping = FUSED && drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = FUSED ? svpc_drop_key(drop) : 0u;
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  float ag[NCH][8], ab[FUSED ? 1 : NCH][8];
#pragma unroll
  for (int i = 0; i < NCH; ++i)
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      ag[i][e] = 0.f;
      if (!FUSED) ab[i][e] = 0.f;
      else { bsum[(i * 8 + e) * 32 + lane] = 0.f; zsum[(i * 8 + e) * 32 + lane] = 0.f; }     // own slots only: no barrier
    }
  for (int r = blockIdx.x * 8 + warp; r < rows; r += nw) {
    const long long base = (long long)r * D;
    const float mu = mean[r], rs = rstd[r];
    float d[NCH][8], xh[NCH][8];
    uint32_t pos = 0;                             // FUSED + ReLU: bit (i * 8 + e) = LN input > 0
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = i * 256 + lane * 8;
      float gm[8];
      ld8(dy + base + c, d[i]);
      ld8(pre + base + c, xh[i]);
      ld8(gamma + c, gm);                       // L1-resident
#pragma unroll
      for (int e = 0; e < 8; ++e) {
        if (FUSED) pos |= (xh[i][e] > 0.f ? 1u : 0u) << (i * 8 + e);
        xh[i][e] = (xh[i][e] - mu) * rs;
        ag[i][e] = fmaf(d[i][e], xh[i][e], ag[i][e]);
        if (FUSED) bsum[(i * 8 + e) * 32 + lane] += d[i][e];
        else ab[i][e] += d[i][e];
        d[i][e] *= gm[e];                       // g = dy * gamma
        s1 += d[i][e];
        s2 = fmaf(d[i][e], xh[i][e], s2);
      }
    }
    const float m1 = warp_sum(s1) * (1.0f / D), m2 = warp_sum(s2) * (1.0f / D);
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = i * 256 + lane * 8;
      float o[8];
#pragma unroll
      for (int e = 0; e < 8; ++e) o[e] = rs * (d[i][e] - m1 - xh[i][e] * m2);
      st8(dpre + base + c, o);
      if (FUSED) {
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = rnd_as(o[e], dpre);          // the stored gradient is what the un-fused pass would read
        if (dropping) {
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            float s0, s1_;
            svpc_drop_scale2(dkey, drop.p, dinv, (unsigned long long)(base + c + e), s0, s1_);
            o[e] *= s0; o[e + 1] *= s1_;
          }
        }
        if (act == SVPC_ACT_RELU && !ref) {
#pragma unroll
          for (int e = 0; e < 8; ++e) o[e] = ((pos >> (i * 8 + e)) & 1u) ? o[e] : 0.f;
        } else if (act != SVPC_ACT_NONE) {
          float y[8];
          ld8(ref + base + c, y);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            if (act == SVPC_ACT_RELU) o[e] = y[e] > 0.f ? o[e] : 0.f;
            else if (act == SVPC_ACT_GELU) o[e] *= svpc_gelu_grad(y[e]);
            else o[e] *= y[e] * (1.f - y[e]);
          }
        }
        st8(dz + base + c, o);
#pragma unroll
        for (int e = 0; e < 8; ++e) zsum[(i * 8 + e) * 32 + lane] += rnd_as(o[e], dz);
      }
    }
  }
  // column sums: across the CTA's warps through shared memory, then one red.global per column per CTA
  __syncthreads();
#pragma unroll
  for (int i = 0; i < NCH; ++i)
#pragma unroll
    for (int e = 0; e < 8; ++e) part[warp][i * 256 + lane * 8 + e] = ag[i][e];
  __syncthreads();
  for (int c = threadIdx.x; c < D; c += 256) {
    float t = 0.f;
#pragma unroll
    for (int w = 0; w < 8; ++w) t += part[w][c];
    atomicAdd(dgamma + c, t);
  }
  if (!FUSED) {
    __syncthreads();
#pragma unroll
    for (int i = 0; i < NCH; ++i)
#pragma unroll
      for (int e = 0; e < 8; ++e) part[warp][i * 256 + lane * 8 + e] = ab[i][e];
    __syncthreads();
    for (int c = threadIdx.x; c < D; c += 256) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += part[w][c];
      atomicAdd(dbeta + c, t);
    }
  } else {
    for (int c = threadIdx.x; c < D; c += 256) {       // (bsum / zsum were complete before the barriers above)
      const int slot = ((c >> 8) * 8 + (c & 7)) * 32 + ((c & 255) >> 3);      // column c = i * 256 + lane * 8 + e
      float tb = 0.f, tz = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) { tb += ln_sm[8 * D + w * D + slot]; tz += ln_sm[16 * D + w * D + slot]; }
      atomicAdd(dbeta + c, tb);
      if (dbias) atomicAdd(dbias + c, tz);
    }
  }
}

// activation backward + bias gradient, 8 columns (16 B of bf16) per thread: a warp covers 256 contiguous columns of a row
template <typename T>
__global__ void __launch_bounds__(256) act_bwd_vec_kernel(int rows, int cols, const T* __restrict__ dy,
                                                          const T* __restrict__ ref, int act, T* __restrict__ dz,
                                                          float* __restrict__ dbias, const SvpcDrop drop) {
  svpc_pdl_sync();
  __shared__ float red[8][256];
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(drop);
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c0 = (blockIdx.x * 32 + tx) * 8;
  float acc[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) acc[e] = 0.f;
  if (c0 < cols) {
    const int rstep = gridDim.y * 8;
    constexpr int NR = 4;     // rows per iteration: every 16-byte load of the iteration is issued (row index clamped) before any is used
    for (int r0 = blockIdx.y * 8 + ty; r0 < rows; r0 += NR * rstep) {
      typename Raw8<T>::type gr[NR], yr[NR];
#pragma unroll
      for (int u = 0; u < NR; ++u) {
        const long long o = (long long)min(r0 + u * rstep, rows - 1) * cols + c0;
        ldraw8(dy + o, gr[u]);
        if (act != SVPC_ACT_NONE) ldraw8(ref + o, yr[u]);
      }
#pragma unroll
      for (int u = 0; u < NR; ++u) {
        const int r = r0 + u * rstep;
        if (r >= rows) break;
        const long long o = (long long)r * cols + c0;
        float g[8], y[8];
        cvt8(gr[u], g);
        if (dropping) {                    // cols % 8 == 0 and c0 % 8 == 0: o is even, (o + e, o + e + 1) are aligned pairs
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            float s0, s1;
            svpc_drop_scale2(dkey, drop.p, dinv, (unsigned long long)(o + e), s0, s1);
            g[e] *= s0; g[e + 1] *= s1;
          }
        }
        if (act != SVPC_ACT_NONE) {
          cvt8(yr[u], y);
#pragma unroll
          for (int e = 0; e < 8; ++e) {
            if (act == SVPC_ACT_RELU) g[e] = y[e] > 0.f ? g[e] : 0.f;
            else if (act == SVPC_ACT_GELU) g[e] *= svpc_gelu_grad(y[e]);
            else g[e] *= y[e] * (1.f - y[e]);
          }
        }
        if (dz) st8(dz + o, g);
#pragma unroll
        for (int e = 0; e < 8; ++e) acc[e] += g[e];
      }
    }
  }
#pragma unroll
  for (int e = 0; e < 8; ++e) red[ty][tx * 8 + e] = acc[e];
  __syncthreads();
  const int c = blockIdx.x * 256 + threadIdx.x;
  if (dbias && c < cols) {
    float t = 0.f;
#pragma unroll
    for (int i = 0; i < 8; ++i) t += red[i][threadIdx.x];
    atomicAdd(dbias + c, t);
  }
}

static inline bool rw_al16(const void* p) { return ((uintptr_t)p & 15) == 0; }

static int act_bwd_vec_launch(int rows, int cols, const void* dy, const void* ref, int dt, int act, void* dz, float* dbias,
                              SvpcDrop drop, cudaStream_t st) {
  const int gx = svpc_cdiv(cols, 256);
  dim3 grid(gx, svpc_imin(svpc_cdiv(rows, 8), svpc_cdiv(SVPC_NUM_SMS * 4, gx)));
  SVPC_DISPATCH(dt, T, (svpc_launch(act_bwd_vec_kernel<T>, grid, 256, 0, st, rows, cols, (const T*)dy, (const T*)ref, act, (T*)dz, dbias, drop)));
  return svpc_check_launch("act_bwd");
}

// ------------------------------------------------------------------------------------------------
// LayerNorm forward: y = ((x+res) - mean) * rstd * gamma + beta (+ pe[row % period]) (+ tt[tt_ids[row]])
// ------------------------------------------------------------------------------------------------
template <typename TI, typename TO>
__global__ void __launch_bounds__(RW_THREADS) ln_fwd_kernel(int rows, int D, const TI* __restrict__ x,
                                                            const TI* __restrict__ res, const float* __restrict__ gamma,
                                                            const float* __restrict__ beta, float eps,
                                                            const float* __restrict__ pe, int pe_period,
                                                            const float* __restrict__ tt, const int* __restrict__ tt_ids,
                                                            TO* __restrict__ y, TI* __restrict__ pre,
                                                            float* __restrict__ mean, float* __restrict__ rstd) {
  svpc_pdl_sync();
  __shared__ float red[64];
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    const long long base = (long long)r * D;
    float v[RW_MAXPT];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < RW_MAXPT; ++i) {
      const int c = threadIdx.x + i * RW_THREADS;
      v[i] = 0.f;
      if (c < D) {
        float t = ldf(x + base + c);
        if (res) {
          t += ldf(res + base + c);
          stf(pre + base + c, t);
          t = ldf(pre + base + c);      // statistics on the stored (possibly bf16-rounded) value the backward will read
        }
        v[i] = t;
        s += t;
      }
    }
    const float mu = block_sum(s, red) / D;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < RW_MAXPT; ++i) {
      const int c = threadIdx.x + i * RW_THREADS;
      if (c < D) { const float d = v[i] - mu; q += d * d; }
    }
    const float var = block_sum(q, red) / D;
    const float rs = 1.0f / sqrtf(var + eps);
    if (threadIdx.x == 0) { mean[r] = mu; rstd[r] = rs; }
    const float* per = pe ? pe + (long long)(r % pe_period) * D : nullptr;
    const float* ttr = tt ? tt + (long long)tt_ids[r] * D : nullptr;
#pragma unroll
    for (int i = 0; i < RW_MAXPT; ++i) {
      const int c = threadIdx.x + i * RW_THREADS;
      if (c < D) {
        float o = (v[i] - mu) * rs * gamma[c] + beta[c];
        if (per) o += per[c];
        if (ttr) o += ttr[c];
        stf(y + base + c, o);
      }
    }
  }
}

extern "C" int svpc_ln_fwd(int rows, int D, const void* x, const void* res, int dt_in, const float* gamma,
                           const float* beta, float eps, const float* pe, int pe_period, const float* tt,
                           const int* tt_ids, void* y, int dt_out, void* pre, float* mean, float* rstd,
                           svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  SVPC_REQUIRE(D <= RW_THREADS * RW_MAXPT, "ln_fwd: D=%d too wide", D);
  SVPC_REQUIRE(!res || pre, "ln_fwd: residual given without a pre buffer");
  cudaStream_t st = (cudaStream_t)stream;
  if (D == 768 && rw_al16(x) && rw_al16(res) && rw_al16(y) && rw_al16(pre) && rw_al16(gamma) && rw_al16(beta) && rw_al16(pe) &&
      rw_al16(tt)) {
    const int grid = svpc_imin(svpc_cdiv(rows, 8), SVPC_NUM_SMS * 8);
    SVPC_DISPATCH(dt_in, TI, SVPC_DISPATCH(dt_out, TO, (svpc_launch(ln_fwd_warp_kernel<TI, TO, 3>, grid, 256, 0, st, 
        rows, (const TI*)x, (const TI*)res, gamma, beta, eps, pe, pe_period, tt, tt_ids, (TO*)y, (TI*)pre, mean, rstd))));
    return svpc_check_launch("ln_fwd");
  }
  SVPC_DISPATCH(dt_in, TI, SVPC_DISPATCH(dt_out, TO, (svpc_launch(ln_fwd_kernel<TI, TO>, rw_grid(rows), RW_THREADS, 0, st, 
      rows, D, (const TI*)x, (const TI*)res, gamma, beta, eps, pe, pe_period, tt, tt_ids, (TO*)y, (TI*)pre, mean, rstd))));
  return svpc_check_launch("ln_fwd");
}

// ------------------------------------------------------------------------------------------------
// LayerNorm backward
// ------------------------------------------------------------------------------------------------
template <typename TG, typename TP>
__global__ void __launch_bounds__(RW_THREADS) ln_bwd_kernel(int rows, int D, const TG* __restrict__ dy,
                                                            const TP* __restrict__ pre, const float* __restrict__ mean,
                                                            const float* __restrict__ rstd, const float* __restrict__ gamma,
                                                            TP* __restrict__ dpre, float* __restrict__ dgamma,
                                                            float* __restrict__ dbeta, TP* __restrict__ dz, int act,
                                                            const TP* __restrict__ ref, float* __restrict__ dbias,
                                                            const SvpcDrop drop) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const bool dropping = dz != nullptr && drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = dz ? svpc_drop_key(drop) : 0u;
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  float ag[RW_MAXPT], ab[RW_MAXPT], az[RW_MAXPT];
#pragma unroll
  for (int i = 0; i < RW_MAXPT; ++i) { ag[i] = 0.f; ab[i] = 0.f; az[i] = 0.f; }
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    const long long base = (long long)r * D;
    const float mu = mean[r], rs = rstd[r];
    float g[RW_MAXPT], xh[RW_MAXPT], raw[RW_MAXPT];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < RW_MAXPT; ++i) {
      const int c = threadIdx.x + i * RW_THREADS;
      g[i] = 0.f; xh[i] = 0.f; raw[i] = 0.f;
      if (c < D) {
        const float d = ldf(dy + base + c);
        raw[i] = ldf(pre + base + c);
        xh[i] = (raw[i] - mu) * rs;
        g[i] = d * gamma[c];
        ag[i] += d * xh[i];
        ab[i] += d;
        s1 += g[i];
        s2 += g[i] * xh[i];
      }
    }
    const float2 ss = block_sum2(s1, s2, red);
    const float m1 = ss.x / D, m2 = ss.y / D;
#pragma unroll
    for (int i = 0; i < RW_MAXPT; ++i) {
      const int c = threadIdx.x + i * RW_THREADS;
      if (c < D) {
        float o = rs * (g[i] - m1 - xh[i] * m2);
        stf(dpre + base + c, o);
        if (dz) {
          o = rnd_as(o, dpre);
          if (dropping) o *= svpc_drop_scale(dkey, drop.p, dinv, (unsigned long long)(base + c));
          if (act != SVPC_ACT_NONE) {
            const float y = ref ? ldf(ref + base + c) : raw[i];
            if (act == SVPC_ACT_RELU) o = y > 0.f ? o : 0.f;
            else if (act == SVPC_ACT_GELU) o *= svpc_gelu_grad(y);
            else o *= y * (1.f - y);
          }
          stf(dz + base + c, o);
          az[i] += rnd_as(o, dz);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < RW_MAXPT; ++i) {
    const int c = threadIdx.x + i * RW_THREADS;
    if (c < D) {
      atomicAdd(dgamma + c, ag[i]);
      atomicAdd(dbeta + c, ab[i]);
      if (dz && dbias) atomicAdd(dbias + c, az[i]);
    }
  }
}

// dz != NULL: fused activation / dropout backward of the Linear in front of the LayerNorm (see ln_bwd_warp_kernel): dz has
// the layout and dtype of dpre, ref (optional) of pre; dbias (optional, D floats) receives the column sums of dz.
extern "C" int svpc_ln_bwd(int rows, int D, const void* dy, int dt_dy, const void* pre, int dt_pre, const float* mean,
                           const float* rstd, const float* gamma, void* dpre, float* dgamma, float* dbeta,
                           void* dz, int act, const void* ref, float* dbias, const unsigned int* drop_seed,
                           unsigned int drop_site, float drop_p, svpc_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(dgamma, 0, sizeof(float) * D, st);
  cudaMemsetAsync(dbeta, 0, sizeof(float) * D, st);
  if (dz && dbias) cudaMemsetAsync(dbias, 0, sizeof(float) * D, st);
  if (rows <= 0) return SVPC_OK;
  SVPC_REQUIRE(D <= RW_THREADS * RW_MAXPT, "ln_bwd: D=%d too wide", D);
  SVPC_REQUIRE(dz || (act == SVPC_ACT_NONE && !ref), "ln_bwd: act / ref need the dz output");
  SVPC_REQUIRE(drop_p >= 0.f && drop_p < 1.f, "ln_bwd: dropout p=%f outside [0, 1)", drop_p);
  const SvpcDrop drop{(dz && drop_p > 0.f) ? drop_seed : nullptr, drop_site, drop_p};
  if (D == 768 && rw_al16(dy) && rw_al16(pre) && rw_al16(dpre) && rw_al16(gamma) && rw_al16(dz) && rw_al16(ref)) {
    const int g2 = svpc_imin(svpc_cdiv(rows, 8), SVPC_NUM_SMS * 4);
    if (dz) {
      const size_t smem = sizeof(float) * 24 * 768;
      SVPC_DISPATCH(dt_dy, TG, SVPC_DISPATCH(dt_pre, TP, {
        static bool attr = false;
        if (!attr) { cudaFuncSetAttribute(ln_bwd_warp_kernel<TG, TP, 3, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem); attr = true; }
        svpc_launch(ln_bwd_warp_kernel<TG, TP, 3, true>, g2, 256, smem, st, rows, (const TG*)dy, (const TP*)pre, mean, rstd, gamma,
                    (TP*)dpre, dgamma, dbeta, (TP*)dz, act, (const TP*)ref, dbias, drop);
      }));
    } else {
      SVPC_DISPATCH(dt_dy, TG, SVPC_DISPATCH(dt_pre, TP, (svpc_launch(ln_bwd_warp_kernel<TG, TP, 3, false>, g2, 256, sizeof(float) * 8 * 768, st,
          rows, (const TG*)dy, (const TP*)pre, mean, rstd, gamma, (TP*)dpre, dgamma, dbeta, (TP*)nullptr, 0, (const TP*)nullptr,
          (float*)nullptr, drop))));
    }
    return svpc_check_launch("ln_bwd");
  }
  const int grid = svpc_imin(rows, SVPC_NUM_SMS * 4);
  SVPC_DISPATCH(dt_dy, TG, SVPC_DISPATCH(dt_pre, TP, (svpc_launch(ln_bwd_kernel<TG, TP>, grid, RW_THREADS, 0, st,
      rows, D, (const TG*)dy, (const TP*)pre, mean, rstd, gamma, (TP*)dpre, dgamma, dbeta, (TP*)dz, act, (const TP*)ref, dbias, drop))));
  return svpc_check_launch("ln_bwd");
}

// out[ids[r], c] += dy[r, c]  (nbins small: token types)
template <typename T>
__global__ void index_colsum_kernel(int rows, int D, const T* __restrict__ dy, const int* __restrict__ ids, int nbins,
                                    float* __restrict__ out) {
  svpc_pdl_sync();
  // blockDim.x threads over columns (grid.x strips), grid.y row slices; up to 8 bins kept in registers
  const int c = blockIdx.x * blockDim.x + threadIdx.x;
  if (c >= D) return;
  float acc[8];
#pragma unroll
  for (int b = 0; b < 8; ++b) acc[b] = 0.f;
  for (int r = blockIdx.y; r < rows; r += gridDim.y) {
    const int id = ids[r];
    const float v = ldf(dy + (long long)r * D + c);
#pragma unroll
    for (int b = 0; b < 8; ++b) acc[b] += (id == b) ? v : 0.f;
  }
#pragma unroll
  for (int b = 0; b < 8; ++b)
    if (b < nbins) atomicAdd(out + (long long)b * D + c, acc[b]);
}

// 8 columns (one 16-byte load of bf16) per thread, 8 row lanes per CTA, up to NB bins in registers: the token-type gradient
// reads the (19200, 768) gradient once at streaming speed instead of two bytes per thread per row
template <typename T, int NB>
__global__ void __launch_bounds__(256) index_colsum_vec_kernel(int rows, int D, const T* __restrict__ dy,
                                                               const int* __restrict__ ids, int nbins,
                                                               float* __restrict__ out) {
  svpc_pdl_sync();
  __shared__ float red[8][256];
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c0 = (blockIdx.x * 32 + tx) * 8;
  float acc[NB][8];
#pragma unroll
  for (int b = 0; b < NB; ++b)
#pragma unroll
    for (int e = 0; e < 8; ++e) acc[b][e] = 0.f;
  if (c0 < D) {
    const int rstep = gridDim.y * 8;
    for (int r0 = blockIdx.y * 8 + ty; r0 < rows; r0 += 2 * rstep) {
      float g0[8], g1[8];
      const int r1 = r0 + rstep;
      const bool has1 = r1 < rows;
      ld8(dy + (long long)r0 * D + c0, g0);
      if (has1) ld8(dy + (long long)r1 * D + c0, g1);
      const int id0 = ids[r0], id1 = has1 ? ids[r1] : -1;
#pragma unroll
      for (int b = 0; b < NB; ++b)
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          acc[b][e] += (id0 == b) ? g0[e] : 0.f;
          if (has1) acc[b][e] += (id1 == b) ? g1[e] : 0.f;
        }
    }
  }
#pragma unroll
  for (int b = 0; b < NB; ++b) {
    if (b >= nbins) break;
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 8; ++e) red[ty][tx * 8 + e] = acc[b][e];
    __syncthreads();
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c < D) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
      atomicAdd(out + (long long)b * D + c, t);
    }
  }
}

extern "C" int svpc_index_colsum(int rows, int D, const void* dy, int dt, const int* ids, int nbins, float* out,
                                 svpc_stream_t stream) {
  SVPC_REQUIRE(nbins <= 8, "index_colsum: nbins=%d > 8", nbins);
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(out, 0, sizeof(float) * nbins * D, st);
  if (rows <= 0) return SVPC_OK;
  if (D % 8 == 0 && nbins <= 4 && rw_al16(dy)) {
    const int gx = svpc_cdiv(D, 256);
    dim3 grid(gx, svpc_imin(svpc_cdiv(rows, 16), svpc_cdiv(SVPC_NUM_SMS * 4, gx)));
    SVPC_DISPATCH(dt, T, (svpc_launch(index_colsum_vec_kernel<T, 4>, grid, 256, 0, st, rows, D, (const T*)dy, ids, nbins, out)));
    return svpc_check_launch("index_colsum");
  }
  dim3 grid(svpc_cdiv(D, 128), svpc_imin(rows, 256));
  SVPC_DISPATCH(dt, T, (svpc_launch(index_colsum_kernel<T>, grid, 128, 0, st, rows, D, (const T*)dy, ids, nbins, out)));
  return svpc_check_launch("index_colsum");
}

// ------------------------------------------------------------------------------------------------
// embedding gather + LN(W)
// ------------------------------------------------------------------------------------------------
#define EM_THREADS 128
#define EM_MAXPT 4      // W <= 512
template <typename T>
__global__ void __launch_bounds__(EM_THREADS) embed_ln_fwd_kernel(int rows, int W, const float* __restrict__ table,
                                                                  const int* __restrict__ ids,
                                                                  const float* __restrict__ gamma,
                                                                  const float* __restrict__ beta, float eps,
                                                                  T* __restrict__ y, long long ldy,
                                                                  float* __restrict__ mean, float* __restrict__ rstd,
                                                                  const SvpcDrop drop) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(drop);
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    const float* src = table + (long long)ids[r] * W;
    float v[EM_MAXPT];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < EM_MAXPT; ++i) {
      const int c = threadIdx.x + i * EM_THREADS;
      v[i] = c < W ? src[c] : 0.f;
      s += v[i];
    }
    const float mu = block_sum(s, red) / W;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < EM_MAXPT; ++i) {
      const int c = threadIdx.x + i * EM_THREADS;
      if (c < W) { const float d = v[i] - mu; q += d * d; }
    }
    const float rs = 1.0f / sqrtf(block_sum(q, red) / W + eps);
    if (threadIdx.x == 0) { mean[r] = mu; rstd[r] = rs; }
#pragma unroll
    for (int i = 0; i < EM_MAXPT; ++i) {
      const int c = threadIdx.x + i * EM_THREADS;
      if (c < W) {
        float o = (v[i] - mu) * rs * gamma[c] + beta[c];
        if (dropping) o *= svpc_drop_scale(dkey, drop.p, dinv, (unsigned long long)r * W + c);
        stf(y + (long long)r * ldy + c, o);
      }
    }
  }
}

extern "C" int svpc_embed_ln_fwd(int rows, int W, const float* table, const int* ids, const float* gamma,
                                 const float* beta, float eps, void* y, long long ldy, int dt, float* mean, float* rstd,
                                 const unsigned int* drop_seed, unsigned int drop_site, float drop_p,
                                 svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  SVPC_REQUIRE(W <= EM_THREADS * EM_MAXPT, "embed_ln: W=%d too wide", W);
  SVPC_REQUIRE(ldy >= W, "embed_ln: ldy=%lld < W=%d", ldy, W);
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  cudaStream_t st = (cudaStream_t)stream;
  SVPC_DISPATCH(dt, T, (svpc_launch(embed_ln_fwd_kernel<T>, rw_grid(rows), EM_THREADS, 0, st, rows, W, table, ids, gamma, beta, eps,
                                                                                    (T*)y, ldy, mean, rstd, drop)));
  return svpc_check_launch("embed_ln_fwd");
}

template <typename T>
__global__ void __launch_bounds__(EM_THREADS) embed_ln_bwd_kernel(int rows, int W, const T* __restrict__ dy,
                                                                  const float* __restrict__ table,
                                                                  const int* __restrict__ ids,
                                                                  const float* __restrict__ mean,
                                                                  const float* __restrict__ rstd,
                                                                  const float* __restrict__ gamma, int padding_idx,
                                                                  float* __restrict__ dtable, float* __restrict__ dgamma,
                                                                  float* __restrict__ dbeta, const SvpcDrop drop) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(drop);
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  float ag[EM_MAXPT], ab[EM_MAXPT];
#pragma unroll
  for (int i = 0; i < EM_MAXPT; ++i) { ag[i] = 0.f; ab[i] = 0.f; }
  for (int r = blockIdx.x; r < rows; r += gridDim.x) {
    const int id = ids[r];
    const float* src = table + (long long)id * W;
    const float mu = mean[r], rs = rstd[r];
    float g[EM_MAXPT], xh[EM_MAXPT];
    float s1 = 0.f, s2 = 0.f;
#pragma unroll
    for (int i = 0; i < EM_MAXPT; ++i) {
      const int c = threadIdx.x + i * EM_THREADS;
      g[i] = 0.f; xh[i] = 0.f;
      if (c < W) {
        float d = ldf(dy + (long long)r * W + c);
        if (dropping) d *= svpc_drop_scale(dkey, drop.p, dinv, (unsigned long long)r * W + c);
        xh[i] = (src[c] - mu) * rs;
        g[i] = d * gamma[c];
        ag[i] += d * xh[i];
        ab[i] += d;
        s1 += g[i];
        s2 += g[i] * xh[i];
      }
    }
    const float2 ss = block_sum2(s1, s2, red);
    const float m1 = ss.x / W, m2 = ss.y / W;
    if (id != padding_idx) {
#pragma unroll
      for (int i = 0; i < EM_MAXPT; ++i) {
        const int c = threadIdx.x + i * EM_THREADS;
        if (c < W) atomicAdd(dtable + (long long)id * W + c, rs * (g[i] - m1 - xh[i] * m2));
      }
    }
  }
#pragma unroll
  for (int i = 0; i < EM_MAXPT; ++i) {
    const int c = threadIdx.x + i * EM_THREADS;
    if (c < W) { atomicAdd(dgamma + c, ag[i]); atomicAdd(dbeta + c, ab[i]); }
  }
}

extern "C" int svpc_embed_ln_bwd(int rows, int W, int V, const void* dy, int dt, const float* table, const int* ids,
                                 const float* mean, const float* rstd, const float* gamma, int padding_idx,
                                 float* dtable, float* dgamma, float* dbeta, const unsigned int* drop_seed,
                                 unsigned int drop_site, float drop_p, svpc_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  cudaMemsetAsync(dtable, 0, sizeof(float) * (size_t)V * W, st);
  cudaMemsetAsync(dgamma, 0, sizeof(float) * W, st);
  cudaMemsetAsync(dbeta, 0, sizeof(float) * W, st);
  if (rows <= 0) return SVPC_OK;
  SVPC_REQUIRE(W <= EM_THREADS * EM_MAXPT, "embed_ln: W=%d too wide", W);
  const int grid = svpc_imin(rows, SVPC_NUM_SMS * 4);
  SVPC_DISPATCH(dt, T, (svpc_launch(embed_ln_bwd_kernel<T>, grid, EM_THREADS, 0, st, rows, W, (const T*)dy, table, ids, mean, rstd,
                                                                           gamma, padding_idx, dtable, dgamma, dbeta, drop)));
  return svpc_check_launch("embed_ln_bwd");
}

// ------------------------------------------------------------------------------------------------
// video features: gather the first Lv rows of every clip from the per-step (B, L, F) tensors, LN statistics,
// write the normalised rows in the compute dtype (the only pass over the 1.23 MB/clip of fp32 features)
// ------------------------------------------------------------------------------------------------
// four consecutive feature values (features arrive as fp32 or, on the bf16 upload path, as bf16)
__device__ __forceinline__ float4 ldfeat4(const float* p) { return __ldg(reinterpret_cast<const float4*>(p)); }
__device__ __forceinline__ float4 ldfeat4(const bf16* p) {
  const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x), b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
  return make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
}

template <typename T, typename TF>
__global__ void __launch_bounds__(RW_THREADS) video_ln_pack_kernel(int n_rows, int B, int L, int Lv, int F,
                                                                   const long long* __restrict__ feat_ptrs,
                                                                   const int* __restrict__ clip_t,
                                                                   const int* __restrict__ clip_b, float eps,
                                                                   T* __restrict__ xhat, float* __restrict__ mean,
                                                                   float* __restrict__ rstd, const float* __restrict__ gamma,
                                                                   const float* __restrict__ beta, const SvpcDrop drop) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const int F4 = F >> 2;
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(drop);
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  for (int r = blockIdx.x; r < n_rows; r += gridDim.x) {
    const int n = r / Lv, row = r - n * Lv;
    const TF* src = reinterpret_cast<const TF*>(feat_ptrs[clip_t[n]]) + ((long long)clip_b[n] * L + row) * F;
    float4 v[RW_MAXPT / 4];
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < RW_MAXPT / 4; ++i) {
      const int c = threadIdx.x + i * RW_THREADS;
      v[i] = make_float4(0.f, 0.f, 0.f, 0.f);
      if (c < F4) { v[i] = ldfeat4(src + 4 * c); s += v[i].x + v[i].y + v[i].z + v[i].w; }
    }
    const float mu = block_sum(s, red) / F;
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < RW_MAXPT / 4; ++i) {
      const int c = threadIdx.x + i * RW_THREADS;
      if (c < F4) {
        const float a = v[i].x - mu, b = v[i].y - mu, cc = v[i].z - mu, d = v[i].w - mu;
        q += a * a + b * b + cc * cc + d * d;
      }
    }
    const float rs = 1.0f / sqrtf(block_sum(q, red) / F + eps);
    if (threadIdx.x == 0) { mean[r] = mu; rstd[r] = rs; }
    T* dst = xhat + (long long)r * F;
#pragma unroll
    for (int i = 0; i < RW_MAXPT / 4; ++i) {
      const int c = threadIdx.x + i * RW_THREADS;
      if (c < F4) {
        float o[4] = {(v[i].x - mu) * rs, (v[i].y - mu) * rs, (v[i].z - mu) * rs, (v[i].w - mu) * rs};
        if (gamma) {                       // train mode: LN affine and dropout applied here (model.py:549-550)
          const float4 g4 = *reinterpret_cast<const float4*>(gamma + 4 * c), b4 = *reinterpret_cast<const float4*>(beta + 4 * c);
          o[0] = o[0] * g4.x + b4.x; o[1] = o[1] * g4.y + b4.y; o[2] = o[2] * g4.z + b4.z; o[3] = o[3] * g4.w + b4.w;
          if (dropping) {                  // F is a multiple of 4: (4c, 4c+1) and (4c+2, 4c+3) are aligned pairs
            float s0, s1;
            svpc_drop_scale2(dkey, drop.p, dinv, (unsigned long long)r * F + 4 * c, s0, s1);
            o[0] *= s0; o[1] *= s1;
            svpc_drop_scale2(dkey, drop.p, dinv, (unsigned long long)r * F + 4 * c + 2, s0, s1);
            o[2] *= s0; o[3] *= s1;
          }
        }
        st4v(dst + 4 * c, o);                // one 8-byte (bf16) / 16-byte (fp32) store per thread (F % 4 == 0)
      }
    }
  }
}

// bf16 features -> bf16 rows, one WARP per row (F = 3072: 12 x 16 bytes per lane kept packed in 48 registers): no block
// barriers, 8 independent rows per CTA in flight.  Two-pass statistics (mean, then centred second moment) as above.
template <int NCH>
__global__ void __launch_bounds__(256) video_ln_pack_warp_kernel(int n_rows, int B, int L, int Lv,
                                                                 const long long* __restrict__ feat_ptrs,
                                                                 const int* __restrict__ clip_t, const int* __restrict__ clip_b,
                                                                 float eps, bf16* __restrict__ xhat, float* __restrict__ mean,
                                                                 float* __restrict__ rstd, const float* __restrict__ gamma,
                                                                 const float* __restrict__ beta, const SvpcDrop drop) {
  svpc_pdl_sync();
  constexpr int F = NCH * 256;
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(drop);
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  const int lane = threadIdx.x & 31;
  const int nw = gridDim.x * 8;
  for (int r = blockIdx.x * 8 + (threadIdx.x >> 5); r < n_rows; r += nw) {
    const int n = r / Lv, row = r - n * Lv;
    const bf16* src = reinterpret_cast<const bf16*>(feat_ptrs[clip_t[n]]) + ((long long)clip_b[n] * L + row) * F;
    uint4 raw[NCH];
#pragma unroll
    for (int i = 0; i < NCH; ++i) raw[i] = __ldg(reinterpret_cast<const uint4*>(src + i * 256 + lane * 8));
    float s = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[i]);
#pragma unroll
      for (int e = 0; e < 4; ++e) s += __low2float(h[e]) + __high2float(h[e]);
    }
    const float mu = warp_sum(s) * (1.0f / F);
    float q = 0.f;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[i]);
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float a = __low2float(h[e]) - mu, b = __high2float(h[e]) - mu;
        q = fmaf(a, a, q);
        q = fmaf(b, b, q);
      }
    }
    const float rs = 1.0f / sqrtf(warp_sum(q) * (1.0f / F) + eps);
    if (lane == 0) { mean[r] = mu; rstd[r] = rs; }
    bf16* dst = xhat + (long long)r * F;
#pragma unroll
    for (int i = 0; i < NCH; ++i) {
      const int c = i * 256 + lane * 8;
      const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&raw[i]);
      float o[8];
#pragma unroll
      for (int e = 0; e < 4; ++e) { o[2 * e] = (__low2float(h[e]) - mu) * rs; o[2 * e + 1] = (__high2float(h[e]) - mu) * rs; }
      if (gamma) {                       // train mode: LN affine and dropout applied here (model.py:549-550)
        float g[8], bb[8];
        ld8(gamma + c, g);
        ld8(beta + c, bb);
#pragma unroll
        for (int e = 0; e < 8; ++e) o[e] = o[e] * g[e] + bb[e];
        if (dropping) {
#pragma unroll
          for (int e = 0; e < 8; e += 2) {
            float s0, s1;
            svpc_drop_scale2(dkey, drop.p, dinv, (unsigned long long)r * F + c + e, s0, s1);
            o[e] *= s0; o[e + 1] *= s1;
          }
        }
      }
      st8(dst + c, o);
    }
  }
}

extern "C" int svpc_video_ln_pack(int n_clips, int B, int L, int Lv, int F, const long long* feat_ptrs, int dt_feat,
                                  const int* clip_t, const int* clip_b, float eps, void* xhat, int dt, float* mean,
                                  float* rstd, const float* gamma, const float* beta, const unsigned int* drop_seed,
                                  unsigned int drop_site, float drop_p, svpc_stream_t stream) {
  if (n_clips <= 0) return SVPC_OK;
  SVPC_REQUIRE((gamma == nullptr) == (beta == nullptr), "video_ln_pack: gamma and beta go together");
  const SvpcDrop drop{(drop_p > 0.f && gamma) ? drop_seed : nullptr, drop_site, drop_p};
  SVPC_REQUIRE(F % 4 == 0 && F <= RW_THREADS * RW_MAXPT, "video_ln_pack: F=%d must be a multiple of 4 and <= 4096", F);
  cudaStream_t st = (cudaStream_t)stream;
  const int n_rows = n_clips * Lv;
  if (dt == SVPC_BF16 && dt_feat == SVPC_BF16 && F == 3072 && rw_al16(xhat) && rw_al16(gamma) && rw_al16(beta)) {
    svpc_launch(video_ln_pack_warp_kernel<12>, svpc_imin(svpc_cdiv(n_rows, 8), SVPC_NUM_SMS * 8), 256, 0, st, n_rows, B, L, Lv,
                feat_ptrs, clip_t, clip_b, eps, (bf16*)xhat, mean, rstd, gamma, beta, drop);
    return svpc_check_launch("video_ln_pack");
  }
  SVPC_DISPATCH(dt, T, SVPC_DISPATCH(dt_feat, TF, (svpc_launch(video_ln_pack_kernel<T, TF>, svpc_imin(n_rows, SVPC_NUM_SMS * 16),
      RW_THREADS, 0, st, n_rows, B, L, Lv, F, feat_ptrs, clip_t, clip_b, eps, (T*)xhat, mean, rstd, gamma, beta, drop))));
  return svpc_check_launch("video_ln_pack");
}

// Backward of the train-mode staging y = drop(xhat * gamma + beta): dgamma[c] = sum_r dy*scale*xhat, dbeta[c] = sum_r dy*scale
// (the features are inputs: no gradient flows further).  xhat is recomputed from the raw features and the saved statistics.
// Layout: a CTA owns a 256-column slab (lane = 8 contiguous columns: one 16-byte bf16 load of dy, one or two 16-byte loads of
// the features) for a strided set of rows (8 row lanes, four rows per iteration with all their loads issued up front); column sums over the row lanes through shared memory, one red.global
// per column per CTA.
template <typename T, typename TF>
__global__ void __launch_bounds__(256, 3) video_ln_affine_bwd_kernel(int n_rows, int B, int L, int Lv, int F,
                                                                  const long long* __restrict__ feat_ptrs,
                                                                  const int* __restrict__ clip_t,
                                                                  const int* __restrict__ clip_b,
                                                                  const float* __restrict__ mean,
                                                                  const float* __restrict__ rstd,
                                                                  const T* __restrict__ dy, float* __restrict__ dgamma,
                                                                  float* __restrict__ dbeta, const SvpcDrop drop) {
  svpc_pdl_sync();
  __shared__ float red[8][256];
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(drop);
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const int c0 = (blockIdx.x * 32 + tx) * 8;
  float ag[8], ab[8];
#pragma unroll
  for (int e = 0; e < 8; ++e) { ag[e] = 0.f; ab[e] = 0.f; }
  if (c0 < F) {
    const int rstep = gridDim.y * 8;
    constexpr int NR = 4;       // rows per iteration: 2 * NR 16-byte loads issued (unconditionally, row index clamped) before any is used
    for (int r0 = blockIdx.y * 8 + ty; r0 < n_rows; r0 += NR * rstep) {
      typename Raw8<T>::type dr[NR];
      typename Raw8<TF>::type xr[NR];
      float mu[NR], rs[NR];
#pragma unroll
      for (int u = 0; u < NR; ++u) {
        const int r = min(r0 + u * rstep, n_rows - 1);
        const int n = r / Lv, row = r - n * Lv;
        const TF* src = reinterpret_cast<const TF*>(feat_ptrs[clip_t[n]]) + ((long long)clip_b[n] * L + row) * F + c0;
        mu[u] = mean[r]; rs[u] = rstd[r];
        ldraw8(dy + (long long)r * F + c0, dr[u]);
        ldraw8(src, xr[u]);
      }
#pragma unroll
      for (int u = 0; u < NR; ++u) {
        const int r = r0 + u * rstep;
        if (r < n_rows) {
          float d[8], x[8];
          cvt8(dr[u], d);
          cvt8(xr[u], x);
          if (dropping) {
#pragma unroll
            for (int e = 0; e < 8; e += 2) {
              float s0, s1;
              svpc_drop_scale2(dkey, drop.p, dinv, (unsigned long long)r * F + c0 + e, s0, s1);
              d[e] *= s0; d[e + 1] *= s1;
            }
          }
#pragma unroll
          for (int e = 0; e < 8; ++e) { ag[e] = fmaf(d[e], (x[e] - mu[u]) * rs[u], ag[e]); ab[e] += d[e]; }
        }
      }
    }
  }
#pragma unroll
  for (int pass = 0; pass < 2; ++pass) {
    __syncthreads();
#pragma unroll
    for (int e = 0; e < 8; ++e) red[ty][tx * 8 + e] = pass == 0 ? ag[e] : ab[e];
    __syncthreads();
    const int c = blockIdx.x * 256 + threadIdx.x;
    if (c < F) {
      float t = 0.f;
#pragma unroll
      for (int w = 0; w < 8; ++w) t += red[w][threadIdx.x];
      atomicAdd((pass == 0 ? dgamma : dbeta) + c, t);
    }
  }
}

extern "C" int svpc_video_ln_affine_bwd(int n_clips, int B, int L, int Lv, int F, const long long* feat_ptrs, int dt_feat,
                                        const int* clip_t, const int* clip_b, const float* mean, const float* rstd,
                                        const void* dy, int dt, float* dgamma, float* dbeta, const unsigned int* drop_seed,
                                        unsigned int drop_site, float drop_p, svpc_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(dgamma, 0, sizeof(float) * F, st);
  cudaMemsetAsync(dbeta, 0, sizeof(float) * F, st);
  if (n_clips <= 0) return SVPC_OK;
  SVPC_REQUIRE(F % 8 == 0 && (((uintptr_t)dy) & 15) == 0, "video_ln_affine_bwd: F=%d must be a multiple of 8, dy 16-byte aligned", F);
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  const int n_rows = n_clips * Lv;
  const int gx = svpc_cdiv(F, 256);
  dim3 grid(gx, svpc_imin(svpc_cdiv(n_rows, 8), svpc_cdiv(SVPC_NUM_SMS * 3, gx)));      // one wave at 3 CTAs per SM
  SVPC_DISPATCH(dt, T, SVPC_DISPATCH(dt_feat, TF, (svpc_launch(video_ln_affine_bwd_kernel<T, TF>, grid, 256, 0, st,
      n_rows, B, L, Lv, F, feat_ptrs, clip_t, clip_b, mean, rstd, (const T*)dy, dgamma, dbeta, drop))));
  return svpc_check_launch("video_ln_affine_bwd");
}

// ------------------------------------------------------------------------------------------------
// ingredient vectors: segment means between [SEP]s (+ pe[j])
// ------------------------------------------------------------------------------------------------
__global__ void seg_scan_kernel(int B, int Li, const int* __restrict__ sep_mask, const int* __restrict__ cu_ingr,
                                int* __restrict__ row_seg, int* __restrict__ seg_lohi) {
  svpc_pdl_sync();
  const int b = blockIdx.x * blockDim.x + threadIdx.x;
  if (b >= B) return;
  const int g0 = cu_ingr[b], ni = cu_ingr[b + 1] - g0;
  int prev = 0, j = 0;
  for (int r = 0; r < Li; ++r) row_seg[b * Li + r] = -1;
  for (int r = 0; r < Li && j < ni; ++r) {
    if (sep_mask[b * Li + r] == 1) {
      seg_lohi[2 * (g0 + j)] = prev;
      seg_lohi[2 * (g0 + j) + 1] = r;
      for (int t = prev; t < r; ++t) row_seg[b * Li + t] = g0 + j;
      prev = r + 1;
      ++j;
    }
  }
  for (; j < ni; ++j) { seg_lohi[2 * (g0 + j)] = 0; seg_lohi[2 * (g0 + j) + 1] = 0; }   // fewer [SEP]s than announced
}

template <typename T>
__global__ void segmean_fwd_kernel(int B, int Li, int D, const T* __restrict__ x, const int* __restrict__ cu_ingr,
                                   const int* __restrict__ seg_lohi, const float* __restrict__ pe,
                                   float* __restrict__ E0) {
  svpc_pdl_sync();
  const int g = blockIdx.x;
  // video of this ingredient
  int b = 0;
  while (b + 1 < B && cu_ingr[b + 1] <= g) ++b;
  const int j = g - cu_ingr[b];
  const int lo = seg_lohi[2 * g], hi = seg_lohi[2 * g + 1];
  const float inv = 1.0f / (float)(hi - lo);      // empty segment -> inf * 0 = NaN, as torch's mean of empty
  for (int c = threadIdx.x; c < D; c += blockDim.x) {
    float s = 0.f;
    for (int r = lo; r < hi; ++r) s += ldf(x + ((long long)b * Li + r) * D + c);
    E0[(long long)g * D + c] = s * inv + pe[(long long)j * D + c];
  }
}

template <typename T>
__global__ void segmean_bwd_kernel(int rows, int D, const float* __restrict__ dE0, const int* __restrict__ row_seg,
                                   const int* __restrict__ seg_lohi, T* __restrict__ dx) {
  svpc_pdl_sync();
  const int r = blockIdx.x;
  const int g = row_seg[r];
  float inv = 0.f;
  if (g >= 0) inv = 1.0f / (float)(seg_lohi[2 * g + 1] - seg_lohi[2 * g]);
  for (int c = threadIdx.x; c < D; c += blockDim.x)
    stf(dx + (long long)r * D + c, g >= 0 ? dE0[(long long)g * D + c] * inv : 0.f);
}

extern "C" int svpc_segmean_fwd(int B, int Li, int D, int NI, const void* x, int dt, const int* sep_mask,
                                const int* cu_ingr, const float* pe, float* E0, int* scratch_row_seg,
                                int* scratch_seg_lohi, svpc_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (B <= 0 || NI <= 0) return SVPC_OK;
  svpc_launch(seg_scan_kernel, svpc_cdiv(B, 64), 64, 0, st, B, Li, sep_mask, cu_ingr, scratch_row_seg, scratch_seg_lohi);
  int rc = svpc_check_launch("seg_scan");
  if (rc) return rc;
  SVPC_DISPATCH(dt, T, (svpc_launch(segmean_fwd_kernel<T>, NI, 256, 0, st, B, Li, D, (const T*)x, cu_ingr, scratch_seg_lohi, pe, E0)));
  return svpc_check_launch("segmean_fwd");
}

extern "C" int svpc_segmean_bwd(int B, int Li, int D, int NI, const float* dE0, const int* sep_mask, const int* cu_ingr,
                                void* dx, int dt, int* scratch_row_seg, int* scratch_seg_lohi, svpc_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (B <= 0) return SVPC_OK;
  svpc_launch(seg_scan_kernel, svpc_cdiv(B, 64), 64, 0, st, B, Li, sep_mask, cu_ingr, scratch_row_seg, scratch_seg_lohi);
  int rc = svpc_check_launch("seg_scan");
  if (rc) return rc;
  SVPC_DISPATCH(dt, T, (svpc_launch(segmean_bwd_kernel<T>, B * Li, 256, 0, st, B * Li, D, dE0, scratch_row_seg, scratch_seg_lohi, (T*)dx)));
  return svpc_check_launch("segmean_bwd");
}

// ------------------------------------------------------------------------------------------------
// svpc_segmean_fwd without caller-provided scratch and without any allocation (no counterpart in the reference: for callers
// that cannot allocate: CUDA-graph capture, serving loops).  The two scratch tables live in a file-scope device array that
// every GPU loading the library holds until exit (320 KB).  All calls share it and it keeps the last call's contents:
// calls that overlap on the GPU (different streams, different host threads) race on it.
// ------------------------------------------------------------------------------------------------
#define SEGMEAN_STATIC_MAX_ROWS 65536     // B * Li
#define SEGMEAN_STATIC_MAX_SEGS 8192      // NI
static __device__ __align__(256) int g_segmean_scratch[SEGMEAN_STATIC_MAX_ROWS + 2 * SEGMEAN_STATIC_MAX_SEGS];

#define SEGMEAN_MAX_DEVICES 64
static std::mutex g_segmean_mu;
static int* g_segmean_ptr[SEGMEAN_MAX_DEVICES];    // device address of g_segmean_scratch on each device (host code cannot use the symbol itself)

static int segmean_scratch_ptr(int** out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= SEGMEAN_MAX_DEVICES) {
    svpc_set_error("segmean_fwd_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return SVPC_ERR_CUDA;
  }
  std::lock_guard<std::mutex> lock(g_segmean_mu);
  if (!g_segmean_ptr[dev]) {      // once per device, with that device current
    void* p = nullptr;
    e = cudaGetSymbolAddress(&p, g_segmean_scratch);
    if (e != cudaSuccess || !p) {
      svpc_set_error("segmean_fwd_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return SVPC_ERR_CUDA;
    }
    g_segmean_ptr[dev] = (int*)p;
  }
  *out = g_segmean_ptr[dev];
  return SVPC_OK;
}

extern "C" int svpc_segmean_fwd_static(int B, int Li, int D, int NI, const void* x, int dt, const int* sep_mask,
                                       const int* cu_ingr, const float* pe, float* E0, svpc_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (B <= 0 || NI <= 0) return SVPC_OK;
  SVPC_REQUIRE(Li > 0 && (long long)B * Li <= SEGMEAN_STATIC_MAX_ROWS,
               "segmean_fwd_static: B * Li = %lld exceeds the built-in scratch of %d rows (use svpc_segmean_fwd with caller scratch)",
               (long long)B * Li, SEGMEAN_STATIC_MAX_ROWS);
  SVPC_REQUIRE(NI <= SEGMEAN_STATIC_MAX_SEGS,
               "segmean_fwd_static: NI = %d exceeds the built-in scratch of %d segments (use svpc_segmean_fwd with caller scratch)",
               NI, SEGMEAN_STATIC_MAX_SEGS);
  int* scratch = nullptr;
  int rc = segmean_scratch_ptr(&scratch);
  if (rc) return rc;
  int* row_seg = scratch;
  int* seg_lohi = scratch + SEGMEAN_STATIC_MAX_ROWS;
  // the array keeps the previous call's tables: segments the scan does not reach must not read them
  const cudaError_t e = cudaMemsetAsync(seg_lohi, 0, sizeof(int) * 2 * (size_t)NI, st);
  if (e != cudaSuccess) {
    svpc_set_error("segmean_fwd_static: cudaMemsetAsync: %s", cudaGetErrorString(e));
    return SVPC_ERR_CUDA;
  }
  svpc_launch(seg_scan_kernel, svpc_cdiv(B, 64), 64, 0, st, B, Li, sep_mask, cu_ingr, row_seg, seg_lohi);
  rc = svpc_check_launch("seg_scan");
  if (rc) return rc;
  SVPC_DISPATCH(dt, T, (svpc_launch(segmean_fwd_kernel<T>, NI, 256, 0, st, B, Li, D, (const T*)x, cu_ingr, seg_lohi, pe, E0)));
  return svpc_check_launch("segmean_fwd_static");
}

__global__ void ragged_mean_fwd_kernel(int D, const float* __restrict__ x, const int* __restrict__ cu, float* __restrict__ y) {
  svpc_pdl_sync();
  const int b = blockIdx.x, lo = cu[b], hi = cu[b + 1];
  const float inv = 1.0f / (float)(hi - lo);
  for (int c = threadIdx.x; c < D; c += blockDim.x) {
    float s = 0.f;
    for (int r = lo; r < hi; ++r) s += x[(long long)r * D + c];
    y[(long long)b * D + c] = s * inv;
  }
}
__global__ void ragged_mean_bwd_kernel(int D, const float* __restrict__ dy, const int* __restrict__ cu, float* __restrict__ dx) {
  svpc_pdl_sync();
  const int b = blockIdx.x, lo = cu[b], hi = cu[b + 1];
  const float inv = 1.0f / (float)(hi - lo);
  for (int c = threadIdx.x; c < D; c += blockDim.x) {
    const float g = dy[(long long)b * D + c] * inv;
    for (int r = lo; r < hi; ++r) dx[(long long)r * D + c] = g;
  }
}
extern "C" int svpc_ragged_mean_fwd(int B, int D, const float* x, const int* cu, float* y, svpc_stream_t stream) {
  if (B <= 0) return SVPC_OK;
  svpc_launch(ragged_mean_fwd_kernel, B, 256, 0, (cudaStream_t)stream, D, x, cu, y);
  return svpc_check_launch("ragged_mean_fwd");
}
extern "C" int svpc_ragged_mean_bwd(int B, int D, const float* dy, const int* cu, float* dx, svpc_stream_t stream) {
  if (B <= 0) return SVPC_OK;
  svpc_launch(ragged_mean_bwd_kernel, B, 256, 0, (cudaStream_t)stream, D, dy, cu, dx);
  return svpc_check_launch("ragged_mean_bwd");
}

// y = x / sum(x)  (model.py:798)
__global__ void row_normalize_fwd_kernel(int cols, const float* __restrict__ x, float* __restrict__ y) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const long long base = (long long)blockIdx.x * cols;
  float s = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) s += x[base + c];
  s = block_sum(s, red);
  for (int c = threadIdx.x; c < cols; c += blockDim.x) y[base + c] = x[base + c] / s;
}
// dx = (dy - sum(dy * x) / s) / s
__global__ void row_normalize_bwd_kernel(int cols, const float* __restrict__ dy, const float* __restrict__ x,
                                         float* __restrict__ dx) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const long long base = (long long)blockIdx.x * cols;
  float s = 0.f, t = 0.f;
  for (int c = threadIdx.x; c < cols; c += blockDim.x) { s += x[base + c]; t += dy[base + c] * x[base + c]; }
  const float2 r = block_sum2(s, t, red);
  for (int c = threadIdx.x; c < cols; c += blockDim.x) dx[base + c] = (dy[base + c] - r.y / r.x) / r.x;
}
extern "C" int svpc_row_normalize_fwd(int rows, int cols, const float* x, float* y, svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  svpc_launch(row_normalize_fwd_kernel, rows, 128, 0, (cudaStream_t)stream, cols, x, y);
  return svpc_check_launch("row_normalize_fwd");
}
extern "C" int svpc_row_normalize_bwd(int rows, int cols, const float* dy, const float* x, float* dx, svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  svpc_launch(row_normalize_bwd_kernel, rows, 128, 0, (cudaStream_t)stream, cols, dy, x, dx);
  return svpc_check_launch("row_normalize_bwd");
}
