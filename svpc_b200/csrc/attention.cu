// Fused masked softmax attention over short ragged sequences (reference BertSelfAttention core, model.py:204-216):
// clip encoder L=100 (key padding mask), caption decoder L=25 (causal + padding), cross attention to <=3 context
// tokens, step-wise encoder L<=12.  One CTA per (sequence, head): Q, K, V of that head are staged once in shared
// memory (fp32, row pitch dh+1 so a warp reading 32 different rows hits 32 banks), scores / probabilities never leave
// the SM.  The backward recomputes the probabilities instead of storing (N*H*L*L) of them.
// Semantics kept from the reference: scale 1/sqrt(dh) BEFORE the additive mask (1-m)*-10000 (not -inf), fp32 softmax.
#include "common.cuh"

#define AT_THREADS 256

// 8 consecutive elements of a row -> fp32 shared memory (16-byte global loads when the row slice is aligned)
__device__ __forceinline__ void at_ld8(float* dst, const bf16* src) {
  const uint4 u = *reinterpret_cast<const uint4*>(src);
  const __nv_bfloat162* h = reinterpret_cast<const __nv_bfloat162*>(&u);
#pragma unroll
  for (int e = 0; e < 4; ++e) { dst[2 * e] = __low2float(h[e]); dst[2 * e + 1] = __high2float(h[e]); }
}
__device__ __forceinline__ void at_ld8(float* dst, const float* src) {
  const float4 a = *reinterpret_cast<const float4*>(src), b = *reinterpret_cast<const float4*>(src + 4);
  dst[0] = a.x; dst[1] = a.y; dst[2] = a.z; dst[3] = a.w; dst[4] = b.x; dst[5] = b.y; dst[6] = b.z; dst[7] = b.w;
}

template <typename T>
__device__ __forceinline__ void at_load_rows(float* dst, int pitch, const T* src, long long ld, int rows, int dh) {
  if ((dh & 7) == 0 && (ld & 7) == 0 && ((uintptr_t)src & 15) == 0) {
    const int cpr = dh >> 3;                       // 8-element chunks per row
    for (int i = threadIdx.x; i < rows * cpr; i += AT_THREADS) {
      const int r = i / cpr, c = i - r * cpr;
      at_ld8(dst + r * pitch + c * 8, src + (long long)r * ld + c * 8);
    }
    return;
  }
  for (int i = threadIdx.x; i < rows * dh; i += AT_THREADS) {
    const int r = i / dh, d = i - r * dh;
    dst[r * pitch + d] = ldf(src + (long long)r * ld + d);
  }
}

// S[i][j] = (Q_i . K_j) * scale + (1 - m_ij) * -10000 ; then row softmax in place
__device__ __forceinline__ void at_scores_softmax(float* S, const float* Q, const float* K, int lq, int lk, int dh,
                                                  int pitch, float scale, const float* kmask, int causal) {
  for (int idx = threadIdx.x; idx < lq * lk; idx += AT_THREADS) {
    const int i = idx / lk, j = idx - i * lk;
    const float* qi = Q + i * pitch;
    const float* kj = K + j * pitch;
    float s = 0.f;
    for (int d = 0; d < dh; ++d) s = fmaf(qi[d], kj[d], s);
    float m = kmask ? kmask[j] : 1.f;
    if (causal && j > i) m = 0.f;
    S[idx] = s * scale + (1.f - m) * -10000.f;
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = warp; i < lq; i += AT_THREADS / 32) {
    float* row = S + i * lk;
    float mx = -INFINITY;
    for (int j = lane; j < lk; j += 32) mx = fmaxf(mx, row[j]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int j = lane; j < lk; j += 32) { const float e = expf(row[j] - mx); row[j] = e; sum += e; }
    sum = warp_sum(sum);
    const float inv = 1.f / sum;
    for (int j = lane; j < lk; j += 32) row[j] *= inv;
  }
  __syncthreads();
}

template <typename T>
__global__ void __launch_bounds__(AT_THREADS) attn_fwd_kernel(int H, int dh, const T* __restrict__ q, long long ldq,
                                                              const T* __restrict__ k, long long ldk,
                                                              const T* __restrict__ v, long long ldv,
                                                              const int* __restrict__ q_off, const int* __restrict__ q_len,
                                                              const int* __restrict__ k_off, const int* __restrict__ k_len,
                                                              const float* __restrict__ key_mask, int causal, int max_lq,
                                                              int max_lk, T* __restrict__ ctx, long long ldo,
                                                              const SvpcDrop drop) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  const int n = blockIdx.x / H, h = blockIdx.x - n * H;
  const int lq = q_len[n], lk = k_len[n], qo = q_off[n], ko = k_off[n];
  const int pitch = dh + 1;
  float* Q = sm;
  float* K = Q + max_lq * pitch;
  float* V = K + max_lk * pitch;
  float* S = V + max_lk * pitch;
  float* km = S + max_lq * max_lk;
  at_load_rows(Q, pitch, q + (long long)qo * ldq + h * dh, ldq, lq, dh);
  at_load_rows(K, pitch, k + (long long)ko * ldk + h * dh, ldk, lk, dh);
  at_load_rows(V, pitch, v + (long long)ko * ldv + h * dh, ldv, lk, dh);
  if (key_mask)
    for (int j = threadIdx.x; j < lk; j += AT_THREADS) km[j] = key_mask[ko + j];
  __syncthreads();
  at_scores_softmax(S, Q, K, lq, lk, dh, pitch, rsqrtf((float)dh), key_mask ? km : nullptr, causal);
  if (drop.seed != nullptr && drop.p > 0.f) {          // dropout on the probabilities (model.py:214)
    const uint32_t dkey = svpc_drop_key(drop);
    const float dinv = 1.f / (1.f - drop.p);
    const unsigned long long base = (unsigned long long)blockIdx.x * max_lq;
    for (int idx = threadIdx.x; idx < lq * lk; idx += AT_THREADS) {
      const int i = idx / lk, j = idx - i * lk;
      S[idx] *= svpc_drop_scale(dkey, drop.p, dinv, (base + i) * ((max_lk + 1) & ~1) + j);
    }
    __syncthreads();
  }
  for (int idx = threadIdx.x; idx < lq * dh; idx += AT_THREADS) {
    const int i = idx / dh, d = idx - i * dh;
    const float* p = S + i * lk;
    float o = 0.f;
    for (int j = 0; j < lk; ++j) o = fmaf(p[j], V[j * pitch + d], o);
    stf(ctx + (long long)(qo + i) * ldo + h * dh + d, o);
  }
}

template <typename T>
__global__ void __launch_bounds__(AT_THREADS) attn_bwd_kernel(int H, int dh, const T* __restrict__ q, long long ldq,
                                                              const T* __restrict__ k, long long ldk,
                                                              const T* __restrict__ v, long long ldv,
                                                              const int* __restrict__ q_off, const int* __restrict__ q_len,
                                                              const int* __restrict__ k_off, const int* __restrict__ k_len,
                                                              const float* __restrict__ key_mask, int causal, int max_lq,
                                                              int max_lk, const T* __restrict__ dctx, long long lddo,
                                                              T* __restrict__ dq, long long lddq, T* __restrict__ dk,
                                                              long long lddk, T* __restrict__ dv, long long lddv,
                                                              const SvpcDrop drop) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(drop);
  const float dinv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  const unsigned long long dbase = (unsigned long long)blockIdx.x * max_lq;
  const int n = blockIdx.x / H, h = blockIdx.x - n * H;
  const int lq = q_len[n], lk = k_len[n], qo = q_off[n], ko = k_off[n];
  const int pitch = dh + 1;
  const float scale = rsqrtf((float)dh);
  float* Q = sm;
  float* K = Q + max_lq * pitch;
  float* V = K + max_lk * pitch;
  float* dO = V + max_lk * pitch;
  float* P = dO + max_lq * pitch;
  float* dS = P + max_lq * max_lk;
  float* km = dS + max_lq * max_lk;
  at_load_rows(Q, pitch, q + (long long)qo * ldq + h * dh, ldq, lq, dh);
  at_load_rows(K, pitch, k + (long long)ko * ldk + h * dh, ldk, lk, dh);
  at_load_rows(V, pitch, v + (long long)ko * ldv + h * dh, ldv, lk, dh);
  at_load_rows(dO, pitch, dctx + (long long)qo * lddo + h * dh, lddo, lq, dh);
  if (key_mask)
    for (int j = threadIdx.x; j < lk; j += AT_THREADS) km[j] = key_mask[ko + j];
  __syncthreads();
  at_scores_softmax(P, Q, K, lq, lk, dh, pitch, scale, key_mask ? km : nullptr, causal);
  // dV[j][d] = sum_i P[i][j] dO[i][d]
  for (int idx = threadIdx.x; idx < lk * dh; idx += AT_THREADS) {
    const int j = idx / dh, d = idx - j * dh;
    float a = 0.f;
    for (int i = 0; i < lq; ++i) {
      float pij = P[i * lk + j];
      if (dropping) pij *= svpc_drop_scale(dkey, drop.p, dinv, (dbase + i) * ((max_lk + 1) & ~1) + j);
      a = fmaf(pij, dO[i * pitch + d], a);
    }
    stf(dv + (long long)(ko + j) * lddv + h * dh + d, a);
  }
  // dP[i][j] = dO_i . V_j
  for (int idx = threadIdx.x; idx < lq * lk; idx += AT_THREADS) {
    const int i = idx / lk, j = idx - i * lk;
    const float* a = dO + i * pitch;
    const float* b = V + j * pitch;
    float s = 0.f;
    for (int d = 0; d < dh; ++d) s = fmaf(a[d], b[d], s);
    if (dropping) s *= svpc_drop_scale(dkey, drop.p, dinv, (dbase + i) * ((max_lk + 1) & ~1) + j);
    dS[idx] = s;
  }
  __syncthreads();
  // dS = P * (dP - rowsum(dP * P)) * scale
  {
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    for (int i = warp; i < lq; i += AT_THREADS / 32) {
      float t = 0.f;
      for (int j = lane; j < lk; j += 32) t = fmaf(dS[i * lk + j], P[i * lk + j], t);
      t = warp_sum(t);
      for (int j = lane; j < lk; j += 32) dS[i * lk + j] = P[i * lk + j] * (dS[i * lk + j] - t) * scale;
    }
  }
  __syncthreads();
  // dQ[i][d] = sum_j dS[i][j] K[j][d]
  for (int idx = threadIdx.x; idx < lq * dh; idx += AT_THREADS) {
    const int i = idx / dh, d = idx - i * dh;
    float a = 0.f;
    for (int j = 0; j < lk; ++j) a = fmaf(dS[i * lk + j], K[j * pitch + d], a);
    stf(dq + (long long)(qo + i) * lddq + h * dh + d, a);
  }
  // dK[j][d] = sum_i dS[i][j] Q[i][d]
  for (int idx = threadIdx.x; idx < lk * dh; idx += AT_THREADS) {
    const int j = idx / dh, d = idx - j * dh;
    float a = 0.f;
    for (int i = 0; i < lq; ++i) a = fmaf(dS[i * lk + j], Q[i * pitch + d], a);
    stf(dk + (long long)(ko + j) * lddk + h * dh + d, a);
  }
}

// ------------------------------------------------------------------------------------------------------------------
// Warp-per-(sequence, head) kernels for the two tiny attentions of the step: the decoder's cross attention (25 queries
// against <= 3 context rows, model.py:650-653) and the [CLS]-only last encoder layer (1 query against 100 keys).  One
// CTA with its staging and barriers per (sequence, head) costs ~100x the arithmetic there; here a warp owns the pair and
// nothing goes through shared memory.  Same dropout element indexing as the kernels above.
//   K4  (<= 4 keys): K and V rows live in registers (2 of the 64 head dims per lane) for all query rows; the <= 4 scores
//        of a row are warp-reduced independently and the softmax is computed redundantly by every lane.
//   Q1  (1 query, <= 128 keys): scores / dP with one KEY per lane (whole 64-dim rows, no reductions), then the weighted
//        sums with one dim pair per lane and the probabilities broadcast by shuffle.
template <typename T> __device__ __forceinline__ float2 at_ld2(const T* p);
template <> __device__ __forceinline__ float2 at_ld2<float>(const float* p) { return *reinterpret_cast<const float2*>(p); }
template <> __device__ __forceinline__ float2 at_ld2<bf16>(const bf16* p) {
  return __bfloat1622float2(*reinterpret_cast<const __nv_bfloat162*>(p));
}
__device__ __forceinline__ void at_st2(float* p, float2 v) { *reinterpret_cast<float2*>(p) = v; }
__device__ __forceinline__ void at_st2(bf16* p, float2 v) { *reinterpret_cast<__nv_bfloat162*>(p) = __float22bfloat162_rn(v); }

struct AtwCommon {
  int n, h, lq, lk, qo, ko, col, lkp;
  bool dropping;
  uint32_t dkey;
  float dinv;
};
__device__ __forceinline__ AtwCommon atw_setup(int pair, int H, const int* q_off, const int* q_len, const int* k_off,
                                               const int* k_len, int max_lk, const SvpcDrop& drop, int lane) {
  AtwCommon c;
  c.n = pair / H;
  c.h = pair - c.n * H;
  c.lq = q_len[c.n]; c.lk = k_len[c.n]; c.qo = q_off[c.n]; c.ko = k_off[c.n];
  c.col = c.h * 64 + 2 * lane;
  c.lkp = (max_lk + 1) & ~1;
  c.dropping = drop.seed != nullptr && drop.p > 0.f;
  c.dkey = svpc_drop_key(drop);
  c.dinv = c.dropping ? 1.f / (1.f - drop.p) : 1.f;
  return c;
}

// ---- K4: probabilities and dropout scales of one query row, identical in every lane
__device__ __forceinline__ void atw_k4_probs(float (&p)[4], float (&ds)[4], float2 qv, const float2 (&kr)[4],
                                             const float (&km)[4], int lk, int causal, int i, const AtwCommon& c, float dp,
                                             unsigned long long dropbase) {
  float sc[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) sc[j] = qv.x * kr[j].x + qv.y * kr[j].y;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
    for (int j = 0; j < 4; ++j) sc[j] += __shfl_xor_sync(0xffffffffu, sc[j], o);
  }
  float mx = -INFINITY;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    float m = km[j];
    if (causal && j > i) m = 0.f;
    sc[j] = j < lk ? sc[j] * 0.125f + (1.f - m) * -10000.f : -INFINITY;
    mx = fmaxf(mx, sc[j]);
  }
  float sum = 0.f;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    p[j] = j < lk ? expf(sc[j] - mx) : 0.f;
    sum += p[j];
  }
  const float inv = 1.f / sum;
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    p[j] *= inv;
    ds[j] = (c.dropping && j < lk) ? svpc_drop_scale(c.dkey, dp, c.dinv, dropbase + j) : 1.f;
  }
}

template <typename T>
__global__ void __launch_bounds__(256) attn_k4_fwd_kernel(int n_pairs, int H, const T* __restrict__ q, long long ldq,
                                                          const T* __restrict__ k, long long ldk, const T* __restrict__ v,
                                                          long long ldv, const int* __restrict__ q_off,
                                                          const int* __restrict__ q_len, const int* __restrict__ k_off,
                                                          const int* __restrict__ k_len, const float* __restrict__ key_mask,
                                                          int causal, int max_lq, int max_lk, T* __restrict__ ctx,
                                                          long long ldo, const SvpcDrop drop) {
  svpc_pdl_sync();
  const int pair = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (pair >= n_pairs) return;
  const AtwCommon c = atw_setup(pair, H, q_off, q_len, k_off, k_len, max_lk, drop, lane);
  float2 kr[4], vr[4];
  float km[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool ok = j < c.lk;
    kr[j] = ok ? at_ld2(k + (long long)(c.ko + j) * ldk + c.col) : make_float2(0.f, 0.f);
    vr[j] = ok ? at_ld2(v + (long long)(c.ko + j) * ldv + c.col) : make_float2(0.f, 0.f);
    km[j] = (ok && key_mask) ? key_mask[c.ko + j] : 1.f;
  }
  float2 qn = c.lq > 0 ? at_ld2(q + (long long)c.qo * ldq + c.col) : make_float2(0.f, 0.f);
  for (int i = 0; i < c.lq; ++i) {
    const float2 qv = qn;
    if (i + 1 < c.lq) qn = at_ld2(q + (long long)(c.qo + i + 1) * ldq + c.col);
    float p[4], ds[4];
    atw_k4_probs(p, ds, qv, kr, km, c.lk, causal, i, c, drop.p, ((unsigned long long)pair * max_lq + i) * c.lkp);
    float2 o = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float pj = p[j] * ds[j];
      o.x = fmaf(pj, vr[j].x, o.x);
      o.y = fmaf(pj, vr[j].y, o.y);
    }
    at_st2(ctx + (long long)(c.qo + i) * ldo + c.col, o);
  }
}

template <typename T>
__global__ void __launch_bounds__(256) attn_k4_bwd_kernel(int n_pairs, int H, const T* __restrict__ q, long long ldq,
                                                          const T* __restrict__ k, long long ldk, const T* __restrict__ v,
                                                          long long ldv, const int* __restrict__ q_off,
                                                          const int* __restrict__ q_len, const int* __restrict__ k_off,
                                                          const int* __restrict__ k_len, const float* __restrict__ key_mask,
                                                          int causal, int max_lq, int max_lk, const T* __restrict__ dctx,
                                                          long long lddo, T* __restrict__ dq, long long lddq,
                                                          T* __restrict__ dk, long long lddk, T* __restrict__ dv,
                                                          long long lddv, const SvpcDrop drop) {
  svpc_pdl_sync();
  const int pair = blockIdx.x * 8 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (pair >= n_pairs) return;
  const AtwCommon c = atw_setup(pair, H, q_off, q_len, k_off, k_len, max_lk, drop, lane);
  float2 kr[4], vr[4], dka[4], dva[4];
  float km[4];
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const bool ok = j < c.lk;
    kr[j] = ok ? at_ld2(k + (long long)(c.ko + j) * ldk + c.col) : make_float2(0.f, 0.f);
    vr[j] = ok ? at_ld2(v + (long long)(c.ko + j) * ldv + c.col) : make_float2(0.f, 0.f);
    km[j] = (ok && key_mask) ? key_mask[c.ko + j] : 1.f;
    dka[j] = dva[j] = make_float2(0.f, 0.f);
  }
  float2 qn = make_float2(0.f, 0.f), dn = make_float2(0.f, 0.f);
  if (c.lq > 0) {
    qn = at_ld2(q + (long long)c.qo * ldq + c.col);
    dn = at_ld2(dctx + (long long)c.qo * lddo + c.col);
  }
  for (int i = 0; i < c.lq; ++i) {
    const float2 qv = qn, dov = dn;
    if (i + 1 < c.lq) {
      qn = at_ld2(q + (long long)(c.qo + i + 1) * ldq + c.col);
      dn = at_ld2(dctx + (long long)(c.qo + i + 1) * lddo + c.col);
    }
    float p[4], ds[4], dp[4];
    atw_k4_probs(p, ds, qv, kr, km, c.lk, causal, i, c, drop.p, ((unsigned long long)pair * max_lq + i) * c.lkp);
#pragma unroll
    for (int j = 0; j < 4; ++j) dp[j] = dov.x * vr[j].x + dov.y * vr[j].y;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
#pragma unroll
      for (int j = 0; j < 4; ++j) dp[j] += __shfl_xor_sync(0xffffffffu, dp[j], o);
    }
    float t = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      dp[j] *= ds[j];
      t = fmaf(dp[j], p[j], t);
    }
    float2 dqv = make_float2(0.f, 0.f);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const float dS = p[j] * (dp[j] - t) * 0.125f, pd = p[j] * ds[j];
      dqv.x = fmaf(dS, kr[j].x, dqv.x); dqv.y = fmaf(dS, kr[j].y, dqv.y);
      dka[j].x = fmaf(dS, qv.x, dka[j].x); dka[j].y = fmaf(dS, qv.y, dka[j].y);
      dva[j].x = fmaf(pd, dov.x, dva[j].x); dva[j].y = fmaf(pd, dov.y, dva[j].y);
    }
    at_st2(dq + (long long)(c.qo + i) * lddq + c.col, dqv);
  }
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    if (j < c.lk) {
      at_st2(dk + (long long)(c.ko + j) * lddk + c.col, dka[j]);
      at_st2(dv + (long long)(c.ko + j) * lddv + c.col, dva[j]);
    }
  }
}

// ---- Q1: one query row.  dots[s] = x . M_j for key j = s * 32 + lane (whole rows per lane; rows must be 16-byte aligned)
template <typename T>
__device__ __forceinline__ void atw_q1_dots(float (&dots)[4], const T* __restrict__ xrow, const T* __restrict__ mbase,
                                            long long ldm, int lk, int lane) {
  float x[64];
#pragma unroll
  for (int ch = 0; ch < 8; ++ch) at_ld8(x + ch * 8, xrow + ch * 8);
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int j = s * 32 + lane;
    float a = 0.f;
    if (j < lk) {
      const T* row = mbase + (long long)j * ldm;
#pragma unroll
      for (int ch = 0; ch < 8; ++ch) {
        float t[8];
        at_ld8(t, row + ch * 8);
#pragma unroll
        for (int e = 0; e < 8; ++e) a = fmaf(t[e], x[ch * 8 + e], a);
      }
    }
    dots[s] = a;
  }
}
// softmax over the <= 128 scores held one per (lane, slot); returns probabilities and dropout scales
__device__ __forceinline__ void atw_q1_probs(float (&p)[4], float (&ds)[4], const float* __restrict__ km, int lk, int lane,
                                             const AtwCommon& c, float dp, unsigned long long dropbase) {
  float mx = -INFINITY;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int j = s * 32 + lane;
    const float m = (km && j < lk) ? km[j] : 1.f;
    p[s] = j < lk ? p[s] * 0.125f + (1.f - m) * -10000.f : -INFINITY;
    mx = fmaxf(mx, p[s]);
  }
  mx = warp_max(mx);
  float sum = 0.f;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    p[s] = (s * 32 + lane < lk) ? expf(p[s] - mx) : 0.f;
    sum += p[s];
  }
  sum = warp_sum(sum);
  const float inv = 1.f / sum;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    p[s] *= inv;
    ds[s] = (c.dropping && s * 32 + lane < lk) ? svpc_drop_scale(c.dkey, dp, c.dinv, dropbase + s * 32 + lane) : 1.f;
  }
}

template <typename T>
__global__ void __launch_bounds__(128) attn_q1_fwd_kernel(int n_pairs, int H, const T* __restrict__ q, long long ldq,
                                                          const T* __restrict__ k, long long ldk, const T* __restrict__ v,
                                                          long long ldv, const int* __restrict__ q_off,
                                                          const int* __restrict__ q_len, const int* __restrict__ k_off,
                                                          const int* __restrict__ k_len, const float* __restrict__ key_mask,
                                                          int max_lq, int max_lk, T* __restrict__ ctx, long long ldo,
                                                          const SvpcDrop drop) {
  svpc_pdl_sync();
  const int pair = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (pair >= n_pairs) return;
  const AtwCommon c = atw_setup(pair, H, q_off, q_len, k_off, k_len, max_lk, drop, lane);
  if (c.lq < 1) return;
  float p[4], ds[4];
  atw_q1_dots(p, q + (long long)c.qo * ldq + c.h * 64, k + (long long)c.ko * ldk + c.h * 64, ldk, c.lk, lane);
  atw_q1_probs(p, ds, key_mask ? key_mask + c.ko : nullptr, c.lk, lane, c, drop.p, (unsigned long long)pair * max_lq * c.lkp);
  const T* vb = v + (long long)c.ko * ldv + c.col;
  float2 o = make_float2(0.f, 0.f);
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int jn = min(32, c.lk - s * 32);
    const float pd = p[s] * ds[s];
#pragma unroll 8
    for (int jj = 0; jj < jn; ++jj) {
      const float pj = __shfl_sync(0xffffffffu, pd, jj);
      const float2 vv = at_ld2(vb + (long long)(s * 32 + jj) * ldv);
      o.x = fmaf(pj, vv.x, o.x);
      o.y = fmaf(pj, vv.y, o.y);
    }
  }
  at_st2(ctx + (long long)c.qo * ldo + c.col, o);
}

template <typename T>
__global__ void __launch_bounds__(128) attn_q1_bwd_kernel(int n_pairs, int H, const T* __restrict__ q, long long ldq,
                                                          const T* __restrict__ k, long long ldk, const T* __restrict__ v,
                                                          long long ldv, const int* __restrict__ q_off,
                                                          const int* __restrict__ q_len, const int* __restrict__ k_off,
                                                          const int* __restrict__ k_len, const float* __restrict__ key_mask,
                                                          int max_lq, int max_lk, const T* __restrict__ dctx, long long lddo,
                                                          T* __restrict__ dq, long long lddq, T* __restrict__ dk,
                                                          long long lddk, T* __restrict__ dv, long long lddv,
                                                          const SvpcDrop drop) {
  svpc_pdl_sync();
  const int pair = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (pair >= n_pairs) return;
  const AtwCommon c = atw_setup(pair, H, q_off, q_len, k_off, k_len, max_lk, drop, lane);
  if (c.lq < 1) return;
  float p[4], ds[4], dp[4];
  atw_q1_dots(p, q + (long long)c.qo * ldq + c.h * 64, k + (long long)c.ko * ldk + c.h * 64, ldk, c.lk, lane);
  atw_q1_probs(p, ds, key_mask ? key_mask + c.ko : nullptr, c.lk, lane, c, drop.p, (unsigned long long)pair * max_lq * c.lkp);
  atw_q1_dots(dp, dctx + (long long)c.qo * lddo + c.h * 64, v + (long long)c.ko * ldv + c.h * 64, ldv, c.lk, lane);
  float t = 0.f;
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    dp[s] *= ds[s];
    t = fmaf(dp[s], p[s], t);
  }
  t = warp_sum(t);
  const float2 qv = at_ld2(q + (long long)c.qo * ldq + c.col), dov = at_ld2(dctx + (long long)c.qo * lddo + c.col);
  const T* kb = k + (long long)c.ko * ldk + c.col;
  float2 dqv = make_float2(0.f, 0.f);
#pragma unroll
  for (int s = 0; s < 4; ++s) {
    const int jn = min(32, c.lk - s * 32);
    const float dS = p[s] * (dp[s] - t) * 0.125f, pd = p[s] * ds[s];
#pragma unroll 8
    for (int jj = 0; jj < jn; ++jj) {
      const int j = s * 32 + jj;
      const float dSj = __shfl_sync(0xffffffffu, dS, jj), pj = __shfl_sync(0xffffffffu, pd, jj);
      const float2 kv = at_ld2(kb + (long long)j * ldk);
      dqv.x = fmaf(dSj, kv.x, dqv.x); dqv.y = fmaf(dSj, kv.y, dqv.y);
      at_st2(dk + (long long)(c.ko + j) * lddk + c.col, make_float2(dSj * qv.x, dSj * qv.y));
      at_st2(dv + (long long)(c.ko + j) * lddv + c.col, make_float2(pj * dov.x, pj * dov.y));
    }
  }
  at_st2(dq + (long long)c.qo * lddq + c.col, dqv);
}

// eligibility of the warp kernels: head dim 64, rows addressable as aligned element pairs
static bool atw_layout_ok(int dh, int dt, long long a, long long b, long long c, long long d, const void* p0, const void* p1,
                          const void* p2, const void* p3) {
  const uintptr_t al = dt == SVPC_BF16 ? 3 : 7;
  return dh == 64 && ((a | b | c | d) & 1) == 0 &&
         ((((uintptr_t)p0 | (uintptr_t)p1 | (uintptr_t)p2 | (uintptr_t)p3) & al) == 0);
}

// whole 64-element rows loadable as 16-byte chunks (the Q1 kernels read one key row per lane)
static bool atw_rows16(int dt, long long a, long long b, long long c, const void* p0, const void* p1, const void* p2) {
  const long long q = dt == SVPC_BF16 ? 8 : 4;
  return (a % q) == 0 && (b % q) == 0 && (c % q) == 0 && ((((uintptr_t)p0 | (uintptr_t)p1 | (uintptr_t)p2) & 15) == 0);
}

static size_t at_smem_fwd(int max_lq, int max_lk, int dh) {
  return sizeof(float) * ((size_t)(max_lq + 2 * max_lk) * (dh + 1) + (size_t)max_lq * max_lk + max_lk);
}
static size_t at_smem_bwd(int max_lq, int max_lk, int dh) {
  return sizeof(float) * ((size_t)(2 * max_lq + 2 * max_lk) * (dh + 1) + 2 * (size_t)max_lq * max_lk + max_lk);
}

// tensor-core route (attention_tc.cu): bf16, head dim 64, 16 <= L <= 128
int svpc_attn_tc_eligible(int dh, int dt, int max_lq, int max_lk, long long ldq, long long ldk, long long ldv, long long ldo,
                          const void* q, const void* k, const void* v, const void* o);
int svpc_attn_tc_fwd(int n_seq, int H, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                     long long ldv, const int* q_off, const int* q_len, const int* k_off, const int* k_len,
                     const float* key_mask, int causal, int max_lq, int max_lk, void* ctx, long long ldo, SvpcDrop drop,
                     cudaStream_t st);
int svpc_attn_tc_bwd(int n_seq, int H, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                     long long ldv, const int* q_off, const int* q_len, const int* k_off, const int* k_len,
                     const float* key_mask, int causal, int max_lq, int max_lk, const void* dctx, long long lddo, void* dq,
                     long long lddq, void* dk, long long lddk, void* dv, long long lddv, SvpcDrop drop, cudaStream_t st);

extern "C" int svpc_attn_fwd(int n_seq, int H, int dh, const void* q, long long ldq, const void* k, long long ldk,
                             const void* v, long long ldv, int dt, const int* q_off, const int* q_len, const int* k_off,
                             const int* k_len, const float* key_mask, int causal, int max_lq, int max_lk, void* ctx,
                             long long ldo, const unsigned int* drop_seed, unsigned int drop_site, float drop_p,
                             svpc_stream_t stream) {
  if (n_seq <= 0) return SVPC_OK;
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  if (svpc_attn_tc_eligible(dh, dt, max_lq, max_lk, ldq, ldk, ldv, ldo, q, k, v, ctx))
    return svpc_attn_tc_fwd(n_seq, H, q, ldq, k, ldk, v, ldv, q_off, q_len, k_off, k_len, key_mask, causal, max_lq, max_lk,
                            ctx, ldo, drop, (cudaStream_t)stream);
  if (atw_layout_ok(dh, dt, ldq, ldk, ldv, ldo, q, k, v, ctx) && (max_lk <= 4 || (max_lq == 1 && max_lk <= 128 && !causal &&
                                                                 atw_rows16(dt, ldq, ldk, ldv, q, k, v)))) {
    const int pairs = n_seq * H;
    cudaStream_t st_ = (cudaStream_t)stream;
    if (max_lk <= 4) {
      SVPC_DISPATCH(dt, T, (svpc_launch(attn_k4_fwd_kernel<T>, svpc_cdiv(pairs, 8), 256, 0, st_, pairs, H, (const T*)q, ldq,
                                        (const T*)k, ldk, (const T*)v, ldv, q_off, q_len, k_off, k_len, key_mask, causal, max_lq,
                                        max_lk, (T*)ctx, ldo, drop)));
    } else {
      SVPC_DISPATCH(dt, T, (svpc_launch(attn_q1_fwd_kernel<T>, svpc_cdiv(pairs, 4), 128, 0, st_, pairs, H, (const T*)q, ldq,
                                        (const T*)k, ldk, (const T*)v, ldv, q_off, q_len, k_off, k_len, key_mask, max_lq, max_lk,
                                        (T*)ctx, ldo, drop)));
    }
    return svpc_check_launch("attn_warp_fwd");
  }
  const size_t smem = at_smem_fwd(max_lq, max_lk, dh);
  SVPC_REQUIRE(smem <= 227 * 1024, "attn_fwd: sequence too long for the short-sequence kernel (lq=%d lk=%d)", max_lq, max_lk);
  cudaStream_t st = (cudaStream_t)stream;
  SVPC_DISPATCH(dt, T, {
    cudaFuncSetAttribute(attn_fwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    svpc_launch(attn_fwd_kernel<T>, n_seq * H, AT_THREADS, smem, st, H, dh, (const T*)q, ldq, (const T*)k, ldk, (const T*)v, ldv, q_off,
                                                            q_len, k_off, k_len, key_mask, causal, max_lq, max_lk, (T*)ctx, ldo, drop);
  });
  return svpc_check_launch("attn_fwd");
}

extern "C" int svpc_attn_bwd(int n_seq, int H, int dh, const void* q, long long ldq, const void* k, long long ldk,
                             const void* v, long long ldv, int dt, const int* q_off, const int* q_len, const int* k_off,
                             const int* k_len, const float* key_mask, int causal, int max_lq, int max_lk, const void* dctx,
                             long long lddo, void* dq, long long lddq, void* dk, long long lddk, void* dv, long long lddv,
                             const unsigned int* drop_seed, unsigned int drop_site, float drop_p, svpc_stream_t stream) {
  if (n_seq <= 0) return SVPC_OK;
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  if (svpc_attn_tc_eligible(dh, dt, max_lq, max_lk, ldq, ldk, ldv, lddo, q, k, v, dctx) && (lddq | lddk | lddv) % 2 == 0 &&
      (((uintptr_t)dq | (uintptr_t)dk | (uintptr_t)dv) & 3) == 0 && lddo % 8 == 0 && ((uintptr_t)dctx & 15) == 0)
    return svpc_attn_tc_bwd(n_seq, H, q, ldq, k, ldk, v, ldv, q_off, q_len, k_off, k_len, key_mask, causal, max_lq, max_lk,
                            dctx, lddo, dq, lddq, dk, lddk, dv, lddv, drop, (cudaStream_t)stream);
  if (atw_layout_ok(dh, dt, ldq, ldk, ldv, lddo, q, k, v, dctx) && atw_layout_ok(dh, dt, lddq, lddk, lddv, 0, dq, dk, dv, nullptr) &&
      (max_lk <= 4 || (max_lq == 1 && max_lk <= 128 && !causal && atw_rows16(dt, ldq, ldk, ldv, q, k, v) &&
                       atw_rows16(dt, lddo, 0, 0, dctx, nullptr, nullptr)))) {
    const int pairs = n_seq * H;
    cudaStream_t st_ = (cudaStream_t)stream;
    if (max_lk <= 4) {
      SVPC_DISPATCH(dt, T, (svpc_launch(attn_k4_bwd_kernel<T>, svpc_cdiv(pairs, 8), 256, 0, st_, pairs, H, (const T*)q, ldq,
                                        (const T*)k, ldk, (const T*)v, ldv, q_off, q_len, k_off, k_len, key_mask, causal, max_lq,
                                        max_lk, (const T*)dctx, lddo, (T*)dq, lddq, (T*)dk, lddk, (T*)dv, lddv, drop)));
    } else {
      SVPC_DISPATCH(dt, T, (svpc_launch(attn_q1_bwd_kernel<T>, svpc_cdiv(pairs, 4), 128, 0, st_, pairs, H, (const T*)q, ldq,
                                        (const T*)k, ldk, (const T*)v, ldv, q_off, q_len, k_off, k_len, key_mask, max_lq, max_lk,
                                        (const T*)dctx, lddo, (T*)dq, lddq, (T*)dk, lddk, (T*)dv, lddv, drop)));
    }
    return svpc_check_launch("attn_warp_bwd");
  }
  const size_t smem = at_smem_bwd(max_lq, max_lk, dh);
  SVPC_REQUIRE(smem <= 227 * 1024, "attn_bwd: sequence too long for the short-sequence kernel (lq=%d lk=%d)", max_lq, max_lk);
  cudaStream_t st = (cudaStream_t)stream;
  SVPC_DISPATCH(dt, T, {
    cudaFuncSetAttribute(attn_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    svpc_launch(attn_bwd_kernel<T>, n_seq * H, AT_THREADS, smem, st, H, dh, (const T*)q, ldq, (const T*)k, ldk, (const T*)v, ldv, q_off,
                                                            q_len, k_off, k_len, key_mask, causal, max_lq, max_lk,
                                                            (const T*)dctx, lddo, (T*)dq, lddq, (T*)dk, lddk, (T*)dv, lddv, drop);
  });
  return svpc_check_launch("attn_bwd");
}
