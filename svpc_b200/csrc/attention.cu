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
