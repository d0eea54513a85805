// Entity-state recurrence of the visual simulator / textual re-simulator (reference EntitiyReasoningNetwork.forward,
// model.py:792-820).  Everything that depends only on the step vector (action selector, W1, W2, W3, W4) is batched
// into GEMMs over all clips by the host; what remains is sequential over the <=12 clips of a video and touches only
// the (Ni, D) entity matrix, so it is ONE persistent CTA per video with E resident in shared memory:
//     e   = sigmoid(E w2v_t)                       (Eq. 2)
//     a   = c0 e + c1 e_prev, c = softmax(clog_t)  (Eq. 3; the third mixture weight multiplies zeros)
//     eb  = (a / sum a) E                          (Eq. 4-5)
//     k   = relu(w4s_t * eb)                       (Eq. 6)
//     E  <- a k^T + (1 - a) (.) E                  (Eq. 7);   e_prev <- e   (model.py:813: ent_prob, not attn_dist)
// The backward walks the clips in reverse with dE resident in shared memory.
#include "common.cuh"

#define RS_THREADS 1024

__global__ void __launch_bounds__(RS_THREADS) reasoner_fwd_kernel(int D, int max_ni, const float* __restrict__ E0,
                                                                  const int* __restrict__ cu_ingr,
                                                                  const int* __restrict__ cu_steps,
                                                                  const int* __restrict__ clip_ent_off,
                                                                  const float* __restrict__ w2v,
                                                                  const float* __restrict__ clog,
                                                                  const float* __restrict__ w4s, float* __restrict__ ent,
                                                                  float* __restrict__ ebar_out, float* __restrict__ Es,
                                                                  float* __restrict__ a_saved) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  float* E = sm;                       // (max_ni, D)
  float* kv = E + (size_t)max_ni * D;  // (D)
  float* e = kv + D;                   // (max_ni)
  float* ep = e + max_ni;              // (max_ni)
  float* a = ep + max_ni;              // (max_ni)
  float* an = a + max_ni;              // (max_ni) a / sum a
  __shared__ float red[64];
  const int b = blockIdx.x;
  const int g0 = cu_ingr[b], ni = cu_ingr[b + 1] - g0;
  const int n0 = cu_steps[b], n1 = cu_steps[b + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < ni * D; i += RS_THREADS) E[i] = E0[(long long)g0 * D + i];
  for (int i = threadIdx.x; i < ni; i += RS_THREADS) ep[i] = 0.f;
  __syncthreads();
  for (int n = n0; n < n1; ++n) {
    const float* w = w2v + (long long)n * D;
    for (int i = warp; i < ni; i += RS_THREADS / 32) {
      float s = 0.f;
      for (int d = lane; d < D; d += 32) s = fmaf(E[i * D + d], w[d], s);
      s = warp_sum(s);
      if (lane == 0) e[i] = svpc_sigmoid(s);
    }
    __syncthreads();
    // softmax over the 3 choice logits
    const float l0 = clog[3 * n], l1 = clog[3 * n + 1], l2 = clog[3 * n + 2];
    const float mx = fmaxf(l0, fmaxf(l1, l2));
    const float x0 = expf(l0 - mx), x1 = expf(l1 - mx), x2 = expf(l2 - mx);
    const float c0 = x0 / (x0 + x1 + x2), c1 = x1 / (x0 + x1 + x2);
    float part = 0.f;
    for (int i = threadIdx.x; i < ni; i += RS_THREADS) {
      const float ai = c0 * e[i] + c1 * ep[i];
      a[i] = ai;
      part += ai;
    }
    const float w4 = w4s[n];                   // (issued before the reduction: its latency hides behind the barriers)
    const long long eo = clip_ent_off[n];
    const float asum = block_sum(part, red);   // (syncs: a[] visible)
    for (int i = threadIdx.x; i < ni; i += RS_THREADS) an[i] = a[i] / asum;   // one division per ingredient, not per element
    __syncthreads();
    for (int d = threadIdx.x; d < D; d += RS_THREADS) {
      float s = 0.f;
      for (int i = 0; i < ni; ++i) s = fmaf(an[i], E[i * D + d], s);
      ebar_out[(long long)n * D + d] = s;
      const float kk = w4 * s;
      kv[d] = kk > 0.f ? kk : 0.f;
    }
    __syncthreads();
    for (int i = warp; i < ni; i += RS_THREADS / 32) {           // one warp per entity row: no index division
      const float ai = a[i], bi = 1.f - ai;
      float* Ei = E + i * D;
      float* Eo = Es + (eo + i) * D;
      for (int d = lane; d < D; d += 32) {
        const float v = ai * kv[d] + bi * Ei[d];
        Ei[d] = v;
        Eo[d] = v;
      }
    }
    for (int i = threadIdx.x; i < ni; i += RS_THREADS) {
      ent[eo + i] = e[i];
      a_saved[eo + i] = a[i];
      ep[i] = e[i];
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(RS_THREADS) reasoner_bwd_kernel(
    int D, int max_ni, const float* __restrict__ E0, const int* __restrict__ cu_ingr, const int* __restrict__ cu_steps,
    const int* __restrict__ clip_ent_off, const float* __restrict__ w2v, const float* __restrict__ clog,
    const float* __restrict__ w4s, const float* __restrict__ ent, const float* __restrict__ ebar,
    const float* __restrict__ Es, const float* __restrict__ a_saved, const float* __restrict__ d_ent,
    const float* __restrict__ d_ebar, const float* __restrict__ d_Es, float* __restrict__ dE0, float* __restrict__ dw2v,
    float* __restrict__ dclog, float* __restrict__ dw4s) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  float* dE = sm;                        // (max_ni, D) gradient w.r.t. the current entity matrix
  float* Ep = dE + (size_t)max_ni * D;   // (max_ni, D) entity matrix BEFORE this step
  float* kv = Ep + (size_t)max_ni * D;   // (D) k
  float* dkv = kv + D;                   // (D) dk, then d(pre-relu)
  float* deb = dkv + D;                  // (D) d ebar
  float* da = deb + D;                   // (max_ni)
  float* dan = da + max_ni;              // (max_ni)
  float* de_next = dan + max_ni;         // (max_ni) gradient flowing into e_t from step t+1 (through e_prev)
  float* dz = de_next + max_ni;          // (max_ni)
  float* avs = dz + max_ni;              // (max_ni) a of this step
  __shared__ float red[64];
  const int b = blockIdx.x;
  const int g0 = cu_ingr[b], ni = cu_ingr[b + 1] - g0;
  const int n0 = cu_steps[b], n1 = cu_steps[b + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < ni * D; i += RS_THREADS) dE[i] = 0.f;
  for (int i = threadIdx.x; i < ni; i += RS_THREADS) de_next[i] = 0.f;
  __syncthreads();
  for (int n = n1 - 1; n >= n0; --n) {
    const long long eo = clip_ent_off[n];
    const float* Eprev = (n == n0) ? E0 + (long long)g0 * D : Es + (long long)clip_ent_off[n - 1] * D;
    const float* av = a_saved + eo;
    const float* ev = ent + eo;
    const float* epv = (n == n0) ? nullptr : ent + clip_ent_off[n - 1];
    const float w4 = w4s[n];
    const float l0 = clog[3 * n], l1 = clog[3 * n + 1], l2 = clog[3 * n + 2];
    const float mx = fmaxf(l0, fmaxf(l1, l2));
    const float x0 = expf(l0 - mx), x1 = expf(l1 - mx), x2 = expf(l2 - mx);
    const float c0 = x0 / (x0 + x1 + x2), c1 = x1 / (x0 + x1 + x2), c2 = x2 / (x0 + x1 + x2);
    float part = 0.f;
    for (int i = threadIdx.x; i < ni; i += RS_THREADS) {
      const float ai = av[i];
      avs[i] = ai;                             // the per-element loops below read a from shared memory, not from L2
      part += ai;
    }
    const float asum = block_sum(part, red);   // (syncs: avs visible)
    // stage E_{t-1}, k ; dE += dEs[n]
    for (int idx = threadIdx.x; idx < ni * D; idx += RS_THREADS) {
      Ep[idx] = Eprev[idx];
      dE[idx] += d_Es[eo * D + idx];
    }
    for (int d = threadIdx.x; d < D; d += RS_THREADS) {
      const float kk = w4 * ebar[(long long)n * D + d];
      kv[d] = kk > 0.f ? kk : 0.f;
    }
    __syncthreads();
    // da_i = sum_d dE[i][d] (k[d] - Ep[i][d])
    for (int i = warp; i < ni; i += RS_THREADS / 32) {
      float s = 0.f;
      for (int d = lane; d < D; d += 32) s = fmaf(dE[i * D + d], kv[d] - Ep[i * D + d], s);
      s = warp_sum(s);
      if (lane == 0) da[i] = s;
    }
    // dk[d] = sum_i a_i dE[i][d] ; dpre = dk * [k>0] ; d ebar = dpre * w4 + d_ebar_out
    float pw4 = 0.f;
    for (int d = threadIdx.x; d < D; d += RS_THREADS) {
      float s = 0.f;
      for (int i = 0; i < ni; ++i) s = fmaf(avs[i], dE[i * D + d], s);
      const float eb = ebar[(long long)n * D + d];
      const float dpre = (w4 * eb > 0.f) ? s : 0.f;
      pw4 = fmaf(dpre, eb, pw4);
      deb[d] = dpre * w4 + d_ebar[(long long)n * D + d];
    }
    const float dw4 = block_sum(pw4, red);     // (syncs: da, deb visible)
    if (threadIdx.x == 0) dw4s[n] = dw4;
    // dan_i = deb . Ep_i
    for (int i = warp; i < ni; i += RS_THREADS / 32) {
      float s = 0.f;
      for (int d = lane; d < D; d += 32) s = fmaf(deb[d], Ep[i * D + d], s);
      s = warp_sum(s);
      if (lane == 0) dan[i] = s;
    }
    __syncthreads();
    float pt = 0.f;
    for (int i = threadIdx.x; i < ni; i += RS_THREADS) pt = fmaf(dan[i], av[i] / asum, pt);
    const float tdot = block_sum(pt, red);
    // da += (dan - tdot) / asum ; mixture / sigmoid backward
    float p0 = 0.f, p1 = 0.f;
    for (int i = threadIdx.x; i < ni; i += RS_THREADS) {
      const float dai = da[i] + (dan[i] - tdot) / asum;
      da[i] = dai;
      const float ei = ev[i];
      const float epi = epv ? epv[i] : 0.f;
      p0 = fmaf(dai, ei, p0);
      p1 = fmaf(dai, epi, p1);
      const float dei = c0 * dai + d_ent[eo + i] + de_next[i];
      de_next[i] = c1 * dai;
      dz[i] = dei * ei * (1.f - ei);
    }
    const float2 dc = block_sum2(p0, p1, red);   // (syncs: dz visible)
    if (threadIdx.x == 0) {
      const float dot = c0 * dc.x + c1 * dc.y;   // dc2 = 0
      dclog[3 * n + 0] = c0 * (dc.x - dot);
      dclog[3 * n + 1] = c1 * (dc.y - dot);
      dclog[3 * n + 2] = c2 * (0.f - dot);
    }
    // dw2v[d] = sum_i dz_i Ep[i][d] ; dE_prev = (1 - a_i) dE + an_i deb + dz_i w2v
    const float* w = w2v + (long long)n * D;
    for (int d = threadIdx.x; d < D; d += RS_THREADS) {
      float s = 0.f;
      for (int i = 0; i < ni; ++i) s = fmaf(dz[i], Ep[i * D + d], s);
      dw2v[(long long)n * D + d] = s;
    }
    for (int i = warp; i < ni; i += RS_THREADS / 32) {           // one warp per entity row: no index division
      const float ai = avs[i], bi = 1.f - ai, ani = ai / asum, dzi = dz[i];
      float* dEi = dE + i * D;
      for (int d = lane; d < D; d += 32) dEi[d] = bi * dEi[d] + ani * deb[d] + dzi * w[d];
    }
    __syncthreads();
  }
  for (int i = threadIdx.x; i < ni * D; i += RS_THREADS) dE0[(long long)g0 * D + i] = dE[i];
}

extern "C" int svpc_reasoner_fwd(int B, int D, int max_ni, const float* E0, const int* cu_ingr, const int* cu_steps,
                                 const int* clip_ent_off, const float* w2v, const float* clog, const float* w4s,
                                 float* ent, float* ebar, float* Es, float* a_saved, svpc_stream_t stream) {
  if (B <= 0) return SVPC_OK;
  const size_t smem = sizeof(float) * ((size_t)max_ni * D + D + 4 * (size_t)max_ni);
  SVPC_REQUIRE(smem <= 220 * 1024, "reasoner_fwd: %d ingredients x %d do not fit in shared memory", max_ni, D);
  cudaFuncSetAttribute(reasoner_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  svpc_launch(reasoner_fwd_kernel, B, RS_THREADS, smem, (cudaStream_t)stream, D, max_ni, E0, cu_ingr, cu_steps, clip_ent_off, w2v, clog,
                                                                    w4s, ent, ebar, Es, a_saved);
  return svpc_check_launch("reasoner_fwd");
}

extern "C" int svpc_reasoner_bwd(int B, int D, int max_ni, const float* E0, const int* cu_ingr, const int* cu_steps,
                                 const int* clip_ent_off, const float* w2v, const float* clog, const float* w4s,
                                 const float* ent, const float* ebar, const float* Es, const float* a_saved,
                                 const float* d_ent, const float* d_ebar, const float* d_Es, float* dE0, float* dw2v,
                                 float* dclog, float* dw4s, svpc_stream_t stream) {
  if (B <= 0) return SVPC_OK;
  const size_t smem = sizeof(float) * (2 * (size_t)max_ni * D + 3 * (size_t)D + 5 * (size_t)max_ni);
  SVPC_REQUIRE(smem <= 220 * 1024, "reasoner_bwd: %d ingredients x %d do not fit in shared memory", max_ni, D);
  cudaFuncSetAttribute(reasoner_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  svpc_launch(reasoner_bwd_kernel, B, RS_THREADS, smem, (cudaStream_t)stream, D, max_ni, E0, cu_ingr, cu_steps, clip_ent_off, w2v, clog,
                                                                    w4s, ent, ebar, Es, a_saved, d_ent, d_ebar, d_Es, dE0,
                                                                    dw2v, dclog, dw4s);
  return svpc_check_launch("reasoner_bwd");
}
