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


// ------------------------------------------------------------------------------------------------
// Pipelined variants (used whenever their shared memory fits): the kernels above spend most of every step waiting -
// a dozen block barriers and a chain of L2 round trips (w2v row, choice logits, the (Ni, D) matrices of the backward).
// Here every global read of step t+1 is in flight (cp.async into a second buffer) while step t computes, the small
// per-ingredient quantities (mixture, normalisation, their gradients) are recomputed redundantly by every thread from
// shared memory instead of being block-reduced, and a thread OWNS a column d of the entity matrix through the
// column-wise phases, so a step needs two block barriers.
// ------------------------------------------------------------------------------------------------
#define RS_MAXSTEPS 64      // steps of a video whose per-step scalars the pipelined kernels keep in shared memory
__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gsrc) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"((uint32_t)__cvta_generic_to_shared(smem_dst)), "l"(gsrc) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
__device__ __forceinline__ void cp_async_wait_all() { asm volatile("cp.async.wait_group 0;" ::: "memory"); }
// n floats (n % 4 == 0, both sides 16-byte aligned) global -> shared, spread over the CTA
__device__ __forceinline__ void stage_async(float* dst, const float* src, int n) {
  for (int i = threadIdx.x * 4; i < n; i += RS_THREADS * 4) cp_async16(dst + i, src + i);
}

__global__ void __launch_bounds__(RS_THREADS) reasoner_fwd_pipe_kernel(int D, int max_ni, int max_steps, const float* __restrict__ E0,
                                                                       const int* __restrict__ cu_ingr,
                                                                       const int* __restrict__ cu_steps,
                                                                       const int* __restrict__ clip_ent_off,
                                                                       const float* __restrict__ w2v,
                                                                       const float* __restrict__ clog,
                                                                       const float* __restrict__ w4s, float* __restrict__ ent,
                                                                       float* __restrict__ ebar_out, float* __restrict__ Es,
                                                                       float* __restrict__ a_saved) {
  svpc_pdl_sync();
  extern __shared__ __align__(16) float sm[];
  float* E = sm;                              // (max_ni, D)
  float* wb = E + (size_t)max_ni * D;         // (steps, D): the w2v rows of ALL steps of this video, staged up front
  float* eb = wb + (size_t)max_steps * D;     // 2 x (max_ni): e of this / the previous step
  const int b = blockIdx.x;
  const int g0 = cu_ingr[b], ni = cu_ingr[b + 1] - g0;
  const int n0 = cu_steps[b], n1 = cu_steps[b + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (n1 <= n0) return;
  stage_async(E, E0 + (long long)g0 * D, ni * D);
  stage_async(wb, w2v + (long long)n0 * D, (n1 - n0) * D);
  cp_async_commit();
  // per-step scalars of the whole video up front (a global load inside the step loop is an exposed L2 round trip per step)
  __shared__ float s_cl[3 * RS_MAXSTEPS], s_w4[RS_MAXSTEPS];
  __shared__ int s_eo[RS_MAXSTEPS];
  for (int k = threadIdx.x; k < min(n1 - n0, RS_MAXSTEPS); k += RS_THREADS) {
    s_cl[3 * k] = clog[3 * (n0 + k)]; s_cl[3 * k + 1] = clog[3 * (n0 + k) + 1]; s_cl[3 * k + 2] = clog[3 * (n0 + k) + 2];
    s_w4[k] = w4s[n0 + k];
    s_eo[k] = clip_ent_off[n0 + k];
  }
  for (int i = threadIdx.x; i < 2 * max_ni; i += RS_THREADS) eb[i] = 0.f;
  cp_async_wait_all();
  __syncthreads();
  for (int n = n0; n < n1; ++n) {
    const int cur = (n - n0) & 1, k = n - n0;
    const float* w = wb + (size_t)(n - n0) * D;
    float* e = eb + cur * max_ni;
    const float* ep = eb + (cur ^ 1) * max_ni;
    const bool st = k < RS_MAXSTEPS;
    const float l0 = st ? s_cl[3 * k] : clog[3 * n], l1 = st ? s_cl[3 * k + 1] : clog[3 * n + 1], l2 = st ? s_cl[3 * k + 2] : clog[3 * n + 2];
    const float w4 = st ? s_w4[k] : w4s[n];
    const long long eo = st ? s_eo[k] : clip_ent_off[n];
    // (1) e_i = sigmoid(E_i . w): one warp per entity row
    for (int i = warp; i < ni; i += RS_THREADS / 32) {
      float s = 0.f;
      for (int d = lane; d < D; d += 32) s = fmaf(E[i * D + d], w[d], s);
      s = warp_sum(s);
      if (lane == 0) e[i] = svpc_sigmoid(s);
    }
    __syncthreads();
    // (2) mixture a = c0 e + c1 e_prev and its sum, recomputed by every thread (ni broadcast reads)
    const float mx = fmaxf(l0, fmaxf(l1, l2));
    const float x0 = expf(l0 - mx), x1 = expf(l1 - mx), x2 = expf(l2 - mx);
    const float c0 = x0 / (x0 + x1 + x2), c1 = x1 / (x0 + x1 + x2);
    float asum = 0.f;
    for (int i = 0; i < ni; ++i) asum += c0 * e[i] + c1 * ep[i];
    const float inv_asum = 1.f / asum;
    if (threadIdx.x < ni) {
      const int i = threadIdx.x;
      ent[eo + i] = e[i];
      a_saved[eo + i] = c0 * e[i] + c1 * ep[i];
    }
    // (3) the thread that owns column d: ebar_d, k_d, and the update of E[:, d]
    for (int d = threadIdx.x; d < D; d += RS_THREADS) {
      float s = 0.f;
      for (int i = 0; i < ni; ++i) s = fmaf((c0 * e[i] + c1 * ep[i]) * inv_asum, E[i * D + d], s);
      ebar_out[(long long)n * D + d] = s;
      const float kk = w4 * s;
      const float kd = kk > 0.f ? kk : 0.f;
      for (int i = 0; i < ni; ++i) {
        const float ai = c0 * e[i] + c1 * ep[i];
        const float v = ai * kd + (1.f - ai) * E[i * D + d];
        E[i * D + d] = v;
        Es[(eo + i) * D + d] = v;
      }
    }
    __syncthreads();
  }
}

__global__ void __launch_bounds__(RS_THREADS) reasoner_bwd_pipe_kernel(
    int D, int max_ni, const float* __restrict__ E0, const int* __restrict__ cu_ingr, const int* __restrict__ cu_steps,
    const int* __restrict__ clip_ent_off, const float* __restrict__ w2v, const float* __restrict__ clog,
    const float* __restrict__ w4s, const float* __restrict__ ent, const float* __restrict__ ebar,
    const float* __restrict__ Es, const float* __restrict__ a_saved, const float* __restrict__ d_ent,
    const float* __restrict__ d_ebar, const float* __restrict__ d_Es, float* __restrict__ dE0, float* __restrict__ dw2v,
    float* __restrict__ dclog, float* __restrict__ dw4s) {
  svpc_pdl_sync();
  extern __shared__ __align__(16) float sm[];
  const size_t ND = (size_t)max_ni * D;
  float* dE = sm;                        // (max_ni, D) gradient w.r.t. the current entity matrix
  float* Epb = dE + ND;                  // 2 x (max_ni, D): entity matrix BEFORE this step / before the next (earlier) one
  float* dEsb = Epb + 2 * ND;            // 2 x (max_ni, D): d_Es of this / the next step
  float* vecb = dEsb + 2 * ND;           // 2 x (3 D): ebar row, d_ebar row, w2v row
  float* deb = vecb + 6 * D;             // (D) d ebar
  float* kvs = deb + D;                  // (D) k
  float* da = kvs + D;                   // (max_ni)
  float* dan = da + max_ni;              // (max_ni)
  float* de_next = dan + max_ni;         // 2 x (max_ni): gradient flowing into e_t from step t+1, this / next
  float* svb = de_next + 2 * max_ni;     // 2 x (4 max_ni): a, e, e_prev, d_ent of this / the next step
  float* pw4 = svb + 8 * max_ni;         // (32) per-warp partial sums of d w4
  const int b = blockIdx.x;
  const int g0 = cu_ingr[b], ni = cu_ingr[b + 1] - g0;
  const int n0 = cu_steps[b], n1 = cu_steps[b + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  if (n1 <= n0) {
    for (int i = threadIdx.x; i < ni * D; i += RS_THREADS) dE0[(long long)g0 * D + i] = 0.f;
    return;
  }
  // entity-row offsets of the video's steps up front: the staging addresses of every step depend on them
  __shared__ int s_eo[RS_MAXSTEPS];
  for (int k = threadIdx.x; k < min(n1 - n0, RS_MAXSTEPS); k += RS_THREADS) s_eo[k] = clip_ent_off[n0 + k];
  __syncthreads();
  auto eo_of = [&](int n) -> long long { return (n - n0 < RS_MAXSTEPS) ? s_eo[n - n0] : clip_ent_off[n]; };
  auto stage_step = [&](int n, int buf) {
    const float* Eprev = (n == n0) ? E0 + (long long)g0 * D : Es + eo_of(n - 1) * D;
    stage_async(Epb + buf * ND, Eprev, ni * D);
    stage_async(dEsb + buf * ND, d_Es + eo_of(n) * D, ni * D);
    stage_async(vecb + buf * 3 * D, ebar + (long long)n * D, D);
    stage_async(vecb + buf * 3 * D + D, d_ebar + (long long)n * D, D);
    stage_async(vecb + buf * 3 * D + 2 * D, w2v + (long long)n * D, D);
  };
  // per-ingredient vectors of a step (a, e, e_prev, d_ent): thread i < ni loads them a step ahead into registers
  float r_a = 0.f, r_e = 0.f, r_ep = 0.f, r_de = 0.f;
  auto load_small = [&](int n) {
    if ((int)threadIdx.x < ni) {
      const long long eo = eo_of(n);
      r_a = a_saved[eo + threadIdx.x];
      r_e = ent[eo + threadIdx.x];
      r_ep = (n == n0) ? 0.f : ent[eo_of(n - 1) + threadIdx.x];
      r_de = d_ent[eo + threadIdx.x];
    }
  };
  auto store_small = [&](int buf) {
    if ((int)threadIdx.x < ni) {
      float* v = svb + buf * 4 * max_ni;
      v[threadIdx.x] = r_a; v[max_ni + threadIdx.x] = r_e; v[2 * max_ni + threadIdx.x] = r_ep; v[3 * max_ni + threadIdx.x] = r_de;
    }
  };
  stage_step(n1 - 1, 0);
  cp_async_commit();
  load_small(n1 - 1);
  // scalars of a step, carried one step ahead in registers
  float w4 = w4s[n1 - 1], l0 = clog[3 * (n1 - 1)], l1 = clog[3 * (n1 - 1) + 1], l2 = clog[3 * (n1 - 1) + 2];
  for (int i = threadIdx.x; i < ni * D; i += RS_THREADS) dE[i] = 0.f;
  for (int i = threadIdx.x; i < 2 * max_ni; i += RS_THREADS) de_next[i] = 0.f;
  store_small(0);
  cp_async_wait_all();
  __syncthreads();
  for (int n = n1 - 1; n >= n0; --n) {
    const int cur = (n1 - 1 - n) & 1;
    const float* Ep = Epb + cur * ND;
    const float* dEs = dEsb + cur * ND;
    const float* ebr = vecb + cur * 3 * D;
    const float* debo = ebr + D;
    const float* w = ebr + 2 * D;
    const float* av = svb + cur * 4 * max_ni;          // a, e, e_prev, d_ent of this step
    const float* ev = av + max_ni;
    const float* epv = ev + max_ni;
    const float* dev_ = epv + max_ni;
    float* den_in = de_next + cur * max_ni;            // written by the previous iteration (step n + 1)
    float* den_out = de_next + (cur ^ 1) * max_ni;
    float w4n = 0.f, l0n = 0.f, l1n = 0.f, l2n = 0.f;
    if (n > n0) {                                      // everything step n - 1 reads from global: in flight during this step
      stage_step(n - 1, cur ^ 1);
      load_small(n - 1);
      w4n = w4s[n - 1]; l0n = clog[3 * (n - 1)]; l1n = clog[3 * (n - 1) + 1]; l2n = clog[3 * (n - 1) + 2];
    }
    cp_async_commit();
    // (A) column owner: dE[:, d] += dEs[:, d]; k_d; dk_d = sum_i a_i dE[i][d]; d ebar_d
    float p4 = 0.f;
    for (int d = threadIdx.x; d < D; d += RS_THREADS) {
      const float ebd = ebr[d];
      const float kk = w4 * ebd;
      kvs[d] = kk > 0.f ? kk : 0.f;
      float s = 0.f;
      for (int i = 0; i < ni; ++i) {
        const float g = dE[i * D + d] + dEs[i * D + d];
        dE[i * D + d] = g;
        s = fmaf(av[i], g, s);
      }
      const float dpre = kk > 0.f ? s : 0.f;
      p4 = fmaf(dpre, ebd, p4);
      deb[d] = dpre * w4 + debo[d];
    }
    p4 = warp_sum(p4);
    if (lane == 0) pw4[warp] = p4;
    __syncthreads();
    // (B) one warp per entity row: da_i = dE_i . (k - Ep_i), dan_i = d ebar . Ep_i
    for (int i = warp; i < ni; i += RS_THREADS / 32) {
      float s = 0.f, t = 0.f;
      for (int d = lane; d < D; d += 32) {
        const float epd = Ep[i * D + d];
        s = fmaf(dE[i * D + d], kvs[d] - epd, s);
        t = fmaf(deb[d], epd, t);
      }
      s = warp_sum(s);
      t = warp_sum(t);
      if (lane == 0) { da[i] = s; dan[i] = t; }
    }
    __syncthreads();
    // (C) the small per-ingredient chain, recomputed by every thread: asum, tdot, then dz_i on the fly
    const float mx = fmaxf(l0, fmaxf(l1, l2));
    const float x0 = expf(l0 - mx), x1 = expf(l1 - mx), x2 = expf(l2 - mx);
    const float c0 = x0 / (x0 + x1 + x2), c1 = x1 / (x0 + x1 + x2), c2 = x2 / (x0 + x1 + x2);
    float asum = 0.f;
    for (int i = 0; i < ni; ++i) asum += av[i];
    const float inv_asum = 1.f / asum;
    float tdot = 0.f;
    for (int i = 0; i < ni; ++i) tdot = fmaf(dan[i], av[i] * inv_asum, tdot);
    if (threadIdx.x == 0) {
      float p0 = 0.f, p1 = 0.f, t4 = 0.f;
      for (int i = 0; i < ni; ++i) {
        const float dai = da[i] + (dan[i] - tdot) * inv_asum;
        p0 = fmaf(dai, ev[i], p0);
        p1 = fmaf(dai, epv[i], p1);
      }
      const float dot = c0 * p0 + c1 * p1;       // dc2 = 0
      dclog[3 * n + 0] = c0 * (p0 - dot);
      dclog[3 * n + 1] = c1 * (p1 - dot);
      dclog[3 * n + 2] = c2 * (0.f - dot);
      for (int k = 0; k < RS_THREADS / 32; ++k) t4 += pw4[k];
      dw4s[n] = t4;
    }
    if ((int)threadIdx.x < ni) {
      const int i = threadIdx.x;
      den_out[i] = c1 * (da[i] + (dan[i] - tdot) * inv_asum);   // read by step n - 1 (other buffer: no hazard with den_in)
    }
    // (D) column owner: dw2v_d = sum_i dz_i Ep[i][d]; dE_prev[:, d] = (1 - a_i) dE + an_i d ebar_d + dz_i w_d
    for (int d = threadIdx.x; d < D; d += RS_THREADS) {
      const float debd = deb[d], wd = w[d];
      float s = 0.f;
      for (int i = 0; i < ni; ++i) {
        const float ai = av[i], ei = ev[i];
        const float dai = da[i] + (dan[i] - tdot) * inv_asum;
        const float dei = c0 * dai + dev_[i] + den_in[i];
        const float dzi = dei * ei * (1.f - ei);
        s = fmaf(dzi, Ep[i * D + d], s);
        dE[i * D + d] = (1.f - ai) * dE[i * D + d] + (ai * inv_asum) * debd + dzi * wd;
      }
      dw2v[(long long)n * D + d] = s;
    }
    store_small(cur ^ 1);
    w4 = w4n; l0 = l0n; l1 = l1n; l2 = l2n;
    cp_async_wait_all();
    __syncthreads();
  }
  for (int i = threadIdx.x; i < ni * D; i += RS_THREADS) dE0[(long long)g0 * D + i] = dE[i];
}

// ------------------------------------------------------------------------------------------------
// Cluster variants: ONE 8-CTA thread-block cluster per video, the D columns of the entity matrix split over the cluster
// (96 columns per CTA at D = 768).  Everything a CTA reads from global memory - its column slice of E0, of the w2v / ebar /
// d_ebar rows and (backward) of E_{t-1} and dEs of EVERY step - is staged in shared memory before the first step, so the
// recurrence itself never waits for L2.  The only cross-column quantities are the per-ingredient dot products (e_i forward;
// da_i, dan_i backward): every CTA pushes its partial sums into the shared memory of all eight CTAs (st.shared::cluster,
// double-buffered by step parity) and one cluster barrier per step publishes them; each CTA then adds the eight partials
// in rank order, so all of them carry identical copies of the per-ingredient chain.  128 SMs work on 16 videos instead of 16.
// ------------------------------------------------------------------------------------------------
#define RC_CL 8
#define RC_THREADS 128
__device__ __forceinline__ uint32_t rc_rank() {
  uint32_t r;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
  return r;
}
__device__ __forceinline__ void rc_cluster_sync() {
  asm volatile("barrier.cluster.arrive.release.aligned;\n\tbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}
// store v into the same shared-memory location of CTA `rank` of this cluster
__device__ __forceinline__ void rc_push(float* local, uint32_t rank, float v) {
  uint32_t remote;
  asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"((uint32_t)__cvta_generic_to_shared(local)), "r"(rank));
  asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(v) : "memory");
}
// rows x Dc floats, row r at src + r * ld  ->  dst (rows, Dc) contiguous   (Dc % 4 == 0, 16-byte aligned)
__device__ __forceinline__ void rc_stage(float* dst, const float* src, long long ld, int rows, int Dc) {
  const int q = Dc >> 2;
  for (int i = threadIdx.x; i < rows * q; i += RC_THREADS) {
    const int r = i / q, c = (i - r * q) * 4;
    cp_async16(dst + r * Dc + c, src + (long long)r * ld + c);
  }
}

__global__ void __cluster_dims__(RC_CL, 1, 1) __launch_bounds__(RC_THREADS)
    reasoner_fwd_cluster_kernel(int D, int max_ni, int max_steps, const float* __restrict__ E0, const int* __restrict__ cu_ingr,
                                const int* __restrict__ cu_steps, const int* __restrict__ clip_ent_off,
                                const float* __restrict__ w2v, const float* __restrict__ clog, const float* __restrict__ w4s,
                                float* __restrict__ ent, float* __restrict__ ebar_out, float* __restrict__ Es,
                                float* __restrict__ a_saved) {
  svpc_pdl_sync();
  extern __shared__ __align__(16) float sm[];
  const int Dc = D / RC_CL;
  float* E = sm;                                   // (max_ni, Dc)
  float* wb = E + (size_t)max_ni * Dc;             // (max_steps, Dc)
  float* xb = wb + (size_t)max_steps * Dc;         // 2 x RC_CL x max_ni: partial dot products, [parity][source rank][i]
  float* eb = xb + 2 * RC_CL * max_ni;             // 2 x max_ni: e of this / the previous step (identical in every thread's view)
  const int b = blockIdx.x / RC_CL;
  const uint32_t rank = rc_rank();
  const int d0 = rank * Dc;
  const int g0 = cu_ingr[b], ni = cu_ingr[b + 1] - g0;
  const int n0 = cu_steps[b], n1 = cu_steps[b + 1];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  rc_stage(E, E0 + (long long)g0 * D + d0, D, ni, Dc);
  rc_stage(wb, w2v + (long long)n0 * D + d0, D, n1 - n0, Dc);
  cp_async_commit();
  for (int i = threadIdx.x; i < 2 * max_ni; i += RC_THREADS) eb[i] = 0.f;
  cp_async_wait_all();
  rc_cluster_sync();                               // every CTA of the cluster is running (and this CTA's tiles are visible)
  for (int n = n0; n < n1; ++n) {
    const int cur = (n - n0) & 1;
    const float* w = wb + (size_t)(n - n0) * Dc;
    float* xcur = xb + cur * RC_CL * max_ni;
    float* e = eb + cur * max_ni;
    const float* ep = eb + (cur ^ 1) * max_ni;
    const float l0 = clog[3 * n], l1 = clog[3 * n + 1], l2 = clog[3 * n + 2];
    const float w4 = w4s[n];
    const long long eo = clip_ent_off[n];
    // (1) partial e_i over this CTA's columns, pushed to every CTA of the cluster
    for (int i = warp; i < ni; i += RC_THREADS / 32) {
      float s = 0.f;
      for (int d = lane; d < Dc; d += 32) s = fmaf(E[i * Dc + d], w[d], s);
      s = warp_sum(s);
      if (lane < RC_CL) rc_push(xcur + rank * max_ni + i, lane, s);
    }
    rc_cluster_sync();
    // (2) the per-ingredient chain, recomputed by every thread (the eight partials are added in rank order everywhere)
    const float mx = fmaxf(l0, fmaxf(l1, l2));
    const float x0 = expf(l0 - mx), x1 = expf(l1 - mx), x2 = expf(l2 - mx);
    const float c0 = x0 / (x0 + x1 + x2), c1 = x1 / (x0 + x1 + x2);
    float asum = 0.f;
    for (int i = 0; i < ni; ++i) {
      float s = 0.f;
#pragma unroll
      for (int r = 0; r < RC_CL; ++r) s += xcur[r * max_ni + i];
      const float ei = svpc_sigmoid(s);
      e[i] = ei;                                   // (every thread writes the same value; each reads back its own write)
      asum += c0 * ei + c1 * ep[i];
    }
    if (rank == 0 && (int)threadIdx.x < ni) {
      const int i = threadIdx.x;
      ent[eo + i] = e[i];
      a_saved[eo + i] = c0 * e[i] + c1 * ep[i];
    }
    // (3) column owner: ebar_d, k_d and the update of E[:, d]
    for (int d = threadIdx.x; d < Dc; d += RC_THREADS) {
      float s = 0.f;
      for (int i = 0; i < ni; ++i) s = fmaf((c0 * e[i] + c1 * ep[i]) / asum, E[i * Dc + d], s);
      ebar_out[(long long)n * D + d0 + d] = s;
      const float kk = w4 * s;
      const float kd = kk > 0.f ? kk : 0.f;
      for (int i = 0; i < ni; ++i) {
        const float ai = c0 * e[i] + c1 * ep[i];
        const float v = ai * kd + (1.f - ai) * E[i * Dc + d];
        E[i * Dc + d] = v;
        Es[(eo + i) * D + d0 + d] = v;
      }
    }
    __syncthreads();                               // E[:, :] of this CTA is updated before the next step's row products
  }
  rc_cluster_sync();                               // nobody leaves while a neighbour may still push into its shared memory
}

__global__ void __cluster_dims__(RC_CL, 1, 1) __launch_bounds__(RC_THREADS)
    reasoner_bwd_cluster_kernel(int D, int max_ni, int max_steps, const float* __restrict__ E0, const int* __restrict__ cu_ingr,
                                const int* __restrict__ cu_steps, const int* __restrict__ clip_ent_off,
                                const float* __restrict__ w2v, const float* __restrict__ clog, const float* __restrict__ w4s,
                                const float* __restrict__ ent, const float* __restrict__ ebar, const float* __restrict__ Es,
                                const float* __restrict__ a_saved, const float* __restrict__ d_ent,
                                const float* __restrict__ d_ebar, const float* __restrict__ d_Es, float* __restrict__ dE0,
                                float* __restrict__ dw2v, float* __restrict__ dclog, float* __restrict__ dw4s) {
  svpc_pdl_sync();
  extern __shared__ __align__(16) float sm[];
  const int Dc = D / RC_CL;
  const size_t NDc = (size_t)max_ni * Dc;
  const int XW = 2 * max_ni + 1;                   // da_i, dan_i partials and the d w4 partial of one CTA
  float* dE = sm;                                  // (max_ni, Dc)
  float* Eps = dE + NDc;                           // (max_steps, max_ni, Dc): E_{t-1} slices of all steps
  float* dEss = Eps + (size_t)max_steps * NDc;     // (max_steps, max_ni, Dc): dEs slices
  float* vec = dEss + (size_t)max_steps * NDc;     // (max_steps, 3, Dc): ebar, d_ebar, w2v slices
  float* deb = vec + (size_t)max_steps * 3 * Dc;   // (Dc)
  float* kvs = deb + Dc;                           // (Dc)
  float* sv = kvs + Dc;                            // (max_steps, 4, max_ni): a, e, e_prev, d_ent
  float* sc = sv + (size_t)max_steps * 4 * max_ni; // (max_steps, 4): w4, l0, l1, l2
  float* xb = sc + (size_t)max_steps * 4;          // 2 x RC_CL x XW
  float* dab = xb + 2 * RC_CL * XW;                // (2 max_ni): da_i | dan_i summed over the cluster
  float* den = dab + 2 * max_ni;                   // 2 x max_ni: gradient into e_t from step t+1, this / next
  float* pw4 = den + 2 * max_ni;                   // (RC_THREADS / 32)
  const int b = blockIdx.x / RC_CL;
  const uint32_t rank = rc_rank();
  const int d0 = rank * Dc;
  const int g0 = cu_ingr[b], ni = cu_ingr[b + 1] - g0;
  const int n0 = cu_steps[b], n1 = cu_steps[b + 1];
  const int S = n1 - n0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int t = 0; t < S; ++t) {
    const int n = n0 + t;
    const float* Eprev = (t == 0) ? E0 + (long long)g0 * D : Es + (long long)clip_ent_off[n - 1] * D;
    rc_stage(Eps + t * NDc, Eprev + d0, D, ni, Dc);
    rc_stage(dEss + t * NDc, d_Es + (long long)clip_ent_off[n] * D + d0, D, ni, Dc);
    rc_stage(vec + (size_t)t * 3 * Dc, ebar + (long long)n * D + d0, D, 1, Dc);
    rc_stage(vec + (size_t)t * 3 * Dc + Dc, d_ebar + (long long)n * D + d0, D, 1, Dc);
    rc_stage(vec + (size_t)t * 3 * Dc + 2 * Dc, w2v + (long long)n * D + d0, D, 1, Dc);
  }
  cp_async_commit();
  for (int i = threadIdx.x; i < S * ni; i += RC_THREADS) {
    const int t = i / ni, k = i - t * ni, n = n0 + t;
    const long long eo = clip_ent_off[n];
    float* v = sv + (size_t)t * 4 * max_ni;
    v[k] = a_saved[eo + k];
    v[max_ni + k] = ent[eo + k];
    v[2 * max_ni + k] = (t == 0) ? 0.f : ent[clip_ent_off[n - 1] + k];
    v[3 * max_ni + k] = d_ent[eo + k];
  }
  for (int t = threadIdx.x; t < S; t += RC_THREADS) {
    const int n = n0 + t;
    sc[t * 4] = w4s[n]; sc[t * 4 + 1] = clog[3 * n]; sc[t * 4 + 2] = clog[3 * n + 1]; sc[t * 4 + 3] = clog[3 * n + 2];
  }
  for (int i = threadIdx.x; i < ni * Dc; i += RC_THREADS) dE[i] = 0.f;
  for (int i = threadIdx.x; i < 2 * max_ni; i += RC_THREADS) den[i] = 0.f;
  cp_async_wait_all();
  rc_cluster_sync();
  for (int t = S - 1; t >= 0; --t) {
    const int n = n0 + t;
    const int cur = (S - 1 - t) & 1;
    const float* Ep = Eps + t * NDc;
    const float* dEs = dEss + t * NDc;
    const float* ebr = vec + (size_t)t * 3 * Dc;
    const float* debo = ebr + Dc;
    const float* w = ebr + 2 * Dc;
    const float* av = sv + (size_t)t * 4 * max_ni;
    const float* ev = av + max_ni;
    const float* epv = ev + max_ni;
    const float* dev_ = epv + max_ni;
    const float w4 = sc[t * 4], l0 = sc[t * 4 + 1], l1 = sc[t * 4 + 2], l2 = sc[t * 4 + 3];
    float* xcur = xb + cur * RC_CL * XW;
    float* den_in = den + cur * max_ni;
    float* den_out = den + (cur ^ 1) * max_ni;
    // (A) column owner: dE[:, d] += dEs[:, d]; k_d; dk_d = sum_i a_i dE[i][d]; d ebar_d
    float p4 = 0.f;
    for (int d = threadIdx.x; d < Dc; d += RC_THREADS) {
      const float ebd = ebr[d];
      const float kk = w4 * ebd;
      kvs[d] = kk > 0.f ? kk : 0.f;
      float s = 0.f;
      for (int i = 0; i < ni; ++i) {
        const float g = dE[i * Dc + d] + dEs[i * Dc + d];
        dE[i * Dc + d] = g;
        s = fmaf(av[i], g, s);
      }
      const float dpre = kk > 0.f ? s : 0.f;
      p4 = fmaf(dpre, ebd, p4);
      deb[d] = dpre * w4 + debo[d];
    }
    p4 = warp_sum(p4);
    if (lane == 0) pw4[warp] = p4;
    __syncthreads();
    // (B) one warp per entity row: partial da_i, dan_i over this CTA's columns, pushed to the whole cluster
    for (int i = warp; i < ni; i += RC_THREADS / 32) {
      float s = 0.f, u = 0.f;
      for (int d = lane; d < Dc; d += 32) {
        const float epd = Ep[i * Dc + d];
        s = fmaf(dE[i * Dc + d], kvs[d] - epd, s);
        u = fmaf(deb[d], epd, u);
      }
      s = warp_sum(s);
      u = warp_sum(u);
      if (lane < RC_CL) { rc_push(xcur + rank * XW + i, lane, s); rc_push(xcur + rank * XW + max_ni + i, lane, u); }
    }
    if (warp == 0 && lane < RC_CL) {
      float tot = 0.f;
#pragma unroll
      for (int k = 0; k < RC_THREADS / 32; ++k) tot += pw4[k];
      rc_push(xcur + rank * XW + 2 * max_ni, lane, tot);
    }
    rc_cluster_sync();
    // (C) the per-ingredient chain, recomputed by every thread from the eight partials (rank order)
    const float mx = fmaxf(l0, fmaxf(l1, l2));
    const float x0 = expf(l0 - mx), x1 = expf(l1 - mx), x2 = expf(l2 - mx);
    const float c0 = x0 / (x0 + x1 + x2), c1 = x1 / (x0 + x1 + x2), c2 = x2 / (x0 + x1 + x2);
    float asum = 0.f, tdot_n = 0.f;
    for (int i = 0; i < ni; ++i) asum += av[i];
    for (int i = 0; i < ni; ++i) {
      float s = 0.f, u = 0.f;
#pragma unroll
      for (int r = 0; r < RC_CL; ++r) { s += xcur[r * XW + i]; u += xcur[r * XW + max_ni + i]; }
      dab[i] = s;                                  // (identical in every thread; each reads back its own write)
      dab[max_ni + i] = u;
      tdot_n = fmaf(u, av[i] / asum, tdot_n);
    }
    const float tdot = tdot_n;
    if (rank == 0 && threadIdx.x == 0) {
      float p0 = 0.f, p1 = 0.f, t4 = 0.f;
      for (int i = 0; i < ni; ++i) {
        const float dai = dab[i] + (dab[max_ni + i] - tdot) / asum;
        p0 = fmaf(dai, ev[i], p0);
        p1 = fmaf(dai, epv[i], p1);
      }
      const float dot = c0 * p0 + c1 * p1;         // dc2 = 0
      dclog[3 * n + 0] = c0 * (p0 - dot);
      dclog[3 * n + 1] = c1 * (p1 - dot);
      dclog[3 * n + 2] = c2 * (0.f - dot);
#pragma unroll
      for (int r = 0; r < RC_CL; ++r) t4 += xcur[r * XW + 2 * max_ni];
      dw4s[n] = t4;
    }
    for (int i = 0; i < ni; ++i) den_out[i] = c1 * (dab[i] + (dab[max_ni + i] - tdot) / asum);   // same value from every thread
    // (D) column owner: dw2v_d = sum_i dz_i Ep[i][d]; dE_prev[:, d] = (1 - a_i) dE + an_i d ebar_d + dz_i w_d
    for (int d = threadIdx.x; d < Dc; d += RC_THREADS) {
      const float debd = deb[d], wd = w[d];
      float s = 0.f;
      for (int i = 0; i < ni; ++i) {
        const float ai = av[i], ei = ev[i];
        const float dai = dab[i] + (dab[max_ni + i] - tdot) / asum;
        const float dei = c0 * dai + dev_[i] + den_in[i];
        const float dzi = dei * ei * (1.f - ei);
        s = fmaf(dzi, Ep[i * Dc + d], s);
        dE[i * Dc + d] = (1.f - ai) * dE[i * Dc + d] + (ai / asum) * debd + dzi * wd;
      }
      dw2v[(long long)n * D + d0 + d] = s;
    }
    __syncthreads();                               // den_in of this step may be overwritten (as den_out) two steps on; dab reuse
  }
  for (int i = threadIdx.x; i < ni * Dc; i += RC_THREADS) {
    const int r = i / Dc, c = i - r * Dc;
    dE0[(long long)(g0 + r) * D + d0 + c] = dE[i];
  }
  rc_cluster_sync();
}

// SVPC_REASONER_PIPE: 1 (default) single-CTA pipelined kernels, 2 the 8-CTA cluster kernels, 0 the un-pipelined ones; each
// level is also the fallback of the one above when its shared-memory staging does not fit.  Measured at the headline shape
// (tools/reasoner_bench.py, 50 launches in a graph): forward 46 / 43 / 53 us, backward 92 / 90 / 81 us for 0 / 1 / 2 - the
// recurrence is bound by its two barriers and the per-ingredient chain of every step, not by memory, and inside the training
// step the cluster kernels (128 SMs busy instead of 16) take SMs from the weight-gradient stream: 4.99 against 4.95 ms/step.
static int reasoner_mode() {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("SVPC_REASONER_PIPE");
    on = e ? atoi(e) : 1;
  }
  return on;
}
static bool reasoner_pipelined() { return reasoner_mode() >= 1; }

extern "C" int svpc_reasoner_fwd(int B, int D, int max_ni, int max_steps_hint, const float* E0, const int* cu_ingr, const int* cu_steps,
                                 const int* clip_ent_off, const float* w2v, const float* clog, const float* w4s,
                                 float* ent, float* ebar, float* Es, float* a_saved, svpc_stream_t stream) {
  if (B <= 0) return SVPC_OK;
  const int max_steps = max_steps_hint > 0 ? max_steps_hint : 0;
  {
    const int Dc = D / RC_CL;
    const size_t smem_c = sizeof(float) * ((size_t)max_ni * Dc + (size_t)max_steps * Dc + 2 * RC_CL * (size_t)max_ni + 2 * (size_t)max_ni);
    if (reasoner_mode() >= 2 && D % (RC_CL * 4) == 0 && max_steps > 0 && max_ni <= RC_THREADS && smem_c <= 200 * 1024) {
      cudaFuncSetAttribute(reasoner_fwd_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_c);
      svpc_launch(reasoner_fwd_cluster_kernel, B * RC_CL, RC_THREADS, smem_c, (cudaStream_t)stream, D, max_ni, max_steps, E0, cu_ingr,
                  cu_steps, clip_ent_off, w2v, clog, w4s, ent, ebar, Es, a_saved);
      return svpc_check_launch("reasoner_fwd");
    }
  }
  const size_t smem_p = sizeof(float) * ((size_t)max_ni * D + (size_t)max_steps * D + 2 * (size_t)max_ni);
  if (reasoner_pipelined() && D % 4 == 0 && max_steps > 0 && smem_p <= 220 * 1024) {
    cudaFuncSetAttribute(reasoner_fwd_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p);
    svpc_launch(reasoner_fwd_pipe_kernel, B, RS_THREADS, smem_p, (cudaStream_t)stream, D, max_ni, max_steps, E0, cu_ingr, cu_steps,
                clip_ent_off, w2v, clog, w4s, ent, ebar, Es, a_saved);
    return svpc_check_launch("reasoner_fwd");
  }
  const size_t smem = sizeof(float) * ((size_t)max_ni * D + D + 4 * (size_t)max_ni);
  SVPC_REQUIRE(smem <= 220 * 1024, "reasoner_fwd: %d ingredients x %d do not fit in shared memory", max_ni, D);
  cudaFuncSetAttribute(reasoner_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  svpc_launch(reasoner_fwd_kernel, B, RS_THREADS, smem, (cudaStream_t)stream, D, max_ni, E0, cu_ingr, cu_steps, clip_ent_off, w2v, clog,
                                                                    w4s, ent, ebar, Es, a_saved);
  return svpc_check_launch("reasoner_fwd");
}

extern "C" int svpc_reasoner_bwd(int B, int D, int max_ni, int max_steps_hint, const float* E0, const int* cu_ingr, const int* cu_steps,
                                 const int* clip_ent_off, const float* w2v, const float* clog, const float* w4s,
                                 const float* ent, const float* ebar, const float* Es, const float* a_saved,
                                 const float* d_ent, const float* d_ebar, const float* d_Es, float* dE0, float* dw2v,
                                 float* dclog, float* dw4s, svpc_stream_t stream) {
  if (B <= 0) return SVPC_OK;
  {
    const int Dc = D / RC_CL;
    const int ms = max_steps_hint;
    const size_t smem_c = sizeof(float) * ((size_t)max_ni * Dc * (1 + 2 * (size_t)ms) + (size_t)ms * 3 * Dc + 2 * (size_t)Dc +
                                           (size_t)ms * 4 * max_ni + (size_t)ms * 4 + 2 * RC_CL * (2 * (size_t)max_ni + 1) +
                                           4 * (size_t)max_ni + RC_THREADS / 32 + 8);
    if (reasoner_mode() >= 2 && D % (RC_CL * 4) == 0 && ms > 0 && max_ni <= RC_THREADS && smem_c <= 200 * 1024) {
      cudaFuncSetAttribute(reasoner_bwd_cluster_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_c);
      svpc_launch(reasoner_bwd_cluster_kernel, B * RC_CL, RC_THREADS, smem_c, (cudaStream_t)stream, D, max_ni, ms, E0, cu_ingr, cu_steps,
                  clip_ent_off, w2v, clog, w4s, ent, ebar, Es, a_saved, d_ent, d_ebar, d_Es, dE0, dw2v, dclog, dw4s);
      return svpc_check_launch("reasoner_bwd");
    }
  }
  const size_t smem_p = sizeof(float) * (5 * (size_t)max_ni * D + 8 * (size_t)D + 12 * (size_t)max_ni + 32);
  if (reasoner_pipelined() && D % 4 == 0 && smem_p <= 220 * 1024) {
    cudaFuncSetAttribute(reasoner_bwd_pipe_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_p);
    svpc_launch(reasoner_bwd_pipe_kernel, B, RS_THREADS, smem_p, (cudaStream_t)stream, D, max_ni, E0, cu_ingr, cu_steps,
                clip_ent_off, w2v, clog, w4s, ent, ebar, Es, a_saved, d_ent, d_ebar, d_Es, dE0, dw2v, dclog, dw4s);
    return svpc_check_launch("reasoner_bwd");
  }
  const size_t smem = sizeof(float) * (2 * (size_t)max_ni * D + 3 * (size_t)D + 5 * (size_t)max_ni);
  SVPC_REQUIRE(smem <= 220 * 1024, "reasoner_bwd: %d ingredients x %d do not fit in shared memory", max_ni, D);
  cudaFuncSetAttribute(reasoner_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
  svpc_launch(reasoner_bwd_kernel, B, RS_THREADS, smem, (cudaStream_t)stream, D, max_ni, E0, cu_ingr, cu_steps, clip_ent_off, w2v, clog,
                                                                    w4s, ent, ebar, Es, a_saved, d_ent, d_ebar, d_Es, dE0,
                                                                    dw2v, dclog, dw4s);
  return svpc_check_launch("reasoner_bwd");
}
