// Output head of the caption decoder and the front end of the textual re-simulator:
//   * pointer-generator mixing + scatter of the copy mass into the (extended) vocabulary + label-smoothed KL loss on
//     PROBABILITIES (reference model.py:896-923, 37-55), one CTA per (clip, text position); the per-ingredient
//     gradients are reduced over a clip's positions by a second, per-clip kernel;
//   * plain softmax + the same loss for the 'video' mode (model.py:1011-1014);
//   * hard Gumbel-softmax with injected noise -> word-vector lookup -> masked mean (model.py:1018-1021).
// Quirks kept: softmax over the INGREDIENT axis; copy mass of ingredient i split equally over its token ids, added in
// dictionary order by one thread (same accumulation order as the reference's Python loop); log(p + 1e-12); smoothing
// mass s/(Vx-1) with the LAST column zeroed (one_hot[ignore_index=-1] = 0) unless it is the target.
#include "common.cuh"
#include <stdlib.h>

#define PG_THREADS 256
// The per-(clip, position) pointer-generator kernels are chains of short dependent phases (a dozen barriers, L2 round
// trips): their throughput is rows in flight per SM / chain latency, so they run with 128-thread CTAs (16 rows per SM).
// With few rows (greedy decoding: one position per launch) the chain latency of a single row is what counts: 256 threads.
static inline int ptr_threads(int Lt, int n_clips) { return (long long)Lt * n_clips >= 8 * SVPC_NUM_SMS ? 128 : PG_THREADS; }

// multi-row kernels (training shapes) unless SVPC_POINTER_ROWS=0
static inline bool ptr_rows_route(int Lt, int n_clips, int V) {
  static int on = -1;
  if (on < 0) { const char* e = getenv("SVPC_POINTER_ROWS"); on = e ? atoi(e) : 1; }
  return on && Lt > 1 && V <= 1024 && (long long)Lt * n_clips >= 8 * SVPC_NUM_SMS;
}

__device__ __forceinline__ float ls_q(int v, int label, int vx, float s, float conf) {
  if (v == label) return conf;
  if (v == vx - 1) return 0.f;
  return s;
}
__device__ __forceinline__ float ls_const(int label, int vx, float s, float conf) {
  const int n_s = vx - 1 - (label != vx - 1 ? 1 : 0);
  float c = conf > 0.f ? conf * logf(conf) : 0.f;
  if (s > 0.f) c += (float)n_s * s * logf(s);
  return c;
}

// ------------------------------------------------------------------------------------------------
template <typename T>
__global__ void __launch_bounds__(PG_THREADS) pointer_fwd_kernel(
    int Lt, int D, int V, int max_ni, int max_vx, const T* __restrict__ dec, const float* __restrict__ G,
    const float* __restrict__ Esrc, const int* __restrict__ clip_src_off, const int* __restrict__ clip_ni,
    const int* __restrict__ clip_gi, const int* __restrict__ clip_vx, const int* __restrict__ clip_P_off,
    const int* __restrict__ tok_ptr, const int* __restrict__ tok_ids, const float* __restrict__ wpgen,
    const float* __restrict__ bpgen, const float* __restrict__ logits, long long ldl, const int* __restrict__ labels,
    float smoothing, float* __restrict__ P, float* __restrict__ loss_rows, float* __restrict__ alpha_out,
    float* __restrict__ pgen_out, float* __restrict__ ctx_out) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  float* sdec = sm;               // D
  float* sctx = sdec + D;         // D
  float* prow = sctx + D;         // max_vx
  float* alpha = prow + max_vx;   // max_ni
  __shared__ float red[64];
  const int n = blockIdx.y, l = blockIdx.x;     // one CTA per (clip, text position): rows are independent
  const int src = clip_src_off[n], ni = clip_ni[n], gi = clip_gi[n], vx = clip_vx[n];
  const long long poff = clip_P_off[n];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float bp = bpgen[0];
  {
    const long long r = (long long)n * Lt + l;
    for (int d = threadIdx.x; d < D; d += (int)blockDim.x) sdec[d] = ldf(dec + r * D + d);
    __syncthreads();
    for (int i = warp; i < ni; i += ((int)blockDim.x >> 5)) {
      const float* g = G + (long long)(src + i) * D;
      float s = 0.f;
      for (int d = lane; d < D; d += 32) s = fmaf(g[d], sdec[d], s);
      s = warp_sum(s);
      if (lane == 0) alpha[i] = s;
    }
    __syncthreads();
    if (warp == 0) {   // softmax over the ingredient axis (model.py:900)
      float mx = -INFINITY;
      for (int i = lane; i < ni; i += 32) mx = fmaxf(mx, alpha[i]);
      mx = warp_max(mx);
      float sum = 0.f;
      for (int i = lane; i < ni; i += 32) { const float e = expf(alpha[i] - mx); alpha[i] = e; sum += e; }
      sum = warp_sum(sum);
      for (int i = lane; i < ni; i += 32) {
        const float a = alpha[i] / sum;
        alpha[i] = a;
        alpha_out[r * max_ni + i] = a;
      }
    }
    __syncthreads();
    float pu = 0.f;
    for (int d = threadIdx.x; d < D; d += (int)blockDim.x) {
      float c = 0.f;
      for (int i = 0; i < ni; ++i) c = fmaf(alpha[i], Esrc[(long long)(src + i) * D + d], c);
      sctx[d] = c;
      ctx_out[r * D + d] = c;
      pu += sdec[d] * wpgen[d] + c * wpgen[D + d];
    }
    const float pg = svpc_sigmoid(block_sum(pu, red) + bp);
    if (threadIdx.x == 0) pgen_out[r] = pg;
    const float* lrow = logits + r * ldl;
    float mx = -INFINITY;
    for (int v = threadIdx.x; v < V; v += (int)blockDim.x) mx = fmaxf(mx, lrow[v]);
    mx = block_max(mx, red);
    float se = 0.f;
    for (int v = threadIdx.x; v < V; v += (int)blockDim.x) { const float e = expf(lrow[v] - mx); prow[v] = e; se += e; }
    se = block_sum(se, red);
    for (int v = threadIdx.x; v < vx; v += (int)blockDim.x) prow[v] = v < V ? pg * (prow[v] / se) : 0.f;
    __syncthreads();
    if (threadIdx.x == 0) {
      for (int i = 0; i < ni; ++i) {
        const int lo = tok_ptr[gi + i], hi = tok_ptr[gi + i + 1];
        const float add = alpha[i] * (1.f - pg) / (float)(hi - lo);
        for (int t = lo; t < hi; ++t) prow[tok_ids[t]] += add;
      }
    }
    __syncthreads();
    const int label = labels[r];
    const float s = smoothing / (float)(vx - 1), conf = 1.f - smoothing;
    float acc = 0.f;
    for (int v = threadIdx.x; v < vx; v += (int)blockDim.x) {
      const float p = prow[v];
      P[poff + (long long)l * vx + v] = p;
      if (label >= 0) acc += ls_q(v, label, vx, s, conf) * logf(p + 1e-12f);
    }
    acc = block_sum(acc, red);
    if (threadIdx.x == 0) loss_rows[r] = label >= 0 ? ls_const(label, vx, s, conf) - acc : 0.f;
  }
}

// Training shapes (thousands of rows): one CTA per (clip, group of up to PR_ROWS text positions).  The per-row kernel above
// is a chain of dependent L2 loads (every row re-reads the clip's ni rows of G and of Esrc: 60 KB per row, 295 MB per launch
// at the benchmark shape, all of it latency: ncu shows half of the warp samples on stall_long_sb at the FFMAs that consume
// them).  Here the CTA reads G / Esrc once for its rows (loads batched eight deep), and the vocabulary phase runs one WARP
// per row (warp shuffles only, no block barrier).
#define PR_ROWS 8
#define PR_THREADS 256
#define PR_NV 32            // logits per lane held in registers by the backward's one-warp-per-row vocabulary phase (V <= 1024)
template <typename T>
__global__ void __launch_bounds__(PR_THREADS) pointer_fwd_rows_kernel(
    int Lt, int R, int D, int V, int max_ni, int max_vx, const T* __restrict__ dec, const float* __restrict__ G,
    const float* __restrict__ Esrc, const int* __restrict__ clip_src_off, const int* __restrict__ clip_ni,
    const int* __restrict__ clip_gi, const int* __restrict__ clip_vx, const int* __restrict__ clip_P_off,
    const int* __restrict__ tok_ptr, const int* __restrict__ tok_ids, const float* __restrict__ wpgen,
    const float* __restrict__ bpgen, const float* __restrict__ logits, long long ldl, const int* __restrict__ labels,
    float smoothing, float* __restrict__ P, float* __restrict__ loss_rows, float* __restrict__ alpha_out,
    float* __restrict__ pgen_out, float* __restrict__ ctx_out) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  float* sdec = sm;                                   // PR_ROWS x D   (rows >= nr are zero)
  float* prow = sdec + (size_t)PR_ROWS * D;           // PR_ROWS x max_vx
  float* alpha = prow + (size_t)PR_ROWS * max_vx;     // PR_ROWS x max_ni
  __shared__ float spu[PR_ROWS][PR_THREADS / 32];
  __shared__ float spg[PR_ROWS];
  const int n = blockIdx.y, l0 = blockIdx.x * R;
  const int nr = min(R, Lt - l0);
  const int src = clip_src_off[n], ni = clip_ni[n], gi = clip_gi[n], vx = clip_vx[n];
  const long long poff = clip_P_off[n];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long r0 = (long long)n * Lt + l0;
  for (int e = threadIdx.x; e < PR_ROWS * D; e += PR_THREADS) sdec[e] = e < nr * D ? ldf(dec + r0 * D + e) : 0.f;
  __syncthreads();
  // scores[r][i] = dec_r . G_i : one warp per ingredient, the G row read once for all rows
  for (int i = warp; i < ni; i += PR_THREADS / 32) {
    const float* g = G + (long long)(src + i) * D;
    float acc[PR_ROWS];
#pragma unroll
    for (int r = 0; r < PR_ROWS; ++r) acc[r] = 0.f;
    for (int d0 = lane; d0 < D; d0 += 256) {
      float gv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) gv[u] = d0 + 32 * u < D ? g[d0 + 32 * u] : 0.f;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int dd = min(d0 + 32 * u, D - 1);
#pragma unroll
        for (int r = 0; r < PR_ROWS; ++r) acc[r] = fmaf(gv[u], sdec[r * D + dd], acc[r]);
      }
    }
#pragma unroll
    for (int r = 0; r < PR_ROWS; ++r) {
      const float s = warp_sum(acc[r]);
      if (lane == 0) alpha[r * max_ni + i] = s;
    }
  }
  __syncthreads();
  if (warp < nr) {   // softmax over the ingredient axis (model.py:900), one warp per row
    float* a = alpha + warp * max_ni;
    float mx = -INFINITY;
    for (int i = lane; i < ni; i += 32) mx = fmaxf(mx, a[i]);
    mx = warp_max(mx);
    float sum = 0.f;
    for (int i = lane; i < ni; i += 32) { const float e = expf(a[i] - mx); a[i] = e; sum += e; }
    sum = warp_sum(sum);
    for (int i = lane; i < ni; i += 32) {
      const float v = a[i] / sum;
      a[i] = v;
      alpha_out[(r0 + warp) * max_ni + i] = v;
    }
  } else {
    for (int i = lane; i < ni; i += 32) alpha[warp * max_ni + i] = 0.f;
  }
  __syncthreads();
  // context rows and the p_gen pre-activation: each Esrc element read once per CTA, four rows of Esrc in flight
  float pu[PR_ROWS];
#pragma unroll
  for (int r = 0; r < PR_ROWS; ++r) pu[r] = 0.f;
  for (int d = threadIdx.x; d < D; d += PR_THREADS) {
    float c[PR_ROWS];
#pragma unroll
    for (int r = 0; r < PR_ROWS; ++r) c[r] = 0.f;
    for (int i0 = 0; i0 < ni; i0 += 4) {
      float ev[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) ev[u] = i0 + u < ni ? Esrc[(long long)(src + i0 + u) * D + d] : 0.f;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int ii = min(i0 + u, ni - 1);
#pragma unroll
        for (int r = 0; r < PR_ROWS; ++r) c[r] = fmaf(alpha[r * max_ni + ii], ev[u], c[r]);
      }
    }
    const float w1 = wpgen[d], w2 = wpgen[D + d];
#pragma unroll
    for (int r = 0; r < PR_ROWS; ++r) {
      if (r < nr) ctx_out[(r0 + r) * D + d] = c[r];
      pu[r] += sdec[r * D + d] * w1 + c[r] * w2;
    }
  }
#pragma unroll
  for (int r = 0; r < PR_ROWS; ++r) {
    const float s = warp_sum(pu[r]);
    if (lane == 0) spu[r][warp] = s;
  }
  __syncthreads();
  if (threadIdx.x < nr) {
    float s = 0.f;
#pragma unroll
    for (int w = 0; w < PR_THREADS / 32; ++w) s += spu[threadIdx.x][w];
    const float pg = svpc_sigmoid(s + bpgen[0]);
    spg[threadIdx.x] = pg;
    pgen_out[r0 + threadIdx.x] = pg;
  }
  __syncthreads();
  if (warp >= nr) return;
  // vocabulary phase: one warp per row
  const long long r = r0 + warp;
  const int l = l0 + warp;
  const float pg = spg[warp];
  const float* a = alpha + warp * max_ni;
  float* pr = prow + (size_t)warp * max_vx;
  const float* lrow = logits + r * ldl;
  // stage the row's logits in shared memory eight loads deep, then max / exp / normalise out of shared memory
  float mx = -INFINITY;
  for (int v0 = lane; v0 < V; v0 += 256) {
    float t[8];
#pragma unroll
    for (int u = 0; u < 8; ++u) t[u] = v0 + 32 * u < V ? lrow[v0 + 32 * u] : -INFINITY;
#pragma unroll
    for (int u = 0; u < 8; ++u) {
      if (v0 + 32 * u < V) pr[v0 + 32 * u] = t[u];
      mx = fmaxf(mx, t[u]);
    }
  }
  mx = warp_max(mx);
  float se = 0.f;
#pragma unroll 4
  for (int v = lane; v < V; v += 32) { const float e = expf(pr[v] - mx); pr[v] = e; se += e; }
  se = warp_sum(se);
#pragma unroll 4
  for (int v = lane; v < V; v += 32) pr[v] = pg * (pr[v] / se);
  for (int v = V + lane; v < vx; v += 32) pr[v] = 0.f;
  __syncwarp();
  if (lane == 0) {     // copy mass in dictionary order (the reference's accumulation order)
    for (int i = 0; i < ni; ++i) {
      const int lo = tok_ptr[gi + i], hi = tok_ptr[gi + i + 1];
      const float add = a[i] * (1.f - pg) / (float)(hi - lo);
      for (int t = lo; t < hi; ++t) pr[tok_ids[t]] += add;
    }
  }
  __syncwarp();
  const int label = labels[r];
  const float s = smoothing / (float)(vx - 1), conf = 1.f - smoothing;
  float acc = 0.f;
#pragma unroll 4
  for (int v = lane; v < vx; v += 32) {
    const float p = pr[v];
    P[poff + (long long)l * vx + v] = p;
    if (label >= 0) acc += ls_q(v, label, vx, s, conf) * logf(p + 1e-12f);
  }
  acc = warp_sum(acc);
  if (lane == 0) loss_rows[r] = label >= 0 ? ls_const(label, vx, s, conf) - acc : 0.f;
}

// Backward, part 1: one CTA per (clip, text position).  Writes d logits, d dec, du (gradient of the p_gen
// pre-activation) and the score gradients dsc[r][i]; the per-ingredient sums over the clip's positions are part 2.
template <typename T>
__global__ void __launch_bounds__(PG_THREADS) pointer_bwd_kernel(
    int Lt, int D, int V, int max_ni, int max_vx, const T* __restrict__ dec, const float* __restrict__ G,
    const float* __restrict__ Esrc, const int* __restrict__ clip_src_off, const int* __restrict__ clip_ni,
    const int* __restrict__ clip_gi, const int* __restrict__ clip_vx, const int* __restrict__ clip_P_off,
    const int* __restrict__ tok_ptr, const int* __restrict__ tok_ids, const float* __restrict__ wpgen,
    const float* __restrict__ logits, long long ldl, const int* __restrict__ labels, float smoothing,
    const float* __restrict__ P, const float* __restrict__ alpha_in, const float* __restrict__ pgen_in,
    const float* __restrict__ dP, const float* __restrict__ dloss, T* __restrict__ ddec, float* __restrict__ du_out,
    float* __restrict__ dlogits, float* __restrict__ dsc_out) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  float* gP = sm;                    // max_vx
  float* dsc = gP + max_vx;          // max_ni   d score[i]
  float* alpha = dsc + max_ni;       // max_ni
  float* ci = alpha + max_ni;        // max_ni
  float* wE = ci + max_ni;           // max_ni   w2 . E_i
  __shared__ float red[64];
  __shared__ float s_du;
  const int n = blockIdx.y, l = blockIdx.x;
  const int src = clip_src_off[n], ni = clip_ni[n], gi = clip_gi[n], vx = clip_vx[n];
  const long long poff = clip_P_off[n];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const float gl = dloss[0];
  for (int i = warp; i < ni; i += ((int)blockDim.x >> 5)) {
    const float* e = Esrc + (long long)(src + i) * D;
    float s = 0.f;
    for (int d = lane; d < D; d += 32) s = fmaf(wpgen[D + d], e[d], s);
    s = warp_sum(s);
    if (lane == 0) wE[i] = s;
  }
  const long long r = (long long)n * Lt + l;
  const float pg = pgen_in[r];
  const int label = labels[r];
  const float s = smoothing / (float)(vx - 1), conf = 1.f - smoothing;
  const float* lrow = logits + r * ldl;
  for (int i = threadIdx.x; i < ni; i += (int)blockDim.x) alpha[i] = alpha_in[r * max_ni + i];
  float mx = -INFINITY;
  for (int v = threadIdx.x; v < V; v += (int)blockDim.x) mx = fmaxf(mx, lrow[v]);
  mx = block_max(mx, red);
  float se = 0.f;
  for (int v = threadIdx.x; v < V; v += (int)blockDim.x) se += expf(lrow[v] - mx);
  se = block_sum(se, red);
  // gradient w.r.t. the mixed probabilities
  float pa = 0.f;
  for (int v = threadIdx.x; v < vx; v += (int)blockDim.x) {
    float g = dP ? dP[poff + (long long)l * vx + v] : 0.f;
    if (label >= 0) g -= gl * ls_q(v, label, vx, s, conf) / (P[poff + (long long)l * vx + v] + 1e-12f);
    gP[v] = g;
    if (v < V) pa = fmaf(g, expf(lrow[v] - mx) / se, pa);
  }
  const float A = block_sum(pa, red);      // sum_v gP_v sm_v   (syncs: gP, wE, alpha visible)
  for (int i = threadIdx.x; i < ni; i += (int)blockDim.x) {
    const int lo = tok_ptr[gi + i], hi = tok_ptr[gi + i + 1];
    float c = 0.f;
    for (int t = lo; t < hi; ++t) c += gP[tok_ids[t]];
    ci[i] = c / (float)(hi - lo);
  }
  __syncthreads();
  if (warp == 0) {
    float t = 0.f;
    for (int i = lane; i < ni; i += 32) t = fmaf(alpha[i], ci[i], t);
    t = warp_sum(t);
    const float dpg = A - t;
    const float du = dpg * pg * (1.f - pg);
    // d alpha_i = (1 - pg) c_i + du * (w2 . E_i);  d score = alpha (d alpha - sum alpha d alpha)
    float dot = 0.f;
    for (int i = lane; i < ni; i += 32) dot = fmaf(alpha[i], (1.f - pg) * ci[i] + du * wE[i], dot);
    dot = warp_sum(dot);
    for (int i = lane; i < ni; i += 32) {
      const float dal = (1.f - pg) * ci[i] + du * wE[i];
      const float ds = alpha[i] * (dal - dot);
      dsc[i] = ds;
      dsc_out[r * max_ni + i] = ds;
    }
    if (lane == 0) { s_du = du; du_out[r] = du; }
  }
  __syncthreads();
  const float du = s_du;
  for (int v = threadIdx.x; v < V; v += (int)blockDim.x)
    dlogits[r * ldl + v] = (expf(lrow[v] - mx) / se) * pg * (gP[v] - A);
  for (int d = threadIdx.x; d < D; d += (int)blockDim.x) {
    float a = du * wpgen[d];
    for (int i = 0; i < ni; ++i) a = fmaf(dsc[i], G[(long long)(src + i) * D + d], a);
    stf(ddec + r * D + d, a);
  }
}

// Backward, part 1 for training shapes: one CTA per (clip, group of up to PR_ROWS positions), same outputs as
// pointer_bwd_kernel.  w2 . E_i is computed once per CTA, the vocabulary phase runs one warp per row with the row's
// softmax in registers, and the d dec rows read each G element once per CTA.
template <typename T>
__global__ void __launch_bounds__(PR_THREADS) pointer_bwd_rows_kernel(
    int Lt, int R, int D, int V, int max_ni, int max_vx, const float* __restrict__ G,
    const float* __restrict__ Esrc, const int* __restrict__ clip_src_off, const int* __restrict__ clip_ni,
    const int* __restrict__ clip_gi, const int* __restrict__ clip_vx, const int* __restrict__ clip_P_off,
    const int* __restrict__ tok_ptr, const int* __restrict__ tok_ids, const float* __restrict__ wpgen,
    const float* __restrict__ logits, long long ldl, const int* __restrict__ labels, float smoothing,
    const float* __restrict__ P, const float* __restrict__ alpha_in, const float* __restrict__ pgen_in,
    const float* __restrict__ dP, const float* __restrict__ dloss, T* __restrict__ ddec, float* __restrict__ du_out,
    float* __restrict__ dlogits, float* __restrict__ dsc_out) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  float* gPs = sm;                                   // PR_ROWS x max_vx
  float* dscs = gPs + (size_t)PR_ROWS * max_vx;      // PR_ROWS x max_ni   d score (zero for rows >= nr)
  float* cis = dscs + PR_ROWS * max_ni;              // PR_ROWS x max_ni
  float* wE = cis + PR_ROWS * max_ni;                // max_ni   w2 . E_i
  __shared__ float sdu[PR_ROWS];
  const int n = blockIdx.y, l0 = blockIdx.x * R;
  const int nr = min(R, Lt - l0);
  const int src = clip_src_off[n], ni = clip_ni[n], gi = clip_gi[n], vx = clip_vx[n];
  const long long poff = clip_P_off[n];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long r0 = (long long)n * Lt + l0;
  const float gl = dloss[0];
  for (int i = warp; i < ni; i += PR_THREADS / 32) {
    const float* e = Esrc + (long long)(src + i) * D;
    float acc = 0.f;
    for (int d0 = lane; d0 < D; d0 += 256) {
      float ev[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) ev[u] = d0 + 32 * u < D ? e[d0 + 32 * u] : 0.f;
#pragma unroll
      for (int u = 0; u < 8; ++u) acc = fmaf(wpgen[D + min(d0 + 32 * u, D - 1)], ev[u], acc);
    }
    acc = warp_sum(acc);
    if (lane == 0) wE[i] = acc;
  }
  if (threadIdx.x < PR_ROWS) sdu[threadIdx.x] = 0.f;
  for (int e = threadIdx.x; e < PR_ROWS * max_ni; e += PR_THREADS) dscs[e] = 0.f;
  __syncthreads();
  if (warp < nr) {
    const long long r = r0 + warp;
    const int l = l0 + warp;
    const float pg = pgen_in[r];
    const int label = labels[r];
    const float s = smoothing / (float)(vx - 1), conf = 1.f - smoothing;
    const float* lrow = logits + r * ldl;
    const float* al = alpha_in + r * max_ni;
    float* gP = gPs + (size_t)warp * max_vx;
    float* ci = cis + warp * max_ni;
    float lv[PR_NV];            // logits, then softmax probabilities, of this lane's columns
#pragma unroll
    for (int k = 0; k < PR_NV; ++k) lv[k] = lane + 32 * k < V ? lrow[lane + 32 * k] : -INFINITY;
    float mx = -INFINITY;
#pragma unroll
    for (int k = 0; k < PR_NV; ++k) mx = fmaxf(mx, lv[k]);
    mx = warp_max(mx);
    float se = 0.f;
#pragma unroll
    for (int k = 0; k < PR_NV; ++k) { lv[k] = expf(lv[k] - mx); se += lv[k]; }
    se = warp_sum(se);
    const float* Prow = P + poff + (long long)l * vx;
    const float* dProw = dP ? dP + poff + (long long)l * vx : nullptr;
    float pa = 0.f;
#pragma unroll
    for (int k0 = 0; k0 < PR_NV; k0 += 8) {     // gradient w.r.t. the mixed probabilities, eight columns per lane in flight
      float pv[8], gv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = lane + 32 * (k0 + u);
        pv[u] = v < vx ? Prow[v] : 1.f;
        gv[u] = (dProw && v < vx) ? dProw[v] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = lane + 32 * (k0 + u);
        lv[k0 + u] /= se;
        if (v < vx) {
          float g = gv[u];
          if (label >= 0) g -= gl * ls_q(v, label, vx, s, conf) / (pv[u] + 1e-12f);
          gP[v] = g;
          if (v < V) pa = fmaf(g, lv[k0 + u], pa);
        }
      }
    }
    for (int v = 32 * PR_NV + lane; v < vx; v += 32) {      // OOV columns beyond the register window (never vocabulary columns)
      float g = dProw ? dProw[v] : 0.f;
      if (label >= 0) g -= gl * ls_q(v, label, vx, s, conf) / (Prow[v] + 1e-12f);
      gP[v] = g;
    }
    const float A = warp_sum(pa);
    __syncwarp();
    float t = 0.f;
    for (int i = lane; i < ni; i += 32) {
      const int lo = tok_ptr[gi + i], hi = tok_ptr[gi + i + 1];
      float c = 0.f;
      for (int k = lo; k < hi; ++k) c += gP[tok_ids[k]];
      c /= (float)(hi - lo);
      ci[i] = c;
      t = fmaf(al[i], c, t);
    }
    t = warp_sum(t);
    const float dpg = A - t;
    const float du = dpg * pg * (1.f - pg);
    float dot = 0.f;
    for (int i = lane; i < ni; i += 32) dot = fmaf(al[i], (1.f - pg) * ci[i] + du * wE[i], dot);
    dot = warp_sum(dot);
    for (int i = lane; i < ni; i += 32) {
      const float dal = (1.f - pg) * ci[i] + du * wE[i];
      const float ds = al[i] * (dal - dot);
      dscs[warp * max_ni + i] = ds;
      dsc_out[r * max_ni + i] = ds;
    }
    if (lane == 0) { sdu[warp] = du; du_out[r] = du; }
#pragma unroll
    for (int k = 0; k < PR_NV; ++k) {
      const int v = lane + 32 * k;
      if (v < V) dlogits[r * ldl + v] = lv[k] * pg * (gP[v] - A);
    }
  }
  __syncthreads();
  for (int d = threadIdx.x; d < D; d += PR_THREADS) {
    const float w1 = wpgen[d];
    float acc[PR_ROWS];
#pragma unroll
    for (int r = 0; r < PR_ROWS; ++r) acc[r] = sdu[r] * w1;
    for (int i0 = 0; i0 < ni; i0 += 4) {
      float gv[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) gv[u] = i0 + u < ni ? G[(long long)(src + i0 + u) * D + d] : 0.f;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int ii = min(i0 + u, ni - 1);
#pragma unroll
        for (int r = 0; r < PR_ROWS; ++r) acc[r] = fmaf(dscs[r * max_ni + ii], gv[u], acc[r]);
      }
    }
#pragma unroll
    for (int r = 0; r < PR_ROWS; ++r)
      if (r < nr) stf(ddec + (r0 + r) * D + d, acc[r]);
  }
}

// Backward, part 2: one CTA per clip.  dG_i[d] += sum_l dscore[l][i] dec[l][d] ;  dE_i[d] += (sum_l alpha_il du_l) w2[d]
template <typename T>
__global__ void __launch_bounds__(PG_THREADS) pointer_bwd_reduce_kernel(
    int Lt, int D, int max_ni, const T* __restrict__ dec, const int* __restrict__ clip_src_off,
    const int* __restrict__ clip_ni, const float* __restrict__ wpgen, const float* __restrict__ alpha_in,
    const float* __restrict__ du, const float* __restrict__ dsc_in, float* __restrict__ dG, float* __restrict__ dE) {
  svpc_pdl_sync();
  extern __shared__ float sm[];
  float* dsc = sm;                   // Lt * max_ni
  float* beta = dsc + Lt * max_ni;   // max_ni
  const int n = blockIdx.x;
  const int src = clip_src_off[n], ni = clip_ni[n];
  for (int e = threadIdx.x; e < Lt * max_ni; e += PG_THREADS) dsc[e] = dsc_in[(long long)n * Lt * max_ni + e];
  for (int i = threadIdx.x; i < ni; i += PG_THREADS) {
    float b = 0.f;
    for (int l = 0; l < Lt; ++l) {
      const long long r = (long long)n * Lt + l;
      b = fmaf(alpha_in[r * max_ni + i], du[r], b);
    }
    beta[i] = b;
  }
  __syncthreads();
  if (ni <= 16) {          // usual case: every ingredient's sum in registers, ONE pass over the clip's decoder rows
    for (int d = blockIdx.y * PG_THREADS + threadIdx.x; d < D; d += gridDim.y * PG_THREADS) {
      const float w2 = wpgen[D + d];
      float a[16];
#pragma unroll
      for (int k = 0; k < 16; ++k) a[k] = 0.f;
#pragma unroll 5
      for (int l = 0; l < Lt; ++l) {
        const float x = ldf(dec + ((long long)n * Lt + l) * D + d);
#pragma unroll
        for (int k = 0; k < 16; ++k)
          if (k < ni) a[k] = fmaf(dsc[l * max_ni + k], x, a[k]);
      }
#pragma unroll
      for (int k = 0; k < 16; ++k)
        if (k < ni) {
          atomicAdd(dG + (long long)(src + k) * D + d, a[k]);
          atomicAdd(dE + (long long)(src + k) * D + d, beta[k] * w2);
        }
    }
    return;
  }
  for (int d = blockIdx.y * PG_THREADS + threadIdx.x; d < D; d += gridDim.y * PG_THREADS) {
    const float w2 = wpgen[D + d];
    for (int i0 = 0; i0 < ni; i0 += 4) {        // 4 ingredients per pass over the clip's decoder rows
      float a[4] = {0.f, 0.f, 0.f, 0.f};
      for (int l = 0; l < Lt; ++l) {
        const float x = ldf(dec + ((long long)n * Lt + l) * D + d);
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (i0 + k < ni) a[k] = fmaf(dsc[l * max_ni + i0 + k], x, a[k]);
      }
#pragma unroll
      for (int k = 0; k < 4; ++k)
        if (i0 + k < ni) {
          atomicAdd(dG + (long long)(src + i0 + k) * D + d, a[k]);
          atomicAdd(dE + (long long)(src + i0 + k) * D + d, beta[i0 + k] * w2);
        }
    }
  }
}

int svpc_weighted_colsum(int rows, int cols, const void* x, int dt, const float* w, float* out, cudaStream_t st);

extern "C" int svpc_pointer_fwd(int n_clips, int Lt, int D, int V, int max_ni, int max_vx, const void* dec, int dt,
                                const float* G, const float* Esrc, const int* clip_src_off, const int* clip_ni,
                                const int* clip_gi, const int* clip_vx, const int* clip_P_off, const int* tok_ptr,
                                const int* tok_ids, const float* wpgen, const float* bpgen, const float* logits,
                                long long ldl, const int* labels, float smoothing, float* P, float* loss_rows,
                                float* alpha, float* pgen, float* ctx, svpc_stream_t stream) {
  if (n_clips <= 0) return SVPC_OK;
  const size_t smem = sizeof(float) * (2 * (size_t)D + max_vx + max_ni);
  SVPC_REQUIRE(smem <= 200 * 1024, "pointer_fwd: shared memory");
  cudaStream_t st = (cudaStream_t)stream;
  const size_t smem_rows = sizeof(float) * PR_ROWS * ((size_t)D + max_vx + max_ni);
  if (ptr_rows_route(Lt, n_clips, V) && smem_rows <= 100 * 1024) {
    const int groups = (Lt + PR_ROWS - 1) / PR_ROWS, R = (Lt + groups - 1) / groups;
    SVPC_DISPATCH(dt, T, {
      cudaFuncSetAttribute(pointer_fwd_rows_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_rows);
      svpc_launch(pointer_fwd_rows_kernel<T>, dim3((Lt + R - 1) / R, n_clips), PR_THREADS, smem_rows, st, Lt, R, D, V, max_ni,
                  max_vx, (const T*)dec, G, Esrc, clip_src_off, clip_ni, clip_gi, clip_vx, clip_P_off, tok_ptr, tok_ids, wpgen,
                  bpgen, logits, ldl, labels, smoothing, P, loss_rows, alpha, pgen, ctx);
    });
    return svpc_check_launch("pointer_fwd_rows");
  }
  SVPC_DISPATCH(dt, T, {
    cudaFuncSetAttribute(pointer_fwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    svpc_launch(pointer_fwd_kernel<T>, dim3(Lt, n_clips), ptr_threads(Lt, n_clips), smem, st, Lt, D, V, max_ni, max_vx, (const T*)dec, G, Esrc, clip_src_off,
                                                            clip_ni, clip_gi, clip_vx, clip_P_off, tok_ptr, tok_ids, wpgen,
                                                            bpgen, logits, ldl, labels, smoothing, P, loss_rows, alpha, pgen, ctx);
  });
  return svpc_check_launch("pointer_fwd");
}

extern "C" int svpc_pointer_bwd(int n_clips, int Lt, int D, int V, int max_ni, int max_vx, const void* dec, int dt,
                                const float* G, const float* Esrc, const int* clip_src_off, const int* clip_ni,
                                const int* clip_gi, const int* clip_vx, const int* clip_P_off, const int* tok_ptr,
                                const int* tok_ids, const float* wpgen, const float* logits, long long ldl,
                                const int* labels, float smoothing, const float* P, const float* alpha, const float* pgen,
                                const float* ctx, const float* dP, const float* dloss, void* ddec, float* dG, float* dE,
                                float* du, float* dwpgen, float* dbpgen, float* dlogits, float* dscore_ws,
                                svpc_stream_t stream) {
  if (n_clips <= 0) return SVPC_OK;
  SVPC_REQUIRE(dscore_ws != nullptr, "pointer_bwd: dscore_ws (n_clips * Lt * max_ni floats) is required");
  const size_t smem = sizeof(float) * ((size_t)max_vx + 4 * (size_t)max_ni);
  const size_t smem2 = sizeof(float) * ((size_t)max_ni * Lt + (size_t)max_ni);
  SVPC_REQUIRE(smem <= 200 * 1024 && smem2 <= 200 * 1024, "pointer_bwd: shared memory");
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(dwpgen, 0, sizeof(float) * 2 * D, st);
  const size_t smem_rows = sizeof(float) * (PR_ROWS * ((size_t)max_vx + 2 * (size_t)max_ni) + max_ni);
  const bool rows_route = ptr_rows_route(Lt, n_clips, V) && smem_rows <= 100 * 1024;
  const int groups = (Lt + PR_ROWS - 1) / PR_ROWS, RG = (Lt + groups - 1) / groups;
  SVPC_DISPATCH(dt, T, {
    if (rows_route) {
      cudaFuncSetAttribute(pointer_bwd_rows_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_rows);
      svpc_launch(pointer_bwd_rows_kernel<T>, dim3((Lt + RG - 1) / RG, n_clips), PR_THREADS, smem_rows, st, Lt, RG, D, V, max_ni,
          max_vx, G, Esrc, clip_src_off, clip_ni, clip_gi, clip_vx, clip_P_off, tok_ptr, tok_ids, wpgen, logits, ldl, labels,
          smoothing, P, alpha, pgen, dP, dloss, (T*)ddec, du, dlogits, dscore_ws);
    } else {
      cudaFuncSetAttribute(pointer_bwd_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
      svpc_launch(pointer_bwd_kernel<T>, dim3(Lt, n_clips), ptr_threads(Lt, n_clips), smem, st,
          Lt, D, V, max_ni, max_vx, (const T*)dec, G, Esrc, clip_src_off, clip_ni, clip_gi, clip_vx, clip_P_off, tok_ptr,
          tok_ids, wpgen, logits, ldl, labels, smoothing, P, alpha, pgen, dP, dloss, (T*)ddec, du, dlogits, dscore_ws);
    }
    cudaFuncSetAttribute(pointer_bwd_reduce_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem2);
    svpc_launch(pointer_bwd_reduce_kernel<T>, dim3(n_clips, svpc_cdiv(D, PG_THREADS)), PG_THREADS, smem2, st, Lt, D, max_ni, (const T*)dec, clip_src_off, clip_ni,
                                                                   wpgen, alpha, du, dscore_ws, dG, dE);
  });
  int rc = svpc_check_launch("pointer_bwd");
  if (rc) return rc;
  const int R = n_clips * Lt;
  rc = svpc_weighted_colsum(R, D, dec, dt, du, dwpgen, st);
  if (rc) return rc;
  rc = svpc_weighted_colsum(R, D, ctx, SVPC_F32, du, dwpgen + D, st);
  if (rc) return rc;
  return svpc_sum(R, du, dbpgen, stream);
}

// ------------------------------------------------------------------------------------------------
// 'video' mode: P = softmax(logits), label-smoothed KL on P
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(PG_THREADS) softmax_ls_fwd_kernel(int V, const float* __restrict__ logits, long long ldl,
                                                                    const int* __restrict__ labels, float smoothing,
                                                                    float* __restrict__ P, float* __restrict__ loss_rows) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const long long r = blockIdx.x;
  const float* lrow = logits + r * ldl;
  float mx = -INFINITY;
  for (int v = threadIdx.x; v < V; v += PG_THREADS) mx = fmaxf(mx, lrow[v]);
  mx = block_max(mx, red);
  float se = 0.f;
  for (int v = threadIdx.x; v < V; v += PG_THREADS) se += expf(lrow[v] - mx);
  se = block_sum(se, red);
  const int label = labels[r];
  const float s = smoothing / (float)(V - 1), conf = 1.f - smoothing;
  float acc = 0.f;
  for (int v = threadIdx.x; v < V; v += PG_THREADS) {
    const float p = expf(lrow[v] - mx) / se;
    P[r * V + v] = p;
    if (label >= 0) acc += ls_q(v, label, V, s, conf) * logf(p + 1e-12f);
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) loss_rows[r] = label >= 0 ? ls_const(label, V, s, conf) - acc : 0.f;
}

__global__ void __launch_bounds__(PG_THREADS) softmax_ls_bwd_kernel(int V, const float* __restrict__ dP,
                                                                    const float* __restrict__ dloss,
                                                                    const float* __restrict__ P,
                                                                    const int* __restrict__ labels, float smoothing,
                                                                    float* __restrict__ dlogits, long long ldl) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const long long r = blockIdx.x;
  const int label = labels[r];
  const float gl = dloss[0];
  const float s = smoothing / (float)(V - 1), conf = 1.f - smoothing;
  float dot = 0.f;
  for (int v = threadIdx.x; v < V; v += PG_THREADS) {
    const float p = P[r * V + v];
    float g = dP ? dP[r * V + v] : 0.f;
    if (label >= 0) g -= gl * ls_q(v, label, V, s, conf) / (p + 1e-12f);
    dot = fmaf(g, p, dot);
  }
  dot = block_sum(dot, red);
  for (int v = threadIdx.x; v < V; v += PG_THREADS) {
    const float p = P[r * V + v];
    float g = dP ? dP[r * V + v] : 0.f;
    if (label >= 0) g -= gl * ls_q(v, label, V, s, conf) / (p + 1e-12f);
    dlogits[r * ldl + v] = p * (g - dot);
  }
}

extern "C" int svpc_softmax_ls_fwd(int rows, int V, const float* logits, long long ldl, const int* labels,
                                   float smoothing, float* P, float* loss_rows, svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  svpc_launch(softmax_ls_fwd_kernel, rows, PG_THREADS, 0, (cudaStream_t)stream, V, logits, ldl, labels, smoothing, P, loss_rows);
  return svpc_check_launch("softmax_ls_fwd");
}
extern "C" int svpc_softmax_ls_bwd(int rows, int V, const float* dP, const float* dloss, const float* P,
                                   const int* labels, float smoothing, float* dlogits, long long ldl,
                                   svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  svpc_launch(softmax_ls_bwd_kernel, rows, PG_THREADS, 0, (cudaStream_t)stream, V, dP, dloss, P, labels, smoothing, dlogits, ldl);
  return svpc_check_launch("softmax_ls_bwd");
}

// ------------------------------------------------------------------------------------------------
// textual re-simulator front end
// ------------------------------------------------------------------------------------------------
// One CTA per (clip, text position): hard Gumbel-softmax of one row -> argmax index and straight-through value.
__global__ void __launch_bounds__(PG_THREADS) gumbel_fwd_kernel(int Lt, int V, const float* __restrict__ P,
                                                                const float* __restrict__ g,
                                                                const int* __restrict__ clip_P_off,
                                                                const int* __restrict__ clip_vx, float tau,
                                                                int* __restrict__ idx_out, float* __restrict__ yval_out) {
  svpc_pdl_sync();
  __shared__ float red[64];
  __shared__ float smax[PG_THREADS / 32];
  __shared__ int sidx[PG_THREADS / 32];
  __shared__ int s_best;
  __shared__ float s_mx;
  const int n = blockIdx.y, l = blockIdx.x, vx = clip_vx[n];
  const long long poff = clip_P_off[n];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const long long r = (long long)n * Lt + l;
  const float* prow = P + poff + (long long)l * vx;
  const float* grow = g + poff + (long long)l * vx;
  float mx = -INFINITY;
  int best = 0x7fffffff;
  float zloc[4];                                  // this thread's z values (vx <= 4 * PG_THREADS on the fast path)
  const bool cache = vx <= 4 * PG_THREADS;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int v = threadIdx.x + k * PG_THREADS;
    zloc[k] = -INFINITY;
    if (cache && v < vx) {
      const float z = (logf(prow[v] + 1e-12f) + grow[v]) / tau;
      zloc[k] = z;
      if (z > mx) { mx = z; best = v; }
    }
  }
  if (!cache)
    for (int v = threadIdx.x; v < vx; v += PG_THREADS) {
      const float z = (logf(prow[v] + 1e-12f) + grow[v]) / tau;
      if (z > mx) { mx = z; best = v; }
    }
  for (int o = 16; o > 0; o >>= 1) {
    const float om = __shfl_xor_sync(0xffffffffu, mx, o);
    const int ob = __shfl_xor_sync(0xffffffffu, best, o);
    if (om > mx || (om == mx && ob < best)) { mx = om; best = ob; }
  }
  if (lane == 0) { smax[warp] = mx; sidx[warp] = best; }
  __syncthreads();
  if (threadIdx.x == 0) {
    float m = smax[0]; int bi = sidx[0];
    for (int w = 1; w < PG_THREADS / 32; ++w)
      if (smax[w] > m || (smax[w] == m && sidx[w] < bi)) { m = smax[w]; bi = sidx[w]; }
    if (bi >= vx) bi = 0;          // a row of NaNs compares false everywhere: never hand out an index outside the row
    s_best = bi; s_mx = m;
  }
  __syncthreads();
  mx = s_mx; best = s_best;
  float se = 0.f;
  if (cache) {
#pragma unroll
    for (int k = 0; k < 4; ++k)
      if (threadIdx.x + k * PG_THREADS < vx) se += expf(zloc[k] - mx);
  } else {
    for (int v = threadIdx.x; v < vx; v += PG_THREADS) se += expf((logf(prow[v] + 1e-12f) + grow[v]) / tau - mx);
  }
  se = block_sum(se, red);
  const float y = 1.0f / se;                 // y_soft at its argmax
  const float yv = (1.0f - y) + y;           // y_hard - y_soft.detach() + y_soft
  if (threadIdx.x == 0) { idx_out[r] = best; yval_out[r] = yv; }
}

// Training shapes: one WARP per (clip, text position) row, eight rows per CTA.  The per-row CTA above keeps 8 rows in flight
// per SM and pays three dependent L2 round trips plus two block barriers per row; here 64 rows are in flight per SM, the row
// is read eight columns per lane at a time, and the reductions are warp shuffles.
__global__ void __launch_bounds__(PG_THREADS) gumbel_fwd_rows_kernel(int n_rows, int Lt, int V, const float* __restrict__ P,
                                                                     const float* __restrict__ g,
                                                                     const int* __restrict__ clip_P_off,
                                                                     const int* __restrict__ clip_vx, float tau,
                                                                     int* __restrict__ idx_out, float* __restrict__ yval_out) {
  svpc_pdl_sync();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * (PG_THREADS / 32) + warp;
  if (r >= n_rows) return;
  const int n = r / Lt, l = r - n * Lt, vx = clip_vx[n];
  const long long poff = clip_P_off[n];
  const float* prow = P + poff + (long long)l * vx;
  const float* grow = g + poff + (long long)l * vx;
  float mx = -INFINITY;
  int best = 0x7fffffff;
  float se = 0.f;
  if (vx <= 32 * PR_NV) {
    // the whole row in registers: every load of the row is issued at once, log / scale computed once per element (the
    // kernel is bound by its instruction count, not by memory: ~35 instead of ~70 instructions per element)
    float z[PR_NV];
#pragma unroll
    for (int k0 = 0; k0 < PR_NV; k0 += 8) {
      float pv[8], gv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = lane + 32 * (k0 + u);
        pv[u] = v < vx ? prow[v] : 1.f;
        gv[u] = v < vx ? grow[v] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = lane + 32 * (k0 + u);
        z[k0 + u] = v < vx ? (logf(pv[u] + 1e-12f) + gv[u]) / tau : -INFINITY;     // same expression as the reference's
        if (z[k0 + u] > mx) { mx = z[k0 + u]; best = v; }
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int ob = __shfl_xor_sync(0xffffffffu, best, o);
      if (om > mx || (om == mx && ob < best)) { mx = om; best = ob; }
    }
#pragma unroll
    for (int k2 = 0; k2 < PR_NV; ++k2) se += expf(z[k2] - mx);        // exp(-inf) = 0 for the padding
  } else {
    for (int v0 = lane; v0 < vx; v0 += 256) {
      float pv[8], gv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = v0 + 32 * u;
        pv[u] = v < vx ? prow[v] : 1.f;
        gv[u] = v < vx ? grow[v] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = v0 + 32 * u;
        const float zz = (logf(pv[u] + 1e-12f) + gv[u]) / tau;
        if (v < vx && zz > mx) { mx = zz; best = v; }
      }
    }
    for (int o = 16; o > 0; o >>= 1) {
      const float om = __shfl_xor_sync(0xffffffffu, mx, o);
      const int ob = __shfl_xor_sync(0xffffffffu, best, o);
      if (om > mx || (om == mx && ob < best)) { mx = om; best = ob; }
    }
    for (int v = lane; v < vx; v += 32) se += expf((logf(prow[v] + 1e-12f) + grow[v]) / tau - mx);
  }
  if (best >= vx) best = 0;          // a row of NaNs compares false everywhere: never hand out an index outside the row
  se = warp_sum(se);
  const float y = 1.0f / se;                 // y_soft at its argmax
  const float yv = (1.0f - y) + y;           // y_hard - y_soft.detach() + y_soft
  if (lane == 0) { idx_out[r] = best; yval_out[r] = yv; }
}

// One CTA per clip: masked mean over the text positions of yv * Emb[argmax] (positions in order, as the reference sums)
__global__ void __launch_bounds__(PG_THREADS) gumbel_mean_kernel(int Lt, int V, int W, const float* __restrict__ emb,
                                                                 const float* __restrict__ text_mask,
                                                                 const int* __restrict__ idx, const float* __restrict__ yval,
                                                                 float* __restrict__ xmean) {
  svpc_pdl_sync();
  const int n = blockIdx.x;
  float cnt = 0.f;
  for (int l = 0; l < Lt; ++l) cnt += (text_mask[(long long)n * Lt + l] == 1.0f) ? 1.f : 0.f;
  for (int w = threadIdx.x; w < W; w += PG_THREADS) {
    float acc = 0.f;
    for (int l = 0; l < Lt; ++l) {
      const long long r = (long long)n * Lt + l;
      const int best = idx[r];
      if (text_mask[r] == 1.0f && best < V) acc = fmaf(yval[r], emb[(long long)best * W + w], acc);
    }
    xmean[(long long)n * W + w] = acc / cnt;
  }
}

// tv[n][v] = Emb[v] . dxmean[n]: one warp per 8 vocabulary entries
__global__ void __launch_bounds__(PG_THREADS) gumbel_tv_kernel(int V, int W, const float* __restrict__ emb,
                                                               const float* __restrict__ dxmean, float* __restrict__ tv) {
  svpc_pdl_sync();
  extern __shared__ float dxm[];   // W
  const int n = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int w = threadIdx.x; w < W; w += PG_THREADS) dxm[w] = dxmean[(long long)n * W + w];
  __syncthreads();
  const int v0 = blockIdx.x * 64 + warp * 8;
  for (int k = 0; k < 8; ++k) {
    const int v = v0 + k;
    if (v >= V) break;
    const float* e = emb + (long long)v * W;
    float s = 0.f;
    for (int w = lane; w < W; w += 32) s = fmaf(e[w], dxm[w], s);
    s = warp_sum(s);
    if (lane == 0) tv[(long long)n * V + v] = s;
  }
}

// tv[n][v] = Emb[v] . dxmean[n] for training shapes: a CTA owns 32 clips (one per LANE, their dxmean rows in shared memory
// with an odd pitch) and TV_VPB vocabulary entries (their Emb rows in shared memory too); a warp reads one Emb row as
// broadcast 16-byte reads and every lane
// accumulates its own clip's dot product - each Emb row is read once per 32 clips instead of once per clip, and there is no
// shuffle reduction.
#define TV_CLIPS 32
#define TV_VPB 40
__global__ void __launch_bounds__(PG_THREADS) gumbel_tv_clips_kernel(int n_clips, int V, int W, const float* __restrict__ emb,
                                                                     const float* __restrict__ dxmean, float* __restrict__ tv) {
  svpc_pdl_sync();
  extern __shared__ __align__(16) float dxm[];   // TV_CLIPS x (W + 1), then the CTA's Emb rows
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int c0 = blockIdx.y * TV_CLIPS, pitch = W + 1;
  float* semb = dxm + (((size_t)TV_CLIPS * pitch + 3) & ~(size_t)3);      // TV_VPB x W (W % 4 == 0), 16-byte aligned rows
  for (int i = threadIdx.x; i < TV_CLIPS * W; i += PG_THREADS) {
    const int c = i / W, w = i - c * W;
    dxm[c * pitch + w] = c0 + c < n_clips ? dxmean[(long long)(c0 + c) * W + w] : 0.f;
  }
  {   // this CTA's Emb rows: contiguous in global memory, copied with every thread's loads in flight
    const int v0 = blockIdx.x * TV_VPB, nv = min(V - v0, TV_VPB);
    const float4* src = reinterpret_cast<const float4*>(emb + (long long)v0 * W);
    float4* dst = reinterpret_cast<float4*>(semb);
    for (int i = threadIdx.x; i < nv * W / 4; i += PG_THREADS) dst[i] = __ldg(src + i);
  }
  __syncthreads();
  const float* mine = dxm + lane * pitch;
  const int vbeg = blockIdx.x * TV_VPB, vend = min(V, vbeg + TV_VPB);
  for (int v = vbeg + warp; v < vend; v += PG_THREADS / 32) {
    const float* e = semb + (size_t)(v - vbeg) * W;
    float s0 = 0.f, s1 = 0.f, s2 = 0.f, s3 = 0.f;
    for (int w = 0; w < W; w += 4) {
      const float4 ev = *reinterpret_cast<const float4*>(e + w);          // broadcast read
      s0 = fmaf(ev.x, mine[w], s0); s1 = fmaf(ev.y, mine[w + 1], s1);
      s2 = fmaf(ev.z, mine[w + 2], s2); s3 = fmaf(ev.w, mine[w + 3], s3);
    }
    if (c0 + lane < n_clips) tv[(long long)(c0 + lane) * V + v] = (s0 + s1) + (s2 + s3);
  }
}

// Backward, one warp per (clip, text position) row (see gumbel_fwd_rows_kernel)
__global__ void __launch_bounds__(PG_THREADS) gumbel_bwd_rows_kernel(int n_rows, int Lt, int V, int W, const float* __restrict__ P,
                                                                     const float* __restrict__ g,
                                                                     const int* __restrict__ clip_P_off,
                                                                     const int* __restrict__ clip_vx, float tau,
                                                                     const float* __restrict__ text_mask,
                                                                     const int* __restrict__ idx_in,
                                                                     const float* __restrict__ yval_in,
                                                                     const float* __restrict__ dxmean,
                                                                     const float* __restrict__ tv_all, float* __restrict__ dP,
                                                                     float* __restrict__ demb) {
  svpc_pdl_sync();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r = blockIdx.x * (PG_THREADS / 32) + warp;
  if (r >= n_rows) return;
  const int n = r / Lt, l = r - n * Lt, vx = clip_vx[n];
  const long long poff = clip_P_off[n];
  const float* prow = P + poff + (long long)l * vx;
  const float* grow = g + poff + (long long)l * vx;
  const float* tv = tv_all + (long long)n * V;
  float* drow = dP + poff + (long long)l * vx;
  const float msk = text_mask[r];
  const int best = idx_in[r];
  float cnt = 0.f;
  for (int k = lane; k < Lt; k += 32) cnt += (text_mask[(long long)n * Lt + k] == 1.0f) ? 1.f : 0.f;
  cnt = warp_sum(cnt);
  if (msk != 1.0f) {
    for (int v = lane; v < vx; v += 32) drow[v] = 0.f;
    return;
  }
  const float mx = (logf(prow[best] + 1e-12f) + grow[best]) / tau;
  const float inv_cnt = 1.f / cnt, inv_tau = 1.f / tau;
  float se = 0.f, sd = 0.f;
  if (vx <= 32 * PR_NV) {
    // the row's e = exp(z - max) and P in registers: one log / exp per element instead of two (instruction-bound kernel)
    float ev[PR_NV], pc[PR_NV];
#pragma unroll
    for (int k0 = 0; k0 < PR_NV; k0 += 8) {
      float gv[8], tvv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = lane + 32 * (k0 + u);
        pc[k0 + u] = v < vx ? prow[v] : 1.f;
        gv[u] = v < vx ? grow[v] : 0.f;
        tvv[u] = v < V ? tv[v] : 0.f;
      }
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = lane + 32 * (k0 + u);
        const float e = v < vx ? expf((logf(pc[k0 + u] + 1e-12f) + gv[u]) / tau - mx) : 0.f;
        ev[k0 + u] = e;
        se += e;
        sd = fmaf(e, tvv[u], sd);
      }
    }
    se = warp_sum(se);
    sd = warp_sum(sd);
    const float inv = 1.f / se;
    const float sdot = sd * inv * inv_cnt;              // sum_v y_v dy_v
#pragma unroll
    for (int k0 = 0; k0 < PR_NV; k0 += 8) {
      float tvv[8];
#pragma unroll
      for (int u = 0; u < 8; ++u) tvv[u] = lane + 32 * (k0 + u) < V ? tv[lane + 32 * (k0 + u)] : 0.f;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int v = lane + 32 * (k0 + u);
        if (v < vx) drow[v] = ev[k0 + u] * inv * (tvv[u] * inv_cnt - sdot) * inv_tau / (pc[k0 + u] + 1e-12f);
      }
    }
  } else {
    for (int v = lane; v < vx; v += 32) {
      const float e = expf((logf(prow[v] + 1e-12f) + grow[v]) / tau - mx);
      se += e;
      if (v < V) sd = fmaf(e, tv[v], sd);
    }
    se = warp_sum(se);
    sd = warp_sum(sd);
    const float inv = 1.f / se;
    const float sdot = sd * inv * inv_cnt;
    for (int v = lane; v < vx; v += 32) {
      const float y = expf((logf(prow[v] + 1e-12f) + grow[v]) / tau - mx) * inv;
      const float dy = v < V ? tv[v] * inv_cnt : 0.f;
      drow[v] = y * (dy - sdot) * inv_tau / (prow[v] + 1e-12f);
    }
  }
  if (best < V) {
    const float c = yval_in[r] / cnt;
    for (int w = lane; w < W; w += 32) atomicAdd(demb + (long long)best * W + w, c * dxmean[(long long)n * W + w]);
  }
}

// One CTA per (clip, text position)
__global__ void __launch_bounds__(PG_THREADS) gumbel_bwd_kernel(int Lt, int V, int W, const float* __restrict__ P,
                                                                const float* __restrict__ g,
                                                                const int* __restrict__ clip_P_off,
                                                                const int* __restrict__ clip_vx, float tau,
                                                                const float* __restrict__ text_mask,
                                                                const int* __restrict__ idx_in,
                                                                const float* __restrict__ yval_in,
                                                                const float* __restrict__ dxmean,
                                                                const float* __restrict__ tv_all, float* __restrict__ dP,
                                                                float* __restrict__ demb) {
  svpc_pdl_sync();
  __shared__ float red[64];
  const int n = blockIdx.y, l = blockIdx.x, vx = clip_vx[n];
  const long long poff = clip_P_off[n];
  const long long r = (long long)n * Lt + l;
  const float* prow = P + poff + (long long)l * vx;
  const float* grow = g + poff + (long long)l * vx;
  const float* tv = tv_all + (long long)n * V;
  float* drow = dP + poff + (long long)l * vx;
  if (text_mask[r] != 1.0f) {
    for (int v = threadIdx.x; v < vx; v += PG_THREADS) drow[v] = 0.f;
    return;
  }
  float cnt = 0.f;
  for (int k = 0; k < Lt; ++k) cnt += (text_mask[(long long)n * Lt + k] == 1.0f) ? 1.f : 0.f;
  const int best = idx_in[r];
  const float mx = (logf(prow[best] + 1e-12f) + grow[best]) / tau;
  float se = 0.f, sd = 0.f;
  float eloc[4], ploc[4];
  const bool cache = vx <= 4 * PG_THREADS;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const int v = threadIdx.x + k * PG_THREADS;
    eloc[k] = 0.f; ploc[k] = 0.f;
    if (cache && v < vx) {
      const float p = prow[v];
      const float e = expf((logf(p + 1e-12f) + grow[v]) / tau - mx);
      eloc[k] = e; ploc[k] = p;
      se += e;
      if (v < V) sd = fmaf(e, tv[v], sd);
    }
  }
  if (!cache)
    for (int v = threadIdx.x; v < vx; v += PG_THREADS) {
      const float e = expf((logf(prow[v] + 1e-12f) + grow[v]) / tau - mx);
      se += e;
      if (v < V) sd = fmaf(e, tv[v], sd);
    }
  const float2 rr = block_sum2(se, sd, red);
  const float inv = 1.f / rr.x;
  const float sdot = rr.y * inv / cnt;              // sum_v y_v dy_v
  if (cache) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
      const int v = threadIdx.x + k * PG_THREADS;
      if (v < vx) {
        const float y = eloc[k] * inv;
        const float dy = v < V ? tv[v] / cnt : 0.f;
        drow[v] = y * (dy - sdot) / tau / (ploc[k] + 1e-12f);
      }
    }
  } else {
    for (int v = threadIdx.x; v < vx; v += PG_THREADS) {
      const float y = expf((logf(prow[v] + 1e-12f) + grow[v]) / tau - mx) * inv;
      const float dy = v < V ? tv[v] / cnt : 0.f;
      drow[v] = y * (dy - sdot) / tau / (prow[v] + 1e-12f);
    }
  }
  if (best < V) {
    const float c = yval_in[r] / cnt;
    for (int w = threadIdx.x; w < W; w += PG_THREADS) atomicAdd(demb + (long long)best * W + w, c * dxmean[(long long)n * W + w]);
  }
}

extern "C" int svpc_gumbel_fwd(int n_clips, int Lt, int V, int W, const float* P, const float* g, const int* clip_P_off,
                               const int* clip_vx, float tau, const float* emb, const float* text_mask, float* xmean,
                               int* idx, float* yval, svpc_stream_t stream) {
  if (n_clips <= 0) return SVPC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int n_rows = n_clips * Lt;
  if (ptr_rows_route(Lt, n_clips, V))
    svpc_launch(gumbel_fwd_rows_kernel, svpc_cdiv(n_rows, PG_THREADS / 32), PG_THREADS, 0, st, n_rows, Lt, V, P, g, clip_P_off, clip_vx,
                tau, idx, yval);
  else
    svpc_launch(gumbel_fwd_kernel, dim3(Lt, n_clips), PG_THREADS, 0, st, Lt, V, P, g, clip_P_off, clip_vx, tau, idx, yval);
  int rc = svpc_check_launch("gumbel_fwd");
  if (rc) return rc;
  svpc_launch(gumbel_mean_kernel, n_clips, PG_THREADS, 0, st, Lt, V, W, emb, text_mask, idx, yval, xmean);
  return svpc_check_launch("gumbel_mean");
}

extern "C" int svpc_gumbel_bwd(int n_clips, int Lt, int V, int W, const float* P, const float* g, const int* clip_P_off,
                               const int* clip_vx, float tau, const float* emb, const float* text_mask, const int* idx,
                               const float* yval, const float* dxmean, float* dP, float* demb, float* tv_ws,
                               svpc_stream_t stream) {
  cudaStream_t st = (cudaStream_t)stream;
  cudaMemsetAsync(demb, 0, sizeof(float) * (size_t)V * W, st);
  if (n_clips <= 0) return SVPC_OK;
  SVPC_REQUIRE(tv_ws != nullptr, "gumbel_bwd: tv_ws (n_clips * V floats) is required");
  const bool rows = ptr_rows_route(Lt, n_clips, V);
  const size_t smem_tv = sizeof(float) * (TV_CLIPS * ((size_t)W + 1) + 4 + (size_t)TV_VPB * W);
  const bool tv_clips = rows && smem_tv <= 160 * 1024 && W % 4 == 0 && (((uintptr_t)emb) & 15) == 0;
  if (tv_clips) cudaFuncSetAttribute(gumbel_tv_clips_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem_tv);
  if (tv_clips)
    svpc_launch(gumbel_tv_clips_kernel, dim3(svpc_cdiv(V, TV_VPB), svpc_cdiv(n_clips, TV_CLIPS)), PG_THREADS, smem_tv, st, n_clips, V, W,
                emb, dxmean, tv_ws);
  else
    svpc_launch(gumbel_tv_kernel, dim3(svpc_cdiv(V, 64), n_clips), PG_THREADS, sizeof(float) * (size_t)W, st, V, W, emb, dxmean, tv_ws);
  int rc = svpc_check_launch("gumbel_tv");
  if (rc) return rc;
  const int n_rows = n_clips * Lt;
  if (rows)
    svpc_launch(gumbel_bwd_rows_kernel, svpc_cdiv(n_rows, PG_THREADS / 32), PG_THREADS, 0, st, n_rows, Lt, V, W, P, g, clip_P_off,
                clip_vx, tau, text_mask, idx, yval, dxmean, tv_ws, dP, demb);
  else
    svpc_launch(gumbel_bwd_kernel, dim3(Lt, n_clips), PG_THREADS, 0, st, Lt, V, W, P, g, clip_P_off, clip_vx, tau, text_mask, idx, yval,
                dxmean, tv_ws, dP, demb);
  return svpc_check_launch("gumbel_bwd");
}
