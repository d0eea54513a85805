// The two multi-label losses taken on probabilities (the LSTM recurrence lives in lstm.cu): nn.BCELoss(sum) (model.py:871) and AsymmetricLoss
// (libs/ASL/src/loss_functions/losses.py:15-50).
#include "common.cuh"

// ------------------------------------------------------------------------------------------------
// nn.BCELoss(reduction='sum'): -(y * max(log p, -100) + (1 - y) * max(log(1 - p), -100)); the clamp and the
// sigmoid-saturation behaviour of the reference are kept on purpose.
// backward as ATen: g * (p - y) / max((1 - p) * p, 1e-12)
// ------------------------------------------------------------------------------------------------
__global__ void bce_fwd_kernel(long long n, const float* __restrict__ p, const float* __restrict__ y, float* __restrict__ loss) {
  svpc_pdl_sync();
  __shared__ float red[64];
  float acc = 0.f;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) {
    const float lp = fmaxf(logf(p[i]), -100.f), lq = fmaxf(logf(1.f - p[i]), -100.f);
    acc -= y[i] * lp + (1.f - y[i]) * lq;
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) *loss = acc;
}
__global__ void bce_bwd_kernel(long long n, const float* __restrict__ p, const float* __restrict__ y,
                               const float* __restrict__ g, float* __restrict__ dp) {
  svpc_pdl_sync();
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dp[i] = g[0] * (p[i] - y[i]) / fmaxf((1.f - p[i]) * p[i], 1e-12f);
}
extern "C" int svpc_bce_fwd(long long n, const float* p, const float* y, float* loss, svpc_stream_t stream) {
  svpc_launch(bce_fwd_kernel, 1, 1024, 0, (cudaStream_t)stream, n, p, y, loss);
  return svpc_check_launch("bce_fwd");
}
extern "C" int svpc_bce_bwd(long long n, const float* p, const float* y, const float* g, float* dp, svpc_stream_t stream) {
  if (n <= 0) return SVPC_OK;
  svpc_launch(bce_bwd_kernel, svpc_cdiv(n, 256), 256, 0, (cudaStream_t)stream, n, p, y, g, dp);
  return svpc_check_launch("bce_bwd");
}

// ------------------------------------------------------------------------------------------------
// AsymmetricLoss on probabilities, gamma_neg 4, gamma_pos 1, clip 0.05, eps 1e-8, over rows with >= 1 positive.
// ------------------------------------------------------------------------------------------------
#define ASL_GN 4.0f
#define ASL_GP 1.0f
#define ASL_CLIP 0.05f
#define ASL_EPS 1e-8f

__device__ __forceinline__ float asl_elem(float p, float y, float* dldp) {
  const float xs_raw = 1.f - p + ASL_CLIP;
  const float xs = fminf(xs_raw, 1.f);
  const float dxs = xs_raw <= 1.f ? -1.f : 0.f;                 // clamp(max=1) passes gradient where xs_raw <= 1 (ATen: <=)
  const float lp = logf(fmaxf(p, ASL_EPS)), ln = logf(fmaxf(xs, ASL_EPS));
  const float l = y * lp + (1.f - y) * ln;
  const float pt = p * y + xs * (1.f - y);
  const float gam = ASL_GP * y + ASL_GN * (1.f - y);
  const float base = 1.f - pt;
  const float w = powf(base, gam);
  if (dldp) {
    const float dl = y * (p >= ASL_EPS ? 1.f / p : 0.f) + (1.f - y) * (xs >= ASL_EPS ? dxs / xs : 0.f);
    const float dpt = y + (1.f - y) * dxs;
    const float dw = gam * powf(base, gam - 1.f) * (-dpt);
    *dldp = -(dl * w + l * dw);
  }
  return -(l * w);
}

__global__ void asl_rows_kernel(int rows, int A, const float* __restrict__ y, int* __restrict__ active) {
  svpc_pdl_sync();
  const int r = blockIdx.x;
  __shared__ int any;
  if (threadIdx.x == 0) any = 0;
  __syncthreads();
  int a = 0;
  for (int c = threadIdx.x; c < A; c += blockDim.x) a |= (y[(long long)r * A + c] == 1.0f);
  if (a) any = 1;
  __syncthreads();
  if (threadIdx.x == 0) active[r] = any;
}

// gamma is 1 on positives and 4 on negatives for 0/1 targets (the only targets the path produces); other y take powf
__device__ __forceinline__ float asl_elem_fwd(float p, float y) {
  if (y != 0.f && y != 1.f) return asl_elem(p, y, nullptr);
  const float xs = fminf(1.f - p + ASL_CLIP, 1.f);
  if (y == 1.f) return -(logf(fmaxf(p, ASL_EPS)) * (1.f - p));
  const float b = 1.f - xs, b2 = b * b;
  return -(logf(fmaxf(xs, ASL_EPS)) * (b2 * b2));
}

// One thread-block cluster of ASL_CTAS CTAs: rows are dealt round-robin to the cluster's warps, every CTA reduces its share,
// CTA 0 adds the partial sums through distributed shared memory in rank order (deterministic, no scratch buffer, one launch).
#define ASL_CTAS 8
#define ASL_THREADS 512
__global__ void __cluster_dims__(ASL_CTAS, 1, 1) __launch_bounds__(ASL_THREADS)
    asl_fwd_kernel(int rows, int A, const float* __restrict__ p, const float* __restrict__ y, float* __restrict__ loss) {
  svpc_pdl_sync();
  __shared__ float red[64];
  __shared__ float part;
  uint32_t rank;
  asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(rank));
  float acc = 0.f;
  const int lane = threadIdx.x & 31, nw = ASL_THREADS >> 5;
  const int gw = rank * nw + (threadIdx.x >> 5);
  for (int r = gw; r < rows; r += nw * ASL_CTAS) {
    const float* yr = y + (long long)r * A;
    const float* pr = p + (long long)r * A;
    float rowacc = 0.f;
    int a = 0;
    for (int c = lane; c < A; c += 32) {
      const float yv = yr[c];
      a |= (yv == 1.0f);
      rowacc += asl_elem_fwd(pr[c], yv);
    }
    if (__any_sync(0xffffffffu, a)) acc += rowacc;       // rows without a positive are dropped (model.py:1103-1104)
  }
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) part = acc;
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
  if (rank == 0 && threadIdx.x == 0) {
    float tot = 0.f;
    const uint32_t local = (uint32_t)__cvta_generic_to_shared(&part);
    for (uint32_t k = 0; k < ASL_CTAS; ++k) {
      uint32_t remote;
      float v;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(local), "r"(k));
      asm volatile("ld.shared::cluster.f32 %0, [%1];" : "=f"(v) : "r"(remote) : "memory");
      tot += v;
    }
    *loss = tot;
  }
  asm volatile("barrier.cluster.arrive.release.aligned;\nbarrier.cluster.wait.acquire.aligned;" ::: "memory");
}

__global__ void asl_bwd_kernel(int rows, int A, const float* __restrict__ p, const float* __restrict__ y,
                               const float* __restrict__ g, float* __restrict__ dp) {
  svpc_pdl_sync();
  const int r = blockIdx.x;
  __shared__ int any;
  if (threadIdx.x == 0) any = 0;
  __syncthreads();
  int a = 0;
  for (int c = threadIdx.x; c < A; c += blockDim.x) a |= (y[(long long)r * A + c] == 1.0f);
  if (a) any = 1;
  __syncthreads();
  const int act = any;
  for (int c = threadIdx.x; c < A; c += blockDim.x) {
    float d = 0.f;
    if (act) asl_elem(p[(long long)r * A + c], y[(long long)r * A + c], &d);
    dp[(long long)r * A + c] = act ? g[0] * d : 0.f;
  }
}

extern "C" int svpc_asl_fwd(int rows, int A, const float* p, const float* y, float* loss, svpc_stream_t stream) {
  svpc_launch(asl_fwd_kernel, ASL_CTAS, ASL_THREADS, 0, (cudaStream_t)stream, rows, A, p, y, loss);
  return svpc_check_launch("asl_fwd");
}
extern "C" int svpc_asl_bwd(int rows, int A, const float* p, const float* y, const float* g, float* dp, svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  svpc_launch(asl_bwd_kernel, rows, 128, 0, (cudaStream_t)stream, rows, A, p, y, g, dp);
  return svpc_check_launch("asl_bwd");
}
