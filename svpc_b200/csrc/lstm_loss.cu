// LSTM cell of the recipe encoder (reference nn.LSTM, model.py:865, 1022-1024; gate order i, f, g, o) and the two
// multi-label losses taken on probabilities: nn.BCELoss(sum) (model.py:871) and AsymmetricLoss
// (libs/ASL/src/loss_functions/losses.py:15-50).
#include "common.cuh"

// One time step for all videos.  rows[b] = row of video b's current clip in the (Nc, .) arrays, or -1 when the video
// has no clip at this step.  xg = x W_ih^T + b_ih + b_hh for every clip (batched GEMM), hg = h W_hh^T for this step.
__global__ void lstm_cell_fwd_kernel(int B, int D, const float* __restrict__ xg, const float* __restrict__ hg,
                                     const int* __restrict__ rows, float* __restrict__ c, float* __restrict__ h,
                                     float* __restrict__ out, int accumulate_out, float* __restrict__ gates_saved,
                                     float* __restrict__ c_saved, float* __restrict__ hprev_saved,
                                     float* __restrict__ cprev_saved) {
  const int b = blockIdx.y;
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= D) return;
  const int r = rows[b];
  if (r < 0) return;
  const float* x = xg + (long long)r * 4 * D;
  const float* hh = hg + (long long)b * 4 * D;
  const float gi = svpc_sigmoid(x[d] + hh[d]);
  const float gf = svpc_sigmoid(x[D + d] + hh[D + d]);
  const float gg = tanhf(x[2 * D + d] + hh[2 * D + d]);
  const float go = svpc_sigmoid(x[3 * D + d] + hh[3 * D + d]);
  const float cp = c[(long long)b * D + d], hp = h[(long long)b * D + d];
  const float cn = gf * cp + gi * gg;
  const float hn = go * tanhf(cn);
  c[(long long)b * D + d] = cn;
  h[(long long)b * D + d] = hn;
  float* g = gates_saved + (long long)r * 4 * D;
  g[d] = gi; g[D + d] = gf; g[2 * D + d] = gg; g[3 * D + d] = go;
  c_saved[(long long)r * D + d] = cn;
  hprev_saved[(long long)r * D + d] = hp;
  cprev_saved[(long long)r * D + d] = cp;
  float* o = out + (long long)r * D + d;
  *o = accumulate_out ? *o + hn : hn;
}

// Reverse of one time step.  dh_rec (B, D) = gradient w.r.t. h of this step coming from the later step (through W_hh),
// dc (B, D) carried cell gradient (in/out).  Writes the pre-activation gate gradients both at the clip's row
// (dgates_all, for the batched weight / input gradients) and per video (dgates_step, for the recurrent GEMM).
__global__ void lstm_cell_bwd_kernel(int B, int D, const int* __restrict__ rows, const float* __restrict__ dout,
                                     const float* __restrict__ dh_rec, float* __restrict__ dc,
                                     const float* __restrict__ gates_saved, const float* __restrict__ c_saved,
                                     const float* __restrict__ cprev_saved, float* __restrict__ dgates_all,
                                     float* __restrict__ dgates_step) {
  const int b = blockIdx.y;
  const int d = blockIdx.x * blockDim.x + threadIdx.x;
  if (d >= D) return;
  const int r = rows[b];
  float* ds = dgates_step + (long long)b * 4 * D;
  if (r < 0) { ds[d] = 0.f; ds[D + d] = 0.f; ds[2 * D + d] = 0.f; ds[3 * D + d] = 0.f; return; }
  const float* g = gates_saved + (long long)r * 4 * D;
  const float gi = g[d], gf = g[D + d], gg = g[2 * D + d], go = g[3 * D + d];
  const float tc = tanhf(c_saved[(long long)r * D + d]);
  const float dh = dout[(long long)r * D + d] + dh_rec[(long long)b * D + d];
  const float dct = dc[(long long)b * D + d] + dh * go * (1.f - tc * tc);
  const float di = dct * gg * gi * (1.f - gi);
  const float df = dct * cprev_saved[(long long)r * D + d] * gf * (1.f - gf);
  const float dg = dct * gi * (1.f - gg * gg);
  const float dO = dh * tc * go * (1.f - go);
  dc[(long long)b * D + d] = dct * gf;
  float* da = dgates_all + (long long)r * 4 * D;
  da[d] = di; da[D + d] = df; da[2 * D + d] = dg; da[3 * D + d] = dO;
  ds[d] = di; ds[D + d] = df; ds[2 * D + d] = dg; ds[3 * D + d] = dO;
}

extern "C" int svpc_lstm_cell_fwd(int B, int D, const float* xg, const float* hg, const int* rows, float* c, float* h,
                                  float* out, int accumulate_out, float* gates_saved, float* c_saved, float* hprev_saved,
                                  float* cprev_saved, svpc_stream_t stream) {
  if (B <= 0) return SVPC_OK;
  dim3 grid(svpc_cdiv(D, 128), B);
  lstm_cell_fwd_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(B, D, xg, hg, rows, c, h, out, accumulate_out, gates_saved,
                                                              c_saved, hprev_saved, cprev_saved);
  return svpc_check_launch("lstm_cell_fwd");
}

extern "C" int svpc_lstm_cell_bwd(int B, int D, const int* rows, const float* dout, const float* dh_rec, float* dc,
                                  const float* gates_saved, const float* c_saved, const float* cprev_saved,
                                  float* dgates_all, float* dgates_step, svpc_stream_t stream) {
  if (B <= 0) return SVPC_OK;
  dim3 grid(svpc_cdiv(D, 128), B);
  lstm_cell_bwd_kernel<<<grid, 128, 0, (cudaStream_t)stream>>>(B, D, rows, dout, dh_rec, dc, gates_saved, c_saved, cprev_saved,
                                                              dgates_all, dgates_step);
  return svpc_check_launch("lstm_cell_bwd");
}

// ------------------------------------------------------------------------------------------------
// nn.BCELoss(reduction='sum'): -(y * max(log p, -100) + (1 - y) * max(log(1 - p), -100)); the clamp and the
// sigmoid-saturation behaviour of the reference are kept on purpose.
// backward as ATen: g * (p - y) / max((1 - p) * p, 1e-12)
// ------------------------------------------------------------------------------------------------
__global__ void bce_fwd_kernel(long long n, const float* __restrict__ p, const float* __restrict__ y, float* __restrict__ loss) {
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
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dp[i] = g[0] * (p[i] - y[i]) / fmaxf((1.f - p[i]) * p[i], 1e-12f);
}
extern "C" int svpc_bce_fwd(long long n, const float* p, const float* y, float* loss, svpc_stream_t stream) {
  bce_fwd_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(n, p, y, loss);
  return svpc_check_launch("bce_fwd");
}
extern "C" int svpc_bce_bwd(long long n, const float* p, const float* y, const float* g, float* dp, svpc_stream_t stream) {
  if (n <= 0) return SVPC_OK;
  bce_bwd_kernel<<<svpc_cdiv(n, 256), 256, 0, (cudaStream_t)stream>>>(n, p, y, g, dp);
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

__global__ void asl_fwd_kernel(int rows, int A, const float* __restrict__ p, const float* __restrict__ y,
                               float* __restrict__ loss) {
  __shared__ float red[64];
  __shared__ int any_s[32];
  float acc = 0.f;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nw = blockDim.x >> 5;
  for (int r = warp; r < rows; r += nw) {
    int a = 0;
    for (int c = lane; c < A; c += 32) a |= (y[(long long)r * A + c] == 1.0f);
    a = __any_sync(0xffffffffu, a);
    if (!a) continue;
    for (int c = lane; c < A; c += 32) acc += asl_elem(p[(long long)r * A + c], y[(long long)r * A + c], nullptr);
  }
  (void)any_s;
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) *loss = acc;
}

__global__ void asl_bwd_kernel(int rows, int A, const float* __restrict__ p, const float* __restrict__ y,
                               const float* __restrict__ g, float* __restrict__ dp) {
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
  asl_fwd_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>(rows, A, p, y, loss);
  return svpc_check_launch("asl_fwd");
}
extern "C" int svpc_asl_bwd(int rows, int A, const float* p, const float* y, const float* g, float* dp, svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  asl_bwd_kernel<<<rows, 128, 0, (cudaStream_t)stream>>>(rows, A, p, y, g, dp);
  return svpc_check_launch("asl_bwd");
}
