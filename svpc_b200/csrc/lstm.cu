// Recurrence of the recipe encoder (reference nn.LSTM(300 -> 768, bidirectional), model.py:865, 1022-1024; gate order
// i, f, g, o).  The input projections x W_ih^T + b of all clips are one batched GEMM per direction (svpc_gemm); what is
// left is sequential over a video's clips with a batch of only `videos` rows, i.e. a weight-streaming problem:
// every step reads W_hh (4D x D fp32 = 9.4 MB per direction at D = 768, L2-resident after the first step) once.
//
// One launch per time step covers both directions and all videos, and fuses the recurrent product with the cell:
//   forward   gates = xg[row] + h_prev W_hh^T ; (i, f, g, o) -> c, h ; out[row] += h           (lstm_step_fwd_kernel)
//   backward  dh = dout[row] + dgates[next row] W_hh ; cell backward -> dgates[row], dc        (lstm_step_bwd_kernel)
// A warp owns one hidden unit d: it streams the 4 gate rows of W_hh (forward) / row d of W_hh^T (backward) with
// coalesced float4 loads against the h (or dgates) rows of up to 16 videos staged in shared memory, reduces the
// 16 x 4 partial sums with a transposing shuffle tree, and lane 2b finishes video b.
#include "common.cuh"

#define LSTM_BV 16          // videos per CTA (smem rows)
#define LSTM_WARPS 8        // hidden units per CTA

// Transposing warp reduction: every lane holds N partial sums v[0..N); afterwards lane l holds the full sums of
// elements 2l*N/64.. (see callers).  N halves per round; static indexing keeps v[] in registers.
template <int N>
__device__ __forceinline__ void warp_reduce_transpose(float (&v)[N], int lane) {
  // rounds: (off, n) = (16, N/2), (8, N/4), ...
#pragma unroll
  for (int off = 16, n = N / 2; n >= 1 && off >= 1; off >>= 1, n >>= 1) {
    const bool up = (lane & off) != 0;
#pragma unroll
    for (int i = 0; i < N / 2; ++i) {
      if (i < n) {
        const float send = up ? v[i] : v[i + n];
        const float keep = up ? v[i + n] : v[i];
        v[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
      }
    }
  }
}

struct LstmStepParams {
  int B, D, Nc, Smax, j;
  const float* xg;      // (2, Nc, 4D)  x W_ih^T + b_ih + b_hh per direction
  const float* whh;     // fwd: (2, 4D, D) W_hh ; bwd: (2, D, 4D) W_hh^T
  const int* rows;      // (2, Smax, B) row of video b's clip at step j of direction dir, -1 when none
  float* hs;            // (2, Nc, D) h per clip row
  float* cs;            // (2, Nc, D) c per clip row
  float* gates;         // (2, Nc, 4D) activated gates per clip row
  float* out;           // (Nc, D)    sum of both directions' h (zero-initialised, atomically accumulated)
  float* hprev;         // (2, Nc, D) h entering each clip's step (operand of the W_hh weight gradient)
  // backward only
  const float* dout;    // (Nc, D)
  float* dgates;        // (2, Nc, 4D) gradient w.r.t. the gate pre-activations
  float* dc;            // (2, B, D)  carried cell gradient
};

__global__ void __launch_bounds__(LSTM_WARPS * 32) lstm_step_fwd_kernel(const LstmStepParams p) {
  svpc_pdl_sync();
  extern __shared__ __align__(16) float hsm[];          // [LSTM_BV][D]
  const int dir = blockIdx.y, b0 = blockIdx.z * LSTM_BV;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int D = p.D, j = p.j;
  const int* rows_j = p.rows + ((long long)dir * p.Smax + j) * p.B;
  const int* rows_p = rows_j - p.B;                      // step j-1 (valid when j > 0)
  const float* hs = p.hs + (long long)dir * p.Nc * D;
  if (j > 0) {
    for (int e = threadIdx.x * 4; e < LSTM_BV * D; e += blockDim.x * 4) {
      const int b = e / D, k = e - b * D;
      int rp = -1;
      if (b0 + b < p.B && rows_j[b0 + b] >= 0) rp = rows_p[b0 + b];
      float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
      if (rp >= 0) v = *reinterpret_cast<const float4*>(hs + (long long)rp * D + k);
      *reinterpret_cast<float4*>(hsm + e) = v;
    }
    __syncthreads();
  }
  const int d = blockIdx.x * LSTM_WARPS + warp;
  if (d >= D) return;
  float acc[LSTM_BV * 4];                                 // [video][gate]
#pragma unroll
  for (int i = 0; i < LSTM_BV * 4; ++i) acc[i] = 0.f;
  if (j > 0) {
    const float* w = p.whh + (long long)dir * 4 * D * D + (long long)d * D;
    for (int k = lane * 4; k < D; k += 128) {
      float4 wv[4];
#pragma unroll
      for (int g = 0; g < 4; ++g) wv[g] = __ldg(reinterpret_cast<const float4*>(w + (long long)g * D * D + k));
#pragma unroll
      for (int b = 0; b < LSTM_BV; ++b) {
        const float4 h = *reinterpret_cast<const float4*>(hsm + b * D + k);
#pragma unroll
        for (int g = 0; g < 4; ++g) {
          float a = acc[b * 4 + g];
          a = fmaf(h.x, wv[g].x, a); a = fmaf(h.y, wv[g].y, a); a = fmaf(h.z, wv[g].z, a); a = fmaf(h.w, wv[g].w, a);
          acc[b * 4 + g] = a;
        }
      }
    }
    warp_reduce_transpose<LSTM_BV * 4>(acc, lane);        // lane l: elements 2l, 2l+1 = video l/2, gates 2(l&1), +1
  }
  // exchange so that both lanes of a pair hold all four gates of video lane/2
  const float o0 = __shfl_xor_sync(0xffffffffu, acc[0], 1), o1 = __shfl_xor_sync(0xffffffffu, acc[1], 1);
  if (lane & 1) return;
  const int b = b0 + (lane >> 1);
  if (b >= p.B) return;
  const int r = rows_j[b];
  if (r < 0) return;
  const int rp = j > 0 ? rows_p[b] : -1;
  const float* x = p.xg + ((long long)dir * p.Nc + r) * 4 * D;
  const float gi = svpc_sigmoid(x[d] + acc[0]);
  const float gf = svpc_sigmoid(x[D + d] + acc[1]);
  const float gg = tanhf(x[2 * D + d] + o0);
  const float go = svpc_sigmoid(x[3 * D + d] + o1);
  float* cs = p.cs + (long long)dir * p.Nc * D;
  const float cp = rp >= 0 ? cs[(long long)rp * D + d] : 0.f;
  const float cn = gf * cp + gi * gg;
  const float hn = go * tanhf(cn);
  cs[(long long)r * D + d] = cn;
  p.hs[((long long)dir * p.Nc + r) * D + d] = hn;
  p.hprev[((long long)dir * p.Nc + r) * D + d] = rp >= 0 ? hs[(long long)rp * D + d] : 0.f;
  float* g = p.gates + ((long long)dir * p.Nc + r) * 4 * D;
  g[d] = gi; g[D + d] = gf; g[2 * D + d] = gg; g[3 * D + d] = go;
  atomicAdd(p.out + (long long)r * D + d, hn);
}

// Step j of the reverse sweep (called with j = Smax-1 ... 0).  rows of step j+1 give the gate gradients already computed.
__global__ void __launch_bounds__(LSTM_WARPS * 32) lstm_step_bwd_kernel(const LstmStepParams p) {
  svpc_pdl_sync();
  extern __shared__ __align__(16) float gsm[];          // [LSTM_BV][D] chunk of the next step's gate gradients
  const int dir = blockIdx.y, b0 = blockIdx.z * LSTM_BV;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int D = p.D, j = p.j;
  const int* rows_j = p.rows + ((long long)dir * p.Smax + j) * p.B;
  const int* rows_n = rows_j + p.B;                      // step j+1 (valid when j + 1 < Smax)
  const bool has_next = j + 1 < p.Smax;
  const int d = blockIdx.x * LSTM_WARPS + warp;          // d >= D warps still help staging
  float acc[LSTM_BV];
#pragma unroll
  for (int i = 0; i < LSTM_BV; ++i) acc[i] = 0.f;
  if (has_next) {
    const float* dg = p.dgates + (long long)dir * p.Nc * 4 * D;
    const float* wt = p.whh + (long long)dir * 4 * D * D + (long long)(d < D ? d : 0) * 4 * D;
    for (int c = 0; c < 4; ++c) {                        // chunks of D along the 4D reduction
      __syncthreads();
      for (int e = threadIdx.x * 4; e < LSTM_BV * D; e += blockDim.x * 4) {
        const int b = e / D, k = e - b * D;
        const int rn = (b0 + b < p.B) ? rows_n[b0 + b] : -1;
        float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
        if (rn >= 0) v = *reinterpret_cast<const float4*>(dg + (long long)rn * 4 * D + c * D + k);
        *reinterpret_cast<float4*>(gsm + e) = v;
      }
      __syncthreads();
      if (d < D) {
        for (int k = lane * 4; k < D; k += 128) {
          const float4 wv = __ldg(reinterpret_cast<const float4*>(wt + c * D + k));
#pragma unroll
          for (int b = 0; b < LSTM_BV; ++b) {
            const float4 g = *reinterpret_cast<const float4*>(gsm + b * D + k);
            float a = acc[b];
            a = fmaf(g.x, wv.x, a); a = fmaf(g.y, wv.y, a); a = fmaf(g.z, wv.z, a); a = fmaf(g.w, wv.w, a);
            acc[b] = a;
          }
        }
      }
    }
    warp_reduce_transpose<LSTM_BV>(acc, lane);            // 4 rounds: lane l holds half-sums of video l/2
    acc[0] += __shfl_xor_sync(0xffffffffu, acc[0], 1);
  }
  if (d >= D || (lane & 1)) return;
  const int b = b0 + (lane >> 1);
  if (b >= p.B) return;
  const int r = rows_j[b];
  if (r < 0) return;
  const int rp = j > 0 ? rows_j[b - p.B] : -1;            // rows of step j-1
  const float* g = p.gates + ((long long)dir * p.Nc + r) * 4 * D;
  const float* cs = p.cs + (long long)dir * p.Nc * D;
  const float gi = g[d], gf = g[D + d], gg = g[2 * D + d], go = g[3 * D + d];
  const float tc = tanhf(cs[(long long)r * D + d]);
  const float cp = rp >= 0 ? cs[(long long)rp * D + d] : 0.f;
  const float dh = p.dout[(long long)r * D + d] + acc[0];
  float* dcp = p.dc + ((long long)dir * p.B + b) * D + d;
  const float dct = *dcp + dh * go * (1.f - tc * tc);
  *dcp = dct * gf;
  float* da = p.dgates + ((long long)dir * p.Nc + r) * 4 * D;
  da[d] = dct * gg * gi * (1.f - gi);
  da[D + d] = dct * cp * gf * (1.f - gf);
  da[2 * D + d] = dct * gi * (1.f - gg * gg);
  da[3 * D + d] = dh * tc * go * (1.f - go);
}

static int lstm_launch(bool bwd, const LstmStepParams& p, cudaStream_t st) {
  SVPC_REQUIRE(p.D % 4 == 0, "lstm_step: hidden size %d must be a multiple of 4", p.D);
  if (p.B <= 0 || p.Nc <= 0) return SVPC_OK;
  const size_t smem = (size_t)LSTM_BV * p.D * sizeof(float);
  SVPC_REQUIRE(smem <= 200 * 1024, "lstm_step: hidden size %d too large", p.D);
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(lstm_step_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    cudaFuncSetAttribute(lstm_step_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    attr = true;
  }
  dim3 grid(svpc_cdiv(p.D, LSTM_WARPS), 2, svpc_cdiv(p.B, LSTM_BV));
  if (bwd) svpc_launch(lstm_step_bwd_kernel, grid, LSTM_WARPS * 32, smem, st, p);
  else svpc_launch(lstm_step_fwd_kernel, grid, LSTM_WARPS * 32, smem, st, p);
  return svpc_check_launch(bwd ? "lstm_step_bwd" : "lstm_step_fwd");
}

extern "C" int svpc_lstm_step_fwd(int B, int D, int Nc, int Smax, int j, const float* xg, const float* whh, const int* rows,
                                  float* hs, float* cs, float* gates, float* out, float* hprev, svpc_stream_t stream) {
  LstmStepParams p{B, D, Nc, Smax, j, xg, whh, rows, hs, cs, gates, out, hprev, nullptr, nullptr, nullptr};
  return lstm_launch(false, p, (cudaStream_t)stream);
}

extern "C" int svpc_lstm_step_bwd(int B, int D, int Nc, int Smax, int j, const float* whh_t, const int* rows,
                                  const float* cs, const float* gates, const float* dout, float* dgates, float* dc,
                                  svpc_stream_t stream) {
  LstmStepParams p{B, D, Nc, Smax, j, nullptr, whh_t, rows, nullptr, const_cast<float*>(cs), const_cast<float*>(gates),
                   nullptr, nullptr, dout, dgates, dc};
  return lstm_launch(true, p, (cudaStream_t)stream);
}

// ------------------------------------------------------------------------------------------------------------------
// Persistent recurrence: ONE launch for all Smax steps.  The grid is one CTA per SM, half of the CTAs per direction;
// CTA c of a direction owns hidden units [c D / CPD, (c+1) D / CPD) (10-11 units at D = 768 on 148 SMs) and keeps their
// slice of W_hh (forward: 4 gate rows per unit; backward: one row of W_hh^T per unit; 135 KB fp32) resident in shared
// memory for the whole sweep, so a step costs one barrier among the CTAs of a direction + the h (or gate-gradient) rows of
// the videos from L2, instead of a kernel launch and 9.4 MB of weights per direction.  The barrier is a monotonic
// counter per direction in global memory (`sync`, zeroed by the caller); all CTAs are co-resident (grid <= #SMs, 1 CTA/SM).
#define LSTM_PW 12          // warps per CTA = most hidden units a CTA can own
#define LSTM_BB 8           // videos per pass of the backward product (register tile LSTM_PW x LSTM_BB)
#define LSTM_SRN 64         // videos whose next-step rows the backward stages ahead of the barrier
#define LSTM_SEQ_EXTRA 2048 // floats of shared memory behind the staging area for small per-step tables (dynamic: the kernels use all of it)
// offset (floats) of the small tables: behind the resident weights and the larger of the forward / backward staging areas
__host__ __device__ __forceinline__ size_t lstm_seq_tail_off(int D) {
  const size_t fwd = (size_t)LSTM_BV * D, bwd = (size_t)2 * LSTM_PW * LSTM_PW * LSTM_BB;
  return (size_t)LSTM_PW * 4 * D + (fwd > bwd ? fwd : bwd);
}

struct LstmSeqParams {
  LstmStepParams s;
  unsigned int* sync;    // [2] arrival counters, one per direction (zeroed by the caller); [2] = error flag
  int cpd;               // CTAs per direction
};

__device__ __forceinline__ unsigned int ld_acquire_u32(const unsigned int* p) {
  unsigned int v;
  asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ void red_release_add(unsigned int* p, unsigned int v) {
  asm volatile("red.release.gpu.global.add.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
// wait until `target` arrivals were counted (bounded: a lost arrival must not hang the GPU - it raises the error flag)
__device__ __forceinline__ void dir_barrier_wait(unsigned int* cnt, unsigned int target, unsigned int* err) {
  if (threadIdx.x == 0) {
    unsigned int spins = 0;
    while (ld_acquire_u32(cnt) < target) {
      __nanosleep(32);
      if (++spins > (1u << 23)) {
        atomicExch(err, 1u);
        break;
      }
    }
  }
  __syncthreads();
}
__device__ __forceinline__ void dir_barrier_arrive(unsigned int* cnt) {
  __syncthreads();                                       // every thread's global writes of this step are issued
  if (threadIdx.x == 0) {
    __threadfence();
    red_release_add(cnt, 1u);
  }
}

__global__ void __launch_bounds__(LSTM_PW * 32, 1) lstm_seq_fwd_kernel(const LstmSeqParams q) {
  svpc_pdl_sync();
  const LstmStepParams& p = q.s;
  extern __shared__ __align__(16) float smem[];
  const int D = p.D;
  const int dir = blockIdx.x / q.cpd, c = blockIdx.x - dir * q.cpd;
  const int u0 = (int)((long long)c * D / q.cpd), u1 = (int)((long long)(c + 1) * D / q.cpd);
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* wsm = smem;                                     // [LSTM_PW][4][D]
  float* hsm = smem + (size_t)LSTM_PW * 4 * D;           // [LSTM_BV][D]
  // resident weights: gate rows g*D + d of W_hh[dir] for the owned units
  for (int e = threadIdx.x * 4; e < (u1 - u0) * 4 * D; e += blockDim.x * 4) {
    const int u = e / (4 * D), r = e - u * 4 * D, g = r / D, k = r - g * D;
    *reinterpret_cast<float4*>(wsm + e) =
        __ldg(reinterpret_cast<const float4*>(p.whh + ((long long)dir * 4 * D + (long long)g * D + u0 + u) * D + k));
  }
  const float* hs = p.hs + (long long)dir * p.Nc * D;
  float* cs = p.cs + (long long)dir * p.Nc * D;
  unsigned int* cnt = q.sync + dir;
  const int d = u0 + warp;
  const bool own = warp < u1 - u0;
  int* srp = reinterpret_cast<int*>(smem + lstm_seq_tail_off(D));   // row of the previous step per video of the current pass (-1: none)
  const int eb = lane >> 1;                               // video (within a pass) whose state this lane finishes
  const bool epi = own && !(lane & 1);
  const int n4 = LSTM_BV * D / 4;
  for (int j = 0; j < p.Smax; ++j) {
    const int* rows_j = p.rows + ((long long)dir * p.Smax + j) * p.B;
    const int* rows_p = rows_j - p.B;
    // everything that does not depend on the other CTAs is fetched BEFORE the barrier wait: the row indices of the first
    // pass and this lane's input-gate pre-activations and previous cell state
    if (threadIdx.x < LSTM_BV)
      srp[threadIdx.x] = (j > 0 && threadIdx.x < p.B && rows_j[threadIdx.x] >= 0) ? rows_p[threadIdx.x] : -1;
    int pr = -1, prp = -1;
    float px0 = 0.f, px1 = 0.f, px2 = 0.f, px3 = 0.f, pcp = 0.f;
    if (epi && eb < p.B) {
      pr = rows_j[eb];
      if (pr >= 0) {
        prp = j > 0 ? rows_p[eb] : -1;
        const float* x = p.xg + ((long long)dir * p.Nc + pr) * 4 * D;
        px0 = x[d]; px1 = x[D + d]; px2 = x[2 * D + d]; px3 = x[3 * D + d];
        if (prp >= 0) pcp = cs[(long long)prp * D + d];              // written by this very thread one step earlier
      }
    }
    if (j > 0) dir_barrier_wait(cnt, (unsigned int)(q.cpd * j), q.sync + 2);     // h of step j-1 is complete
    else __syncthreads();                                                        // weights staged
    for (int b0 = 0; b0 < p.B; b0 += LSTM_BV) {
      if (j > 0) {
        if (b0 > 0) {
          __syncthreads();
          if (threadIdx.x < LSTM_BV)
            srp[threadIdx.x] = (b0 + threadIdx.x < p.B && rows_j[b0 + threadIdx.x] >= 0) ? rows_p[b0 + threadIdx.x] : -1;
          __syncthreads();
        }
        // h rows of the pass: eight 16-byte loads per thread in flight (written by other SMs: L2 loads)
        for (int base = threadIdx.x; base < n4; base += 8 * (int)blockDim.x) {
          float4 v[8];
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int i = base + u * (int)blockDim.x;
            v[u] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (i < n4) {
              const int e = i * 4, b = e / D, k = e - b * D;
              const int rp = srp[b];
              if (rp >= 0) v[u] = __ldcg(reinterpret_cast<const float4*>(hs + (long long)rp * D + k));
            }
          }
#pragma unroll
          for (int u = 0; u < 8; ++u) {
            const int i = base + u * (int)blockDim.x;
            if (i < n4) *reinterpret_cast<float4*>(hsm + (size_t)i * 4) = v[u];
          }
        }
        __syncthreads();
      }
      if (!own) continue;
      float acc[LSTM_BV * 4];
#pragma unroll
      for (int i = 0; i < LSTM_BV * 4; ++i) acc[i] = 0.f;
      if (j > 0) {
        const float* w = wsm + (size_t)warp * 4 * D;
        for (int k = lane * 4; k < D; k += 128) {
          float4 wv[4];
#pragma unroll
          for (int g = 0; g < 4; ++g) wv[g] = *reinterpret_cast<const float4*>(w + g * D + k);
#pragma unroll
          for (int b = 0; b < LSTM_BV; ++b) {
            const float4 h = *reinterpret_cast<const float4*>(hsm + b * D + k);
#pragma unroll
            for (int g = 0; g < 4; ++g) {
              float a = acc[b * 4 + g];
              a = fmaf(h.x, wv[g].x, a); a = fmaf(h.y, wv[g].y, a); a = fmaf(h.z, wv[g].z, a); a = fmaf(h.w, wv[g].w, a);
              acc[b * 4 + g] = a;
            }
          }
        }
        warp_reduce_transpose<LSTM_BV * 4>(acc, lane);
      }
      const float o0 = __shfl_xor_sync(0xffffffffu, acc[0], 1), o1 = __shfl_xor_sync(0xffffffffu, acc[1], 1);
      const int b = b0 + eb;
      if ((lane & 1) || b >= p.B) continue;
      int r, rp;
      float x0, x1, x2, x3, cp;
      if (b0 == 0) {
        r = pr; rp = prp; x0 = px0; x1 = px1; x2 = px2; x3 = px3; cp = pcp;
        if (r < 0) continue;
      } else {
        r = rows_j[b];
        if (r < 0) continue;
        rp = j > 0 ? rows_p[b] : -1;
        const float* x = p.xg + ((long long)dir * p.Nc + r) * 4 * D;
        x0 = x[d]; x1 = x[D + d]; x2 = x[2 * D + d]; x3 = x[3 * D + d];
        cp = rp >= 0 ? cs[(long long)rp * D + d] : 0.f;
      }
      const float gi = svpc_sigmoid(x0 + acc[0]);
      const float gf = svpc_sigmoid(x1 + acc[1]);
      const float gg = tanhf(x2 + o0);
      const float go = svpc_sigmoid(x3 + o1);
      const float cn = gf * cp + gi * gg;
      const float hn = go * tanhf(cn);
      cs[(long long)r * D + d] = cn;
      p.hs[((long long)dir * p.Nc + r) * D + d] = hn;
      p.hprev[((long long)dir * p.Nc + r) * D + d] = rp >= 0 ? hsm[eb * D + d] : 0.f;
      float* g = p.gates + ((long long)dir * p.Nc + r) * 4 * D;
      g[d] = gi; g[D + d] = gf; g[2 * D + d] = gg; g[3 * D + d] = go;
      atomicAdd(p.out + (long long)r * D + d, hn);
    }
    if (j + 1 < p.Smax) dir_barrier_arrive(cnt);
  }
}

// Backward sweep.  Per step the CTA needs dh[b][u] = sum_e dgates[row_next(b)][e] * W_hh^T[u][e] over ALL e < 4D for its
// (up to LSTM_PW) hidden units u and the B videos: a (B x 4D) x (4D x LSTM_PW) product whose cost is operand delivery, not
// flops.  The contraction index e is split over the CTA's threads (thread t owns e = 4t, 4t + 4*blockDim, ...): every
// gate-gradient element is loaded from L2 exactly once per CTA, straight into registers (no staging, no barriers), every
// resident weight is read from shared memory once per pass of LSTM_BB videos, and each thread carries an LSTM_PW x LSTM_BB
// register tile of partial sums, reduced by a transposing warp reduction and one cross-warp pass through shared memory.
__global__ void __launch_bounds__(LSTM_PW * 32, 1) lstm_seq_bwd_kernel(const LstmSeqParams q) {
  svpc_pdl_sync();
  const LstmStepParams& p = q.s;
  extern __shared__ __align__(16) float smem[];
  constexpr int NACC = LSTM_PW * LSTM_BB;
  const int D = p.D, E4 = 4 * p.D;
  const int dir = blockIdx.x / q.cpd, c = blockIdx.x - dir * q.cpd;
  const int u0 = (int)((long long)c * D / q.cpd), u1 = (int)((long long)(c + 1) * D / q.cpd);
  const int nu = u1 - u0;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  float* wsm = smem;                                     // [LSTM_PW][4D]  rows u0.. of W_hh^T[dir], zero beyond nu
  float* red = smem + (size_t)LSTM_PW * E4;              // [2][LSTM_PW warps][NACC] partial sums
  int* srn = reinterpret_cast<int*>(smem + lstm_seq_tail_off(D));   // [LSTM_SRN] rows of step j+1
  float* sepi = smem + lstm_seq_tail_off(D) + LSTM_SRN;             // [2 passes][8 operands][NACC] epilogue operands
  static_assert(LSTM_SRN + 2 * 8 * NACC <= LSTM_SEQ_EXTRA, "small tables exceed their shared-memory window");
  for (int e = threadIdx.x * 4; e < LSTM_PW * E4; e += blockDim.x * 4) {
    float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
    if (e < nu * E4) v = __ldg(reinterpret_cast<const float4*>(p.whh + ((long long)dir * D + u0) * E4 + e));
    *reinterpret_cast<float4*>(wsm + e) = v;
  }
  const float* dg = p.dgates + (long long)dir * p.Nc * E4;
  const float* cs = p.cs + (long long)dir * p.Nc * D;
  unsigned int* cnt = q.sync + dir;
  for (int j = p.Smax - 1, it = 0; j >= 0; --j, ++it) {
    const int* rows_j = p.rows + ((long long)dir * p.Smax + j) * p.B;
    const int* rows_n = rows_j + p.B;
    const bool has_next = j + 1 < p.Smax;
    // operands that do not depend on the other CTAs are staged in shared memory BEFORE the barrier wait, by warps that do not
    // poll: the next-step rows of the first LSTM_SRN videos and the epilogue operands of the first two passes
    if (has_next && threadIdx.x >= 256 && threadIdx.x < 256 + LSTM_SRN) {
      const int b = threadIdx.x - 256;
      srn[b] = b < p.B ? rows_n[b] : -1;
    }
    if (threadIdx.x >= 128 && threadIdx.x < 128 + NACC) {
      const int t = threadIdx.x - 128, u = t / LSTM_BB, d = u0 + u;
#pragma unroll
      for (int pass = 0; pass < 2; ++pass) {
        const int b = pass * LSTM_BB + (t - u * LSTM_BB);
        float* e = sepi + (size_t)pass * 8 * NACC + t;
        int r = -1;
        if (u < nu && b < p.B) r = rows_j[b];
        e[0] = __int_as_float(r);
        if (r >= 0) {
          const int rp = j > 0 ? rows_j[b - p.B] : -1;
          const float* g = p.gates + ((long long)dir * p.Nc + r) * E4;
          e[1 * NACC] = g[d]; e[2 * NACC] = g[D + d]; e[3 * NACC] = g[2 * D + d]; e[4 * NACC] = g[3 * D + d];
          e[5 * NACC] = cs[(long long)r * D + d];
          e[6 * NACC] = rp >= 0 ? cs[(long long)rp * D + d] : 0.f;
          e[7 * NACC] = p.dout[(long long)r * D + d];
        }
      }
    }
    if (it > 0) dir_barrier_wait(cnt, (unsigned int)(q.cpd * it), q.sync + 2);   // gate gradients of step j+1 complete
    else __syncthreads();
    for (int b0 = 0, pass = 0; b0 < p.B; b0 += LSTM_BB, ++pass) {
      float* rbuf = red + (size_t)(pass & 1) * (LSTM_PW * NACC);
      if (has_next) {
        float acc[NACC];
#pragma unroll
        for (int i = 0; i < NACC; ++i) acc[i] = 0.f;
        const float* gp[LSTM_BB];
#pragma unroll
        for (int b = 0; b < LSTM_BB; ++b) {
          const int rn = (b0 + b < LSTM_SRN) ? srn[b0 + b] : (b0 + b < p.B) ? rows_n[b0 + b] : -1;
          gp[b] = rn >= 0 ? dg + (long long)rn * E4 : nullptr;
        }
        for (int e = threadIdx.x * 4; e < E4; e += blockDim.x * 4) {
          float4 g[LSTM_BB];
#pragma unroll
          for (int b = 0; b < LSTM_BB; ++b)
            g[b] = gp[b] ? __ldcg(reinterpret_cast<const float4*>(gp[b] + e)) : make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
          for (int u = 0; u < LSTM_PW; ++u) {
            const float4 wv = *reinterpret_cast<const float4*>(wsm + (size_t)u * E4 + e);
#pragma unroll
            for (int b = 0; b < LSTM_BB; ++b) {
              float a = acc[u * LSTM_BB + b];
              a = fmaf(g[b].x, wv.x, a); a = fmaf(g[b].y, wv.y, a); a = fmaf(g[b].z, wv.z, a); a = fmaf(g[b].w, wv.w, a);
              acc[u * LSTM_BB + b] = a;
            }
          }
        }
        warp_reduce_transpose<NACC>(acc, lane);            // lane l holds the warp's sums of elements 3l, 3l+1, 3l+2
        static_assert(NACC == 96, "three elements per lane after the transposing reduction");
        float* rw = rbuf + warp * NACC + 3 * lane;
        rw[0] = acc[0]; rw[1] = acc[1]; rw[2] = acc[2];
      }
      __syncthreads();
      const int t = threadIdx.x;
      const int u = t / LSTM_BB, b = b0 + (t - u * LSTM_BB);
      if (t >= NACC || u >= nu || b >= p.B) continue;
      const int d = u0 + u;
      int r;
      float gi, gf, gg, go, cr, cp, dob;
      if (pass < 2) {
        const float* e = sepi + (size_t)pass * 8 * NACC + t;
        r = __float_as_int(e[0]);
        if (r < 0) continue;
        gi = e[1 * NACC]; gf = e[2 * NACC]; gg = e[3 * NACC]; go = e[4 * NACC]; cr = e[5 * NACC]; cp = e[6 * NACC]; dob = e[7 * NACC];
      } else {
        r = rows_j[b];
        if (r < 0) continue;
        const int rp = j > 0 ? rows_j[b - p.B] : -1;
        const float* g = p.gates + ((long long)dir * p.Nc + r) * E4;
        gi = g[d]; gf = g[D + d]; gg = g[2 * D + d]; go = g[3 * D + d];
        cr = cs[(long long)r * D + d];
        cp = rp >= 0 ? cs[(long long)rp * D + d] : 0.f;
        dob = p.dout[(long long)r * D + d];
      }
      float sum = 0.f;
      if (has_next) {
#pragma unroll
        for (int w = 0; w < LSTM_PW; ++w) sum += rbuf[w * NACC + t];
      }
      const float tc = tanhf(cr);
      const float dh = dob + sum;
      float* dcp = p.dc + ((long long)dir * p.B + b) * D + d;           // owned by this thread for the whole sweep
      const float dct = *dcp + dh * go * (1.f - tc * tc);
      *dcp = dct * gf;
      float* da = p.dgates + ((long long)dir * p.Nc + r) * E4;
      da[d] = dct * gg * gi * (1.f - gi);
      da[D + d] = dct * cp * gf * (1.f - gf);
      da[2 * D + d] = dct * gi * (1.f - gg * gg);
      da[3 * D + d] = dh * tc * go * (1.f - go);
    }
    if (j > 0) dir_barrier_arrive(cnt);
  }
}

// SM count of the CURRENT device (cached per device ordinal)
static int lstm_device_sms() {
  static int sms[64] = {0};
  int dev = 0;
  cudaGetDevice(&dev);
  if (dev < 0 || dev >= 64) dev = 0;
  if (!sms[dev]) cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev);
  return sms[dev];
}

static size_t lstm_seq_smem(int D) {
  return (lstm_seq_tail_off(D) + LSTM_SEQ_EXTRA) * sizeof(float);   // resident weights; fwd: h staging, bwd: reduction buffers; small tables
}

static void lstm_seq_attrs() {
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(lstm_seq_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    cudaFuncSetAttribute(lstm_seq_bwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    attr = true;
  }
}

// The persistent sweep synchronises its CTAs through global counters, so the whole grid (2 x sms/2 CTAs of one per SM)
// must be co-resident: 1 when the hidden size fits the per-CTA slice, the shared memory fits, and the occupancy
// calculator gives both kernels at least one CTA per SM on this device; 0 -> the caller uses the per-step kernels.
extern "C" int svpc_lstm_seq_supported(int D) {
  const int sms = lstm_device_sms();
  const int cpd = sms / 2;
  if (D <= 0 || D % 4 != 0 || cpd < 1 || (D + cpd - 1) / cpd > LSTM_PW) return 0;
  const size_t smem = lstm_seq_smem(D);
  if (smem > 227 * 1024) return 0;
  lstm_seq_attrs();
  int nf = 0, nb = 0;
  if (cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nf, lstm_seq_fwd_kernel, LSTM_PW * 32, smem) != cudaSuccess ||
      cudaOccupancyMaxActiveBlocksPerMultiprocessor(&nb, lstm_seq_bwd_kernel, LSTM_PW * 32, smem) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return (nf >= 1 && nb >= 1) ? 1 : 0;
}

static int lstm_seq_launch(bool bwd, LstmSeqParams& q, cudaStream_t st) {
  const LstmStepParams& p = q.s;
  SVPC_REQUIRE(p.D % 4 == 0, "lstm_seq: hidden size %d must be a multiple of 4", p.D);
  SVPC_REQUIRE(q.sync != nullptr, "lstm_seq: needs the barrier counters");
  if (p.B <= 0 || p.Nc <= 0 || p.Smax <= 0) return SVPC_OK;
  q.cpd = lstm_device_sms() / 2;
  SVPC_REQUIRE(q.cpd >= 1 && (p.D + q.cpd - 1) / q.cpd <= LSTM_PW,
               "lstm_seq: %d hidden units over %d CTAs per direction exceed %d units per CTA", p.D, q.cpd, LSTM_PW);
  const size_t smem = lstm_seq_smem(p.D);
  SVPC_REQUIRE(smem <= 227 * 1024, "lstm_seq: hidden size %d needs %zu bytes of shared memory", p.D, smem);
  lstm_seq_attrs();
  // the two arrival counters start at zero for every sweep; sync[2] (barrier time-out) is STICKY: only the host clears it
  cudaError_t e = cudaMemsetAsync(q.sync, 0, 2 * sizeof(unsigned int), st);
  if (e != cudaSuccess) { svpc_set_error("lstm_seq: memset: %s", cudaGetErrorString(e)); return SVPC_ERR_CUDA; }
  if (bwd) svpc_launch(lstm_seq_bwd_kernel, 2 * q.cpd, LSTM_PW * 32, smem, st, q);
  else svpc_launch(lstm_seq_fwd_kernel, 2 * q.cpd, LSTM_PW * 32, smem, st, q);
  return svpc_check_launch(bwd ? "lstm_seq_bwd" : "lstm_seq_fwd");
}

extern "C" int svpc_lstm_seq_fwd(int B, int D, int Nc, int Smax, const float* xg, const float* whh, const int* rows, float* hs,
                                 float* cs, float* gates, float* out, float* hprev, unsigned int* sync,
                                 svpc_stream_t stream) {
  LstmSeqParams q{{B, D, Nc, Smax, 0, xg, whh, rows, hs, cs, gates, out, hprev, nullptr, nullptr, nullptr}, sync, 0};
  return lstm_seq_launch(false, q, (cudaStream_t)stream);
}

extern "C" int svpc_lstm_seq_bwd(int B, int D, int Nc, int Smax, const float* whh_t, const int* rows, const float* cs,
                                 const float* gates, const float* dout, float* dgates, float* dc, unsigned int* sync,
                                 svpc_stream_t stream) {
  LstmSeqParams q{{B, D, Nc, Smax, 0, nullptr, whh_t, rows, nullptr, const_cast<float*>(cs), const_cast<float*>(gates),
                   nullptr, nullptr, dout, dgates, dc},
                  sync, 0};
  return lstm_seq_launch(true, q, (cudaStream_t)stream);
}

// out[c, r] = in[r, c]  (fp32; W_hh^T for the reverse sweep, once per backward)
__global__ void transpose_f32_kernel(int R, int Cc, const float* __restrict__ in, float* __restrict__ out) {
  svpc_pdl_sync();
  __shared__ float t[32][33];
  const int c0 = blockIdx.x * 32, r0 = blockIdx.y * 32;
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int r = r0 + i, c = c0 + threadIdx.x;
    if (r < R && c < Cc) t[i][threadIdx.x] = in[(long long)r * Cc + c];
  }
  __syncthreads();
  for (int i = threadIdx.y; i < 32; i += blockDim.y) {
    const int c = c0 + i, r = r0 + threadIdx.x;
    if (r < R && c < Cc) out[(long long)c * R + r] = t[threadIdx.x][i];
  }
}

extern "C" int svpc_transpose_f32(int R, int Cc, const float* in, float* out, svpc_stream_t stream) {
  if (R <= 0 || Cc <= 0) return SVPC_OK;
  dim3 grid(svpc_cdiv(Cc, 32), svpc_cdiv(R, 32));
  svpc_launch(transpose_f32_kernel, grid, dim3(32, 8), 0, (cudaStream_t)stream, R, Cc, in, out);
  return svpc_check_launch("transpose_f32");
}
