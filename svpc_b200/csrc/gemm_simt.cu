// fp32-accumulate FFMA GEMM with arbitrary element strides: the parity path (fp32 operands, true fp32 products like the
// reference's sgemm) and the route for shapes the tcgen05 kernel does not take (N of 1 or 3, K = 300, M = 16).
// D[M,N] = act(A[M,K] . B[N,K]^T + bias) + residual (+ D).
#include "common.cuh"

#define GS_BM 64
#define GS_BN 64
#define GS_BK 16
#define GS_THREADS 256

struct GemmSimtParams {
  int M, N, K;
  const void* A; long long sa_m, sa_k;
  const void* B; long long sb_n, sb_k;
  void* D; long long ldd;
  const float* bias; int act;
  const void* residual; long long ldr;
  void* pre; long long ldp;
  int accumulate;
  int k_per_split;         // grid.z slices of K (split-K: fp32 atomics into a zeroed / accumulated D)
  int atomic;
  SvpcDrop drop;           // dropout after the activation, before the residual
  const void* gref; long long ldg; int gact;   // last: D = (...) * gact'(gref[m, n])  (fused activation backward of a dgrad)
};

__device__ __forceinline__ float gs_act_grad(int act, float ref) {
  if (act == SVPC_ACT_RELU) return ref > 0.f ? 1.f : 0.f;
  if (act == SVPC_ACT_GELU) return svpc_gelu_grad(ref);
  if (act == SVPC_ACT_SIGMOID) return ref * (1.f - ref);
  return 1.f;
}

template <typename TA, typename TB, typename TD>
__global__ void __launch_bounds__(GS_THREADS) gemm_simt_kernel(const GemmSimtParams p) {
  svpc_pdl_sync();
  __shared__ float As[GS_BK][GS_BM + 4];
  __shared__ float Bs[GS_BK][GS_BN + 4];
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;
  const int m0 = blockIdx.y * GS_BM, n0 = blockIdx.x * GS_BN;
  const TA* __restrict__ A = (const TA*)p.A;
  const TB* __restrict__ B = (const TB*)p.B;
  const bool a_kfast = p.sa_k == 1, b_kfast = p.sb_k == 1;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  const int kbeg = blockIdx.z * p.k_per_split, kend = min(p.K, kbeg + p.k_per_split);
  for (int k0 = kbeg; k0 < kend; k0 += GS_BK) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const int e = tid + i * GS_THREADS;
      int mm, kk;
      if (a_kfast) { kk = e & 15; mm = e >> 4; } else { mm = e & 63; kk = e >> 6; }
      const int gm = m0 + mm, gk = k0 + kk;
      As[kk][mm] = (gm < p.M && gk < kend) ? ldf(A + (long long)gm * p.sa_m + (long long)gk * p.sa_k) : 0.f;
      int nn, kb;
      if (b_kfast) { kb = e & 15; nn = e >> 4; } else { nn = e & 63; kb = e >> 6; }
      const int gn = n0 + nn, gkb = k0 + kb;
      Bs[kb][nn] = (gn < p.N && gkb < kend) ? ldf(B + (long long)gn * p.sb_n + (long long)gkb * p.sb_k) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < GS_BK; ++kk) {
      const float4 a = *reinterpret_cast<const float4*>(&As[kk][ty * 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w};
      const float bv[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], bv[j], acc[i][j]);
    }
    __syncthreads();
  }

  const bool dropping = p.drop.seed != nullptr && p.drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(p.drop);
  const float dinv = dropping ? 1.f / (1.f - p.drop.p) : 1.f;
  TD* __restrict__ D = (TD*)p.D;
  const TD* __restrict__ R = (const TD*)p.residual;
  TD* __restrict__ Pre = (TD*)p.pre;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int m = m0 + ty * 4 + i;
    if (m >= p.M) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = n0 + tx * 4 + j;
      if (n >= p.N) continue;
      float v = acc[i][j];
      if (p.atomic) { atomicAdd(reinterpret_cast<float*>(D) + (long long)m * p.ldd + n, v); continue; }
      if (p.bias) v += p.bias[n];
      if (Pre) stf(Pre + (long long)m * p.ldp + n, v);
      v = svpc_act(v, p.act);
      if (dropping) v *= svpc_drop_scale(dkey, p.drop.p, dinv, (unsigned long long)m * p.N + n);
      if (R) v += ldf(R + (long long)m * p.ldr + n);
      TD* d = D + (long long)m * p.ldd + n;
      if (p.accumulate) v += ldf(d);
      if (p.gref) v *= gs_act_grad(p.gact, ldf((const TD*)p.gref + (long long)m * p.ldg + n));
      stf(d, v);
    }
  }
}

// shared epilogue of the thin kernels: one output element
template <typename TD>
__device__ __forceinline__ void gs_epilogue(const GemmSimtParams& p, int m, int n, float v, bool dropping, uint32_t dkey,
                                            float dinv) {
  TD* __restrict__ D = (TD*)p.D;
  const TD* __restrict__ R = (const TD*)p.residual;
  TD* __restrict__ Pre = (TD*)p.pre;
  if (p.bias) v += p.bias[n];
  if (Pre) stf(Pre + (long long)m * p.ldp + n, v);
  v = svpc_act(v, p.act);
  if (dropping) v *= svpc_drop_scale(dkey, p.drop.p, dinv, (unsigned long long)m * p.N + n);
  if (R) v += ldf(R + (long long)m * p.ldr + n);
  TD* d = D + (long long)m * p.ldd + n;
  if (p.accumulate) v += ldf(d);
  if (p.gref) v *= gs_act_grad(p.gact, ldf((const TD*)p.gref + (long long)m * p.ldg + n));
  stf(d, v);
}

// N <= 8 outputs per row, both operands K-major (the reasoner's 768 -> 3 and 300 -> 1 projections, model.py:804, 811):
// one warp per row of A, lanes along K, N accumulators, shuffle reduction.  The 64 x 64 tile kernel would put such a
// problem on 3 CTAs.
#define GS_ROWDOT_MAXN 8
template <typename TA, typename TB, typename TD>
__global__ void __launch_bounds__(256) gemm_rowdot_kernel(const GemmSimtParams p) {
  svpc_pdl_sync();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m = blockIdx.x * 8 + warp;
  if (m >= p.M) return;
  const TA* __restrict__ a = (const TA*)p.A + (long long)m * p.sa_m;
  const TB* __restrict__ B = (const TB*)p.B;
  float acc[GS_ROWDOT_MAXN];
#pragma unroll
  for (int n = 0; n < GS_ROWDOT_MAXN; ++n) acc[n] = 0.f;
  for (int k = lane; k < p.K; k += 32) {
    const float x = ldf(a + k);
#pragma unroll
    for (int n = 0; n < GS_ROWDOT_MAXN; ++n)
      if (n < p.N) acc[n] = fmaf(x, ldf(B + (long long)n * p.sb_n + k), acc[n]);
  }
  const bool dropping = p.drop.seed != nullptr && p.drop.p > 0.f;
  const uint32_t dkey = svpc_drop_key(p.drop);
  const float dinv = dropping ? 1.f / (1.f - p.drop.p) : 1.f;
#pragma unroll
  for (int n = 0; n < GS_ROWDOT_MAXN; ++n) {
    if (n >= p.N) break;
    const float v = warp_sum(acc[n]);
    if (lane == 0) gs_epilogue<TD>(p, m, n, v, dropping, dkey, dinv);
  }
}

// B MN-major (N contiguous) and a short loop: M <= 8 (weight gradients of those projections: few rows of dz^T against
// all of x) or K <= 8 (their input gradients).  One thread per output element, n fastest: B reads coalesce, A reads
// broadcast.
template <typename TA, typename TB, typename TD>
__global__ void __launch_bounds__(256) gemm_pointwise_kernel(const GemmSimtParams p) {
  svpc_pdl_sync();
  const long long e = (long long)blockIdx.x * 256 + threadIdx.x;
  if (e >= (long long)p.M * p.N) return;
  const int m = (int)(e / p.N), n = (int)(e - (long long)m * p.N);
  const TA* __restrict__ a = (const TA*)p.A + (long long)m * p.sa_m;
  const TB* __restrict__ b = (const TB*)p.B + n;
  // 16 independent load pairs in flight per round trip (the loop is latency-bound: few CTAs, one element per thread)
  float acc[4] = {0.f, 0.f, 0.f, 0.f};
  int k = 0;
  for (; k + 16 <= p.K; k += 16) {
    float av[16], bv[16];
#pragma unroll
    for (int u = 0; u < 16; ++u) {
      av[u] = ldf(a + (long long)(k + u) * p.sa_k);
      bv[u] = ldf(b + (long long)(k + u) * p.sb_k);
    }
#pragma unroll
    for (int u = 0; u < 16; ++u) acc[u & 3] = fmaf(av[u], bv[u], acc[u & 3]);
  }
  for (; k < p.K; ++k) acc[0] = fmaf(ldf(a + (long long)k * p.sa_k), ldf(b + (long long)k * p.sb_k), acc[0]);
  const float a0 = acc[0], a1 = acc[1], a2 = acc[2], a3 = acc[3];
  const bool dropping = p.drop.seed != nullptr && p.drop.p > 0.f;
  gs_epilogue<TD>(p, m, n, (a0 + a1) + (a2 + a3), dropping, svpc_drop_key(p.drop), dropping ? 1.f / (1.f - p.drop.p) : 1.f);
}

template <typename TA, typename TB>
static int launch_td(GemmSimtParams p, int dtD, cudaStream_t st) {
  cudaStream_t st_ = st;
  if (p.N <= GS_ROWDOT_MAXN && p.sa_k == 1 && p.sb_k == 1 && p.K >= 64) {
    SVPC_DISPATCH(dtD, TD, (svpc_launch(gemm_rowdot_kernel<TA, TB, TD>, svpc_cdiv(p.M, 8), 256, 0, st_, p)));
    return svpc_check_launch("gemm_rowdot");
  }
  if (p.sb_n == 1 && (p.M <= 8 || p.K <= 8) && p.K <= 4096) {
    SVPC_DISPATCH(dtD, TD, (svpc_launch(gemm_pointwise_kernel<TA, TB, TD>, svpc_cdiv((long long)p.M * p.N, 256), 256, 0, st_, p)));
    return svpc_check_launch("gemm_pointwise");
  }
  const int tiles = svpc_cdiv(p.N, GS_BN) * svpc_cdiv(p.M, GS_BM);
  int split = 1;
  const bool plain = !p.bias && p.act == SVPC_ACT_NONE && !p.residual && !p.pre && dtD == SVPC_F32 && !p.drop.seed && !p.gref;
  if (plain && tiles < SVPC_NUM_SMS / 2 && p.K >= 512) split = svpc_imin(SVPC_NUM_SMS / tiles, p.K / 128);
  if (split < 1) split = 1;
  p.k_per_split = svpc_cdiv(svpc_cdiv(p.K, split), GS_BK) * GS_BK;
  split = svpc_cdiv(p.K, p.k_per_split);
  p.atomic = split > 1;
  if (p.atomic && !p.accumulate) {
    cudaError_t e = cudaMemset2DAsync(p.D, (size_t)p.ldd * 4, 0, (size_t)p.N * 4, (size_t)p.M, st);
    if (e != cudaSuccess) { svpc_set_error("gemm_simt: memset2D: %s", cudaGetErrorString(e)); return SVPC_ERR_CUDA; }
  }
  dim3 grid(svpc_cdiv(p.N, GS_BN), svpc_cdiv(p.M, GS_BM), split);
  SVPC_DISPATCH(dtD, TD, (svpc_launch(gemm_simt_kernel<TA, TB, TD>, grid, GS_THREADS, 0, st, p)));
  return svpc_check_launch("gemm_simt");
}

int svpc_gemm_simt_launch(int M, int N, int K, const void* A, int dtA, long long sa_m, long long sa_k, const void* B,
                          int dtB, long long sb_n, long long sb_k, void* D, int dtD, long long ldd, const float* bias,
                          int act, const void* residual, long long ldr, void* pre, long long ldp, int accumulate,
                          SvpcDrop drop, cudaStream_t st, const void* gref, long long ldg, int gact) {
  GemmSimtParams p{M, N, K, A, sa_m, sa_k, B, sb_n, sb_k, D, ldd, bias, act, residual, ldr, pre, ldp, accumulate, K, 0, drop,
                   gref, ldg, gact};
  SVPC_DISPATCH(dtA, TA, SVPC_DISPATCH(dtB, TB, return (launch_td<TA, TB>(p, dtD, st))));
  return SVPC_OK;
}

// ------------------------------------------------------------------------------------------------
// public entry: route to the tcgen05 kernel when the problem is eligible, else to the FFMA kernel
// ------------------------------------------------------------------------------------------------
int svpc_gemm_tc_eligible(int M, int N, int K, int dtA, long long sa_m, long long sa_k, int dtB, long long sb_n,
                          long long sb_k, int dtD, long long ldd);
int svpc_gemm_tc_launch(int M, int N, int K, const void* A, long long sa_m, long long sa_k, const void* B, long long sb_n,
                        long long sb_k, void* D, int dtD, long long ldd, const float* bias, int act, const void* residual,
                        long long ldr, void* pre, long long ldp, int accumulate, SvpcDrop drop, cudaStream_t st,
                        const void* A_lo, const void* B_lo, const void* gref, long long ldg, int gact);

extern "C" int svpc_gemm_route(int M, int N, int K, int dtA, long long sa_m, long long sa_k, int dtB, long long sb_n,
                               long long sb_k, int dtD, long long ldd) {
  return svpc_gemm_tc_eligible(M, N, K, dtA, sa_m, sa_k, dtB, sb_n, sb_k, dtD, ldd) ? SVPC_GEMM_TC : SVPC_GEMM_SIMT;
}

extern "C" int svpc_gemm(int M, int N, int K, const void* A, int dtA, long long sa_m, long long sa_k, const void* B,
                         int dtB, long long sb_n, long long sb_k, void* D, int dtD, long long ldd, const float* bias,
                         int act, const void* residual, long long ldr, void* pre, long long ldp, int accumulate,
                         int route, const unsigned int* drop_seed, unsigned int drop_site, float drop_p,
                         const void* gref, long long ldg, int gact, svpc_stream_t stream) {
  SVPC_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm: negative size");
  if (M == 0 || N == 0) return SVPC_OK;
  SVPC_REQUIRE(A && B && D, "gemm: null operand");
  SVPC_REQUIRE(sa_m == 1 || sa_k == 1, "gemm: A needs a unit stride (sa_m=%lld sa_k=%lld)", sa_m, sa_k);
  SVPC_REQUIRE(sb_n == 1 || sb_k == 1, "gemm: B needs a unit stride (sb_n=%lld sb_k=%lld)", sb_n, sb_k);
  SVPC_REQUIRE(!(accumulate && (act != SVPC_ACT_NONE || residual)), "gemm: accumulate excludes act / residual");
  cudaStream_t st = (cudaStream_t)stream;
  SVPC_REQUIRE(drop_p >= 0.f && drop_p < 1.f, "gemm: dropout p=%f outside [0, 1)", drop_p);
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  const int ok = svpc_gemm_tc_eligible(M, N, K, dtA, sa_m, sa_k, dtB, sb_n, sb_k, dtD, ldd) &&
                 (((uintptr_t)A | (uintptr_t)B) & 15) == 0;   // TMA: 16-byte aligned global base
  if (route == SVPC_GEMM_TC && !ok) {
    svpc_set_error("gemm: problem %dx%dx%d not eligible for the tcgen05 route", M, N, K);
    return SVPC_ERR_UNSUPPORTED;
  }
  if (ok && route != SVPC_GEMM_SIMT)
    return svpc_gemm_tc_launch(M, N, K, A, sa_m, sa_k, B, sb_n, sb_k, D, dtD, ldd, bias, act, residual, ldr, pre, ldp,
                               accumulate, drop, st, nullptr, nullptr, gref, ldg, gact);
  return svpc_gemm_simt_launch(M, N, K, A, dtA, sa_m, sa_k, B, dtB, sb_n, sb_k, D, dtD, ldd, bias, act, residual, ldr,
                               pre, ldp, accumulate, drop, st, gref, ldg, gact);
}

// ------------------------------------------------------------------------------------------------
// fp32 operands on the tensor cores ("bf16 x 3"): x = hi + lo with hi = bf16(x), lo = bf16(x - hi) (|x - hi - lo| <= 2^-18 |x|);
//   D = A_hi.B_hi^T + A_lo.B_hi^T + A_hi.B_lo^T   accumulated in ONE fp32 TMEM accumulator by gemm_tc_kernel (nseg = 3);
// the dropped A_lo.B_lo^T term is O(2^-18) relative.  This is the parity (1e-4) path's route for every contraction that is
// eligible for the tcgen05 kernel; everything else of that path stays fp32 FFMA (gemm_simt_kernel).
// ------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) split_bf16_kernel(int rows, int cols, const float* __restrict__ src, long long lds,
                                                         bf16* __restrict__ hi, bf16* __restrict__ lo, long long ldd) {
  svpc_pdl_sync();
  const int c4 = cols >> 2;                       // groups of 4 columns (vector path needs lds % 4 == 0, ldd % 4 == 0)
  const bool vec = (lds % 4 == 0) && (ldd % 4 == 0) && (((uintptr_t)src | (uintptr_t)hi | (uintptr_t)lo) & 15) == 0;
  const long long stride = (long long)gridDim.x * blockDim.x;
  if (vec) {
    const long long n = (long long)rows * c4;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      const long long r = i / c4;
      const int c = (int)(i - r * c4) * 4;
      const float4 v = *reinterpret_cast<const float4*>(src + r * lds + c);
      const float x[4] = {v.x, v.y, v.z, v.w};
      __align__(8) bf16 h[4], l[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        h[e] = __float2bfloat16_rn(x[e]);
        l[e] = __float2bfloat16_rn(x[e] - __bfloat162float(h[e]));
      }
      *reinterpret_cast<uint2*>(hi + r * ldd + c) = *reinterpret_cast<const uint2*>(h);
      *reinterpret_cast<uint2*>(lo + r * ldd + c) = *reinterpret_cast<const uint2*>(l);
    }
  }
  const int c0 = vec ? c4 * 4 : 0;                // tail columns (or everything on the unaligned path)
  const int ct = cols - c0;
  if (ct > 0) {
    const long long n = (long long)rows * ct;
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += stride) {
      const long long r = i / ct;
      const int c = c0 + (int)(i - r * ct);
      const float x = src[r * lds + c];
      const bf16 h = __float2bfloat16_rn(x);
      hi[r * ldd + c] = h;
      lo[r * ldd + c] = __float2bfloat16_rn(x - __bfloat162float(h));
    }
  }
}

extern "C" int svpc_split_bf16(int rows, int cols, const float* src, long long lds, void* hi, void* lo, long long ldd,
                               svpc_stream_t stream) {
  if (rows <= 0 || cols <= 0) return SVPC_OK;
  SVPC_REQUIRE(src && hi && lo && lds >= cols && ldd >= cols, "split_bf16: bad arguments");
  const int blocks = svpc_imin(svpc_cdiv((long long)rows * cols, 1024), SVPC_NUM_SMS * 16);
  svpc_launch(split_bf16_kernel, blocks, 256, 0, (cudaStream_t)stream, rows, cols, src, lds, (bf16*)hi, (bf16*)lo, ldd);
  return svpc_check_launch("split_bf16");
}

extern "C" int svpc_gemm_split_eligible(int M, int N, int K, long long sa_m, long long sa_k, long long sb_n, long long sb_k) {
  return svpc_gemm_tc_eligible(M, N, K, SVPC_BF16, sa_m, sa_k, SVPC_BF16, sb_n, sb_k, SVPC_F32, 0);
}

extern "C" int svpc_gemm_split(int M, int N, int K, const void* A_hi, const void* A_lo, long long sa_m, long long sa_k,
                               const void* B_hi, const void* B_lo, long long sb_n, long long sb_k, void* D, int dtD,
                               long long ldd, const float* bias, int act, const void* residual, long long ldr, void* pre,
                               long long ldp, int accumulate, const unsigned int* drop_seed, unsigned int drop_site,
                               float drop_p, const void* gref, long long ldg, int gact, svpc_stream_t stream) {
  SVPC_REQUIRE(M >= 0 && N >= 0 && K >= 0, "gemm_split: negative size");
  if (M == 0 || N == 0) return SVPC_OK;
  SVPC_REQUIRE(A_hi && A_lo && B_hi && B_lo && D, "gemm_split: null operand");
  SVPC_REQUIRE(!(accumulate && (act != SVPC_ACT_NONE || residual)), "gemm_split: accumulate excludes act / residual");
  SVPC_REQUIRE(drop_p >= 0.f && drop_p < 1.f, "gemm_split: dropout p=%f outside [0, 1)", drop_p);
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  const bool ok = svpc_gemm_split_eligible(M, N, K, sa_m, sa_k, sb_n, sb_k) &&
                  (((uintptr_t)A_hi | (uintptr_t)A_lo | (uintptr_t)B_hi | (uintptr_t)B_lo) & 15) == 0;
  if (!ok) {
    svpc_set_error("gemm_split: problem %dx%dx%d / strides not eligible for the tcgen05 route", M, N, K);
    return SVPC_ERR_UNSUPPORTED;
  }
  return svpc_gemm_tc_launch(M, N, K, A_hi, sa_m, sa_k, B_hi, sb_n, sb_k, D, dtD, ldd, bias, act, residual, ldr, pre, ldp,
                             accumulate, drop, (cudaStream_t)stream, A_lo, B_lo, gref, ldg, gact);
}
