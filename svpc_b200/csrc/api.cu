// Error plumbing, launch counter and small utility kernels of libsvpc_b200.so.
#include "common.cuh"
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <atomic>

static thread_local char g_err[512] = "";
static std::atomic<long long> g_launches{0};

void svpc_set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int svpc_check_launch(const char* what) {
  g_launches.fetch_add(1, std::memory_order_relaxed);
  cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    svpc_set_error("%s: %s", what, cudaGetErrorString(e));
    return SVPC_ERR_CUDA;
  }
  return SVPC_OK;
}

extern "C" const char* svpc_last_error(void) { return g_err; }
extern "C" int svpc_abi_version(void) { return 1; }
extern "C" long long svpc_launch_count(void) { return g_launches.load(); }
extern "C" void svpc_reset_launch_count(void) { g_launches.store(0); }

// ------------------------------------------------------------------------------------------------
template <typename TS, typename TD>
__global__ void cast_kernel(long long n, const TS* __restrict__ s, TD* __restrict__ d) {
  svpc_pdl_sync();
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) stf(d + i, ldf(s + i));
}

extern "C" int svpc_cast(long long n, const void* src, int dts, void* dst, int dtd, svpc_stream_t stream) {
  if (n <= 0) return SVPC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int blocks = svpc_imin(svpc_cdiv(n, 256), SVPC_NUM_SMS * 16);
  SVPC_DISPATCH(dts, TS, SVPC_DISPATCH(dtd, TD, (svpc_launch(cast_kernel<TS, TD>, blocks, 256, 0, st, n, (const TS*)src, (TD*)dst))));
  return svpc_check_launch("cast");
}

// strided 2-D cast: dst[r, c] = src[r, c] with independent row pitches (padded bf16 GEMM operands: TMA wants row pitches
// that are multiples of 16 bytes, the model's word-vector size 300 and vocabulary 951 are not multiples of 8)
template <typename TS, typename TD>
__global__ void cast2d_kernel(int rows, int cols, const TS* __restrict__ s, long long lds, TD* __restrict__ d, long long ldd) {
  svpc_pdl_sync();
  const long long n = (long long)rows * cols;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long stride = (long long)gridDim.x * blockDim.x;
  for (; i < n; i += stride) {
    const long long r = i / cols, c = i - r * cols;
    stf(d + r * ldd + c, ldf(s + r * lds + c));
  }
}

extern "C" int svpc_cast2d(int rows, int cols, const void* src, int dts, long long lds, void* dst, int dtd, long long ldd,
                           svpc_stream_t stream) {
  if (rows <= 0 || cols <= 0) return SVPC_OK;
  cudaStream_t st = (cudaStream_t)stream;
  int blocks = svpc_imin(svpc_cdiv((long long)rows * cols, 256), SVPC_NUM_SMS * 16);
  SVPC_DISPATCH(dts, TS, SVPC_DISPATCH(dtd, TD, (svpc_launch(cast2d_kernel<TS, TD>, blocks, 256, 0, st, rows, cols, (const TS*)src, lds,
                                                                                             (TD*)dst, ldd))));
  return svpc_check_launch("cast2d");
}

__global__ void dropout_mask_kernel(long long n, const SvpcDrop drop, float* __restrict__ out) {
  svpc_pdl_sync();
  const bool dropping = drop.seed != nullptr && drop.p > 0.f;
  const uint32_t key = svpc_drop_key(drop);
  const float inv = dropping ? 1.f / (1.f - drop.p) : 1.f;
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < n; i += (long long)gridDim.x * blockDim.x)
    out[i] = dropping ? svpc_drop_scale(key, drop.p, inv, (unsigned long long)i) : 1.f;
}

extern "C" int svpc_dropout_mask(long long n, const unsigned int* drop_seed, unsigned int drop_site, float drop_p, float* out,
                                 svpc_stream_t stream) {
  if (n <= 0) return SVPC_OK;
  const SvpcDrop drop{drop_p > 0.f ? drop_seed : nullptr, drop_site, drop_p};
  svpc_launch(dropout_mask_kernel, svpc_imin(svpc_cdiv(n, 256), SVPC_NUM_SMS * 16), 256, 0, (cudaStream_t)stream, n, drop, out);
  return svpc_check_launch("dropout_mask");
}

// Gumbel(0, 1) noise for the hard Gumbel-softmax of the textual re-simulator (reference model.py:1018, F.gumbel_softmax:
// -log(Exp(1)) drawn on the GPU inside the step).  Counter-based like the dropout masks: element i of the step with device
// seed *seed is -log(-log(u)), u = (23 bits of hash(seed, site, i) + 0.5) / 2^23, so a captured graph draws fresh noise on
// every replay (the seed lives on the device) and nothing has to be uploaded.
__global__ void gumbel_noise_kernel(long long n, const SvpcDrop key_src, float* __restrict__ out) {
  svpc_pdl_sync();
  const uint32_t key = svpc_drop_key(key_src);
  long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  for (; i < n; i += (long long)gridDim.x * blockDim.x) {
    const uint32_t h = svpc_mix32((uint32_t)i * 0x9e3779b1u + (uint32_t)((unsigned long long)i >> 32) * 0x85ebca6bu + key);
    // 23 bits + 0.5: every value is exact in fp32 and strictly inside (0, 1) (with 24 bits the largest one rounds to 1.0 and
    // -log(-log(1)) is +inf: one element in 2^24, i.e. every few steps at 4.6 M elements per step)
    const float u = ((float)(h >> 9) + 0.5f) * (1.0f / 8388608.0f);
    out[i] = -logf(-logf(u));
  }
}

extern "C" int svpc_gumbel_noise(long long n, const unsigned int* seed, unsigned int site, float* out, svpc_stream_t stream) {
  if (n <= 0) return SVPC_OK;
  SVPC_REQUIRE(seed != nullptr && out != nullptr, "gumbel_noise: needs the device seed and the output");
  const SvpcDrop key_src{seed, site, 1.f};
  svpc_launch(gumbel_noise_kernel, svpc_imin(svpc_cdiv(n, 256), SVPC_NUM_SMS * 16), 256, 0, (cudaStream_t)stream, n, key_src, out);
  return svpc_check_launch("gumbel_noise");
}

// deterministic single-CTA sum (loss reductions: a few thousand elements)
__global__ void sum_kernel(long long n, const float* __restrict__ x, float* __restrict__ out) {
  svpc_pdl_sync();
  __shared__ float red[64];
  float acc = 0.f;
  for (long long i = threadIdx.x; i < n; i += blockDim.x) acc += x[i];
  acc = block_sum(acc, red);
  if (threadIdx.x == 0) *out = acc;
}

extern "C" int svpc_sum(long long n, const float* x, float* out, svpc_stream_t stream) {
  svpc_launch(sum_kernel, 1, 1024, 0, (cudaStream_t)stream, n, x, out);
  return svpc_check_launch("sum");
}

// Host -> device input upload that skips rows the path never reads: the loader's clip tensors are
// (B, 125, 3072) but only the first 100 rows of each clip hold video features (model.py:889 `[:, :max_v_len]`), so the
// upload is a pitched copy of `height` blocks of `width` bytes.  `src` is PINNED HOST memory for kind 1 (the one host
// pointer of this ABI), device memory for kind 3; nothing is synchronised.
extern "C" int svpc_memcpy2d_async(void* dst, long long dpitch, const void* src, long long spitch, long long width,
                                   long long height, int kind, svpc_stream_t stream) {
  if (width <= 0 || height <= 0) return SVPC_OK;
  if (kind != 1 && kind != 3) {
    svpc_set_error("memcpy2d: kind must be 1 (pinned host -> device) or 3 (device -> device)");
    return SVPC_ERR_ARG;
  }
  const cudaMemcpyKind k = kind == 1 ? cudaMemcpyHostToDevice : cudaMemcpyDeviceToDevice;
  // Diagnostic switch SVPC_H2D_ROWS=1: one 1-D copy per row instead of the single pitched call (measured on two B200 hosts:
  // no difference at 192 rows x 1.2 MB, e2e 6.517 against 6.511 ms/step).
  static int rows_mode = -1;
  if (rows_mode < 0) {
    const char* ev = getenv("SVPC_H2D_ROWS");
    rows_mode = (ev && ev[0] == '1') ? 1 : 0;
  }
  cudaError_t e = cudaSuccess;
  if (height == 1) {
    e = cudaMemcpyAsync(dst, src, (size_t)width, k, (cudaStream_t)stream);
  } else if (kind == 1 && rows_mode && width >= (1 << 18)) {
    for (long long r = 0; r < height && e == cudaSuccess; ++r)
      e = cudaMemcpyAsync((char*)dst + r * dpitch, (const char*)src + r * spitch, (size_t)width, k, (cudaStream_t)stream);
  } else {
    e = cudaMemcpy2DAsync(dst, (size_t)dpitch, src, (size_t)spitch, (size_t)width, (size_t)height, k, (cudaStream_t)stream);
  }
  if (e != cudaSuccess) {
    svpc_set_error("memcpy2d: %s", cudaGetErrorString(e));
    return SVPC_ERR_CUDA;
  }
  return SVPC_OK;
}

// ------------------------------------------------------------------------------------------------
// Refresh of the compute-dtype weight copies in ONE launch (the step used ~30 separate casts, each a few microseconds of
// work behind a launch): tensor t has numel[t] fp32 elements in rows of cols[t]; element i goes to
// dst[t][(i / cols[t]) * dst_ld[t] + i % cols[t]] (dst_ld > cols for row-padded TMA operands, slices of fused QKV
// buffers are plain offsets).  Chunk table as in optim.cu.
template <typename TD>
__global__ void __launch_bounds__(256) cast_multi_kernel(const int* __restrict__ chunk_tensor,
                                                         const long long* __restrict__ chunk_start, int chunk_elems,
                                                         const float* const* __restrict__ src, TD* const* __restrict__ dst,
                                                         const long long* __restrict__ numel, const int* __restrict__ cols,
                                                         const long long* __restrict__ dst_ld) {
  svpc_pdl_sync();
  const int c = blockIdx.x, t = chunk_tensor[c];
  const long long s0 = chunk_start[c];
  const long long rem = numel[t] - s0;
  const int n = (int)(rem < chunk_elems ? rem : chunk_elems);
  const float* sp = src[t] + s0;
  TD* dp = dst[t];
  const int cl = cols[t];
  const long long ld = dst_ld[t];
  if (ld == cl && (((uintptr_t)sp | (uintptr_t)(dp + s0)) & 15) == 0) {
    TD* d = dp + s0;
    const int n4 = n >> 2;
    for (int i = threadIdx.x; i < n4; i += 256) {
      const float4 v = __ldg(reinterpret_cast<const float4*>(sp) + i);
      stf(d + 4 * i, v.x); stf(d + 4 * i + 1, v.y); stf(d + 4 * i + 2, v.z); stf(d + 4 * i + 3, v.w);
    }
    for (int i = (n4 << 2) + threadIdx.x; i < n; i += 256) stf(d + i, sp[i]);
  } else {
    for (int i = threadIdx.x; i < n; i += 256) {
      const long long e = s0 + i;
      const long long r = e / cl;
      stf(dp + r * ld + (e - r * cl), sp[i]);
    }
  }
}

extern "C" int svpc_cast_multi(int n_chunks, const int* chunk_tensor, const long long* chunk_start, int chunk_elems,
                               const void* const* src, void* const* dst, const long long* numel, const int* cols,
                               const long long* dst_ld, int dt_dst, svpc_stream_t stream) {
  if (n_chunks <= 0) return SVPC_OK;
  if (chunk_elems <= 0) {
    svpc_set_error("cast_multi: chunk_elems must be positive");
    return SVPC_ERR_ARG;
  }
  cudaStream_t st = (cudaStream_t)stream;
  SVPC_DISPATCH(dt_dst, TD, (svpc_launch(cast_multi_kernel<TD>, n_chunks, 256, 0, st, chunk_tensor, chunk_start, chunk_elems,
                                         (const float* const*)src, (TD* const*)dst, numel, cols, dst_ld)));
  return svpc_check_launch("cast_multi");
}
