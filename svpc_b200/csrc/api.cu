// Error plumbing, launch counter and small utility kernels of libsvpc_b200.so.
#include "common.cuh"
#include <stdarg.h>
#include <stdio.h>
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
  cudaError_t e = height == 1 ? cudaMemcpyAsync(dst, src, (size_t)width, k, (cudaStream_t)stream)
                              : cudaMemcpy2DAsync(dst, (size_t)dpitch, src, (size_t)spitch, (size_t)width, (size_t)height, k,
                                                  (cudaStream_t)stream);
  if (e != cudaSuccess) {
    svpc_set_error("memcpy2d: %s", cudaGetErrorString(e));
    return SVPC_ERR_CUDA;
  }
  return SVPC_OK;
}
