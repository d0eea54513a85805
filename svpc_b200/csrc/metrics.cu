// Training-loop bookkeeping on the device: the reference reads ~6 scalars per video per step with
// .item() (src/train.py:32-49, 153-173: cal_performance, calculate_f1), each a stream synchronisation.  Here the counts
// of a whole step are accumulated into one device buffer by two launches and read once per epoch.
#include "common.cuh"

// acc[0] += #{rows with gold != ignore and argmax(scores row) == gold}, acc[1] += #{rows with gold != ignore}
// (cal_performance, train.py:32-38; n_word, train.py:156-157).  Row r of video b = scores[b] + r * width[b]; one warp per
// row; ties resolve to the lowest index (torch.max on equal values is unspecified; exact ties do not occur on real scores).
__global__ void __launch_bounds__(256) caption_accuracy_kernel(int n_videos, const float* const* __restrict__ scores,
                                                               const long long* const* __restrict__ gold,
                                                               const int* __restrict__ rows, const int* __restrict__ width,
                                                               const int* __restrict__ row_start, int total_rows,
                                                               long long ignore, double* __restrict__ acc) {
  svpc_pdl_sync();
  const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5, lane = threadIdx.x & 31;
  if (warp >= total_rows) return;
  int b = 0;
  while (b + 1 < n_videos && row_start[b + 1] <= warp) ++b;     // <= a few dozen videos
  const int r = warp - row_start[b], w = width[b];
  const long long g = gold[b][r];
  if (g == ignore) return;
  const float* s = scores[b] + (long long)r * w;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int j = lane; j < w; j += 32) {
    const float v = s[j];
    if (v > best || (v == best && j < bi)) { best = v; bi = j; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) {
    atomicAdd(acc + 1, 1.0);
    if ((long long)bi == g) atomicAdd(acc, 1.0);
  }
}

// calculate_f1 (train.py:40-49) over n elements: acc[0] += sum gold[p > 0.5], acc[1] += sum gold, acc[2] += #{p > 0.5}
__global__ void __launch_bounds__(256) f1_counts_kernel(long long n, const float* __restrict__ p, const float* __restrict__ y,
                                                        double* __restrict__ acc) {
  svpc_pdl_sync();
  __shared__ float red[96];
  float c = 0.f, r = 0.f, q = 0.f;
  for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
    const bool pos = p[i] > 0.5f;
    const float g = y[i];
    c += pos ? g : 0.f;
    r += g;
    q += pos ? 1.f : 0.f;
  }
  c = block_sum(c, red);
  r = block_sum(r, red + 32);
  q = block_sum(q, red + 64);
  if (threadIdx.x == 0) {
    atomicAdd(acc, (double)c);
    atomicAdd(acc + 1, (double)r);
    atomicAdd(acc + 2, (double)q);
  }
}

extern "C" int svpc_caption_accuracy(int n_videos, const void* const* scores, const void* const* gold, const int* rows,
                                     const int* width, const int* row_start, int total_rows, long long ignore, double* acc,
                                     svpc_stream_t stream) {
  if (n_videos <= 0 || total_rows <= 0) return SVPC_OK;
  svpc_launch(caption_accuracy_kernel, svpc_cdiv((long long)total_rows * 32, 256), 256, 0, (cudaStream_t)stream, n_videos,
              (const float* const*)scores, (const long long* const*)gold, rows, width, row_start, total_rows, ignore, acc);
  return svpc_check_launch("caption_accuracy");
}

extern "C" int svpc_f1_counts(long long n, const float* p, const float* y, double* acc, svpc_stream_t stream) {
  if (n <= 0) return SVPC_OK;
  svpc_launch(f1_counts_kernel, svpc_imin(svpc_cdiv(n, 256), SVPC_NUM_SMS * 4), 256, 0, (cudaStream_t)stream, n, p, y, acc);
  return svpc_check_launch("f1_counts");
}
