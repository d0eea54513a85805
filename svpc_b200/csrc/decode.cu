// Greedy token selection of the caption decoder (reference src/translator.py:106-112): for every clip take the argmax of
// the current position's scores with the [UNK] column forced to -1e10; the extended id (possibly >= V: a copied OOV
// word) is recorded in the output sequence, and ids >= Vx - n_oov are fed back as [UNK] because the text embedding has
// no row for them.  One warp per clip, lowest index wins ties.
#include "common.cuh"

__global__ void greedy_pick_kernel(int rows, const float* __restrict__ scores, const int* __restrict__ row_off, long long ld,
                                   const int* __restrict__ row_width, int width, const int* __restrict__ row_noov, int unk,
                                   int* __restrict__ next_in, long long* __restrict__ ext_out, long long ld_ext, int pos) {
  svpc_pdl_sync();
  const int r = blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= rows) return;
  const float* s = scores + (row_off ? (long long)row_off[r] : (long long)r * ld);
  const int w = row_width ? row_width[r] : width;
  float best = -INFINITY;
  int bi = 0x7fffffff;
  for (int c = lane; c < w; c += 32) {
    const float v = c == unk ? -1e10f : s[c];
    if (v > best) { best = v; bi = c; }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const float ov = __shfl_xor_sync(0xffffffffu, best, o);
    const int oi = __shfl_xor_sync(0xffffffffu, bi, o);
    if (ov > best || (ov == best && oi < bi)) { best = ov; bi = oi; }
  }
  if (lane == 0) {
    const int noov = row_noov ? row_noov[r] : 0;
    ext_out[(long long)r * ld_ext + pos] = bi;
    next_in[r] = (bi >= w - noov) ? unk : bi;
  }
}

extern "C" int svpc_greedy_pick(int rows, const float* scores, const int* row_off, long long ld, const int* row_width,
                                int width, const int* row_noov, int unk, int* next_in, long long* ext_out, long long ld_ext,
                                int pos, svpc_stream_t stream) {
  if (rows <= 0) return SVPC_OK;
  svpc_launch(greedy_pick_kernel, svpc_cdiv(rows, 8), 256, 0, (cudaStream_t)stream, rows, scores, row_off, ld, row_width, width,
                                                                          row_noov, unk, next_in, ext_out, ld_ext, pos);
  return svpc_check_launch("greedy_pick");
}
