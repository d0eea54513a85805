// Tensor-core version of the fused short-sequence attention (reference BertSelfAttention core, model.py:204-216) for
// the bf16 path: head dim 64, sequences of up to 128 tokens (clip encoder L = 100, caption decoder L = 25).
//
// One CTA per (sequence, head); warp w owns query rows [16w, 16w+16).  Q, K, V (and dO in the backward) of the head
// are staged once in shared memory as bf16 with a 144-byte row pitch (conflict-free ldmatrix); S = QK^T, the softmax
// and P never leave the registers in the forward (flash-attention style, the whole key range fits one pass).  The
// products run on the warp-level bf16 MMA (mma.sync.m16n8k16, fp32 accumulate): the tiles are 16 x <=128 x 64 per
// warp and the softmax sits between the two products in registers, so there is nothing for a 128-row tcgen05 tile /
// TMEM round trip to win here; the large contractions of the step (gemm_tc.cu) are the tcgen05 ones.
// Semantics kept from the reference: scale 1/sqrt(dh) BEFORE the additive mask (1-m)*-10000 (not -inf), fp32 softmax.
#include "common.cuh"

#define ATC_DH 64
#ifndef ATC_FWD_BLOCKS
#define ATC_FWD_BLOCKS 4   // CTAs per SM the L <= 112 forward is compiled for (measured at L = 100: 2 -> 83 us without spills, 3 -> 75 us, 4 -> 73 us: occupancy wins over the spills)
#endif
#define ATC_PITCH 72   // bf16 elements per smem row (64 + 8 pad)

__device__ __forceinline__ uint32_t atc_smem(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void ldsm4t(uint32_t (&r)[4], uint32_t addr) {
  asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0,%1,%2,%3}, [%4];"
               : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(addr));
}
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                         uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
               : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi) {
  const __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
  return *reinterpret_cast<const uint32_t*>(&v);
}

// rows [0, len) of one head (64 bf16 = 8 x 16 B per row) -> smem, rows [len, rows_pad) zero-filled.  Asynchronous
// 16-byte copies (cp.async, zero-fill through src-size 0): every load of the CTA's operands is in flight at once instead
// of one global-load round trip per loop iteration; atc_load_wait() before the barrier that publishes the tiles.
__device__ __forceinline__ void atc_load(bf16* dst, const bf16* src, long long ld, int len, int rows_pad) {
  for (int i = threadIdx.x; i < rows_pad * 8; i += blockDim.x) {
    const int r = i >> 3, c = i & 7;
    const bf16* g = src + (long long)(r < len ? r : 0) * ld + c * 8;
    const uint32_t nbytes = r < len ? 16u : 0u;
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(atc_smem(dst + r * ATC_PITCH + c * 8)), "l"(g), "r"(nbytes)
                 : "memory");
  }
}
__device__ __forceinline__ void atc_load_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// A fragments (16 rows x 64) of a row-major smem tile starting at row r0
__device__ __forceinline__ void atc_load_a(uint32_t (&f)[4][4], const bf16* base, int r0, int lane) {
  const int row = r0 + (lane & 7) + ((lane >> 3) & 1) * 8, col = (lane >> 4) * 8;
#pragma unroll
  for (int ks = 0; ks < 4; ++ks) ldsm4(f[ks], atc_smem(base + row * ATC_PITCH + ks * 16 + col));
}

// one 16 x 8 tile nt of A(16 x 64) . Bt^T, Bt row-major [n][64] in smem ("col-major B": K in QK^T, V in dO V^T)
__device__ __forceinline__ void atc_tile_abt(float (&acc)[4], const uint32_t (&a)[4][4], const bf16* bt, int nt, int lane) {
  acc[0] = acc[1] = acc[2] = acc[3] = 0.f;
  uint32_t b[4];
  const bf16* p = bt + (nt * 8 + (lane & 7)) * ATC_PITCH + (lane >> 3) * 8;
  ldsm4(b, atc_smem(p));               // k = 0..31
  mma16816(acc, a[0][0], a[0][1], a[0][2], a[0][3], b[0], b[1]);
  mma16816(acc, a[1][0], a[1][1], a[1][2], a[1][3], b[2], b[3]);
  ldsm4(b, atc_smem(p + 32));          // k = 32..63
  mma16816(acc, a[2][0], a[2][1], a[2][2], a[2][3], b[0], b[1]);
  mma16816(acc, a[3][0], a[3][1], a[3][2], a[3][3], b[2], b[3]);
}

template <int NT>
__device__ __forceinline__ void atc_mma_abt(float (&acc)[NT][4], const uint32_t (&a)[4][4], const bf16* bt, int n_pad,
                                            int lane) {
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    if (nt * 8 < n_pad) atc_tile_abt(acc[nt], a, bt, nt, lane);
    else acc[nt][0] = acc[nt][1] = acc[nt][2] = acc[nt][3] = 0.f;
  }
}

// S -> P in place (fp32): scale, additive mask, softmax over the key axis.  Thread owns rows g = lane/4 and g+8 of the
// warp's 16-row block, columns nt*8 + 2*(lane%4) + {0,1}.  madd[j] = (1 - key_mask[j]) * -10000.
template <int NT>
__device__ __forceinline__ void atc_softmax(float (&s)[NT][4], int i0, int lk, int lk_pad, const float* madd, int causal,
                                            float scale, int lane, bool normalise, float (&inv)[2]) {
  const int t2 = (lane & 3) * 2;
  float mx[2] = {-INFINITY, -INFINITY};
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    if (nt * 8 < lk_pad) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int j = nt * 8 + t2 + (e & 1), i = i0 + (e >> 1) * 8;
        float v = -INFINITY;
        if (j < lk) v = s[nt][e] * scale + ((causal && j > i) ? -10000.f : madd[j]);
        s[nt][e] = v;
        mx[e >> 1] = fmaxf(mx[e >> 1], v);
      }
    }
  }
  float sum[2] = {0.f, 0.f};
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 1));
    mx[r] = fmaxf(mx[r], __shfl_xor_sync(0xffffffffu, mx[r], 2));
  }
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    if (nt * 8 < lk_pad) {
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float p = __expf(s[nt][e] - mx[e >> 1]);
        s[nt][e] = p;
        sum[e >> 1] += p;
      }
    }
  }
#pragma unroll
  for (int r = 0; r < 2; ++r) {
    sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], 1);
    sum[r] += __shfl_xor_sync(0xffffffffu, sum[r], 2);
    inv[r] = 1.f / sum[r];
  }
  if (normalise) {
#pragma unroll
    for (int nt = 0; nt < NT; ++nt)
#pragma unroll
      for (int e = 0; e < 4; ++e) s[nt][e] *= inv[e >> 1];
  }
}

// dropout scale (0 or 1/(1-p), or 0/1 when inv == 1) of the thread's elements of tile nt: element index
// ((seq * H + head) * max_lq + i) * lks + j with lks = max_lk rounded up to even; the same in the forward, the backward
// and the FFMA kernel
struct AtcDrop {
  bool on;
  uint32_t key;
  float p, inv;
  unsigned long long base;     // (seq * H + head) * max_lq
  int lks;                     // row stride of the index space: max_lk rounded up to even, so (j, j+1) with j even is one pair
  // scales of columns j (even) and j + 1 of row i
  __device__ __forceinline__ void scale2(int i, int j, float& s0, float& s1) const {
    if (on) svpc_drop_scale2(key, p, inv, (base + (unsigned long long)i) * lks + j, s0, s1);
    else { s0 = 1.f; s1 = 1.f; }
  }
};
__device__ __forceinline__ AtcDrop atc_make_drop(const SvpcDrop& d, int seq_head, int max_lq, int max_lk) {
  AtcDrop a;
  a.on = d.seed != nullptr && d.p > 0.f;
  a.key = svpc_drop_key(d);
  a.p = d.p;
  a.inv = a.on ? 1.f / (1.f - d.p) : 1.f;
  a.base = (unsigned long long)seq_head * max_lq;
  a.lks = (max_lk + 1) & ~1;
  return a;
}
// multiply the thread's elements of every tile by their dropout scale: v[nt][0..1] row i0, v[nt][2..3] row i0 + 8
template <int NT>
__device__ __forceinline__ void atc_apply_drop(float (&s)[NT][4], const AtcDrop& dr, int i0, int lk_pad, int lane) {
  if (!dr.on) return;
  const int t2 = (lane & 3) * 2;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt)
    if (nt * 8 < lk_pad) {
      float a, b;
      dr.scale2(i0, nt * 8 + t2, a, b);
      s[nt][0] *= a; s[nt][1] *= b;
      dr.scale2(i0 + 8, nt * 8 + t2, a, b);
      s[nt][2] *= a; s[nt][3] *= b;
    }
}

// acc[dt] (16 x 8 tile over the head dim) += A(16 x k_pad, bf16 pairs taken from fp32 tiles p[NT][4]) . B, B row-major
// [k][64] in smem (V in P.V, K in dS.K)
template <int NT>
__device__ __forceinline__ void atc_mma_pb(float (&acc)[8][4], const float (&p)[NT][4], const bf16* b, int k_pad, int lane) {
#pragma unroll
  for (int ks = 0; ks < NT / 2; ++ks) {
    if (ks * 16 < k_pad) {
      const uint32_t a0 = pack_bf16(p[2 * ks][0], p[2 * ks][1]), a1 = pack_bf16(p[2 * ks][2], p[2 * ks][3]);
      const uint32_t a2 = pack_bf16(p[2 * ks + 1][0], p[2 * ks + 1][1]), a3 = pack_bf16(p[2 * ks + 1][2], p[2 * ks + 1][3]);
      const bf16* row = b + (ks * 16 + (lane & 7) + ((lane >> 3) & 1) * 8) * ATC_PITCH + (lane >> 4) * 8;
#pragma unroll
      for (int dp = 0; dp < 4; ++dp) {
        uint32_t bb[4];
        ldsm4t(bb, atc_smem(row + dp * 16));
        mma16816(acc[2 * dp], a0, a1, a2, a3, bb[0], bb[1]);
        mma16816(acc[2 * dp + 1], a0, a1, a2, a3, bb[2], bb[3]);
      }
    }
  }
}

// store a warp's 16 x 64 fp32 tile (rows r0 + g, r0 + g + 8) as bf16, rows < len only
__device__ __forceinline__ void atc_store(bf16* dst, long long ld, const float (&acc)[8][4], int r0, int len, int lane,
                                          float s0, float s1) {
  const int g = lane >> 2, t2 = (lane & 3) * 2;
#pragma unroll
  for (int dt = 0; dt < 8; ++dt) {
    if (r0 + g < len)
      *reinterpret_cast<uint32_t*>(dst + (long long)(r0 + g) * ld + dt * 8 + t2) = pack_bf16(acc[dt][0] * s0, acc[dt][1] * s0);
    if (r0 + g + 8 < len)
      *reinterpret_cast<uint32_t*>(dst + (long long)(r0 + g + 8) * ld + dt * 8 + t2) = pack_bf16(acc[dt][2] * s1, acc[dt][3] * s1);
  }
}

struct AtcParams {
  int H;
  const bf16 *q, *k, *v;
  long long ldq, ldk, ldv;
  const int *q_off, *q_len, *k_off, *k_len;
  const float* key_mask;
  int causal, lq_pad, lk_pad, max_lq, max_lk;
  SvpcDrop drop;                 // dropout on the attention probabilities (model.py:214)
  bf16* ctx; long long ldo;
  const bf16* dctx; long long lddo;
  bf16 *dq, *dk, *dv;
  long long lddq, lddk, lddv;
};

// ------------------------------------------------------------------------------------------------ forward
// FULL: both padded lengths equal NT * 8 (the usual case: L = 100 -> 112, L = 25 -> 32), so that every column-block bound
// check of the unrolled loops folds at compile time
template <int NT, bool FULL>
__global__ void __launch_bounds__(NT <= 14 ? 224 : 256, NT <= 4 ? 4 : (NT <= 14 ? ATC_FWD_BLOCKS : 1)) attn_tc_fwd_kernel(const AtcParams p) {
  svpc_pdl_sync();
  const int lq_pad = FULL ? NT * 8 : p.lq_pad, lk_pad = FULL ? NT * 8 : p.lk_pad;
  extern __shared__ __align__(16) uint8_t atc_raw[];
  bf16* sQ = reinterpret_cast<bf16*>(atc_raw);
  bf16* sK = sQ + lq_pad * ATC_PITCH;
  bf16* sV = sK + lk_pad * ATC_PITCH;
  float* madd = reinterpret_cast<float*>(sV + lk_pad * ATC_PITCH);
  const int n = blockIdx.x / p.H, h = blockIdx.x - n * p.H;
  const int lq = p.q_len[n], lk = p.k_len[n], qo = p.q_off[n], ko = p.k_off[n];
  atc_load(sQ, p.q + (long long)qo * p.ldq + h * ATC_DH, p.ldq, lq, lq_pad);
  atc_load(sK, p.k + (long long)ko * p.ldk + h * ATC_DH, p.ldk, lk, lk_pad);
  atc_load(sV, p.v + (long long)ko * p.ldv + h * ATC_DH, p.ldv, lk, lk_pad);
  for (int j = threadIdx.x; j < lk_pad; j += blockDim.x)
    madd[j] = (j < lk && p.key_mask) ? (1.f - p.key_mask[ko + j]) * -10000.f : 0.f;
  atc_load_wait();
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int r0 = warp * 16;
  if (r0 >= lq) return;
  uint32_t qf[4][4];
  atc_load_a(qf, sQ, r0, lane);
  float s[NT][4];
  atc_mma_abt<NT>(s, qf, sK, lk_pad, lane);
  float inv[2];
  atc_softmax<NT>(s, r0 + (lane >> 2), lk, lk_pad, madd, p.causal, 0.125f, lane, false, inv);
  const AtcDrop dr = atc_make_drop(p.drop, blockIdx.x, p.max_lq, p.max_lk);
  atc_apply_drop<NT>(s, dr, r0 + (lane >> 2), lk_pad, lane);     // scale 1/(1-p) rides on the kept probabilities
  float o[8][4];
#pragma unroll
  for (int dt = 0; dt < 8; ++dt) o[dt][0] = o[dt][1] = o[dt][2] = o[dt][3] = 0.f;
  atc_mma_pb<NT>(o, s, sV, lk_pad, lane);
  atc_store(p.ctx + (long long)qo * p.ldo + h * ATC_DH, p.ldo, o, r0, lq, lane, inv[0], inv[1]);
}

// ------------------------------------------------------------------------------------------------ backward
// acc (16 keys x 64) = X^T(keys x rows) . B(rows x 64): X row-major [row i][key j] in smem (pitch xp), B = dO or Q
__device__ __forceinline__ void atc_mma_xtb(float (&acc)[8][4], const bf16* X, int xp, const bf16* B, int j0, int rows_pad,
                                            int lane) {
#pragma unroll
  for (int dt = 0; dt < 8; ++dt) acc[dt][0] = acc[dt][1] = acc[dt][2] = acc[dt][3] = 0.f;
  for (int i0 = 0; i0 < rows_pad; i0 += 16) {
    uint32_t a[4];
    // A[m = key][k = row] = X[row][key]: transposed load, matrices (i0, j0), (i0, j0+8), (i0+8, j0), (i0+8, j0+8)
    ldsm4t(a, atc_smem(X + (i0 + (lane & 7) + (lane >> 4) * 8) * xp + j0 + ((lane >> 3) & 1) * 8));
    const bf16* row = B + (i0 + (lane & 7) + ((lane >> 3) & 1) * 8) * ATC_PITCH + (lane >> 4) * 8;
#pragma unroll
    for (int dp = 0; dp < 4; ++dp) {
      uint32_t bb[4];
      ldsm4t(bb, atc_smem(row + dp * 16));
      mma16816(acc[2 * dp], a[0], a[1], a[2], a[3], bb[0], bb[1]);
      mma16816(acc[2 * dp + 1], a[0], a[1], a[2], a[3], bb[2], bb[3]);
    }
  }
}

template <int NT>
__device__ __forceinline__ void atc_write_x(bf16* X, int xp, const float (&s)[NT][4], int r0, int lk_pad, int lane,
                                            const AtcDrop* dr) {
  const int g = lane >> 2, t2 = (lane & 3) * 2;
#pragma unroll
  for (int nt = 0; nt < NT; ++nt) {
    if (nt * 8 < lk_pad) {
      float v[4] = {s[nt][0], s[nt][1], s[nt][2], s[nt][3]};
      if (dr) {                          // P with the dropout decision in its sign (see attn_tc_bwd_kernel): dropped -> 0
#pragma unroll
        for (int e = 0; e < 4; ++e) v[e] = fmaxf(v[e], 0.f) * dr->inv;
      }
      *reinterpret_cast<uint32_t*>(X + (r0 + g) * xp + nt * 8 + t2) = pack_bf16(v[0], v[1]);
      *reinterpret_cast<uint32_t*>(X + (r0 + g + 8) * xp + nt * 8 + t2) = pack_bf16(v[2], v[3]);
    }
  }
}

template <int NT, bool FULL>
__global__ void __launch_bounds__(NT <= 14 ? 224 : 256, NT <= 14 ? 2 : 1) attn_tc_bwd_kernel(const AtcParams p) {
  svpc_pdl_sync();
  const int lq_pad = FULL ? NT * 8 : p.lq_pad, lk_pad = FULL ? NT * 8 : p.lk_pad;
  extern __shared__ __align__(16) uint8_t atc_raw[];
  bf16* sQ = reinterpret_cast<bf16*>(atc_raw);
  bf16* sdO = sQ + lq_pad * ATC_PITCH;
  bf16* sK = sdO + lq_pad * ATC_PITCH;
  bf16* sV = sK + lk_pad * ATC_PITCH;
  const int xp = lk_pad + 8;
  bf16* sX = sV + lk_pad * ATC_PITCH;                  // [lq_pad][xp]: P, then dS
  float* madd = reinterpret_cast<float*>(sX + lq_pad * xp);
  const int n = blockIdx.x / p.H, h = blockIdx.x - n * p.H;
  const int lq = p.q_len[n], lk = p.k_len[n], qo = p.q_off[n], ko = p.k_off[n];
  atc_load(sQ, p.q + (long long)qo * p.ldq + h * ATC_DH, p.ldq, lq, lq_pad);
  atc_load(sdO, p.dctx + (long long)qo * p.lddo + h * ATC_DH, p.lddo, lq, lq_pad);
  atc_load(sK, p.k + (long long)ko * p.ldk + h * ATC_DH, p.ldk, lk, lk_pad);
  atc_load(sV, p.v + (long long)ko * p.ldv + h * ATC_DH, p.ldv, lk, lk_pad);
  for (int j = threadIdx.x; j < lk_pad; j += blockDim.x)
    madd[j] = (j < lk && p.key_mask) ? (1.f - p.key_mask[ko + j]) * -10000.f : 0.f;
  atc_load_wait();
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, nwarps = blockDim.x >> 5;
  const int r0 = warp * 16;
  const bool active = r0 < lq_pad;                       // every warp of the CTA owns one padded row block
  const float scale = 0.125f;
  const AtcDrop dr = atc_make_drop(p.drop, blockIdx.x, p.max_lq, p.max_lk);
  const int gi = r0 + (lane >> 2), t2 = (lane & 3) * 2;
  float s[NT][4];
  if (active) {
    uint32_t af[4][4];
    atc_load_a(af, sQ, r0, lane);
    atc_mma_abt<NT>(s, af, sK, lk_pad, lane);
    float inv[2];
    atc_softmax<NT>(s, r0 + (lane >> 2), lk, lk_pad, madd, p.causal, scale, lane, true, inv);
    // dropout: the keep / drop decision of every probability is computed ONCE and kept in the sign of P (P >= 0; a dropped
    // P is stored negated), instead of re-hashing the element index in each of the three places that need the mask
    if (dr.on) {
#pragma unroll
      for (int nt = 0; nt < NT; ++nt) {
        if (nt * 8 < lk_pad) {
          float m[4];
          dr.scale2(gi, nt * 8 + t2, m[0], m[1]);
          dr.scale2(gi + 8, nt * 8 + t2, m[2], m[3]);
#pragma unroll
          for (int e = 0; e < 4; ++e) s[nt][e] = m[e] == 0.f ? -s[nt][e] : s[nt][e];
        }
      }
    }
    atc_write_x<NT>(sX, xp, s, r0, lk_pad, lane, dr.on ? &dr : nullptr);   // dropped P (bf16) for dV
    // D_i = sum_j P_ij dP_ij, dP = dO V^T (recomputed below instead of kept: 4*NT registers saved)
    atc_load_a(af, sdO, r0, lane);
    float dsum[2] = {0.f, 0.f};
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      if (nt * 8 < lk_pad) {
        float dp[4];
        atc_tile_abt(dp, af, sV, nt, lane);
#pragma unroll
        for (int e = 0; e < 4; ++e)                                          // (a dropped probability carries a minus sign)
          dsum[e >> 1] = fmaf(fmaxf(s[nt][e], 0.f), dp[e], dsum[e >> 1]);
      }
    }
#pragma unroll
    for (int r = 0; r < 2; ++r) {
      dsum[r] += __shfl_xor_sync(0xffffffffu, dsum[r], 1);
      dsum[r] += __shfl_xor_sync(0xffffffffu, dsum[r], 2);
      dsum[r] *= dr.inv * scale;                            // D_i * scale (the kept probabilities carry 1 / (1 - p))
    }
    const float c1 = dr.inv * scale;
    // dS = P * (dP - D) * scale, in place of P (dP tile recomputed: two MMAs cost less than 4*NT live registers)
#pragma unroll
    for (int nt = 0; nt < NT; ++nt) {
      if (nt * 8 < lk_pad) {
        float dp[4];
        atc_tile_abt(dp, af, sV, nt, lane);
#pragma unroll
        for (int e = 0; e < 4; ++e)     // P (m dP - D) scale = relu(s) dP (inv scale) - |s| (D scale)
          s[nt][e] = fmaf(fmaxf(s[nt][e], 0.f) * dp[e], c1, -fabsf(s[nt][e]) * dsum[e >> 1]);
      }
    }
  }
  __syncthreads();                                          // P complete in sX; sV no longer needed as an operand
  // dV[j] = sum_i P[i][j] dO[i]
  for (int j0 = warp * 16; j0 < lk_pad; j0 += nwarps * 16) {
    float acc[8][4];
    atc_mma_xtb(acc, sX, xp, sdO, j0, lq_pad, lane);
    atc_store(p.dv + (long long)ko * p.lddv + h * ATC_DH, p.lddv, acc, j0, lk, lane, 1.f, 1.f);
  }
  // dQ[i] = sum_j dS[i][j] K[j]   (dS still in registers)
  if (active && r0 < lq) {
    float acc[8][4];
#pragma unroll
    for (int dt = 0; dt < 8; ++dt) acc[dt][0] = acc[dt][1] = acc[dt][2] = acc[dt][3] = 0.f;
    atc_mma_pb<NT>(acc, s, sK, lk_pad, lane);
    atc_store(p.dq + (long long)qo * p.lddq + h * ATC_DH, p.lddq, acc, r0, lq, lane, 1.f, 1.f);
  }
  __syncthreads();                                          // every warp is done reading P
  if (active) atc_write_x<NT>(sX, xp, s, r0, lk_pad, lane, nullptr);
  __syncthreads();
  // dK[j] = sum_i dS[i][j] Q[i]
  for (int j0 = warp * 16; j0 < lk_pad; j0 += nwarps * 16) {
    float acc[8][4];
    atc_mma_xtb(acc, sX, xp, sQ, j0, lq_pad, lane);
    atc_store(p.dk + (long long)ko * p.lddk + h * ATC_DH, p.lddk, acc, j0, lk, lane, 1.f, 1.f);
  }
}

// ------------------------------------------------------------------------------------------------ host side
static inline int atc_pad16(int x) { return (x + 15) / 16 * 16; }

int svpc_attn_tc_eligible(int dh, int dt, int max_lq, int max_lk, long long ldq, long long ldk, long long ldv, long long ldo,
                          const void* q, const void* k, const void* v, const void* o) {
  if (dt != SVPC_BF16 || dh != ATC_DH) return 0;
  if (max_lq < 16 || max_lk < 16 || max_lq > 128 || max_lk > 128) return 0;
  if ((ldq | ldk | ldv) % 8 != 0 || ldo % 2 != 0) return 0;
  if ((((uintptr_t)q | (uintptr_t)k | (uintptr_t)v) & 15) != 0 || ((uintptr_t)o & 3) != 0) return 0;
  return 1;
}

int svpc_attn_tc_fwd_launch(int n_seq, const AtcParams& p0, int max_lq, int max_lk, cudaStream_t st) {
  AtcParams p = p0;
  p.lq_pad = atc_pad16(max_lq);
  p.lk_pad = atc_pad16(max_lk);
  p.max_lq = max_lq; p.max_lk = max_lk;
  const size_t smem = (size_t)(p.lq_pad + 2 * p.lk_pad) * ATC_PITCH * 2 + (size_t)p.lk_pad * 4;
  const int threads = 32 * (p.lq_pad / 16);
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(attn_tc_fwd_kernel<14, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(attn_tc_fwd_kernel<14, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(attn_tc_fwd_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    cudaFuncSetAttribute(attn_tc_fwd_kernel<16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 100 * 1024);
    attr = true;
  }
  const int grid = n_seq * p.H;
  const bool sq = p.lq_pad == p.lk_pad;
  if (p.lk_pad <= 32) {
    if (sq && p.lk_pad == 32) svpc_launch(attn_tc_fwd_kernel<4, true>, grid, threads, smem, st, p);
    else svpc_launch(attn_tc_fwd_kernel<4, false>, grid, threads, smem, st, p);
  } else if (p.lk_pad <= 112) {
    if (sq && p.lk_pad == 112) svpc_launch(attn_tc_fwd_kernel<14, true>, grid, threads, smem, st, p);
    else svpc_launch(attn_tc_fwd_kernel<14, false>, grid, threads, smem, st, p);
  } else {
    if (sq && p.lk_pad == 128) svpc_launch(attn_tc_fwd_kernel<16, true>, grid, threads, smem, st, p);
    else svpc_launch(attn_tc_fwd_kernel<16, false>, grid, threads, smem, st, p);
  }
  return svpc_check_launch("attn_tc_fwd");
}

int svpc_attn_tc_bwd_launch(int n_seq, const AtcParams& p0, int max_lq, int max_lk, cudaStream_t st) {
  AtcParams p = p0;
  p.lq_pad = atc_pad16(max_lq);
  p.lk_pad = atc_pad16(max_lk);
  p.max_lq = max_lq; p.max_lk = max_lk;
  const size_t smem = (size_t)(2 * p.lq_pad + 2 * p.lk_pad) * ATC_PITCH * 2 + (size_t)p.lq_pad * (p.lk_pad + 8) * 2 +
                      (size_t)p.lk_pad * 4;
  const int threads = 32 * (p.lq_pad / 16);
  static bool attr = false;
  if (!attr) {
    cudaFuncSetAttribute(attn_tc_bwd_kernel<4, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    cudaFuncSetAttribute(attn_tc_bwd_kernel<4, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    cudaFuncSetAttribute(attn_tc_bwd_kernel<14, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    cudaFuncSetAttribute(attn_tc_bwd_kernel<14, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    cudaFuncSetAttribute(attn_tc_bwd_kernel<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    cudaFuncSetAttribute(attn_tc_bwd_kernel<16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 120 * 1024);
    attr = true;
  }
  const int grid = n_seq * p.H;
  const bool sq = p.lq_pad == p.lk_pad;
  if (p.lk_pad <= 32) {
    if (sq && p.lk_pad == 32) svpc_launch(attn_tc_bwd_kernel<4, true>, grid, threads, smem, st, p);
    else svpc_launch(attn_tc_bwd_kernel<4, false>, grid, threads, smem, st, p);
  } else if (p.lk_pad <= 112 && p.lq_pad <= 112) {
    if (sq && p.lk_pad == 112) svpc_launch(attn_tc_bwd_kernel<14, true>, grid, threads, smem, st, p);
    else svpc_launch(attn_tc_bwd_kernel<14, false>, grid, threads, smem, st, p);
  } else {
    if (sq && p.lk_pad == 128) svpc_launch(attn_tc_bwd_kernel<16, true>, grid, threads, smem, st, p);
    else svpc_launch(attn_tc_bwd_kernel<16, false>, grid, threads, smem, st, p);
  }
  return svpc_check_launch("attn_tc_bwd");
}

// plain-argument entry points used by the dispatch in attention.cu
int svpc_attn_tc_fwd(int n_seq, int H, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                     long long ldv, const int* q_off, const int* q_len, const int* k_off, const int* k_len,
                     const float* key_mask, int causal, int max_lq, int max_lk, void* ctx, long long ldo, SvpcDrop drop,
                     cudaStream_t st) {
  AtcParams p{};
  p.drop = drop;
  p.H = H; p.q = (const bf16*)q; p.k = (const bf16*)k; p.v = (const bf16*)v;
  p.ldq = ldq; p.ldk = ldk; p.ldv = ldv;
  p.q_off = q_off; p.q_len = q_len; p.k_off = k_off; p.k_len = k_len;
  p.key_mask = key_mask; p.causal = causal;
  p.ctx = (bf16*)ctx; p.ldo = ldo;
  return svpc_attn_tc_fwd_launch(n_seq, p, max_lq, max_lk, st);
}

int svpc_attn_tc_bwd(int n_seq, int H, const void* q, long long ldq, const void* k, long long ldk, const void* v,
                     long long ldv, const int* q_off, const int* q_len, const int* k_off, const int* k_len,
                     const float* key_mask, int causal, int max_lq, int max_lk, const void* dctx, long long lddo, void* dq,
                     long long lddq, void* dk, long long lddk, void* dv, long long lddv, SvpcDrop drop, cudaStream_t st) {
  AtcParams p{};
  p.drop = drop;
  p.H = H; p.q = (const bf16*)q; p.k = (const bf16*)k; p.v = (const bf16*)v;
  p.ldq = ldq; p.ldk = ldk; p.ldv = ldv;
  p.q_off = q_off; p.q_len = q_len; p.k_off = k_off; p.k_len = k_len;
  p.key_mask = key_mask; p.causal = causal;
  p.dctx = (const bf16*)dctx; p.lddo = lddo;
  p.dq = (bf16*)dq; p.dk = (bf16*)dk; p.dv = (bf16*)dv;
  p.lddq = lddq; p.lddk = lddk; p.lddv = lddv;
  return svpc_attn_tc_bwd_launch(n_seq, p, max_lq, max_lk, st);
}
