// tcgen05 / TMEM / TMA GEMM for the dense contractions of the SVPC step (bf16 operands, fp32 accumulate in TMEM).
//   D[M,N] = act(A[M,K] . B[N,K]^T + bias) + residual   (optionally += D, optionally a copy of the pre-activation)
// Both operands may be K-major (row-major with K contiguous: activations x weights^T, the forward Linear) or MN-major
// (the M / N index contiguous: dgrad uses W as an MN-major B, wgrad uses dY and X as MN-major A and B), so no
// transposed copies of weights or activations are ever materialised.
//
// Structure: PERSISTENT kernel, one CTA of 320 threads (10 warps) per SM, walking a list of work units (128 x BN tiles, BN <= 256;
// split-K slices for weight gradients; narrower column slices for the tiles of the last partial round):
//   warp 0      TMA producer: cp.async.bulk.tensor 2D boxes, 128B swizzle, into a 4-deep shared-memory ring
//               (A 16 KB + B up to 32 KB per stage), completion on mbarriers (expect_tx); runs ahead across units
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma.cta_group::1.kind::f16 (UMMA 128 x width x 16) from
//               shared-memory descriptors into one of TWO 256-column TMEM accumulators; tcgen05.commit releases ring
//               slots / signals the epilogue
//   warps 2-9   epilogue (two warps per TMEM lane quarter, alternating 32-column chunks): tcgen05.ld (32 lanes x 32
//               columns) -> per-warp 32 x 32 fp32 XOR-swizzled smem transposition buffer -> bias / activation / dropout /
//               residual -> row-contiguous global stores, overlapped with the next unit's main loop; warp 2 also owns
//               the TMEM allocation
// Split-K with fp32 red.global.add.v4 for weight-gradient shapes whose output has too few tiles to fill 148 SMs.
#include "common.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <mutex>

#define TC_BM 128
#define TC_BK 64                 // 64 bf16 = 128 B = one swizzle row
#define TC_BN_MAX 256
#define TC_STAGES 4              // ring depth at the full tile width (256 columns); narrower launches cut the same 192 KB into more stages
#define TC_MAX_STAGES 8
#define TC_EPI_WARPS 8                                  // two per TMEM lane quarter
#define TC_THREADS (64 + 32 * TC_EPI_WARPS)
#define TC_UMMA_K 16
#define TC_A_BYTES (TC_BM * TC_BK * 2)                  // 16 KB
#define TC_B_BYTES (TC_BN_MAX * TC_BK * 2)              // 32 KB
#define TC_STAGE_BYTES (TC_A_BYTES + TC_B_BYTES)
#define TC_EPI_BYTES (TC_EPI_WARPS * 32 * 32 * 4)        // one 32 x 32 fp32 transposition buffer (XOR-swizzled) per warp
#define TC_SMEM_BYTES (TC_STAGES * TC_STAGE_BYTES + TC_EPI_BYTES + 256 + 1024)

// ------------------------------------------------------------------------------------------------ PTX wrappers
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t done;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(done)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!done);
}
__device__ __forceinline__ void tma_load_2d(void* dst, const CUtensorMap* map, uint64_t* bar, int c0, int c1) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(smem_u32(dst)), "l"((uint64_t)map), "r"(smem_u32(bar)), "r"(c0), "r"(c1)
      : "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t* dst_smem, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(dst_smem)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void umma_f16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* v) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
      "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
}
// the registers written by tmem_ld32 may only be read after this (the first reader must be a memory operation: the
// "memory" clobber is what keeps the compiler from hoisting it)
__device__ __forceinline__ void tmem_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// 4-element row-contiguous loads / stores of the output type
__device__ __forceinline__ float4 ld4(const float* p) { return *reinterpret_cast<const float4*>(p); }
__device__ __forceinline__ float4 ld4(const bf16* p) {
  const uint2 u = *reinterpret_cast<const uint2*>(p);
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x), b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
  return make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
}
__device__ __forceinline__ void st4(float* p, float4 v) { *reinterpret_cast<float4*>(p) = v; }
__device__ __forceinline__ void st4(bf16* p, float4 v) {
  const __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
  uint2 u;
  u.x = *reinterpret_cast<const uint32_t*>(&a);
  u.y = *reinterpret_cast<const uint32_t*>(&b);
  *reinterpret_cast<uint2*>(p) = u;
}

// explicit shared-space 16-byte accesses (a pointer derived from the dynamic shared base through integer arithmetic is
// "generic" to the compiler, which then emits slower generic ST / LD)
__device__ __forceinline__ void sts128(uint32_t addr, uint32_t a, uint32_t b, uint32_t c, uint32_t d) {
  asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(addr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ float4 lds128(uint32_t addr) {
  float4 v;
  asm volatile("ld.shared.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "r"(addr) : "memory");
  return v;
}
__device__ __forceinline__ void red_add_f32x4(float* gptr, float4 v) {
  asm volatile("red.global.add.v4.f32 [%0], {%1, %2, %3, %4};" ::"l"(gptr), "f"(v.x), "f"(v.y), "f"(v.z), "f"(v.w) : "memory");
}
__device__ __forceinline__ void red_add_f32(float* gptr, float v) {
  asm volatile("red.global.add.f32 [%0], %1;" ::"l"(gptr), "f"(v) : "memory");
}

// four consecutive elements in their storage type
template <typename T> struct Raw4;
template <> struct Raw4<float> { typedef float4 type; };
template <> struct Raw4<bf16> { typedef uint2 type; };
__device__ __forceinline__ float4 raw4_to_float4(float4 v) { return v; }
__device__ __forceinline__ float4 raw4_to_float4(uint2 u) {
  const __nv_bfloat162 a = *reinterpret_cast<const __nv_bfloat162*>(&u.x), b = *reinterpret_cast<const __nv_bfloat162*>(&u.y);
  return make_float4(__low2float(a), __high2float(a), __low2float(b), __high2float(b));
}

// shared-memory matrix descriptor (cute::UMMA::SmemDescriptor), 128B swizzle, version 1 (Blackwell)
__device__ __forceinline__ uint64_t make_sdesc(uint32_t saddr, uint32_t lbo_bytes, uint32_t sbo_bytes) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo_bytes >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)((sbo_bytes >> 4) & 0x3FFF) << 32;
  d |= (uint64_t)1 << 46;   // version
  d |= (uint64_t)2 << 61;   // SWIZZLE_128B
  return d;
}

struct TcParams {
  int M, N, K;
  void* D; long long ldd;
  const float* bias; int act;
  const void* residual; long long ldr;
  void* pre; long long ldp;
  int accumulate;          // D += result
  int atomic;              // split-K: red.global.add.f32
  SvpcDrop drop;           // dropout after the activation, before the residual (model.py:231, 287, 757)
  const void* gref; long long ldg; int gact;   // last step of the epilogue: D = (...) * gact'(gref[m, n])  (fused activation backward)
  // ---- work list (see tc_unit): every role of every CTA walks units blockIdx.x, blockIdx.x + gridDim.x, ...
  int BN;                  // tile width (64 / 128 / 256)
  int tiles_n;             // tiles along N
  int split, kb_per;       // split-K slices per tile, K blocks (of TC_BK) per slice
  int kb_total;
  int main_tiles;          // tiles handled whole (x split); the remaining `tiles - main_tiles` are cut along N
  int tail_s;              // ... into tail_s sub-tiles of width BN / tail_s each, so the last round is short
  int total_units;
  int nseg;                // 1: D = A.B^T;  3: split-fp32 operands, D = A.B^T + Alo.B^T + A.Blo^T (same accumulator)
  int nstages, stage_bytes;   // shared-memory ring: nstages slots of A (16 KB) + the launch's widest B tile
};

struct TcUnit { int m0, n0, width, kb0, kb1; };

// unit index -> (row block, column range, K-block range).  Full 128 x BN tiles first (split-K slices adjacent); the
// tiles of the last, partial round are cut into narrower column slices so that every SM gets a (shorter) piece.
__device__ __forceinline__ TcUnit tc_unit(const TcParams& p, int u) {
  TcUnit w;
  const int main_units = p.main_tiles * p.split;
  int tile, ks = 0, sub = 0;
  w.width = p.BN;
  if (u < main_units) {
    tile = u / p.split;
    ks = u - tile * p.split;
  } else {
    const int r = u - main_units;
    tile = p.main_tiles + r / p.tail_s;
    sub = r % p.tail_s;
    w.width = p.BN / p.tail_s;
  }
  const int tm = tile / p.tiles_n;
  w.m0 = tm * TC_BM;
  w.n0 = (tile - tm * p.tiles_n) * p.BN + sub * w.width;
  w.kb0 = ks * p.kb_per;
  w.kb1 = min(p.kb_total, w.kb0 + p.kb_per);
  return w;
}

__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}

template <int ACT>
__device__ __forceinline__ float tc_act(float v) {
  if (ACT == SVPC_ACT_RELU) return v > 0.f ? v : 0.f;
  if (ACT == SVPC_ACT_GELU) return svpc_gelu(v);
  if (ACT == SVPC_ACT_SIGMOID) return svpc_sigmoid(v);
  return v;
}

__device__ __forceinline__ float tc_act_grad(int act, float ref) {
  if (act == SVPC_ACT_RELU) return ref > 0.f ? 1.f : 0.f;
  if (act == SVPC_ACT_GELU) return svpc_gelu_grad(ref);
  if (act == SVPC_ACT_SIGMOID) return ref * (1.f - ref);
  return 1.f;
}

// The warp's 32 x 32 fp32 transposition buffer: 16-byte slot (row r, column group g) lives at r * 8 + (g ^ (r & 7)), so
// both the writes (lane = row, one column group at a time) and the reads (8 lanes per row) are bank-conflict free.
__device__ __forceinline__ uint32_t epi_slot(int r, int g) { return (uint32_t)(r * 8 + (g ^ (r & 7))) * 16u; }   // byte offset

// Epilogue of one 32 x 32 chunk: this lane owns column group g = lane & 7 (columns n .. n+3) of rows rsub, rsub + 4, ...
// Vector path: all four columns in range, 16-byte aligned rows.  Loads (residual, accumulate) are issued for all 8 rows
// before the first store so their latencies overlap.
template <int ACT, bool DROP, typename TD>
__device__ __forceinline__ void epi_vec(const TcParams& p, uint32_t stage, int mw, int n, int rsub, int g,
                                        uint32_t dkey, float dinv) {
  float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.bias) bv = *reinterpret_cast<const float4*>(p.bias + n);
  const int m_first = mw + rsub;
  const int rows = min(8, (p.M - m_first + 3) >> 2);       // rows m_first, m_first + 4, ... below M
  TD* __restrict__ d0 = (TD*)p.D + (long long)m_first * p.ldd + n;
  const long long dstep = 4 * p.ldd;
#pragma unroll 1
  for (int j0 = 0; j0 < 8; j0 += 4) {       // two batches of 4 rows: bounded register use, 4 loads in flight
    typename Raw4<TD>::type rr[4];         // residual rows, still in the storage type (2 registers each for bf16)
    typename Raw4<TD>::type gg[4];         // activation-gradient reference rows
    if (p.residual) {
      const TD* r0 = (const TD*)p.residual + (long long)(m_first + 4 * j0) * p.ldr + n;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j0 + j < rows) rr[j] = *reinterpret_cast<const typename Raw4<TD>::type*>(r0 + (long long)j * 4 * p.ldr);
    }
    if (p.gref) {
      const TD* g0 = (const TD*)p.gref + (long long)(m_first + 4 * j0) * p.ldg + n;
#pragma unroll
      for (int j = 0; j < 4; ++j)
        if (j0 + j < rows) gg[j] = *reinterpret_cast<const typename Raw4<TD>::type*>(g0 + (long long)j * 4 * p.ldg);
    }
#pragma unroll
    for (int jj = 0; jj < 4; ++jj) {
      const int j = j0 + jj;
      if (j >= rows) break;
      const int r = j * 4 + rsub;
      float4 x = lds128(stage + epi_slot(r, g));
      x.x += bv.x; x.y += bv.y; x.z += bv.z; x.w += bv.w;
      if (p.pre) st4((TD*)p.pre + (long long)(mw + r) * p.ldp + n, x);
      x.x = tc_act<ACT>(x.x); x.y = tc_act<ACT>(x.y); x.z = tc_act<ACT>(x.z); x.w = tc_act<ACT>(x.w);
      if (DROP) {
        const unsigned long long e0 = (unsigned long long)(mw + r) * p.N + n;
        if ((e0 & 1) == 0) {             // aligned pairs: one hash per two elements
          float s0, s1;
          svpc_drop_scale2(dkey, p.drop.p, dinv, e0, s0, s1);
          x.x *= s0; x.y *= s1;
          svpc_drop_scale2(dkey, p.drop.p, dinv, e0 + 2, s0, s1);
          x.z *= s0; x.w *= s1;
        } else {
          x.x *= svpc_drop_scale(dkey, p.drop.p, dinv, e0); x.y *= svpc_drop_scale(dkey, p.drop.p, dinv, e0 + 1);
          x.z *= svpc_drop_scale(dkey, p.drop.p, dinv, e0 + 2); x.w *= svpc_drop_scale(dkey, p.drop.p, dinv, e0 + 3);
        }
      }
      if (p.residual) { const float4 r4 = raw4_to_float4(rr[jj]); x.x += r4.x; x.y += r4.y; x.z += r4.z; x.w += r4.w; }
      if (p.accumulate) { const float4 dd = ld4(d0 + j * dstep); x.x += dd.x; x.y += dd.y; x.z += dd.z; x.w += dd.w; }
      if (p.gref) {
        const float4 g4 = raw4_to_float4(gg[jj]);
        x.x *= tc_act_grad(p.gact, g4.x); x.y *= tc_act_grad(p.gact, g4.y);
        x.z *= tc_act_grad(p.gact, g4.z); x.w *= tc_act_grad(p.gact, g4.w);
      }
      st4(d0 + j * dstep, x);
    }
  }
}

// bf16 outputs: the residual / activation-gradient reference rows of the whole chunk (8 rows x 8 bytes per lane) are loaded
// BEFORE the accumulator chunk is fetched from TMEM, so their L2 latency hides behind tcgen05.ld + the transposition instead
// of stalling the row loop twice per chunk (the residual epilogues ran at 0.36 of peak against 0.60 plain).
struct EpiPre { uint2 rr[8]; uint2 gg[8]; float4 bias; };
__device__ __forceinline__ void epi_prefetch_bf16(const TcParams& p, int mw, int n, int rsub, EpiPre& o) {
  const int m_first = mw + rsub;
  const int rows = min(8, (p.M - m_first + 3) >> 2);        // >= 1 (the caller checked mw < M ... for rsub = 0; clamped below)
  // unconditional loads (rows past M re-read the last valid one): a predicated load would be merged into its register with
  // a PRMT right here, and that PRMT is where the warp would then wait for the data instead of after the TMEM fetch
  const int last = max(rows - 1, 0);
  const int m0c = min(m_first, p.M - 1);
  o.bias = make_float4(0.f, 0.f, 0.f, 0.f);
  if (p.bias) o.bias = __ldg(reinterpret_cast<const float4*>(p.bias + n));
  if (p.residual) {
    const bf16* r0 = (const bf16*)p.residual + (long long)m0c * p.ldr + n;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.rr[j] = __ldg(reinterpret_cast<const uint2*>(r0 + (long long)min(j, last) * 4 * p.ldr));
  }
  if (p.gref) {
    const bf16* g0 = (const bf16*)p.gref + (long long)m0c * p.ldg + n;
#pragma unroll
    for (int j = 0; j < 8; ++j) o.gg[j] = __ldg(reinterpret_cast<const uint2*>(g0 + (long long)min(j, last) * 4 * p.ldg));
  }
}

template <int ACT, bool DROP>
__device__ __forceinline__ void epi_vec_bf16(const TcParams& p, uint32_t stage, int mw, int n, int rsub, int g,
                                             uint32_t dkey, float dinv, const EpiPre& o) {
  const float4 bv = o.bias;
  const int m_first = mw + rsub;
  const int rows = min(8, (p.M - m_first + 3) >> 2);
  bf16* __restrict__ d0 = (bf16*)p.D + (long long)m_first * p.ldd + n;
  const long long dstep = 4 * p.ldd;
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    if (j >= rows) break;
    const int r = j * 4 + rsub;
    float4 x = lds128(stage + epi_slot(r, g));
    x.x += bv.x; x.y += bv.y; x.z += bv.z; x.w += bv.w;
    if (p.pre) st4((bf16*)p.pre + (long long)(mw + r) * p.ldp + n, x);
    x.x = tc_act<ACT>(x.x); x.y = tc_act<ACT>(x.y); x.z = tc_act<ACT>(x.z); x.w = tc_act<ACT>(x.w);
    if (DROP) {
      const unsigned long long e0 = (unsigned long long)(mw + r) * p.N + n;
      if ((e0 & 1) == 0) {             // aligned pairs: one hash per two elements
        float s0, s1;
        svpc_drop_scale2(dkey, p.drop.p, dinv, e0, s0, s1);
        x.x *= s0; x.y *= s1;
        svpc_drop_scale2(dkey, p.drop.p, dinv, e0 + 2, s0, s1);
        x.z *= s0; x.w *= s1;
      } else {
        x.x *= svpc_drop_scale(dkey, p.drop.p, dinv, e0); x.y *= svpc_drop_scale(dkey, p.drop.p, dinv, e0 + 1);
        x.z *= svpc_drop_scale(dkey, p.drop.p, dinv, e0 + 2); x.w *= svpc_drop_scale(dkey, p.drop.p, dinv, e0 + 3);
      }
    }
    if (p.residual) { const float4 r4 = raw4_to_float4(o.rr[j]); x.x += r4.x; x.y += r4.y; x.z += r4.z; x.w += r4.w; }
    if (p.accumulate) { const float4 dd = ld4(d0 + j * dstep); x.x += dd.x; x.y += dd.y; x.z += dd.z; x.w += dd.w; }
    if (p.gref) {
      const float4 g4 = raw4_to_float4(o.gg[j]);
      x.x *= tc_act_grad(p.gact, g4.x); x.y *= tc_act_grad(p.gact, g4.y);
      x.z *= tc_act_grad(p.gact, g4.z); x.w *= tc_act_grad(p.gact, g4.w);
    }
    st4(d0 + j * dstep, x);
  }
}

// Edge / unaligned path (N not a multiple of 4, odd row pitch, last partial columns): rare, kept out of line.
template <typename TD>
__device__ __forceinline__ void epi_scalar(const TcParams& p, uint32_t stage, int mw, int n, int rsub, int g, bool dropping,
                                        uint32_t dkey, float dinv) {
  TD* __restrict__ Dp = (TD*)p.D;
  const TD* __restrict__ Rp = (const TD*)p.residual;
  TD* __restrict__ Pp = (TD*)p.pre;
#pragma unroll 1
  for (int j = 0; j < 8; ++j) {
    const int r = j * 4 + rsub;
    const int m = mw + r;
    if (m >= p.M) break;
    const float4 x = lds128(stage + epi_slot(r, g));
#pragma unroll 1
    for (int e = 0; e < 4 && n + e < p.N; ++e) {
      float y = e == 0 ? x.x : (e == 1 ? x.y : (e == 2 ? x.z : x.w));
      if (p.bias) y += p.bias[n + e];
      if (Pp) stf(Pp + (long long)m * p.ldp + n + e, y);
      y = svpc_act(y, p.act);
      if (dropping) y *= svpc_drop_scale(dkey, p.drop.p, dinv, (unsigned long long)m * p.N + n + e);
      if (Rp) y += ldf(Rp + (long long)m * p.ldr + n + e);
      TD* d = Dp + (long long)m * p.ldd + n + e;
      if (p.accumulate) y += ldf(d);
      if (p.gref) y *= tc_act_grad(p.gact, ldf((const TD*)p.gref + (long long)m * p.ldg + n + e));
      stf(d, y);
    }
  }
}

// split-K partial sums: fp32 reductions into D (plain epilogue only)
__device__ __forceinline__ void epi_atomic(const TcParams& p, uint32_t stage, int mw, int n, int rsub, int g, bool vec) {
  float* Dp = reinterpret_cast<float*>(p.D);
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const int r = j * 4 + rsub;
    const int m = mw + r;
    if (m >= p.M) break;
    const float4 x = lds128(stage + epi_slot(r, g));
    float* d = Dp + (long long)m * p.ldd + n;
    if (vec) {
      red_add_f32x4(d, x);
    } else {
      red_add_f32(d, x.x);
      if (n + 1 < p.N) red_add_f32(d + 1, x.y);
      if (n + 2 < p.N) red_add_f32(d + 2, x.z);
      if (n + 3 < p.N) red_add_f32(d + 3, x.w);
    }
  }
}

// Persistent kernel, one CTA per SM.  Shared memory: TC_STAGES (4) x (A 128x64 + B 256x64) bf16 ring (48 KB per stage),
// one 32 x 32 fp32 transposition buffer per epilogue warp (8 x 4 KB), barriers.  TMEM: 2 accumulators of 256 fp32 columns
// (all 512 columns), so the epilogue of unit i drains one while the MMAs of unit i+1 fill the other.
// ACT / DROP: the epilogue is specialised at compile time (a runtime switch over activations inside the row loop made
// the epilogue, not the MMAs, the bottleneck of the K = 768 GEMMs).
template <bool A_MN, bool B_MN, typename TD, int ACT, bool DROP>
__global__ void __launch_bounds__(TC_THREADS, 1) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB,
                                                                const __grid_constant__ CUtensorMap tmAlo,
                                                                const __grid_constant__ CUtensorMap tmBlo,
                                                                const __grid_constant__ TcParams p) {
  // Programmatic dependent launch: the NEXT grid may become resident right away, and this CTA's prologue (barrier
  // initialisation, the TMEM allocation, the tensor-map prefetch - ~2 us that touch no global data) runs while the grid in
  // front of it is still draining; griddepcontrol.wait below is the point after which its results may be read.
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  float* epi_stage = reinterpret_cast<float*>(smem + TC_STAGES * TC_STAGE_BYTES);
  uint64_t* full_bar = (uint64_t*)(smem + TC_STAGES * TC_STAGE_BYTES + TC_EPI_BYTES);
  uint64_t* empty_bar = full_bar + TC_MAX_STAGES;
  uint64_t* tmem_full = empty_bar + TC_MAX_STAGES;   // [2]
  uint64_t* tmem_empty = tmem_full + 2;              // [2]
  uint32_t* tmem_slot = (uint32_t*)(tmem_empty + 2);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmB) : "memory");
    if (p.nseg > 1) {
      asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmAlo) : "memory");
      asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmBlo) : "memory");
    }
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < p.nstages; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    for (int a = 0; a < 2; ++a) { mbar_init(tmem_full + a, 1); mbar_init(tmem_empty + a, TC_EPI_WARPS); }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) tmem_alloc(tmem_slot, 512);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  asm volatile("griddepcontrol.wait;" ::: "memory");

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      uint32_t s = 0, ph = 0;                            // ring slot and its phase
      for (int u = blockIdx.x; u < p.total_units; u += gridDim.x) {
        const TcUnit w = tc_unit(p, u);
        if (w.n0 >= p.N) continue;
        const uint32_t bytes = TC_A_BYTES + (uint32_t)w.width * (TC_BK * 2);
        const int nbox = w.width >> 6;
        // split-fp32 operands: the K loop runs three times over the same accumulator - (A_hi, B_hi), (A_lo, B_hi), (A_hi, B_lo)
        for (int seg = 0; seg < p.nseg; ++seg) {
          const CUtensorMap* ma = seg == 1 ? &tmAlo : &tmA;
          const CUtensorMap* mb = seg == 2 ? &tmBlo : &tmB;
          for (int kb = w.kb0; kb < w.kb1; ++kb) {
            mbar_wait(empty_bar + s, ph ^ 1);
            uint8_t* sa = smem + s * p.stage_bytes;
            uint8_t* sb = sa + TC_A_BYTES;
            mbar_expect_tx(full_bar + s, bytes);
            const int k = kb * TC_BK;
            if (A_MN) {
#pragma unroll
              for (int j = 0; j < TC_BM / 64; ++j) tma_load_2d(sa + j * (64 * TC_BK * 2), ma, full_bar + s, w.m0 + j * 64, k);
            } else {
              tma_load_2d(sa, ma, full_bar + s, k, w.m0);
            }
            // B in 64-column boxes (8 KB each): K-major boxes stack into one [width][64] swizzled tile, MN-major boxes are
            // the 64-wide N blocks the descriptor strides over
            for (int j = 0; j < nbox; ++j) {
              if (B_MN) tma_load_2d(sb + j * (64 * TC_BK * 2), mb, full_bar + s, w.n0 + j * 64, k);
              else tma_load_2d(sb + j * (64 * TC_BK * 2), mb, full_bar + s, k, w.n0 + j * 64);
            }
            if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; }
          }
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      uint32_t s = 0, ph = 0, ui = 0;
      for (int u = blockIdx.x; u < p.total_units; u += gridDim.x) {
        const TcUnit w = tc_unit(p, u);
        if (w.n0 >= p.N) continue;
        const uint32_t acc = ui & 1, aph = (ui >> 1) & 1;
        ++ui;
        mbar_wait(tmem_empty + acc, aph ^ 1);            // epilogue has drained this accumulator
        tc_fence_after();
        // instruction descriptor: fp32 accumulate, bf16 x bf16, majors, N>>3, M>>4 (cute::UMMA::InstrDescriptor)
        const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                               ((uint32_t)(w.width >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
        const uint32_t tacc = tmem_base + acc * TC_BN_MAX;
        const int nkb = (w.kb1 - w.kb0) * p.nseg;
        for (int kq = 0; kq < nkb; ++kq) {
          mbar_wait(full_bar + s, ph);
          tc_fence_after();
          const uint32_t sa = smem_u32(smem + s * p.stage_bytes);
          const uint32_t sb = sa + TC_A_BYTES;
#pragma unroll
          for (int k = 0; k < TC_BK / TC_UMMA_K; ++k) {
            // K-major: 8-row groups 1024 B apart, +32 B per K step inside the swizzle row.
            // MN-major: 64-wide MN blocks (TC_BK * 128 B apart, LBO), 8-k groups 1024 B apart (SBO), +2048 B per K step.
            const uint64_t ad = A_MN ? make_sdesc(sa + k * (TC_UMMA_K * 128), TC_BK * 128, 1024)
                                     : make_sdesc(sa + k * (TC_UMMA_K * 2), 16, 1024);
            const uint64_t bd = B_MN ? make_sdesc(sb + k * (TC_UMMA_K * 128), TC_BK * 128, 1024)
                                     : make_sdesc(sb + k * (TC_UMMA_K * 2), 16, 1024);
            umma_f16(tacc, ad, bd, idesc, (kq > 0 || k > 0) ? 1u : 0u);
          }
          umma_commit(empty_bar + s);      // ring slot free once these MMAs have read it
          if (++s == (uint32_t)p.nstages) { s = 0; ph ^= 1; }
        }
        umma_commit(tmem_full + acc);      // accumulator complete
      }
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2 .. 2 + TC_EPI_WARPS)
    // Two warps per TMEM lane quarter, each taking every other 32-column chunk.  Per chunk: TMEM -> registers
    // (lane = row) -> swizzled shared memory -> re-read with lanes along the columns, so bias / activation / dropout /
    // residual / accumulate and the stores are row-contiguous (4 rows x 128 B (fp32) or 64 B (bf16) per warp
    // instruction); the other warp of the scheduler covers the TMEM / memory latencies.
    const int q = warp & 3;              // TMEM lane quarter this warp may read
    const int ew = warp - 2;             // epilogue warp index
    const int half = ew >> 2;            // which chunks of the quarter: half, half + 2, ...
    constexpr int CSTEP = TC_EPI_WARPS / 4;
    const uint32_t stage = smem_u32(epi_stage) + (uint32_t)ew * (32 * 32 * 4);
    const bool vec_all = (p.ldd % 4 == 0) && (((uintptr_t)p.D & 15) == 0) &&
                         (!p.residual || (p.ldr % 4 == 0 && ((uintptr_t)p.residual & 15) == 0)) &&
                         (!p.gref || (p.ldg % 4 == 0 && ((uintptr_t)p.gref & 15) == 0)) &&
                         (!p.pre || (p.ldp % 4 == 0 && ((uintptr_t)p.pre & 15) == 0)) && (((uintptr_t)p.bias & 15) == 0);
    const bool dropping = DROP;
    const uint32_t dkey = DROP ? svpc_drop_key(p.drop) : 0u;
    const float dinv = dropping ? 1.f / (1.f - p.drop.p) : 1.f;
    const int g = lane & 7;              // this lane's column group (4 columns) inside a chunk
    const int rsub = lane >> 3;          // row inside a 4-row pass
    uint32_t ui = 0;
    for (int u = blockIdx.x; u < p.total_units; u += gridDim.x) {
      const TcUnit w = tc_unit(p, u);
      if (w.n0 >= p.N) continue;
      const uint32_t acc = ui & 1, aph = (ui >> 1) & 1;
      ++ui;
      mbar_wait(tmem_full + acc, aph);
      tc_fence_after();
      const int mw = w.m0 + q * 32;
      const int nchunks = w.width >> 5;
      const uint32_t trow = tmem_base + acc * TC_BN_MAX + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int c = half; c < nchunks; c += CSTEP) {
        const int nb = w.n0 + c * 32;
        const int n = nb + g * 4;
        const bool in_range = n < p.N && mw < p.M;
        const bool vec = vec_all && (n + 4 <= p.N);
        EpiPre opnd;
        if (sizeof(TD) == 2 && in_range && vec && !p.atomic) epi_prefetch_bf16(p, mw, n, rsub, opnd);
        {
          uint32_t v[32];
          tmem_ld32(trow + (uint32_t)(c * 32), v);
          tmem_ld_wait();
#pragma unroll
          for (int jj = 0; jj < 8; ++jj)
            sts128(stage + epi_slot(lane, jj), v[4 * jj], v[4 * jj + 1], v[4 * jj + 2], v[4 * jj + 3]);
        }
        __syncwarp();
        if (in_range) {
          if (p.atomic) {
            epi_atomic(p, stage, mw, n, rsub, g, vec);
          } else if (vec) {
            if (sizeof(TD) == 2) epi_vec_bf16<ACT, DROP>(p, stage, mw, n, rsub, g, dkey, dinv, opnd);
            else epi_vec<ACT, DROP, TD>(p, stage, mw, n, rsub, g, dkey, dinv);
          } else {
            epi_scalar<TD>(p, stage, mw, n, rsub, g, dropping, dkey, dinv);
          }
        }
        __syncwarp();                      // the transposition buffer is rewritten by the next chunk
      }
      // every column this warp reads of the accumulator has left TMEM (the last tmem_ld_wait above): hand it back
      tc_fence_before();
      __syncwarp();
      if (lane == 0) mbar_arrive(tmem_empty + acc);
    }
  }
  tc_fence_before();
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 512);
  }
}

// ------------------------------------------------------------------------------------------------ host side
typedef CUresult (*PFN_encodeTiled)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                                    const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static PFN_encodeTiled get_encode() {
  static PFN_encodeTiled fn = nullptr;
  if (!fn) {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = (PFN_encodeTiled)p;
  }
  return fn;
}

// Encoded tensor maps are cached (direct-mapped, keyed on every argument of the encoding): on the eager / ragged path the
// same few hundred operand geometries recur every step and cuTensorMapEncodeTiled costs ~1 us of host time per call.
struct MapKey { const void* base; long long inner, outer, ld; int bi, bo; };
struct MapSlot { MapKey k; CUtensorMap m; bool used; };
#define TC_MAP_SLOTS 2048
static MapSlot* map_cache() {
  static MapSlot* c = (MapSlot*)calloc(TC_MAP_SLOTS, sizeof(MapSlot));
  return c;
}

// 2-D bf16 tensor map: inner (contiguous) extent `inner`, outer extent `outer`, outer stride `ld` elements
static int make_map(CUtensorMap* m, const void* base, long long inner, long long outer, long long ld, int box_inner,
                    int box_outer) {
  MapKey key;
  memset(&key, 0, sizeof(key));
  key.base = base; key.inner = inner; key.outer = outer; key.ld = ld; key.bi = box_inner; key.bo = box_outer;
  unsigned long long h = (unsigned long long)(uintptr_t)base * 0x9E3779B97F4A7C15ull ^ (unsigned long long)inner * 0xC2B2AE3D27D4EB4Full ^
                         (unsigned long long)outer * 0x165667B19E3779F9ull ^ (unsigned long long)ld * 0x27D4EB2F165667C5ull ^
                         (unsigned long long)(box_inner * 131 + box_outer);
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  MapSlot* slot = map_cache() ? map_cache() + ((h >> 20) % TC_MAP_SLOTS) : nullptr;
  if (slot && slot->used && memcmp(&slot->k, &key, sizeof(key)) == 0) { *m = slot->m; return SVPC_OK; }
  PFN_encodeTiled enc = get_encode();
  if (!enc) { svpc_set_error("cuTensorMapEncodeTiled not available"); return SVPC_ERR_CUDA; }
  cuuint64_t dims[2] = {(cuuint64_t)inner, (cuuint64_t)outer};
  cuuint64_t strides[1] = {(cuuint64_t)ld * 2};
  cuuint32_t box[2] = {(cuuint32_t)box_inner, (cuuint32_t)box_outer};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = enc(m, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                   CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                   CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) { svpc_set_error("cuTensorMapEncodeTiled failed (%d)", (int)r); return SVPC_ERR_CUDA; }
  if (slot) { slot->k = key; slot->m = *m; slot->used = true; }
  return SVPC_OK;
}

static bool operand_ok(long long s_mn, long long s_k, int mn, int K) {
  // one unit stride, the other a multiple of 8 elements (16 B) as TMA requires
  if (s_k == 1) return s_mn % 8 == 0 && s_mn >= K;
  if (s_mn == 1) return s_k % 8 == 0 && s_k >= mn;
  return false;
}

int svpc_gemm_tc_eligible(int M, int N, int K, int dtA, long long sa_m, long long sa_k, int dtB, long long sb_n,
                          long long sb_k, int dtD, long long ldd) {
  if (dtA != SVPC_BF16 || dtB != SVPC_BF16) return 0;
  if (M < 16 || N < 16 || K < 32) return 0;        // (rows past M are zero-filled by TMA and never stored)
  if (!operand_ok(sa_m, sa_k, M, K) || !operand_ok(sb_n, sb_k, N, K)) return 0;
  (void)dtD; (void)ldd;
  return 1;
}

template <bool A_MN, bool B_MN, typename TD, int ACT, bool DROP>
static int launch_cfg(const CUtensorMap* tm, const TcParams& p, int grid, cudaStream_t st) {
  auto kern = gemm_tc_kernel<A_MN, B_MN, TD, ACT, DROP>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)TC_SMEM_BYTES);
    attr_set = true;
  }
  svpc_launch(kern, grid, TC_THREADS, TC_SMEM_BYTES, st, tm[0], tm[1], tm[2], tm[3], p);
  return svpc_check_launch("gemm_tc");
}

template <bool A_MN, bool B_MN, typename TD>
static int launch_epi(const CUtensorMap* tm, const TcParams& p, int grid, cudaStream_t st) {
  const bool drop = p.drop.seed != nullptr && p.drop.p > 0.f;
#define TC_CASE(A) \
  case A: return drop ? launch_cfg<A_MN, B_MN, TD, A, true>(tm, p, grid, st) : launch_cfg<A_MN, B_MN, TD, A, false>(tm, p, grid, st);
  switch (p.act) {
    TC_CASE(SVPC_ACT_RELU)
    TC_CASE(SVPC_ACT_GELU)
    TC_CASE(SVPC_ACT_SIGMOID)
    default: break;
  }
  return drop ? launch_cfg<A_MN, B_MN, TD, SVPC_ACT_NONE, true>(tm, p, grid, st)
              : launch_cfg<A_MN, B_MN, TD, SVPC_ACT_NONE, false>(tm, p, grid, st);
#undef TC_CASE
}

template <typename TD>
static int launch_major(bool a_mn, bool b_mn, const CUtensorMap* tm, const TcParams& p, int grid, cudaStream_t st) {
  if (!a_mn && !b_mn) return launch_epi<false, false, TD>(tm, p, grid, st);
  if (!a_mn && b_mn) return launch_epi<false, true, TD>(tm, p, grid, st);
  if (a_mn && !b_mn) return launch_epi<true, false, TD>(tm, p, grid, st);
  return launch_epi<true, true, TD>(tm, p, grid, st);
}

// tile width: the padded width ceil(N / BN) * BN, weighted by how much worse a narrower UMMA uses the operand bandwidth
static int choose_bn(int N) {
  const int cand[3] = {256, 128, 64};
  const double pen[3] = {1.0, 1.05, 1.25};
  int best = 256;
  double best_cost = 1e30;
  for (int i = 0; i < 3; ++i) {
    const double c = (double)svpc_cdiv(N, cand[i]) * cand[i] * pen[i];
    if (c < best_cost) { best_cost = c; best = cand[i]; }
  }
  return best;
}

// A_lo / B_lo: the low-order bf16 planes of split-fp32 operands (same strides as A / B), or NULL for a plain bf16 GEMM.
int svpc_gemm_tc_launch(int M, int N, int K, const void* A, long long sa_m, long long sa_k, const void* B, long long sb_n,
                        long long sb_k, void* D, int dtD, long long ldd, const float* bias, int act, const void* residual,
                        long long ldr, void* pre, long long ldp, int accumulate, SvpcDrop drop, cudaStream_t st,
                        const void* A_lo, const void* B_lo, const void* gref, long long ldg, int gact) {
  const bool a_mn = sa_k != 1, b_mn = sb_k != 1;
  // tuning switches (read once): SVPC_TC_BN forces the tile width, SVPC_TC_SPLIT_MAX caps the split-K factor
  static int env_bn = -1, env_split_max = -1;
  if (env_bn < 0) {
    const char* e = getenv("SVPC_TC_BN");
    env_bn = e ? atoi(e) : 0;
    e = getenv("SVPC_TC_SPLIT_MAX");
    env_split_max = e ? atoi(e) : 32;
  }
  const int BN = (env_bn == 64 || env_bn == 128 || env_bn == 256) ? env_bn : choose_bn(N);
  const int nseg = (A_lo && B_lo) ? 3 : 1;
  CUtensorMap tm[4];
  int rc;
  const void* ops[4] = {A, B, nseg == 3 ? A_lo : A, nseg == 3 ? B_lo : B};
  for (int i = 0; i < 4; i += 2) {
    if (a_mn) rc = make_map(&tm[i], ops[i], M, K, sa_k, 64, TC_BK); else rc = make_map(&tm[i], ops[i], K, M, sa_m, TC_BK, TC_BM);
    if (rc) return rc;
    if (b_mn) rc = make_map(&tm[i + 1], ops[i + 1], N, K, sb_k, 64, TC_BK); else rc = make_map(&tm[i + 1], ops[i + 1], K, N, sb_n, TC_BK, 64);
    if (rc) return rc;
    if (nseg == 1) { tm[2] = tm[0]; tm[3] = tm[1]; break; }
  }
  const int G = SVPC_NUM_SMS;
  const int tiles_m = svpc_cdiv(M, TC_BM), tiles_n = svpc_cdiv(N, BN), tiles = tiles_m * tiles_n;
  const int kb_total = svpc_cdiv(K, TC_BK);
  SVPC_REQUIRE(kb_total > 0, "gemm_tc: K == 0");
  int split = 1;
  const bool plain = !bias && act == SVPC_ACT_NONE && !residual && !pre && dtD == SVPC_F32 && !drop.seed && !gref;
  if (plain && tiles * 2 <= G && kb_total >= 16) {
    // split-K: rounds of G units x K blocks per unit, plus ~4 K-block-times of atomic epilogue per unit
    long long best = (long long)svpc_cdiv(tiles, G) * (kb_total + 4);
    for (int sp = 2; sp <= env_split_max && kb_total / sp >= 8; ++sp) {
      const long long c = (long long)svpc_cdiv((long long)tiles * sp, G) * (svpc_cdiv(kb_total, sp) + 4);
      if (c < best) { best = c; split = sp; }
    }
  }
  const int kb_per = svpc_cdiv(kb_total, split);
  split = svpc_cdiv(kb_total, kb_per);
  int main_tiles = tiles, tail_s = 1;
  if (split == 1) {
    const int rem = tiles % G;
    if (rem > 0) {
      while (tail_s * 2 <= BN / 64 && (long long)rem * tail_s * 2 <= G) tail_s *= 2;
      if (tail_s > 1) main_tiles = tiles - rem;
    }
  }
  const int total_units = main_tiles * split + (tiles - main_tiles) * tail_s;
  TcParams p{M, N, K, D, ldd, bias, act, residual, ldr, pre, ldp, accumulate, split > 1 ? 1 : 0, drop, gref, ldg, gact,
             BN, tiles_n, split, kb_per, kb_total, main_tiles, tail_s, total_units, nseg, 0, 0};
  {
    // the 192 KB ring holds 4 stages of a 256-wide tile, 6 of a 128-wide, 8 of a 64-wide one: a launch of narrow tiles
    // (the 192-row step-level modules, K = 768 = 12 blocks) then has most of its K loop in flight at once
    const int stage_w = main_tiles > 0 ? BN : BN / tail_s;
    p.stage_bytes = TC_A_BYTES + stage_w * (TC_BK * 2);
    p.nstages = svpc_imin(TC_MAX_STAGES, (TC_STAGES * TC_STAGE_BYTES) / p.stage_bytes);
  }
  if (split > 1 && !accumulate) {
    cudaError_t e = cudaMemset2DAsync(D, (size_t)ldd * 4, 0, (size_t)N * 4, (size_t)M, st);
    if (e != cudaSuccess) { svpc_set_error("gemm_tc: memset2D: %s", cudaGetErrorString(e)); return SVPC_ERR_CUDA; }
  }
  const int grid = svpc_imin(total_units, G);
  if (dtD == SVPC_F32) return launch_major<float>(a_mn, b_mn, tm, p, grid, st);
  if (dtD == SVPC_BF16) return launch_major<bf16>(a_mn, b_mn, tm, p, grid, st);
  svpc_set_error("gemm_tc: bad output dtype %d", dtD);
  return SVPC_ERR_ARG;
}
