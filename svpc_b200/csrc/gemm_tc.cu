// tcgen05 / TMEM / TMA GEMM for the dense contractions of the SVPC step (bf16 operands, fp32 accumulate in TMEM).
//   D[M,N] = act(A[M,K] . B[N,K]^T + bias) + residual   (optionally += D, optionally a copy of the pre-activation)
// Both operands may be K-major (row-major with K contiguous: activations x weights^T, the forward Linear) or MN-major
// (the M / N index contiguous: dgrad uses W as an MN-major B, wgrad uses dY and X as MN-major A and B), so no
// transposed copies of weights or activations are ever materialised.
//
// Structure (one 128 x BN output tile per CTA, 192 threads):
//   warp 0      TMA producer: cp.async.bulk.tensor 2D boxes, 128B swizzle, into a KSTAGES-deep shared-memory ring,
//               completion on mbarriers (expect_tx)
//   warp 1      MMA issuer: one elected lane issues tcgen05.mma.cta_group::1.kind::f16 (UMMA 128 x BN x 16) from
//               shared-memory descriptors into a TMEM accumulator; tcgen05.commit releases ring slots / signals the
//               epilogue
//   warps 2-5   epilogue: tcgen05.ld (32 lanes x 32 columns per warp) -> bias / activation / residual -> global;
//               warp 2 also owns the TMEM allocation
// Split-K (grid.z) with fp32 red.global for weight-gradient shapes whose output has too few tiles to fill 148 SMs.
#include "common.cuh"
#include <cuda.h>

#define TC_BM 128
#define TC_BK 64                 // 64 bf16 = 128 B = one swizzle row
#define TC_THREADS 192
#define TC_UMMA_K 16

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
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}

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
  int kb_per_split;        // K blocks (of TC_BK) per grid.z slice
  void* D; long long ldd;
  const float* bias; int act;
  const void* residual; long long ldr;
  void* pre; long long ldp;
  int accumulate;          // D += result
  int atomic;              // split-K: red.global.add.f32
  SvpcDrop drop;           // dropout after the activation, before the residual (model.py:231, 287, 757)
};

template <int BN, bool A_MN, bool B_MN, int KSTAGES, typename TD>
__global__ void __launch_bounds__(TC_THREADS, 2) gemm_tc_kernel(const __grid_constant__ CUtensorMap tmA,
                                                                const __grid_constant__ CUtensorMap tmB,
                                                                const TcParams p) {
  extern __shared__ __align__(1024) uint8_t smem_raw[];
  constexpr uint32_t A_BYTES = TC_BM * TC_BK * 2;   // 16 KB
  constexpr uint32_t B_BYTES = BN * TC_BK * 2;
  constexpr uint32_t STAGE_BYTES = A_BYTES + B_BYTES;
  uint8_t* smem = (uint8_t*)(((uintptr_t)smem_raw + 1023) & ~(uintptr_t)1023);
  uint64_t* full_bar = (uint64_t*)(smem + KSTAGES * STAGE_BYTES);
  uint64_t* empty_bar = full_bar + KSTAGES;
  uint64_t* tmem_full = empty_bar + KSTAGES;
  uint32_t* tmem_slot = (uint32_t*)(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int m0 = blockIdx.y * TC_BM, n0 = blockIdx.x * BN;
  const int kb_total = (p.K + TC_BK - 1) / TC_BK;
  const int kb0 = blockIdx.z * p.kb_per_split;
  const int kb1 = min(kb_total, kb0 + p.kb_per_split);
  const int nkb = kb1 - kb0;

  if (warp == 0 && lane == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmA) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"((uint64_t)&tmB) : "memory");
  }
  if (warp == 1 && lane == 0) {
    for (int s = 0; s < KSTAGES; ++s) { mbar_init(full_bar + s, 1); mbar_init(empty_bar + s, 1); }
    mbar_init(tmem_full, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 2) tmem_alloc(tmem_slot, BN);
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;

  if (warp == 0) {
    // ------------------------------------------------------------------ TMA producer
    if (lane == 0) {
      for (int i = 0; i < nkb; ++i) {
        const int s = i % KSTAGES, ph = (i / KSTAGES) & 1;
        mbar_wait(empty_bar + s, ph ^ 1);
        uint8_t* sa = smem + s * STAGE_BYTES;
        uint8_t* sb = sa + A_BYTES;
        mbar_expect_tx(full_bar + s, STAGE_BYTES);
        const int k = (kb0 + i) * TC_BK;
        if (A_MN) {
#pragma unroll
          for (int j = 0; j < TC_BM / 64; ++j) tma_load_2d(sa + j * (64 * TC_BK * 2), &tmA, full_bar + s, m0 + j * 64, k);
        } else {
          tma_load_2d(sa, &tmA, full_bar + s, k, m0);
        }
        if (B_MN) {
#pragma unroll
          for (int j = 0; j < BN / 64; ++j) tma_load_2d(sb + j * (64 * TC_BK * 2), &tmB, full_bar + s, n0 + j * 64, k);
        } else {
          tma_load_2d(sb, &tmB, full_bar + s, k, n0);
        }
      }
    }
  } else if (warp == 1) {
    // ------------------------------------------------------------------ MMA issuer
    if (lane == 0) {
      // instruction descriptor: fp32 accumulate, bf16 x bf16, majors, N>>3, M>>4 (cute::UMMA::InstrDescriptor)
      const uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((A_MN ? 1u : 0u) << 15) | ((B_MN ? 1u : 0u) << 16) |
                             ((uint32_t)(BN >> 3) << 17) | ((uint32_t)(TC_BM >> 4) << 24);
      for (int i = 0; i < nkb; ++i) {
        const int s = i % KSTAGES, ph = (i / KSTAGES) & 1;
        mbar_wait(full_bar + s, ph);
        tc_fence_after();
        const uint32_t sa = smem_u32(smem + s * STAGE_BYTES);
        const uint32_t sb = sa + A_BYTES;
#pragma unroll
        for (int k = 0; k < TC_BK / TC_UMMA_K; ++k) {
          // K-major: 8-row groups 1024 B apart, +32 B per K step inside the swizzle row.
          // MN-major: 64-wide MN blocks (TC_BK * 128 B apart, LBO), 8-k groups 1024 B apart (SBO), +2048 B per K step.
          const uint64_t ad = A_MN ? make_sdesc(sa + k * (TC_UMMA_K * 128), TC_BK * 128, 1024)
                                   : make_sdesc(sa + k * (TC_UMMA_K * 2), 16, 1024);
          const uint64_t bd = B_MN ? make_sdesc(sb + k * (TC_UMMA_K * 128), TC_BK * 128, 1024)
                                   : make_sdesc(sb + k * (TC_UMMA_K * 2), 16, 1024);
          umma_f16(tmem_base, ad, bd, idesc, (i > 0 || k > 0) ? 1u : 0u);
        }
        umma_commit(empty_bar + s);      // ring slot free once these MMAs have read it
      }
      umma_commit(tmem_full);            // accumulator complete
    }
  } else {
    // ------------------------------------------------------------------ epilogue (warps 2..5)
    // Phase 1: TMEM -> registers (lane = row) -> shared memory (the ring is free once tmem_full has fired).
    // Phase 2: each warp re-reads its 32 rows with lanes along the columns, so bias / activation / residual /
    //          accumulate and the stores are row-contiguous: 512 B (fp32) or 256 B (bf16) per warp instruction.
    const int q = warp & 3;              // TMEM lane quarter this warp may read
    if (nkb > 0) {
      mbar_wait(tmem_full, 0);
      tc_fence_after();
    }
    constexpr int LDS = BN + 4;          // padded row stride (floats): conflict-free float4 writes by 8-lane phases
    float* stage = reinterpret_cast<float*>(smem) + (size_t)q * 32 * LDS;
#pragma unroll 1
    for (int c = 0; c < BN / 32; ++c) {
      uint32_t v[32];
      if (nkb > 0) {
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(c * 32), v);
      } else {
#pragma unroll
        for (int jj = 0; jj < 32; ++jj) v[jj] = 0u;
      }
      float* dst = stage + (size_t)lane * LDS + c * 32;
#pragma unroll
      for (int jj = 0; jj < 32; jj += 4)
        *reinterpret_cast<uint4*>(dst + jj) = make_uint4(v[jj], v[jj + 1], v[jj + 2], v[jj + 3]);
    }
    tc_fence_before();
    __syncwarp();
    TD* __restrict__ Dp = (TD*)p.D;
    const TD* __restrict__ Rp = (const TD*)p.residual;
    TD* __restrict__ Pp = (TD*)p.pre;
    const int mw = m0 + q * 32;
    constexpr int VEC_ROWS = 128 / BN;   // rows covered by one warp pass in the vector path (1 for BN=128, 2 for BN=64)
    const bool vec_ok = !p.atomic && (n0 + BN <= p.N) && (p.ldd % 4 == 0) && (((uintptr_t)Dp & 15) == 0) &&
                        (!Rp || (p.ldr % 4 == 0 && ((uintptr_t)Rp & 15) == 0)) &&
                        (!Pp || (p.ldp % 4 == 0 && ((uintptr_t)Pp & 15) == 0)) && (((uintptr_t)p.bias & 15) == 0);
    const bool dropping = p.drop.seed != nullptr && p.drop.p > 0.f;
    const uint32_t dkey = svpc_drop_key(p.drop);
    const float dinv = dropping ? 1.f / (1.f - p.drop.p) : 1.f;
    if (vec_ok) {
      const int cl = (lane % (BN / 4)) * 4;            // this lane's 4 columns inside the tile
      const int rsub = lane / (BN / 4);                // row inside the pass
      float4 bv = make_float4(0.f, 0.f, 0.f, 0.f);
      if (p.bias) bv = *reinterpret_cast<const float4*>(p.bias + n0 + cl);
#pragma unroll 4
      for (int r = rsub; r < 32; r += VEC_ROWS) {
        const int m = mw + r;
        if (m >= p.M) break;
        float4 x = *reinterpret_cast<const float4*>(stage + (size_t)r * LDS + cl);
        x.x += bv.x; x.y += bv.y; x.z += bv.z; x.w += bv.w;
        if (Pp) st4(Pp + (long long)m * p.ldp + n0 + cl, x);
        x.x = svpc_act(x.x, p.act); x.y = svpc_act(x.y, p.act); x.z = svpc_act(x.z, p.act); x.w = svpc_act(x.w, p.act);
        if (dropping) {
          const unsigned long long e0 = (unsigned long long)m * p.N + n0 + cl;
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
        if (Rp) { const float4 rr = ld4(Rp + (long long)m * p.ldr + n0 + cl); x.x += rr.x; x.y += rr.y; x.z += rr.z; x.w += rr.w; }
        TD* d = Dp + (long long)m * p.ldd + n0 + cl;
        if (p.accumulate) { const float4 dd = ld4(d); x.x += dd.x; x.y += dd.y; x.z += dd.z; x.w += dd.w; }
        st4(d, x);
      }
    } else {
      for (int r = 0; r < 32; ++r) {
        const int m = mw + r;
        if (m >= p.M) break;
#pragma unroll
        for (int cc = 0; cc < BN / 32; ++cc) {
          const int n = n0 + cc * 32 + lane;
          if (n >= p.N) continue;
          float x = stage[(size_t)r * LDS + cc * 32 + lane];
          if (p.atomic) {
            atomicAdd(reinterpret_cast<float*>(Dp) + (long long)m * p.ldd + n, x);
          } else {
            if (p.bias) x += p.bias[n];
            if (Pp) stf(Pp + (long long)m * p.ldp + n, x);
            x = svpc_act(x, p.act);
            if (dropping) x *= svpc_drop_scale(dkey, p.drop.p, dinv, (unsigned long long)m * p.N + n);
            if (Rp) x += ldf(Rp + (long long)m * p.ldr + n);
            TD* d = Dp + (long long)m * p.ldd + n;
            if (p.accumulate) x += ldf(d);
            stf(d, x);
          }
        }
      }
    }
  }
  __syncthreads();
  if (warp == 2) {
    tc_fence_after();
    tmem_dealloc(tmem_base, BN);
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

// 2-D bf16 tensor map: inner (contiguous) extent `inner`, outer extent `outer`, outer stride `ld` elements
static int make_map(CUtensorMap* m, const void* base, long long inner, long long outer, long long ld, int box_inner,
                    int box_outer) {
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
  if (M < 64 || N < 16 || K < 32) return 0;
  if (!operand_ok(sa_m, sa_k, M, K) || !operand_ok(sb_n, sb_k, N, K)) return 0;
  (void)dtD; (void)ldd;
  return 1;
}

template <int BN, bool A_MN, bool B_MN, typename TD>
static int launch_cfg(const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, dim3 grid, cudaStream_t st) {
  constexpr int KSTAGES = BN == 128 ? 3 : 4;   // 96 KB per CTA: two CTAs per SM, one's epilogue under the other's main loop
  constexpr size_t smem = (size_t)KSTAGES * (TC_BM * TC_BK * 2 + BN * TC_BK * 2) + 1024 + 256;
  auto kern = gemm_tc_kernel<BN, A_MN, B_MN, KSTAGES, TD>;
  static bool attr_set = false;
  if (!attr_set) {
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    cudaFuncSetAttribute(kern, cudaFuncAttributePreferredSharedMemoryCarveout, cudaSharedmemCarveoutMaxShared);
    attr_set = true;
  }
  kern<<<grid, TC_THREADS, smem, st>>>(ta, tb, p);
  return svpc_check_launch("gemm_tc");
}

template <int BN, typename TD>
static int launch_major(bool a_mn, bool b_mn, const CUtensorMap& ta, const CUtensorMap& tb, const TcParams& p, dim3 grid,
                        cudaStream_t st) {
  if (!a_mn && !b_mn) return launch_cfg<BN, false, false, TD>(ta, tb, p, grid, st);
  if (!a_mn && b_mn) return launch_cfg<BN, false, true, TD>(ta, tb, p, grid, st);
  if (a_mn && !b_mn) return launch_cfg<BN, true, false, TD>(ta, tb, p, grid, st);
  return launch_cfg<BN, true, true, TD>(ta, tb, p, grid, st);
}

int svpc_gemm_tc_launch(int M, int N, int K, const void* A, long long sa_m, long long sa_k, const void* B, long long sb_n,
                        long long sb_k, void* D, int dtD, long long ldd, const float* bias, int act, const void* residual,
                        long long ldr, void* pre, long long ldp, int accumulate, SvpcDrop drop, cudaStream_t st) {
  const bool a_mn = sa_k != 1, b_mn = sb_k != 1;
  const int BN = (N <= 64) ? 64 : 128;
  CUtensorMap ta, tb;
  int rc;
  if (a_mn) rc = make_map(&ta, A, M, K, sa_k, 64, TC_BK); else rc = make_map(&ta, A, K, M, sa_m, TC_BK, TC_BM);
  if (rc) return rc;
  if (b_mn) rc = make_map(&tb, B, N, K, sb_k, 64, TC_BK); else rc = make_map(&tb, B, K, N, sb_n, TC_BK, BN);
  if (rc) return rc;
  const int tiles_m = svpc_cdiv(M, TC_BM), tiles_n = svpc_cdiv(N, BN);
  const int kb_total = svpc_cdiv(K, TC_BK);
  int split = 1;
  const bool plain = !bias && act == SVPC_ACT_NONE && !residual && !pre && dtD == SVPC_F32 && !drop.seed;
  if (plain && tiles_m * tiles_n < SVPC_NUM_SMS / 2 && kb_total >= 16) {
    split = svpc_imin(svpc_imin(SVPC_NUM_SMS / (tiles_m * tiles_n), kb_total / 8), 32);
    if (split < 1) split = 1;
  }
  int kb_per = svpc_cdiv(kb_total, split);
  split = svpc_cdiv(kb_total, kb_per);
  TcParams p{M, N, K, kb_per, D, ldd, bias, act, residual, ldr, pre, ldp, accumulate, split > 1 ? 1 : 0, drop};
  if (split > 1 && !accumulate) {
    cudaError_t e = cudaMemset2DAsync(D, (size_t)ldd * 4, 0, (size_t)N * 4, (size_t)M, st);
    if (e != cudaSuccess) { svpc_set_error("gemm_tc: memset2D: %s", cudaGetErrorString(e)); return SVPC_ERR_CUDA; }
  }
  dim3 grid(tiles_n, tiles_m, split);
  if (dtD == SVPC_F32) {
    return BN == 64 ? launch_major<64, float>(a_mn, b_mn, ta, tb, p, grid, st) : launch_major<128, float>(a_mn, b_mn, ta, tb, p, grid, st);
  } else if (dtD == SVPC_BF16) {
    return BN == 64 ? launch_major<64, bf16>(a_mn, b_mn, ta, tb, p, grid, st) : launch_major<128, bf16>(a_mn, b_mn, ta, tb, p, grid, st);
  }
  svpc_set_error("gemm_tc: bad output dtype %d", dtD);
  return SVPC_ERR_ARG;
}
