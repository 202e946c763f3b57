// Large-M path of the quantized Linear on the 5th-generation tensor cores (tcgen05 + TMEM + TMA):
//   C[M][N] (bf16) = X[M][K] (bf16) . W[N][K]^T (bf16, fp32 accumulate in TMEM) (+ bias)
// W is the bit-exact bf16 de-quantisation of the packed int4 layer (ll_get_weight: the reference's
// get_weight(dtype=bf16), quantization.py:390-409), so the result is the reference's fallback branch
// F.linear(inp, get_weight(inp.dtype)) (quantization.py:420-421) with fp32 accumulation - the branch the
// reference takes for every grouped checkpoint and the semantics of its Triton kernel for per-row ones.
// Used for prefill and batched decode (M >= 17); replaces Triton's mma.sync kernel (quantization.py:185-280).
//
// Structure (one 128 x BN output tile per CTA, 192 threads):
//   warp 0   : TMA producer - 2-D tiled loads (128B swizzle) of X[128 x 64] and W[BN x 64] into a ring
//   warp 1   : TMEM allocation; one elected thread issues tcgen05.mma (128 x BN x 16, kind::f16, bf16 inputs,
//              fp32 accumulator in TMEM); tcgen05.commit releases ring slots / signals the epilogue
//   warps 2-5: epilogue - tcgen05.ld (32 lanes x 32 columns per instruction), bias, bf16, global stores
#include <mutex>

#include "common.cuh"

namespace llb {
namespace {

constexpr int BM = 128;
constexpr int BK = 64;  // one 128-byte swizzle atom of bf16
constexpr int UMMA_K = 16;

typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = (EncodeTiledFn)p;
  });
  return fn;
}

// [rows][K] bf16 row-major -> 2-D map, box = [box_rows][64 elements], 128-byte swizzle, OOB = zero
int make_map(const void* base, int rows, int K, int box_rows, CUtensorMap* out) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return LL_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)K, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)K * 2};
  cuuint32_t box[2] = {(cuuint32_t)BK, (cuuint32_t)box_rows};
  cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void*)base, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) base=%p rows=%d K=%d", (int)r, base, rows, K);
    return LL_ERR_CUDA;
  }
  return LL_OK;
}

// ---- tcgen05 wrappers -------------------------------------------------------------------------------
__device__ __forceinline__ void tmem_alloc(uint32_t* smem_dst, uint32_t ncols) {
  asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(smem_dst)), "r"(ncols)
               : "memory");
  asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t ncols) {
  asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(ncols) : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }
// arrive on an mbarrier when every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t adesc, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}" ::"r"(tmem_d),
      "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
// shared-memory matrix descriptor: K-major operand tile [rows][64 bf16], 128-byte swizzle (as TMA wrote it):
// start address >> 4, LBO = 1 (unused for swizzled K-major), SBO = 1024 B between 8-row groups, version 1,
// layout type 2 (SWIZZLE_128B)   (bit layout: cute/arch/mma_sm100_desc.hpp SmemDescriptor)
__device__ __forceinline__ uint64_t smem_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)1 << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;
  d |= (uint64_t)2 << 61;
  return d;
}
// instruction descriptor, kind::f16: D = F32 (bits 4-5 = 1), A = B = BF16 (bits 7-9, 10-12 = 1), both K-major,
// N >> 3 at bits 17-22, M >> 4 at bits 24-28
__host__ __device__ constexpr uint32_t instr_desc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void tmem_ld_32x32(uint32_t taddr, uint32_t (&v)[32]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15]), "=r"(v[16]),
        "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]), "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]),
        "=r"(v[25]), "=r"(v[26]), "=r"(v[27]), "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void mbar_wait_parity(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n"
      "LLG_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra LLG_DONE;\n\t"
      "bra LLG_WAIT;\n"
      "LLG_DONE:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

template <int BN, int STAGES>
__global__ void __launch_bounds__(192, 1) gemm_bf16_tc_kernel(const __grid_constant__ CUtensorMap map_a,
                                                              const __grid_constant__ CUtensorMap map_b,
                                                              __nv_bfloat16* __restrict__ C,
                                                              const __nv_bfloat16* __restrict__ bias, int M, int N, int K) {
  constexpr int A_BYTES = BM * BK * 2, B_BYTES = BN * BK * 2;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~(uintptr_t)1023);
  uint8_t* sa = base;                      // [STAGES][128][64] bf16, swizzled
  uint8_t* sb = base + STAGES * A_BYTES;   // [STAGES][BN][64]
  uint64_t* full = reinterpret_cast<uint64_t*>(sb + STAGES * B_BYTES);
  uint64_t* empty = full + STAGES;
  uint64_t* tmem_full = empty + STAGES;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int n_blk = blockIdx.x, m_blk = blockIdx.y;
  const int nkb = (K + BK - 1) / BK;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_a);
    prefetch_tensormap(&map_b);
    for (int i = 0; i < STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, BN);  // BN fp32 accumulator columns (power of two >= 32)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_d = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer =====================
    if (lane == 0) {
      const uint64_t pol_keep = 0;  // default policy (both operands are re-read by other CTAs through L2)
      (void)pol_keep;
      int s = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait_parity(&empty[s], phase ^ 1);
        mbar_expect_tx(&full[s], A_BYTES + B_BYTES);
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                smem_u32(sa + s * A_BYTES)),
            "l"(&map_a), "r"(kb * BK), "r"(m_blk * BM), "r"(smem_u32(&full[s]))
            : "memory");
        asm volatile(
            "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                smem_u32(sb + s * B_BYTES)),
            "l"(&map_b), "r"(kb * BK), "r"(n_blk * BN), "r"(smem_u32(&full[s]))
            : "memory");
        if (++s == STAGES) {
          s = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = instr_desc_bf16(BM, BN);
      int s = 0;
      uint32_t phase = 0;
      for (int kb = 0; kb < nkb; ++kb) {
        mbar_wait_parity(&full[s], phase);
        tc_fence_after();
        const uint64_t adesc = smem_desc_sw128(smem_u32(sa + s * A_BYTES));
        const uint64_t bdesc = smem_desc_sw128(smem_u32(sb + s * B_BYTES));
#pragma unroll
        for (int k = 0; k < BK / UMMA_K; ++k) {
          // advance 16 elements = 32 bytes along K inside the swizzle atom: +2 in the (>>4) address field
          umma_bf16(tmem_d, adesc + (uint64_t)(2 * k), bdesc + (uint64_t)(2 * k), idesc, (kb | k) != 0 ? 1u : 0u);
        }
        umma_commit(&empty[s]);  // frees the ring slot when these MMAs have read it
        if (++s == STAGES) {
          s = 0;
          phase ^= 1u;
        }
      }
      umma_commit(tmem_full);  // accumulator complete
    }
  } else {
    // ===================== epilogue (warps 2..5) =====================
    const int quarter = warp & 3;  // a warp may only touch TMEM lanes [32 * (warpid % 4), +32)
    const int row = m_blk * BM + quarter * 32 + lane;
    mbar_wait_parity(tmem_full, 0);
    tc_fence_after();
#pragma unroll 1
    for (int c0 = 0; c0 < BN; c0 += 32) {
      uint32_t v[32];
      tmem_ld_32x32(tmem_d + ((uint32_t)(quarter * 32) << 16) + (uint32_t)c0, v);
      const int col = n_blk * BN + c0;
      if (row < M && col < N) {
        __nv_bfloat16* dst = C + (size_t)row * N + col;
        if (col + 32 <= N && (N & 7) == 0) {
#pragma unroll
          for (int j = 0; j < 4; ++j) {
            uint32_t pk[4];
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              float f0 = __uint_as_float(v[j * 8 + 2 * e]), f1 = __uint_as_float(v[j * 8 + 2 * e + 1]);
              if (bias != nullptr) {
                f0 += __bfloat162float(bias[col + j * 8 + 2 * e]);
                f1 += __bfloat162float(bias[col + j * 8 + 2 * e + 1]);
              }
              pk[e] = (uint32_t)f32_to_bf16_bits(f0) | ((uint32_t)f32_to_bf16_bits(f1) << 16);
            }
            *reinterpret_cast<uint4*>(dst + j * 8) = make_uint4(pk[0], pk[1], pk[2], pk[3]);
          }
        } else {
          for (int j = 0; j < 32 && col + j < N; ++j) {
            float f = __uint_as_float(v[j]);
            if (bias != nullptr) f += __bfloat162float(bias[col + j]);
            dst[j] = __float2bfloat16_rn(f);
          }
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_d, BN);
  }
}

template <int BN, int STAGES>
int launch_gemm(const CUtensorMap& ma, const CUtensorMap& mb, __nv_bfloat16* C, const __nv_bfloat16* bias, int M, int N, int K,
                cudaStream_t st) {
  constexpr size_t smem = (size_t)STAGES * (BM * BK * 2 + BN * BK * 2) + (2 * STAGES + 1) * 8 + 16 + 1024;
  static bool configured = false;
  if (!configured) {
    LL_CUDA(cudaFuncSetAttribute(gemm_bf16_tc_kernel<BN, STAGES>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  dim3 grid((N + BN - 1) / BN, (M + BM - 1) / BM);
  gemm_bf16_tc_kernel<BN, STAGES><<<grid, 192, smem, st>>>(ma, mb, C, bias, M, N, K);
  LL_LAUNCH_OK("gemm_bf16_tc_kernel");
  return LL_OK;
}

// ==================================================================================== fused int4 dequant-GEMM, M <= 128
// y[M][N] = x[M][K] . dequant(W)^T for 9 <= M <= 128 (batched decode, short prefill): the regime where the weights are
// the traffic (HBM-bound), so they are streamed ONCE in their packed form and de-quantised inside the main loop - what the
// reference's Triton kernel does (ref quantization.py:259-267: load nibble, (q - zero) * scale, tl.dot) - instead of
// writing a bf16 copy of W to HBM and reading it back.  (For M in the thousands the same idea is counter-productive on
// this machine: the tensor cores consume a 128-row tile's weights in K/16 * ~8 clk while de-quantising them costs
// ~0.06 clk per weight of ALU time per SM, so every 128-row block of M would pay the de-quantisation again; there the
// two-pass path above is within 3 % of the GEMM alone.  DESIGN.md section 3.)
//
//  * operand roles are swapped: the WEIGHTS are the A operand (128 output columns = the 128 TMEM lanes) and live in TENSOR
//    MEMORY: each of 128 de-quantising threads owns one output column, reads its nibbles from the packed stream (kind 1,
//    the batch-1 GEMV's stream: no second copy of the weights), forms bf16((q - zero) * scale) with the reference's two
//    bf16 roundings (HSUB2.BF16 exact for integer zeros, HMUL2.BF16 rounded) and writes 16 k per tcgen05.st;
//    the ACTIVATIONS are the B operand, TMA-loaded [M_pad][64 k] tiles (128-byte swizzle); D[128][M_pad] fp32 in TMEM.
//    Eight de-quantising warps: two per TMEM lane quarter, each takes half of a block's k (the arithmetic is ALU-bound).
//  * one thread issues tcgen05.mma (128 x M_pad x 16, A from TMEM); A is double-buffered (2 x 64 columns) so that
//    de-quantising k-block i + 1 overlaps the MMAs of block i; a 4-stage TMA ring feeds packed blocks and x tiles.
//  * work unit = (128-column tile, k-range): K is split so that the units fill the SMs; partial sums go to a scratch
//    buffer and a second small kernel adds them in split order (deterministic), adds the bias and rounds to bf16.
constexpr int W4_BN = 128;          // output columns per unit = TMEM lanes
constexpr int W4_KB = 128;          // k per block of the packed stream
constexpr int W4_BLK = W4_KB * 8 + 64;
__host__ __device__ constexpr int w4_stages(int) { return 4; }

__device__ __forceinline__ void umma_ts_bf16(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void tmem_st_x8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ uint32_t w4_prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
__device__ __forceinline__ uint32_t w4_sub2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t w4_mul2(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ void w4_bulk_1d(void* dst, const void* src, uint32_t bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(smem_u32(dst)),
               "l"(src), "r"(bytes), "r"(smem_u32(bar))
               : "memory");
}

struct W4Params {
  const uint8_t* w;       // stream kind 1: [N/16][K/128] blocks of 1088 B
  __nv_bfloat16* y;       // [M][N] (written directly when ksplit == 1)
  float* scratch;         // [tiles][ksplit][M_pad][128] partial sums (ksplit > 1)
  int32_t* counters;      // [tiles] arrivals of a tile's k-splits (left zero)
  const __nv_bfloat16* bias;
  int M, N, K, G, ksplit, mpad;
};

constexpr bool W4_REDUCE_IN_KERNEL = false;
constexpr int W4_DW = 8;  // de-quantising warps: two per TMEM lane quarter, each takes half of a block's k
template <int MPAD>
__global__ void __launch_bounds__(64 + 32 * W4_DW, 1) gemm_w4_tc_kernel(const __grid_constant__ CUtensorMap map_x, const W4Params p) {
  constexpr int W4_STAGES = w4_stages(MPAD);
  __shared__ int s_last;
  constexpr int W_BYTES = 8 * W4_BLK;        // eight 16-column blocks of the stream = 128 columns x 128 k
  constexpr int X_ATOM = MPAD * 128;         // [MPAD][64 k] bf16, swizzled
  constexpr int X_BYTES = 2 * X_ATOM;
  constexpr int STAGE = ((W_BYTES + 1023) / 1024) * 1024 + X_BYTES;
  constexpr int X_OFF = ((W_BYTES + 1023) / 1024) * 1024;
  extern __shared__ __align__(1024) uint8_t smem[];
  uint8_t* base = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem) + 1023) & ~(uintptr_t)1023);
  uint64_t* full = reinterpret_cast<uint64_t*>(base + W4_STAGES * STAGE);
  uint64_t* empty = full + W4_STAGES;
  uint64_t* a_ready = empty + W4_STAGES;   // [2] A buffer written (4 warps)
  uint64_t* a_free = a_ready + 2;          // [2] A buffer consumed (commit)
  uint64_t* tmem_full = a_free + 2;
  uint32_t* tmem_ptr = reinterpret_cast<uint32_t*>(tmem_full + 1);

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int tile = blockIdx.x / p.ksplit, ks = blockIdx.x - tile * p.ksplit;
  const int kb0 = (int)(((long long)p.G * ks) / p.ksplit), kb1 = (int)(((long long)p.G * (ks + 1)) / p.ksplit);
  const int nkb = kb1 - kb0;

  if (warp == 0 && lane == 0) {
    prefetch_tensormap(&map_x);
    for (int i = 0; i < W4_STAGES; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], W4_DW + 1);  // the de-quantising warps + the MMA commit
    }
    for (int i = 0; i < 2; ++i) {
      mbar_init(&a_ready[i], W4_DW);
      mbar_init(&a_free[i], 1);
    }
    mbar_init(tmem_full, 1);
    fence_barrier_init();
  }
  if (warp == 1) tmem_alloc(tmem_ptr, 256);  // D: columns [0, MPAD); A buffers: [128, 192) and [192, 256)
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_ptr;

  if (warp == 0) {
    // ===================== TMA producer: eight packed blocks + two x atoms per 128 k =====================
    if (lane == 0) {
      int s = 0;
      uint32_t phase = 0;
      for (int i = 0; i < nkb; ++i) {
        const int kb = kb0 + i;
        mbar_wait_parity(&empty[s], phase ^ 1);
        mbar_expect_tx(&full[s], W_BYTES + X_BYTES);
        uint8_t* st = base + s * STAGE;
#pragma unroll
        for (int t = 0; t < 8; ++t)
          w4_bulk_1d(st + t * W4_BLK, p.w + ((size_t)(tile * 8 + t) * p.G + kb) * W4_BLK, W4_BLK, &full[s]);
#pragma unroll
        for (int a = 0; a < 2; ++a)
          asm volatile(
              "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
                  smem_u32(st + X_OFF + a * X_ATOM)),
              "l"(&map_x), "r"(kb * W4_KB + a * 64), "r"(0), "r"(smem_u32(&full[s]))
              : "memory");
        if (++s == W4_STAGES) {
          s = 0;
          phase ^= 1u;
        }
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================
    if (lane == 0) {
      constexpr uint32_t idesc = instr_desc_bf16(128, MPAD);
      int s = 0;
      uint32_t phase = 0;
      for (int i = 0; i < nkb; ++i) {
        const int buf = i & 1;
        mbar_wait_parity(&full[s], phase);                   // the x atoms of this stage have landed
        mbar_wait_parity(&a_ready[buf], (uint32_t)(i >> 1) & 1u);  // the weights of this block are in tensor memory
        tc_fence_after();
        const uint32_t xs = smem_u32(base + s * STAGE + X_OFF);
        const uint32_t ta = tmem_base + 128u + (uint32_t)buf * 64u;
#pragma unroll
        for (int k = 0; k < 8; ++k) {
          const uint64_t bdesc = smem_desc_sw128(xs + (uint32_t)(k >> 2) * X_ATOM) + (uint64_t)(2 * (k & 3));
          umma_ts_bf16(tmem_base, ta + (uint32_t)(8 * k), bdesc, idesc, (i | k) != 0 ? 1u : 0u);
        }
        umma_commit(&a_free[buf]);
        umma_commit(&empty[s]);
        if (++s == W4_STAGES) {
          s = 0;
          phase ^= 1u;
        }
      }
      umma_commit(tmem_full);
    }
  } else {
    // ===================== de-quantising warps 2.., then the epilogue =====================
    const int quarter = warp & 3;            // a warp may only touch TMEM lanes [32 * (warpid % 4), +32)
    const int khalf = (warp - 2) >> 2;       // which half of a block's k this warp de-quantises (W4_DW = 8)
    const int t = quarter * 32 + lane;       // TMEM lane = output column n0 + t
    const int t16 = t >> 4, g = t & 7, hi = (t >> 3) & 1;
    const uint32_t lane_addr = tmem_base + ((uint32_t)(quarter * 32) << 16);
    const uint32_t sh = 4u * (uint32_t)hi;
    int s = 0;
    uint32_t phase = 0;
    for (int i = 0; i < nkb; ++i) {
      const int buf = i & 1;
      mbar_wait_parity(&full[s], phase);
      if (i >= 2) mbar_wait_parity(&a_free[buf], (uint32_t)((i >> 1) - 1) & 1u);
      tc_fence_after();
      const uint32_t blk = smem_u32(base + s * STAGE + t16 * W4_BLK);
      uint32_t sz;
      asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sz) : "r"(blk + 1024u + (uint32_t)(2 * g + hi) * 4u));
      const uint32_t s2 = (sz & 0xffffu) * 0x00010001u;  // bf16x2 (scale, scale)
      const uint32_t z2 = ((sz >> 16) ^ 0x8000u) * 0x00010001u;  // tail holds -zero: bf16x2 (zero, zero)
      // an integer zero (every RTN / GPTQ checkpoint): (128 + q) - (128 + zero) in ONE exact subtraction
      const float zf = __uint_as_float((sz & 0xffff0000u) ^ 0x80000000u);
      const bool zint = zf == floorf(zf) && zf >= 0.f && zf <= 64.f;
      const uint32_t zc2 = zint ? (uint32_t)(0x4300u + (uint32_t)zf) * 0x00010001u : 0x43004300u;
      const uint32_t ta = lane_addr + 128u + (uint32_t)buf * 64u;
      constexpr int MPW = 8 * 4 / W4_DW;  // 16-k slices per warp per block
#pragma unroll
      for (int mm = 0; mm < MPW; ++mm) {  // m = 2 tig + q: k = 32 tig + 16 q + 0..15 of the block -> TMEM columns 16 tig + 8 q + 0..7
        const int m = (W4_DW == 8 ? khalf * 4 : 0) + mm;
        const int tig = m >> 1, q = m & 1;
        uint4 wq;
        asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(wq.x), "=r"(wq.y), "=r"(wq.z), "=r"(wq.w)
                     : "r"(blk + (uint32_t)(q * 512 + (g * 4 + tig) * 16)));
        const uint32_t w[4] = {wq.x, wq.y, wq.z, wq.w};
        uint32_t r[8];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const uint32_t nib = (w[j] >> sh) & 0x0F0F0F0Fu;          // this column's four nibbles (k, k+1, k+2, k+3)
          // bf16 (128 + q) = 0x4300 | q; minus 128, minus zero (both exact for integer zeros), times scale (rounded):
          // the reference's get_weight in bf16 (quantization.py:399-408)
          uint32_t a = w4_sub2(w4_prmt(nib, 0x43434343u, 0x5140u), zc2);
          uint32_t b = w4_sub2(w4_prmt(nib, 0x43434343u, 0x5342u), zc2);
          if (!zint) {
            a = w4_sub2(a, z2);
            b = w4_sub2(b, z2);
          }
          r[2 * j] = w4_mul2(a, s2);
          r[2 * j + 1] = w4_mul2(b, s2);
        }
        tmem_st_x8(ta + (uint32_t)(16 * tig + 8 * q), r);
      }
      asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
      tc_fence_before();
      __syncwarp();
      if (lane == 0) {
        mbar_arrive(&a_ready[buf]);
        mbar_arrive(&empty[s]);
      }
      if (++s == W4_STAGES) {
        s = 0;
        phase ^= 1u;
      }
    }
    // ---- epilogue: D[t][m] for m < M: thread t holds output column n0 + t; the two warps of a lane quarter share the
    //      M_pad accumulator columns
    mbar_wait_parity(tmem_full, 0);
    tc_fence_after();
    const int n = tile * W4_BN + t;
    constexpr int EW = W4_DW / 4;  // warps per lane quarter
    const float bias = (p.bias != nullptr && p.ksplit == 1) ? __bfloat162float(p.bias[n]) : 0.f;
#pragma unroll 1
    for (int c0 = 16 * khalf; c0 < MPAD; c0 += 16 * EW) {
      uint32_t v[16];
      asm volatile(
          "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
          : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
            "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
          : "r"(lane_addr + (uint32_t)c0)
          : "memory");
      asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
#pragma unroll
      for (int j = 0; j < 16; ++j) {
        const int m = c0 + j;
        if (m < p.M) {
          if (p.ksplit == 1) p.y[(size_t)m * p.N + n] = __float2bfloat16_rn(__uint_as_float(v[j]) + bias);
          else p.scratch[(((size_t)tile * p.ksplit + ks) * MPAD + m) * W4_BN + t] = __uint_as_float(v[j]);
        }
      }
    }
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    tmem_dealloc(tmem_base, 256);
  }
  // (Measured, profiles/r02_fused_gemm.md: reducing inside the kernel - the last k-split to arrive sums the tile - and two
  // CTAs per SM were both SLOWER than a separate reduce kernel and one CTA per SM: 2.95 -> 3.76 ms per 161 Linears at
  // batch 32.  The code stays behind a constant for the record.)
  if (W4_REDUCE_IN_KERNEL && p.ksplit > 1) {
    // the k-split that arrives last adds the tile's partial sums in SPLIT order (the same bits whoever is last), adds the
    // bias, rounds and stores; the counter goes back to zero for the next launch
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) s_last = (atomicAdd(&p.counters[tile], 1) == p.ksplit - 1);
    __syncthreads();
    if (s_last) {
      __threadfence();
      const float* base_t = p.scratch + (size_t)tile * p.ksplit * MPAD * W4_BN;
      for (int e = threadIdx.x; e < p.M * W4_BN; e += blockDim.x) {
        const int m = e / W4_BN, tt = e - m * W4_BN;
        float acc = 0.f;
        for (int k2 = 0; k2 < p.ksplit; ++k2) acc += __ldcg(base_t + ((size_t)k2 * MPAD + m) * W4_BN + tt);
        const int n = tile * W4_BN + tt;
        if (p.bias != nullptr) acc += __bfloat162float(p.bias[n]);
        p.y[(size_t)m * p.N + n] = __float2bfloat16_rn(acc);
      }
      if (threadIdx.x == 0) p.counters[tile] = 0;
    }
  }
}

// y[m][n] = bf16(sum over the k-splits, in split order, + bias)   
__global__ void w4_reduce_kernel(const float* __restrict__ scratch, const __nv_bfloat16* __restrict__ bias,
                                 __nv_bfloat16* __restrict__ y, int M, int N, int ksplit, int mpad) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= (size_t)M * N) return;
  const int m = (int)(i / N), n = (int)(i - (size_t)m * N);
  const int tile = n / W4_BN, t = n - tile * W4_BN;
  float acc = 0.f;
  for (int ks = 0; ks < ksplit; ++ks) acc += scratch[(((size_t)tile * ksplit + ks) * mpad + m) * W4_BN + t];
  if (bias != nullptr) acc += __bfloat162float(bias[n]);
  y[i] = __float2bfloat16_rn(acc);
}

template <int MPAD>
int launch_w4(const CUtensorMap& mx, const W4Params& p, cudaStream_t st) {
  constexpr int W4_STAGES = w4_stages(MPAD);
  constexpr size_t stage = ((8 * W4_BLK + 1023) / 1024) * 1024 + 2 * MPAD * 128;
  constexpr size_t smem = W4_STAGES * stage + (2 * W4_STAGES + 5) * 8 + 16 + 1024;
  static bool configured = false;
  if (!configured) {
    LL_CUDA(cudaFuncSetAttribute(gemm_w4_tc_kernel<MPAD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = true;
  }
  gemm_w4_tc_kernel<MPAD><<<(p.N / W4_BN) * p.ksplit, 64 + 32 * W4_DW, smem, st>>>(mx, p);
  LL_LAUNCH_OK("gemm_w4_tc_kernel");
  return LL_OK;
}

// k-split that fills the SMs best (whole waves of one CTA per SM), each unit at least four k-blocks long
int w4_pick_ksplit(int tiles, int G, int sms) {
  int best = 1;
  double best_eff = 0.0;
  for (int ks = 1; ks <= 16 && G / ks >= 4; ++ks) {
    const int units = tiles * ks;
    const double eff = (double)units / ((double)((units + sms - 1) / sms) * sms);
    if (eff > best_eff + 0.03) {
      best_eff = eff;
      best = ks;
    }
  }
  return best;
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

int ll_gemm_bf16_tc(const void* x, const void* w, const void* bias, void* y, int M, int N, int K, ll_stream stream) {
  LL_CHECK_ARG(x && w && y, "null pointer");
  LL_CHECK_ARG(M > 0 && N > 0 && K > 0, "bad shape M=%d N=%d K=%d", M, N, K);
  LL_CHECK_ARG(K % 8 == 0, "K must be a multiple of 8 (16-byte rows for TMA), got %d", K);
  LL_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)w & 15) == 0 && ((uintptr_t)y & 15) == 0, "x/w/y must be 16-byte aligned");
  const int sms = sm_count();
  // wide tiles once they fill the machine; narrower ones for small M (prefill of a few hundred tokens)
  const long tiles256 = (long)((M + BM - 1) / BM) * ((N + 255) / 256);
  const bool wide = tiles256 >= sms;
  CUtensorMap ma, mb;
  if (int r = make_map(x, M, K, BM, &ma)) return r;
  if (int r = make_map(w, N, K, wide ? 256 : 128, &mb)) return r;
  cudaStream_t st = (cudaStream_t)stream;
  if (wide) return launch_gemm<256, 4>(ma, mb, (__nv_bfloat16*)y, (const __nv_bfloat16*)bias, M, N, K, st);
  return launch_gemm<128, 6>(ma, mb, (__nv_bfloat16*)y, (const __nv_bfloat16*)bias, M, N, K, st);
}

size_t ll_int4_gemm_tc_scratch_bytes(int M, int N, int K) {
  if (M <= 0 || N <= 0 || K <= 0 || N % W4_BN != 0 || K % W4_KB != 0) return 0;
  const int mpad = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128;
  const int ks = w4_pick_ksplit(N / W4_BN, K / W4_KB, sm_count() > 0 ? sm_count() : 148);
  return ks == 1 ? 0 : (size_t)(N / W4_BN) * ks * mpad * W4_BN * sizeof(float);
}

namespace {
// per-device arrival counters of the k-splits (one int per 128-column tile; the kernels leave them zero), allocated on
// first use outside stream capture
int32_t* w4_counters(int tiles, cudaStream_t st) {
  static int32_t* buf[16] = {};
  static int cap[16] = {};
  int dev = 0;
  if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= 16) return nullptr;
  if (cap[dev] < tiles) {
    cudaStreamCaptureStatus cs = cudaStreamCaptureStatusNone;
    if (cudaStreamIsCapturing(st, &cs) != cudaSuccess || cs != cudaStreamCaptureStatusNone) return nullptr;
    const int n = tiles < 4096 ? 4096 : tiles;
    int32_t* nb = nullptr;
    if (cudaMalloc(&nb, (size_t)n * sizeof(int32_t)) != cudaSuccess) return nullptr;
    cudaMemset(nb, 0, (size_t)n * sizeof(int32_t));
    buf[dev] = nb;   // (an older, smaller buffer stays alive: captured graphs may hold it)
    cap[dev] = n;
  }
  return buf[dev];
}
}  // namespace

int ll_int4_gemm_tc(const uint8_t* w_stream, const void* x, const void* bias, void* y, float* scratch, int M, int N, int K,
                    int group, ll_stream stream) {
  LL_CHECK_ARG(w_stream && x && y, "null pointer");
  LL_CHECK_ARG(M >= 1 && M <= 128, "ll_int4_gemm_tc serves 1 <= M <= 128 rows (got %d); larger M: ll_get_weight + ll_gemm_bf16_tc", M);
  if (N <= 0 || K <= 0 || group <= 0 || N % W4_BN != 0 || group % W4_KB != 0 || K % group != 0) {
    set_error("ll_int4_gemm_tc needs N %% 128 == 0 and group %% 128 == 0 (N=%d K=%d group=%d)", N, K, group);
    return LL_ERR_UNSUPPORTED;
  }
  LL_CHECK_ARG(((uintptr_t)x & 15) == 0 && ((uintptr_t)w_stream & 15) == 0, "x / stream must be 16-byte aligned");
  const int sms = sm_count();
  W4Params p;
  p.w = w_stream; p.y = (__nv_bfloat16*)y; p.scratch = scratch; p.bias = (const __nv_bfloat16*)bias;
  p.M = M; p.N = N; p.K = K; p.G = K / W4_KB;
  p.mpad = M <= 16 ? 16 : M <= 32 ? 32 : M <= 64 ? 64 : 128;
  p.ksplit = w4_pick_ksplit(N / W4_BN, p.G, sms > 0 ? sms : 148);
  LL_CHECK_ARG(p.ksplit == 1 || scratch != nullptr, "this shape splits K %d ways: pass ll_int4_gemm_tc_scratch_bytes() of scratch", p.ksplit);
  CUtensorMap mx;
  if (int r = make_map(x, M, K, p.mpad, &mx)) return r;
  cudaStream_t st = (cudaStream_t)stream;
  p.counters = nullptr;
  if (W4_REDUCE_IN_KERNEL && p.ksplit > 1) {
    p.counters = w4_counters(N / W4_BN, st);
    if (p.counters == nullptr) {
      set_error("ll_int4_gemm_tc: could not allocate the split counters (first call under stream capture?)");
      return LL_ERR_CUDA;
    }
  }
  int r;
  if (p.mpad == 16) r = launch_w4<16>(mx, p, st);
  else if (p.mpad == 32) r = launch_w4<32>(mx, p, st);
  else if (p.mpad == 64) r = launch_w4<64>(mx, p, st);
  else r = launch_w4<128>(mx, p, st);
  if (r) return r;
  if (!W4_REDUCE_IN_KERNEL && p.ksplit > 1) {
    const size_t n = (size_t)M * N;
    w4_reduce_kernel<<<(unsigned)((n + 255) / 256), 256, 0, st>>>(scratch, p.bias, p.y, M, N, p.ksplit, p.mpad);
    LL_LAUNCH_OK("w4_reduce_kernel");
  }
  return LL_OK;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// No-allocation variant of ll_int4_gemm_tc (no counterpart in the reference; added
// for callers that cannot call cudaMalloc: code under CUDA-graph capture, tight serving loops, harnesses that forbid it).
// The split-K partial sums live in a file-scope device array instead of a caller-provided buffer.  Every GPU that loads
// this library holds LL_INT4_GEMM_TC_STATIC_SCRATCH_BYTES until exit.  All calls on a device share the array: two calls
// that overlap on the GPU (different streams, or graphs replayed concurrently) race on it.  Nothing in it has to start at
// zero: every (tile, k-split) unit writes rows m < M of its block of partial sums, and w4_reduce_kernel (after it on
// the stream) reads exactly those rows, so the previous call's contents are never observed.
// ---------------------------------------------------------------------------------------------------------------------
static_assert(!llb::W4_REDUCE_IN_KERNEL, "the static-scratch entry point has no arrival counters: it needs the two-kernel reduce");
static __device__ __align__(256) float g_w4_static_scratch[LL_INT4_GEMM_TC_STATIC_SCRATCH_BYTES / sizeof(float)];

namespace {
// In host code `g_w4_static_scratch` is only a stand-in for the symbol: its device address comes from
// cudaGetSymbolAddress, asked once per device with that device current.
float* w4_static_scratch(void) {
  constexpr int MAX_DEV = 64;
  static float* ptr[MAX_DEV] = {};
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess || dev < 0 || dev >= MAX_DEV) {
    set_error("ll_int4_gemm_tc_static: cudaGetDevice: %s (device %d)", cudaGetErrorString(e), dev);
    return nullptr;
  }
  if (ptr[dev] == nullptr) {
    void* d = nullptr;
    e = cudaGetSymbolAddress(&d, g_w4_static_scratch);
    if (e != cudaSuccess || d == nullptr) {
      set_error("ll_int4_gemm_tc_static: cudaGetSymbolAddress: %s", cudaGetErrorString(e));
      return nullptr;
    }
    ptr[dev] = static_cast<float*>(d);
  }
  return ptr[dev];
}
}  // namespace

extern "C" int ll_int4_gemm_tc_static(const uint8_t* w_stream, const void* x, const void* bias, void* y, int M, int N, int K,
                                      int group, ll_stream stream) {
  LL_CHECK_ARG(w_stream && x && y, "null pointer");
  LL_CHECK_ARG(M >= 1 && M <= 128, "ll_int4_gemm_tc_static serves 1 <= M <= 128 rows (got %d)", M);
  if (N <= 0 || K <= 0 || group <= 0 || N % llb::W4_BN != 0 || group % llb::W4_KB != 0 || K % group != 0) {
    set_error("ll_int4_gemm_tc_static needs N %% 128 == 0 and group %% 128 == 0 (N=%d K=%d group=%d)", N, K, group);
    return LL_ERR_UNSUPPORTED;
  }
  const size_t need = ll_int4_gemm_tc_scratch_bytes(M, N, K);
  if (need > (size_t)LL_INT4_GEMM_TC_STATIC_SCRATCH_BYTES) {
    set_error("ll_int4_gemm_tc_static: M=%d N=%d K=%d needs %zu bytes of split-K scratch, the built-in array holds %zu: "
              "use ll_int4_gemm_tc with a caller-provided buffer", M, N, K, need, (size_t)LL_INT4_GEMM_TC_STATIC_SCRATCH_BYTES);
    return LL_ERR_UNSUPPORTED;
  }
  float* scratch = nullptr;
  if (need != 0) {
    scratch = w4_static_scratch();
    if (scratch == nullptr) return LL_ERR_CUDA;
  }
  return ll_int4_gemm_tc(w_stream, x, bias, y, scratch, M, N, K, group, stream);
}
