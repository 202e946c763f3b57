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

}  // extern "C"
