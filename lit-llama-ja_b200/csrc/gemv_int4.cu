// int4 (GPTQ column-block) decode GEMV for sm_100a: y[M<=8][N] = x[M][K] . dequant(W)^T
//
// Replaces the reference's Triton launcher (ref: lit_llama/quantization.py:282-331, which pads
// M=1 to 256 rows and reads every packed byte twice) and its dequantise-then-F.linear fallback
// (ref: quantization.py:390-421) for batch-1..8 decode.
//
// Design (B200-first; HBM-bound byte work, no reshaping into a big GEMM):
//  * One-time repack (ll_int4_prepare) of the checkpoint bytes into a fragment-ordered STREAM of
//    blocks: 16 output columns x KB input features (KB = 128/64/32), k-blocks fastest.  A block is
//    the 32 lanes' mma.m16n8k16 A-fragment words in issue order (one 32-bit word = one MMA's
//    A fragment = 8 nibbles placed at bit offsets 3+4i so that `(w >> 4i) & 0x00780078 | 0x41804180`
//    IS the bf16x2 register holding (16+q, 16+q')), followed by the block's 16 (scale, -(16+zero))
//    bf16 pairs.  7 ALU instructions per 8 weights, no per-weight scale/zero arithmetic.
//  * Work item = one block.  Whole 16-column tiles are split contiguously over one CTA per SM, so a
//    CTA's share is one contiguous byte range per matrix and no tile is shared between CTAs: one
//    producer thread streams the range with 1-D bulk TMA copies (L2 evict-first) into a
//    shared-memory ring of chunks gated by mbarriers.  Weights do not depend on the previous kernel:
//    under programmatic dependent launch the next GEMV's CTA co-resides (<= 110 KB smem each) and
//    fills its ring while this one computes.
//  * Consumer warps take BPW consecutive blocks per chunk each: LDS.128 of fragment words, dequant to
//    bf16 by LOP3/SHF, mma.sync m16n8k16 (bf16 x bf16 -> fp32) with the weights as the A operand and
//    up to 8 batch rows of x as the B columns.  Per block: sum (16+q) x, then one FFMA pair folds
//    -(16+zero) * sum(x_block) and the scale into the running column sums.  The blocks a warp takes
//    from one chunk share their tile, so the chunk body is straight-line code behind ONE tile test.
//  * x is staged once per CTA in shared memory (batched ld.global.cg, one L2 round trip); the optional
//    fused RMSNorm (ref model.py:276-283 rounding steps) is computed from the loaded row (its sum of
//    squares included: every CTA holds the whole row), together with the per-block sums of x.
//  * Fixed-order reduction across warps in shared memory, then the fused epilogue, one warp per tile:
//    plain store, residual add, RoPE + KV-cache write (c_attn), SwiGLU (c_fc1/c_fc2 as two matrices
//    of one launch), or peer stores for the tensor-parallel push.  The decode chain's configuration
//    gets one instantiation per (epilogue, norm, batch 1): less cold code per launch.
//  * Measurements, ablations and rejected variants: profiles/r01_gemv_experiments.md, DESIGN.md section 3.
//    Tuning switches (-D): LLB_XPRED, LLB_SHIFT_FMA, LLB_XSTAGE_TMA, LLB_TRACE, LLB_ABL_* (timing
//    ablations, wrong sums) - build with `build.py --variant`, select with LLB200_LIB.
#include <stdlib.h>

#include "common.cuh"

namespace llb {
namespace {

constexpr int MAXM = 8;
constexpr int TILE_N = 16;
constexpr int TAIL_BYTES = TILE_N * 4;  // 16 x (bf16 scale, bf16 -(16+zero))

__host__ __device__ inline int pick_kb(int K, int group) {
  if (group <= 0 || K <= 0 || K % group != 0) return 0;
  if (group % 128 == 0) return 128;
  if (group % 64 == 0) return 64;
  if (group % 32 == 0) return 32;
  return 0;
}
__host__ __device__ inline int block_bytes(int kb) { return kb * 8 + TAIL_BYTES; }

struct Params {
  const uint8_t* w[2];
  int nmat, N, K, group, M;
  int G;       // k-blocks per (tile, matrix) = K / KB
  int GT;      // items per tile = nmat * G
  int ntiles;  // N / 16
  int I;       // total items
  const __nv_bfloat16* x;
  const __nv_bfloat16* norm_w;
  float eps;
  int epilogue;
  __nv_bfloat16* y;
  const __nv_bfloat16* bias;
  const float* rope;
  const int32_t* pos;
  __nv_bfloat16* k_cache;
  __nv_bfloat16* v_cache;
  int n_head, head_size, cache_len;
  __nv_bfloat16* push_dst[8];  // LL_EPI_PUSH: every rank's receive buffer for this rank's partial sums
  uint32_t* push_cnt[8];       //              and its arrival counter
  int push_world;
  // shared-memory plan
  int S;          // ring slots
  int xs_stride;  // bytes per staged x row
  int ts_max;     // (tile, matrix) slots of the cross-warp reduction
  int off_xs, off_sx, off_red, off_wv, off_bar, off_nw;
  int prefetch_delay_ns;  // producer start delay under PDL (lets the running kernel's x staging go first)
  int helpers;            // batch-1 kernel, ODD footprint, no fused norm: the parked warps convert x too before they park
  const uint8_t* xdg_planes;  // batch-1 kernel, no fused norm: x already as four digit planes [4][K + 16] (ll_decode_attn_digits)
  const float* xdg_rec;       //   and per-block records [K / 128][4]: the staging is one bulk copy
  unsigned long long* trace;  // debug: [P][8] globaltimer stamps of this launch, or nullptr
};

__device__ __forceinline__ unsigned long long gtimer() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
// stage stamps for profiles/gemv_trace.py: only in the tuning build (build.py --variant trace LLB_TRACE=1) - the
// stamps are code every launch would have to fetch
#ifdef LLB_TRACE
#define TRACE(i)                                                                   \
  do {                                                                             \
    if (p.trace != nullptr && threadIdx.x == 0) {                                  \
      p.trace[blockIdx.x * 16 + (i)] = gtimer();                                   \
      if ((i) == 0) {                                                              \
        unsigned sm_;                                                              \
        asm volatile("mov.u32 %0, %smid;" : "=r"(sm_));                            \
        p.trace[blockIdx.x * 16 + 14] = sm_ + 1;                                   \
      }                                                                            \
    }                                                                              \
  } while (0)
#else
#define TRACE(i) do { } while (0)
#endif

__device__ __forceinline__ void consumer_sync(int n) { asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory"); }

__device__ __forceinline__ uint32_t lop_and_or(uint32_t a, uint32_t mask, uint32_t orv) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(r) : "r"(a), "r"(mask), "r"(orv));  // (a & mask) | orv
  return r;
}
__device__ __forceinline__ uint32_t rotr(uint32_t a, uint32_t n) {
  uint32_t r;
  asm("shf.r.wrap.b32 %0, %1, %1, %2;" : "=r"(r) : "r"(a), "r"(n));
  return r;
}
template <int S>
__device__ __forceinline__ uint32_t shr_fma(uint32_t a) {  // a >> S, issued to the FMA pipe
  uint32_t r;
  asm("mul.hi.u32 %0, %1, %2;" : "=r"(r) : "r"(a), "n"(1u << (32 - S)));
  return r;
}
// 1-D bulk copy global -> shared with an L2 cache policy, completion on an mbarrier.
__device__ __forceinline__ void bulk_load_1d_hint(void* dst, const void* src, uint32_t bytes, uint64_t* bar,
                                                  uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(dst)),
      "l"(src), "r"(bytes), "r"(smem_u32(bar)), "l"(policy)
      : "memory");
}

// first tile of CTA c when I tiles are split over P CTAs
// (c * I fits 32 bits: c <= SM count, I = tiles of one layer; the host checks it)
__device__ __forceinline__ int range_start(int c, int I, int P) { return (int)(((unsigned)c * (unsigned)I) / (unsigned)P); }
constexpr uint32_t NIB_MASK = 0x00780078u;   // nibble at bits 3..6 of each bf16 half
constexpr uint32_t NIB_MAGIC = 0x41804180u;  // bf16 16.0: 16 + q exact for q in 0..15

// spin on an mbarrier phase (try_wait suspends in hardware up to a time limit, so this is not a hot loop)
__device__ __forceinline__ void mbar_wait_u32(uint32_t bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n"
      "LLB_WAIT:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra LLB_DONE;\n\t"
      "bra LLB_WAIT;\n"
      "LLB_DONE:\n\t}" ::"r"(bar),
      "r"(parity)
      : "memory");
}
// W <= 8: two CTAs per SM (this GEMV and the next one co-reside under programmatic dependent launch, <= 112
// registers).  W >= 12 (tuning configurations): one CTA per SM with the whole shared memory as the ring.
// EPI / NORM >= 0: the epilogue / the fused RMSNorm are compile-time constants (the decode chain's hot
// configuration gets one specialised kernel per layer kind: less code to fetch cold at every launch - the kernel
// runs its prologue and epilogue once - and no epilogue dispatch); -1: taken from the arguments.
// MM > 0: the batch (rows of x) is the compile-time constant MM (batch-1 decode: every per-row loop folds away).
// (A neighbour stream-K split of the boundary tiles was built and measured in round 2: q 311 -> 361, p 131 -> 199, f 423 -> 519,
// d 277 -> 325 us per 32 layers - slower than the 7-14 % tile imbalance it removes; taken out, profiles/r02_gemv_experiments.md.)
template <int W, int KSP, int BPW, int EPI = -1, int NORM = -1, int MM = 0>
__global__ void __launch_bounds__((W + 1) * 32, (W >= 12 ? 1 : 2)) gemv_int4_kernel(const __grid_constant__ Params p) {
  const int epilogue = EPI >= 0 ? EPI : p.epilogue;
  constexpr int KB = 32 * KSP;
  constexpr int BLK = KB * 8 + TAIL_BYTES;
  constexpr int CH = W * BPW;  // items per chunk: BPW consecutive blocks per consumer warp
  constexpr int CHB = CH * BLK;
  constexpr int NCONS = W * 32;
  constexpr int LPB = KB / 8;  // lanes staging one k-block of x (8 elements each)
  constexpr int MAX_SLOTS_C = 12;

  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* ring = smem;
  uint8_t* xs = smem + p.off_xs;                            // [M+1][xs_stride] bf16, last row zero
  float* sx = reinterpret_cast<float*>(smem + p.off_sx);    // [G][8]
  float* red = reinterpret_cast<float*>(smem + p.off_red);  // [ts_max][W][M][16]
  float* wv = reinterpret_cast<float*>(smem + p.off_wv);    // [W][nmat][M][16]
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  uint64_t* empty = full + p.S;
  __shared__ float s_rnorm[W][MAXM];
  __shared__ float s_part[W][MAXM];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int P = gridDim.x, c = blockIdx.x;
  // whole tiles per CTA: no tile is shared between CTAs, so no cross-CTA exchange on the critical path
  const int i0 = range_start(c, p.ntiles, P) * p.GT;
  const int i1 = range_start(c + 1, p.ntiles, P) * p.GT;
  const int nch = (i1 - i0 + CH - 1) / CH;
  const int G = p.G, S = p.S, M = MM > 0 ? MM : p.M;

  TRACE(0);
  if (threadIdx.x == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], W);
    }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();

  if (warp == W) {
    // =============================== producer warp ===============================
    if (lane == 0) {
      const uint64_t pol = policy_evict_first();
      if (p.prefetch_delay_ns > 0) __nanosleep(p.prefetch_delay_ns);
      int slot = 0;
      uint32_t phase = 1;  // parity of the PREVIOUS use of the slot once the ring wraps
      int it = i0;
      int vt = it / G, g = it - vt * G;
      for (int ch = 0; ch < nch; ++ch, ++slot) {
        if (slot == S) {
          slot = 0;
          phase ^= 1u;
        }
        if (ch >= S) mbar_wait_u32(smem_u32(&empty[slot]), phase);
        const int b = min(it + CH, i1);
        mbar_expect_tx(&full[slot], (uint32_t)(b - it) * BLK);
        uint8_t* dst = ring + (size_t)slot * CHB;
        while (it < b) {  // contiguous runs: a run ends where the (tile, matrix) changes
          const int run = min(G - g, b - it);
          const int mat = p.nmat == 2 ? (vt & 1) : 0, tile = p.nmat == 2 ? (vt >> 1) : vt;
          const uint8_t* src = p.w[mat] + ((size_t)tile * G + g) * BLK;
          bulk_load_1d_hint(dst, src, (uint32_t)run * BLK, &full[slot], pol);
          dst += (size_t)run * BLK;
          it += run;
          g += run;
          if (g == G) {
            g = 0;
            ++vt;
          }
        }
      }
    }
    return;
  }

  // ================================= consumer warps =================================
  const int g = lane >> 2, tig = lane & 3;
  const int ctid = threadIdx.x;  // 0..NCONS-1
  const int vt_first = i0 / G, vt_last = (i1 - 1) / G;
  const int nvt = vt_last - vt_first + 1;
  const int MT = M * TILE_N;  // floats per (tile, matrix) result
  const bool has_norm = NORM >= 0 ? (NORM != 0) : (p.norm_w != nullptr);
  const int vpr = p.K >> 3;

  // norm weights are constants: their bulk copy into shared memory starts before the previous kernel ends
  uint8_t* nws = smem + p.off_nw;  // [K] bf16 (only with the fused RMSNorm)
  uint64_t* xbar = full + 2 * MAX_SLOTS_C;  // [0]: x rows landed, [1]: norm weights landed
  if (ctid == 0) {
    mbar_init(&xbar[0], 1);
    mbar_init(&xbar[1], 1);
    fence_barrier_init();
    if (has_norm) {
      mbar_expect_tx(&xbar[1], (uint32_t)p.K * 2);
      bulk_load_1d(nws, p.norm_w, (uint32_t)p.K * 2, &xbar[1]);
    }
  }
  for (int i = ctid; i < nvt * W * MT; i += NCONS) red[i] = 0.f;
  for (int i = ctid; i < G * 8; i += NCONS) sx[i] = 0.f;
  for (int i = ctid; i < (p.K >> 2); i += NCONS) *reinterpret_cast<uint2*>(xs + (size_t)M * p.xs_stride + i * 8) = make_uint2(0u, 0u);
  consumer_sync(NCONS);

  TRACE(1);
  pdl_wait();  // activations (x, ssq, residual, caches, scratch) come from earlier kernels
  TRACE(2);
  int rope_pos = 0, kv_slot = 0;
  if (epilogue == LL_EPI_ROPE_KV) {
    rope_pos = p.pos[0];
    kv_slot = p.pos[1];
  }

#ifndef LLB_XSTAGE_TMA
  // ---- stage x: every thread fetches its 16-byte pieces of the row with up to XB loads in flight (one L2 round
  //      trip for K <= 8192 with 256 consumer threads; eight in flight measured slower; 0.1-0.3 us per kernel faster than one bulk-TMA copy
  //      of the row into shared memory, -DLLB_XSTAGE_TMA), accumulates the sum of squares from the registers (fused RMSNorm; every CTA holds the
  //      whole row, fixed summation order) and parks the raw row in shared memory for the second pass
  for (int m = 0; m < M; ++m) {
    const uint4* src = reinterpret_cast<const uint4*>(p.x + (size_t)m * p.K);
    uint8_t* row = xs + (size_t)m * p.xs_stride;
    float ss = 0.f;
    constexpr int XB = 4;
    for (int v0 = ctid; v0 < vpr; v0 += XB * NCONS) {
      uint4 raw[XB];
#pragma unroll
      for (int j = 0; j < XB; ++j) {
        const int v = v0 + j * NCONS;
        raw[j] = v < vpr ? __ldcg(src + v) : make_uint4(0u, 0u, 0u, 0u);
      }
#pragma unroll
      for (int j = 0; j < XB; ++j) {
        const int v = v0 + j * NCONS;
        if (v < vpr) {
          if (has_norm) {
            const uint32_t wd[4] = {raw[j].x, raw[j].y, raw[j].z, raw[j].w};
#pragma unroll
            for (int e = 0; e < 4; ++e) {
              const float x0 = bf16_bits_to_f32(wd[e] & 0xffffu), x1 = bf16_bits_to_f32(wd[e] >> 16);
              ss += bf16r(__fmul_rn(x0, x0)) + bf16r(__fmul_rn(x1, x1));
            }
          }
          *reinterpret_cast<uint4*>(row + v * 16) = raw[j];
        }
      }
    }
    if (has_norm) {
      ss = warp_sum(ss);
      if (lane == 0) s_part[warp][m] = ss;
    }
  }
  TRACE(9);
  if (has_norm) {
    consumer_sync(NCONS);
    mbar_wait_u32(smem_u32(&xbar[1]), 0);
    TRACE(7);
  }
#else
  // ---- stage x: one bulk copy per row straight into shared memory (a single memory round trip whatever
  //      K is), then normalise in place (optional fused RMSNorm) and form the per-k-block sums
  if (ctid == 0) {
    mbar_expect_tx(&xbar[0], (uint32_t)M * p.K * 2);
    for (int m = 0; m < M; ++m) bulk_load_1d(xs + (size_t)m * p.xs_stride, p.x + (size_t)m * p.K, (uint32_t)p.K * 2, &xbar[0]);
  }
  if (has_norm) mbar_wait_u32(smem_u32(&xbar[1]), 0);
  TRACE(8);
  mbar_wait_u32(smem_u32(&xbar[0]), 0);
  TRACE(9);
  if (has_norm) {
    // sum of squares of each row straight from the staged copy (every CTA holds the whole row): no
    // partial sums handed over from the previous kernel, fixed summation order
    for (int m = 0; m < M; ++m) {
      float ss = 0.f;
      for (int v = ctid; v < vpr; v += NCONS) {
        const uint4 raw = *reinterpret_cast<const uint4*>(xs + (size_t)m * p.xs_stride + v * 16);
        const uint32_t wd[4] = {raw.x, raw.y, raw.z, raw.w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          const float x0 = bf16_bits_to_f32(wd[e] & 0xffffu), x1 = bf16_bits_to_f32(wd[e] >> 16);
          ss += bf16r(__fmul_rn(x0, x0)) + bf16r(__fmul_rn(x1, x1));
        }
      }
      ss = warp_sum(ss);
      if (lane == 0) s_part[warp][m] = ss;
    }
    consumer_sync(NCONS);
    if (lane < M) {
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < W; ++w) tot += s_part[w][lane];
      const float mean = bf16r(tot * (1.0f / (float)p.K));
      s_rnorm[warp][lane] = bf16r(rsqrtf(bf16r(mean + p.eps)));
    }
    __syncwarp();
    TRACE(7);
  }
#endif
  for (int m = 0; m < M; ++m) {
#ifndef LLB_XSTAGE_TMA
    float rn = 0.f;
    if (has_norm) {  // every thread forms the row's scale itself (same order of the warps' partial sums)
      float tot = 0.f;
#pragma unroll
      for (int w = 0; w < W; ++w) tot += s_part[w][m];
      const float mean = bf16r(tot * (1.0f / (float)p.K));
      rn = bf16r(rsqrtf(bf16r(mean + p.eps)));
    }
#else
    const float rn = has_norm ? s_rnorm[warp][m] : 0.f;
#endif
    uint8_t* row = xs + (size_t)m * p.xs_stride;
#pragma unroll 2
    for (int v0 = 0; v0 < vpr; v0 += NCONS) {
      const int v = v0 + ctid;
      const bool ok = v < vpr;
      float part = 0.f;
      if (ok) {
        const uint4 raw = *reinterpret_cast<const uint4*>(row + v * 16);
        uint32_t wd[4] = {raw.x, raw.y, raw.z, raw.w};
        if (has_norm) {
          const uint4 nraw = *reinterpret_cast<const uint4*>(nws + v * 16);
          const uint32_t nw[4] = {nraw.x, nraw.y, nraw.z, nraw.w};
#pragma unroll
          for (int e = 0; e < 4; ++e) {
            const float x0 = bf16r(__fmul_rn(bf16_bits_to_f32(wd[e] & 0xffffu), rn));
            const float x1 = bf16r(__fmul_rn(bf16_bits_to_f32(wd[e] >> 16), rn));
            const uint32_t o0 = f32_to_bf16_bits(__fmul_rn(bf16_bits_to_f32(nw[e] & 0xffffu), x0));
            const uint32_t o1 = f32_to_bf16_bits(__fmul_rn(bf16_bits_to_f32(nw[e] >> 16), x1));
            wd[e] = o0 | (o1 << 16);
          }
          *reinterpret_cast<uint4*>(row + v * 16) = make_uint4(wd[0], wd[1], wd[2], wd[3]);
        }
#pragma unroll
        for (int e = 0; e < 4; ++e) part += bf16_bits_to_f32(wd[e] & 0xffffu) + bf16_bits_to_f32(wd[e] >> 16);
      }
#pragma unroll
      for (int o = LPB / 2; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
      if (ok && (lane % LPB) == 0) sx[(v / LPB) * 8 + m] = part;
    }
  }
  TRACE(10);
  consumer_sync(NCONS);
  TRACE(3);

  // epilogue operands that do not depend on this kernel's sums (residual stream, bias, RoPE row) are
  // fetched now for this warp's first tile, behind the main loop
  const int nmat = p.nmat;
  const int T_first = vt_first / nmat, T_last = vt_last / nmat;
  auto load_pre = [&](int T, float (&pre)[4], float2 (&cs)[2]) {
    const int n0 = T * TILE_N;
    if (epilogue == LL_EPI_RESIDUAL) {
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e = lane + 32 * j;
        if (e < MT) pre[j] = __bfloat162float(p.y[(size_t)(e >> 4) * p.N + n0 + (e & 15)]);
      }
    } else if (epilogue == LL_EPI_STORE && p.bias != nullptr) {
      pre[0] = __bfloat162float(p.bias[n0 + (lane & 15)]);
    } else if (epilogue == LL_EPI_ROPE_KV) {
      const int E = p.N / 3, hs = p.head_size;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int e = lane + 32 * j;
        const int n = n0 + 2 * (e & 7);
        const int part = n / E, d = (n - part * E) % hs;
        if (e < M * (TILE_N / 2) && part < 2)
          cs[j] = *reinterpret_cast<const float2*>(p.rope + ((size_t)rope_pos * (hs / 2) + d / 2) * 2);
      }
    }
  };
  float pre0[4] = {0.f, 0.f, 0.f, 0.f};
  float2 cs0[2] = {make_float2(1.f, 0.f), make_float2(1.f, 0.f)};
  if (T_first + warp <= T_last) load_pre(T_first + warp, pre0, cs0);

  // ---- stream this CTA's items: BPW consecutive blocks per warp per chunk, computed as independent
  //      MMA chains (ILP), then folded into the running column sums in item order
  float out[4] = {0.f, 0.f, 0.f, 0.f};
  int cur_vt = -1;
  const uint32_t xs_lane = smem_u32(xs) + (uint32_t)(min(g, M) * p.xs_stride + tig * 16);  // rows >= M read zeros
#if defined(LLB_ABL_NO_XLD)  // timing ablation (wrong sums): no B-fragment loads
#define LLB_XLD(addr) make_uint4(0x3f803f80u, 0x3f803f80u, 0x3f803f80u, 0x3f803f80u)
#elif defined(LLB_XPRED)
  // only the lanes holding a real batch row fetch their B fragment (for M = 1 that is one quarter-warp: one
  // shared-memory wavefront per LDS.128 instead of four); the others keep zeros in registers
  const bool xl = g < M;
#define LLB_XLD(addr) (xl ? lds128(addr) : make_uint4(0u, 0u, 0u, 0u))
#else
#define LLB_XLD(addr) lds128(addr)
#endif
  const uint32_t sx_lane = smem_u32(sx) + (uint32_t)(tig * 8);
  const uint32_t ring_u = smem_u32(ring) + (uint32_t)(warp * BPW * BLK);
  const uint32_t full_u = smem_u32(full), empty_u = smem_u32(empty);

  auto flush = [&](int vt) {
    float* r = red + ((size_t)(vt - vt_first) * W + warp) * MT;
    const int m0 = 2 * tig, m1 = m0 + 1;
    if (m0 < M) {
      r[m0 * TILE_N + g] = out[0];
      r[m0 * TILE_N + g + 8] = out[2];
    }
    if (m1 < M) {
      r[m1 * TILE_N + g] = out[1];
      r[m1 * TILE_N + g + 8] = out[3];
    }
  };
  auto mma_word = [&](float (&a)[4], uint32_t w, uint32_t b0, uint32_t b1) {
#if defined(LLB_ABL_NO_DEQ)  // timing ablation (wrong sums): no de-quantisation arithmetic
    mma_bf16_16816(a, w, w, w, w, b0, b1);
#else
    const uint32_t a0 = lop_and_or(w, NIB_MASK, NIB_MAGIC);
#ifdef LLB_SHIFT_FMA
    // the two plain right shifts as IMAD.HI (w * 2^(32-s) >> 32) on the FMA pipe.  Measured SLOWER on B200
    // (qpfd 1426 -> 1596 us per 32 layers: IMAD.HI is not a half-rate instruction); kept as a tuning switch.
    const uint32_t a1 = lop_and_or(shr_fma<4>(w), NIB_MASK, NIB_MAGIC);
    const uint32_t a2 = lop_and_or(shr_fma<8>(w), NIB_MASK, NIB_MAGIC);
#else
    const uint32_t a1 = lop_and_or(w >> 4, NIB_MASK, NIB_MAGIC);
    const uint32_t a2 = lop_and_or(w >> 8, NIB_MASK, NIB_MAGIC);
#endif
    const uint32_t a3 = lop_and_or(rotr(w, 12), NIB_MASK, NIB_MAGIC);
#if defined(LLB_ABL_NO_MMA)  // timing ablation (wrong sums): de-quantise, no tensor-core instruction
    asm volatile("" ::"r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
    a[0] += 1.f;
#else
    mma_bf16_16816(a, a0, a1, a2, a3, b0, b1);
#endif
#endif
  };

#ifdef LLB_LOOPCLK  // tuning build: cycles each consumer warp spends waiting for its chunk vs in the whole loop
  long long lc_wait = 0;
  const long long lc_t0 = clock64();
#endif
  int it = i0 + warp * BPW;  // first item of this warp in the current chunk
  int vt = it / G, gk = it - vt * G;
  int slot = 0;
  uint32_t phase = 0;
  for (int ch = 0; ch < nch; ++ch) {
    // The BPW blocks a warp takes from a chunk share their (tile, matrix) and lie inside the CTA's range together
    // (the host guarantees G % BPW == 0).  So the tile-change test is ONE warp-uniform test per chunk, ahead of the
    // loads: the body below is straight-line code in which the fold of one block overlaps the MMAs of the next.
    // (Measured: a per-block test between the MMAs and the fold cost a third of the kernel's time.)
    const bool valid = it < i1;
    if (valid && vt != cur_vt) {
      if (cur_vt >= 0) flush(cur_vt);
      cur_vt = vt;
      out[0] = out[1] = out[2] = out[3] = 0.f;
    }
#ifdef LLB_LOOPCLK
    const long long lc_a = clock64();
#endif
    mbar_wait_u32(full_u + slot * 8, phase);
#ifdef LLB_LOOPCLK
    lc_wait += clock64() - lc_a;
#endif
    if (valid) {
      const uint32_t blk0 = ring_u + (uint32_t)(slot * CHB);
      float t[BPW][4];
      uint2 sz[BPW], sxv[BPW];
#pragma unroll
      for (int b = 0; b < BPW; ++b) {
        const uint32_t blk = blk0 + b * BLK;
        const int gg = gk + b;
        sz[b] = lds64(blk + KB * 8 + g * 8);
        sxv[b] = lds64(sx_lane + gg * 32);
        const uint32_t xk = xs_lane + (uint32_t)(gg * KB * 2);
        float acc[4][4];
#pragma unroll
        for (int i = 0; i < 4; ++i) acc[0][i] = acc[1][i] = acc[2][i] = acc[3][i] = 0.f;
        if constexpr (KSP == 1) {
          const uint2 wq = lds64(blk + lane * 8);
          const uint4 xv = LLB_XLD(xk);
          mma_word(acc[0], wq.x, xv.x, xv.y);
          mma_word(acc[1], wq.y, xv.z, xv.w);
        } else {
#pragma unroll
          for (int h = 0; h < KSP / 2; ++h) {
            const uint4 wq = lds128(blk + h * 512 + lane * 16);
            const uint4 xa = LLB_XLD(xk + (2 * h) * 64);
            const uint4 xb = LLB_XLD(xk + (2 * h + 1) * 64);
            mma_word(acc[0], wq.x, xa.x, xa.y);
            mma_word(acc[1], wq.y, xa.z, xa.w);
            mma_word(acc[2], wq.z, xb.x, xb.y);
            mma_word(acc[3], wq.w, xb.z, xb.w);
          }
        }
#pragma unroll
        for (int i = 0; i < 4; ++i) t[b][i] = (acc[0][i] + acc[1][i]) + (acc[2][i] + acc[3][i]);
      }
      // the folds come after all of the chunk's MMAs (measured: folding block b behind block b+1's MMAs and
      // handing the slot back before the last fold is 7 % slower, profiles/r01_gemv_experiments.md)
#pragma unroll
      for (int b = 0; b < BPW; ++b) {
        const float s0 = __uint_as_float(sz[b].x << 16), zn0 = __uint_as_float(sz[b].x & 0xffff0000u);
        const float s1 = __uint_as_float(sz[b].y << 16), zn1 = __uint_as_float(sz[b].y & 0xffff0000u);
        const float sx0 = __uint_as_float(sxv[b].x), sx1 = __uint_as_float(sxv[b].y);
        // sum (16+q) x  -  (16+z) sum x  =  sum (q - z) x ; then the group scale
        out[0] = fmaf(s0, fmaf(zn0, sx0, t[b][0]), out[0]);
        out[1] = fmaf(s0, fmaf(zn0, sx1, t[b][1]), out[1]);
        out[2] = fmaf(s1, fmaf(zn1, sx0, t[b][2]), out[2]);
        out[3] = fmaf(s1, fmaf(zn1, sx1, t[b][3]), out[3]);
      }
    }
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty_u + slot * 8) : "memory");
    it += CH;
    gk += CH;
    while (gk >= G) {
      gk -= G;
      ++vt;
    }
    if (++slot == S) {
      slot = 0;
      phase ^= 1u;
    }
  }
  if (cur_vt >= 0) flush(cur_vt);
#ifdef LLB_LOOPCLK
  if (p.trace != nullptr && lane == 0) {
    atomicAdd(&p.trace[blockIdx.x * 16 + 11], (unsigned long long)lc_wait);
    atomicAdd(&p.trace[blockIdx.x * 16 + 12], (unsigned long long)(clock64() - lc_t0));
    if (warp == 0) p.trace[blockIdx.x * 16 + 13] = (unsigned long long)nch;
  }
#endif
  TRACE(4);
  consumer_sync(NCONS);

  TRACE(5);
  // ---- fused epilogue, one warp per tile (every tile is complete inside its CTA)
  float* vals = wv + (size_t)warp * nmat * MT;  // [nmat][M][16]
  for (int T = T_first + warp; T <= T_last; T += W) {
    const int n0 = T * TILE_N;
    float pre[4] = {pre0[0], pre0[1], pre0[2], pre0[3]};
    float2 cs[2] = {cs0[0], cs0[1]};
    if (T != T_first + warp) load_pre(T, pre, cs);
    // cross-warp reduction of this tile's sums in fixed warp order (deterministic)
    for (int e = lane; e < nmat * MT; e += 32) {
      const float* r = red + (size_t)(T * nmat - vt_first) * W * MT + (e / MT) * W * MT + (e % MT);
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < W; ++w) v += r[w * MT];
      vals[e] = v;
    }
    __syncwarp();
    if (epilogue == LL_EPI_PUSH) {
      // row-parallel layer under tensor parallelism: this rank's partial sums go straight into every rank's
      // receive buffer over NVLink (fire-and-forget stores; the arrival is signalled once per CTA below)
      for (int e = lane; e < MT; e += 32) {
        const int m = e >> 4, n = n0 + (e & 15);
        const __nv_bfloat16 v = __float2bfloat16_rn(vals[e]);
        for (int r = 0; r < p.push_world; ++r) p.push_dst[r][(size_t)m * p.N + n] = v;
      }
    } else if (epilogue == LL_EPI_STORE) {
      const float bias = pre[0];
      for (int e = lane; e < MT; e += 32) {
        const int m = e >> 4, n = n0 + (e & 15);
        p.y[(size_t)m * p.N + n] = __float2bfloat16_rn(vals[e] + bias);
      }
    } else if (epilogue == LL_EPI_RESIDUAL) {
      // x = bf16(x + bf16(h))   (ref model.py:172-173)
#pragma unroll
      for (int j = 0; j < 4; ++j) {
        const int e = lane + 32 * j;
        if (e < MT) {
          const float h = bf16r(vals[e]);
          p.y[(size_t)(e >> 4) * p.N + n0 + (e & 15)] = __float2bfloat16_rn(__fadd_rn(pre[j], h));
        }
      }
    } else if (epilogue == LL_EPI_ROPE_KV) {
      const int E = p.N / 3, hs = p.head_size;
#pragma unroll
      for (int j = 0; j < 2; ++j) {
        const int e = lane + 32 * j;
        if (e < M * (TILE_N / 2)) {
          const int m = e >> 3, cc = 2 * (e & 7), n = n0 + cc;
          const int part = n / E, ee = n - part * E, head = ee / hs, d = ee - head * hs;
          float x0 = bf16r(vals[m * TILE_N + cc]), x1 = bf16r(vals[m * TILE_N + cc + 1]);
          if (part < 2) {
            const float o0 = __fsub_rn(__fmul_rn(x0, cs[j].x), __fmul_rn(x1, cs[j].y));
            const float o1 = __fadd_rn(__fmul_rn(x1, cs[j].x), __fmul_rn(x0, cs[j].y));
            x0 = o0; x1 = o1;
          }
          const __nv_bfloat162 o = __floats2bfloat162_rn(x0, x1);
          if (part == 0) {
            *reinterpret_cast<__nv_bfloat162*>(p.y + (size_t)m * E + ee) = o;
          } else {
            __nv_bfloat16* cache = part == 1 ? p.k_cache : p.v_cache;
            *reinterpret_cast<__nv_bfloat162*>(cache + (((size_t)m * p.n_head + head) * p.cache_len + kv_slot) * hs + d) = o;
          }
        }
      }
    } else {  // LL_EPI_SWIGLU
      for (int e = lane; e < MT; e += 32) {
        const int m = e >> 4, n = n0 + (e & 15);
        const float a = bf16r(vals[e]), b = bf16r(vals[MT + e]);
        const float sl = bf16r(a / (1.0f + expf(-a)));
        p.y[(size_t)m * p.N + n] = __float2bfloat16_rn(__fmul_rn(sl, b));
      }
    }
    __syncwarp();
  }
  if (epilogue == LL_EPI_PUSH) {
    // all of this CTA's peer stores happen-before the barrier; one thread per rank then publishes them: the
    // release at system scope is cumulative over what the barrier made visible to that thread
    consumer_sync(NCONS);
    if (ctid < p.push_world) {
      asm volatile("fence.acq_rel.sys;" ::: "memory");
      asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(p.push_cnt[ctid]) : "memory");
    }
  }
  TRACE(6);
}

// ------------------------------------------------------------------------------------ repack
// One thread per 32-bit fragment word (and per tail entry).  ref layout: quantization.py:349-357, 385-388.
template <typename P>
__global__ void int4_repack_kernel(const uint8_t* __restrict__ qw, const P* __restrict__ scales, const P* __restrict__ zeros,
                                   uint8_t* __restrict__ out, int N, int K, int group, int KB, int32_t* status) {
  const int G = K / KB, Gq = K / group;
  const int BLK = block_bytes(KB);
  const int words = KB * 2;          // fragment words per block
  const int per_blk = words + TILE_N;  // + tail entries
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nblk = (long long)(N / TILE_N) * G;
  if (idx >= nblk * per_blk) return;
  const long long b = idx / per_blk;
  const int r = (int)(idx - b * per_blk);
  const int tile = (int)(b / G), gk = (int)(b - (long long)tile * G);
  uint8_t* blk = out + (size_t)b * BLK;
  if (r >= words) {  // tail: entry e -> column tile*16 + (e>>1) + 8*(e&1)
    const int e = r - words;
    const int n = tile * TILE_N + (e >> 1) + 8 * (e & 1);
    const int gq = (gk * KB) / group;
    const float s = dt<P>::ld(scales, (size_t)n * Gq + gq);
    const float z = dt<P>::ld(zeros, (size_t)n * Gq + gq);
    const float zn = -(16.f + z);
    if (!(bf16r(zn) == zn) || !(bf16r(s) == s)) atomicExch(status, 1);
    reinterpret_cast<uint32_t*>(blk + KB * 8)[e] = (uint32_t)f32_to_bf16_bits(s) | ((uint32_t)f32_to_bf16_bits(zn) << 16);
    return;
  }
  // word r: lane, step pair, which half of the pair
  int lane, sp, e0;
  if (KB == 32) {
    lane = r >> 1; sp = 0; e0 = (r & 1) * 4;
  } else {
    const int h = r >> 7, rr = r & 127;  // 128 words (512 B) per LDS.128 row of the warp
    lane = rr >> 2;
    const int j = rr & 3;
    sp = 2 * h + (j >> 1);
    e0 = (j & 1) * 4;
  }
  const int g = lane >> 2, tig = lane & 3;
  const int kbase = gk * KB + sp * 32 + tig * 8 + e0;
  const int n_lo = tile * TILE_N + g, n_hi = n_lo + 8;
  auto nib = [&](int n, int k) -> uint32_t { return (qw[(size_t)(k >> 1) * N + n] >> ((k & 1) * 4)) & 15u; };
  auto rotl = [](uint32_t v, int s) -> uint32_t { return (v << s) | (v >> (32 - s)); };
  uint32_t w = 0;
  w |= rotl(nib(n_lo, kbase + 0), 3) | rotl(nib(n_lo, kbase + 1), 19);
  w |= rotl(nib(n_hi, kbase + 0), 7) | rotl(nib(n_hi, kbase + 1), 23);
  w |= rotl(nib(n_lo, kbase + 2), 11) | rotl(nib(n_lo, kbase + 3), 27);
  w |= rotl(nib(n_hi, kbase + 2), 15) | rotl(nib(n_hi, kbase + 3), 31);
  const int off = (KB == 32) ? r * 4 : ((r >> 7) * 512 + (r & 127) * 4);
  *reinterpret_cast<uint32_t*>(blk + off) = w;
}


// ==================================================================================== batch-1 kernel, integer MMA
// The decode chain's kernel (M = 1, scale groups of a multiple of 128): the same ring / tile split / epilogues as above with
// a cheaper main loop.  profiles/r02_gemv_experiments.md: the bf16 loop above is bound by instruction ISSUE (ALU-pipe
// instructions hold the issue port for two cycles: 7 of them per 8 weights + bookkeeping = 24 slots per 32-bit word), not
// by HBM or the tensor pipe.  Here:
//  * stream kind 1 (ll_int4_prepare kind = 1): a 32-bit word holds 4 k-consecutive nibbles of output column g in the LOW
//    nibbles of its bytes and the same 4 k of column g + 8 in the HIGH nibbles, so `w & 0x0F0F0F0F` and
//    `(w >> 4) & 0x0F0F0F0F` ARE two A-fragment registers of mma.m16n8k32.u8.s8: 3 ALU instructions per 8 weights, one
//    MMA per 16 weights per lane.  Tail per block as before, with -zero in place of -(16 + zero).
//  * the activation row is turned into FOUR signed base-256 digit planes of x * 2^sh (sh per 128-element k-block from the
//    block's largest magnitude; exact for every element within 2^22 of it) and the digits are the first four B COLUMNS of
//    the same MMA: Sum_k q x = Sum_d 256^d (q . digit_d) with every product and sum exact in int32; the digits of a lane pair
//    are combined in integers, converted once per block and folded with the group scale and zero in fp32.  No (16 + q)
//    offset, so no cancellation: this path is closer to the exact sum than the bf16 one.
//  * x staging costs one more shuffle reduction per block (its exponent) and the digit transposition; no extra barrier.
__device__ __forceinline__ void mma_u8s8_16832(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                               uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint32_t bf16x2_mul(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t bf16x2_max(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("max.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
// c + a.lo * b.lo and c + a.hi * b.hi with bf16 factors and an fp32 sum: one FHFMA each (native on sm_100a)
__device__ __forceinline__ float fhfma_lo(uint32_t a, uint32_t b, float c) {
  float r;
  asm("{\n\t.reg .b16 al, ah, bl, bh;\n\tmov.b32 {al, ah}, %1;\n\tmov.b32 {bl, bh}, %2;\n\tfma.rn.f32.bf16 %0, al, bl, %3;\n\t}"
      : "=f"(r) : "r"(a), "r"(b), "f"(c));
  return r;
}
__device__ __forceinline__ float fhfma_hi(uint32_t a, uint32_t b, float c) {
  float r;
  asm("{\n\t.reg .b16 al, ah, bl, bh;\n\tmov.b32 {al, ah}, %1;\n\tmov.b32 {bl, bh}, %2;\n\tfma.rn.f32.bf16 %0, ah, bh, %3;\n\t}"
      : "=f"(r) : "r"(a), "r"(b), "f"(c));
  return r;
}
__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}

constexpr int I8_W = 8, I8_BPW = 2, I8_KB = 128;
constexpr int I8_BLK = I8_KB * 8 + TAIL_BYTES;
constexpr int I8_CH = I8_W * I8_BPW, I8_CHB = I8_CH * I8_BLK, I8_NCONS = I8_W * 32;
constexpr int I8_BLK8 = I8_KB * 16 + TAIL_BYTES;  // 8-bit weights: 16 columns x 128 k bytes + tail
constexpr uint32_t NIB_LO = 0x0F0F0F0Fu;

// WB = 8: the same kernel for GPTQ 8-bit weights (`gptq.int8`, ref utils.py:182-184 / quantization.py:343-347, for which
// the reference has only its de-quantise + F.linear branch).  Stream kind 2: a block is 16 columns x 128 k bytes; its words
// ARE the u8 A-fragment registers (lane's uint4 number i = {row g, row g + 8, row g k + 16, row g + 8 k + 16} of MMA i), so
// the main loop has no extraction arithmetic at all: 4 LDS.128 + 4 IMMA per block.  One block per warp per chunk (a chunk
// is again 17 KB), so any number of k-blocks per row is fine.
template <int EPI, int NORM, int WB = 4>
__global__ void __launch_bounds__((I8_W + 9) * 32, 1) gemv_w4a8_kernel(const __grid_constant__ Params p) {
  constexpr int W = I8_W, BPW = WB == 8 ? 1 : I8_BPW, KB = I8_KB, BLK = WB == 8 ? I8_BLK8 : I8_BLK, CH = W * BPW,
                CHB = CH * BLK, NCONS = I8_NCONS, TAIL_OFF = WB == 8 ? KB * 16 : KB * 8;
  constexpr int MAX_SLOTS_C = 12;
  constexpr int MT = TILE_N;  // floats per (tile, matrix) result at batch 1
  constexpr bool has_norm = NORM != 0;
  constexpr int epilogue = EPI;

  extern __shared__ __align__(128) uint8_t smem[];
  uint8_t* ring = smem;
  uint8_t* xs = smem + p.off_xs;                            // 4 digit planes [K + 16] + 128 zero bytes
  float* rec = reinterpret_cast<float*>(smem + p.off_sx);   // [G][2][2] (scale of the digits, zero-point sum) + 4 zeros
  float* red = reinterpret_cast<float*>(smem + p.off_red);  // [ts_max][W][16]
  float* wv = reinterpret_cast<float*>(smem + p.off_wv);    // [W][nmat][16]
  uint64_t* full = reinterpret_cast<uint64_t*>(smem + p.off_bar);
  uint64_t* empty = full + p.S;
  __shared__ float s_part[W];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int P = gridDim.x, c = blockIdx.x;
  const int i0 = range_start(c, p.ntiles, P) * p.GT;
  const int i1 = range_start(c + 1, p.ntiles, P) * p.GT;
  const int nch = (i1 - i0 + CH - 1) / CH;
  const int G = p.G, S = p.S;

  TRACE(0);
  if (threadIdx.x == 0) {
    for (int i = 0; i < S; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], W);
    }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();
  // idle warps of an ODD-parity launch: they only widen the CTA's register footprint, so they must stay resident (an
  // exited warp gives its registers back) - parked on a named barrier that warp 0 releases at the end.  p.helpers (kernels
  // without the fused norm): before parking they take their share of the x conversion, which runs at 0.45 instructions per
  // clock with two warps per scheduler (dependent-instruction latency): more warps shorten it.
  const bool helper = warp > W;
  const bool helping = !has_norm && p.helpers != 0;  // uniform over the CTA
  if (helper && !helping) {
    asm volatile("bar.sync 2, %0;" ::"r"((int)blockDim.x - W * 32) : "memory");
    return;
  }

  if (warp == W) {
    // =============================== producer warp (as in the bf16 kernel) ===============================
    if (lane == 0) {
      const uint64_t pol = policy_evict_first();
      int slot = 0;
      uint32_t phase = 1;
      int it = i0;
      int vt = it / G, g = it - vt * G;
      for (int ch = 0; ch < nch; ++ch, ++slot) {
        if (slot == S) {
          slot = 0;
          phase ^= 1u;
        }
        if (ch >= S) mbar_wait_u32(smem_u32(&empty[slot]), phase);
        const int b = min(it + CH, i1);
        mbar_expect_tx(&full[slot], (uint32_t)(b - it) * BLK);
        uint8_t* dst = ring + (size_t)slot * CHB;
        while (it < b) {
          const int run = min(G - g, b - it);
          const int mat = p.nmat == 2 ? (vt & 1) : 0, tile = p.nmat == 2 ? (vt >> 1) : vt;
          const uint8_t* src = p.w[mat] + ((size_t)tile * G + g) * BLK;
          bulk_load_1d_hint(dst, src, (uint32_t)run * BLK, &full[slot], pol);
          dst += (size_t)run * BLK;
          it += run;
          g += run;
          if (g == G) {
            g = 0;
            ++vt;
          }
        }
      }
    }
    return;
  }

  // ================================= consumer warps =================================
  const int g = lane >> 2, tig = lane & 3;
  const int ctid = threadIdx.x;
  const int vt_first = i0 / G, vt_last = (i1 - 1) / G;
  const int nvt = vt_last - vt_first + 1;
  const int vpr = p.K >> 3;
  const int pstride = p.xs_stride;  // bytes between digit planes (K + 16: neighbouring planes land in different banks)

  uint8_t* nws = smem + p.off_nw;
  uint64_t* xbar = full + 2 * MAX_SLOTS_C;
  // threads that convert x: the 256 consumer threads, plus the helper warps (numbered after them; the producer warp is not one)
  const int stid = helper ? NCONS + ((int)threadIdx.x - (W + 1) * 32) : ctid;
  const int nst = helping ? (int)blockDim.x - 32 : NCONS;
  if (ctid == 0) {
    mbar_init(&xbar[0], 1);
    mbar_init(&xbar[1], 1);
    fence_barrier_init();
    if (has_norm) {
      mbar_expect_tx(&xbar[1], (uint32_t)p.K * 2);
      bulk_load_1d(nws, p.norm_w, (uint32_t)p.K * 2, &xbar[1]);
    }
  }
  if (!helper) {
    for (int i = ctid; i < nvt * W * MT; i += NCONS) red[i] = 0.f;
    if (ctid < 32) *reinterpret_cast<uint32_t*>(xs + 4 * (size_t)pstride + ctid * 4) = 0u;  // zero area of the B fragments
    if (ctid < 4) rec[(size_t)G * 4 + ctid] = 0.f;                                           // and of the records
  }
  // every converting thread meets here once before the staging (also what tells ptxas that the warps are converged: without
  // a barrier in front of it every shuffle of the staging loop is compiled behind a WARPSYNC)
  if (helping) asm volatile("bar.sync 3, %0;" ::"r"(nst) : "memory");
  else consumer_sync(NCONS);

  TRACE(1);
  pdl_wait();
  TRACE(2);
  int rope_pos = 0, kv_slot = 0;
  if (epilogue == LL_EPI_ROPE_KV) {
    rope_pos = p.pos[0];
    kv_slot = p.pos[1];
  }

  // ---- stage x.  One 16-byte piece (8 elements) per thread per pass of 256 pieces; the 16 threads of a half-warp hold
  //      one 128-element k-block.  A thread's pieces are fetched TOGETHER (one L2 round trip for K <= 16384; longer rows
  //      in groups of XR pieces).  With the fused RMSNorm the whole row is fetched first (its sum of squares needs all of
  //      it); the first group stays in registers for the second pass, later groups are fetched again (an L2 hit).
  //      Arithmetic on packed bf16 pairs (HMUL2.BF16 rounds exactly where the reference's bf16 ops round) and
  //      fp32 += bf16 * bf16 (FHFMA): a fifth of the instructions of the scalar fp32 form.
  bool staged = false;
  if constexpr (!has_norm) {
    if (p.xdg_planes != nullptr) {
      // the producer (the attention kernel's leader CTAs) already wrote x as digit planes and records in this kernel's
      // shared-memory layout: two bulk copies instead of a conversion
      if (ctid == 0) {
        mbar_expect_tx(&xbar[0], (uint32_t)(4 * pstride + G * 16));
        bulk_load_1d(xs, p.xdg_planes, (uint32_t)(4 * pstride), &xbar[0]);
        bulk_load_1d(rec, p.xdg_rec, (uint32_t)(G * 16), &xbar[0]);
      }
      mbar_wait_u32(smem_u32(&xbar[0]), 0);
      TRACE(10);
      consumer_sync(NCONS);
      staged = true;
    }
  }
  if (!staged) {
    constexpr int XR = 8;
    constexpr uint32_t ONE2 = 0x3f803f80u;  // bf16x2 (1, 1)
    uint4 raw[XR];
    uint32_t rn2 = 0u;
    const uint4* xsrc = reinterpret_cast<const uint4*>(p.x);
    auto fetch = [&](int v0) {
  #pragma unroll
      for (int j = 0; j < XR; ++j) {
        const int v = v0 + j * nst + stid;
        raw[j] = v < vpr ? __ldcg(xsrc + v) : make_uint4(0u, 0u, 0u, 0u);
      }
    };
    fetch(0);
    if (has_norm) {
      float ss = 0.f;
      for (int v0 = 0; v0 < vpr; v0 += XR * NCONS) {
        uint4 r4[XR];
        if (v0 == 0) {
  #pragma unroll
          for (int j = 0; j < XR; ++j) r4[j] = raw[j];
        } else {
  #pragma unroll
          for (int j = 0; j < XR; ++j) {
            const int v = v0 + j * NCONS + ctid;
            r4[j] = v < vpr ? __ldcg(xsrc + v) : make_uint4(0u, 0u, 0u, 0u);
          }
        }
  #pragma unroll
        for (int j = 0; j < XR; ++j) {
          if (v0 + j * NCONS < vpr) {  // uniform
            const uint32_t wd[4] = {r4[j].x, r4[j].y, r4[j].z, r4[j].w};
  #pragma unroll
            for (int e = 0; e < 4; ++e) {
              const uint32_t sq = bf16x2_mul(wd[e], wd[e]);  // bf16(x * x), ref model.py:280
              ss = fhfma_lo(sq, ONE2, ss);
              ss = fhfma_hi(sq, ONE2, ss);
            }
          }
        }
      }
      ss = warp_sum(ss);
      if (lane == 0) s_part[warp] = ss;
      TRACE(9);
      consumer_sync(NCONS);
      float tot = 0.f;
  #pragma unroll
      for (int w = 0; w < W; ++w) tot += s_part[w];
      const float mean = bf16r(tot * (1.0f / (float)p.K));
      const float rn = bf16r(rsqrtf(bf16r(mean + p.eps)));
      rn2 = (__float_as_uint(rn) >> 16) * 0x00010001u;
      mbar_wait_u32(smem_u32(&xbar[1]), 0);
      TRACE(7);
    }
    // The conversion of one piece is a chain of dependent instructions (normalise, block maximum and sum over 16 lanes,
    // exponent, digits, byte transposition): NP pieces of a thread go through it side by side - two with the fused norm
    // (K = 4096: exactly the thread's two pieces), one otherwise (with helper warps most threads have a single piece).
    constexpr int NP = has_norm ? 2 : 1;
    for (int v0 = 0; v0 < vpr; v0 += XR * nst) {
      if (v0 > 0) fetch(v0);
  #pragma unroll
      for (int j = 0; j < XR; j += NP) {
        if (v0 + j * nst >= vpr) continue;  // uniform over the CTA
        int v[NP];
        bool ok[NP];
        uint32_t wd[NP][4];
  #pragma unroll
        for (int q = 0; q < NP; ++q) {
          v[q] = v0 + (j + q) * nst + stid;
          ok[q] = v[q] < vpr;  // uniform over a half-warp: K is a multiple of 128
          wd[q][0] = raw[j + q].x; wd[q][1] = raw[j + q].y; wd[q][2] = raw[j + q].z; wd[q][3] = raw[j + q].w;
        }
        if (has_norm) {  // ref model.py:276-283 rounding steps: bf16(w * bf16(x * r))
          uint4 nraw[NP];
  #pragma unroll
          for (int q = 0; q < NP; ++q) {
            nraw[q] = make_uint4(0u, 0u, 0u, 0u);
            if (ok[q]) nraw[q] = *reinterpret_cast<const uint4*>(nws + (size_t)v[q] * 16);
          }
  #pragma unroll
          for (int q = 0; q < NP; ++q) {
            const uint32_t nw[4] = {nraw[q].x, nraw[q].y, nraw[q].z, nraw[q].w};
  #pragma unroll
            for (int e = 0; e < 4; ++e) wd[q][e] = bf16x2_mul(nw[e], bf16x2_mul(wd[q][e], rn2));
          }
        }
        // the block's exponent from the largest magnitude of its 128 elements (16 lanes); the sum of the block for the zero term
        uint32_t m2[NP];
        float sum[NP];
  #pragma unroll
        for (int q = 0; q < NP; ++q) {
          m2[q] = bf16x2_max(bf16x2_max(wd[q][0] & 0x7fff7fffu, wd[q][1] & 0x7fff7fffu),
                             bf16x2_max(wd[q][2] & 0x7fff7fffu, wd[q][3] & 0x7fff7fffu));
          float s0 = 0.f, s1 = 0.f;  // two chains
  #pragma unroll
          for (int e = 0; e < 4; e += 2) {
            s0 = fhfma_lo(wd[q][e], ONE2, s0);
            s1 = fhfma_lo(wd[q][e + 1], ONE2, s1);
            s0 = fhfma_hi(wd[q][e], ONE2, s0);
            s1 = fhfma_hi(wd[q][e + 1], ONE2, s1);
          }
          sum[q] = s0 + s1;
        }
  #pragma unroll
        for (int o = 8; o > 0; o >>= 1) {
  #pragma unroll
          for (int q = 0; q < NP; ++q) {
            m2[q] = bf16x2_max(m2[q], __shfl_xor_sync(0xffffffffu, m2[q], o));
            sum[q] += __shfl_xor_sync(0xffffffffu, sum[q], o);
          }
        }
        uint32_t u[NP][8];
        float inv[NP];
  #pragma unroll
        for (int q = 0; q < NP; ++q) {
          int ex = (int)max((m2[q] >> 7) & 0xffu, (m2[q] >> 23) & 0xffu);  // biased exponent of the largest magnitude: am < 2^(ex - 126)
          ex = max(ex, 32);                                                  // all-zero / denormal blocks: any finite scale will do
          const uint32_t scale2 = ((uint32_t)(283 - ex) << 7) * 0x00010001u;  // bf16x2 2^(156 - ex): am * scale < 2^30
          inv[q] = __uint_as_float((uint32_t)(ex - 29) << 23);                // 2^-(156 - ex)
  #pragma unroll
          for (int e = 0; e < 4; ++e) {
            const int v0i = __float2int_rn(fhfma_lo(wd[q][e], scale2, 0.f)), v1i = __float2int_rn(fhfma_hi(wd[q][e], scale2, 0.f));
            u[q][2 * e] = ((uint32_t)v0i + 0x00808080u) ^ 0x00808080u;  // bytes = four signed base-256 digits
            u[q][2 * e + 1] = ((uint32_t)v1i + 0x00808080u) ^ 0x00808080u;
          }
        }
  #pragma unroll
        for (int q = 0; q < NP; ++q) {
          if (ok[q]) {
            // 4 x 4 byte transpositions: plane d gets byte d of eight consecutive elements
  #pragma unroll
            for (int h = 0; h < 2; ++h) {
              const uint32_t a = prmt(u[q][4 * h], u[q][4 * h + 1], 0x5140u), b = prmt(u[q][4 * h], u[q][4 * h + 1], 0x7362u);
              const uint32_t cc = prmt(u[q][4 * h + 2], u[q][4 * h + 3], 0x5140u), d = prmt(u[q][4 * h + 2], u[q][4 * h + 3], 0x7362u);
              uint8_t* dst = xs + (size_t)v[q] * 8 + h * 4;
              *reinterpret_cast<uint32_t*>(dst) = prmt(a, cc, 0x5410u);
              *reinterpret_cast<uint32_t*>(dst + pstride) = prmt(a, cc, 0x7632u);
              *reinterpret_cast<uint32_t*>(dst + 2 * (size_t)pstride) = prmt(b, d, 0x5410u);
              *reinterpret_cast<uint32_t*>(dst + 3 * (size_t)pstride) = prmt(b, d, 0x7632u);
            }
            if ((lane & 15) == 0) {  // record of the block: lanes tig = 0 take (inv, sum x), lanes tig = 1 take (inv * 2^16, 0)
              float4 r;
              r.x = inv[q];
              r.y = sum[q];
              r.z = inv[q] * 65536.f;
              r.w = 0.f;
              *reinterpret_cast<float4*>(rec + (size_t)(v[q] >> 4) * 4) = r;
            }
          }
        }
      }
    }
    TRACE(10);
    if (helping) {
      // the consumers wait for every converting thread on named barrier 3; the helper warps only arrive, then park
      if (helper) {
        asm volatile("bar.arrive 3, %0;" ::"r"(nst) : "memory");
        asm volatile("bar.sync 2, %0;" ::"r"((int)blockDim.x - W * 32) : "memory");
        return;
      }
      asm volatile("bar.sync 3, %0;" ::"r"(nst) : "memory");
    } else {
      consumer_sync(NCONS);
    }
}
  TRACE(3);

  // epilogue operands that do not depend on this kernel's sums, fetched for this warp's first tile behind the main loop
  const int nmat = p.nmat;
  const int T_first = vt_first / nmat, T_last = vt_last / nmat;
  auto load_pre = [&](int T, float& pre, float2& cs) {
    const int n0 = T * TILE_N;
    if (epilogue == LL_EPI_RESIDUAL) {
      if (lane < MT) pre = __bfloat162float(p.y[n0 + lane]);
    } else if (epilogue == LL_EPI_STORE && p.bias != nullptr) {
      pre = __bfloat162float(p.bias[n0 + (lane & 15)]);
    } else if (epilogue == LL_EPI_ROPE_KV) {
      const int E = p.N / 3, hs = p.head_size;
      const int n = n0 + 2 * (lane & 7);
      const int part = n / E, d = (n - part * E) % hs;
      if (lane < TILE_N / 2 && part < 2)
        cs = *reinterpret_cast<const float2*>(p.rope + ((size_t)rope_pos * (hs / 2) + d / 2) * 2);
    }
  };
  float pre0 = 0.f;
  float2 cs0 = make_float2(1.f, 0.f);
  if (T_first + warp <= T_last) load_pre(T_first + warp, pre0, cs0);

  // ---- main loop
  // per-lane addressing: lanes with g < 4 read digit plane g (32 bytes per block at tig * 32), the others a zero area with
  // stride 0; lanes with tig < 2 read their record, the others a zero record with stride 0
  const uint32_t xs_lane = smem_u32(xs) + (g < 4 ? (uint32_t)(g * pstride + tig * 32) : (uint32_t)(4 * pstride + tig * 32));
  const uint32_t xs_step = g < 4 ? (uint32_t)KB : 0u;
  const uint32_t rec_lane = smem_u32(rec) + (tig < 2 ? (uint32_t)(tig * 8) : (uint32_t)(G * 16));
  const uint32_t rec_step = tig < 2 ? 16u : 0u;
  const uint32_t ring_u = smem_u32(ring) + (uint32_t)(warp * BPW * BLK);
  const uint32_t full_u = smem_u32(full), empty_u = smem_u32(empty);

  float out0 = 0.f, out1 = 0.f;  // columns g and g + 8 of the current tile, this lane's digit pair
  int cur_vt = -1;
  auto flush = [&](int vt) {
    float v0 = out0, v1 = out1;
    v0 += __shfl_xor_sync(0xffffffffu, v0, 1);
    v1 += __shfl_xor_sync(0xffffffffu, v1, 1);
    v0 += __shfl_xor_sync(0xffffffffu, v0, 2);
    v1 += __shfl_xor_sync(0xffffffffu, v1, 2);
    if (tig == 0) {
      float* r = red + ((size_t)(vt - vt_first) * W + warp) * MT;
      r[g] = v0;
      r[g + 8] = v1;
    }
  };
#ifdef LLB_LOOPCLK
  long long lc_wait = 0;
  const long long lc_t0 = clock64();
#endif
  int it = i0 + warp * BPW;
  int vt = it / G, gk = it - vt * G;
  int slot = 0;
  uint32_t phase = 0;
  for (int ch = 0; ch < nch; ++ch) {
    const bool valid = it < i1;
    if (valid && vt != cur_vt) {  // the two blocks a warp takes from a chunk share their tile (G is even)
      if (cur_vt >= 0) flush(cur_vt);
      cur_vt = vt;
      out0 = out1 = 0.f;
    }
#ifdef LLB_LOOPCLK
    const long long lc_a = clock64();
#endif
    mbar_wait_u32(full_u + slot * 8, phase);
#ifdef LLB_LOOPCLK
    lc_wait += clock64() - lc_a;
#endif
    if (valid) {
      const uint32_t blk0 = ring_u + (uint32_t)(slot * CHB);
      int d[BPW][4];
      uint2 sz[BPW], rc[BPW];
#pragma unroll
      for (int b = 0; b < BPW; ++b) {
        const uint32_t blk = blk0 + b * BLK;
        const uint32_t xk = xs_lane + (uint32_t)(gk + b) * xs_step;
        const uint4 xa = lds128(xk), xb = lds128(xk + 16);
        sz[b] = lds64(blk + TAIL_OFF + g * 8);
        rc[b] = lds64(rec_lane + (uint32_t)(gk + b) * rec_step);
        d[b][0] = d[b][1] = d[b][2] = d[b][3] = 0;
        if constexpr (WB == 8) {
          const uint4 w0 = lds128(blk + lane * 16), w1 = lds128(blk + 512 + lane * 16);
          const uint4 w2 = lds128(blk + 1024 + lane * 16), w3 = lds128(blk + 1536 + lane * 16);
          mma_u8s8_16832(d[b], w0.x, w0.y, w0.z, w0.w, xa.x, xa.y);
          mma_u8s8_16832(d[b], w1.x, w1.y, w1.z, w1.w, xa.z, xa.w);
          mma_u8s8_16832(d[b], w2.x, w2.y, w2.z, w2.w, xb.x, xb.y);
          mma_u8s8_16832(d[b], w3.x, w3.y, w3.z, w3.w, xb.z, xb.w);
        } else {
          const uint4 w0 = lds128(blk + lane * 16), w1 = lds128(blk + 512 + lane * 16);
          mma_u8s8_16832(d[b], w0.x & NIB_LO, (w0.x >> 4) & NIB_LO, w0.y & NIB_LO, (w0.y >> 4) & NIB_LO, xa.x, xa.y);
          mma_u8s8_16832(d[b], w0.z & NIB_LO, (w0.z >> 4) & NIB_LO, w0.w & NIB_LO, (w0.w >> 4) & NIB_LO, xa.z, xa.w);
          mma_u8s8_16832(d[b], w1.x & NIB_LO, (w1.x >> 4) & NIB_LO, w1.y & NIB_LO, (w1.y >> 4) & NIB_LO, xb.x, xb.y);
          mma_u8s8_16832(d[b], w1.z & NIB_LO, (w1.z >> 4) & NIB_LO, w1.w & NIB_LO, (w1.w >> 4) & NIB_LO, xb.z, xb.w);
        }
      }
#pragma unroll
      for (int b = 0; b < BPW; ++b) {
        // this lane's two digits of each column, combined in integers (|sum| < 2^26), converted once
        const float f0 = (float)(d[b][0] + (d[b][1] << 8)), f1 = (float)(d[b][2] + (d[b][3] << 8));
        const float s0 = __uint_as_float(sz[b].x << 16), zn0 = __uint_as_float(sz[b].x & 0xffff0000u);
        const float s1 = __uint_as_float(sz[b].y << 16), zn1 = __uint_as_float(sz[b].y & 0xffff0000u);
        const float e = __uint_as_float(rc[b].x), zs = __uint_as_float(rc[b].y);
        // scale * (2^-sh 256^(2 tig) digits  -  zero * sum x)
        out0 = fmaf(s0, fmaf(zn0, zs, e * f0), out0);
        out1 = fmaf(s1, fmaf(zn1, zs, e * f1), out1);
      }
    }
    __syncwarp();
    if (lane == 0) asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(empty_u + slot * 8) : "memory");
    it += CH;
    gk += CH;
    while (gk >= G) {
      gk -= G;
      ++vt;
    }
    if (++slot == S) {
      slot = 0;
      phase ^= 1u;
    }
  }
  if (cur_vt >= 0) flush(cur_vt);
#ifdef LLB_LOOPCLK
  if (p.trace != nullptr && lane == 0) {
    atomicAdd(&p.trace[blockIdx.x * 16 + 11], (unsigned long long)lc_wait);
    atomicAdd(&p.trace[blockIdx.x * 16 + 12], (unsigned long long)(clock64() - lc_t0));
    if (warp == 0) p.trace[blockIdx.x * 16 + 13] = (unsigned long long)nch;
  }
#endif
  TRACE(4);
  consumer_sync(NCONS);

  TRACE(5);
  // ---- fused epilogue, one warp per tile (every tile is complete inside its CTA); rounding steps as in the bf16 kernel
  float* vals = wv + (size_t)warp * nmat * MT;
  for (int T = T_first + warp; T <= T_last; T += W) {
    const int n0 = T * TILE_N;
    float pre = pre0;
    float2 cs = cs0;
    if (T != T_first + warp) load_pre(T, pre, cs);
    if (lane < nmat * MT) {
      const float* r = red + (size_t)(T * nmat - vt_first) * W * MT + (lane / MT) * W * MT + (lane % MT);
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < W; ++w) v += r[w * MT];
      vals[lane] = v;
    }
    __syncwarp();
    if (epilogue == LL_EPI_STORE) {
      if (lane < MT) p.y[n0 + lane] = __float2bfloat16_rn(vals[lane] + pre);
    } else if (epilogue == LL_EPI_RESIDUAL) {
      if (lane < MT) p.y[n0 + lane] = __float2bfloat16_rn(__fadd_rn(pre, bf16r(vals[lane])));   // ref model.py:172-173
    } else if (epilogue == LL_EPI_ROPE_KV) {
      const int E = p.N / 3, hs = p.head_size;
      if (lane < TILE_N / 2) {
        const int cc = 2 * lane, n = n0 + cc;
        const int part = n / E, ee = n - part * E, head = ee / hs, d = ee - head * hs;
        float x0 = bf16r(vals[cc]), x1 = bf16r(vals[cc + 1]);
        if (part < 2) {
          const float o0 = __fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y));
          const float o1 = __fadd_rn(__fmul_rn(x1, cs.x), __fmul_rn(x0, cs.y));
          x0 = o0; x1 = o1;
        }
        const __nv_bfloat162 o = __floats2bfloat162_rn(x0, x1);
        if (part == 0) {
          *reinterpret_cast<__nv_bfloat162*>(p.y + ee) = o;
        } else {
          __nv_bfloat16* cache = part == 1 ? p.k_cache : p.v_cache;
          *reinterpret_cast<__nv_bfloat162*>(cache + ((size_t)head * p.cache_len + kv_slot) * hs + d) = o;
        }
      }
    } else if (epilogue == LL_EPI_SWIGLU) {
      if (lane < MT) {
        const float a = bf16r(vals[lane]), b = bf16r(vals[MT + lane]);
        const float sl = bf16r(a / (1.0f + expf(-a)));
        p.y[n0 + lane] = __float2bfloat16_rn(__fmul_rn(sl, b));
      }
    }
    __syncwarp();
  }
  TRACE(6);
  if (blockDim.x > (W + 1) * 32 && warp == 0) asm volatile("bar.arrive 2, %0;" ::"r"((int)blockDim.x - W * 32) : "memory");
}

// stream kind 1: one thread per 32-bit word (and per tail entry)
template <typename P>
__global__ void int4_repack_u8_kernel(const uint8_t* __restrict__ qw, const P* __restrict__ scales, const P* __restrict__ zeros,
                                      uint8_t* __restrict__ out, int N, int K, int group, int32_t* status) {
  constexpr int KB = I8_KB, BLK = I8_BLK;
  const int G = K / KB, Gq = K / group;
  const int words = KB * 2;
  const int per_blk = words + TILE_N;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nblk = (long long)(N / TILE_N) * G;
  if (idx >= nblk * per_blk) return;
  const long long b = idx / per_blk;
  const int r = (int)(idx - b * per_blk);
  const int tile = (int)(b / G), gk = (int)(b - (long long)tile * G);
  uint8_t* blk = out + (size_t)b * BLK;
  if (r >= words) {  // tail: entry e -> column tile*16 + (e>>1) + 8*(e&1): (scale, -zero)
    const int e = r - words;
    const int n = tile * TILE_N + (e >> 1) + 8 * (e & 1);
    const int gq = (gk * KB) / group;
    const float s = dt<P>::ld(scales, (size_t)n * Gq + gq);
    const float zn = -dt<P>::ld(zeros, (size_t)n * Gq + gq);
    if (!(bf16r(zn) == zn) || !(bf16r(s) == s)) atomicExch(status, 1);
    reinterpret_cast<uint32_t*>(blk + KB * 8)[e] = (uint32_t)f32_to_bf16_bits(s) | ((uint32_t)f32_to_bf16_bits(zn) << 16);
    return;
  }
  // word r = (q, lane, j): the lane's word 4 q + j covers k = 32 tig + 4 (4 q + j) + 0..3 of the block
  const int q = r >> 7, lane = (r & 127) >> 2, j = r & 3;
  const int g = lane >> 2, tig = lane & 3;
  const int kbase = gk * KB + 32 * tig + 4 * (4 * q + j);
  const int n_lo = tile * TILE_N + g, n_hi = n_lo + 8;
  auto nib = [&](int n, int k) -> uint32_t { return (qw[(size_t)(k >> 1) * N + n] >> ((k & 1) * 4)) & 15u; };
  uint32_t w = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) w |= (nib(n_lo, kbase + i) << (8 * i)) | (nib(n_hi, kbase + i) << (8 * i + 4));
  *reinterpret_cast<uint32_t*>(blk + q * 512 + lane * 16 + j * 4) = w;
}

// ------------------------------------------------------------------------------------ host side
// stream kind 2 (8-bit weights, checkpoint memory [K][N] bytes): one thread per 32-bit word (and per tail entry)
template <typename P>
__global__ void q8_repack_kernel(const uint8_t* __restrict__ qw, const P* __restrict__ scales, const P* __restrict__ zeros,
                                 uint8_t* __restrict__ out, int N, int K, int group, int32_t* status) {
  constexpr int KB = I8_KB, BLK = I8_BLK8;
  const int G = K / KB, Gq = K / group;
  const int words = KB * 4;
  const int per_blk = words + TILE_N;
  const long long idx = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const long long nblk = (long long)(N / TILE_N) * G;
  if (idx >= nblk * per_blk) return;
  const long long b = idx / per_blk;
  const int r = (int)(idx - b * per_blk);
  const int tile = (int)(b / G), gk = (int)(b - (long long)tile * G);
  uint8_t* blk = out + (size_t)b * BLK;
  if (r >= words) {  // tail: entry e -> column tile*16 + (e>>1) + 8*(e&1): (scale, -zero)
    const int e = r - words;
    const int n = tile * TILE_N + (e >> 1) + 8 * (e & 1);
    const int gq = (gk * KB) / group;
    const float s = dt<P>::ld(scales, (size_t)n * Gq + gq);
    const float zn = -dt<P>::ld(zeros, (size_t)n * Gq + gq);
    if (!(bf16r(zn) == zn) || !(bf16r(s) == s)) atomicExch(status, 1);
    reinterpret_cast<uint32_t*>(blk + KB * 16)[e] = (uint32_t)f32_to_bf16_bits(s) | ((uint32_t)f32_to_bf16_bits(zn) << 16);
    return;
  }
  // word r = (MMA i, lane, c): c & 1 = row g or g + 8, c >> 1 = first or second 4 k of the lane's 8 k of that MMA;
  // a lane's 32 k of the block are k = 32 tig .. 32 tig + 31 (the digit planes' order, as in stream kind 1)
  const int i = r >> 7, lane = (r & 127) >> 2, c = r & 3;
  const int g = lane >> 2, tig = lane & 3;
  const int n = tile * TILE_N + g + 8 * (c & 1);
  const int kbase = gk * KB + 32 * tig + 8 * i + 4 * (c >> 1);
  uint32_t w = 0;
#pragma unroll
  for (int j = 0; j < 4; ++j) w |= (uint32_t)qw[(size_t)(kbase + j) * N + n] << (8 * j);
  *reinterpret_cast<uint32_t*>(blk + i * 512 + lane * 16 + c * 4) = w;
}

struct Plan {
  int KB, KSP, BLK, W, BPW;
  int G, GT, ntiles, I, P;
  int ts_max, S;
  int xs_stride, off_xs, off_sx, off_red, off_wv, off_bar, off_nw;
  size_t smem;
};

unsigned long long* g_trace = nullptr;  // debug: device buffer [launches][256][8]
int g_trace_launches = 0, g_trace_idx = 0;

int specialise_level();
constexpr int SMEM_CORESIDENT = 110 * 1024;  // two such CTAs (this GEMV + the next one, under PDL) fit one SM
constexpr int SMEM_SOLO = 220 * 1024;
constexpr int MAX_SLOTS = 12;

// consumer warps x blocks per warp per chunk; LLB200_GEMV_CFG="WxB" overrides the M <= 2 default (tuning)
void consumer_config(int M, int KSP, int* W, int* BPW) {
  static int env_w = -1, env_b = -1;
  if (env_w < 0) {
    env_w = 0; env_b = 0;
    if (const char* e = getenv("LLB200_GEMV_CFG")) {
      int w = 0, b = 0;
      if (sscanf(e, "%dx%d", &w, &b) == 2 && (w == 4 || w == 8 || w == 12 || w == 16) && (b == 1 || b == 2)) {
        env_w = w; env_b = b;
      }
    }
  }
  // measured on the 7B decode chain (profiles/r01_step_probe.md): 8 warps x 2 blocks beats 12 x 1 by 6 %
  if (M <= 2) { *W = 8; *BPW = KSP == 4 ? 2 : 1; } else if (M <= 4) { *W = 8; *BPW = 1; } else { *W = 4; *BPW = 1; }
  if (M <= 2 && KSP == 4 && env_w > 0) { *W = env_w; *BPW = env_b; }
}

// Fills `pl`; returns LL_OK, or LL_ERR_UNSUPPORTED with a message when the shape does not fit.
int make_plan(int N, int K, int group, int nmat, int M, int grid, bool has_norm, Plan* pl) {
  const int KB = pick_kb(K, group);
  if (KB == 0 || N % TILE_N != 0 || K % 8 != 0) {
    set_error("int4 GEMV needs N %% 16 == 0, K %% group == 0 and group %% 32 == 0 (N=%d K=%d group=%d)", N, K, group);
    return LL_ERR_UNSUPPORTED;
  }
  const int sms = sm_count();
  if (sms <= 0) {
    set_error("no CUDA device");
    return LL_ERR_CUDA;
  }
  Plan& q = *pl;
  q.KB = KB; q.KSP = KB / 32; q.BLK = block_bytes(KB);
  consumer_config(M, q.KSP, &q.W, &q.BPW);
  while ((K / KB) % q.BPW != 0) q.BPW >>= 1;  // the blocks a warp takes from one chunk must share their tile (kernel main loop)
  q.G = K / KB; q.GT = nmat * q.G; q.ntiles = N / TILE_N;
  const long long items = (long long)q.ntiles * q.GT;
  if (items > 0x3fffffff) {
    set_error("int4 GEMV: too many work items (%lld)", items);
    return LL_ERR_UNSUPPORTED;
  }
  q.I = (int)items;
  q.P = grid > 0 ? grid : sms;
  if (q.P > q.ntiles) q.P = q.ntiles;  // whole tiles per CTA
  if ((long long)q.ntiles * (q.P + 1) > 0xffffffffLL) {
    set_error("int4 GEMV: N too large (%d)", N);
    return LL_ERR_UNSUPPORTED;
  }
  const int tiles_max = (q.ntiles + q.P - 1) / q.P;
  const int ipc_max = tiles_max * q.GT;
  q.ts_max = tiles_max * nmat;
  const int MT = M * TILE_N;
  q.xs_stride = K * 2 + 64;
  size_t fixed = 0;
  auto take = [&](size_t bytes) { size_t o = fixed; fixed += (bytes + 127) & ~(size_t)127; return (int)o; };
  q.off_xs = take((size_t)(M + 1) * q.xs_stride);
  q.off_sx = take((size_t)q.G * 8 * 4);
  q.off_red = take((size_t)q.ts_max * q.W * MT * 4);
  q.off_wv = take((size_t)q.W * nmat * MT * 4);
  q.off_bar = take((size_t)(2 * MAX_SLOTS + 2) * 8);
  q.off_nw = take(has_norm ? (size_t)K * 2 : 0);
  const int ch = q.W * q.BPW;
  const size_t chb = (size_t)ch * q.BLK;
  const int nch_max = (ipc_max + ch - 1) / ch;
  long s = ((long)(q.W >= 12 ? SMEM_SOLO : SMEM_CORESIDENT) - (long)fixed) / (long)chb;
  if (s < 3 && nch_max > s) s = ((long)SMEM_SOLO - (long)fixed) / (long)chb;
  if (s > MAX_SLOTS) s = MAX_SLOTS;
  if (s > nch_max) s = nch_max;
  if (s < 1 || (s < 2 && nch_max > 1)) {
    set_error("int4 GEMV: M*K too large for shared memory (M=%d K=%d); use ll_qlinear_simt", M, K);
    return LL_ERR_UNSUPPORTED;
  }
  q.S = (int)s;
  const size_t ring = (size_t)q.S * chb;
  q.off_xs += (int)ring; q.off_sx += (int)ring; q.off_red += (int)ring;
  q.off_wv += (int)ring; q.off_bar += (int)ring; q.off_nw += (int)ring;
  q.smem = ring + fixed;
  return LL_OK;
}

template <int W, int KSP, int BPW, int EPI = -1, int NORM = -1, int MM = 0>
int launch_cfg(const Params& p, const Plan& pl, bool pdl, cudaStream_t st) {
  static size_t configured = 0;
  if (pl.smem > configured) {
    if (configured == 0) prefer_max_smem(gemv_int4_kernel<W, KSP, BPW, EPI, NORM, MM>);
    LL_CUDA(cudaFuncSetAttribute(gemv_int4_kernel<W, KSP, BPW, EPI, NORM, MM>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pl.smem));
    configured = pl.smem;
  }
  LL_CUDA(launch(gemv_int4_kernel<W, KSP, BPW, EPI, NORM, MM>, dim3(pl.P), dim3((W + 1) * 32), pl.smem, st, pdl, p));
  return LL_OK;
}

// LLB200_GEMV_SPECIALISE: 0 = the generic kernel, 1 = epilogue / norm as constants, 2 (default) = and batch 1 as a constant
int specialise_level() {
  static int lv = -1;
  if (lv < 0) {
    const char* e = getenv("LLB200_GEMV_SPECIALISE");
    lv = e ? atoi(e) : 2;
    if (lv < 0 || lv > 2) lv = 2;
  }
  return lv;
}

// Plan of the batch-1 integer-MMA kernel (stream kind 1): same tile split and ring as the bf16 kernel; shared memory holds
// four digit planes of x instead of bf16 rows.
// kind-1 stream: any whole number of 128-element k-blocks per scale group; the batch-1 GEMV additionally wants an even
// number of k-blocks (the two blocks a warp takes from a chunk share their tile)
bool i8_stream_ok(int N, int K, int group) {
  return group > 0 && group % I8_KB == 0 && K > 0 && K % group == 0 && N > 0 && N % TILE_N == 0;
}
bool i8_shape_ok(int N, int K, int group, int M) { return M == 1 && i8_stream_ok(N, K, group) && (K / I8_KB) % I8_BPW == 0; }
// 8-bit weights (stream kind 2): one block per warp per chunk, so any whole number of k-blocks
bool q8_shape_ok(int N, int K, int group, int M) { return M == 1 && i8_stream_ok(N, K, group); }

int make_plan_i8(int N, int K, int group, int nmat, int grid, bool has_norm, Plan* pl, int budget = SMEM_CORESIDENT, int wbits = 4) {
  if (!(wbits == 8 ? q8_shape_ok(N, K, group, 1) : i8_shape_ok(N, K, group, 1))) {
    set_error("int4 GEMV (stream kind 1) needs M = 1, N %% 16 == 0, group %% 128 == 0, K %% 256 == 0 (N=%d K=%d group=%d)", N, K, group);
    return LL_ERR_UNSUPPORTED;
  }
  const int sms = sm_count();
  if (sms <= 0) {
    set_error("no CUDA device");
    return LL_ERR_CUDA;
  }
  Plan& q = *pl;
  q.KB = I8_KB; q.KSP = 4; q.BLK = wbits == 8 ? I8_BLK8 : I8_BLK; q.W = I8_W; q.BPW = wbits == 8 ? 1 : I8_BPW;
  q.G = K / I8_KB; q.GT = nmat * q.G; q.ntiles = N / TILE_N;
  const long long items = (long long)q.ntiles * q.GT;
  if (items > 0x3fffffff) {
    set_error("int4 GEMV: too many work items (%lld)", items);
    return LL_ERR_UNSUPPORTED;
  }
  q.I = (int)items;
  q.P = grid > 0 ? grid : sms;
  if (q.P > q.ntiles) q.P = q.ntiles;
  if ((long long)q.ntiles * (q.P + 1) > 0xffffffffLL) {
    set_error("int4 GEMV: N too large (%d)", N);
    return LL_ERR_UNSUPPORTED;
  }
  const int tiles_max = (q.ntiles + q.P - 1) / q.P;
  const int ipc_max = tiles_max * q.GT;
  q.ts_max = tiles_max * nmat;
  q.xs_stride = K + 16;
  size_t fixed = 0;
  auto take = [&](size_t bytes) { size_t o = fixed; fixed += (bytes + 127) & ~(size_t)127; return (int)o; };
  q.off_xs = take((size_t)4 * q.xs_stride + 128);
  q.off_sx = take((size_t)q.G * 16 + 16);
  q.off_red = take((size_t)q.ts_max * q.W * TILE_N * 4);
  q.off_wv = take((size_t)q.W * nmat * TILE_N * 4);
  q.off_bar = take((size_t)(2 * MAX_SLOTS + 2) * 8);
  q.off_nw = take(has_norm ? (size_t)K * 2 : 0);
  const int ch = q.W * q.BPW;
  const size_t chb = (size_t)ch * q.BLK;
  const int nch_max = (ipc_max + ch - 1) / ch;
  long s = ((long)budget - (long)fixed) / (long)chb;
  if (s < 3 && nch_max > s) s = ((long)SMEM_SOLO - (long)fixed) / (long)chb;
  if (s > MAX_SLOTS) s = MAX_SLOTS;
  if (s > nch_max) s = nch_max;
  if (s < 1 || (s < 2 && nch_max > 1)) {
    set_error("int4 GEMV: K too large for shared memory (K=%d)", K);
    return LL_ERR_UNSUPPORTED;
  }
  q.S = (int)s;
  const size_t ring = (size_t)q.S * chb;
  q.off_xs += (int)ring; q.off_sx += (int)ring; q.off_red += (int)ring;
  q.off_wv += (int)ring; q.off_bar += (int)ring; q.off_nw += (int)ring;
  q.smem = ring + fixed;
  return LL_OK;
}

constexpr int SMEM_EVEN = 116 * 1024;  // two EVEN CTAs (2 x 116 + 2 KB) exceed an SM's 228 KB; EVEN + ODD (116 + 108 + 2) fit
constexpr int SMEM_ODD = 108 * 1024;
int g_parity = 0;  // of the launch being issued: 0 none, LL_GEMV_EVEN, LL_GEMV_ODD

template <int EPI, int NORM, int WB = 4>
int launch_i8(const Params& p, const Plan& pl, bool pdl, cudaStream_t st) {
  static size_t configured = 0;
  size_t smem = pl.smem;
  int threads = (I8_W + 1) * 32;
  if (g_parity == LL_GEMV_EVEN && smem < (size_t)SMEM_EVEN) smem = SMEM_EVEN;  // the footprint, not the ring, is the point
  if (g_parity == LL_GEMV_ODD && smem <= (size_t)SMEM_ODD) {
    // idle warps until two of these CTAs need more than the 64 K registers of an SM (and one still fits beside an EVEN CTA
    // of 288 threads x <= 96 registers)
    static int odd_threads = 0;
    if (odd_threads == 0) {
      cudaFuncAttributes fa;
      LL_CUDA(cudaFuncGetAttributes(&fa, gemv_w4a8_kernel<EPI, NORM, WB>));
      const int regs = ((fa.numRegs + 7) / 8) * 8;
      int t = ((32768 / regs) / 32 + 1) * 32;
      if (t < threads + 64) t = threads + 64;
      if (t > (I8_W + 9) * 32) t = (I8_W + 9) * 32;
      odd_threads = t;
    }
    threads = odd_threads;
  }
  if (smem > configured) {
    if (configured == 0) prefer_max_smem(gemv_w4a8_kernel<EPI, NORM, WB>);
    LL_CUDA(cudaFuncSetAttribute(gemv_w4a8_kernel<EPI, NORM, WB>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  LL_CUDA(launch(gemv_w4a8_kernel<EPI, NORM, WB>, dim3(pl.P), dim3(threads), smem, st, pdl, p));
  return LL_OK;
}

// debug: launches per specialised kernel [epilogue][fused norm][batch-1 constant] and of the generic kernels (last entry);
// read by the tests to assert WHICH instantiations a model-level run exercised (ll_debug_gemv_counts, not in the public header)
int g_spec_counts[5][2][2] = {};
int g_generic_count = 0;

int launch_plan(const Params& p, const Plan& pl, bool pdl, cudaStream_t st) {
  // the decode chain's configuration (batch <= 2, group % 128 == 0): one kernel per layer kind
  if (pl.W == 8 && pl.KSP == 4 && pl.BPW == 2 && specialise_level() > 0) {
    const bool norm = p.norm_w != nullptr;
    const bool m1 = p.M == 1 && specialise_level() > 1;
#define SPEC(e, n)                                                              \
  if (p.epilogue == e && norm == (n != 0)) {                                    \
    ++g_spec_counts[e][n][m1 ? 1 : 0];                                          \
    return m1 ? launch_cfg<8, 4, 2, e, n, 1>(p, pl, pdl, st) : launch_cfg<8, 4, 2, e, n>(p, pl, pdl, st); \
  }
    SPEC(LL_EPI_ROPE_KV, 1); SPEC(LL_EPI_RESIDUAL, 0); SPEC(LL_EPI_SWIGLU, 1); SPEC(LL_EPI_STORE, 1); SPEC(LL_EPI_STORE, 0);
#undef SPEC
  }
  ++g_generic_count;
#define CFG(w, k, b) \
  if (pl.W == w && pl.KSP == k && pl.BPW == b) return launch_cfg<w, k, b>(p, pl, pdl, st)
  CFG(12, 4, 1); CFG(8, 4, 2); CFG(12, 4, 2); CFG(8, 4, 1); CFG(16, 4, 1); CFG(16, 4, 2); CFG(4, 4, 1); CFG(4, 4, 2);
  CFG(12, 2, 1); CFG(8, 2, 1); CFG(4, 2, 1);
  CFG(12, 1, 1); CFG(8, 1, 1); CFG(4, 1, 1);
#undef CFG
  set_error("int4 GEMV: no kernel for W=%d KSP=%d BPW=%d", pl.W, pl.KSP, pl.BPW);
  return LL_ERR_UNSUPPORTED;
}

int launch_plan_i8(const Params& p, const Plan& pl, bool pdl, cudaStream_t st) {
  const bool norm = p.norm_w != nullptr;
#define SPEC_I8(e, n)                                                                               \
  if (p.epilogue == e && norm == (n != 0)) {                                                        \
    ++g_spec_counts[e][n][1];                                                                       \
    return pl.BLK == I8_BLK8 ? launch_i8<e, n, 8>(p, pl, pdl, st) : launch_i8<e, n>(p, pl, pdl, st); \
  }
  SPEC_I8(LL_EPI_ROPE_KV, 1); SPEC_I8(LL_EPI_RESIDUAL, 0); SPEC_I8(LL_EPI_SWIGLU, 1); SPEC_I8(LL_EPI_STORE, 1); SPEC_I8(LL_EPI_STORE, 0);
  SPEC_I8(LL_EPI_SWIGLU, 0); SPEC_I8(LL_EPI_RESIDUAL, 1); SPEC_I8(LL_EPI_ROPE_KV, 0);
#undef SPEC_I8
  set_error("int4 GEMV (stream kind 1): no kernel for epilogue %d", p.epilogue);
  return LL_ERR_UNSUPPORTED;
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

size_t ll_int4_stream_bytes(int N, int K, int group) {
  const int KB = pick_kb(K, group);
  if (KB == 0 || N <= 0 || N % TILE_N != 0 || K % 8 != 0) return 0;
  return (size_t)(N / TILE_N) * (K / KB) * block_bytes(KB);
}

int ll_int4_prepare(const uint8_t* qw, const void* scales, const void* zeros, int pdt, uint8_t* out, int N, int K,
                    int group, int kind, int32_t* status, ll_stream stream) {
  LL_CHECK_ARG(qw && scales && zeros && out && status, "null pointer");
  LL_CHECK_ARG(kind == LL_STREAM_BF16 || kind == LL_STREAM_U8, "bad stream kind %d", kind);
  if (kind == LL_STREAM_U8) {
    if (!i8_stream_ok(N, K, group)) {
      set_error("int4_prepare: stream kind 1 needs group %% 128 == 0, K %% group == 0, N %% 16 == 0 (N=%d K=%d group=%d)", N, K, group);
      return LL_ERR_UNSUPPORTED;
    }
    LL_CHECK_ARG(((uintptr_t)out & 15) == 0, "stream buffer must be 16-byte aligned");
    const long long nblk8 = (long long)(N / TILE_N) * (K / I8_KB);
    const long long total8 = nblk8 * (I8_KB * 2 + TILE_N);
    const unsigned blocks8 = (unsigned)((total8 + 255) / 256);
    cudaStream_t st8 = (cudaStream_t)stream;
    if (pdt == LL_F32) int4_repack_u8_kernel<float><<<blocks8, 256, 0, st8>>>(qw, (const float*)scales, (const float*)zeros, out, N, K, group, status);
    else if (pdt == LL_BF16) int4_repack_u8_kernel<__nv_bfloat16><<<blocks8, 256, 0, st8>>>(qw, (const __nv_bfloat16*)scales, (const __nv_bfloat16*)zeros, out, N, K, group, status);
    else {
      set_error("int4_prepare: unsupported param dtype %d", pdt);
      return LL_ERR_UNSUPPORTED;
    }
    LL_LAUNCH_OK("int4_repack_u8_kernel");
    return LL_OK;
  }
  const int KB = pick_kb(K, group);
  if (KB == 0 || N <= 0 || N % TILE_N != 0 || K % 8 != 0) {
    set_error("int4_prepare: unsupported shape N=%d K=%d group=%d", N, K, group);
    return LL_ERR_UNSUPPORTED;
  }
  LL_CHECK_ARG(((uintptr_t)out & 15) == 0, "stream buffer must be 16-byte aligned");
  const long long nblk = (long long)(N / TILE_N) * (K / KB);
  const long long total = nblk * (KB * 2 + TILE_N);
  const unsigned blocks = (unsigned)((total + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (pdt == LL_F32) int4_repack_kernel<float><<<blocks, 256, 0, st>>>(qw, (const float*)scales, (const float*)zeros, out, N, K, group, KB, status);
  else if (pdt == LL_BF16) int4_repack_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(qw, (const __nv_bfloat16*)scales, (const __nv_bfloat16*)zeros, out, N, K, group, KB, status);
  else {
    set_error("int4_prepare: unsupported param dtype %d", pdt);
    return LL_ERR_UNSUPPORTED;
  }
  LL_LAUNCH_OK("int4_repack_kernel");
  return LL_OK;
}

/* GPTQ 8-bit weights (`gptq.int8`): stream kind 2 of the batch-1 GEMV */
size_t ll_q8_stream_bytes(int N, int K, int group) {
  return i8_stream_ok(N, K, group) ? (size_t)(N / TILE_N) * (K / I8_KB) * I8_BLK8 : 0;
}

int ll_q8_gemv_supported(int N, int K, int group, int nmat, int M) {
  if (!(nmat == 1 || nmat == 2) || !q8_shape_ok(N, K, group, M)) return 0;
  Plan pl;
  return make_plan_i8(N, K, group, nmat, 0, true, &pl, SMEM_CORESIDENT, 8) == LL_OK ? 1 : 0;
}

int ll_q8_prepare(const uint8_t* qw, const void* scales, const void* zeros, int pdt, uint8_t* out, int N, int K, int group,
                  int32_t* status, ll_stream stream) {
  LL_CHECK_ARG(qw && scales && zeros && out && status, "null pointer");
  if (!i8_stream_ok(N, K, group)) {
    set_error("q8_prepare: needs group %% 128 == 0, K %% group == 0, N %% 16 == 0 (N=%d K=%d group=%d)", N, K, group);
    return LL_ERR_UNSUPPORTED;
  }
  LL_CHECK_ARG(((uintptr_t)out & 15) == 0, "stream buffer must be 16-byte aligned");
  const long long nblk = (long long)(N / TILE_N) * (K / I8_KB);
  const long long total = nblk * (I8_KB * 4 + TILE_N);
  const unsigned blocks = (unsigned)((total + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (pdt == LL_F32) q8_repack_kernel<float><<<blocks, 256, 0, st>>>(qw, (const float*)scales, (const float*)zeros, out, N, K, group, status);
  else if (pdt == LL_BF16) q8_repack_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>(qw, (const __nv_bfloat16*)scales, (const __nv_bfloat16*)zeros, out, N, K, group, status);
  else {
    set_error("q8_prepare: unsupported param dtype %d", pdt);
    return LL_ERR_UNSUPPORTED;
  }
  LL_LAUNCH_OK("q8_repack_kernel");
  return LL_OK;
}

/* debug (not in the public header): stage time stamps of the next `launches` GEMV launches */
__attribute__((visibility("default"))) void ll_debug_gemv_trace(unsigned long long* buf, int launches) {
  g_trace = buf; g_trace_launches = launches; g_trace_idx = 0;
}

/* debug (not in the public header): out[21] = launches per specialised kernel [epilogue 0..4][norm 0/1][batch-1 0/1], then
 * the generic kernels; counted since the last reset */
__attribute__((visibility("default"))) void ll_debug_gemv_counts(int32_t* out, int reset) {
  if (out != nullptr) {
    memcpy(out, g_spec_counts, sizeof(g_spec_counts));
    out[20] = g_generic_count;
  }
  if (reset) {
    memset(g_spec_counts, 0, sizeof(g_spec_counts));
    g_generic_count = 0;
  }
}

int ll_int4_stream_kind(int N, int K, int group, int M) {
  static int force = -1;  // LLB200_GEMV_KIND=0 keeps the bf16 kernel for batch 1 (tuning / A-B measurements)
  if (force < 0) {
    const char* e = getenv("LLB200_GEMV_KIND");
    force = e ? atoi(e) : 1;
  }
  return (force != 0 && i8_shape_ok(N, K, group, M)) ? LL_STREAM_U8 : LL_STREAM_BF16;
}

int ll_int4_gemv_plan(int N, int K, int group, int nmat, int M, int grid, size_t* smem_bytes, int32_t* ctas) {
  LL_CHECK_ARG(N > 0 && K > 0 && group > 0 && (nmat == 1 || nmat == 2) && M >= 1 && M <= MAXM, "bad plan arguments");
  Plan pl;
  if (ll_int4_stream_kind(N, K, group, M) == LL_STREAM_U8) {
    if (int r = make_plan_i8(N, K, group, nmat, grid, true, &pl)) return r;
  } else if (int r = make_plan(N, K, group, nmat, M, grid, true, &pl)) return r;
  if (smem_bytes) *smem_bytes = pl.smem;
  if (ctas) *ctas = pl.P;
  return LL_OK;
}

int ll_int4_gemv(const ll_gemv_args* a, ll_stream stream) {
  LL_CHECK_ARG(a != nullptr, "null args");
  LL_CHECK_ARG(a->M >= 1 && a->M <= MAXM, "M must be in 1..8 (got %d)", a->M);
  LL_CHECK_ARG(a->N > 0 && a->K > 0 && a->group > 0, "bad shape N=%d K=%d group=%d", a->N, a->K, a->group);
  LL_CHECK_ARG(a->epilogue >= 0 && a->epilogue <= 4, "bad epilogue %d", a->epilogue);
  LL_CHECK_ARG(a->w && a->x && (a->y || a->epilogue == LL_EPI_PUSH), "null pointer in gemv args");
  if (a->epilogue == LL_EPI_PUSH) {
    LL_CHECK_ARG(a->push_world >= 1 && a->push_world <= 8, "LL_EPI_PUSH: push_world must be 1..8");
    for (int r = 0; r < a->push_world; ++r) LL_CHECK_ARG(a->push_dst[r] && a->push_cnt[r], "LL_EPI_PUSH: null destination for rank %d", r);
  }
  const int nmat = a->epilogue == LL_EPI_SWIGLU ? 2 : 1;
  if (nmat == 2) LL_CHECK_ARG(a->w2 != nullptr, "SWIGLU epilogue needs w2");
  LL_CHECK_ARG(((uintptr_t)a->w & 15) == 0 && ((uintptr_t)a->w2 & 15) == 0 && ((uintptr_t)a->x & 15) == 0,
               "w/w2/x must be 16-byte aligned");
  if (a->norm_weight) {
    LL_CHECK_ARG(((uintptr_t)a->norm_weight & 15) == 0, "norm_weight must be 16-byte aligned");
  }
  if (a->epilogue == LL_EPI_ROPE_KV) {
    LL_CHECK_ARG(a->rope && a->pos && a->k_cache && a->v_cache, "ROPE_KV epilogue needs rope/pos/caches");
    LL_CHECK_ARG(a->N % 3 == 0 && (a->N / 3) % 2 == 0 && a->head_size % 2 == 0 && a->n_head * a->head_size == a->N / 3,
                 "ROPE_KV: N must be 3*n_head*head_size");
  }
  cudaStream_t st = (cudaStream_t)stream;
  Plan pl;
  const bool w8 = a->stream_kind == LL_STREAM_W8;
  const bool i8 = a->stream_kind == LL_STREAM_U8 || w8;
  LL_CHECK_ARG(a->stream_kind == LL_STREAM_BF16 || i8, "bad stream kind %d", a->stream_kind);
  if (i8) {
    LL_CHECK_ARG(a->M == 1 && a->epilogue != LL_EPI_PUSH, "stream kinds 1 and 2 serve batch 1 with the store / residual / RoPE / SwiGLU epilogues");
    const int par = (a->flags & LL_GEMV_PDL) ? (a->flags & (LL_GEMV_EVEN | LL_GEMV_ODD)) : 0;
    g_parity = par == LL_GEMV_EVEN || par == LL_GEMV_ODD ? par : 0;
    const int budget = g_parity == LL_GEMV_EVEN ? SMEM_EVEN : g_parity == LL_GEMV_ODD ? SMEM_ODD : SMEM_CORESIDENT;
    if (int r = make_plan_i8(a->N, a->K, a->group, nmat, a->grid, a->norm_weight != nullptr, &pl, budget, w8 ? 8 : 4)) return r;
    if (pl.smem > (size_t)budget) g_parity = 0;  // a row too long for the co-resident budget runs alone anyway
  } else if (int r = make_plan(a->N, a->K, a->group, nmat, a->M, a->grid, a->norm_weight != nullptr, &pl)) return r;

  Params p;
  memset(&p, 0, sizeof(p));
  p.w[0] = a->w; p.w[1] = a->w2;
  p.nmat = nmat; p.N = a->N; p.K = a->K; p.group = a->group; p.M = a->M;
  p.G = pl.G; p.GT = pl.GT; p.ntiles = pl.ntiles; p.I = pl.I;
  p.x = (const __nv_bfloat16*)a->x;
  p.norm_w = (const __nv_bfloat16*)a->norm_weight;
  p.eps = a->eps;
  p.epilogue = a->epilogue;
  p.y = (__nv_bfloat16*)a->y;
  p.bias = (const __nv_bfloat16*)a->bias;
  p.rope = a->rope; p.pos = a->pos;
  p.k_cache = (__nv_bfloat16*)a->k_cache; p.v_cache = (__nv_bfloat16*)a->v_cache;
  p.n_head = a->n_head; p.head_size = a->head_size; p.cache_len = a->cache_len;
  if (a->epilogue == LL_EPI_PUSH) {
    p.push_world = a->push_world;
    for (int r = 0; r < a->push_world; ++r) {
      p.push_dst[r] = (__nv_bfloat16*)(uintptr_t)a->push_dst[r];
      p.push_cnt[r] = (uint32_t*)(uintptr_t)a->push_cnt[r];
    }
  }
  
  p.S = pl.S; p.xs_stride = pl.xs_stride; p.ts_max = pl.ts_max;
  p.off_xs = pl.off_xs; p.off_sx = pl.off_sx; p.off_red = pl.off_red;
  p.off_wv = pl.off_wv; p.off_bar = pl.off_bar; p.off_nw = pl.off_nw;

  if (g_trace != nullptr && g_trace_idx < g_trace_launches && pl.P <= 256) p.trace = g_trace + (size_t)(g_trace_idx++) * 256 * 16;
  const bool pdl = (a->flags & LL_GEMV_PDL) != 0;
  static int delay_ns = -1;
  if (delay_ns < 0) {
    const char* e = getenv("LLB200_GEMV_DELAY_NS");
    delay_ns = e ? atoi(e) : 0;
  }
  p.prefetch_delay_ns = pdl ? delay_ns : 0;
  {
    // LLB200_GEMV_HELPERS=0: the parked warps of an ODD launch do not take part in the x conversion (A/B measurements);
    // read per launch: the chain is captured into a graph once
    const char* e = getenv("LLB200_GEMV_HELPERS");
    p.helpers = (i8 && g_parity == LL_GEMV_ODD && a->norm_weight == nullptr && (e ? atoi(e) != 0 : true)) ? 1 : 0;
  }
  if (a->x_planes != nullptr) {
    LL_CHECK_ARG(i8 && a->norm_weight == nullptr && a->x_rec != nullptr, "x_planes needs stream kind 1 / 2, no fused norm, and x_rec");
    LL_CHECK_ARG(((uintptr_t)a->x_planes & 15) == 0 && ((uintptr_t)a->x_rec & 15) == 0, "x_planes / x_rec must be 16-byte aligned");
    p.xdg_planes = a->x_planes;
    p.xdg_rec = a->x_rec;
    p.helpers = 0;
  }
  return i8 ? launch_plan_i8(p, pl, pdl, st) : launch_plan(p, pl, pdl, st);
}

}  // extern "C"
