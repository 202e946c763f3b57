// llm.int8 (LLM.int8(), arXiv:2208.07339) Linear for sm_100a: row-wise int8 weight quantisation and the
// outlier-decomposed int8 x int8 -> int32 matmul.
// ref: /root/reference/lit_llama/quantization.py:36-75 (Linear8bitLt over bitsandbytes, has_fp16_weights=False,
// threshold=6.0).  bitsandbytes itself is un-vendored and unpinned in the reference, so the arithmetic follows
// the published algorithm as restated in oracle/quant.py (int8_quantize_weight, int8_linear_forward):
//   A = x.half(); outlier columns = features with any |A| >= threshold;
//   SCA[m] = absmax over the entries of row m below the threshold; CA = rint(127 A / SCA), outlier columns 0;
//   C32 = CA . CB^T (int32, exact); out = fp16(C32 * SCA * SCB / 127^2 (+ bias));
//   out = fp16(out + fp16(A[:, idx] . fp16(CB[:, idx] * SCB / 127)^T)); cast to the input dtype.
// Decode-shaped (M <= 64 per pass): HBM-bound on the int8 weight read (N*K bytes per layer).  The integer
// part runs on the tensor cores (mma.sync m16n8k32 s8) with the weights as the A operand straight from
// global memory in 16-byte loads; K is split across CTAs and combined with int32 atomics (integer addition
// is associative: bit-deterministic).
#include <stdlib.h>

#include "common.cuh"

namespace llb {
namespace {

template <typename T> __device__ __forceinline__ float to_half_f32(T v);
template <> __device__ __forceinline__ float to_half_f32<float>(float v) { return __half2float(__float2half_rn(v)); }
template <> __device__ __forceinline__ float to_half_f32<__nv_bfloat16>(__nv_bfloat16 v) {
  return __half2float(__float2half_rn(__bfloat162float(v)));
}
template <> __device__ __forceinline__ float to_half_f32<__half>(__half v) { return __half2float(v); }

// ---------------------------------------------------------------- weight quantisation (load time)
template <typename T>
__global__ void int8_quant_weight_kernel(const T* __restrict__ w, int8_t* __restrict__ cb, float* __restrict__ scb, int K) {
  __shared__ float red[32];
  const T* row = w + (size_t)blockIdx.x * K;
  float mx = 0.f;
  for (int k = threadIdx.x; k < K; k += blockDim.x) mx = fmaxf(mx, fabsf(to_half_f32<T>(row[k])));
  mx = warp_max(mx);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = mx;
  __syncthreads();
  if (threadIdx.x < 32) {
    float v = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    v = warp_max(v);
    if (threadIdx.x == 0) red[0] = v;
  }
  __syncthreads();
  mx = red[0];
  if (threadIdx.x == 0) scb[blockIdx.x] = mx;
  const float inv = __fdiv_rn(127.f, mx == 0.f ? 1.f : mx);
  for (int k = threadIdx.x; k < K; k += blockDim.x) {
    const float q = rintf(__fmul_rn(to_half_f32<T>(row[k]), inv));
    cb[(size_t)blockIdx.x * K + k] = (int8_t)fminf(fmaxf(q, -127.f), 127.f);
  }
}

// ---------------------------------------------------------------- activation statistics
// a16 = x.half(); colflag[k] = any |a16[:,k]| >= thr; sca_bits[m] = max over entries below thr (float bits, >= 0)
template <typename T>
__global__ void int8_act_stats_kernel(const T* __restrict__ x, __half* __restrict__ a16, int32_t* __restrict__ colflag,
                                      int32_t* __restrict__ sca_bits, int M, int K, float thr) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int m0 = blockIdx.y * 8;
  bool flag = false;
  for (int m = m0; m < min(m0 + 8, M); ++m) {
    float v = 0.f;
    if (k < K) {
      const __half h = __float2half_rn(to_half_f32<T>(x[(size_t)m * K + k]));
      a16[(size_t)m * K + k] = h;
      v = fabsf(__half2float(h));
    }
    const bool big = v >= thr;
    flag |= big;
    float mx = big ? 0.f : v;
    mx = warp_max(mx);
    if ((threadIdx.x & 31) == 0 && mx > 0.f) atomicMax(&sca_bits[m], __float_as_int(mx));
  }
  if (flag && k < K) colflag[k] = 1;
}

// ascending list of outlier columns (single block; deterministic order)
__global__ void int8_compact_cols_kernel(const int32_t* __restrict__ colflag, int32_t* __restrict__ idx, int32_t* __restrict__ count,
                                         int K) {
  __shared__ int s_base;
  __shared__ int s_warp[32];
  if (threadIdx.x == 0) s_base = 0;
  __syncthreads();
  for (int k0 = 0; k0 < K; k0 += blockDim.x) {
    const int k = k0 + threadIdx.x;
    const int f = (k < K && colflag[k] != 0) ? 1 : 0;
    const unsigned bal = __ballot_sync(0xffffffffu, f);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (lane == 0) s_warp[warp] = __popc(bal);
    __syncthreads();
    int off = s_base;
    for (int w = 0; w < warp; ++w) off += s_warp[w];
    if (f) idx[off + __popc(bal & ((1u << lane) - 1u))] = k;
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_warp[w];
      s_base += t;
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) *count = s_base;
}

// CA[m][k] = colflag[k] ? 0 : clip(rint(a16 * 127 / SCA[m]), -127, 127), entries >= thr are 0 as well
__global__ void int8_act_quant_kernel(const __half* __restrict__ a16, const int32_t* __restrict__ colflag,
                                      const int32_t* __restrict__ sca_bits, int8_t* __restrict__ ca, int M, int K, float thr) {
  const int k = blockIdx.x * blockDim.x + threadIdx.x;
  const int m = blockIdx.y;
  if (k >= K) return;
  const float sca = __int_as_float(sca_bits[m]);
  const float inv = __fdiv_rn(127.f, sca == 0.f ? 1.f : sca);
  const float a = __half2float(a16[(size_t)m * K + k]);
  float q = 0.f;
  if (colflag[k] == 0 && fabsf(a) < thr) q = fminf(fmaxf(rintf(__fmul_rn(a, inv)), -127.f), 127.f);
  ca[(size_t)m * K + k] = (int8_t)q;
}

// ---------------------------------------------------------------- activation statistics + outlier list + quantisation, one kernel
// What int8_act_stats_kernel, int8_compact_cols_kernel and int8_act_quant_kernel do in three launches (+ the clearing of the
// int32 accumulator) as ONE: a CTA owns 64 feature columns for all rows of the pass.
//   phase 1: a16 = x.half(); per column "any |a16| >= thr"; per row the absmax of the entries below thr (atomicMax);
//            this CTA's number of outlier columns -> cta_count[]; its share of c32 cleared;
//   grid barrier (arrival counter in the workspace; the grid is K / 64 CTAs, all resident);
//   phase 2: CA from the final row maxima; the CTA's outlier columns appended in ascending order at the prefix sum of the
//            counts of the CTAs before it (so the list is ascending overall, as the single-block compaction produced it).
constexpr int PREP_T = 256, PREP_C = 64, PREP_RG = PREP_T / PREP_C;  // a CTA: 64 columns x 4 interleaved row groups
template <typename T>
__global__ void __launch_bounds__(PREP_T) int8_act_prep_kernel(const T* __restrict__ x, __half* __restrict__ a16, int8_t* __restrict__ ca,
                                                              int32_t* __restrict__ sca_bits, int32_t* __restrict__ idx,
                                                              int32_t* __restrict__ count, int32_t* __restrict__ cta_count,
                                                              int32_t* __restrict__ arrive, int32_t* __restrict__ c32, size_t c32_words,
                                                              int M, int K, float thr) {
  __shared__ int s_row[64];      // per-row maxima of this CTA (float bits, >= 0), later the final row scales
  __shared__ int s_flag[PREP_C];
  __shared__ int s_cnt[2];
  __shared__ int s_red[PREP_T / 32];
  __shared__ int s_base;
  const int tx = threadIdx.x & (PREP_C - 1), ty = threadIdx.x / PREP_C;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const int k = blockIdx.x * PREP_C + tx;
  const bool k_ok = k < K;
  if (threadIdx.x < 64) s_row[threadIdx.x] = 0;
  if (threadIdx.x < PREP_C) s_flag[threadIdx.x] = 0;
  // this CTA's share of the accumulator the matmul adds into
  for (size_t i = (size_t)blockIdx.x * PREP_T + threadIdx.x; i < c32_words; i += (size_t)gridDim.x * PREP_T) c32[i] = 0;
  __syncthreads();
  bool flag = false;
  for (int m = ty; m < M; m += PREP_RG) {
    float v = 0.f;
    if (k_ok) {
      const __half h = __float2half_rn(to_half_f32<T>(x[(size_t)m * K + k]));
      a16[(size_t)m * K + k] = h;
      v = fabsf(__half2float(h));
    }
    const bool big = v >= thr;
    flag |= big;
    float mx = big ? 0.f : v;
    mx = warp_max(mx);
    if (lane == 0 && mx > 0.f) atomicMax(&s_row[m], __float_as_int(mx));
  }
  if (flag && k_ok) s_flag[tx] = 1;   // (several row groups may store the same 1)
  __syncthreads();
  if (threadIdx.x < M && s_row[threadIdx.x] > 0) atomicMax(&sca_bits[threadIdx.x], s_row[threadIdx.x]);
  const bool col_out = s_flag[tx] != 0;
  unsigned bal = 0;
  if (ty == 0) {   // the first 64 threads count the CTA's outlier columns
    bal = __ballot_sync(0xffffffffu, col_out);
    if (lane == 0) s_cnt[warp] = __popc(bal);
  }
  __syncthreads();
  const int mine = s_cnt[0] + s_cnt[1];
  if (threadIdx.x == 0) {
    cta_count[blockIdx.x] = mine;
    __threadfence();
    atomicAdd(arrive, 1);
    while (*reinterpret_cast<volatile int32_t*>(arrive) < (int32_t)gridDim.x) {
    }
    __threadfence();
  }
  __syncthreads();
  // ---- every CTA's statistics are in: prefix of the counts before this CTA, final row scales
  int part = 0;
  for (int c = threadIdx.x; c < (int)blockIdx.x; c += PREP_T) part += __ldcg(cta_count + c);
  part = __reduce_add_sync(0xffffffffu, part);
  if (lane == 0) s_red[warp] = part;
  if (threadIdx.x < M) s_row[threadIdx.x] = __ldcg(sca_bits + threadIdx.x);
  __syncthreads();
  if (threadIdx.x == 0) {
    int base = 0;
    for (int w = 0; w < PREP_T / 32; ++w) base += s_red[w];
    s_base = base;
    if (blockIdx.x == gridDim.x - 1) *count = base + mine;
  }
  __syncthreads();
  if (ty == 0 && col_out && k_ok) idx[s_base + (warp == 1 ? s_cnt[0] : 0) + __popc(bal & ((1u << lane) - 1u))] = k;
  if (k_ok) {
    for (int m = ty; m < M; m += PREP_RG) {
      const float sca = __int_as_float(s_row[m]);
      const float inv = __fdiv_rn(127.f, sca == 0.f ? 1.f : sca);
      const float a = __half2float(a16[(size_t)m * K + k]);
      float q = 0.f;
      if (!col_out && fabsf(a) < thr) q = fminf(fmaxf(rintf(__fmul_rn(a, inv)), -127.f), 127.f);
      ca[(size_t)m * K + k] = (int8_t)q;
    }
  }
}

// ---------------------------------------------------------------- int8 tensor-core matmul, split-K
__device__ __forceinline__ void mma_s8_16832(int (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0,
                                             uint32_t b1) {
  asm volatile(
      "mma.sync.aligned.m16n8k32.row.col.s32.s8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
      : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3])
      : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// grid (N/64, ksplit); block = 4 warps = four 16-column tiles walking the SAME 128-k blocks of the CTA's k range: the
// quantised activations of a k-block ([rows][128] bytes) are staged once per CTA in shared memory (double-buffered, one
// barrier per k-block) instead of being fetched from L2 by every 16-column tile (that was 2 x the weight bytes in L2
// traffic); the next k-block's weights are requested before this one's MMAs.
// c32[m][n] += sum_k CA[m][k] * CB[n][k]   (M <= 8 * MG rows; weights are the MMA A operand)
template <int MG>
__global__ void __launch_bounds__(128) int8_mm_kernel(const int8_t* __restrict__ cb, const int8_t* __restrict__ ca,
                                                      int32_t* __restrict__ c32, int M, int N, int K, int kblocks_per_split) {
  constexpr int ROWS = MG * 8, STR = 144;  // 128 bytes of a row + 16: neighbouring rows start four banks apart
  __shared__ __align__(16) uint8_t a_s[2][ROWS][STR];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, g = lane >> 2, tig = lane & 3;
  const int n0 = (blockIdx.x * 4 + warp) * 16;
  const int nkb = K / 128;
  const int kb0 = blockIdx.y * kblocks_per_split, kb1 = min(kb0 + kblocks_per_split, nkb);
  int acc[MG][4];
#pragma unroll
  for (int i = 0; i < MG; ++i) acc[i][0] = acc[i][1] = acc[i][2] = acc[i][3] = 0;
  const int8_t* w_lo = cb + (size_t)(n0 + g) * K + 16 * tig;
  const int8_t* w_hi = cb + (size_t)(n0 + g + 8) * K + 16 * tig;
  const bool n_lo_ok = n0 + g < N, n_hi_ok = n0 + g + 8 < N;
  const uint4 z4 = make_uint4(0u, 0u, 0u, 0u);
  auto stage = [&](int kb, int buf) {
#pragma unroll
    for (int e = threadIdx.x; e < ROWS * 8; e += 128) {
      const int row = e >> 3, seg = e & 7;
      uint4 v = z4;
      if (row < M) v = __ldg(reinterpret_cast<const uint4*>(ca + (size_t)row * K + (size_t)kb * 128 + seg * 16));
      *reinterpret_cast<uint4*>(&a_s[buf][row][seg * 16]) = v;
    }
  };
  auto load_w = [&](int kb, uint4 (&w)[4]) {
    const int k = kb * 128;
    // lane (g, tig) holds k + 16 tig + [0,16) (slots 0..15 of four MMAs) and k + 64 + 16 tig + [0,16)
    w[0] = n_lo_ok ? __ldcs(reinterpret_cast<const uint4*>(w_lo + k)) : z4;
    w[1] = n_hi_ok ? __ldcs(reinterpret_cast<const uint4*>(w_hi + k)) : z4;
    w[2] = n_lo_ok ? __ldcs(reinterpret_cast<const uint4*>(w_lo + k + 64)) : z4;
    w[3] = n_hi_ok ? __ldcs(reinterpret_cast<const uint4*>(w_hi + k + 64)) : z4;
  };
  uint4 wcur[4], wnext[4];
  if (kb0 < kb1) {
    load_w(kb0, wcur);
    stage(kb0, 0);
  }
  __syncthreads();
  int buf = 0;
  for (int kb = kb0; kb < kb1; ++kb) {
    if (kb + 1 < kb1) {
      load_w(kb + 1, wnext);
      stage(kb + 1, buf ^ 1);
    }
    const uint32_t al0[4] = {wcur[0].x, wcur[0].y, wcur[0].z, wcur[0].w}, ah0[4] = {wcur[1].x, wcur[1].y, wcur[1].z, wcur[1].w};
    const uint32_t al1[4] = {wcur[2].x, wcur[2].y, wcur[2].z, wcur[2].w}, ah1[4] = {wcur[3].x, wcur[3].y, wcur[3].z, wcur[3].w};
#pragma unroll
    for (int mg = 0; mg < MG; ++mg) {
      const int m = mg * 8 + g;
      const uint4 b0 = *reinterpret_cast<const uint4*>(&a_s[buf][m][16 * tig]);
      const uint4 b1 = *reinterpret_cast<const uint4*>(&a_s[buf][m][64 + 16 * tig]);
      const uint32_t bb0[4] = {b0.x, b0.y, b0.z, b0.w}, bb1[4] = {b1.x, b1.y, b1.z, b1.w};
#pragma unroll
      for (int j = 0; j < 4; ++j) mma_s8_16832(acc[mg], al0[j], ah0[j], al1[j], ah1[j], bb0[j], bb1[j]);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) wcur[j] = wnext[j];
    __syncthreads();
    buf ^= 1;
  }
  // D: c0,c1 -> (n = n0+g, m = 2tig, 2tig+1); c2,c3 -> (n = n0+g+8, ...)
#pragma unroll
  for (int mg = 0; mg < MG; ++mg) {
    const int m0 = mg * 8 + 2 * tig;
    if (m0 < M) {
      if (n_lo_ok && acc[mg][0] != 0) atomicAdd(&c32[(size_t)m0 * N + n0 + g], acc[mg][0]);
      if (n_hi_ok && acc[mg][2] != 0) atomicAdd(&c32[(size_t)m0 * N + n0 + g + 8], acc[mg][2]);
    }
    if (m0 + 1 < M) {
      if (n_lo_ok && acc[mg][1] != 0) atomicAdd(&c32[(size_t)(m0 + 1) * N + n0 + g], acc[mg][1]);
      if (n_hi_ok && acc[mg][3] != 0) atomicAdd(&c32[(size_t)(m0 + 1) * N + n0 + g + 8], acc[mg][3]);
    }
  }
}

// ---------------------------------------------------------------- dequantise + outlier part + store
// The outlier part  sum_i a16[m][k_i] * fp16(CB[n][k_i] * SCB[n] / 127)  in fp32 with the fp16 roundings of the oracle.
constexpr int FIN_ROWS = 16, FIN_COLS = 64, FIN_SL = 4;
// 256 threads = 64 output columns x FIN_SL slices of the outlier list: slice s takes the chunks of 64 outlier columns number
// s, s + 4, ... (a thread per column walking the whole list alone left two warps per SM with hundreds of dependent steps:
// 78 us per Linear in a batch-32 step whose activations have hundreds of outlier columns, 43 % of the step); the slices'
// partial sums are added in slice order through shared memory (deterministic), then slice 0 finishes the row group.
template <typename T>
__global__ void __launch_bounds__(FIN_COLS * FIN_SL) int8_finish_kernel(const int32_t* __restrict__ c32, const int32_t* __restrict__ sca_bits,
                                                          const float* __restrict__ scb, const T* __restrict__ bias,
                                                          const __half* __restrict__ a16, const int8_t* __restrict__ cb,
                                                          const int8_t* __restrict__ cbt,
                                                          const int32_t* __restrict__ idx, const int32_t* __restrict__ count,
                                                          T* __restrict__ y, int M, int N, int K) {
  __shared__ __align__(16) float a_s[FIN_SL][FIN_COLS][FIN_ROWS];
  __shared__ int k_s[FIN_SL][FIN_COLS];
  __shared__ float part[FIN_SL - 1][FIN_ROWS][FIN_COLS];
  const int col = threadIdx.x & (FIN_COLS - 1), sl = threadIdx.x / FIN_COLS;
  const int n = blockIdx.x * FIN_COLS + col;
  const bool n_ok = n < N;
  const float sb = n_ok ? scb[n] : 0.f;
  const float bv = (n_ok && bias != nullptr) ? to_half_f32<T>(bias[n]) : 0.f;
  // fp16(CB * SCB / 127) as fp16(CB * (SCB / 127)): one multiplication per weight instead of a multiplication and an IEEE
  // division (25 instructions); the two fp32 values differ by at most one ulp, i.e. the fp16 result differs only where the
  // product sits within 2^-13 of a rounding boundary - far inside the stated tolerance (the int8 oracle is unpinned anyway)
  const float sb127 = __fdiv_rn(sb, 127.f);
  const int cnt = *count;
  const int8_t* wrow = cb + (size_t)(n_ok ? n : 0) * K;
  const int m0 = blockIdx.y * FIN_ROWS;   // one group of FIN_ROWS rows per CTA
  float acc[FIN_ROWS];
#pragma unroll
  for (int r = 0; r < FIN_ROWS; ++r) acc[r] = 0.f;
  for (int base = 0; base < cnt; base += FIN_COLS * FIN_SL) {   // uniform over the CTA
    const int i0 = base + sl * FIN_COLS;
    const int ni = max(0, min(FIN_COLS, cnt - i0));
    __syncthreads();
    if (col < ni) k_s[sl][col] = idx[i0 + col];
    __syncthreads();
    for (int e = col; e < FIN_COLS * FIN_ROWS; e += FIN_COLS) {   // the slice's 64 threads stage its chunk of activations
      const int i = e / FIN_ROWS, r = e - i * FIN_ROWS;
      a_s[sl][i][r] = (i < ni && m0 + r < M) ? __half2float(a16[(size_t)(m0 + r) * K + k_s[sl][i]]) : 0.f;
    }
    __syncthreads();
    // eight weights in flight per thread; entries past ni multiply activations that were staged as zero
    for (int i = 0; i < ni; i += 8) {
      int8_t q[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        // cbt = CB transposed ([K][N]): an outlier column is a contiguous row of it, read coalesced over the 64 columns;
        // without it every thread gathers single bytes out of its own weight row (one 32-byte sector each)
        if (cbt != nullptr) q[j] = (i + j < ni && n_ok) ? cbt[(size_t)k_s[sl][i + j] * N + n] : (int8_t)0;
        else q[j] = (i + j < ni) ? wrow[k_s[sl][i + j]] : (int8_t)0;
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        const float wv = __half2float(__float2half_rn(__fmul_rn((float)q[j], sb127)));
        if (i + j < ni) {
#pragma unroll
          for (int r4 = 0; r4 < FIN_ROWS / 4; ++r4) {   // the outlier's 16 activations in four 16-byte reads (a broadcast)
            const float4 av = *reinterpret_cast<const float4*>(&a_s[sl][i + j][4 * r4]);
            acc[4 * r4] = fmaf(av.x, wv, acc[4 * r4]);
            acc[4 * r4 + 1] = fmaf(av.y, wv, acc[4 * r4 + 1]);
            acc[4 * r4 + 2] = fmaf(av.z, wv, acc[4 * r4 + 2]);
            acc[4 * r4 + 3] = fmaf(av.w, wv, acc[4 * r4 + 3]);
          }
        }
      }
    }
  }
  __syncthreads();
  if (sl > 0) {
#pragma unroll
    for (int r = 0; r < FIN_ROWS; ++r) part[sl - 1][r][col] = acc[r];
  }
  __syncthreads();
  if (sl == 0 && n_ok) {
#pragma unroll
    for (int r = 0; r < FIN_ROWS; ++r) {
      const int m = m0 + r;
      if (m < M) {
        float o = acc[r];
#pragma unroll
        for (int t = 0; t < FIN_SL - 1; ++t) o += part[t][r][col];   // slice order
        const float sca = __int_as_float(sca_bits[m]);
        float out = __fmul_rn(__fmul_rn(__fmul_rn((float)c32[(size_t)m * N + n], sca), sb), 1.0f / (127.0f * 127.0f));
        if (bias != nullptr) out = __fadd_rn(out, bv);
        out = __half2float(__float2half_rn(out));
        if (cnt > 0) out = __half2float(__float2half_rn(__fadd_rn(out, __half2float(__float2half_rn(o)))));
        dt<T>::st(y, (size_t)m * N + n, out);
      }
    }
  }
}

struct Ws {
  __half* a16; int8_t* ca; int32_t* c32; int32_t* colflag; int32_t* idx; int32_t* sca_bits; int32_t* count;
  int32_t* cta_count; int32_t* arrive;
  size_t zero_off, zero_bytes, total;
};
Ws carve(void* base, int M, int N, int K) {
  Ws w;
  size_t o = 0;
  auto take = [&](size_t b) { size_t r = o; o += (b + 255) & ~(size_t)255; return r; };
  uint8_t* p = (uint8_t*)base;
  w.a16 = (__half*)(p + take((size_t)M * K * 2));
  w.ca = (int8_t*)(p + take((size_t)M * K));
  w.idx = (int32_t*)(p + take((size_t)K * 4));
  w.c32 = (int32_t*)(p + take((size_t)M * N * 4));       // cleared by the prep kernel
  w.colflag = (int32_t*)(p + take((size_t)K * 4));       // (unused since the fused prep kernel; kept for the layout)
  w.cta_count = (int32_t*)(p + take((size_t)((K + PREP_C - 1) / PREP_C) * 4));
  w.zero_off = o;  // the few words below are cleared at the start of every call (one small memset)
  w.sca_bits = (int32_t*)(p + take((size_t)M * 4));
  w.count = (int32_t*)(p + take(4));
  w.arrive = (int32_t*)(p + take(4));
  w.zero_bytes = o - w.zero_off;
  w.total = o;
  return w;
}

}  // namespace
}  // namespace llb

using namespace llb;

template <typename T>
static int int8_linear_t(const T* x, const int8_t* cb, const int8_t* cbt, const float* scb, const T* bias, T* y, int M, int N, int K,
                         float thr, void* ws_base, cudaStream_t st) {
  const int sms = sm_count();
  for (int m0 = 0; m0 < M; m0 += 64) {  // decode-shaped passes of up to 64 rows (weights re-read per pass)
    const int Mp = min(64, M - m0);
    Ws w = carve(ws_base, Mp, N, K);
    LL_CUDA(cudaMemsetAsync((uint8_t*)ws_base + w.zero_off, 0, w.zero_bytes, st));
    const T* xp = x + (size_t)m0 * K;
    const int prep_ctas = (K + PREP_C - 1) / PREP_C;
    if (prep_ctas <= 4 * sms) {   // the fused kernel's grid barrier needs every CTA resident (256 threads, 1.5 KB: >= 4 per SM)
      int8_act_prep_kernel<T><<<prep_ctas, PREP_T, 0, st>>>(xp, w.a16, w.ca, w.sca_bits, w.idx, w.count, w.cta_count, w.arrive, w.c32,
                                                          (size_t)Mp * N, Mp, K, thr);
    } else {
      LL_CUDA(cudaMemsetAsync(w.c32, 0, (size_t)((uint8_t*)w.cta_count - (uint8_t*)w.c32), st));   // c32 and colflag (adjacent)
      int8_act_stats_kernel<T><<<dim3((K + 255) / 256, (Mp + 7) / 8), 256, 0, st>>>(xp, w.a16, w.colflag, w.sca_bits, Mp, K, thr);
      int8_compact_cols_kernel<<<1, 1024, 0, st>>>(w.colflag, w.idx, w.count, K);
      int8_act_quant_kernel<<<dim3((K + 255) / 256, Mp), 256, 0, st>>>(w.a16, w.colflag, w.sca_bits, w.ca, Mp, K, thr);
    }
    const int ntiles = (N + 63) / 64, nkb = K / 128;   // CTAs along N: four 16-column tiles each
    static int mult = 0;   // CTAs aimed at, in units of the SM count (LLB200_INT8_KSPLIT_MULT, tuning)
    if (mult == 0) {
      const char* e = getenv("LLB200_INT8_KSPLIT_MULT");
      mult = e ? atoi(e) : 4;
      if (mult < 1 || mult > 16) mult = 4;
    }
    int ksplit = (mult * sms + ntiles - 1) / ntiles;
    if (ksplit > nkb) ksplit = nkb;
    if (ksplit < 1) ksplit = 1;
    const int per = (nkb + ksplit - 1) / ksplit;
    ksplit = (nkb + per - 1) / per;
    dim3 grid(ntiles, ksplit);
    const int mg = (Mp + 7) / 8;
#define MM(G) int8_mm_kernel<G><<<grid, 128, 0, st>>>(cb, w.ca, w.c32, Mp, N, K, per)
    if (mg <= 1) MM(1); else if (mg <= 2) MM(2); else if (mg <= 4) MM(4); else MM(8);
#undef MM
    int8_finish_kernel<T><<<dim3((N + FIN_COLS - 1) / FIN_COLS, (Mp + FIN_ROWS - 1) / FIN_ROWS), FIN_COLS * FIN_SL, 0, st>>>(w.c32, w.sca_bits, scb, bias, w.a16, cb, cbt, w.idx, w.count,
                                                                    y + (size_t)m0 * N, Mp, N, K);
    LL_LAUNCH_OK("int8 linear kernels");
  }
  return LL_OK;
}

extern "C" {

size_t ll_int8_workspace_bytes(int M, int N, int K) { return carve(nullptr, M, N, K).total; }

int ll_int8_quantize_weight(const void* w, int w_dtype, int8_t* cb, float* scb, int N, int K, ll_stream stream) {
  LL_CHECK_ARG(w && cb && scb && N > 0 && K > 0, "bad int8_quantize_weight arguments");
  cudaStream_t st = (cudaStream_t)stream;
  if (w_dtype == LL_F32) int8_quant_weight_kernel<float><<<N, 256, 0, st>>>((const float*)w, cb, scb, K);
  else if (w_dtype == LL_BF16) int8_quant_weight_kernel<__nv_bfloat16><<<N, 256, 0, st>>>((const __nv_bfloat16*)w, cb, scb, K);
  else if (w_dtype == LL_F16) int8_quant_weight_kernel<__half><<<N, 256, 0, st>>>((const __half*)w, cb, scb, K);
  else {
    set_error("int8_quantize_weight: unsupported dtype %d", w_dtype);
    return LL_ERR_UNSUPPORTED;
  }
  LL_LAUNCH_OK("int8_quant_weight_kernel");
  return LL_OK;
}

int ll_int8_linear(const void* x, int act_dtype, const int8_t* cb, const int8_t* cbt, const float* scb, const void* bias, void* y,
                   int M, int N, int K, float threshold, void* workspace, size_t workspace_bytes, ll_stream stream) {
  LL_CHECK_ARG(x && cb && scb && y && workspace, "null pointer");
  LL_CHECK_ARG(M > 0 && N > 0 && K > 0 && K % 128 == 0, "int8 linear needs K %% 128 == 0 (M=%d N=%d K=%d)", M, N, K);
  LL_CHECK_ARG(((uintptr_t)cb & 15) == 0 && ((uintptr_t)workspace & 255) == 0, "cb must be 16-byte, workspace 256-byte aligned");
  LL_CHECK_ARG(workspace_bytes >= ll_int8_workspace_bytes(M < 64 ? M : 64, N, K), "workspace too small");
  cudaStream_t st = (cudaStream_t)stream;
  if (act_dtype == LL_BF16)
    return int8_linear_t<__nv_bfloat16>((const __nv_bfloat16*)x, cb, cbt, scb, (const __nv_bfloat16*)bias, (__nv_bfloat16*)y, M, N, K, threshold, workspace, st);
  if (act_dtype == LL_F16)
    return int8_linear_t<__half>((const __half*)x, cb, cbt, scb, (const __half*)bias, (__half*)y, M, N, K, threshold, workspace, st);
  if (act_dtype == LL_F32)
    return int8_linear_t<float>((const float*)x, cb, cbt, scb, (const float*)bias, (float*)y, M, N, K, threshold, workspace, st);
  set_error("int8_linear: unsupported activation dtype %d", act_dtype);
  return LL_ERR_UNSUPPORTED;
}

}  // extern "C"
