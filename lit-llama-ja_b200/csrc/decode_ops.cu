// model.py kernels around the quantized Linears: RMSNorm, embedding (+ sum-of-squares partials),
// RoPE + in-place KV-cache write, split-KV decode attention, SwiGLU product, residual add.
// ref: /root/reference/lit_llama/model.py:110, 162-175, 192-243, 257-260, 276-283, 312-329.
// Every kernel mirrors torch's rounding points (each elementwise op rounds to the activation dtype).
#include <stdlib.h>

#include "common.cuh"

namespace llb {
namespace {

// ---------------------------------------------------------------- RMSNorm (row N)
template <typename T>
__global__ void rmsnorm_kernel(const T* __restrict__ x, const T* __restrict__ w, T* __restrict__ y, int dim, float eps) {
  __shared__ float red[32];
  __shared__ float s_r;
  const T* xr = x + (size_t)blockIdx.x * dim;
  T* yr = y + (size_t)blockIdx.x * dim;
  float ss = 0.f;
  for (int i = threadIdx.x; i < dim; i += blockDim.x) {
    const float v = dt<T>::ld(xr, i);
    ss += dt<T>::rd(__fmul_rn(v, v));
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x == 0) {
    float t = 0.f;
    for (int i = 0; i < (int)(blockDim.x >> 5); ++i) t += red[i];
    const float mean = dt<T>::rd(t * (1.0f / (float)dim));
    s_r = dt<T>::rd(rsqrtf(dt<T>::rd(__fadd_rn(mean, eps))));
  }
  __syncthreads();
  const float r = s_r;
  for (int i = threadIdx.x; i < dim; i += blockDim.x) {
    const float xn = dt<T>::rd(__fmul_rn(dt<T>::ld(xr, i), r));
    dt<T>::st(yr, i, __fmul_rn(dt<T>::ld(w, i), xn));
  }
}

// bf16 rows whose length is a multiple of 8: one 16-byte piece per thread per trip, kept in registers between the two passes
// (the scalar kernel above re-reads the row and issues one 2-byte load per element: 16 us for 32 rows of 4096 with one
// 256-thread CTA per row).  Same rounding steps; the order of the fp32 sum differs (as it does between any two reductions).
constexpr int RN_PIECES = 4;  // pieces a thread keeps: rows up to threads * 8 * RN_PIECES elements
__global__ void __launch_bounds__(1024) rmsnorm_bf16_vec_kernel(const __nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ w,
                                                                __nv_bfloat16* __restrict__ y, int dim, float eps) {
  __shared__ float red[32];
  __shared__ float s_r;
  const uint4* xr = reinterpret_cast<const uint4*>(x + (size_t)blockIdx.x * dim);
  uint4* yr = reinterpret_cast<uint4*>(y + (size_t)blockIdx.x * dim);
  const uint4* wr = reinterpret_cast<const uint4*>(w);
  const int np = dim >> 3;
  uint4 px[RN_PIECES], pw[RN_PIECES];
  float ss = 0.f;
#pragma unroll
  for (int j = 0; j < RN_PIECES; ++j) {
    const int i = threadIdx.x + j * blockDim.x;
    px[j] = make_uint4(0u, 0u, 0u, 0u);
    pw[j] = px[j];
    if (i < np) {
      px[j] = xr[i];
      pw[j] = __ldg(wr + i);
    }
  }
#pragma unroll
  for (int j = 0; j < RN_PIECES; ++j) {
    const uint32_t q[4] = {px[j].x, px[j].y, px[j].z, px[j].w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float a = bf16_bits_to_f32(q[e] & 0xffffu), b = bf16_bits_to_f32(q[e] >> 16);
      ss += bf16r(__fmul_rn(a, a));
      ss += bf16r(__fmul_rn(b, b));
    }
  }
  ss = warp_sum(ss);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = ss;
  __syncthreads();
  if (threadIdx.x < 32) {
    float t = threadIdx.x < (blockDim.x >> 5) ? red[threadIdx.x] : 0.f;
    t = warp_sum(t);
    if (threadIdx.x == 0) {
      const float mean = bf16r(t * (1.0f / (float)dim));
      s_r = bf16r(rsqrtf(bf16r(__fadd_rn(mean, eps))));
    }
  }
  __syncthreads();
  const float r = s_r;
#pragma unroll
  for (int j = 0; j < RN_PIECES; ++j) {
    const int i = threadIdx.x + j * blockDim.x;
    if (i < np) {
      const uint32_t q[4] = {px[j].x, px[j].y, px[j].z, px[j].w}, ww[4] = {pw[j].x, pw[j].y, pw[j].z, pw[j].w};
      uint32_t o[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const float a = bf16r(__fmul_rn(bf16_bits_to_f32(q[e] & 0xffffu), r)), b = bf16r(__fmul_rn(bf16_bits_to_f32(q[e] >> 16), r));
        o[e] = (uint32_t)f32_to_bf16_bits(__fmul_rn(bf16_bits_to_f32(ww[e] & 0xffffu), a)) |
               ((uint32_t)f32_to_bf16_bits(__fmul_rn(bf16_bits_to_f32(ww[e] >> 16), b)) << 16);
      }
      yr[i] = make_uint4(o[0], o[1], o[2], o[3]);
    }
  }
}

// ---------------------------------------------------------------- embedding + ssq partials
__global__ void embed_ssq_kernel(const __nv_bfloat16* __restrict__ wte, const int64_t* __restrict__ idx,
                                 __nv_bfloat16* __restrict__ x, float* __restrict__ ssq, int dim, int nparts) {
  // one sum-of-squares partial per 16 columns (the int4 GEMV's tile width)
  const int m = blockIdx.y;
  const int n = blockIdx.x * 128 + threadIdx.x;
  float v = 0.f;
  if (n < dim) {
    const __nv_bfloat16 e = wte[(size_t)idx[m] * dim + n];
    x[(size_t)m * dim + n] = e;
    v = __bfloat162float(e);
    v = bf16r(__fmul_rn(v, v));
  }
  v += __shfl_xor_sync(0xffffffffu, v, 8);
  v += __shfl_xor_sync(0xffffffffu, v, 4);
  v += __shfl_xor_sync(0xffffffffu, v, 2);
  v += __shfl_xor_sync(0xffffffffu, v, 1);
  const int part = n >> 4;
  if (ssq != nullptr && (threadIdx.x & 15) == 0 && part < nparts) ssq[m * nparts + part] = v;
}

// ---------------------------------------------------------------- RoPE + KV append (rows R, A)
// grid (T, B*n_head), block head_size/2 threads
template <typename T>
__global__ void rope_kv_kernel(const T* __restrict__ qkv, const float* __restrict__ rope,
                               const int32_t* __restrict__ pos_ids, const int32_t* __restrict__ slots,
                               T* __restrict__ q_out, T* __restrict__ kc, T* __restrict__ vc, int Tn, int n_head,
                               int hs, int S) {
  const int t = blockIdx.x, bh = blockIdx.y, b = bh / n_head, head = bh % n_head;
  const int E = n_head * hs;
  const int i = threadIdx.x;  // pair index
  if (i >= hs / 2) return;
  const T* row = qkv + ((size_t)b * Tn + t) * 3 * E;
  const float2 cs = *reinterpret_cast<const float2*>(rope + ((size_t)pos_ids[t] * (hs / 2) + i) * 2);
  const int slot = slots[t];
  const int d = 2 * i;
  {
    const float x0 = dt<T>::ld(row, head * hs + d), x1 = dt<T>::ld(row, head * hs + d + 1);
    T* o = q_out + (((size_t)b * n_head + head) * Tn + t) * hs;
    dt<T>::st(o, d, __fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y)));
    dt<T>::st(o, d + 1, __fadd_rn(__fmul_rn(x1, cs.x), __fmul_rn(x0, cs.y)));
  }
  {
    const float x0 = dt<T>::ld(row, E + head * hs + d), x1 = dt<T>::ld(row, E + head * hs + d + 1);
    T* o = kc + (((size_t)b * n_head + head) * S + slot) * hs;
    dt<T>::st(o, d, __fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y)));
    dt<T>::st(o, d + 1, __fadd_rn(__fmul_rn(x1, cs.x), __fmul_rn(x0, cs.y)));
  }
  {
    T* o = vc + (((size_t)b * n_head + head) * S + slot) * hs;
    o[d] = row[2 * E + head * hs + d];
    o[d + 1] = row[2 * E + head * hs + d + 1];
  }
}

// ---------------------------------------------------------------- decode attention (row A, T=1)
// grid (B*n_head, nsplit); 4 warps; a position is handled by 8 lanes (HS/8 elements each).
template <typename T> struct vec_ld;
template <> struct vec_ld<__nv_bfloat16> {
  template <int EPL> static __device__ __forceinline__ void ld(const __nv_bfloat16* p, float (&o)[EPL]) {
    if constexpr (EPL % 8 == 0) {
#pragma unroll
      for (int i = 0; i < EPL / 8; ++i) {
        const uint4 r = *reinterpret_cast<const uint4*>(p + 8 * i);
        const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          o[8 * i + 2 * j] = bf16_bits_to_f32(w[j] & 0xffff);
          o[8 * i + 2 * j + 1] = bf16_bits_to_f32(w[j] >> 16);
        }
      }
    } else {
#pragma unroll
      for (int i = 0; i < EPL / 4; ++i) {
        const uint2 r = *reinterpret_cast<const uint2*>(p + 4 * i);
        o[4 * i] = bf16_bits_to_f32(r.x & 0xffff); o[4 * i + 1] = bf16_bits_to_f32(r.x >> 16);
        o[4 * i + 2] = bf16_bits_to_f32(r.y & 0xffff); o[4 * i + 3] = bf16_bits_to_f32(r.y >> 16);
      }
    }
  }
};
template <> struct vec_ld<float> {
  template <int EPL> static __device__ __forceinline__ void ld(const float* p, float (&o)[EPL]) {
#pragma unroll
    for (int i = 0; i < EPL / 4; ++i) {
      const float4 r = *reinterpret_cast<const float4*>(p + 4 * i);
      o[4 * i] = r.x; o[4 * i + 1] = r.y; o[4 * i + 2] = r.z; o[4 * i + 3] = r.w;
    }
  }
};

// Registers: 151 at the natural allocation.  Capping the kernel at 128 (so that two CTAs fit an SM next to a GEMV
// CTA) costs 10 % of its time and does not remove the penalty of overlapping it with the neighbouring GEMVs
// (profiles/r01_gemv_experiments.md).
// CL: the splits of one (batch, head) form a thread-block cluster (nsplit <= 8) and hand their partial states to the
// cluster's first CTA through distributed shared memory - no global scratch, fence, atomic or reload on the way.
template <typename T, int HS, bool CL>
__global__ void __launch_bounds__(128) decode_attn_kernel(const T* __restrict__ q, const T* __restrict__ kc,
                                                          const T* __restrict__ vc, const int32_t* __restrict__ pos,
                                                          T* __restrict__ y, int n_head, int S, int nsplit,
                                                          float* __restrict__ scratch, int32_t* __restrict__ counters,
                                                          int use_pdl, unsigned long long* __restrict__ trace) {
#ifdef LLB_TRACE  // tuning build: %globaltimer stamps per CTA (profiles/attn_trace.py)
#define ATRACE(i)                                                                                     \
  do {                                                                                                \
    if (trace != nullptr && threadIdx.x == 0) {                                                       \
      unsigned long long t_;                                                                          \
      asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t_));                                           \
      trace[((size_t)blockIdx.y * gridDim.x + blockIdx.x) * 4 + (i)] = t_;                            \
    }                                                                                                 \
  } while (0)
#else
#define ATRACE(i) do { } while (0)
#endif
  ATRACE(0);
  constexpr int EPL = HS / 8;
  constexpr int NW = 4;
  __shared__ float s_m[NW * 4], s_l[NW * 4];
  __shared__ float s_o[NW * 4][HS];
  __shared__ float c_st[CL ? 8 : 1][HS + 2];  // leader CTA: (M, L, O[HS]) of every split
  __shared__ int s_last;
  // use_pdl bit 0: launched as a programmatic dependent (wait for the producer of q / the caches);
  // bits 1-2: where this kernel lets the NEXT kernel of the stream (the c_proj GEMV: its weights do not depend on
  // this kernel) start its prologue and ring fill: 0 = at completion, 1 = at the start, 2 = once this CTA's K/V rows
  // have been consumed (the next kernel's weight stream then does not compete with the latency-bound K/V loads)
  const int trig = use_pdl >> 1;
  if (trig == 1 || (use_pdl & 1)) pdl_launch_dependents();
  if (use_pdl & 1) pdl_wait();
  ATRACE(1);
  const int bh = blockIdx.x, split = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lg = lane >> 3, li = lane & 7;
  const int nvalid = min(pos[2], S);
  const int per = (nvalid + nsplit - 1) / nsplit;
  const int s0 = split * per, s1 = min(s0 + per, nvalid);
  const float scale_log2 = rsqrtf((float)HS) * 1.4426950408889634f;
  const unsigned gmask = 0xffu << (lg * 8);

  float qr[EPL];
  vec_ld<T>::template ld<EPL>(q + (size_t)bh * HS + li * EPL, qr);
#pragma unroll
  for (int e = 0; e < EPL; ++e) qr[e] *= scale_log2;

  float m = -INFINITY, l = 0.f, o[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) o[e] = 0.f;
  const T* kbase = kc + (size_t)bh * S * HS + li * EPL;
  const T* vbase = vc + (size_t)bh * S * HS + li * EPL;
  // R rows of K and V per lane group are requested before any of them is used: one memory round trip per R * 16
  // positions of the CTA instead of one per 16 (the loop was latency-bound: 5 dependent trips at 640 positions)
  constexpr int R = 6;
  for (int sb = s0 + warp * 4 + lg; sb < s1; sb += 4 * NW * R) {
    float kr[R][EPL], vr[R][EPL];
#pragma unroll
    for (int r = 0; r < R; ++r) {
      const int s = min(sb + r * 4 * NW, s1 - 1);  // clamped rows are loaded but not used
      vec_ld<T>::template ld<EPL>(kbase + (size_t)s * HS, kr[r]);
      vec_ld<T>::template ld<EPL>(vbase + (size_t)s * HS, vr[r]);
    }
#pragma unroll
    for (int r = 0; r < R; ++r) {
      float sc = 0.f;
#pragma unroll
      for (int e = 0; e < EPL; ++e) sc = fmaf(qr[e], kr[r][e], sc);
      sc += __shfl_xor_sync(gmask, sc, 1);   // the 8 lanes of a group share a trip count; groups do not
      sc += __shfl_xor_sync(gmask, sc, 2);
      sc += __shfl_xor_sync(gmask, sc, 4);
      if (sb + r * 4 * NW < s1) {
        const float mn = fmaxf(m, sc);
        const float corr = exp2f(m - mn), pr = exp2f(sc - mn);
        l = l * corr + pr;
#pragma unroll
        for (int e = 0; e < EPL; ++e) o[e] = o[e] * corr + pr * vr[r][e];
        m = mn;
      }
    }
  }
  if (trig == 2) pdl_launch_dependents();
  ATRACE(2);
  // combine the 16 (warp, lane-group) partial states of this CTA in fixed order
  if (li == 0) { s_m[warp * 4 + lg] = m; s_l[warp * 4 + lg] = l; }
#pragma unroll
  for (int e = 0; e < EPL; ++e) s_o[warp * 4 + lg][li * EPL + e] = o[e];
  __syncthreads();
  float* my = scratch + ((size_t)bh * nsplit + split) * (HS + 2);
  if (threadIdx.x < HS) {
    float M = -INFINITY;
#pragma unroll
    for (int i = 0; i < NW * 4; ++i) M = fmaxf(M, s_m[i]);
    float L = 0.f, O = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int i = 0; i < NW * 4; ++i) {
        const float c = exp2f(s_m[i] - M);
        L += s_l[i] * c;
        O += s_o[i][threadIdx.x] * c;
      }
    }
    if (nsplit == 1) {
      dt<T>::st(y + (size_t)bh * HS, threadIdx.x, O / L);
    } else if constexpr (CL) {
      // write this split's state into the leader's shared memory (cluster rank 0 = split 0)
      uint32_t remote;
      asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(&c_st[split][0])), "r"(0));
      asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote + 4u * (2 + threadIdx.x)), "f"(O) : "memory");
      if (threadIdx.x == 0) {
        asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(M) : "memory");
        asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote + 4u), "f"(L) : "memory");
      }
    } else {
      my[2 + threadIdx.x] = O;
      if (threadIdx.x == 0) { my[0] = M; my[1] = L; }
    }
  }
  if constexpr (CL) {
    if (nsplit > 1) {
      asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
      asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
      ATRACE(3);
      if (split != 0) return;
      if (threadIdx.x < HS) {
        float M = -INFINITY;
        for (int i = 0; i < nsplit; ++i) M = fmaxf(M, c_st[i][0]);
        float L = 0.f, O = 0.f;
        for (int i = 0; i < nsplit; ++i) {  // split order: deterministic
          const float mi = c_st[i][0];
          if (mi > -INFINITY) {
            const float c = exp2f(mi - M);
            L += c_st[i][1] * c;
            O += c_st[i][2 + threadIdx.x] * c;
          }
        }
        dt<T>::st(y + (size_t)bh * HS, threadIdx.x, O / L);
      }
      return;
    }
  }
  if (nsplit == 1) return;
  __threadfence();
  __syncthreads();
  if (threadIdx.x == 0) s_last = (atomicAdd(&counters[bh], 1) == nsplit - 1);
  __syncthreads();
  if (!s_last) return;
  __threadfence();
  if (threadIdx.x < HS) {
    const float* base = scratch + (size_t)bh * nsplit * (HS + 2);
    float M = -INFINITY;
    for (int i = 0; i < nsplit; ++i) M = fmaxf(M, __ldcg(base + i * (HS + 2)));
    float L = 0.f, O = 0.f;
    for (int i = 0; i < nsplit; ++i) {
      const float mi = __ldcg(base + i * (HS + 2));
      if (mi > -INFINITY) {
        const float c = exp2f(mi - M);
        L += __ldcg(base + i * (HS + 2) + 1) * c;
        O += __ldcg(base + i * (HS + 2) + 2 + threadIdx.x) * c;
      }
    }
    dt<T>::st(y + (size_t)bh * HS, threadIdx.x, O / L);
  }
  if (threadIdx.x == 0) counters[bh] = 0;
}

// ---------------------------------------------------------------- decode attention, K/V staged ahead of the wait
// The kernel above is latency-bound: after the c_attn GEMV retires it pays position -> K/V rows -> combine as dependent
// memory round trips (3.2 us of its 9.4 us per layer at position 640, profiles/r02_gemv_experiments.md).  Only ONE row of
// the layer's cache changes in this step - the slot the c_attn epilogue writes - so everything else can be in shared
// memory before the producer finishes.  As a programmatic dependent this kernel
//   1. BEFORE griddepcontrol.wait: reads the position (a guess: the kernel that advanced it may formally still be running),
//      and one thread issues two bulk-TMA copies - this split's K rows and V rows, contiguous in [head][slot][hs] - into
//      shared memory: they stream while the c_attn GEMV still runs;
//   2. AFTER the wait, in one round trip: the true position, q, and the row at the guessed slot (the row c_attn just wrote)
//      - the latter patches shared memory;
//   3. if the guess was the true position (always, except when a step starts within microseconds of the position update)
//      the softmax runs out of shared memory; otherwise the same loop reads global memory (same arithmetic, same bits).
// One (batch, head) = one cluster of nsplit <= 8 CTAs combining through distributed shared memory, as above.
template <typename T> struct piece16;
template <> struct piece16<__nv_bfloat16> {
  static constexpr int PV = 8;
  static __device__ __forceinline__ void cvt(const uint4& r, float* o) {
    const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      o[2 * j] = bf16_bits_to_f32(w[j] & 0xffff);
      o[2 * j + 1] = bf16_bits_to_f32(w[j] >> 16);
    }
  }
};
template <> struct piece16<float> {
  static constexpr int PV = 4;
  static __device__ __forceinline__ void cvt(const uint4& r, float* o) {
    o[0] = __uint_as_float(r.x); o[1] = __uint_as_float(r.y); o[2] = __uint_as_float(r.z); o[3] = __uint_as_float(r.w);
  }
};
#define LL_ATTN_FLAG_MISGUESS 0x100  // test hook (ll_decode_attn flags): a wrong position guess -> the global-memory path
#define LL_ATTN_FLAG_POISON 0x200    // test hook: the staged copy of the producer's row is NaN before it is patched

template <typename T, int HS>
__global__ void __launch_bounds__(128) decode_attn_pf_kernel(const T* __restrict__ q, const T* __restrict__ kc,
                                                             const T* __restrict__ vc, const int32_t* __restrict__ pos,
                                                             T* __restrict__ y, int S, int nsplit, int rows_cap, int use_pdl,
                                                             unsigned long long* __restrict__ trace, uint8_t* __restrict__ dg_planes,
                                                             float* __restrict__ dg_rec, int dg_stride) {
  ATRACE(0);
  constexpr int EPL = HS / 8, PV = piece16<T>::PV, NP = EPL / PV, PPR = HS / PV, NW = 4, RL = 6;
  static_assert(EPL % PV == 0, "a lane takes whole 16-byte pieces");
  extern __shared__ __align__(128) uint8_t dyn[];
  T* ks = reinterpret_cast<T*>(dyn);
  T* vs = ks + (size_t)rows_cap * HS;
  float* s_o = reinterpret_cast<float*>(dyn);  // [NW * 4][HS], aliased onto the K rows once every warp has left the loop
  __shared__ float s_m[NW * 4], s_l[NW * 4];
  __shared__ float c_st[8][HS + 2];
  __shared__ __align__(8) uint64_t bar;
  const int trig = (use_pdl >> 1) & 3;
  if (trig == 1 || (use_pdl & 1)) pdl_launch_dependents();
  const int bh = blockIdx.x, split = blockIdx.y;
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31, lg = lane >> 3, li = lane & 7;

  // ---- 1. the guess and the copies
  int g_nvalid = min(max(__ldcg(pos + 2), 0), S), g_slot = min(max(__ldcg(pos + 1), 0), S - 1);
  if (use_pdl & LL_ATTN_FLAG_MISGUESS) g_slot = (g_slot + 1) % S;
  const int gper = (g_nvalid + nsplit - 1) / nsplit;
  const int gs0 = min(split * gper, g_nvalid), gs1 = min(gs0 + gper, g_nvalid);
  const int rows = gs1 - gs0;  // <= rows_cap = ceil(S / nsplit)
  if (threadIdx.x == 0) {
    mbar_init(&bar, 1);
    fence_barrier_init();
    if (rows > 0) {
      const uint32_t bytes = (uint32_t)rows * HS * (uint32_t)sizeof(T);
      mbar_expect_tx(&bar, 2 * bytes);
      bulk_load_1d(ks, kc + ((size_t)bh * S + gs0) * HS, bytes, &bar);
      bulk_load_1d(vs, vc + ((size_t)bh * S + gs0) * HS, bytes, &bar);
    }
  }
  __syncthreads();
  if (use_pdl & 1) pdl_wait();
  ATRACE(1);

  // ---- 2. one round trip: true position, q, the row the producer wrote
  const int nvalid = min(__ldcg(pos + 2), S), slot = __ldcg(pos + 1);
  uint4 qraw[NP];
#pragma unroll
  for (int j = 0; j < NP; ++j) qraw[j] = __ldcg(reinterpret_cast<const uint4*>(q + (size_t)bh * HS + (j * 8 + li) * PV));
  const bool fix_mine = g_slot >= gs0 && g_slot < gs1 && threadIdx.x < 2 * PPR;
  const int fpiece = threadIdx.x < PPR ? threadIdx.x : threadIdx.x - PPR;
  uint4 fx = make_uint4(0u, 0u, 0u, 0u);
  if (fix_mine)
    fx = __ldcg(reinterpret_cast<const uint4*>((threadIdx.x < PPR ? kc : vc) + ((size_t)bh * S + g_slot) * HS + fpiece * PV));
  const bool ok = nvalid == g_nvalid && slot == g_slot;  // uniform over the cluster
  if (rows > 0) mbar_wait(&bar, 0);
  if (fix_mine && (use_pdl & LL_ATTN_FLAG_POISON))
    *reinterpret_cast<uint4*>((threadIdx.x < PPR ? ks : vs) + (size_t)(g_slot - gs0) * HS + fpiece * PV) =
        make_uint4(0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u, 0x7fc07fc0u);
  if (ok && fix_mine)
    *reinterpret_cast<uint4*>((threadIdx.x < PPR ? ks : vs) + (size_t)(g_slot - gs0) * HS + fpiece * PV) = fx;
  __syncthreads();

  const float scale_log2 = rsqrtf((float)HS) * 1.4426950408889634f;
  float qr[EPL];
#pragma unroll
  for (int j = 0; j < NP; ++j) piece16<T>::cvt(qraw[j], qr + j * PV);
#pragma unroll
  for (int e = 0; e < EPL; ++e) qr[e] *= scale_log2;
  const int per = (nvalid + nsplit - 1) / nsplit;
  const int s0 = min(split * per, nvalid), s1 = min(s0 + per, nvalid);
  const unsigned gmask = 0xffu << (lg * 8);

  // ---- 3. softmax over this split's rows: RL rows per lane group at a time (scores first, then one rescale)
  float m = -INFINITY, l = 0.f, o[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) o[e] = 0.f;
  auto run = [&](auto load_piece) {
    for (int sb = s0 + warp * 4 + lg; sb < s1; sb += 4 * NW * RL) {
      float sc[RL];
#pragma unroll
      for (int r = 0; r < RL; ++r) {
        const int s = min(sb + r * 4 * NW, s1 - 1);  // clamped rows are read but not used
        float a = 0.f;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
          float kr[PV];
          piece16<T>::cvt(load_piece(0, s, j), kr);
#pragma unroll
          for (int i = 0; i < PV; ++i) a = fmaf(qr[j * PV + i], kr[i], a);
        }
        a += __shfl_xor_sync(gmask, a, 1);
        a += __shfl_xor_sync(gmask, a, 2);
        a += __shfl_xor_sync(gmask, a, 4);
        sc[r] = sb + r * 4 * NW < s1 ? a : -INFINITY;
      }
      float mn = m;
#pragma unroll
      for (int r = 0; r < RL; ++r) mn = fmaxf(mn, sc[r]);
      const float corr = exp2f(m - mn);  // sc[0] is finite: mn > -inf
      l *= corr;
#pragma unroll
      for (int e = 0; e < EPL; ++e) o[e] *= corr;
#pragma unroll
      for (int r = 0; r < RL; ++r) {
        const float pr = exp2f(sc[r] - mn);  // 0 for the clamped rows
        const int s = min(sb + r * 4 * NW, s1 - 1);
        l += pr;
#pragma unroll
        for (int j = 0; j < NP; ++j) {
          float vr[PV];
          piece16<T>::cvt(load_piece(1, s, j), vr);
#pragma unroll
          for (int i = 0; i < PV; ++i) o[j * PV + i] = fmaf(pr, vr[i], o[j * PV + i]);
        }
      }
      m = mn;
    }
  };
  if (ok) {
    const uint32_t ku = smem_u32(ks), vu = smem_u32(vs);
    run([&](int which, int s, int j) {
      return lds128((which ? vu : ku) + (uint32_t)(((s - s0) * HS + (j * 8 + li) * PV) * (int)sizeof(T)));
    });
  } else {
    run([&](int which, int s, int j) {
      return __ldcg(reinterpret_cast<const uint4*>((which ? vc : kc) + ((size_t)bh * S + s) * HS + (j * 8 + li) * PV));
    });
  }
  if (trig == 2) pdl_launch_dependents();
  ATRACE(2);
  __syncthreads();  // s_o aliases the K rows
  if (li == 0) { s_m[warp * 4 + lg] = m; s_l[warp * 4 + lg] = l; }
#pragma unroll
  for (int j = 0; j < NP; ++j)
#pragma unroll
    for (int i = 0; i < PV; ++i) s_o[(warp * 4 + lg) * HS + (j * 8 + li) * PV + i] = o[j * PV + i];
  __syncthreads();
  if (threadIdx.x < HS) {
    float M = -INFINITY;
#pragma unroll
    for (int i = 0; i < NW * 4; ++i) M = fmaxf(M, s_m[i]);
    float L = 0.f, O = 0.f;
    if (M > -INFINITY) {
#pragma unroll
      for (int i = 0; i < NW * 4; ++i) {
        const float c = exp2f(s_m[i] - M);
        L += s_l[i] * c;
        O += s_o[i * HS + threadIdx.x] * c;
      }
    }
    uint32_t remote;
    asm volatile("mapa.shared::cluster.u32 %0, %1, %2;" : "=r"(remote) : "r"(smem_u32(&c_st[split][0])), "r"(0));
    asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote + 4u * (2 + threadIdx.x)), "f"(O) : "memory");
    if (threadIdx.x == 0) {
      asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote), "f"(M) : "memory");
      asm volatile("st.shared::cluster.f32 [%0], %1;" ::"r"(remote + 4u), "f"(L) : "memory");
    }
  }
  asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
  asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
  ATRACE(3);
  if (split != 0) return;
  float outv = 0.f;
  if (threadIdx.x < HS) {
    float M = -INFINITY;
    for (int i = 0; i < nsplit; ++i) M = fmaxf(M, c_st[i][0]);
    float L = 0.f, O = 0.f;
    for (int i = 0; i < nsplit; ++i) {  // split order: deterministic
      const float mi = c_st[i][0];
      if (mi > -INFINITY) {
        const float c = exp2f(mi - M);
        L += c_st[i][1] * c;
        O += c_st[i][2 + threadIdx.x] * c;
      }
    }
    outv = O / L;
    dt<T>::st(y + (size_t)bh * HS, threadIdx.x, outv);
  }
  if constexpr (HS == 128 && sizeof(T) == 2) {
    // One head = one 128-element k-block of attn.c_proj's input: the leader also emits the block in the form the batch-1
    // integer-MMA GEMV stages its activations in - four signed base-256 digit planes of x * 2^sh and the block's record
    // (2^-sh, sum x) - so that c_proj's prologue is one bulk copy instead of a conversion every CTA repeats.  Same
    // arithmetic, in the same order, as gemv_w4a8_kernel's own staging (csrc/gemv_int4.cu): identical bits.
    if (dg_planes != nullptr) {
      __shared__ uint32_t s_mx[4];
      __shared__ float s_ps[16];
      const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
      const uint32_t xb = (uint32_t)f32_to_bf16_bits(outv);        // the value as stored
      const float xf = bf16_bits_to_f32(xb);
      const uint32_t mw = __reduce_max_sync(0xffffffffu, xb & 0x7fffu);
      if (lane == 0) s_mx[warp] = mw;
      // piece sums: 8 consecutive elements, two chains (x0 + x1 + x4 + x5) + (x2 + x3 + x6 + x7), as the GEMV's staging
      const int l0 = lane & ~7;
      float e[8];
#pragma unroll
      for (int t = 0; t < 8; ++t) e[t] = __shfl_sync(0xffffffffu, xf, l0 + t);
      float c0 = 0.f, c1 = 0.f;
      c0 = __fadd_rn(c0, e[0]); c1 = __fadd_rn(c1, e[2]);
      c0 = __fadd_rn(c0, e[1]); c1 = __fadd_rn(c1, e[3]);
      c0 = __fadd_rn(c0, e[4]); c1 = __fadd_rn(c1, e[6]);
      c0 = __fadd_rn(c0, e[5]); c1 = __fadd_rn(c1, e[7]);
      if ((lane & 7) == 0) s_ps[warp * 4 + (lane >> 3)] = __fadd_rn(c0, c1);
      __syncthreads();
      const uint32_t mx = max(max(s_mx[0], s_mx[1]), max(s_mx[2], s_mx[3]));
      int ex = (int)((mx >> 7) & 0xffu);
      ex = max(ex, 32);
      const float scale = __uint_as_float(((uint32_t)(283 - ex) << 7) << 16);   // bf16 2^(156 - ex)
      const int vi = __float2int_rn(__fmul_rn(xf, scale));                    // exact product of two bf16 values
      const uint32_t u = ((uint32_t)vi + 0x00808080u) ^ 0x00808080u;
      uint8_t* dst = dg_planes + (size_t)bh * HS + threadIdx.x;
      dst[0] = (uint8_t)(u & 0xffu);
      dst[dg_stride] = (uint8_t)((u >> 8) & 0xffu);
      dst[2 * (size_t)dg_stride] = (uint8_t)((u >> 16) & 0xffu);
      dst[3 * (size_t)dg_stride] = (uint8_t)(u >> 24);
      if (threadIdx.x == 0) {
        float t1[8], t2[4];   // the 16-lane xor tree of the staging loop: partners 8, 4, 2, 1 apart
#pragma unroll
        for (int i = 0; i < 8; ++i) t1[i] = __fadd_rn(s_ps[i], s_ps[i + 8]);
#pragma unroll
        for (int i = 0; i < 4; ++i) t2[i] = __fadd_rn(t1[i], t1[i + 4]);
        const float sum = __fadd_rn(__fadd_rn(t2[0], t2[2]), __fadd_rn(t2[1], t2[3]));
        const float inv = __uint_as_float((uint32_t)(ex - 29) << 23);
        *reinterpret_cast<float4*>(dg_rec + (size_t)bh * 4) = make_float4(inv, sum, inv * 65536.f, 0.f);
      }
    }
  }
}

// The K/V-staging kernel serves the shapes whose split fits shared memory beside a resident GEMV CTA (two attention CTAs
// of <= 54 KB next to a 116 KB one); LLB200_ATTN_STAGE=0 keeps the direct kernel (A/B measurements).
constexpr size_t PF_DYN_MAX = 50 * 1024;
template <typename T, int HS>
size_t pf_dyn_bytes(int rows_cap) {
  const size_t kv = (size_t)2 * rows_cap * HS * sizeof(T), so = (size_t)16 * HS * sizeof(float);
  return kv > so ? kv : so;
}
template <typename T, int HS>
bool pf_ok(int cache_len, int splits) {
  static int on = -1;
  if (on < 0) {
    const char* e = getenv("LLB200_ATTN_STAGE");
    on = (e && atoi(e) == 0) ? 0 : 1;
  }
  return on && pf_dyn_bytes<T, HS>((cache_len + splits - 1) / splits) <= PF_DYN_MAX;
}
// 1: launched, 0: not for this shape, < 0: error
template <typename T, int HS>
int launch_attn_pf(const void* q, const void* k_cache, const void* v_cache, const int32_t* pos, void* y, int cache_len,
                   int splits, dim3 grid, dim3 block, cudaStream_t st, bool pdl, int kflags, unsigned long long* tr,
                   uint8_t* dg_planes, float* dg_rec, int dg_stride) {
  if constexpr ((HS / 8) % piece16<T>::PV != 0) {
    return 0;
  } else {
    if (!pf_ok<T, HS>(cache_len, splits)) return 0;
    const int rows_cap = (cache_len + splits - 1) / splits;
    const size_t dyn = pf_dyn_bytes<T, HS>(rows_cap);
    static size_t configured = 0;
    if (dyn > configured) {
      if (configured == 0) prefer_max_smem(decode_attn_pf_kernel<T, HS>);
      LL_CUDA(cudaFuncSetAttribute(decode_attn_pf_kernel<T, HS>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dyn));
      configured = dyn;
    }
    cudaLaunchConfig_t cfg{};
    cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = dyn; cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = (unsigned)splits; attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = pdl ? 2 : 1;
    LL_CUDA(cudaLaunchKernelEx(&cfg, decode_attn_pf_kernel<T, HS>, (const T*)q, (const T*)k_cache, (const T*)v_cache, pos,
                               (T*)y, cache_len, splits, rows_cap, kflags, tr, dg_planes, dg_rec, dg_stride));
    return 1;
  }
}

// ---------------------------------------------------------------- elementwise
template <typename T>
__global__ void silu_mul_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) {
    const float av = dt<T>::ld(a, i);
    const float s = dt<T>::rd(av / (1.0f + expf(-av)));
    dt<T>::st(y, i, __fmul_rn(s, dt<T>::ld(b, i)));
  }
}
template <typename T>
__global__ void add_kernel(const T* __restrict__ a, const T* __restrict__ b, T* __restrict__ y, size_t n) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) dt<T>::st(y, i, __fadd_rn(dt<T>::ld(a, i), dt<T>::ld(b, i)));
}
// A position past the RoPE table (ref: model.py:101 rope_cache.index_select raises) stops the kernel: the error
// surfaces at the next synchronisation instead of an out-of-bounds read in the c_attn epilogue.
__global__ void advance_pos_kernel(int32_t* pos, int S, int max_pos) {
  const int p = pos[0] + 1;
  if (p < 0 || p >= max_pos) __trap();
  pos[0] = p;
  pos[1] = p % S;
  pos[2] = min(p + 1, S);
}

// standalone apply_rope (ref: model.py:312-329): x,y [B][T][nh][hs], rope [T][hs/2][2] fp32
template <typename T>
__global__ void apply_rope_kernel(const T* __restrict__ x, const float* __restrict__ rope, T* __restrict__ y, int Tn,
                                  int n_head, int hs, size_t npairs) {
  const size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= npairs) return;
  const int half = hs / 2;
  const int pi = (int)(i % half);
  const int t = (int)((i / half / n_head) % Tn);
  const float2 cs = *reinterpret_cast<const float2*>(rope + ((size_t)t * half + pi) * 2);
  const float x0 = dt<T>::ld(x, 2 * i), x1 = dt<T>::ld(x, 2 * i + 1);
  dt<T>::st(y, 2 * i, __fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y)));
  dt<T>::st(y, 2 * i + 1, __fadd_rn(__fmul_rn(x1, cs.x), __fmul_rn(x0, cs.y)));
}

__global__ void set_pos_kernel(int32_t* pos, const int64_t* input_pos, int S, int max_pos) {
  if (input_pos[0] < 0 || input_pos[0] >= (int64_t)max_pos) __trap();
  const int p = (int)input_pos[0];
  pos[0] = p;
  pos[1] = p % S;
  pos[2] = min(p + 1, S);
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

int ll_rmsnorm(const void* x, const void* w, void* y, int dtype, int rows, int dim, float eps, ll_stream stream) {
  LL_CHECK_ARG(x && w && y && rows >= 0 && dim > 0, "bad rmsnorm arguments");
  if (rows == 0) return LL_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int threads = dim >= 1024 ? 256 : 128;
  if (dtype == LL_BF16 && dim % 8 == 0 && dim / 8 <= 1024 * RN_PIECES && ((uintptr_t)x & 15) == 0 && ((uintptr_t)y & 15) == 0 &&
      ((uintptr_t)w & 15) == 0) {
    int vt = ((dim / 8 + 31) / 32) * 32;   // one piece per thread up to 1024 threads, then more pieces per thread
    if (vt > 1024) vt = 1024;
    if (vt > 256 && rows > 1024 && dim / 8 <= 256 * RN_PIECES) vt = 256;   // many rows (prefill): smaller CTAs, several pieces per thread
    rmsnorm_bf16_vec_kernel<<<rows, vt, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w, (__nv_bfloat16*)y, dim, eps);
    LL_LAUNCH_OK("rmsnorm_bf16_vec_kernel");
    return LL_OK;
  }
  if (dtype == LL_F32) rmsnorm_kernel<float><<<rows, threads, 0, st>>>((const float*)x, (const float*)w, (float*)y, dim, eps);
  else if (dtype == LL_BF16) rmsnorm_kernel<__nv_bfloat16><<<rows, threads, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)w, (__nv_bfloat16*)y, dim, eps);
  else {
    set_error("rmsnorm: unsupported dtype %d", dtype);
    return LL_ERR_UNSUPPORTED;
  }
  LL_LAUNCH_OK("rmsnorm_kernel");
  return LL_OK;
}

int ll_embed_ssq(const void* wte, const int64_t* idx, void* x, float* ssq, int M, int dim, ll_stream stream) {
  LL_CHECK_ARG(wte && idx && x && M > 0 && dim > 0 && dim % 16 == 0, "bad embed arguments (dim must be a multiple of 16)");
  const int nchunks = (dim + 127) / 128;
  static bool once = false;
  if (!once) { prefer_max_smem(embed_ssq_kernel); prefer_max_smem(set_pos_kernel); prefer_max_smem(advance_pos_kernel); once = true; }
  embed_ssq_kernel<<<dim3(nchunks, M), 128, 0, (cudaStream_t)stream>>>((const __nv_bfloat16*)wte, idx, (__nv_bfloat16*)x, ssq, dim, dim / 16);
  LL_LAUNCH_OK("embed_ssq_kernel");
  return LL_OK;
}

int ll_rope_kv_append(const void* qkv, const float* rope, const int32_t* pos_ids, const int32_t* slots, void* q_out,
                      void* k_cache, void* v_cache, int dtype, int B, int T, int n_head, int head_size, int cache_len,
                      ll_stream stream) {
  LL_CHECK_ARG(qkv && rope && pos_ids && slots && q_out && k_cache && v_cache, "null pointer");
  LL_CHECK_ARG(B > 0 && T > 0 && n_head > 0 && head_size > 0 && head_size % 2 == 0 && cache_len > 0, "bad shape");
  LL_CHECK_ARG(head_size / 2 <= 1024 && (long)B * n_head <= 65535, "shape too large");
  dim3 grid(T, B * n_head), block(((head_size / 2) + 31) / 32 * 32);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LL_F32) rope_kv_kernel<float><<<grid, block, 0, st>>>((const float*)qkv, rope, pos_ids, slots, (float*)q_out, (float*)k_cache, (float*)v_cache, T, n_head, head_size, cache_len);
  else if (dtype == LL_BF16) rope_kv_kernel<__nv_bfloat16><<<grid, block, 0, st>>>((const __nv_bfloat16*)qkv, rope, pos_ids, slots, (__nv_bfloat16*)q_out, (__nv_bfloat16*)k_cache, (__nv_bfloat16*)v_cache, T, n_head, head_size, cache_len);
  else {
    set_error("rope_kv_append: unsupported dtype %d", dtype);
    return LL_ERR_UNSUPPORTED;
  }
  LL_LAUNCH_OK("rope_kv_kernel");
  return LL_OK;
}

size_t ll_decode_attn_scratch_bytes(int B, int n_head, int head_size, int max_splits) {
  return (size_t)B * n_head * max_splits * (head_size + 2) * sizeof(float);
}

unsigned long long* g_attn_trace = nullptr;  // debug: device buffer [launches][CTAs <= 512][4]
int g_attn_trace_launches = 0, g_attn_trace_idx = 0;
/* debug (not in the public header): stamp the next `launches` decode-attention launches */
__attribute__((visibility("default"))) void ll_debug_attn_trace(unsigned long long* buf, int launches) {
  g_attn_trace = buf; g_attn_trace_launches = launches; g_attn_trace_idx = 0;
}

static int decode_attn_impl(const void* q, const void* k_cache, const void* v_cache, const int32_t* pos, void* y, int dtype, int B,
                            int n_head, int head_size, int cache_len, int max_splits, float* scratch, int32_t* counters,
                            int flags, ll_stream stream, uint8_t* dg_planes, float* dg_rec, int dg_stride) {
  LL_CHECK_ARG(q && k_cache && v_cache && pos && y, "null pointer");
  LL_CHECK_ARG(B > 0 && n_head > 0 && cache_len > 0 && max_splits >= 1 && max_splits <= 64, "bad shape");
  if (max_splits > 8) LL_CHECK_ARG(scratch && counters, "decode attention with more than 8 splits needs scratch and counters");
  dim3 grid(B * n_head, max_splits), block(128);
  cudaStream_t st = (cudaStream_t)stream;
  const bool pdl = (flags & LL_GEMV_PDL) != 0;
  // 2..8 splits: one cluster per (batch, head), partial states exchanged through distributed shared memory
  static int use_cluster = -1;
  if (use_cluster < 0) {
    const char* e = getenv("LLB200_ATTN_CLUSTER");
    use_cluster = (e && atoi(e) == 0) ? 0 : 1;
  }
  static int trig = -1;
  if (trig < 0) {
    const char* e = getenv("LLB200_ATTN_TRIGGER");
    trig = e ? atoi(e) : 0;
    if (trig < 0 || trig > 2) trig = 0;
  }
  const int kflags = (pdl ? 1 : 0) | (trig << 1);
  unsigned long long* tr = nullptr;
  if (g_attn_trace != nullptr && g_attn_trace_idx < g_attn_trace_launches && grid.x * grid.y <= 512)
    tr = g_attn_trace + (size_t)(g_attn_trace_idx++) * 512 * 4;
  const bool cl = use_cluster && max_splits >= 2 && max_splits <= 8;
#define ATTN(TT, HS)                                                                                              \
  do {                                                                                                            \
    static bool once = false;                                                                                     \
    if (!once) {                                                                                                  \
      prefer_max_smem(decode_attn_kernel<TT, HS, false>);                                                         \
      prefer_max_smem(decode_attn_kernel<TT, HS, true>);                                                          \
      once = true;                                                                                                \
    }                                                                                                             \
    const int pfr = cl ? launch_attn_pf<TT, HS>(q, k_cache, v_cache, pos, y, cache_len, max_splits, grid, block, st, pdl, \
                                                kflags | (flags & (LL_ATTN_FLAG_MISGUESS | LL_ATTN_FLAG_POISON)), tr,  \
                                                dg_planes, dg_rec, dg_stride) : 0;                                \
    if (pfr < 0) return pfr;                                                                                      \
    if (pfr == 0 && dg_planes != nullptr) {                                                                       \
      set_error("decode_attn_digits: this shape does not take the K/V-staging kernel");                           \
      return LL_ERR_UNSUPPORTED;                                                                                  \
    }                                                                                                             \
    if (pfr > 0) {                                                                                                \
    } else if (cl) {                                                                                              \
      cudaLaunchConfig_t cfg{};                                                                                   \
      cfg.gridDim = grid; cfg.blockDim = block; cfg.dynamicSmemBytes = 0; cfg.stream = st;                        \
      cudaLaunchAttribute attr[2];                                                                                \
      attr[0].id = cudaLaunchAttributeClusterDimension;                                                           \
      attr[0].val.clusterDim.x = 1; attr[0].val.clusterDim.y = (unsigned)max_splits; attr[0].val.clusterDim.z = 1; \
      attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;                                            \
      attr[1].val.programmaticStreamSerializationAllowed = 1;                                                     \
      cfg.attrs = attr; cfg.numAttrs = pdl ? 2 : 1;                                                               \
      LL_CUDA(cudaLaunchKernelEx(&cfg, decode_attn_kernel<TT, HS, true>, (const TT*)q, (const TT*)k_cache,        \
                                 (const TT*)v_cache, pos, (TT*)y, n_head, cache_len, max_splits, scratch, counters, \
                                 kflags, tr));                                                                 \
    } else {                                                                                                      \
      LL_CUDA(launch(decode_attn_kernel<TT, HS, false>, grid, block, 0, st, pdl, (const TT*)q, (const TT*)k_cache, \
                     (const TT*)v_cache, pos, (TT*)y, n_head, cache_len, max_splits, scratch, counters, kflags, tr)); \
    }                                                                                                             \
  } while (0)
  if (dtype == LL_BF16) {
    if (head_size == 128) ATTN(__nv_bfloat16, 128);
    else if (head_size == 64) ATTN(__nv_bfloat16, 64);
    else if (head_size == 32) ATTN(__nv_bfloat16, 32);
    else { set_error("decode_attn: head_size %d unsupported (32, 64, 128)", head_size); return LL_ERR_UNSUPPORTED; }
  } else if (dtype == LL_F32) {
    if (head_size == 128) ATTN(float, 128);
    else if (head_size == 64) ATTN(float, 64);
    else if (head_size == 32) ATTN(float, 32);
    else { set_error("decode_attn: head_size %d unsupported (32, 64, 128)", head_size); return LL_ERR_UNSUPPORTED; }
  } else {
    set_error("decode_attn: unsupported dtype %d", dtype);
    return LL_ERR_UNSUPPORTED;
  }
#undef ATTN
  return LL_OK;
}

int ll_decode_attn(const void* q, const void* k_cache, const void* v_cache, const int32_t* pos, void* y, int dtype, int B,
                   int n_head, int head_size, int cache_len, int max_splits, float* scratch, int32_t* counters,
                   int flags, ll_stream stream) {
  return decode_attn_impl(q, k_cache, v_cache, pos, y, dtype, B, n_head, head_size, cache_len, max_splits, scratch, counters,
                          flags, stream, nullptr, nullptr, 0);
}

int ll_decode_attn_digits_supported(int dtype, int B, int head_size, int cache_len, int max_splits) {
  static int use_cluster = -1;
  if (use_cluster < 0) {
    const char* e = getenv("LLB200_ATTN_CLUSTER");
    use_cluster = (e && atoi(e) == 0) ? 0 : 1;
  }
  return (dtype == LL_BF16 && B == 1 && head_size == 128 && use_cluster && max_splits >= 2 && max_splits <= 8 &&
          pf_ok<__nv_bfloat16, 128>(cache_len, max_splits)) ? 1 : 0;
}

int ll_decode_attn_digits(const void* q, const void* k_cache, const void* v_cache, const int32_t* pos, void* y, int B, int n_head,
                          int head_size, int cache_len, int max_splits, int flags, uint8_t* planes, float* rec, int plane_stride,
                          ll_stream stream) {
  LL_CHECK_ARG(planes && rec && plane_stride >= n_head * head_size && plane_stride % 16 == 0, "bad digit-plane arguments");
  LL_CHECK_ARG(ll_decode_attn_digits_supported(LL_BF16, B, head_size, cache_len, max_splits) == 1,
               "decode_attn_digits needs bf16, batch 1, head size 128 and a split that fits the K/V-staging kernel");
  return decode_attn_impl(q, k_cache, v_cache, pos, y, LL_BF16, B, n_head, head_size, cache_len, max_splits, nullptr, nullptr, flags,
                          stream, planes, rec, plane_stride);
}

int ll_silu_mul(const void* a, const void* b, void* y, int dtype, size_t n, ll_stream stream) {
  LL_CHECK_ARG(a && b && y, "null pointer");
  if (n == 0) return LL_OK;
  const unsigned blocks = (unsigned)((n + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LL_F32) silu_mul_kernel<float><<<blocks, 256, 0, st>>>((const float*)a, (const float*)b, (float*)y, n);
  else if (dtype == LL_BF16) silu_mul_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)y, n);
  else { set_error("silu_mul: unsupported dtype %d", dtype); return LL_ERR_UNSUPPORTED; }
  LL_LAUNCH_OK("silu_mul_kernel");
  return LL_OK;
}

int ll_add(const void* a, const void* b, void* y, int dtype, size_t n, ll_stream stream) {
  LL_CHECK_ARG(a && b && y, "null pointer");
  if (n == 0) return LL_OK;
  const unsigned blocks = (unsigned)((n + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LL_F32) add_kernel<float><<<blocks, 256, 0, st>>>((const float*)a, (const float*)b, (float*)y, n);
  else if (dtype == LL_BF16) add_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)a, (const __nv_bfloat16*)b, (__nv_bfloat16*)y, n);
  else { set_error("add: unsupported dtype %d", dtype); return LL_ERR_UNSUPPORTED; }
  LL_LAUNCH_OK("add_kernel");
  return LL_OK;
}

int ll_advance_pos(int32_t* pos, int cache_len, int max_pos, ll_stream stream) {
  LL_CHECK_ARG(pos && cache_len > 0 && max_pos > 0, "bad advance_pos arguments");
  advance_pos_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(pos, cache_len, max_pos);
  LL_LAUNCH_OK("advance_pos_kernel");
  return LL_OK;
}

int ll_apply_rope(const void* x, const float* rope, void* y, int dtype, int B, int T, int n_head, int head_size,
                  ll_stream stream) {
  LL_CHECK_ARG(x && rope && y && B > 0 && T > 0 && n_head > 0 && head_size > 0 && head_size % 2 == 0, "bad apply_rope arguments");
  const size_t npairs = (size_t)B * T * n_head * (head_size / 2);
  const unsigned blocks = (unsigned)((npairs + 255) / 256);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LL_F32) apply_rope_kernel<float><<<blocks, 256, 0, st>>>((const float*)x, rope, (float*)y, T, n_head, head_size, npairs);
  else if (dtype == LL_BF16) apply_rope_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)x, rope, (__nv_bfloat16*)y, T, n_head, head_size, npairs);
  else if (dtype == LL_F16) apply_rope_kernel<__half><<<blocks, 256, 0, st>>>((const __half*)x, rope, (__half*)y, T, n_head, head_size, npairs);
  else { set_error("apply_rope: unsupported dtype %d", dtype); return LL_ERR_UNSUPPORTED; }
  LL_LAUNCH_OK("apply_rope_kernel");
  return LL_OK;
}

int ll_set_pos(int32_t* pos, const int64_t* input_pos, int cache_len, int max_pos, ll_stream stream) {
  LL_CHECK_ARG(pos && input_pos && cache_len > 0 && max_pos > 0, "bad set_pos arguments");
  set_pos_kernel<<<1, 1, 0, (cudaStream_t)stream>>>(pos, input_pos, cache_len, max_pos);
  LL_LAUNCH_OK("set_pos_kernel");
  return LL_OK;
}

}  // extern "C"
