// int4 (GPTQ column-block) decode GEMV for sm_100a: y[M<=8][N] = x[M][K] . dequant(qw)^T
//
// Replaces the reference's Triton launcher (ref: lit_llama/quantization.py:282-331, which pads
// M=1 to 256 rows and reads every packed byte twice) and its dequantise-then-F.linear fallback
// (ref: quantization.py:390-421) for batch-1..8 decode.
//
// Design (B200-first; HBM-bound integer/byte work, no reshaping into a big GEMM):
//  * The packed matrix is used IN PLACE in the checkpoint's layout: memory [K/2][N] bytes,
//    N fastest; byte (r, n) holds k=2r (low nibble) and k=2r+1 (high nibble).
//  * Work unit = (matrix, 128-column chunk, K split).  A persistent grid (<= 2 CTAs/SM) walks
//    units round-robin.  One producer warp streams [TILE_ROWS x 128 B] weight tiles with 2-D
//    TMA (128B swizzle, L2 evict-first) plus the group's 512-byte scale/zero row with a 1-D bulk
//    copy into a per-consumer-warp ring of shared-memory slots, gated by mbarriers.  Weights do
//    not depend on the previous kernel, so under programmatic dependent launch the producer
//    starts streaming while the previous kernel is still draining.
//  * Consumer warps: conflict-free LDS.128 of two packed rows, nibble -> bf16 with the
//    0x4300 magic (128+q exact), HSUB2 (128+zero) and HMUL2 scale - bit-identical to
//    get_weight(bf16) - then mma.sync m16n8k16 (bf16 x bf16 -> fp32) with the weights as the
//    A operand: a packed byte's (k, k+1) pair for one column n IS an A-fragment register, so no
//    shuffles.  x sits in shared memory pre-permuted as B fragments; the 8 MMA columns carry up
//    to 8 batch rows for free.
//  * Cross-warp reduction in shared memory (fixed order); split-K partials go to a global
//    scratch and the LAST-arriving CTA of a chunk sums them in split order (deterministic,
//    self-resetting counters) and runs the fused epilogue: plain store, residual add (+ RMSNorm
//    sum-of-squares partials for the next kernel), RoPE + KV-cache write (c_attn), or SwiGLU
//    (c_fc1/c_fc2 as two matrices of one launch).
//  * Optional fused RMSNorm prologue while staging x (ref: model.py:276-283 rounding steps).
#include <mutex>
#include <unordered_map>

#include "common.cuh"

namespace llb {
namespace {

constexpr int CHUNK = 128;  // columns (= packed bytes per row segment) per unit
constexpr int MAXM = 8;

template <int CW_, int SLOTS_, int TILE_ROWS_>
struct Cfg {
  static constexpr int CW = CW_;                // consumer warps
  static constexpr int SLOTS = SLOTS_;          // ring slots per consumer warp
  static constexpr int TILE_ROWS = TILE_ROWS_;  // packed rows per tile (k per tile = 2x)
  static constexpr int TILE_K = 2 * TILE_ROWS_;
  static constexpr int TILE_BYTES = TILE_ROWS_ * CHUNK;
  static constexpr int SZ_BYTES = CHUNK * 4;
  static constexpr int STEPS = TILE_ROWS_ / 8;
  static constexpr int THREADS = (CW_ + 1) * 32;
  static constexpr int NCONS = CW_ * 32;
};

struct Params {
  CUtensorMap map[2];
  const uint32_t* sz[2];
  int nmat;
  int N, K, group, G;
  int M;
  const __nv_bfloat16* x;
  const __nv_bfloat16* norm_w;
  const float* ssq;
  int ssq_parts;
  float eps;
  int epilogue;
  __nv_bfloat16* y;
  const __nv_bfloat16* bias;
  float* ssq_out;
  const float* rope;
  const int32_t* pos;
  __nv_bfloat16* k_cache;
  __nv_bfloat16* v_cache;
  int n_head, head_size, cache_len;
  float* partials;
  int32_t* counters;
  int ksplit;
  int nchunks, ktiles, tiles_per_split, nunits;
  int xs_stride;  // elements per x row in smem (max slice length + 8)
};

struct Unit {
  int mat, chunk, split, t0, t1;
};

__device__ __forceinline__ Unit decode_unit(const Params& p, int u) {
  Unit r;
  r.chunk = u % p.nchunks;
  int rest = u / p.nchunks;
  r.mat = rest % p.nmat;
  r.split = rest / p.nmat;
  r.t0 = r.split * p.tiles_per_split;
  r.t1 = min(r.t0 + p.tiles_per_split, p.ktiles);
  return r;
}

__device__ __forceinline__ void consumer_sync(int n) { asm volatile("bar.sync 1, %0;" ::"r"(n) : "memory"); }

__device__ __forceinline__ uint32_t prmt(uint32_t a, uint32_t b, uint32_t sel) {
  uint32_t r;
  asm("prmt.b32 %0, %1, %2, %3;" : "=r"(r) : "r"(a), "r"(b), "r"(sel));
  return r;
}
__device__ __forceinline__ uint32_t lop_and_or(uint32_t a, uint32_t mask, uint32_t orv) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(r) : "r"(a), "r"(mask), "r"(orv));  // (a & mask) | orv
  return r;
}
__device__ __forceinline__ uint32_t bf2_sub(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("sub.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}
__device__ __forceinline__ uint32_t bf2_mul(uint32_t a, uint32_t b) {
  uint32_t r;
  asm("mul.rn.bf16x2 %0, %1, %2;" : "=r"(r) : "r"(a), "r"(b));
  return r;
}

template <class C>
__global__ void __launch_bounds__(C::THREADS, 2) gemv_int4_kernel(const __grid_constant__ Params p) {
  extern __shared__ __align__(1024) uint8_t smem[];
  // layout: [tiles CW*SLOTS*TILE_BYTES][sz CW*SLOTS*SZ_BYTES][x M*xs_stride*2][red CW*M*CHUNK*4 (also out)][barriers]
  uint8_t* tiles = smem;
  uint8_t* szs = tiles + C::CW * C::SLOTS * C::TILE_BYTES;
  __nv_bfloat16* xs = reinterpret_cast<__nv_bfloat16*>(szs + C::CW * C::SLOTS * C::SZ_BYTES);
  float* red = reinterpret_cast<float*>(reinterpret_cast<uint8_t*>(xs) + ((p.M * p.xs_stride * 2 + 15) & ~15));
  float* outs = red + C::CW * p.M * CHUNK;           // [nmat][M][CHUNK] final sums
  uint64_t* full = reinterpret_cast<uint64_t*>(outs + 2 * p.M * CHUNK);
  uint64_t* empty = full + C::CW * C::SLOTS;
  __shared__ int s_flag;
  __shared__ float s_rnorm[MAXM];

  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

  if (threadIdx.x == 0) {
    for (int i = 0; i < C::CW * C::SLOTS; ++i) {
      mbar_init(&full[i], 1);
      mbar_init(&empty[i], 1);
    }
    fence_barrier_init();
  }
  __syncthreads();
  pdl_launch_dependents();

  if (warp == C::CW) {
    // =============================== producer warp ===============================
    if (lane == 0) {
      prefetch_tensormap(&p.map[0]);
      if (p.nmat > 1) prefetch_tensormap(&p.map[1]);
      const uint64_t pol = policy_evict_first();
      uint32_t fill[C::CW];
#pragma unroll
      for (int w = 0; w < C::CW; ++w) fill[w] = 0;
      for (int u = blockIdx.x; u < p.nunits; u += gridDim.x) {
        const Unit un = decode_unit(p, u);
        const int n0 = un.chunk * CHUNK;
        const uint32_t sz_bytes = (uint32_t)min(CHUNK, p.N - n0) * 4u;
        for (int t = un.t0; t < un.t1; ++t) {
          const int w = (t - un.t0) % C::CW;
          uint32_t cnt = 0;
#pragma unroll
          for (int ww = 0; ww < C::CW; ++ww)
            if (ww == w) cnt = fill[ww]++;
          const int slot = cnt % C::SLOTS;
          const int idx = w * C::SLOTS + slot;
          if (cnt >= (uint32_t)C::SLOTS) mbar_wait(&empty[idx], ((cnt / C::SLOTS) - 1) & 1);
          mbar_expect_tx(&full[idx], C::TILE_BYTES + sz_bytes);
          tma_load_2d(tiles + idx * C::TILE_BYTES, &p.map[un.mat], n0, t * C::TILE_ROWS, &full[idx], pol);
          const int g = (t * C::TILE_K) / p.group;
          bulk_load_1d(szs + idx * C::SZ_BYTES, p.sz[un.mat] + (size_t)g * p.N + n0, sz_bytes, &full[idx]);
        }
      }
    }
    return;
  }

  // ================================= consumer warps =================================
  const int g = lane >> 2, tig = lane & 3;
  const int ctid = threadIdx.x;  // 0..NCONS-1
  pdl_wait();  // activations (x, ssq, residual, caches, scratch) come from earlier kernels

  uint32_t consumed = 0;
  int staged_t0 = -1, staged_t1 = -1;

  for (int u = blockIdx.x; u < p.nunits; u += gridDim.x) {
    const Unit un = decode_unit(p, u);
    const int n0 = un.chunk * CHUNK;
    const int kbeg = un.t0 * C::TILE_K;
    const int klen = (un.t1 - un.t0) * C::TILE_K;

    // ---- stage x[:, kbeg : kbeg+klen] as pre-permuted B fragments (x0,x2,x1,x3 per 4)
    if (staged_t0 != un.t0 || staged_t1 != un.t1) {
      consumer_sync(C::NCONS);  // previous unit's readers are done with xs
      if (p.norm_w != nullptr) {
        if (ctid < p.M) {
          float s = 0.f;
          for (int i = 0; i < p.ssq_parts; ++i) s += p.ssq[ctid * p.ssq_parts + i];
          const float mean = bf16r(s * (1.0f / (float)p.K));
          s_rnorm[ctid] = bf16r(rsqrtf(bf16r(mean + p.eps)));
        }
        consumer_sync(C::NCONS);
      }
      for (int m = 0; m < p.M; ++m) {
        const float rn = p.norm_w != nullptr ? s_rnorm[m] : 0.f;
        for (int q4 = ctid; q4 < klen / 4; q4 += C::NCONS) {
          const int k = kbeg + q4 * 4;
          uint16_t v[4];
          if (k + 3 < p.K) {
            const uint2 raw = *reinterpret_cast<const uint2*>(p.x + (size_t)m * p.K + k);
            v[0] = raw.x & 0xffff; v[1] = raw.x >> 16; v[2] = raw.y & 0xffff; v[3] = raw.y >> 16;
          } else {
#pragma unroll
            for (int j = 0; j < 4; ++j)
              v[j] = (k + j < p.K) ? __bfloat16_as_ushort(p.x[(size_t)m * p.K + k + j]) : (uint16_t)0;
          }
          if (p.norm_w != nullptr) {
#pragma unroll
            for (int j = 0; j < 4; ++j) {
              if (k + j < p.K) {
                const float xn = bf16r(__fmul_rn(bf16_bits_to_f32(v[j]), rn));
                v[j] = f32_to_bf16_bits(__fmul_rn(__bfloat162float(p.norm_w[k + j]), xn));
              }
            }
          }
          uint2 o;
          o.x = (uint32_t)v[0] | ((uint32_t)v[2] << 16);
          o.y = (uint32_t)v[1] | ((uint32_t)v[3] << 16);
          *reinterpret_cast<uint2*>(xs + (size_t)m * p.xs_stride + q4 * 4) = o;
        }
      }
      staged_t0 = un.t0; staged_t1 = un.t1;
      consumer_sync(C::NCONS);
    }

    // ---- stream this warp's tiles
    float acc[8][4];
#pragma unroll
    for (int i = 0; i < 8; ++i)
#pragma unroll
      for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

    for (int t = un.t0 + warp; t < un.t1; t += C::CW) {
      const int slot = consumed % C::SLOTS;
      const int idx = warp * C::SLOTS + slot;
      mbar_wait(&full[idx], (consumed / C::SLOTS) & 1);
      const uint32_t tile = smem_u32(tiles + idx * C::TILE_BYTES);
      const uint32_t szb = smem_u32(szs + idx * C::SZ_BYTES) + g * 64;
      // scale / (128+zero) pairs for this thread's 16 columns, duplicated to both halves
      uint32_t s2[16], z2[16];
#pragma unroll
      for (int q = 0; q < 4; ++q) {
        const uint4 v = lds128(szb + q * 16);
        s2[q * 4 + 0] = prmt(v.x, v.x, 0x1010); z2[q * 4 + 0] = prmt(v.x, v.x, 0x3232);
        s2[q * 4 + 1] = prmt(v.y, v.y, 0x1010); z2[q * 4 + 1] = prmt(v.y, v.y, 0x3232);
        s2[q * 4 + 2] = prmt(v.z, v.z, 0x1010); z2[q * 4 + 2] = prmt(v.z, v.z, 0x3232);
        s2[q * 4 + 3] = prmt(v.w, v.w, 0x1010); z2[q * 4 + 3] = prmt(v.w, v.w, 0x3232);
      }
      const uint32_t xrow = smem_u32(xs) + (uint32_t)(((size_t)g * p.xs_stride + (size_t)(t - un.t0) * C::TILE_K + 4 * tig) * 2);
#pragma unroll
      for (int s = 0; s < C::STEPS; ++s) {
        const int rA = 8 * s + 2 * tig, rB = rA + 1;
        const uint4 A = lds128(tile + rA * CHUNK + ((g ^ (rA & 7)) << 4));
        const uint4 B = lds128(tile + rB * CHUNK + ((g ^ (rB & 7)) << 4));
        uint2 xb = make_uint2(0u, 0u);
        if (g < p.M) xb = lds64(xrow + s * 32);
        const uint32_t Aw[4] = {A.x, A.y, A.z, A.w};
        const uint32_t Bw[4] = {B.x, B.y, B.z, B.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#pragma unroll
          for (int h = 0; h < 2; ++h) {
            // bytes [A.b(2h), A.b(2h+1), B.b(2h), B.b(2h+1)]: columns 4j+2h and 4j+2h+1
            const uint32_t uu = prmt(Aw[j], Bw[j], h == 0 ? 0x5410u : 0x7632u);
            const int c0 = 4 * j + 2 * h, c1 = c0 + 1;
            uint32_t a0 = lop_and_or(uu, 0x000f000fu, 0x43004300u);        // col c0, k (2rA, 2rB)
            uint32_t a2 = lop_and_or(uu >> 4, 0x000f000fu, 0x43004300u);   // col c0, k (2rA+1, 2rB+1)
            uint32_t a1 = lop_and_or(uu >> 8, 0x000f000fu, 0x43004300u);   // col c1, k (2rA, 2rB)
            uint32_t a3 = lop_and_or(uu >> 12, 0x000f000fu, 0x43004300u);  // col c1, k (2rA+1, 2rB+1)
            a0 = bf2_mul(bf2_sub(a0, z2[c0]), s2[c0]);
            a2 = bf2_mul(bf2_sub(a2, z2[c0]), s2[c0]);
            a1 = bf2_mul(bf2_sub(a1, z2[c1]), s2[c1]);
            a3 = bf2_mul(bf2_sub(a3, z2[c1]), s2[c1]);
            mma_bf16_16816(acc[2 * j + h], a0, a1, a2, a3, xb.x, xb.y);
          }
        }
      }
      __syncwarp();
      if (lane == 0) mbar_arrive(&empty[idx]);
      ++consumed;
    }

    // ---- cross-warp reduction (fixed order).  acc[i]: c0,c1 -> column 16g+2i (batch 2tig, 2tig+1);
    //      c2,c3 -> column 16g+2i+1.
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      const int ca = 16 * g + 2 * i, cb = ca + 1;
      const int m0 = 2 * tig, m1 = m0 + 1;
      if (m0 < p.M) {
        red[(warp * p.M + m0) * CHUNK + ca] = acc[i][0];
        red[(warp * p.M + m0) * CHUNK + cb] = acc[i][2];
      }
      if (m1 < p.M) {
        red[(warp * p.M + m1) * CHUNK + ca] = acc[i][1];
        red[(warp * p.M + m1) * CHUNK + cb] = acc[i][3];
      }
    }
    consumer_sync(C::NCONS);
    const int total = p.nmat * p.ksplit;
    bool do_epilogue = true;
    for (int e = ctid; e < p.M * CHUNK; e += C::NCONS) {
      float v = 0.f;
#pragma unroll
      for (int w = 0; w < C::CW; ++w) v += red[w * p.M * CHUNK + e];
      if (total > 1) {
        const int m = e / CHUNK, c = e % CHUNK;
        if (n0 + c < p.N)
          p.partials[(((size_t)un.split * p.nmat + un.mat) * p.M + m) * p.N + n0 + c] = v;
      } else {
        outs[e] = v;
      }
    }
    if (total > 1) {
      __threadfence();
      consumer_sync(C::NCONS);
      if (ctid == 0) {
        const int old = atomicAdd(&p.counters[un.chunk], 1);
        s_flag = (old == total - 1);
      }
      consumer_sync(C::NCONS);
      do_epilogue = s_flag != 0;
      if (do_epilogue) {
        __threadfence();
        for (int e = ctid; e < p.nmat * p.M * CHUNK; e += C::NCONS) {
          const int mat = e / (p.M * CHUNK), m = (e / CHUNK) % p.M, c = e % CHUNK;
          float v = 0.f;
          if (n0 + c < p.N)
            for (int sp = 0; sp < p.ksplit; ++sp)
              v += __ldcg(&p.partials[(((size_t)sp * p.nmat + mat) * p.M + m) * p.N + n0 + c]);
          outs[e] = v;
        }
        if (ctid == 0) p.counters[un.chunk] = 0;
      }
    }
    consumer_sync(C::NCONS);

    // ---- fused epilogue on outs[mat][m][c]
    if (do_epilogue) {
      if (p.epilogue == LL_EPI_STORE) {
        for (int e = ctid; e < p.M * CHUNK; e += C::NCONS) {
          const int m = e / CHUNK, n = n0 + e % CHUNK;
          if (n < p.N) {
            float v = outs[e];
            if (p.bias) v += __bfloat162float(p.bias[n]);
            p.y[(size_t)m * p.N + n] = __float2bfloat16_rn(v);
          }
        }
      } else if (p.epilogue == LL_EPI_RESIDUAL) {
        // x = bf16(x + bf16(h)); partial sum of bf16(x*x) over this chunk for the next RMSNorm
        for (int m = warp; m < p.M; m += C::CW) {
          float ss = 0.f;
#pragma unroll
          for (int q = 0; q < CHUNK / 32; ++q) {
            const int c = lane + 32 * q, n = n0 + c;
            if (n < p.N) {
              const float h = bf16r(outs[m * CHUNK + c]);
              const float xn = bf16r(__fadd_rn(__bfloat162float(p.y[(size_t)m * p.N + n]), h));
              p.y[(size_t)m * p.N + n] = __float2bfloat16_rn(xn);
              ss += bf16r(__fmul_rn(xn, xn));
            }
          }
          ss = warp_sum(ss);
          if (lane == 0 && p.ssq_out) p.ssq_out[m * p.nchunks + un.chunk] = ss;
        }
      } else if (p.epilogue == LL_EPI_ROPE_KV) {
        const int E = p.N / 3, hs = p.head_size;
        const int rope_pos = p.pos[0], slot = p.pos[1];
        for (int e = ctid; e < p.M * (CHUNK / 2); e += C::NCONS) {
          const int m = e / (CHUNK / 2), c = 2 * (e % (CHUNK / 2)), n = n0 + c;
          if (n < p.N) {
            const int part = n / E, ee = n % E, head = ee / hs, d = ee % hs;
            float x0 = bf16r(outs[m * CHUNK + c]), x1 = bf16r(outs[m * CHUNK + c + 1]);
            if (part < 2) {
              const float2 cs = *reinterpret_cast<const float2*>(p.rope + ((size_t)rope_pos * (hs / 2) + d / 2) * 2);
              const float o0 = __fsub_rn(__fmul_rn(x0, cs.x), __fmul_rn(x1, cs.y));
              const float o1 = __fadd_rn(__fmul_rn(x1, cs.x), __fmul_rn(x0, cs.y));
              x0 = o0; x1 = o1;
            }
            const __nv_bfloat162 o = __floats2bfloat162_rn(x0, x1);
            if (part == 0) {
              *reinterpret_cast<__nv_bfloat162*>(p.y + (size_t)m * E + ee) = o;
            } else {
              __nv_bfloat16* cache = part == 1 ? p.k_cache : p.v_cache;
              *reinterpret_cast<__nv_bfloat162*>(cache + (((size_t)m * p.n_head + head) * p.cache_len + slot) * hs + d) = o;
            }
          }
        }
      } else {  // LL_EPI_SWIGLU
        for (int e = ctid; e < p.M * CHUNK; e += C::NCONS) {
          const int m = e / CHUNK, n = n0 + e % CHUNK;
          if (n < p.N) {
            const float a = bf16r(outs[e]), b = bf16r(outs[p.M * CHUNK + e]);
            const float s = bf16r(a / (1.0f + expf(-a)));
            p.y[(size_t)m * p.N + n] = __float2bfloat16_rn(__fmul_rn(s, b));
          }
        }
      }
    }
  }
}

// ------------------------------------------------------------------------------------ host side
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

struct MapKey {
  const void* ptr;
  int N, rows, tile_rows;
  bool operator==(const MapKey& o) const { return ptr == o.ptr && N == o.N && rows == o.rows && tile_rows == o.tile_rows; }
};
struct MapKeyHash {
  size_t operator()(const MapKey& k) const {
    return std::hash<const void*>()(k.ptr) ^ (size_t)k.N * 1000003u ^ (size_t)k.rows * 10007u ^ (size_t)k.tile_rows;
  }
};

int get_map(const uint8_t* qw, int N, int rows, int tile_rows, CUtensorMap* out) {
  static std::mutex mu;
  static std::unordered_map<MapKey, CUtensorMap, MapKeyHash> cache;
  MapKey key{qw, N, rows, tile_rows};
  std::lock_guard<std::mutex> lk(mu);
  auto it = cache.find(key);
  if (it != cache.end()) {
    *out = it->second;
    return LL_OK;
  }
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    set_error("cuTensorMapEncodeTiled entry point not available");
    return LL_ERR_CUDA;
  }
  cuuint64_t dims[2] = {(cuuint64_t)N, (cuuint64_t)rows};
  cuuint64_t strides[1] = {(cuuint64_t)N};
  cuuint32_t box[2] = {(cuuint32_t)CHUNK, (cuuint32_t)tile_rows};
  cuuint32_t estr[2] = {1, 1};
  CUtensorMap m;
  CUresult r = fn(&m, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void*)qw, dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    set_error("cuTensorMapEncodeTiled failed (%d) for qw=%p N=%d rows=%d", (int)r, (const void*)qw, N, rows);
    return LL_ERR_CUDA;
  }
  if (cache.size() > 4096) cache.clear();
  cache.emplace(key, m);
  *out = m;
  return LL_OK;
}

int pick_tile_rows(int group) {
  if (group % 128 == 0) return 64;
  if (group % 64 == 0) return 32;
  if (group % 32 == 0) return 16;
  return 0;
}

constexpr int MAX_KSPLIT = 32;
constexpr int XS_BUDGET = 24 * 1024;  // bytes of shared memory for staged x per CTA

// K split: fill ~2 CTAs per SM, keep the staged x slice within budget, minimise the makespan.
int choose_ksplit(int N, int K, int group, int nmat, int M, int tile_rows, int sms) {
  const int nchunks = (N + CHUNK - 1) / CHUNK;
  const int ktiles = (K / 2 + tile_rows - 1) / tile_rows;
  const int slots = sms * 2;
  const int tile_k = 2 * tile_rows;
  int min_split = 1;
  while ((long)((ktiles + min_split - 1) / min_split) * tile_k * M * 2 > XS_BUDGET && min_split < ktiles) ++min_split;
  int best = min_split;
  double best_cost = 1e30;
  for (int ks = min_split; ks <= ktiles && ks <= MAX_KSPLIT; ++ks) {
    const int tps = (ktiles + ks - 1) / ks;
    const int ks_eff = (ktiles + tps - 1) / tps;
    if (ks_eff != ks) continue;
    const long units = (long)nchunks * nmat * ks;
    const long waves = (units + slots - 1) / slots;
    // per unit: tiles handled 4 warps wide + fixed cost (stage x, reduce, epilogue, split-K fixup)
    const double per_unit = (double)((tps + 3) / 4) * (tile_rows / 64.0) + 0.75 + (ks > 1 ? 0.25 : 0.0);
    const double cost = waves * per_unit;
    if (cost < best_cost - 1e-9) {
      best_cost = cost;
      best = ks;
    }
  }
  return best;
}

template <class C>
int launch_cfg(Params& p, const ll_gemv_args* a, cudaStream_t st) {
  const size_t smem = (size_t)C::CW * C::SLOTS * (C::TILE_BYTES + C::SZ_BYTES) + ((p.M * p.xs_stride * 2 + 15) & ~15) +
                      (size_t)(C::CW + 2) * p.M * CHUNK * 4 + (size_t)2 * C::CW * C::SLOTS * 8;
  static size_t configured = 0;
  if (smem > configured) {
    LL_CUDA(cudaFuncSetAttribute(gemv_int4_kernel<C>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  const int sms = sm_count();
  const int grid = p.nunits < 2 * sms ? p.nunits : 2 * sms;
  LL_CUDA(launch(gemv_int4_kernel<C>, dim3(grid), dim3(C::THREADS), smem, st, (a->flags & LL_GEMV_PDL) != 0, p));
  return LL_OK;
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

int ll_int4_gemv_plan(int N, int K, int group, int nmat, int M, int32_t* ksplit, size_t* partial_bytes,
                      size_t* counter_bytes) {
  LL_CHECK_ARG(N > 0 && K > 0 && group > 0 && (nmat == 1 || nmat == 2) && M >= 1 && M <= MAXM, "bad plan arguments");
  const int tr = pick_tile_rows(group);
  LL_CHECK_ARG(tr != 0, "group size %d is not a multiple of 32", group);
  const int sms = sm_count();
  if (sms < 0) {
    set_error("no CUDA device");
    return LL_ERR_CUDA;
  }
  const int ks = choose_ksplit(N, K, group, nmat, M, tr, sms);
  if (ksplit) *ksplit = ks;
  if (partial_bytes) *partial_bytes = (size_t)ks * nmat * M * N * sizeof(float);
  if (counter_bytes) *counter_bytes = (size_t)((N + CHUNK - 1) / CHUNK) * sizeof(int32_t);
  return LL_OK;
}

int ll_int4_gemv(const ll_gemv_args* a, ll_stream stream) {
  LL_CHECK_ARG(a != nullptr, "null args");
  LL_CHECK_ARG(a->qw && a->sz && a->x && a->y, "null pointer in gemv args");
  LL_CHECK_ARG(a->M >= 1 && a->M <= MAXM, "M must be in 1..8 (got %d)", a->M);
  LL_CHECK_ARG(a->N > 0 && a->N % 16 == 0, "N must be a positive multiple of 16 (got %d)", a->N);
  LL_CHECK_ARG(a->K > 0 && a->K % 8 == 0, "K must be a positive multiple of 8 (got %d)", a->K);
  LL_CHECK_ARG(a->group > 0 && (a->group == a->K || a->K % a->group == 0 || true), "bad group");
  const int tr = pick_tile_rows(a->group);
  LL_CHECK_ARG(tr != 0, "group size %d is not a multiple of 32", a->group);
  LL_CHECK_ARG(((uintptr_t)a->qw & 15) == 0 && ((uintptr_t)a->sz & 15) == 0 && ((uintptr_t)a->x & 7) == 0,
               "qw/sz must be 16-byte aligned, x 8-byte aligned");
  const int nmat = a->epilogue == LL_EPI_SWIGLU ? 2 : 1;
  if (nmat == 2) LL_CHECK_ARG(a->qw2 && a->sz2, "SWIGLU epilogue needs qw2/sz2");
  if (a->norm_weight) LL_CHECK_ARG(a->ssq && a->ssq_parts > 0, "fused RMSNorm needs ssq partials");
  if (a->epilogue == LL_EPI_ROPE_KV) {
    LL_CHECK_ARG(a->rope && a->pos && a->k_cache && a->v_cache, "ROPE_KV epilogue needs rope/pos/caches");
    LL_CHECK_ARG(a->N % 3 == 0 && (a->N / 3) % 2 == 0 && a->head_size % 2 == 0 && a->n_head * a->head_size == a->N / 3,
                 "ROPE_KV: N must be 3*n_head*head_size");
  }
  LL_CHECK_ARG(a->epilogue >= 0 && a->epilogue <= 3, "bad epilogue %d", a->epilogue);

  Params p;
  memset(&p, 0, sizeof(p));
  const int rows = a->K / 2;
  if (int r = get_map(a->qw, a->N, rows, tr, &p.map[0])) return r;
  if (nmat == 2) {
    if (int r = get_map(a->qw2, a->N, rows, tr, &p.map[1])) return r;
  }
  p.sz[0] = a->sz; p.sz[1] = a->sz2;
  p.nmat = nmat; p.N = a->N; p.K = a->K; p.group = a->group;
  p.G = (a->K + a->group - 1) / a->group;
  p.M = a->M;
  p.x = (const __nv_bfloat16*)a->x;
  p.norm_w = (const __nv_bfloat16*)a->norm_weight;
  p.ssq = a->ssq; p.ssq_parts = a->ssq_parts; p.eps = a->eps;
  p.epilogue = a->epilogue;
  p.y = (__nv_bfloat16*)a->y;
  p.bias = (const __nv_bfloat16*)a->bias;
  p.ssq_out = a->ssq_out;
  p.rope = a->rope; p.pos = a->pos;
  p.k_cache = (__nv_bfloat16*)a->k_cache; p.v_cache = (__nv_bfloat16*)a->v_cache;
  p.n_head = a->n_head; p.head_size = a->head_size; p.cache_len = a->cache_len;
  p.partials = a->partials; p.counters = a->counters;
  p.nchunks = (a->N + CHUNK - 1) / CHUNK;
  p.ktiles = (rows + tr - 1) / tr;
  const int sms = sm_count();
  int ks = a->ksplit > 0 ? a->ksplit : choose_ksplit(a->N, a->K, a->group, nmat, a->M, tr, sms);
  if (ks > p.ktiles) ks = p.ktiles;
  p.tiles_per_split = (p.ktiles + ks - 1) / ks;
  ks = (p.ktiles + p.tiles_per_split - 1) / p.tiles_per_split;
  p.ksplit = ks;
  LL_CHECK_ARG((long)p.tiles_per_split * 2 * tr * a->M * 2 <= 64 * 1024, "K slice too large for shared memory; raise ksplit");
  LL_CHECK_ARG(ks <= MAX_KSPLIT, "ksplit must be <= %d", MAX_KSPLIT);
  if (nmat * ks > 1) LL_CHECK_ARG(a->partials && a->counters, "split-K / two-matrix launch needs partials and counters scratch");
  p.nunits = p.nchunks * nmat * ks;
  p.xs_stride = p.tiles_per_split * 2 * tr + 8;

  cudaStream_t st = (cudaStream_t)stream;
  if (tr == 64) return launch_cfg<Cfg<4, 2, 64>>(p, a, st);
  if (tr == 32) return launch_cfg<Cfg<4, 2, 32>>(p, a, st);
  return launch_cfg<Cfg<4, 2, 16>>(p, a, st);
}

}  // extern "C"
