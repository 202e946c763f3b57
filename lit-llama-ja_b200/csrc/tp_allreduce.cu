// One-shot all-reduce over NVLink peer memory for the tensor-parallel decode step, fused with the residual add:
//   x[i] = bf16( x[i] + bf16( sum_r part_r[i] ) )        (r in rank order on every rank: identical bits everywhere)
// ref: none - the reference is single-device (generate.py:122); this is the one exchange step of the sharded path
// (SURVEY 8e): [B, n_embd] bf16 after each row-parallel GEMV (16 KB at batch 1 for 65B), latency-bound.  NCCL's
// all-reduce costs a launch + protocol round per call and does not sit well inside a captured graph next to PDL
// kernels; here every rank pushes nothing - it only raises a flag, then PULLS the peers' partial sums straight out of
// their HBM through the NVSwitch (P2P loads on symmetric-memory mappings), so one small kernel is the whole
// collective.  Double-buffered partials (the caller alternates buffers) make one flag exchange per call enough:
// a peer can only overwrite buffer b two calls later, after it has seen this rank's next flag, which this rank
// raises only when it has finished reading.
#include "common.cuh"

namespace llb {
namespace {

constexpr int MAX_WORLD = 16;

struct TpArgs {
  const __nv_bfloat16* part[MAX_WORLD];  // peer-mapped partial-sum buffers (entry `rank` is the local one)
  int32_t* flags[MAX_WORLD];             // peer-mapped flag arrays: flags[r][s] = last epoch rank s announced to rank r
};

__device__ __forceinline__ void st_release_sys(int32_t* p, int32_t v) {
  asm volatile("st.release.sys.global.s32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ int32_t ld_acquire_sys(const int32_t* p) {
  int32_t v;
  asm volatile("ld.acquire.sys.global.s32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ uint4 ld_relaxed_sys_v4(const void* p) {
  uint4 v;
  asm volatile("ld.relaxed.sys.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
  return v;
}

__global__ void __launch_bounds__(1024) tp_allreduce_add_kernel(const __grid_constant__ TpArgs a, __nv_bfloat16* __restrict__ x,
                                                                int32_t* __restrict__ epoch_counter, int rank, int world, int n,
                                                                int use_pdl) {
  __shared__ int s_epoch;
  // use_pdl (round 2): a link of the programmatic-launch chain.  Round 1 launched this kernel plainly because an early
  // launch behind it cost time - the cause was CTA placement (a 148-CTA GEMV grid released while a fat CTA holds an SM
  // doubles up elsewhere, profiles/r02_gemv_experiments.md section 6), cured by the EVEN / ODD GEMV footprints.  With them
  // the next GEMV's CTAs can start (barriers, norm weights, ring fill from HBM) while the flags and partial sums travel,
  // and this kernel's own launch latency hides behind the row-parallel GEMV.  Everything that touches data comes after
  // the wait.
  if (use_pdl) {
    pdl_launch_dependents();
    pdl_wait();
  }
  // every CTA counts its own calls (epoch_counter[blockIdx.x]): same call count on every rank and CTA -> same epoch
  if (threadIdx.x == 0) s_epoch = ++epoch_counter[blockIdx.x];
  // the local residual vector does not depend on the peers: fetched while the flags travel
  const int first = (blockIdx.x * blockDim.x + threadIdx.x) * 8, stride = gridDim.x * blockDim.x * 8;
  uint4 x_first = make_uint4(0u, 0u, 0u, 0u);
  if (first < n) x_first = *reinterpret_cast<const uint4*>(x + first);
  __syncthreads();
  const int epoch = s_epoch;
  // announce (first CTA): "my partial sums of this epoch are in my buffer" (written by the previous kernel of this stream)
  if (blockIdx.x == 0 && threadIdx.x < world) st_release_sys(a.flags[threadIdx.x] + rank, epoch);
  // wait for every rank's announcement
  if (threadIdx.x < world) {
    const int32_t* f = a.flags[rank] + threadIdx.x;
    while (ld_acquire_sys(f) < epoch) {
    }
  }
  __syncthreads();
  for (int i = first; i < n; i += stride) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    // all peers' vectors are requested before any is used: one NVLink round trip, not `world` of them
    for (int r0 = 0; r0 < world; r0 += 8) {
      uint4 v[8];
#pragma unroll
      for (int j = 0; j < 8; ++j) {
        v[j] = make_uint4(0u, 0u, 0u, 0u);
        if (r0 + j < world) v[j] = ld_relaxed_sys_v4(a.part[r0 + j] + i);
      }
#pragma unroll
      for (int j = 0; j < 8; ++j) {  // rank order; absent ranks add +0
        const uint32_t w[4] = {v[j].x, v[j].y, v[j].z, v[j].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          acc[2 * e] += bf16_bits_to_f32(w[e] & 0xffffu);
          acc[2 * e + 1] += bf16_bits_to_f32(w[e] >> 16);
        }
      }
    }
    const uint4 xv = i == first ? x_first : *reinterpret_cast<const uint4*>(x + i);
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float h0 = bf16r(acc[2 * e]), h1 = bf16r(acc[2 * e + 1]);
      const uint32_t lo = f32_to_bf16_bits(__fadd_rn(bf16_bits_to_f32(xw[e] & 0xffffu), h0));
      const uint32_t hi = f32_to_bf16_bits(__fadd_rn(bf16_bits_to_f32(xw[e] >> 16), h1));
      o[e] = lo | (hi << 16);
    }
    *reinterpret_cast<uint4*>(x + i) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// Second half of the PUSHED exchange: the row-parallel GEMV's epilogue (LL_EPI_PUSH, gemv_int4.cu) has already stored
// every rank's partial sums into every rank's receive buffer and counted its CTAs into cnt[j]; here one CTA waits for
// the expected number of arrivals and reduces from LOCAL memory - no flag round trip and no NVLink load in this
// kernel.  Buffers and counters alternate (j = call parity): arrivals of the next exchange from a fast peer land in the
// other counter and cannot be mistaken for this one's.
__global__ void __launch_bounds__(1024) tp_reduce_add_kernel(__nv_bfloat16* __restrict__ x, const __nv_bfloat16* __restrict__ recv,
                                                             uint32_t* __restrict__ cnt, int32_t* __restrict__ epoch_counter, int world,
                                                             int n, int arrivals) {
  __shared__ int s_epoch;
  if (threadIdx.x == 0) s_epoch = ++(*epoch_counter);
  __syncthreads();
  const int epoch = s_epoch, j = (epoch - 1) & 1;
  const uint32_t target = (uint32_t)((epoch + 1 - j) >> 1) * (uint32_t)arrivals;  // exchanges on buffer j so far
  if (threadIdx.x == 0) {
    const uint32_t* c = cnt + j;
    uint32_t v;
    do {
      asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(c) : "memory");
    } while ((int32_t)(v - target) < 0);
  }
  __syncthreads();
  uint4 x_first = make_uint4(0u, 0u, 0u, 0u);
  if (threadIdx.x * 8 < n) x_first = *reinterpret_cast<const uint4*>(x + threadIdx.x * 8);
  const __nv_bfloat16* base = recv + (size_t)j * world * n;
  for (int i = threadIdx.x * 8; i < n; i += blockDim.x * 8) {
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    for (int r0 = 0; r0 < world; r0 += 8) {
      uint4 v[8];
#pragma unroll
      for (int q = 0; q < 8; ++q) {  // L1 is not coherent with the peers' stores: system-scope loads (L2)
        v[q] = make_uint4(0u, 0u, 0u, 0u);
        if (r0 + q < world) v[q] = ld_relaxed_sys_v4(base + (size_t)(r0 + q) * n + i);
      }
#pragma unroll
      for (int q = 0; q < 8; ++q) {
        const uint32_t w[4] = {v[q].x, v[q].y, v[q].z, v[q].w};
#pragma unroll
        for (int e = 0; e < 4; ++e) {
          acc[2 * e] += bf16_bits_to_f32(w[e] & 0xffffu);
          acc[2 * e + 1] += bf16_bits_to_f32(w[e] >> 16);
        }
      }
    }
    const uint4 xv = i == threadIdx.x * 8 ? x_first : *reinterpret_cast<const uint4*>(x + i);
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float h0 = bf16r(acc[2 * e]), h1 = bf16r(acc[2 * e + 1]);
      const uint32_t lo = f32_to_bf16_bits(__fadd_rn(bf16_bits_to_f32(xw[e] & 0xffffu), h0));
      const uint32_t hi = f32_to_bf16_bits(__fadd_rn(bf16_bits_to_f32(xw[e] >> 16), h1));
      o[e] = lo | (hi << 16);
    }
    *reinterpret_cast<uint4*>(x + i) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// ---------------------------------------------------------------------------------------------------------------
// The exchange in ONE NVLink traversal ("LL": data and flag travel in the same 16-byte store, as in NCCL's low-latency
// protocol).  The pull kernel above costs three traversals per exchange (flag out, request out, data back); here every
// rank PUSHES its partial sums into every peer's receive buffer as {word, epoch, word, epoch} units - a 16-byte store
// carries two payload words and the epoch twice, so a unit whose two epochs both match has arrived whole, whatever the
// order the fabric delivered the stores in - and then reads the peers' units out of its OWN memory, spinning on the
// epochs.  No flag array, no fence, no release / acquire pair.  Receive buffers alternate with the epoch's parity: a
// peer can only send epoch e + 2 after it has received this rank's epoch e + 1, which this rank sends only after the
// kernel that consumed epoch e has completed (stream order), so a unit is never overwritten before it was read.
// Sums in rank order (fp32), one bf16 rounding, residual add: identical bits on every rank.
struct TpLLArgs {
  uint8_t* recv[MAX_WORLD];  // peer-mapped receive buffers: [2 parities][world senders][2 halves][n / 8 pieces][16 bytes]
};
__device__ __forceinline__ void st_ll(uint8_t* p, uint32_t a, uint32_t b, uint32_t flag) {
  asm volatile("st.volatile.global.v4.u32 [%0], {%1, %2, %3, %4};" ::"l"(p), "r"(a), "r"(flag), "r"(b), "r"(flag) : "memory");
}
__device__ __forceinline__ uint4 ld_ll(const uint8_t* p) {
  uint4 v;
  asm volatile("ld.volatile.global.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p) : "memory");
  return v;
}

template <int WORLD>  // 0: any world size (loop), else unrolled: all peers' units are requested before any is examined
__global__ void __launch_bounds__(256) tp_ll_allreduce_add_kernel(const __grid_constant__ TpLLArgs a,
                                                                  const __nv_bfloat16* __restrict__ part,
                                                                  __nv_bfloat16* __restrict__ x, int32_t* __restrict__ epoch_counter,
                                                                  int rank, int world_rt, int n, int use_pdl) {
  const int world = WORLD > 0 ? WORLD : world_rt;
  __shared__ int s_epoch;
  if (use_pdl) {
    pdl_launch_dependents();
    pdl_wait();
  }
  if (threadIdx.x == 0) s_epoch = ++epoch_counter[blockIdx.x];
  __syncthreads();
  const uint32_t epoch = (uint32_t)s_epoch;
  const int npieces = n >> 3;
  // a sender's lane = two arrays of npieces 16-byte units (first / second half of every piece): a warp's store instruction
  // then covers 512 contiguous bytes - whole 128-byte lines on the fabric instead of half-filled ones
  const size_t half_bytes = (size_t)npieces * 16, lane_bytes = 2 * half_bytes;
  const size_t half = (size_t)(epoch & 1u) * world * lane_bytes;
  for (int p = blockIdx.x * blockDim.x + threadIdx.x; p < npieces; p += gridDim.x * blockDim.x) {
    const uint4 mine = __ldcg(reinterpret_cast<const uint4*>(part) + p);
    const uint4 xv = *(reinterpret_cast<const uint4*>(x) + p);
    // send to every peer
    const size_t out_off = half + (size_t)rank * lane_bytes + (size_t)p * 16;
    for (int r = 0; r < world; ++r) {
      if (r == rank) continue;
      st_ll(a.recv[r] + out_off, mine.x, mine.y, epoch);
      st_ll(a.recv[r] + out_off + half_bytes, mine.z, mine.w, epoch);
    }
    // receive: spin until every peer's unit of this epoch is whole
    const uint8_t* in = a.recv[rank] + half + (size_t)p * 16;
    float acc[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    auto add = [&](uint32_t w0, uint32_t w1, uint32_t w2, uint32_t w3) {
      const uint32_t w[4] = {w0, w1, w2, w3};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        acc[2 * e] += bf16_bits_to_f32(w[e] & 0xffffu);
        acc[2 * e + 1] += bf16_bits_to_f32(w[e] >> 16);
      }
    };
    if constexpr (WORLD > 0) {
      uint4 lo[WORLD], hi[WORLD];
      bool all;
      do {
        all = true;
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
          if (r == rank) continue;
          lo[r] = ld_ll(in + (size_t)r * lane_bytes);
          hi[r] = ld_ll(in + (size_t)r * lane_bytes + half_bytes);
        }
#pragma unroll
        for (int r = 0; r < WORLD; ++r) {
          if (r == rank) continue;
          all = all && lo[r].y == epoch && lo[r].w == epoch && hi[r].y == epoch && hi[r].w == epoch;
        }
      } while (!all);
#pragma unroll
      for (int r = 0; r < WORLD; ++r) {  // rank order
        if (r == rank) add(mine.x, mine.y, mine.z, mine.w);
        else add(lo[r].x, lo[r].z, hi[r].x, hi[r].z);
      }
    } else {
      for (int r = 0; r < world; ++r) {
        if (r == rank) {
          add(mine.x, mine.y, mine.z, mine.w);
          continue;
        }
        uint4 lo, hi;
        do {
          lo = ld_ll(in + (size_t)r * lane_bytes);
          hi = ld_ll(in + (size_t)r * lane_bytes + half_bytes);
        } while (!(lo.y == epoch && lo.w == epoch && hi.y == epoch && hi.w == epoch));
        add(lo.x, lo.z, hi.x, hi.z);
      }
    }
    const uint32_t xw[4] = {xv.x, xv.y, xv.z, xv.w};
    uint32_t o[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float h0 = bf16r(acc[2 * e]), h1 = bf16r(acc[2 * e + 1]);
      const uint32_t l = f32_to_bf16_bits(__fadd_rn(bf16_bits_to_f32(xw[e] & 0xffffu), h0));
      const uint32_t h = f32_to_bf16_bits(__fadd_rn(bf16_bits_to_f32(xw[e] >> 16), h1));
      o[e] = l | (h << 16);
    }
    *(reinterpret_cast<uint4*>(x) + p) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

int ll_tp_allreduce_add(void* x, const uint64_t* peer_parts /* host: world pointers */, const uint64_t* peer_flags /* host */,
                        int32_t* epoch_counter, int rank, int world, int n, int flags, ll_stream stream) {
  LL_CHECK_ARG(x && peer_parts && peer_flags && epoch_counter, "null pointer");
  LL_CHECK_ARG(world >= 2 && world <= MAX_WORLD && rank >= 0 && rank < world, "bad rank/world (%d/%d)", rank, world);
  LL_CHECK_ARG(n > 0 && n % 8 == 0 && ((uintptr_t)x & 15) == 0, "n must be a multiple of 8 and x 16-byte aligned");
  TpArgs a;
  memset(&a, 0, sizeof(a));
  for (int r = 0; r < world; ++r) {
    LL_CHECK_ARG(peer_parts[r] != 0 && (peer_parts[r] & 15) == 0 && peer_flags[r] != 0, "bad peer pointer for rank %d", r);
    a.part[r] = (const __nv_bfloat16*)(uintptr_t)peer_parts[r];
    a.flags[r] = (int32_t*)(uintptr_t)peer_flags[r];
  }
  static bool once = false;
  if (!once) {
    prefer_max_smem(tp_allreduce_add_kernel);
    once = true;
  }
  const bool pdl = (flags & LL_GEMV_PDL) != 0;
  // as a link of the programmatic chain the CTAs are kept small (256 threads each, one 16-byte piece per thread, up to
  // LL_TP_MAX_CTAS of them: each must fit beside a resident GEMV CTA to start early, and every piece's peer loads must be
  // in flight at once - one NVLink round trip); launched plainly it is one CTA of 1024 threads
  int ctas = 1;
  if (pdl) {
    ctas = (n / 8 + 255) / 256;
    if (ctas > LL_TP_MAX_CTAS) ctas = LL_TP_MAX_CTAS;
  }
  LL_CUDA(launch(tp_allreduce_add_kernel, dim3(ctas), dim3(pdl ? 256 : 1024), 0, (cudaStream_t)stream, pdl, a, (__nv_bfloat16*)x,
                 epoch_counter, rank, world, n, pdl ? 1 : 0));
  return LL_OK;
}

size_t ll_tp_ll_recv_bytes(int world, int n) { return world > 0 && n > 0 ? (size_t)2 * world * (n / 8) * 32 : 0; }

int ll_tp_ll_allreduce_add(void* x, const void* part, const uint64_t* peer_recv /* host: world pointers */, int32_t* epoch_counter,
                           int rank, int world, int n, int flags, ll_stream stream) {
  LL_CHECK_ARG(x && part && peer_recv && epoch_counter, "null pointer");
  LL_CHECK_ARG(world >= 2 && world <= MAX_WORLD && rank >= 0 && rank < world, "bad rank/world (%d/%d)", rank, world);
  LL_CHECK_ARG(n > 0 && n % 8 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)part & 15) == 0, "n must be a multiple of 8, x / part 16-byte aligned");
  TpLLArgs a;
  memset(&a, 0, sizeof(a));
  for (int r = 0; r < world; ++r) {
    LL_CHECK_ARG(peer_recv[r] != 0 && (peer_recv[r] & 15) == 0, "bad receive buffer pointer for rank %d", r);
    a.recv[r] = (uint8_t*)(uintptr_t)peer_recv[r];
  }
  const bool pdl = (flags & LL_GEMV_PDL) != 0;
  int ctas = (n / 8 + 255) / 256;
  if (ctas > LL_TP_MAX_CTAS) ctas = LL_TP_MAX_CTAS;
  cudaStream_t st = (cudaStream_t)stream;
  const __nv_bfloat16* pp = (const __nv_bfloat16*)part;
  __nv_bfloat16* xx = (__nv_bfloat16*)x;
#define LL_TP_WORLD(W) \
  LL_CUDA(launch(tp_ll_allreduce_add_kernel<W>, dim3(ctas), dim3(256), 0, st, pdl, a, pp, xx, epoch_counter, rank, world, n, pdl ? 1 : 0))
  if (world == 2) LL_TP_WORLD(2);
  else if (world == 4) LL_TP_WORLD(4);
  else if (world == 8) LL_TP_WORLD(8);
  else LL_TP_WORLD(0);
#undef LL_TP_WORLD
  return LL_OK;
}

int ll_tp_reduce_add(void* x, const void* recv, uint32_t* cnt, int32_t* epoch_counter, int world, int n, int arrivals,
                     ll_stream stream) {
  LL_CHECK_ARG(x && recv && cnt && epoch_counter, "null pointer");
  LL_CHECK_ARG(world >= 1 && world <= MAX_WORLD && arrivals > 0, "bad world / arrivals (%d / %d)", world, arrivals);
  LL_CHECK_ARG(n > 0 && n % 8 == 0 && ((uintptr_t)x & 15) == 0 && ((uintptr_t)recv & 15) == 0, "n must be a multiple of 8, buffers 16-byte aligned");
  static bool once = false;
  if (!once) {
    prefer_max_smem(tp_reduce_add_kernel);
    once = true;
  }
  tp_reduce_add_kernel<<<1, 1024, 0, (cudaStream_t)stream>>>((__nv_bfloat16*)x, (const __nv_bfloat16*)recv, cnt, epoch_counter, world, n,
                                                            arrivals);
  LL_LAUNCH_OK("tp_reduce_add_kernel");
  return LL_OK;
}

}  // extern "C"
