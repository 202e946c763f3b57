// Micro-test for the planned tcgen05 form of the int4 GEMV: the A operand of
// tcgen05.mma comes from TENSOR MEMORY, written there by the de-quantising warps with tcgen05.st, the B operand
// (activations) is a no-swizzle K-major tile in shared memory, the accumulator is read back with tcgen05.ld.
// It answers, in one short GPU call, the three things the design rests on:
//   1. the TMEM layout of a bf16 A operand at M = 128 (assumed: row i in lane i, 32-bit column j = k 2j | k 2j+1 << 16),
//   2. which N is accepted at M = 128 (16 by the ISA table; 8 tried when argv[1] = 8) and that the second 8-row group of B
//      may alias the first (SBO = 0), so that a batch-1 x row costs 8, not 16, rows of shared memory,
//   3. the cost per 128 x 16 slice of `tcgen05.st x8 + tcgen05.mma` issued back to back (clock64 around a loop).
// NOT yet run (written when the round's GPU budget was spent).  Build / run:
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tmem_a_mma tmem_a_mma.cu && ./tmem_a_mma [N=16] [sbo_bytes=0]
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_bf16.h>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

// no-swizzle K-major shared-memory descriptor: core matrix = 8 rows x 16 bytes, contiguous (row r at +16 r);
// LBO = bytes between core matrices adjacent in K, SBO = bytes between 8-row groups (bit layout as in
// csrc/gemm_bf16_tc.cu: address >> 4 at bits 0-13, LBO >> 4 at 16-29, SBO >> 4 at 32-45, version 1 at bit 46,
// layout type 0 = no swizzle at bits 61-63)
__device__ __forceinline__ uint64_t smem_desc_noswizzle(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3fff);
  d |= (uint64_t)((lbo >> 4) & 0x3fff) << 16;
  d |= (uint64_t)((sbo >> 4) & 0x3fff) << 32;
  d |= (uint64_t)1 << 46;
  return d;
}
// kind::f16 instruction descriptor: D = F32, A = B = BF16, both K-major, N >> 3 at bits 17-22, M >> 4 at bits 24-28
__host__ __device__ constexpr uint32_t instr_desc_bf16(int M, int N) {
  return (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}

__device__ __forceinline__ void tmem_st_x8(uint32_t taddr, const uint32_t (&r)[8]) {
  asm volatile("tcgen05.st.sync.aligned.32x32b.x8.b32 [%0], {%1,%2,%3,%4,%5,%6,%7,%8};" ::"r"(taddr), "r"(r[0]), "r"(r[1]),
               "r"(r[2]), "r"(r[3]), "r"(r[4]), "r"(r[5]), "r"(r[6]), "r"(r[7])
               : "memory");
}
__device__ __forceinline__ void tmem_ld_x16(uint32_t taddr, uint32_t (&v)[16]) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15}, [%16];"
      : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
        "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
// A from tensor memory ([%1] is a TMEM address), B from a shared-memory descriptor
__device__ __forceinline__ void umma_ts_bf16(uint32_t tmem_d, uint32_t tmem_a, uint64_t bdesc, uint32_t idesc, uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, p;\n\t}" ::"r"(tmem_d),
      "r"(tmem_a), "l"(bdesc), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  asm volatile(
      "{\n\t.reg .pred p;\n"
      "W_%=:\n\t"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
      "@p bra D_%=;\n\t"
      "bra W_%=;\n"
      "D_%=:\n\t}" ::"r"(smem_u32(bar)),
      "r"(parity)
      : "memory");
}

// host and device agree on the test data: small integers, exact in bf16 and in the fp32 sums
__host__ __device__ inline int a_val(int i, int k) { return (i * 3 + k * 5) % 17 - 8; }
__host__ __device__ inline int b_val(int n, int k) { return (n * 7 + k * 3) % 13 - 6; }

__global__ void __launch_bounds__(128, 1) tmem_a_kernel(float* out /*[128][16]*/, long long* cyc /*[2]*/, int N, uint32_t sbo,
                                                        int iters) {
  __shared__ __align__(128) uint8_t bs[512];  // B: group gi (8 rows) at +256 gi, k core matrix j at +128 j, row r at +16 r
  __shared__ __align__(8) uint64_t bar;
  __shared__ uint32_t tmem_base_s;
  const int tid = threadIdx.x, warp = tid >> 5;

  for (int e = tid; e < 16 * 16; e += 128) {
    const int n = e >> 4, k = e & 15;
    const __nv_bfloat16 v = __float2bfloat16_rn((float)b_val(n, k));
    *reinterpret_cast<__nv_bfloat16*>(bs + (n >> 3) * 256 + (k >> 3) * 128 + (n & 7) * 16 + (k & 7) * 2) = v;
  }
  if (tid == 0) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(smem_u32(&bar)));
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");  // B was written through the generic proxy
  if (warp == 0) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&tmem_base_s)), "r"(64u) : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  const uint32_t tbase = tmem_base_s;
  const uint32_t t_a = tbase + ((uint32_t)(32 * warp) << 16);        // this warp's lane quarter, columns 0-7: A
  const uint32_t t_d = tbase + ((uint32_t)(32 * warp) << 16) + 32;   // columns 32..32+N: D

  // ---- 1. A row `tid` -> TMEM lane `tid`: column j = (k = 2j) | (k = 2j+1) << 16
  uint32_t r[8];
#pragma unroll
  for (int j = 0; j < 8; ++j) {
    const uint32_t lo = __bfloat16_as_ushort(__float2bfloat16_rn((float)a_val(tid, 2 * j)));
    const uint32_t hi = __bfloat16_as_ushort(__float2bfloat16_rn((float)a_val(tid, 2 * j + 1)));
    r[j] = lo | (hi << 16);
  }
  tmem_st_x8(t_a, r);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();

  // ---- 2. one MMA: D[128][N] = A[128][16] . B[N][16]^T
  const uint32_t idesc = instr_desc_bf16(128, N);
  const uint64_t bdesc = smem_desc_noswizzle(smem_u32(bs), 128, sbo);
  if (tid == 0) {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    umma_ts_bf16(tbase + 32, tbase, bdesc, idesc, 0);
    umma_commit(&bar);
  }
  mbar_wait(&bar, 0);
  asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
  uint32_t v[16];
  tmem_ld_x16(t_d, v);
#pragma unroll
  for (int j = 0; j < 16; ++j) out[tid * 16 + j] = __uint_as_float(v[j]);

  // ---- 3. rates: `iters` tcgen05.st x8 per warp back to back; then `iters` MMAs from one thread and one commit
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) tmem_st_x8(t_a, r);
  asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
  long long t1 = clock64();
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (tid == 0) {
    cyc[0] = t1 - t0;
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    t0 = clock64();
    for (int it = 0; it < iters; ++it) umma_ts_bf16(tbase + 32, tbase, bdesc, idesc, 1);
    umma_commit(&bar);
    mbar_wait(&bar, 1);
    cyc[1] = clock64() - t0;
  }
  asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
  __syncthreads();
  if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tbase), "r"(64u) : "memory");
}

int main(int argc, char** argv) {
  const int N = argc > 1 ? atoi(argv[1]) : 16;
  const uint32_t sbo = argc > 2 ? (uint32_t)atoi(argv[2]) : 0u;
  const int iters = 1000;
  float* d_out;
  long long* d_cyc;
  cudaMalloc(&d_out, 128 * 16 * sizeof(float));
  cudaMalloc(&d_cyc, 2 * sizeof(long long));
  cudaMemset(d_out, 0, 128 * 16 * sizeof(float));
  tmem_a_kernel<<<1, 128>>>(d_out, d_cyc, N, sbo, iters);
  cudaError_t e = cudaDeviceSynchronize();
  if (e != cudaSuccess) {
    printf("N=%d sbo=%u: CUDA error: %s\n", N, sbo, cudaGetErrorString(e));
    return 1;
  }
  static float out[128 * 16];
  long long cyc[2];
  cudaMemcpy(out, d_out, sizeof(out), cudaMemcpyDeviceToHost);
  cudaMemcpy(cyc, d_cyc, sizeof(cyc), cudaMemcpyDeviceToHost);
  int bad = 0;
  for (int i = 0; i < 128; ++i)
    for (int n = 0; n < N; ++n) {
      const int nb = (sbo == 0 && n >= 8) ? n - 8 : n;  // SBO = 0: the second 8-row group of B aliases the first
      int ref = 0;
      for (int k = 0; k < 16; ++k) ref += a_val(i, k) * b_val(nb, k);
      if (out[i * 16 + n] != (float)ref && bad++ < 8) printf("mismatch D[%d][%d] = %g, expected %d\n", i, n, out[i * 16 + n], ref);
    }
  printf("N=%d sbo=%u: %s (%d mismatches); tcgen05.st x8: %.1f clk each per warp; tcgen05.mma 128x%dx16 from TMEM A: %.1f clk each\n",
         N, sbo, bad ? "LAYOUT ASSUMPTION WRONG" : "A-from-TMEM layout confirmed", bad, (double)cyc[0] / iters, N,
         (double)cyc[1] / iters);
  return bad ? 2 : 0;
}
