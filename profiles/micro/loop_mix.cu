// Candidate int4-GEMV inner loops on sm_100a, isolated from HBM: every warp streams fragment words from a
// shared-memory buffer that is already resident and runs the de-quantise + multiply-accumulate mix of one design.
// Output: clk per 32-bit word (8 weights per lane, 256 per warp) per SM sub-partition, for several warp counts.
// At the measured HBM rate (6.55 TB/s / 148 SMs / 1.965 GHz = 22.5 B/clk/SM) the budget is 22.7 clk per word per
// sub-partition; a main loop that is to catch up after a kernel boundary wants <= 16.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o loop_mix loop_mix.cu ; ./loop_mix
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t lop_and_or(uint32_t a, uint32_t mask, uint32_t orv) {
  uint32_t r;
  asm("lop3.b32 %0, %1, %2, %3, 0xea;" : "=r"(r) : "r"(a), "r"(mask), "r"(orv));
  return r;
}
__device__ __forceinline__ uint32_t rotr(uint32_t a, uint32_t n) {
  uint32_t r;
  asm("shf.r.wrap.b32 %0, %1, %1, %2;" : "=r"(r) : "r"(a), "r"(n));
  return r;
}
__device__ __forceinline__ void fhfma_bf16(float& acc, uint32_t a, uint32_t b) {  // acc += a.lo*b.lo + a.hi*b.hi (two FHFMA)
  asm("{\n\t.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1;\n\tmov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.bf16 %0, al, bl, %0;\n\tfma.rn.f32.bf16 %0, ah, bh, %0;\n\t}"
      : "+f"(acc) : "r"(a), "r"(b));
}
__device__ __forceinline__ void fhfma_f16(float& acc, uint32_t a, uint32_t b) {
  asm("{\n\t.reg .b16 al, ah, bl, bh;\n\t"
      "mov.b32 {al, ah}, %1;\n\tmov.b32 {bl, bh}, %2;\n\t"
      "fma.rn.f32.f16 %0, al, bl, %0;\n\tfma.rn.f32.f16 %0, ah, bh, %0;\n\t}"
      : "+f"(acc) : "r"(a), "r"(b));
}
__device__ __forceinline__ void mma16816(float (&c)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3, uint32_t b0, uint32_t b1) {
  asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
               : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
__device__ __forceinline__ uint4 lds128(uint32_t addr) {
  uint4 v;
  asm volatile("ld.shared.v4.u32 {%0,%1,%2,%3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(addr));
  return v;
}

constexpr uint32_t M7 = 0x00780078u, G7 = 0x41804180u;    // bf16 16+q, nibble at bits 3..6
constexpr uint32_t MA = 0x003c003cu, GA = 0x5c005c00u;    // fp16 256+q, nibble at bits 2..5
constexpr uint32_t MB = 0x03c003c0u, GB = 0x4c004c00u;    // fp16 16+q,  nibble at bits 6..9

// KIND 0: today's loop: per 4 words 1 LDS.128 (weights) + 2 LDS.128 (B fragments of x) + 4 x (7 ALU + HMMA)
// KIND 1: x-stationary HMMA: B fragments stay in registers
// KIND 2: x-stationary SIMT, bf16: 4 x (7 ALU + 8 FHFMA.BF16)
// KIND 3: x-stationary SIMT, fp16 two-nibbles-per-mask: 4 x (5 ALU + 8 FHFMA.F16)
// KIND 4: x-stationary HMMA fp16 two-nibbles-per-mask: 4 x (5 ALU + HMMA.F16)
// KIND 5: like 2 plus the per-(column, group) fold every 2 words (LDS.32 scale pair, 2 ALU, 3 FFMA)
// KIND 6: like 3 plus the fold
template <int KIND>
__global__ void __launch_bounds__(1024, 1) mix_kernel(float* out, long long* cycles, int iters) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  uint32_t* s32 = reinterpret_cast<uint32_t*>(smem);
  for (int i = threadIdx.x; i < 16384; i += blockDim.x) s32[i] = (i * 2654435761u) ^ 0x9e3779b9u;
  __syncthreads();
  const uint32_t base = (uint32_t)__cvta_generic_to_shared(smem);
  const uint32_t wbase = base + (warp & 7) * 4096 + lane * 16;  // 4 KB window per warp: 8 x 512 B
  const uint32_t xbase = base + 32768 + ((lane >> 2) ? 0 : 2048) + (lane & 3) * 16;
  uint32_t xr[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) xr[i] = 0x3f803f80u + i * 0x00010001u + lane;
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  float outv = 0.f, sxg = 1.25f;
  __syncthreads();
  long long t0 = clock64();
#pragma unroll 2
  for (int it = 0; it < iters; ++it) {
    const uint4 wq = lds128(wbase + (it & 7) * 512);
    const uint32_t w[4] = {wq.x, wq.y, wq.z, wq.w};
    if (KIND == 0) {
      const uint4 xa = lds128(xbase + (it & 7) * 128), xb = lds128(xbase + (it & 7) * 128 + 64);
      const uint32_t b[8] = {xa.x, xa.y, xa.z, xa.w, xb.x, xb.y, xb.z, xb.w};
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t a0 = lop_and_or(w[i], M7, G7), a1 = lop_and_or(w[i] >> 4, M7, G7), a2 = lop_and_or(w[i] >> 8, M7, G7),
                       a3 = lop_and_or(rotr(w[i], 12), M7, G7);
        mma16816(acc[i], a0, a1, a2, a3, b[2 * i], b[2 * i + 1]);
      }
    } else if (KIND == 1) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t a0 = lop_and_or(w[i], M7, G7), a1 = lop_and_or(w[i] >> 4, M7, G7), a2 = lop_and_or(w[i] >> 8, M7, G7),
                       a3 = lop_and_or(rotr(w[i], 12), M7, G7);
        mma16816(acc[i], a0, a1, a2, a3, xr[2 * i], xr[2 * i + 1]);
      }
    } else if (KIND == 4) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t r8 = rotr(w[i], 8);
        const uint32_t a0 = lop_and_or(w[i], MA, GA), a1 = lop_and_or(w[i], MB, GB), a2 = lop_and_or(r8, MA, GA), a3 = lop_and_or(r8, MB, GB);
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(acc[i][0]), "+f"(acc[i][1]), "+f"(acc[i][2]), "+f"(acc[i][3])
                     : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(xr[2 * i]), "r"(xr[2 * i + 1]));
      }
    } else if (KIND == 2 || KIND == 5) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t a0 = lop_and_or(w[i], M7, G7), a1 = lop_and_or(w[i] >> 4, M7, G7), a2 = lop_and_or(w[i] >> 8, M7, G7),
                       a3 = lop_and_or(rotr(w[i], 12), M7, G7);
        const int xo = (i & 1) * 4;
        fhfma_bf16(acc[i][0], a0, xr[xo + 0]);
        fhfma_bf16(acc[i][1], a1, xr[xo + 1]);
        fhfma_bf16(acc[i][2], a2, xr[xo + 2]);
        fhfma_bf16(acc[i][3], a3, xr[xo + 3]);
      }
    } else if (KIND == 3 || KIND == 6) {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        const uint32_t r8 = rotr(w[i], 8);
        const uint32_t a0 = lop_and_or(w[i], MA, GA), a1 = lop_and_or(w[i], MB, GB), a2 = lop_and_or(r8, MA, GA), a3 = lop_and_or(r8, MB, GB);
        const int xo = (i & 1) * 4;
        fhfma_f16(acc[i][0], a0, xr[xo + 0]);
        fhfma_f16(acc[i][1], a1, xr[xo + 1]);
        fhfma_f16(acc[i][2], a2, xr[xo + 2]);
        fhfma_f16(acc[i][3], a3, xr[xo + 3]);
      }
    }
    if (KIND == 5 || KIND == 6) {  // two (tile, k-block) folds per 4 words
#pragma unroll
      for (int h = 0; h < 2; ++h) {
        uint32_t sz;
        asm volatile("ld.shared.u32 %0, [%1];" : "=r"(sz) : "r"(base + 40960 + ((it * 2 + h) & 63) * 16 + (lane & 3) * 4));
        const float s = __uint_as_float(sz << 16), zn = __uint_as_float(sz & 0xffff0000u);
        float t = (acc[2 * h][0] + acc[2 * h][1]) + (acc[2 * h][2] + acc[2 * h][3]);
        t += (acc[2 * h + 1][0] + acc[2 * h + 1][1]) + (acc[2 * h + 1][2] + acc[2 * h + 1][3]);
        outv = fmaf(s, fmaf(zn, sxg, t), outv);
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[2 * h][j] = acc[2 * h + 1][j] = 0.f;
      }
    }
  }
  long long t1 = clock64();
  float s = outv;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) s += acc[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

// single-instruction issue rates
template <int KIND>
__global__ void rate_kernel(float* out, long long* cycles, int iters) {
  float c[8];
  uint32_t r[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) { c[i] = 0.f; r[i] = threadIdx.x * 0x01010101u + i; }
  const uint32_t b = 0x3f803f80u + threadIdx.x;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      if (KIND == 0) asm volatile("{.reg .b16 l,h,bl,bh; mov.b32 {l,h}, %1; mov.b32 {bl,bh}, %2; fma.rn.f32.bf16 %0, l, bl, %0;}" : "+f"(c[i]) : "r"(r[i]), "r"(b));
      else if (KIND == 1) asm volatile("{.reg .b16 l,h,bl,bh; mov.b32 {l,h}, %1; mov.b32 {bl,bh}, %2; fma.rn.f32.bf16 %0, h, bh, %0;}" : "+f"(c[i]) : "r"(r[i]), "r"(b));
      else if (KIND == 2) asm volatile("shf.r.wrap.b32 %0, %0, %0, 12;" : "+r"(r[i]));
      else if (KIND == 3) asm volatile("mad.lo.u32 %0, %0, 16, %1;" : "+r"(r[i]) : "r"(b));
      else if (KIND == 4) asm volatile("fma.rn.bf16x2 %0, %0, %1, %1;" : "+r"(r[i]) : "r"(b));
      else if (KIND == 5) asm volatile("prmt.b32 %0, %0, %1, 0x5140;" : "+r"(r[i]) : "r"(b));
      else if (KIND == 6) asm volatile("{.reg .b16 l,h,bl,bh; mov.b32 {l,h}, %1; mov.b32 {bl,bh}, %2; fma.rn.f32.f16 %0, l, bl, %0;}" : "+f"(c[i]) : "r"(r[i]), "r"(b));
      else if (KIND == 7) asm volatile("lop3.b32 %0, %0, %1, %2, 0xea;" : "+r"(r[i]) : "r"(b), "r"(0x41804180u));
      else if (KIND == 8) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(c[i]) : "f"(__uint_as_float(r[i])), "f"(__uint_as_float(b)));
      else if (KIND == 9) asm volatile("sub.bf16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(b));
      else if (KIND == 10) asm volatile("cvt.rni.s32.f32 %0, %1;" : "=r"(r[i]) : "f"(c[i] + __uint_as_float(r[i])));
      else if (KIND == 11) asm volatile("cvt.rn.f32.s32 %0, %1;" : "=f"(c[i]) : "r"(r[i]));
      else if (KIND == 12) asm volatile("max.bf16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(b));
      else if (KIND == 13) asm volatile("mul.rn.bf16x2 %0, %0, %1;" : "+r"(r[i]) : "r"(b));
    }
  }
  long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 8; ++i) s += c[i] + __uint_as_float(r[i]);
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

static float* g_out;
static long long* g_cyc;

template <int KIND>
void run_rate(const char* name) {
  const int iters = 2000;
  for (int warps : {4, 8, 16}) {
    rate_kernel<KIND><<<148, warps * 32>>>(g_out, g_cyc, iters);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, g_cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    printf("%-26s warps/SM=%2d  clk per warp-instr per SMSP = %6.2f\n", name, warps, avg / ((double)iters * 8 * (warps / 4.0)));
  }
  cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) printf("  error: %s\n", cudaGetErrorString(e));
}

template <int KIND>
void run_mix(const char* name) {
  const int iters = 4000;
  cudaFuncSetAttribute(mix_kernel<KIND>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536);
  for (int warps : {4, 8, 12, 16, 24, 32}) {
    mix_kernel<KIND><<<148, warps * 32, 65536>>>(g_out, g_cyc, iters);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, g_cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    const double per_word = avg / ((double)iters * 4 * (warps / 4.0));
    printf("%-44s warps/SM=%2d  clk per word per SMSP = %6.2f  -> %5.2f TB/s at 1.965 GHz x 148 SMs\n", name, warps, per_word,
           128.0 / per_word * 4 * 148 * 1.965e9 / 1e12);
  }
  cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) printf("  error: %s\n", cudaGetErrorString(e));
}

int main() {
  cudaMalloc(&g_out, 148 * 1024 * 4); cudaMalloc(&g_cyc, 148 * 8);
  run_rate<0>("FHFMA.BF16 (lo halves)");
  run_rate<1>("FHFMA.BF16 (hi halves)");
  run_rate<6>("FHFMA.F16");
  run_rate<8>("FFMA");
  run_rate<4>("HFMA2.BF16");
  run_rate<9>("HSUB2.BF16");
  run_rate<3>("IMAD lo");
  run_rate<2>("SHF");
  run_rate<7>("LOP3");
  run_rate<5>("PRMT");
  run_rate<10>("F2I (+1 FADD)");
  run_rate<11>("I2F");
  run_rate<12>("HMNMX2.BF16");
  run_rate<13>("HMUL2.BF16");
  if (getenv("RATES_ONLY")) return 0;
  run_mix<0>("today: LDS w + LDS x frags + 7 ALU + HMMA");
  run_mix<1>("x-stationary: 7 ALU + HMMA");
  run_mix<4>("x-stationary fp16: 5 ALU + HMMA");
  run_mix<2>("x-stationary SIMT bf16: 7 ALU + 8 FHFMA");
  run_mix<3>("x-stationary SIMT fp16: 5 ALU + 8 FHFMA");
  run_mix<5>("SIMT bf16 + fold per 2 words");
  run_mix<6>("SIMT fp16 + fold per 2 words");
  return 0;
}
