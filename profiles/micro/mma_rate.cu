// Issue-rate microbenchmark for the legacy warp-level MMA / dot instructions on sm_100a.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o mma_rate mma_rate.cu ; ./mma_rate
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>

template <int KIND>
__global__ void rate_kernel(float* out, long long* cycles, int iters) {
  float c[4][4];
  int ci[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) { c[i][j] = 0.f; ci[i][j] = 0; }
  uint32_t a0 = threadIdx.x * 0x01010101u, a1 = a0 ^ 0x3c003c00u, a2 = a0 + 7, a3 = a1 + 9, b0 = 0x3c003c00u, b1 = 0x38003800u;
  __syncthreads();
  long long t0 = clock64();
  for (int it = 0; it < iters; ++it) {
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      if (KIND == 0)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
      else if (KIND == 1)
        asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
      else if (KIND == 2)
        asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+r"(ci[i][0]), "+r"(ci[i][1]), "+r"(ci[i][2]), "+r"(ci[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
      else if (KIND == 3)
        asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                     : "+f"(c[i][0]), "+f"(c[i][1]), "+f"(c[i][2]), "+f"(c[i][3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
      else if (KIND == 4) {  // 4 x dp4a (u8 x s8) per "op group"
#pragma unroll
        for (int j = 0; j < 4; ++j) asm volatile("dp4a.u32.s32 %0, %1, %2, %0;" : "+r"(ci[i][j]) : "r"(a0 + j), "r"(b0));
      } else if (KIND == 5) {  // 4 x FFMA
#pragma unroll
        for (int j = 0; j < 4; ++j) asm volatile("fma.rn.f32 %0, %1, %2, %0;" : "+f"(c[i][j]) : "f"(__uint_as_float(a1 + j)), "f"(__uint_as_float(b0)));
      } else if (KIND == 6) {  // 4 x LOP3
#pragma unroll
        for (int j = 0; j < 4; ++j) asm volatile("lop3.b32 %0, %0, %1, %2, 0xea;" : "+r"(ci[i][j]) : "r"(a0 + j), "r"(b0));
      }
    }
  }
  long long t1 = clock64();
  float s = 0.f;
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) s += c[i][j] + (float)ci[i][j];
  out[blockIdx.x * blockDim.x + threadIdx.x] = s;
  if (threadIdx.x == 0) cycles[blockIdx.x] = t1 - t0;
}

template <int KIND>
void run(const char* name, int opsPerGroup) {
  float* out; long long* cyc;
  cudaMalloc(&out, 148 * 1024 * 4); cudaMalloc(&cyc, 148 * 8);
  const int iters = 2000;
  for (int warps : {4, 8, 16, 32}) {
    rate_kernel<KIND><<<148, warps * 32>>>(out, cyc, iters);
    cudaDeviceSynchronize();
    long long h[148]; cudaMemcpy(h, cyc, sizeof(h), cudaMemcpyDeviceToHost);
    double avg = 0; for (int i = 0; i < 148; ++i) avg += h[i]; avg /= 148;
    // per SMSP: warps/4 warps, each issuing iters*4*opsPerGroup instructions
    double per = avg / ((double)iters * 4 * opsPerGroup * (warps / 4.0));
    printf("%-28s warps/SM=%2d  cycles=%9.0f  clk per warp-instr per SMSP = %6.2f\n", name, warps, avg, per);
  }
  cudaError_t e = cudaGetLastError(); if (e != cudaSuccess) printf("  error: %s\n", cudaGetErrorString(e));
  cudaFree(out); cudaFree(cyc);
}

int main() {
  run<0>("mma m16n8k16 bf16->f32", 1);
  run<3>("mma m16n8k16 f16->f32", 1);
  run<1>("mma m16n8k32 e4m3->f32", 1);
  run<2>("mma m16n8k32 u8.s8->s32", 1);
  run<4>("dp4a.u32.s32", 4);
  run<5>("fma.rn.f32", 4);
  run<6>("lop3", 4);
  return 0;
}
