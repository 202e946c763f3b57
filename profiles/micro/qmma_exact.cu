// Is the legacy FP8 warp MMA (mma.sync m16n8k32 e4m3 x e4m3 -> f32) exact enough to carry the int4 GEMV?
//   A = raw nibbles as e4m3 bytes 0x00..0x0F  (= q * 2^-9 exactly: the subnormals and the first binade are linear)
//   B = a bf16 activation split into three e4m3 "planes" y1 + y2 * 2^-8 + y3 * 2^-16 = x * S (exact for 22 binades)
// Checks (1) the plane split reconstructs x*S exactly, (2) the MMA sums of q*y over k = 128 equal the exact sums.
// nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o qmma_exact qmma_exact.cu ; ./qmma_exact
#include <cmath>
#include <cstdint>
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint16_t cvt_e4m3x2(float hi, float lo) {
  uint16_t r;
  asm("cvt.rn.satfinite.e4m3x2.f32 %0, %1, %2;" : "=h"(r) : "f"(hi), "f"(lo));
  return r;
}
__device__ __forceinline__ float2 e4m3x2_to_f32(uint16_t v) {  // .x = low byte
  uint32_t h2;
  asm("cvt.rn.f16x2.e4m3x2 %0, %1;" : "=r"(h2) : "h"(v));
  float lo, hi;
  asm("{.reg .b16 l, h; mov.b32 {l, h}, %2; cvt.f32.f16 %0, l; cvt.f32.f16 %1, h;}" : "=f"(lo), "=f"(hi) : "r"(h2));
  return make_float2(lo, hi);
}

// K = 128, one warp.  q[16][128] bytes, x[128] bf16 bits, planes out [3][128], D out [16][8], mode 0: one chain, 1: four chains
__global__ void qmma_kernel(const uint8_t* q, const uint16_t* x, float S, uint8_t* planes, float* D, int mode) {
  __shared__ __align__(16) uint8_t pl[8][128];
  const int lane = threadIdx.x, g = lane >> 2, tig = lane & 3;
  for (int i = lane; i < 8 * 128; i += 32) (&pl[0][0])[i] = 0;
  __syncwarp();
  for (int k = 2 * lane; k < 128; k += 64) {
    const float v0 = __uint_as_float((uint32_t)x[k] << 16) * S, v1 = __uint_as_float((uint32_t)x[k + 1] << 16) * S;
    const uint16_t y1 = cvt_e4m3x2(v1, v0);
    const float2 f1 = e4m3x2_to_f32(y1);
    const float r0 = (v0 - f1.x) * 256.f, r1 = (v1 - f1.y) * 256.f;
    const uint16_t y2 = cvt_e4m3x2(r1, r0);
    const float2 f2 = e4m3x2_to_f32(y2);
    const float s0 = (r0 - f2.x) * 256.f, s1 = (r1 - f2.y) * 256.f;
    const uint16_t y3 = cvt_e4m3x2(s1, s0);
    *reinterpret_cast<uint16_t*>(&pl[0][k]) = y1;
    *reinterpret_cast<uint16_t*>(&pl[1][k]) = y2;
    *reinterpret_cast<uint16_t*>(&pl[2][k]) = y3;
  }
  __syncwarp();
  for (int i = lane; i < 3 * 128; i += 32) planes[i] = (&pl[0][0])[i];
  float acc[4][4] = {};
#pragma unroll
  for (int j = 0; j < 4; ++j) {
    const int kb = 32 * j;
    const uint32_t a0 = *reinterpret_cast<const uint32_t*>(q + g * 128 + kb + tig * 4);
    const uint32_t a1 = *reinterpret_cast<const uint32_t*>(q + (g + 8) * 128 + kb + tig * 4);
    const uint32_t a2 = *reinterpret_cast<const uint32_t*>(q + g * 128 + kb + 16 + tig * 4);
    const uint32_t a3 = *reinterpret_cast<const uint32_t*>(q + (g + 8) * 128 + kb + 16 + tig * 4);
    const uint32_t b0 = *reinterpret_cast<const uint32_t*>(&pl[g][kb + tig * 4]);
    const uint32_t b1 = *reinterpret_cast<const uint32_t*>(&pl[g][kb + 16 + tig * 4]);
    float* c = acc[mode ? j : 0];
    asm volatile("mma.sync.aligned.m16n8k32.row.col.f32.e4m3.e4m3.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, {%0,%1,%2,%3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
  }
  float o[4];
  for (int i = 0; i < 4; ++i) o[i] = mode ? (acc[0][i] + acc[1][i]) + (acc[2][i] + acc[3][i]) : acc[0][i];
  D[g * 8 + 2 * tig] = o[0];
  D[g * 8 + 2 * tig + 1] = o[1];
  D[(g + 8) * 8 + 2 * tig] = o[2];
  D[(g + 8) * 8 + 2 * tig + 1] = o[3];
}

static double e4m3_val(uint8_t b) {
  const int s = b >> 7, e = (b >> 3) & 15, m = b & 7;
  const double v = e == 0 ? ldexp(m / 8.0, -6) : ldexp(1.0 + m / 8.0, e - 7);
  return s ? -v : v;
}
static uint16_t f2bf(float f) {
  uint32_t u;
  memcpy(&u, &f, 4);
  u += 0x7fffu + ((u >> 16) & 1u);
  return (uint16_t)(u >> 16);
}
static double bf2d(uint16_t b) {
  uint32_t u = (uint32_t)b << 16;
  float f;
  memcpy(&f, &u, 4);
  return f;
}

int main() {
  uint8_t *dq, *dpl;
  uint16_t* dx;
  float* dD;
  cudaMalloc(&dq, 16 * 128);
  cudaMalloc(&dx, 256);
  cudaMalloc(&dpl, 3 * 128);
  cudaMalloc(&dD, 16 * 8 * 4);
  srand(1);
  int bad_planes = 0, trials = 0;
  double worst_rel_mag[2] = {0, 0}, worst_abs_ulp[2] = {0, 0};
  long inexact[2] = {0, 0}, total = 0;
  for (int t = 0; t < 400; ++t) {
    uint8_t q[16 * 128];
    uint16_t x[128];
    for (auto& v : q) v = (uint8_t)(rand() & 15);
    // activations: gaussian-ish with outliers and tiny values, several dynamic ranges
    double amax = 0;
    const double spread = (t % 4 == 0) ? 1.0 : (t % 4 == 1) ? 30.0 : (t % 4 == 2) ? 1e3 : 1e6;
    for (auto& v : x) {
      double u = 0;
      for (int i = 0; i < 6; ++i) u += rand() / (double)RAND_MAX - 0.5;
      double mag = pow(spread, rand() / (double)RAND_MAX - 1.0);
      v = f2bf((float)(u * mag * 3.7));
      amax = fmax(amax, fabs(bf2d(v)));
    }
    int e;
    frexp(amax, &e);                       // amax = f * 2^e, f in [0.5, 1)
    const float S = (float)ldexp(1.0, 8 - e);  // amax * S in [128, 256)
    for (int mode = 0; mode < 2; ++mode) {
      cudaMemcpy(dq, q, sizeof(q), cudaMemcpyHostToDevice);
      cudaMemcpy(dx, x, sizeof(x), cudaMemcpyHostToDevice);
      qmma_kernel<<<1, 32>>>(dq, dx, S, dpl, dD, mode);
      uint8_t pl[3 * 128];
      float D[16 * 8];
      cudaMemcpy(pl, dpl, sizeof(pl), cudaMemcpyDeviceToHost);
      if (cudaMemcpy(D, dD, sizeof(D), cudaMemcpyDeviceToHost) != cudaSuccess) {
        printf("CUDA error %s\n", cudaGetErrorString(cudaGetLastError()));
        return 1;
      }
      if (mode == 0) {
        for (int k = 0; k < 128; ++k) {
          const double rec = e4m3_val(pl[k]) + ldexp(e4m3_val(pl[128 + k]), -8) + ldexp(e4m3_val(pl[256 + k]), -16);
          const double want = bf2d(x[k]) * S;
          ++trials;
          if (rec != want && fabs(want) >= ldexp(1.0, -14)) ++bad_planes;
          else if (fabs(rec - want) > ldexp(1.0, -25)) ++bad_planes;
        }
      }
      for (int r = 0; r < 16; ++r)
        for (int n = 0; n < 3; ++n) {
          double exact = 0, mag = 0;
          for (int k = 0; k < 128; ++k) {
            const double p = ldexp((double)q[r * 128 + k], -9) * e4m3_val(pl[n * 128 + k]);
            exact += p;
            mag += fabs(p);
          }
          const double got = D[r * 8 + n], err = fabs(got - exact);
          ++total;
          if (err != 0) ++inexact[mode];
          if (mag > 0) worst_rel_mag[mode] = fmax(worst_rel_mag[mode], err / mag);
          const double ulp = exact != 0 ? ldexp(1.0, ilogb(exact) - 23) : 1e-300;
          worst_abs_ulp[mode] = fmax(worst_abs_ulp[mode], err / ulp);
        }
    }
  }
  printf("plane split: %d of %d elements not reconstructed exactly\n", bad_planes, trials);
  for (int mode = 0; mode < 2; ++mode)
    printf("mode %d (%s): %ld of %ld sums inexact; worst |err| / sum|q y| = 2^%.1f ; worst |err| in fp32 ulps of the exact sum = %.1f\n", mode,
           mode ? "4 accumulators" : "1 chained accumulator", inexact[mode], total / 2, worst_rel_mag[mode] > 0 ? log2(worst_rel_mag[mode]) : -99.0,
           worst_abs_ulp[mode]);
  return 0;
}
