// GPTQ solver, inner block (ref: GPTQQuantizer.quantize, lit_llama/quantization.py:559-590).
//
// The reference walks the `count` (<= 128) columns of a block one at a time in Python: quantise column i, scale its
// error by 1 / Hinv[i][i], subtract err (x) Hinv[i][i:] from the columns to the right - 128 tiny launches chains per
// block, thousands per layer.  Rows never interact inside a block (Hinv is shared, everything else is per row), so
// the whole walk is ONE kernel: a warp owns a row, its 128 working weights live in registers (4 per lane), the
// block of Hinv sits in shared memory, and the column loop runs in the reference's order with the reference's fp32
// operation order (divide, round-half-even, clamp; product then subtract - no FMA contraction), so that given the
// same Hinv the chosen codes are bit-identical.  The cross-block update W[:, i2:] -= Err @ Hinv[i1:i2, i2:] that
// follows is a plain GEMM and stays in the host code (cuBLAS through torch.matmul).
#include "common.cuh"

namespace llb {
namespace {

constexpr int GB_MAX = 128;       // columns per block
constexpr int GB_WARPS = 8;       // warps per CTA
constexpr int GB_ROWS = 8;        // rows per warp (sequential)

__global__ void __launch_bounds__(GB_WARPS * 32) gptq_block_kernel(float* __restrict__ W, int ldw, const float* __restrict__ Hinv, int ldh,
                                                                   float* __restrict__ scales, float* __restrict__ zeros, int G,
                                                                   float* __restrict__ Q, int ldq, float* __restrict__ Err,
                                                                   float* __restrict__ loss, int N, int i1, int count, int groupsize,
                                                                   float maxq) {
  extern __shared__ float hs[];  // [count][count + 1]: rows i, columns >= i used
  const int ldhs = count + 1;
  for (int e = threadIdx.x; e < count * count; e += blockDim.x) {
    const int i = e / count, j = e - i * count;
    hs[i * ldhs + j] = Hinv[(size_t)(i1 + i) * ldh + i1 + j];
  }
  __syncthreads();
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  const int row0 = (blockIdx.x * GB_WARPS + warp) * GB_ROWS;
  for (int rr = 0; rr < GB_ROWS; ++rr) {
    const int r = row0 + rr;
    if (r >= N) return;
    float w[4], w0[4];
#pragma unroll
    for (int c = 0; c < 4; ++c) {
      const int j = lane + 32 * c;
      w[c] = j < count ? W[(size_t)r * ldw + i1 + j] : 0.f;
      w0[c] = w[c];
    }
    float scale = 1.f, zero = 0.f;
    if (groupsize < 0) {
      scale = scales[(size_t)r * G];
      zero = zeros[(size_t)r * G];
    }
    float lsum = 0.f;
    for (int i = 0; i < count; ++i) {
      if (groupsize > 0 && (i1 + i) % groupsize == 0) {
        // ref 571-577 (with the evident intent of the assignment): parameters of the group starting here, from the
        // weights as they stood when the block was entered (the reference reads the global W, which in-block
        // updates have not touched); find_params_weight, quantization.py:474-511
        float mn = INFINITY, mx = -INFINITY;
#pragma unroll
        for (int c = 0; c < 4; ++c) {
          const int j = lane + 32 * c;
          if (j >= i && j < i + groupsize && j < count) {
            mn = fminf(mn, w0[c]);
            mx = fmaxf(mx, w0[c]);
          }
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
          mn = fminf(mn, __shfl_xor_sync(0xffffffffu, mn, o));
          mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        }
        float xmin = fminf(mn, 0.f), xmax = fmaxf(mx, 0.f);
        if (xmin == 0.f && xmax == 0.f) {
          xmin = -1.f;
          xmax = 1.f;
        }
        scale = __fdiv_rn(__fsub_rn(xmax, xmin), maxq);
        zero = rintf(__fdiv_rn(-xmin, scale));
        if (lane == 0) {
          scales[(size_t)r * G + (i1 + i) / groupsize] = scale;
          zeros[(size_t)r * G + (i1 + i) / groupsize] = zero;
        }
      }
      const int slot = i >> 5, owner = i & 31;
      const float mine = slot == 0 ? w[0] : slot == 1 ? w[1] : slot == 2 ? w[2] : w[3];
      const float wi = __shfl_sync(0xffffffffu, mine, owner);
      const float d = hs[i * ldhs + i];
      // quantize_weight (quantization.py:468-472): scale * (clamp(round(x / scale) + zero, 0, maxq) - zero)
      const float q = __fmul_rn(scale, __fsub_rn(fminf(fmaxf(__fadd_rn(rintf(__fdiv_rn(wi, scale)), zero), 0.f), maxq), zero));
      const float diff = __fsub_rn(wi, q);
      lsum += __fdiv_rn(__fmul_rn(diff, diff), __fmul_rn(d, d));  // Losses1[:, i] = (w - q)^2 / d^2
      const float err = __fdiv_rn(diff, d);
      if (lane == owner) {
        Q[(size_t)r * ldq + i1 + i] = q;
        Err[(size_t)r * count + i] = err;
      }
      // W1[:, i:] -= err1 (x) Hinv1[i, i:]   (product rounded, then subtracted: as torch computes it)
      const float* hrow = hs + i * ldhs;
#pragma unroll
      for (int c = 0; c < 4; ++c) {
        const int j = lane + 32 * c;
        if (j >= i && j < count) w[c] = __fsub_rn(w[c], __fmul_rn(err, hrow[j]));
      }
    }
    if (lane == 0) loss[r] += lsum;
  }
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

int ll_gptq_block(float* W, int ldw, const float* Hinv, int ldh, float* scales, float* zeros, int G, float* Q, int ldq, float* Err,
                  float* loss, int N, int i1, int count, int groupsize, int maxq, ll_stream stream) {
  LL_CHECK_ARG(W && Hinv && scales && zeros && Q && Err && loss, "null pointer");
  LL_CHECK_ARG(N > 0 && count > 0 && count <= GB_MAX && i1 >= 0 && G > 0 && maxq > 0, "bad shape (count must be <= %d)", GB_MAX);
  LL_CHECK_ARG(groupsize < 0 || (groupsize > 0 && groupsize <= count && count % groupsize == 0 && i1 % groupsize == 0),
               "groupsize must be -1 or divide the block (groupsize=%d count=%d)", groupsize, count);
  const size_t smem = (size_t)count * (count + 1) * sizeof(float);
  static size_t configured = 0;
  if (smem > configured) {
    LL_CUDA(cudaFuncSetAttribute(gptq_block_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
    configured = smem;
  }
  const int rows_per_cta = GB_WARPS * GB_ROWS;
  const unsigned blocks = (unsigned)((N + rows_per_cta - 1) / rows_per_cta);
  gptq_block_kernel<<<blocks, GB_WARPS * 32, smem, (cudaStream_t)stream>>>(W, ldw, Hinv, ldh, scales, zeros, G, Q, ldq, Err, loss, N, i1,
                                                                            count, groupsize, (float)maxq);
  LL_LAUNCH_OK("gptq_block_kernel");
  return LL_OK;
}

}  // extern "C"
