// LLaMA-Adapter inference ops (ref: lit_llama/adapter.py:145-165, lit_llama/adapter_v2.py:29-32).
//
// The reference adds, to every attention output of the adapted layers,
//     gating_factor[h] * softmax(q . prefix_k^T / sqrt(hs)) . prefix_v
// where prefix_k / prefix_v are c_attn applied to the learned adaption prompt (10 rows, NOT rotated, the same for every
// batch row and every position) - a second scaled_dot_product_attention call, a mask allocation, a broadcast multiply and
// an add per layer.  Here it is ONE kernel: a warp per (batch row, position, head) walks the prefix rows with an online
// softmax in fp32 and updates the attention output in place, rounding where torch rounds (the SDPA result, the product
// with the gate, the sum - each to the activation dtype).
//
// Adapter-v2 gives every Linear a per-output-column bias and scale applied after the matrix product,
//     y = adapter_scale * (linear(x) + adapter_bias):
// one elementwise kernel over the Linear's output (same roundings: the sum, then the product).
#include "common.cuh"

namespace llb {
namespace {

constexpr int AD_WARPS = 4;
constexpr int AD_MAX_EPL = 8;   // elements per lane: head_size <= 256

// q [B][nh][T][hs]; ak, av [Bk][nh][aT][hs] with batch stride kv_bstride (0: one prefix for every batch row);
// gate [nh]; y [B][T][nh][hs] updated in place.
template <typename T, int EPL>
__global__ void __launch_bounds__(AD_WARPS * 32)
adapter_attn_add_kernel(const T* __restrict__ q, const T* __restrict__ ak, const T* __restrict__ av,
                        const T* __restrict__ gate, T* __restrict__ y, int B, int nh, int Tn, int aT,
                        size_t kv_bstride, float scale) {
  constexpr int HS = EPL * 32;
  const int lane = threadIdx.x & 31;
  const size_t w = (size_t)blockIdx.x * AD_WARPS + (threadIdx.x >> 5);   // (b, t, h), h fastest
  if (w >= (size_t)B * Tn * nh) return;
  const int h = (int)(w % nh);
  const int t = (int)((w / nh) % Tn);
  const int b = (int)(w / ((size_t)nh * Tn));
  const T* qp = q + (((size_t)b * nh + h) * Tn + t) * HS;
  const T* kp = ak + b * kv_bstride + (size_t)h * aT * HS;
  const T* vp = av + b * kv_bstride + (size_t)h * aT * HS;
  float qv[EPL], acc[EPL];
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    qv[e] = dt<T>::ld(qp, lane + 32 * e);
    acc[e] = 0.f;
  }
  float m = -INFINITY, l = 0.f;
  for (int j = 0; j < aT; ++j) {
    float d = 0.f;
#pragma unroll
    for (int e = 0; e < EPL; ++e) d = fmaf(qv[e], dt<T>::ld(kp, (size_t)j * HS + lane + 32 * e), d);
    const float s = warp_sum(d) * scale;
    const float mn = fmaxf(m, s);
    const float c = expf(m - mn);       // 0 on the first row (m = -inf)
    const float p = expf(s - mn);
    l = l * c + p;
#pragma unroll
    for (int e = 0; e < EPL; ++e) acc[e] = fmaf(p, dt<T>::ld(vp, (size_t)j * HS + lane + 32 * e), acc[e] * c);
    m = mn;
  }
  const float g = dt<T>::ld(gate, h);
  const float inv = 1.0f / l;
  T* yp = y + (((size_t)b * Tn + t) * nh + h) * HS;
#pragma unroll
  for (int e = 0; e < EPL; ++e) {
    const float ay = dt<T>::rd(acc[e] * inv);            // the SDPA result in the activation dtype
    const float ga = dt<T>::rd(__fmul_rn(g, ay));        // gating_factor * ay
    dt<T>::st(yp, lane + 32 * e, __fadd_rn(dt<T>::ld(yp, lane + 32 * e), ga));
  }
}

template <typename T>
__global__ void bias_scale_kernel(const T* __restrict__ y, const T* __restrict__ bias, const T* __restrict__ scale,
                                  T* __restrict__ out, size_t rows, int n) {
  const size_t total = rows * (size_t)n;
  for (size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (size_t)gridDim.x * blockDim.x) {
    const int c = (int)(i % n);
    const float s = dt<T>::rd(__fadd_rn(dt<T>::ld(y, i), dt<T>::ld(bias, c)));
    dt<T>::st(out, i, __fmul_rn(dt<T>::ld(scale, c), s));
  }
}

template <typename T>
int launch_adapter_attn(const void* q, const void* ak, const void* av, const void* gate, void* y, int B, int nh, int Tn,
                        int hs, int aT, size_t kv_bstride, cudaStream_t st) {
  const size_t warps = (size_t)B * Tn * nh;
  const unsigned blocks = (unsigned)((warps + AD_WARPS - 1) / AD_WARPS);
  const float scale = 1.0f / sqrtf((float)hs);
#define LL_AD_CASE(E)                                                                                           \
  case E:                                                                                                       \
    adapter_attn_add_kernel<T, E><<<blocks, AD_WARPS * 32, 0, st>>>((const T*)q, (const T*)ak, (const T*)av,     \
                                                                    (const T*)gate, (T*)y, B, nh, Tn, aT,       \
                                                                    kv_bstride, scale);                         \
    break;
  switch (hs / 32) {
    LL_AD_CASE(1)
    LL_AD_CASE(2)
    LL_AD_CASE(3)
    LL_AD_CASE(4)
    LL_AD_CASE(5)
    LL_AD_CASE(6)
    LL_AD_CASE(7)
    LL_AD_CASE(8)
    default:
      set_error("adapter_attn_add: head_size %d not supported (multiples of 32 up to 256)", hs);
      return LL_ERR_UNSUPPORTED;
  }
#undef LL_AD_CASE
  LL_LAUNCH_OK("adapter_attn_add_kernel");
  return LL_OK;
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

int ll_adapter_attn_add(const void* q, const void* prefix_k, const void* prefix_v, const void* gate, void* y, int dtype,
                        int B, int n_head, int T, int head_size, int prompt_len, int prefix_batched, ll_stream stream) {
  LL_CHECK_ARG(q && prefix_k && prefix_v && gate && y, "null pointer");
  LL_CHECK_ARG(B > 0 && n_head > 0 && T > 0 && prompt_len > 0, "bad adapter_attn_add sizes");
  if (head_size <= 0 || head_size % 32 != 0 || head_size > 32 * AD_MAX_EPL) {
    set_error("adapter_attn_add: head_size %d not supported (multiples of 32 up to 256)", head_size);
    return LL_ERR_UNSUPPORTED;
  }
  const size_t bstride = prefix_batched ? (size_t)n_head * prompt_len * head_size : 0;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LL_F32) return launch_adapter_attn<float>(q, prefix_k, prefix_v, gate, y, B, n_head, T, head_size, prompt_len, bstride, st);
  if (dtype == LL_BF16) return launch_adapter_attn<__nv_bfloat16>(q, prefix_k, prefix_v, gate, y, B, n_head, T, head_size, prompt_len, bstride, st);
  if (dtype == LL_F16) return launch_adapter_attn<__half>(q, prefix_k, prefix_v, gate, y, B, n_head, T, head_size, prompt_len, bstride, st);
  set_error("adapter_attn_add: unsupported dtype %d", dtype);
  return LL_ERR_UNSUPPORTED;
}

int ll_bias_scale(const void* y, const void* bias, const void* scale, void* out, int dtype, size_t rows, int n,
                  ll_stream stream) {
  LL_CHECK_ARG(y && bias && scale && out && n > 0, "bad bias_scale arguments");
  if (rows == 0) return LL_OK;
  const size_t total = rows * (size_t)n;
  const unsigned blocks = (unsigned)((total + 255) / 256 < (size_t)sm_count() * 16 ? (total + 255) / 256 : (size_t)sm_count() * 16);
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == LL_F32) bias_scale_kernel<float><<<blocks, 256, 0, st>>>((const float*)y, (const float*)bias, (const float*)scale, (float*)out, rows, n);
  else if (dtype == LL_BF16) bias_scale_kernel<__nv_bfloat16><<<blocks, 256, 0, st>>>((const __nv_bfloat16*)y, (const __nv_bfloat16*)bias, (const __nv_bfloat16*)scale, (__nv_bfloat16*)out, rows, n);
  else if (dtype == LL_F16) bias_scale_kernel<__half><<<blocks, 256, 0, st>>>((const __half*)y, (const __half*)bias, (const __half*)scale, (__half*)out, rows, n);
  else { set_error("bias_scale: unsupported dtype %d", dtype); return LL_ERR_UNSUPPORTED; }
  LL_LAUNCH_OK("bias_scale_kernel");
  return LL_OK;
}

}  // extern "C"
