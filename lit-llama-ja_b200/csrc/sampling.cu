// Fused temperature / top-k / softmax / categorical draw for batch-1 decode (ref: generate.py:62-75).
//
// The reference runs, per token, `logits / temperature`, torch.topk, torch.where, softmax and torch.multinomial -
// a dozen small kernels and as many host launches, 0.25 ms per token on B200 against a 1.9 ms decode step.  Here
// the whole tail is ONE single-CTA kernel that can sit inside the captured decode graph: it reads the previous
// step's logits, writes the chosen token where the next step's embedding lookup and the output sequence expect it.
//
// Semantics, step for step as the reference computes them on a bf16 model:
//   l_i   = bf16(logit_i / temperature)
//   kth   = k-th largest l (ties at the threshold are kept, as `logits < kth -> -inf` keeps them)
//   p_i   = bf16(exp(l_i - max l) / sum_j exp(l_j - max l))        over the kept entries, fp32 inside
//   token ~ Categorical(p / sum p)
// The draw is by inverse CDF in index order from ONE uniform number supplied by the caller (torch's generator, so
// torch.manual_seed still governs the run); torch.multinomial consumes its generator differently (one exponential
// per entry), so the token STREAM differs from the reference's for the same seed while the distribution is the same.
#include "common.cuh"

namespace llb {
namespace {

constexpr int ST = 1024;  // threads

__device__ __forceinline__ uint32_t bf16_key(uint32_t b) { return (b & 0x8000u) ? (~b & 0xffffu) : (b | 0x8000u); }
__device__ __forceinline__ float key_value(uint32_t k) {
  const uint32_t b = (k & 0x8000u) ? (k & 0x7fffu) : (~k & 0xffffu);
  return __uint_as_float(b << 16);
}

// warp 0: the bin (scanning from 255 down) in which the `want`-th largest entry lies, and how many are still
// wanted inside that bin.  Lane l owns bins 255-8l .. 248-8l.
__device__ __forceinline__ void select_bin(const uint32_t* hist, uint32_t want, uint32_t* bin_out, uint32_t* want_out) {
  const int lane = threadIdx.x & 31;
  uint32_t h[8], mine = 0;
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    h[i] = hist[255 - 8 * lane - i];
    mine += h[i];
  }
  uint32_t incl = mine;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const uint32_t n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += n;
  }
  const uint32_t above = incl - mine;
  const bool here = above < want && want <= incl;
  const bool none = __ballot_sync(0xffffffffu, here) == 0u;  // fewer than `want` entries in all: keep everything
  if (here) {
    uint32_t w = want - above;
    int i = 0;
    for (; i < 7; ++i) {
      if (h[i] >= w) break;
      w -= h[i];
    }
    *bin_out = (uint32_t)(255 - 8 * lane - i);
    *want_out = w;
  }
  if (none && lane == 0) {
    *bin_out = 0;
    *want_out = 1;
  }
}

template <int EPT>
__global__ void __launch_bounds__(ST) sample_topk_kernel(const __nv_bfloat16* __restrict__ logits, int V, float temperature, int top_k,
                                                         const float* __restrict__ u, int32_t* __restrict__ step,
                                                         const int32_t* __restrict__ pos, int64_t* __restrict__ out_tokens,
                                                         int64_t* __restrict__ tok) {
  __shared__ uint32_t hist[256];
  __shared__ uint32_t s_sel[4];  // [0] high byte of the threshold, [1] entries still wanted inside it, [2] low byte, [3] max key
  __shared__ float s_warp[32];
  __shared__ float s_tot[2];
  __shared__ int s_pick;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int i0 = t * EPT;  // this thread's entries: i0 .. i0 + EPT - 1 (index order is the order of the CDF)
  uint32_t key[EPT];       // 0 = not an entry (every real key is >= 0x0080 or so; 0 would be -NaN)

  pdl_wait();
  if (t < 256) hist[t] = 0;
  if (t == 0) {
    s_sel[3] = 0;
    s_pick = 0x7fffffff;
  }
  __syncthreads();
  // ---- l_i and their order-preserving 16-bit keys; histogram of the high byte
  uint32_t kmax = 0;
#pragma unroll
  for (int j = 0; j < EPT; ++j) {
    key[j] = 0;
    if (i0 + j < V) {
      const float l = bf16r(__fdiv_rn(__bfloat162float(logits[i0 + j]), temperature));
      key[j] = max(bf16_key(f32_to_bf16_bits(l)), 1u);
      kmax = max(kmax, key[j]);
      atomicAdd(&hist[key[j] >> 8], 1u);
    }
  }
  kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, 16));
  kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, 8));
  kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, 4));
  kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, 2));
  kmax = max(kmax, __shfl_xor_sync(0xffffffffu, kmax, 1));
  if (lane == 0) atomicMax(&s_sel[3], kmax);
  __syncthreads();
  const bool all = top_k <= 0 || top_k >= V;
  if (!all) {
    // ---- radix select of the k-th largest key: high byte, then low byte
    if (warp == 0) select_bin(hist, (uint32_t)top_k, &s_sel[0], &s_sel[1]);
    __syncthreads();
    const uint32_t hb = s_sel[0], want = s_sel[1];
    if (t < 256) hist[t] = 0;
    __syncthreads();
#pragma unroll
    for (int j = 0; j < EPT; ++j)
      if (key[j] != 0 && (key[j] >> 8) == hb) atomicAdd(&hist[key[j] & 255u], 1u);
    __syncthreads();
    if (warp == 0) select_bin(hist, want, &s_sel[2], &s_sel[1]);
    __syncthreads();
  }
  const uint32_t thr = all ? 1u : ((s_sel[0] << 8) | s_sel[2]);
  const float lmax = key_value(s_sel[3]);

  // ---- softmax over the kept entries (fp32 inside, bf16 probabilities as F.softmax returns them)
  auto block_sum = [&](float v, int slot) {
    v = warp_sum(v);
    if (lane == 0) s_warp[warp] = v;
    __syncthreads();
    if (warp == 0) {
      float w = s_warp[lane];  // ST / 32 == 32 warps
      w = warp_sum(w);
      if (lane == 0) s_tot[slot] = w;
    }
    __syncthreads();
    return s_tot[slot];
  };
  float esum = 0.f;
#pragma unroll
  for (int j = 0; j < EPT; ++j)
    if (key[j] >= thr) esum += expf(__fsub_rn(key_value(key[j]), lmax));
  const float denom = block_sum(esum, 0);
  float psum = 0.f;
#pragma unroll
  for (int j = 0; j < EPT; ++j)
    if (key[j] >= thr) psum += bf16r(__fdiv_rn(expf(__fsub_rn(key_value(key[j]), lmax)), denom));
  const float total = block_sum(psum, 1);

  // ---- inverse CDF in index order: exclusive prefix of the per-thread sums, then a serial walk inside the thread
  float incl = psum;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const float n = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += n;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  float base = 0.f;
  for (int w = 0; w < warp; ++w) base += s_warp[w];
  const float target = __fmul_rn(u[step[0]], total);
  float cum = base + (incl - psum);
  int pick = -1, last_kept = -1;
#pragma unroll
  for (int j = 0; j < EPT; ++j)
    if (key[j] >= thr) {
      cum += bf16r(__fdiv_rn(expf(__fsub_rn(key_value(key[j]), lmax)), denom));
      last_kept = i0 + j;
      if (pick < 0 && cum > target) pick = i0 + j;
    }
  // the first index whose inclusive cumulative sum exceeds the target = the smallest candidate over all threads
  if (pick >= 0) atomicMin(&s_pick, pick);
  __syncthreads();
  if (s_pick == 0x7fffffff) {  // u * total rounded up to the total: the last kept entry (uniform branch: read after the sync)
    __syncthreads();
    if (t == 0) s_pick = -1;
    __syncthreads();
    if (last_kept >= 0) atomicMax(&s_pick, last_kept);
    __syncthreads();
  }
  if (t == 0) {
    const int64_t choice = s_pick;
    tok[0] = choice;
    if (out_tokens != nullptr) out_tokens[pos[0] + 1] = choice;
    step[0] = step[0] + 1;
  }
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

int ll_sample_topk(const void* logits, int V, float temperature, int top_k, const float* u, int32_t* step, const int32_t* pos,
                   int64_t* out_tokens, int64_t* tok, ll_stream stream) {
  LL_CHECK_ARG(logits && u && step && tok, "null pointer");
  LL_CHECK_ARG(V > 0 && V <= ST * 64, "vocabulary size %d out of range (max %d)", V, ST * 64);
  LL_CHECK_ARG(temperature > 0.f, "temperature must be positive");
  LL_CHECK_ARG(out_tokens == nullptr || pos != nullptr, "out_tokens needs pos");
  static bool once = false;
  if (!once) {
    prefer_max_smem(sample_topk_kernel<32>);
    prefer_max_smem(sample_topk_kernel<64>);
    once = true;
  }
  cudaStream_t st = (cudaStream_t)stream;
  if (V <= ST * 32)
    sample_topk_kernel<32><<<1, ST, 0, st>>>((const __nv_bfloat16*)logits, V, temperature, top_k, u, step, pos, out_tokens, tok);
  else
    sample_topk_kernel<64><<<1, ST, 0, st>>>((const __nv_bfloat16*)logits, V, temperature, top_k, u, step, pos, out_tokens, tok);
  LL_LAUNCH_OK("sample_topk_kernel");
  return LL_OK;
}

}  // extern "C"
