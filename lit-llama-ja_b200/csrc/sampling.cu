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

// Fused tail of a sampled decode step (engine path): the draw above, then what `advance_pos_kernel` and
// `embed_ssq_kernel` did as two more launches - position update (with the same out-of-range trap) and the copy of the
// chosen token's embedding row into the residual stream - so that a token costs one small kernel before the first GEMV.
struct SampleTail {
  int32_t* pos_rw;               // int32[3] (position, ring slot, valid slots) to advance, or nullptr
  int cache_len, max_pos;
  const __nv_bfloat16* wte;      // [V][dim] embedding table, or nullptr (no embedding copy)
  __nv_bfloat16* x;              // [dim]
  int dim;
};

template <int EPT>
__global__ void __launch_bounds__(ST) sample_topk_kernel(const __nv_bfloat16* __restrict__ logits, int V, float temperature, int top_k,
                                                         const float* __restrict__ u, int32_t* __restrict__ step,
                                                         const int32_t* __restrict__ pos, int64_t* __restrict__ out_tokens,
                                                         int64_t* __restrict__ tok, const SampleTail tail) {
  __shared__ uint32_t s_cnt[32];
  __shared__ uint32_t s_bcast[2];
  __shared__ float s_warp[32];
  __shared__ float s_tot[2];
  __shared__ int s_pick;
  const int t = threadIdx.x, lane = t & 31, warp = t >> 5;
  const int i0 = t * EPT;  // this thread's entries: i0 .. i0 + EPT - 1 (index order is the order of the CDF)
  uint32_t key[EPT];       // 0 = not an entry (every real key is >= 0x0080 or so; 0 would be -NaN)

  pdl_wait();
  if (t == 0) s_pick = 0x7fffffff;
  // ---- l_i and their order-preserving 16-bit keys.  Eight entries per 16-byte load where the row allows it.
  uint32_t kmax = 0;
  const bool vec_ok = (reinterpret_cast<uintptr_t>(logits) & 15) == 0;
#pragma unroll
  for (int j0 = 0; j0 < EPT; j0 += 8) {
    uint32_t raw[8];
    if (vec_ok && i0 + j0 + 8 <= V) {
      const uint4 r = *reinterpret_cast<const uint4*>(logits + i0 + j0);
      const uint32_t w[4] = {r.x, r.y, r.z, r.w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        raw[2 * e] = w[e] & 0xffffu;
        raw[2 * e + 1] = w[e] >> 16;
      }
    } else {
#pragma unroll
      for (int e = 0; e < 8; ++e)
        raw[e] = i0 + j0 + e < V ? (uint32_t)__bfloat16_as_ushort(logits[i0 + j0 + e]) : 0x10000u;  // marker: no entry
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) {
      key[j0 + e] = 0;
      if (raw[e] <= 0xffffu) {
        const float l = bf16r(__fdiv_rn(__uint_as_float(raw[e] << 16), temperature));
        key[j0 + e] = max(bf16_key(f32_to_bf16_bits(l)), 1u);
        kmax = max(kmax, key[j0 + e]);
      }
    }
  }
  auto block_sum_u = [&](uint32_t v) -> uint32_t {  // sum over the CTA, result to every thread (two barriers)
    v = __reduce_add_sync(0xffffffffu, v);
    if (lane == 0) s_cnt[warp] = v;
    __syncthreads();
    if (warp == 0) {
      const uint32_t w = __reduce_add_sync(0xffffffffu, s_cnt[lane]);  // ST / 32 == 32 warps
      if (lane == 0) s_bcast[0] = w;
    }
    __syncthreads();
    return s_bcast[0];
  };
  kmax = __reduce_max_sync(0xffffffffu, kmax);
  if (lane == 0) s_cnt[warp] = kmax;
  __syncthreads();
  if (warp == 0) {
    const uint32_t w = __reduce_max_sync(0xffffffffu, s_cnt[lane]);
    if (lane == 0) s_bcast[1] = w;
  }
  __syncthreads();
  const uint32_t kmax_all = s_bcast[1];
  const bool all = top_k <= 0 || top_k >= V;
  // ---- the k-th largest key = the largest threshold that still keeps at least k entries: a bisection on the 16-bit key
  //      space with the keys in registers (16 counting rounds, no atomics: a histogram in shared memory serialises on the
  //      two or three exponent bins that hold most logits - it was 2/3 of this kernel's 30 us)
  uint32_t thr = 1u;
  if (!all) {
    uint32_t lo = 1u, hi = kmax_all;  // count(key >= lo) >= k holds throughout (k < V, every entry's key >= 1)
    while (lo < hi) {
      const uint32_t mid = (lo + hi + 1u) >> 1;
      uint32_t c = 0;
#pragma unroll
      for (int j = 0; j < EPT; ++j) c += key[j] >= mid ? 1u : 0u;
      if (block_sum_u(c) >= (uint32_t)top_k) lo = mid;
      else hi = mid - 1u;
    }
    thr = lo;
  }
  const float lmax = key_value(kmax_all);

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
  const int64_t choice = s_pick;
  if (t == 0) {
    tok[0] = choice;
    if (out_tokens != nullptr) out_tokens[pos[0] + 1] = choice;
    step[0] = step[0] + 1;
    if (tail.pos_rw != nullptr) {  // advance_pos_kernel's update; a position past the RoPE table stops the kernel (ref model.py:101)
      const int p = tail.pos_rw[0] + 1;
      if (p < 0 || p >= tail.max_pos) __trap();
      tail.pos_rw[0] = p;
      tail.pos_rw[1] = p % tail.cache_len;
      tail.pos_rw[2] = min(p + 1, tail.cache_len);
    }
  }
  if (tail.wte != nullptr && choice >= 0) {  // embed_ssq_kernel's copy: x = wte[token] (ref model.py:110)
    const uint4* src = reinterpret_cast<const uint4*>(tail.wte + (size_t)choice * tail.dim);
    uint4* dst = reinterpret_cast<uint4*>(tail.x);
    for (int i = t; i < tail.dim / 8; i += ST) dst[i] = src[i];
  }
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

static int launch_sample(const void* logits, int V, float temperature, int top_k, const float* u, int32_t* step, const int32_t* pos,
                         int64_t* out_tokens, int64_t* tok, const SampleTail& tail, ll_stream stream) {
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
    sample_topk_kernel<32><<<1, ST, 0, st>>>((const __nv_bfloat16*)logits, V, temperature, top_k, u, step, pos, out_tokens, tok, tail);
  else
    sample_topk_kernel<64><<<1, ST, 0, st>>>((const __nv_bfloat16*)logits, V, temperature, top_k, u, step, pos, out_tokens, tok, tail);
  LL_LAUNCH_OK("sample_topk_kernel");
  return LL_OK;
}

int ll_sample_topk(const void* logits, int V, float temperature, int top_k, const float* u, int32_t* step, const int32_t* pos,
                   int64_t* out_tokens, int64_t* tok, ll_stream stream) {
  SampleTail tail;
  memset(&tail, 0, sizeof(tail));
  return launch_sample(logits, V, temperature, top_k, u, step, pos, out_tokens, tok, tail, stream);
}

int ll_sample_advance_embed(const void* logits, int V, float temperature, int top_k, const float* u, int32_t* step, int32_t* pos,
                            int64_t* out_tokens, int64_t* tok, int cache_len, int max_pos, const void* wte, void* x, int dim,
                            ll_stream stream) {
  LL_CHECK_ARG(pos && cache_len > 0 && max_pos > 0, "bad position arguments");
  LL_CHECK_ARG(wte && x && dim > 0 && dim % 8 == 0 && ((uintptr_t)wte & 15) == 0 && ((uintptr_t)x & 15) == 0,
               "embedding table / row must be 16-byte aligned, dim a multiple of 8");
  SampleTail tail;
  tail.pos_rw = pos; tail.cache_len = cache_len; tail.max_pos = max_pos;
  tail.wte = (const __nv_bfloat16*)wte; tail.x = (__nv_bfloat16*)x; tail.dim = dim;
  return launch_sample(logits, V, temperature, top_k, u, step, pos, out_tokens, tok, tail, stream);
}

}  // extern "C"
