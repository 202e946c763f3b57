/*
 * llb200 - B200-native (sm_100a) kernels for lit-llama-ja's quantized-inference hot path.
 *
 * Drop-in boundary: a C-ABI shared library (plain pointers and sizes, no torch types).
 * Every pointer is a DEVICE pointer unless it says "host".  Every call enqueues work on
 * `stream` (a cudaStream_t passed as void*) and returns immediately; nothing synchronises.
 * Return value: LL_OK or a negative LL_ERR_* code; ll_last_error() then holds a message
 * (thread-local).  Nothing throws across this boundary.
 *
 * "ref:" citations are into /root/reference (if001/lit-llama-ja); each entry point names
 * the reference interface it replaces.  The reference has no native code on this path: its
 * FFI-equivalent is the Triton launcher + the PyTorch fallback inside
 * lit_llama/quantization.py, bitsandbytes for llm.int8, and ATen ops in lit_llama/model.py.
 */
#ifndef LLB200_H
#define LLB200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define LL_ABI_VERSION 11

#define LL_OK 0
#define LL_ERR_ARG (-1)         /* bad shape / dtype / null pointer */
#define LL_ERR_CUDA (-2)        /* a CUDA runtime/driver call failed */
#define LL_ERR_UNSUPPORTED (-3) /* valid request this build has no kernel for */

/* element types of activation / parameter buffers */
#define LL_F32 0
#define LL_BF16 1
#define LL_F16 2

typedef void* ll_stream; /* cudaStream_t */

#if defined(__GNUC__)
#define LL_API __attribute__((visibility("default")))
#else
#define LL_API
#endif

LL_API const char* ll_last_error(void);
LL_API int ll_abi_version(void);
/* number of SMs of the current device (host query; <0 on error) */
LL_API int ll_sm_count(void);

/* ---------------------------------------------------------------------------------------
 * GPTQ column-block storage (ref: lit_llama/quantization.py:339-372).
 *   qw      uint8, logical [N][K/epb], MEMORY [K/epb][N] (N fastest) - the reference's
 *           column-major buffer, used in place.  epb = 8/bits; entry nr of a byte sits at
 *           bit nr*bits (bits=4: even k = low nibble).
 *   scales, zeros  [N][G] row-major, G = ceil(K/tile); tile_cols = -1 means tile = K.
 * ------------------------------------------------------------------------------------- */

/* ref: ColBlockQuantizedLinear.pack_weight, quantization.py:374-388.  Bit-exact: per group,
 * in the weight dtype, `/= scale` then `+= zero`, clamp [0, 2^bits-1], TRUNCATE to uint8. */
LL_API int ll_pack_weight(const void* weight /*[N][K]*/, int weight_dtype, const void* scales, const void* zeros,
                   int param_dtype, uint8_t* qw, int N, int K, int bits, int tile_cols, ll_stream stream);

/* ref: ColBlockQuantizedLinear.get_weight, quantization.py:390-409.  Bit-exact: code as float,
 * `-= zero`, `*= scale`, each rounded to out_dtype.  out is [N][K] row-major. */
LL_API int ll_get_weight(const uint8_t* qw, const void* scales, const void* zeros, int param_dtype, void* out,
                  int out_dtype, int N, int K, int bits, int tile_cols, ll_stream stream);

/* ref: ColBlockQuantizedLinear.forward, quantization.py:411-421 (both branches), any M, bits in
 * {4,8}, any tile, fp32 or bf16 activations: y[M][N] = x[M][K] . W^T (+ bias), W as get_weight
 * would produce it in the activation dtype, fp32 accumulation, one rounding on store.
 * SIMT kernel: the always-correct path for shapes/dtypes the tensor-core kernels do not take. */
LL_API int ll_qlinear_simt(const void* x, int act_dtype, int M, const uint8_t* qw, const void* scales, const void* zeros,
                    int param_dtype, const void* bias, void* y, int N, int K, int bits, int tile_cols,
                    ll_stream stream);

/* ---------------------------------------------------------------------------------------
 * int4 decode GEMV (M <= 8), bf16 activations.  Replaces the Triton launcher
 * qlinear_4bit_weight (quantization.py:282-331: pads M to 256, per-row scales only) and the
 * dequantise-then-F.linear fallback (quantization.py:420-421) for grouped scales.
 *
 * The kernel streams a one-time REPACK of the checkpoint bytes ("stream"): blocks of
 * 16 output columns x KB input features (KB = 128, 64 or 32 = gcd-style divisor of the group
 * size), each block = the 32 lanes' mma.m16n8k16 A-fragments in issue order followed by the
 * block's 16 (scale, -(16+zero)) bf16 pairs.  Blocks are ordered (tile, k-block), k fastest,
 * so any contiguous range of work items is one contiguous byte range (1-D bulk TMA).
 * Arithmetic: y = sum_blocks scale * (sum_k (16+q) x_k - (16+zero) sum_k x_k), products exact,
 * fp32 accumulation: the exact (q - zero) * scale of the reference's Triton kernel
 * (quantization.py:247-251) without its TF32 rounding or the fallback's bf16 weight rounding.
 * ------------------------------------------------------------------------------------- */

/* bytes of the repacked stream for an [N][K] int4 matrix with scale groups of `group` columns
 * (group = K for per-row); 0 if the shape is not supported by the GEMV (use ll_qlinear_simt). */
LL_API size_t ll_int4_stream_bytes(int N, int K, int group);

/* One-time repack: qw (checkpoint layout) + scales/zeros [N][K/group] -> stream.  status[0]
 * (device int32) is set to 1 unless every scale and every (16 + zero) is exactly representable
 * in bf16 (then the caller must use ll_qlinear_simt). */
LL_API int ll_int4_prepare(const uint8_t* qw, const void* scales, const void* zeros, int param_dtype, uint8_t* stream_out,
                    int N, int K, int group, int kind, int32_t* status, ll_stream stream);

/* Stream kinds (same size, same tile / block order; they differ in how a block's nibbles are arranged):
 *   LL_STREAM_BF16: words are mma.m16n8k16 bf16 A fragments (nibble i at bit 3 + 4 i); tail (scale, -(16 + zero)); M = 1..8.
 *   LL_STREAM_U8:   words are mma.m16n8k32 u8 A fragment pairs (column g in the low nibbles, g + 8 in the high nibbles);
 *                   tail (scale, -zero); the batch-1 decode kernel (activations as four exact base-256 digit planes).
 * ll_int4_stream_kind() names the kind the GEMV wants for a shape and batch: prepare that one and pass it in
 * ll_gemv_args.stream_kind. */
#define LL_STREAM_BF16 0
#define LL_STREAM_U8 1
LL_API int ll_int4_stream_kind(int N, int K, int group, int M);

/* GPTQ 8-bit weights (`gptq.int8`: ColBlockQuantizedLinear(bits=8), ref utils.py:182-184, quantization.py:343-347; the
 * reference serves it only through get_weight + F.linear, quantization.py:420-421).  Stream kind LL_STREAM_W8 of the same
 * batch-1 GEMV (ll_int4_gemv with stream_kind = LL_STREAM_W8, M = 1, group % 128 == 0): a block is 16 columns x 128 k
 * bytes laid out as the lanes' u8 mma.m16n8k32 A-fragment registers + the (scale, -zero) tail, so the main loop is loads and
 * MMAs only; activations as the four exact base-256 digit planes.  qw: the checkpoint buffer, logical [N][K], memory [K][N].
 * status as in ll_int4_prepare. */
#define LL_STREAM_W8 2
LL_API size_t ll_q8_stream_bytes(int N, int K, int group);
LL_API int ll_q8_gemv_supported(int N, int K, int group, int nmat, int M); /* 1 / 0 */
LL_API int ll_q8_prepare(const uint8_t* qw, const void* scales, const void* zeros, int param_dtype, uint8_t* stream_out, int N,
                  int K, int group, int32_t* status, ll_stream stream);

#define LL_EPI_STORE 0    /* y = bf16(acc (+bias)) */
#define LL_EPI_RESIDUAL 1 /* y (residual stream, in place) = bf16(y + bf16(acc)) */
#define LL_EPI_ROPE_KV 2  /* c_attn: RoPE on q,k; q -> y, k,v -> KV cache at pos[1] */
#define LL_EPI_SWIGLU 3   /* y = bf16(bf16(silu(bf16(accA))) * bf16(accB)), accB from w2 */
#define LL_EPI_PUSH 4     /* y_r[M][N] = x . W^T on every rank r (peer stores) + arrival counters, see push_dst */

#define LL_GEMV_PDL 1 /* launch with programmatic stream serialization (prologue overlap) */
/* Footprint parity of a batch-1 launch inside a PDL chain (stream kind 1 only).  Consecutive GEMVs of a chain alternate
 * EVEN / ODD: an EVEN launch asks for 116 KB of shared memory (two EVEN CTAs cannot share an SM), an ODD launch carries
 * two idle warps (two ODD CTAs exceed the register file), an EVEN + ODD pair fits - so the block scheduler can never put
 * two CTAs of the SAME GEMV on one SM, whatever else is resident (profiles/r02_gemv_experiments.md section 6). */
#define LL_GEMV_EVEN 4
#define LL_GEMV_ODD 2

typedef struct ll_gemv_args {
  const uint8_t* w;    /* repacked stream from ll_int4_prepare */
  const uint8_t* w2;   /* second matrix (LL_EPI_SWIGLU: c_fc2), same shape; else NULL */
  int32_t N, K, group; /* group = k per scale group (K when per-row) */
  int32_t M;           /* rows of x, 1..8 */
  const void* x;       /* [M][K] bf16 */
  /* optional fused RMSNorm on x (ref: model.py:276-283): x_used = bf16(w * bf16(x * r)),
   * r = bf16(rsqrt(bf16(bf16(sum(bf16(x*x)) / K) + eps))), computed by every CTA from its staged row */
  const void* norm_weight; /* [K] bf16 or NULL */
  float eps;
  int32_t epilogue;
  int32_t flags;
  void* y;          /* see LL_EPI_* ; [M][N] bf16 (ROPE_KV: q [M][N/3]) */
  const void* bias; /* [N] bf16 or NULL (LL_EPI_STORE) */
  /* LL_EPI_ROPE_KV (ref: model.py:204-228, 312-329) */
  const float* rope;  /* [block_size][head_size/2][2] fp32 (cos, sin) */
  const int32_t* pos; /* device int32[3]: rope position, cache slot, #valid slots */
  void* k_cache;      /* [M][n_head][cache_len][head_size] bf16 */
  void* v_cache;
  int32_t n_head, head_size, cache_len;
  int32_t grid; /* 0 = one CTA per SM; >0 forces the CTA count (tests).  CTAs own whole 16-column tiles. */
  /* LL_EPI_PUSH (tensor parallel, row-parallel layer): the bf16 result [M][N] is stored into push_dst[r] for every
   * rank r < push_world (peer-mapped receive buffers, the local one included; y is not used), then every CTA adds 1 to
   * *push_cnt[r] with release semantics at system scope: compute and the first half of the all-reduce in one kernel. */
  uint64_t push_dst[8];
  uint64_t push_cnt[8];
  int32_t push_world;
  int32_t stream_kind; /* LL_STREAM_* of w / w2 (what ll_int4_prepare was asked for) */
  /* Stream kinds 1 / 2 without the fused norm (optional): x already converted by its producer (ll_decode_attn_digits) into the
   * kernel's own staging format - four signed base-256 digit planes [4][K + 16] bytes and [K / 128][4] float records
   * (2^-sh, sum x, 2^(16 - sh), 0) - so that the prologue is two bulk copies.  x is then not read. */
  const uint8_t* x_planes;
  const float* x_rec;
} ll_gemv_args;

/* Host query: LL_OK if the GEMV takes (N, K, group, nmat, M) (LL_ERR_UNSUPPORTED otherwise, e.g. M*K too
 * large for shared memory); *smem_bytes / *ctas (optional) report the launch it would use. */
LL_API int ll_int4_gemv_plan(int N, int K, int group, int nmat, int M, int grid, size_t* smem_bytes, int32_t* ctas);
LL_API int ll_int4_gemv(const ll_gemv_args* args /* host */, ll_stream stream);

/* ---------------------------------------------------------------------------------------
 * Large-M path (prefill, batched decode): tcgen05 / TMEM / TMA tensor-core GEMM.
 * y[M][N] = x[M][K] . w[N][K]^T (+ bias), all bf16 row-major, fp32 accumulation in TMEM.  With
 * w = ll_get_weight(..., LL_BF16) this is the reference's F.linear(inp, get_weight(inp.dtype), bias)
 * (quantization.py:420-421) and what its Triton kernel (quantization.py:185-331) computes for large M.
 * K % 8 == 0; any M, N (tails are zero-filled by TMA and masked on store).
 * ------------------------------------------------------------------------------------- */
LL_API int ll_gemm_bf16_tc(const void* x, const void* w, const void* bias, void* y, int M, int N, int K, ll_stream stream);

/* Fused int4 de-quantise + GEMM on tcgen05 for 1 <= M <= 128 rows (batched decode, short prefill; the HBM-bound regime):
 * y[M][N] (bf16) = x[M][K] (bf16) . W^T (+ bias), W de-quantised INSIDE the main loop from the packed stream
 * (w_stream = ll_int4_prepare(..., LL_STREAM_U8, ...), the batch-1 GEMV's stream) with the reference's bf16 rounding steps
 * bf16(bf16(q - zero) * scale) (get_weight, ref quantization.py:390-409), fp32 accumulation in tensor memory - the
 * arithmetic of F.linear(inp, get_weight(bf16)) (ref quantization.py:420-421) and the structure of the reference's
 * Triton kernel (de-quantise in the K loop, ref quantization.py:259-267).  N % 128 == 0, group % 128 == 0, K % group == 0.
 * K is split over CTAs to fill the SMs; `scratch` (ll_int4_gemm_tc_scratch_bytes(), may be 0 -> NULL allowed) holds the
 * partial sums, added in a fixed order by a second kernel. */
LL_API size_t ll_int4_gemm_tc_scratch_bytes(int M, int N, int K);
LL_API int ll_int4_gemm_tc(const uint8_t* w_stream, const void* x, const void* bias, void* y, float* scratch, int M, int N,
                           int K, int group, ll_stream stream);

/* No counterpart in the reference: added for callers that cannot call cudaMalloc
 * (code under CUDA-graph capture, tight serving loops, harnesses that forbid it).  Same arithmetic, arguments and result
 * as ll_int4_gemm_tc, but the split-K partial sums go to an array built into the library
 * (LL_INT4_GEMM_TC_STATIC_SCRATCH_BYTES per GPU, held until exit); the call allocates nothing.  Shapes whose
 * ll_int4_gemm_tc_scratch_bytes() exceeds that maximum return LL_ERR_UNSUPPORTED (every 7B layer up to M = 128 fits
 * except the 32000-column head above M = 64).  All calls on a device share the array: calls that overlap on the GPU
 * race; calls ordered on one stream do not (nothing in the array has to start at zero). */
#define LL_INT4_GEMM_TC_STATIC_SCRATCH_BYTES (32u << 20)
LL_API int ll_int4_gemm_tc_static(const uint8_t* w_stream, const void* x, const void* bias, void* y, int M, int N, int K,
                                  int group, ll_stream stream);

/* ---------------------------------------------------------------------------------------
 * llm.int8 (ref: Linear8bitLt, quantization.py:36-75; arithmetic = bitsandbytes double_quant / MatMul8bitLt,
 * un-vendored in the reference, restated from arXiv:2208.07339 - see oracle/quant.py).
 * ------------------------------------------------------------------------------------- */

/* ref: Linear8bitLt._quantize_weight, quantization.py:67-75: W.half(), SCB[n] = absmax of row n,
 * CB = rint(127 * W / SCB) int8 [N][K] row-major. */
LL_API int ll_int8_quantize_weight(const void* w /*[N][K]*/, int w_dtype, int8_t* cb, float* scb, int N, int K, ll_stream stream);

/* ref: bnb MatMul8bitLt forward (has_fp16_weights=False, threshold) as reached through quantization.py:36-45:
 * y[M][N] = int8 matmul of the row-quantised non-outlier part + fp16 outlier-column part (+ bias), in the
 * activation dtype.  K % 128 == 0.  workspace: ll_int8_workspace_bytes(min(M, 64), N, K) bytes, 256-byte aligned. */
LL_API size_t ll_int8_workspace_bytes(int M, int N, int K);
/* cbt: CB transposed, int8 [K][N] (optional, NULL = none): with it an outlier column of the weights is one contiguous row
 * read coalesced (the outlier part then costs count * N bytes instead of one 32-byte sector per (row, outlier column)). */
LL_API int ll_int8_linear(const void* x, int act_dtype, const int8_t* cb, const int8_t* cbt, const float* scb, const void* bias,
                   void* y, int M, int N, int K, float threshold, void* workspace, size_t workspace_bytes, ll_stream stream);

/* ---------------------------------------------------------------------------------------
 * GPTQ solver, one block of <= 128 columns (ref GPTQQuantizer.quantize, quantization.py:559-590; fp32).
 * W [N][ldw] working weights (columns i1 .. i1+count-1 are read), Hinv [K][ldh] upper Cholesky factor of the inverse
 * Hessian, scales/zeros [N][G] (read for groupsize == -1; written at group starts otherwise - the reference's grouped
 * branch raises at 573-577, its evident intent is implemented), Q [N][ldq] receives the on-grid weights of the block's
 * columns, Err [N][count] the scaled errors for the caller's cross-block update W[:, i2:] -= Err @ Hinv[i1:i2, i2:],
 * loss[N] += sum_i (w - q)^2 / d^2.  Same fp32 operation order as the reference: identical codes for identical Hinv.
 * ------------------------------------------------------------------------------------- */
LL_API int ll_gptq_block(float* W, int ldw, const float* Hinv, int ldh, float* scales, float* zeros, int G, float* Q, int ldq,
                  float* Err, float* loss, int N, int i1, int count, int groupsize, int maxq, ll_stream stream);

/* ---------------------------------------------------------------------------------------
 * Sampling tail of generate() (ref generate.py:62-75: logits / temperature, top-k mask, softmax, multinomial) as one
 * single-CTA kernel, bf16 logits [V].  l = bf16(logit / temperature); entries below the top_k-th largest l are dropped
 * (ties kept; top_k <= 0 keeps all); p = bf16(softmax(l)); the token is drawn by inverse CDF in index order from the
 * uniform number u[*step]; *step is then incremented.  The token is written to tok[0] (int64: the next step's
 * embedding index) and, when out_tokens is not NULL, to out_tokens[pos[0] + 1].  Same distribution as the reference's
 * torch.multinomial; a different random stream.
 * ------------------------------------------------------------------------------------- */
LL_API int ll_sample_topk(const void* logits, int V, float temperature, int top_k, const float* u, int32_t* step, const int32_t* pos,
                   int64_t* out_tokens, int64_t* tok, ll_stream stream);
/* The same draw followed, in the same kernel, by what ll_advance_pos and ll_embed_ssq do (the engine's sampled step: one
 * launch instead of three before the first GEMV of the next token): pos (int32[3]) is advanced by one position after the
 * token was written to out_tokens[pos[0] + 1] (a position >= max_pos traps, as ll_advance_pos), and the chosen token's row
 * of wte [V][dim] bf16 is copied to x [dim] (ref model.py:110). */
LL_API int ll_sample_advance_embed(const void* logits, int V, float temperature, int top_k, const float* u, int32_t* step,
                            int32_t* pos, int64_t* out_tokens, int64_t* tok, int cache_len, int max_pos, const void* wte,
                            void* x, int dim, ll_stream stream);

/* ---------------------------------------------------------------------------------------
 * Tensor-parallel exchange (new: the reference is single-device, generate.py:122).  One-shot all-reduce over
 * NVLink peer memory fused with the residual add: x = bf16(x + bf16(sum over ranks of part_r)), summed in rank
 * order on every rank.  peer_parts / peer_flags: HOST arrays of `world` device pointers (symmetric-memory mappings
 * of every rank's partial-sum buffer and int32[world] flag array, zeroed once).  The caller alternates between two
 * partial-sum buffers from call to call.  epoch_counter: local device int32[LL_TP_MAX_CTAS], zero before first use.
 * flags: LL_GEMV_PDL = launch as a link of a programmatic-dependent-launch chain (the next kernel of the stream may start
 * its prologue while the exchange runs; needs the EVEN / ODD GEMV footprints around it), 0 = plain launch. */
#define LL_TP_MAX_CTAS 8
LL_API int ll_tp_allreduce_add(void* x, const uint64_t* peer_parts, const uint64_t* peer_flags, int32_t* epoch_counter, int rank,
                        int world, int n, int flags, ll_stream stream);
/* The same exchange in ONE NVLink traversal: every rank pushes its partial sums `part` ([n] bf16, local) into every peer's
 * receive buffer as 16-byte {word, epoch, word, epoch} units (data and flag in one store, NCCL-LL style) and reads the
 * peers' units from its own memory.  peer_recv: HOST array of `world` device pointers - the symmetric-memory mappings of
 * every rank's receive buffer of ll_tp_ll_recv_bytes(world, n) bytes, zeroed once.  epoch_counter, flags: as above.
 * Same result bits as ll_tp_allreduce_add (rank-order fp32 sum, one bf16 rounding, residual add). */
LL_API size_t ll_tp_ll_recv_bytes(int world, int n);
LL_API int ll_tp_ll_allreduce_add(void* x, const void* part, const uint64_t* peer_recv, int32_t* epoch_counter, int rank, int world,
                           int n, int flags, ll_stream stream);
/* Second half of the pushed exchange (LL_EPI_PUSH): wait until `arrivals` more CTAs have signalled the counter of the
 * current buffer (cnt[j], j alternating 0, 1 from call to call), then x = bf16(x + bf16(sum_r recv[j][r])) from LOCAL
 * memory, rank order.  recv: [2][world][n] bf16 in this rank's symmetric memory; cnt: uint32[2], zero before first use. */
LL_API int ll_tp_reduce_add(void* x, const void* recv, uint32_t* cnt, int32_t* epoch_counter, int world, int n, int arrivals,
                     ll_stream stream);

/* ---------------------------------------------------------------------------------------
 * model.py kernels
 * ------------------------------------------------------------------------------------- */

/* ref: RMSNorm.forward, model.py:276-283 - every intermediate rounded to the activation dtype. */
LL_API int ll_rmsnorm(const void* x, const void* weight, void* y, int dtype, int rows, int dim, float eps, ll_stream stream);

/* ref: wte lookup, model.py:110, fused with the sum-of-squares partials the first fused RMSNorm
 * consumes: x[m][:] = wte[idx[m]][:]; ssq (optional, may be NULL): ssq[m][p] = sum over 16-column tile p of
 * bf16(x*x). */
LL_API int ll_embed_ssq(const void* wte, const int64_t* idx, void* x, float* ssq, int M, int dim, ll_stream stream);

/* ref: apply_rope (model.py:312-329) on q and k of a fused qkv row + KV-cache write (model.py:218-228,
 * in place - the ring slot replaces the reference's roll).  qkv [B][T][3E]; q_out [B][nh][T][hs];
 * caches [B][nh][S][hs]; pos_ids[T] rope rows, slots[T] cache slots (device int32). */
LL_API int ll_rope_kv_append(const void* qkv, const float* rope, const int32_t* pos_ids, const int32_t* slots, void* q_out,
                      void* k_cache, void* v_cache, int dtype, int B, int T, int n_head, int head_size, int cache_len,
                      ll_stream stream);

/* ref: apply_rope, model.py:312-329 (standalone form): x, y [B][T][nh][hs]; rope [T][hs/2][2] fp32 rows
 * already selected for the T positions.  fp32 rotation of interleaved pairs, rounded to dtype. */
LL_API int ll_apply_rope(const void* x, const float* rope, void* y, int dtype, int B, int T, int n_head, int head_size,
                  ll_stream stream);

/* ref: the prefix cross-attention of LLaMA-Adapter, lit_llama/adapter.py:145-165 - the second
 * scaled_dot_product_attention (queries against the adaption prompt's keys/values, all-true mask), the product with
 * the per-head gating factor and the sum with the self-attention output, in one launch:
 *   y[b][t][h][:] = dtype(y[b][t][h][:] + dtype(gate[h] * dtype(softmax(q[b][h][t] . K[h]^T / sqrt(hs)) . V[h])))
 * q [B][nh][T][hs] (rotated, as ll_rope_kv_append leaves it); prefix_k / prefix_v [nh][prompt_len][hs] (one prompt for
 * every batch row, prefix_batched = 0) or [B][nh][prompt_len][hs] (prefix_batched = 1, the reference's repeated
 * layout); gate [nh]; y [B][T][nh*hs] updated in place.  head_size: multiples of 32 up to 256. */
LL_API int ll_adapter_attn_add(const void* q, const void* prefix_k, const void* prefix_v, const void* gate, void* y,
                               int dtype, int B, int n_head, int T, int head_size, int prompt_len, int prefix_batched,
                               ll_stream stream);

/* ref: adapter_v2_new_forward, lit_llama/adapter_v2.py:29-32, applied to a Linear's output y [rows][n]:
 *   out = dtype(scale[c] * dtype(y[r][c] + bias[c]));  out may alias y. */
LL_API int ll_bias_scale(const void* y, const void* bias, const void* scale, void* out, int dtype, size_t rows, int n,
                         ll_stream stream);

/* ref: F.scaled_dot_product_attention over the KV cache for ONE query token per sequence
 * (model.py:237 with T=1): attends slots [0, pos[2]).  q [B][nh][hs]; y [B][nh*hs].
 * Split along the cache (flash-decoding) with a deterministic last-arriver combine.
 * scratch: float [B*nh*splits*(hs+2)], counters int32 [B*nh] (zero; left zero). */
LL_API int ll_decode_attn(const void* q, const void* k_cache, const void* v_cache, const int32_t* pos, void* y, int dtype,
                   int B, int n_head, int head_size, int cache_len, int max_splits, float* scratch,
                   int32_t* counters, int flags, ll_stream stream);
LL_API size_t ll_decode_attn_scratch_bytes(int B, int n_head, int head_size, int max_splits);
/* The same attention (bf16, batch 1, head size 128, a split that fits the K/V-staging kernel: ..._supported() says so) whose
 * per-head leader also writes y in the batch-1 GEMV's staging format: planes [4][plane_stride] bytes (plane_stride =
 * n_head * 128 + 16), rec [n_head][4] floats - one head is one 128-element k-block of attn.c_proj's input
 * (ll_gemv_args.x_planes / x_rec).  Same arithmetic and order as the GEMV's own conversion: identical bits downstream. */
LL_API int ll_decode_attn_digits_supported(int dtype, int B, int head_size, int cache_len, int max_splits);
LL_API int ll_decode_attn_digits(const void* q, const void* k_cache, const void* v_cache, const int32_t* pos, void* y, int B,
                          int n_head, int head_size, int cache_len, int max_splits, int flags, uint8_t* planes, float* rec,
                          int plane_stride, ll_stream stream);

/* ref: MLP.forward's silu(a) * b, model.py:258, rounded like torch (silu -> dtype, product -> dtype). */
LL_API int ll_silu_mul(const void* a, const void* b, void* y, int dtype, size_t n, ll_stream stream);
/* y = dtype(a + b) : the residual adds of Block.forward, model.py:172-173 */
LL_API int ll_add(const void* a, const void* b, void* y, int dtype, size_t n, ll_stream stream);
/* pos[0] += 1; pos[1] = pos[0] % cache_len; pos[2] = min(pos[0] + 1, cache_len): one decode step
 * (ref: generate.py:77 input_pos advance + model.py:221-225 sliding window).  A position outside [0, max_pos)
 * (max_pos = rows of the RoPE table = block_size) traps on the device: the reference's
 * rope_cache.index_select(0, input_pos) raises there (model.py:101). */
LL_API int ll_advance_pos(int32_t* pos, int cache_len, int max_pos, ll_stream stream);
/* pos[0] = input_pos[0] (device int64, the caller's position tensor); pos[1], pos[2] as above.
 * Lets LLaMA.forward(idx, max_seq_length, input_pos) honour an arbitrary input_pos without a
 * host sync (ref: model.py:101-104 index_select by input_pos). */
LL_API int ll_set_pos(int32_t* pos, const int64_t* input_pos, int cache_len, int max_pos, ll_stream stream);

#ifdef __cplusplus
}
#endif
#endif /* LLB200_H */
