// GPTQ column-block storage kernels: pack, unpack, reference-semantics linear.
// ref: /root/reference/lit_llama/quantization.py:339-421 (ColBlockQuantizedLinear).
#include <stdarg.h>

#include "common.cuh"

namespace llb {

static thread_local char g_err[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}
int cuda_fail(cudaError_t e, const char* what) {
  set_error("CUDA error %d (%s) at %s", (int)e, cudaGetErrorString(e), what);
  return LL_ERR_CUDA;
}
int sm_count() {
  static int n = 0;
  if (n == 0) {
    int dev = 0;
    if (cudaGetDevice(&dev) != cudaSuccess) return -1;
    if (cudaDeviceGetAttribute(&n, cudaDevAttrMultiProcessorCount, dev) != cudaSuccess) return -1;
  }
  return n;
}

namespace {

// ---------------------------------------------------------------- pack (row P)
// One thread per output byte; threads run along N (the contiguous dimension of qw).
template <typename W, typename P>
__global__ void pack_kernel(const W* __restrict__ w, const P* __restrict__ scales, const P* __restrict__ zeros,
                            uint8_t* __restrict__ qw, int N, int K, int bits, int tile, int G) {
  const int epb = 8 / bits;
  const int n = blockIdx.x * blockDim.x + threadIdx.x;
  const int kb = blockIdx.y;
  if (n >= N) return;
  const float maxq = (float)((1 << bits) - 1);
  uint32_t byte = 0;
  for (int nr = 0; nr < epb; ++nr) {
    const int k = kb * epb + nr;
    const int g = k / tile;
    const float s = dt<P>::ld(scales, (size_t)n * G + g);
    const float z = dt<P>::ld(zeros, (size_t)n * G + g);
    float v = dt<W>::ld(w, (size_t)n * K + k);
    v = dt<W>::rd(__fdiv_rn(v, s));   // weight[:, tile] /= scale   (in the weight dtype)
    v = dt<W>::rd(__fadd_rn(v, z));   // weight[:, tile] += zero
    v = fminf(fmaxf(v, 0.f), maxq);   // clamp_(0, 2^bits - 1)
    const uint32_t code = (v == v) ? (uint32_t)v : 0u;  // .to(uint8): truncation
    byte |= (code & 0xffu) << (nr * bits);
  }
  qw[(size_t)kb * N + n] = (uint8_t)byte;
}

// ---------------------------------------------------------------- unpack (row G)
// Tile [32 n][32 k] through shared memory: reads coalesced along N, writes coalesced along K.
template <typename O, typename P>
__global__ void get_weight_kernel(const uint8_t* __restrict__ qw, const P* __restrict__ scales,
                                  const P* __restrict__ zeros, O* __restrict__ out, int N, int K, int bits, int tile,
                                  int G) {
  __shared__ float t[32][33];
  const int epb = 8 / bits;
  const uint32_t mask = (1u << bits) - 1u;
  const int n0 = blockIdx.x * 32, k0 = blockIdx.y * 32;
  const int kb_per = 32 / epb;  // packed rows covering 32 k
  // load: threadIdx.x along n, threadIdx.y (8) along packed rows
  for (int r = threadIdx.y; r < kb_per; r += blockDim.y) {
    const int n = n0 + threadIdx.x;
    const int kb = k0 / epb + r;
    uint32_t byte = 0;
    if (n < N && kb * epb < K) byte = qw[(size_t)kb * N + n];
    for (int nr = 0; nr < epb; ++nr) t[threadIdx.x][r * epb + nr] = (float)((byte >> (nr * bits)) & mask);
  }
  __syncthreads();
  for (int r = threadIdx.y; r < 32; r += blockDim.y) {
    const int n = n0 + r, k = k0 + threadIdx.x;
    if (n < N && k < K) {
      const int g = k / tile;
      const float s = dt<P>::ld(scales, (size_t)n * G + g);
      const float z = dt<P>::ld(zeros, (size_t)n * G + g);
      float v = dt<O>::rd(t[r][threadIdx.x]);
      v = dt<O>::rd(__fsub_rn(v, z));
      v = dt<O>::rd(__fmul_rn(v, s));
      dt<O>::st(out, (size_t)n * K + k, v);
    }
  }
}

// ---------------------------------------------------------------- unpack, int4 -> bf16 fast path
// Same arithmetic as get_weight_kernel (bit-exact get_weight(dtype=bf16)); tile = 64 columns x 128 k staged
// through shared memory, 16-byte coalesced loads of the packed rows, whole 128-byte lines stored per half-warp.
template <typename P>
__global__ void __launch_bounds__(256) get_weight_int4_bf16_kernel(const uint8_t* __restrict__ qw, const P* __restrict__ scales,
                                                                   const P* __restrict__ zeros, __nv_bfloat16* __restrict__ out,
                                                                   int N, int K, int tile, int G) {
  __shared__ __align__(16) uint8_t raw[64][80];  // [packed row][column], 16 B of padding per row
  const int n0 = blockIdx.x * 64, k0 = blockIdx.y * 128;
  {
    const int r = threadIdx.x >> 2, seg = threadIdx.x & 3;  // 64 packed rows x 4 segments of 16 columns
    const uint4 v = *reinterpret_cast<const uint4*>(qw + (size_t)(k0 / 2 + r) * N + n0 + seg * 16);
    *reinterpret_cast<uint4*>(&raw[r][seg * 16]) = v;
  }
  __syncthreads();
  // a thread produces one 16-byte piece (8 consecutive k of one column) per pass; a half-warp = one column's 128 k =
  // 256 contiguous bytes of the output row, so every store instruction writes whole lines (each thread writing 64 bytes
  // of its own row instead left every line to four separate instructions: 2.6 TB/s)
  const int piece = threadIdx.x & 15, cpass = threadIdx.x >> 4;  // 16 columns per pass
  const int kb = k0 + piece * 8;
  const int g = kb / tile;
#pragma unroll
  for (int pass = 0; pass < 4; ++pass) {
    const int nl = pass * 16 + cpass, n = n0 + nl;
    const float s = dt<P>::ld(scales, (size_t)n * G + g);
    const float z = dt<P>::ld(zeros, (size_t)n * G + g);
    uint32_t o[4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      const uint32_t b = raw[piece * 4 + i][nl];
      const float lo = bf16r(__fmul_rn(bf16r(__fsub_rn((float)(b & 15u), z)), s));
      const float hi = bf16r(__fmul_rn(bf16r(__fsub_rn((float)(b >> 4), z)), s));
      o[i] = (uint32_t)f32_to_bf16_bits(lo) | ((uint32_t)f32_to_bf16_bits(hi) << 16);
    }
    *reinterpret_cast<uint4*>(out + (size_t)n * K + kb) = make_uint4(o[0], o[1], o[2], o[3]);
  }
}

// ---------------------------------------------------------------- reference-semantics linear (row F)
// Block = 32 columns x 8 k-lanes; up to 8 rows of x per pass.  fp32 accumulation.
template <typename A, typename P>
__global__ void qlinear_simt_kernel(const A* __restrict__ x, int M, const uint8_t* __restrict__ qw,
                                    const P* __restrict__ scales, const P* __restrict__ zeros,
                                    const A* __restrict__ bias, A* __restrict__ y, int N, int K, int bits, int tile,
                                    int G) {
  constexpr int MT = 8;
  __shared__ float red[8][MT][33];
  const int epb = 8 / bits;
  const uint32_t mask = (1u << bits) - 1u;
  const int n = blockIdx.x * 32 + threadIdx.x;
  const int m0 = blockIdx.y * MT;
  const int ky = threadIdx.y;
  float acc[MT];
#pragma unroll
  for (int m = 0; m < MT; ++m) acc[m] = 0.f;
  if (n < N) {
    const int KB = K / epb;
    for (int kb = ky; kb < KB; kb += 8) {
      const uint32_t byte = qw[(size_t)kb * N + n];
      for (int nr = 0; nr < epb; ++nr) {
        const int k = kb * epb + nr;
        const int g = k / tile;
        const float s = dt<P>::ld(scales, (size_t)n * G + g);
        const float z = dt<P>::ld(zeros, (size_t)n * G + g);
        float w = (float)((byte >> (nr * bits)) & mask);
        w = dt<A>::rd(__fsub_rn(w, z));
        w = dt<A>::rd(__fmul_rn(w, s));
#pragma unroll
        for (int m = 0; m < MT; ++m)
          if (m0 + m < M) acc[m] = fmaf(dt<A>::ld(x, (size_t)(m0 + m) * K + k), w, acc[m]);
      }
    }
  }
#pragma unroll
  for (int m = 0; m < MT; ++m) red[ky][m][threadIdx.x] = acc[m];
  __syncthreads();
  if (n < N && ky < MT && m0 + ky < M) {
    float v = 0.f;
#pragma unroll
    for (int j = 0; j < 8; ++j) v += red[j][ky][threadIdx.x];
    if (bias) v += dt<A>::ld(bias, n);
    dt<A>::st(y, (size_t)(m0 + ky) * N + n, v);
  }
}

int check_quant_args(int N, int K, int bits, int tile_cols, int* tile, int* G) {
  LL_CHECK_ARG(bits == 1 || bits == 2 || bits == 4 || bits == 8, "bits must be 1, 2, 4 or 8 (got %d)", bits);
  LL_CHECK_ARG(N > 0 && K > 0, "N and K must be positive (N=%d K=%d)", N, K);
  LL_CHECK_ARG(K % (8 / bits) == 0, "in_features (%d) must be a multiple of entries-per-byte (%d)", K, 8 / bits);
  *tile = tile_cols == -1 ? K : tile_cols;
  LL_CHECK_ARG(*tile > 0, "tile_cols must be -1 or positive (got %d)", tile_cols);
  *G = (K + *tile - 1) / *tile;
  return LL_OK;
}

}  // namespace
}  // namespace llb

using namespace llb;

extern "C" {

const char* ll_last_error(void) { return g_err; }
int ll_abi_version(void) { return LL_ABI_VERSION; }
int ll_sm_count(void) {
  int n = sm_count();
  if (n < 0) set_error("no CUDA device");
  return n;
}

int ll_pack_weight(const void* weight, int wdt, const void* scales, const void* zeros, int pdt, uint8_t* qw, int N,
                   int K, int bits, int tile_cols, ll_stream stream) {
  int tile, G;
  if (int r = check_quant_args(N, K, bits, tile_cols, &tile, &G)) return r;
  LL_CHECK_ARG(weight && scales && zeros && qw, "null pointer");
  const int epb = 8 / bits;
  dim3 block(128), grid((N + 127) / 128, K / epb);
  LL_CHECK_ARG(grid.y <= 65535, "K/epb too large for this kernel (%d)", K / epb);
  cudaStream_t st = (cudaStream_t)stream;
#define PACK(WT, PT) pack_kernel<WT, PT><<<grid, block, 0, st>>>((const WT*)weight, (const PT*)scales, (const PT*)zeros, qw, N, K, bits, tile, G)
  if (wdt == LL_F32 && pdt == LL_F32) PACK(float, float);
  else if (wdt == LL_F32 && pdt == LL_BF16) PACK(float, __nv_bfloat16);
  else if (wdt == LL_BF16 && pdt == LL_F32) PACK(__nv_bfloat16, float);
  else if (wdt == LL_BF16 && pdt == LL_BF16) PACK(__nv_bfloat16, __nv_bfloat16);
  else {
    set_error("pack_weight: unsupported dtypes weight=%d params=%d", wdt, pdt);
    return LL_ERR_UNSUPPORTED;
  }
#undef PACK
  LL_LAUNCH_OK("pack_kernel");
  return LL_OK;
}

int ll_get_weight(const uint8_t* qw, const void* scales, const void* zeros, int pdt, void* out, int odt, int N, int K,
                  int bits, int tile_cols, ll_stream stream) {
  int tile, G;
  if (int r = check_quant_args(N, K, bits, tile_cols, &tile, &G)) return r;
  LL_CHECK_ARG(qw && scales && zeros && out, "null pointer");
  cudaStream_t st = (cudaStream_t)stream;
  if (bits == 4 && odt == LL_BF16 && N % 64 == 0 && K % 128 == 0 && tile % 32 == 0 && K / 128 <= 65535 &&
      ((uintptr_t)qw & 15) == 0 && ((uintptr_t)out & 15) == 0 && (pdt == LL_BF16 || pdt == LL_F32)) {
    dim3 fgrid(N / 64, K / 128);
    if (pdt == LL_BF16)
      get_weight_int4_bf16_kernel<__nv_bfloat16><<<fgrid, 256, 0, st>>>(qw, (const __nv_bfloat16*)scales, (const __nv_bfloat16*)zeros, (__nv_bfloat16*)out, N, K, tile, G);
    else
      get_weight_int4_bf16_kernel<float><<<fgrid, 256, 0, st>>>(qw, (const float*)scales, (const float*)zeros, (__nv_bfloat16*)out, N, K, tile, G);
    LL_LAUNCH_OK("get_weight_int4_bf16_kernel");
    return LL_OK;
  }
  dim3 block(32, 8), grid((N + 31) / 32, (K + 31) / 32);
  LL_CHECK_ARG(grid.y <= 65535, "K too large for this kernel (%d)", K);
#define GETW(OT, PT) get_weight_kernel<OT, PT><<<grid, block, 0, st>>>(qw, (const PT*)scales, (const PT*)zeros, (OT*)out, N, K, bits, tile, G)
  if (odt == LL_F32 && pdt == LL_F32) GETW(float, float);
  else if (odt == LL_F32 && pdt == LL_BF16) GETW(float, __nv_bfloat16);
  else if (odt == LL_BF16 && pdt == LL_F32) GETW(__nv_bfloat16, float);
  else if (odt == LL_BF16 && pdt == LL_BF16) GETW(__nv_bfloat16, __nv_bfloat16);
  else if (odt == LL_F16 && pdt == LL_F32) GETW(__half, float);
  else if (odt == LL_F16 && pdt == LL_BF16) GETW(__half, __nv_bfloat16);
  else {
    set_error("get_weight: unsupported dtypes out=%d params=%d", odt, pdt);
    return LL_ERR_UNSUPPORTED;
  }
#undef GETW
  LL_LAUNCH_OK("get_weight_kernel");
  return LL_OK;
}

int ll_qlinear_simt(const void* x, int adt, int M, const uint8_t* qw, const void* scales, const void* zeros, int pdt,
                    const void* bias, void* y, int N, int K, int bits, int tile_cols, ll_stream stream) {
  int tile, G;
  if (int r = check_quant_args(N, K, bits, tile_cols, &tile, &G)) return r;
  LL_CHECK_ARG(x && qw && scales && zeros && y, "null pointer");
  LL_CHECK_ARG(M >= 0, "M must be >= 0");
  if (M == 0) return LL_OK;
  dim3 block(32, 8), grid((N + 31) / 32, (M + 7) / 8);
  LL_CHECK_ARG(grid.y <= 65535, "M too large for this kernel (%d)", M);
  cudaStream_t st = (cudaStream_t)stream;
#define QL(AT, PT) qlinear_simt_kernel<AT, PT><<<grid, block, 0, st>>>((const AT*)x, M, qw, (const PT*)scales, (const PT*)zeros, (const AT*)bias, (AT*)y, N, K, bits, tile, G)
  if (adt == LL_F32 && pdt == LL_F32) QL(float, float);
  else if (adt == LL_F32 && pdt == LL_BF16) QL(float, __nv_bfloat16);
  else if (adt == LL_BF16 && pdt == LL_F32) QL(__nv_bfloat16, float);
  else if (adt == LL_BF16 && pdt == LL_BF16) QL(__nv_bfloat16, __nv_bfloat16);
  else {
    set_error("qlinear_simt: unsupported dtypes act=%d params=%d", adt, pdt);
    return LL_ERR_UNSUPPORTED;
  }
#undef QL
  LL_LAUNCH_OK("qlinear_simt_kernel");
  return LL_OK;
}

}  // extern "C"
