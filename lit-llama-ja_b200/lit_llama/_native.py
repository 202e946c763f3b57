"""ctypes binding of the llb200 C-ABI library (include/llb200.h).

There is no CPU or PyTorch fallback behind these calls: if ``libllb200.so`` is missing the import
of this module raises, and every wrapper raises ``RuntimeError`` with the library's message when a
call fails.  Build the library with ``python lit-llama-ja_b200/build.py``.
"""
from __future__ import annotations

import ctypes as C
from pathlib import Path

import torch

import os

# LLB200_LIB: a tuning build of the same library (build.py --variant); the default is the in-tree libllb200.so
_LIB_PATH = Path(os.environ.get("LLB200_LIB") or Path(__file__).resolve().parent / "libllb200.so")
if not _LIB_PATH.exists():
    raise ImportError(
        f"{_LIB_PATH} not found: the CUDA kernels are not built. Run `python lit-llama-ja_b200/build.py` "
        "(nvcc, sm_100a). There is no CPU fallback."
    )
lib = C.CDLL(str(_LIB_PATH))

LL_F32, LL_BF16, LL_F16 = 0, 1, 2
EPI_STORE, EPI_RESIDUAL, EPI_ROPE_KV, EPI_SWIGLU, EPI_PUSH = 0, 1, 2, 3, 4
GEMV_PDL = 1
GEMV_ODD, GEMV_EVEN = 2, 4

_DT = {torch.float32: LL_F32, torch.bfloat16: LL_BF16, torch.float16: LL_F16}

vp, i32, sz_t = C.c_void_p, C.c_int, C.c_size_t


class GemvArgs(C.Structure):
    """struct ll_gemv_args (include/llb200.h)."""
    _fields_ = [
        ("w", vp), ("w2", vp),
        ("N", C.c_int32), ("K", C.c_int32), ("group", C.c_int32), ("M", C.c_int32),
        ("x", vp), ("norm_weight", vp), ("eps", C.c_float),
        ("epilogue", C.c_int32), ("flags", C.c_int32),
        ("y", vp), ("bias", vp),
        ("rope", vp), ("pos", vp), ("k_cache", vp), ("v_cache", vp),
        ("n_head", C.c_int32), ("head_size", C.c_int32), ("cache_len", C.c_int32),
        ("grid", C.c_int32),
        ("push_dst", C.c_uint64 * 8), ("push_cnt", C.c_uint64 * 8), ("push_world", C.c_int32),
        ("stream_kind", C.c_int32),
        ("x_planes", vp), ("x_rec", vp),
    ]


def _proto(name, argtypes, restype=C.c_int):
    fn = getattr(lib, name)
    fn.argtypes = argtypes
    fn.restype = restype
    return fn


# every symbol include/llb200.h declares
EXPORTS = {
    "ll_last_error": ([], C.c_char_p),
    "ll_abi_version": ([], C.c_int),
    "ll_sm_count": ([], C.c_int),
    "ll_pack_weight": ([vp, i32, vp, vp, i32, vp, i32, i32, i32, i32, vp], C.c_int),
    "ll_get_weight": ([vp, vp, vp, i32, vp, i32, i32, i32, i32, i32, vp], C.c_int),
    "ll_qlinear_simt": ([vp, i32, i32, vp, vp, vp, i32, vp, vp, i32, i32, i32, i32, vp], C.c_int),
    "ll_int4_stream_bytes": ([i32, i32, i32], sz_t),
    "ll_int4_prepare": ([vp, vp, vp, i32, vp, i32, i32, i32, i32, vp, vp], C.c_int),
    "ll_int4_stream_kind": ([i32, i32, i32, i32], C.c_int),
    "ll_q8_stream_bytes": ([i32, i32, i32], sz_t),
    "ll_q8_gemv_supported": ([i32, i32, i32, i32, i32], C.c_int),
    "ll_q8_prepare": ([vp, vp, vp, i32, vp, i32, i32, i32, vp, vp], C.c_int),
    "ll_int4_gemv_plan": ([i32, i32, i32, i32, i32, i32, C.POINTER(sz_t), C.POINTER(C.c_int32)], C.c_int),
    "ll_int4_gemv": ([C.POINTER(GemvArgs), vp], C.c_int),
    "ll_gemm_bf16_tc": ([vp, vp, vp, vp, i32, i32, i32, vp], C.c_int),
    "ll_int4_gemm_tc_scratch_bytes": ([i32, i32, i32], sz_t),
    "ll_int4_gemm_tc": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, vp], C.c_int),
    "ll_int4_gemm_tc_static": ([vp, vp, vp, vp, i32, i32, i32, i32, vp], C.c_int),
    "ll_int8_quantize_weight": ([vp, i32, vp, vp, i32, i32, vp], C.c_int),
    "ll_int8_workspace_bytes": ([i32, i32, i32], sz_t),
    "ll_int8_linear": ([vp, i32, vp, vp, vp, vp, vp, i32, i32, i32, C.c_float, vp, sz_t, vp], C.c_int),
    "ll_gptq_block": ([vp, i32, vp, i32, vp, vp, i32, vp, i32, vp, vp, i32, i32, i32, i32, i32, vp], C.c_int),
    "ll_sample_topk": ([vp, i32, C.c_float, i32, vp, vp, vp, vp, vp, vp], C.c_int),
    "ll_sample_advance_embed": ([vp, i32, C.c_float, i32, vp, vp, vp, vp, vp, i32, i32, vp, vp, i32, vp], C.c_int),
    "ll_tp_reduce_add": ([vp, vp, vp, vp, i32, i32, i32, vp], C.c_int),
    "ll_tp_allreduce_add": ([vp, vp, vp, vp, i32, i32, i32, i32, vp], C.c_int),
    "ll_tp_ll_recv_bytes": ([i32, i32], sz_t),
    "ll_tp_ll_allreduce_add": ([vp, vp, vp, vp, i32, i32, i32, i32, vp], C.c_int),
    "ll_rmsnorm": ([vp, vp, vp, i32, i32, i32, C.c_float, vp], C.c_int),
    "ll_embed_ssq": ([vp, vp, vp, vp, i32, i32, vp], C.c_int),
    "ll_rope_kv_append": ([vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp], C.c_int),
    "ll_decode_attn": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp, i32, vp], C.c_int),
    "ll_decode_attn_scratch_bytes": ([i32, i32, i32, i32], sz_t),
    "ll_decode_attn_digits_supported": ([i32, i32, i32, i32, i32], C.c_int),
    "ll_decode_attn_digits": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, vp, vp, i32, vp], C.c_int),
    "ll_silu_mul": ([vp, vp, vp, i32, sz_t, vp], C.c_int),
    "ll_adapter_attn_add": ([vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, vp], C.c_int),
    "ll_bias_scale": ([vp, vp, vp, vp, i32, sz_t, i32, vp], C.c_int),
    "ll_add": ([vp, vp, vp, i32, sz_t, vp], C.c_int),
    "ll_advance_pos": ([vp, i32, i32, vp], C.c_int),
    "ll_set_pos": ([vp, vp, i32, i32, vp], C.c_int),
    "ll_apply_rope": ([vp, vp, vp, i32, i32, i32, i32, i32, vp], C.c_int),
}
_fn = {name: _proto(name, a, r) for name, (a, r) in EXPORTS.items()}

ABI_VERSION = 11
if _fn["ll_abi_version"]() != ABI_VERSION:
    raise ImportError("libllb200.so ABI version mismatch; rebuild with lit-llama-ja_b200/build.py")


def last_error() -> str:
    return (_fn["ll_last_error"]() or b"").decode()


def check(rc: int, what: str) -> None:
    if rc != 0:
        raise RuntimeError(f"{what} failed ({rc}): {last_error()}")


LAUNCHES = 0   # kernels launched through this binding (bench.py's gpu_launches claim)


def call(name: str, *args) -> None:
    global LAUNCHES
    check(_fn[name](*args), name)
    LAUNCHES += 1


def call_ret(name: str, *args):
    """A query that returns a value (no launch, no error code)."""
    return _fn[name](*args)


def require_cuda(*tensors: torch.Tensor) -> None:
    for t in tensors:
        if t is not None and t.device.type != "cuda":
            raise RuntimeError(
                "lit-llama-ja_b200 computes on CUDA (sm_100a) only - there is no CPU fallback; "
                f"got a tensor on {t.device}"
            )


def dtype_code(dt: torch.dtype) -> int:
    try:
        return _DT[dt]
    except KeyError:
        raise RuntimeError(f"unsupported dtype {dt}") from None


def ptr(t) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def gemv_plan(N: int, K: int, group: int, nmat: int, M: int, grid: int = 0):
    """(shared-memory bytes, CTAs) of the launch the GEMV would use; raises if the shape is unsupported."""
    sm, ctas = sz_t(), C.c_int32()
    check(_fn["ll_int4_gemv_plan"](N, K, group, nmat, M, grid, C.byref(sm), C.byref(ctas)), "ll_int4_gemv_plan")
    return sm.value, ctas.value


def gemv_supported(N: int, K: int, group: int, nmat: int, M: int) -> bool:
    return _fn["ll_int4_gemv_plan"](N, K, group, nmat, M, 0, None, None) == 0


STREAM_BF16, STREAM_U8, STREAM_W8 = 0, 1, 2


def stream_kind(N: int, K: int, group: int, M: int) -> int:
    """The weight-stream kind the int4 GEMV wants for this shape and batch (LL_STREAM_*)."""
    return _fn["ll_int4_stream_kind"](N, K, group, M)


def stream_bytes(N: int, K: int, group: int) -> int:
    return _fn["ll_int4_stream_bytes"](N, K, group)


def gemv(args: GemvArgs) -> None:
    global LAUNCHES
    LAUNCHES += 1
    check(_fn["ll_int4_gemv"](C.byref(args), stream()), "ll_int4_gemv")


def attn_scratch_bytes(B, n_head, hs, splits) -> int:
    return _fn["ll_decode_attn_scratch_bytes"](B, n_head, hs, splits)


class Workspace:
    """Per-(device, stream) scratch for the attention split states.  Kernels on one stream are ordered,
    so one buffer is shared by every layer; counters are zeroed once and the kernels leave them zero."""

    _cache: dict = {}

    def __init__(self, device):
        self.device = device
        self.attn_scratch = torch.empty(1 << 18, dtype=torch.float32, device=device)
        self.attn_counters = torch.zeros(4096, dtype=torch.int32, device=device)

    @classmethod
    def get(cls, device) -> "Workspace":
        key = (device.index if device.index is not None else torch.cuda.current_device(),
               torch.cuda.current_stream().cuda_stream)
        ws = cls._cache.get(key)
        if ws is None:
            ws = cls._cache[key] = Workspace(device)
        return ws

    def ensure(self, attn_bytes=0, attn_counters=0):
        if attn_bytes > self.attn_scratch.numel() * 4:
            self.attn_scratch = torch.empty(attn_bytes // 4 + 1, dtype=torch.float32, device=self.device)
        if attn_counters > self.attn_counters.numel():
            self.attn_counters = torch.zeros(attn_counters, dtype=torch.int32, device=self.device)
        return self
