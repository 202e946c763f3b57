"""Quantized Linear modules backed by hand-written sm_100a kernels (libllb200.so).

Same module surface as the reference's ``lit_llama/quantization.py`` (buffer names, shapes,
strides, state-dict keys, constructor protocol) so GPTQ checkpoints written by the reference's
``quantize/gptq.py`` load unchanged.  All arithmetic runs on CUDA through the C-ABI in
``include/llb200.h``; there is no Triton, no bitsandbytes and no CPU fallback.
"""
from __future__ import annotations

import math

import torch

from . import _native as N


def _entries_per_byte(bits: int) -> int:
    epb = 8 // bits
    assert epb > 0 and epb * bits == 8, f"bits must be 1, 2, 4 or 8, got {bits}"
    return epb


class ColBlockQuantizedLinear(torch.nn.Module):
    """GPTQ column-block quantized Linear (ref: lit_llama/quantization.py:338-421).

    Buffers (identical to the reference): ``quant_weight`` uint8 ``[out, in/epb]`` with strides
    ``(1, out)`` (memory is ``[in/epb][out]``), ``scales`` / ``zeros`` ``[out, ceil(in/tile)]`` in
    the default dtype, ``bias`` ``[out]`` or ``None``.  ``bias`` has no default and must be a bool.
    """

    def __init__(self, in_features, out_features, bias: bool, *, bits, tile_cols):
        super().__init__()
        self.in_features = in_features
        self.out_features = out_features
        self.bits = bits
        self.entries_per_byte = _entries_per_byte(bits)
        assert in_features % self.entries_per_byte == 0
        self.tile_cols = in_features if tile_cols == -1 else tile_cols
        packed_t = torch.empty((in_features // self.entries_per_byte, out_features), dtype=torch.uint8)
        self.register_buffer("quant_weight", packed_t.t())  # logical [out, in/epb], N-fastest storage
        n_groups = (in_features + self.tile_cols - 1) // self.tile_cols
        self.register_buffer("scales", torch.empty((out_features, n_groups)))
        self.register_buffer("zeros", torch.empty((out_features, n_groups)))
        assert isinstance(bias, bool)
        self.register_buffer("bias", torch.empty((out_features,)) if bias else None)
        self._streams = {}     # stream kind -> (key, buffer, ok): fragment-ordered repacks of quant_weight/scales/zeros (lazy)

    # -- state-dict: accept a checkpoint whose group count differs from the constructor's
    # (``quantization("gptq.int4")`` always builds per-row modules, ref utils.py:179-191; a
    # group-128 checkpoint carries ``scales [out, in/128]``).
    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        s = state_dict.get(prefix + "scales")
        if s is not None and tuple(s.shape) != tuple(self.scales.shape) and s.shape[0] == self.out_features:
            g = s.shape[1]
            if g > 0 and self.in_features % g == 0:
                self.tile_cols = self.in_features // g
                self.scales = torch.empty((self.out_features, g), dtype=self.scales.dtype, device=self.scales.device)
                self.zeros = torch.empty_like(self.scales)
        super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)
        self._streams = {}

    def _apply(self, fn, *args, **kwargs):
        # keep the reference's (1, out) strides of quant_weight across .to()/.cuda()
        stride_ok = self.quant_weight.stride() == (1, self.out_features)
        super()._apply(fn, *args, **kwargs)
        if stride_ok and self.quant_weight.stride() != (1, self.out_features):
            self.quant_weight = self.quant_weight.t().contiguous().t()
        self._streams = {}
        return self

    def _packed_ptr_ok(self):
        qw = self.quant_weight
        if qw.stride() != (1, self.out_features):
            self.quant_weight = qw = qw.t().contiguous().t()
        return qw

    # ------------------------------------------------------------------ rows P, G
    def pack_weight(self, weight):
        """ref: quantization.py:374-388 - bit-exact (``/= scale``, ``+= zero``, clamp, truncate)."""
        qw = self._packed_ptr_ok()
        N.require_cuda(qw, self.scales, self.zeros)
        weight = weight.to(device=qw.device).contiguous()
        assert weight.shape == (self.out_features, self.in_features)
        assert self.scales.dtype == self.zeros.dtype
        N.call("ll_pack_weight", N.ptr(weight), N.dtype_code(weight.dtype), N.ptr(self.scales.contiguous()),
               N.ptr(self.zeros.contiguous()), N.dtype_code(self.scales.dtype), N.ptr(qw), self.out_features,
               self.in_features, self.bits, self.tile_cols, N.stream())
        self._streams = {}

    def get_weight(self, dtype=torch.float):
        """ref: quantization.py:390-409 - bit-exact de-quantisation to ``[out, in]`` in ``dtype``."""
        qw = self._packed_ptr_ok()
        N.require_cuda(qw, self.scales, self.zeros)
        out = torch.empty((self.out_features, self.in_features), device=qw.device, dtype=dtype)
        N.call("ll_get_weight", N.ptr(qw), N.ptr(self.scales.contiguous()), N.ptr(self.zeros.contiguous()),
               N.dtype_code(self.scales.dtype), N.ptr(out), N.dtype_code(dtype), self.out_features,
               self.in_features, self.bits, self.tile_cols, N.stream())
        return out

    # ------------------------------------------------------------------ int4 GEMV weight stream
    def _fast_eligible(self) -> bool:
        if self.scales.dtype != self.zeros.dtype:
            return False
        if self.bits == 8:      # gptq.int8: stream kind 2 of the batch-1 GEMV
            return N.call_ret("ll_q8_stream_bytes", self.out_features, self.in_features, self.tile_cols) > 0
        return self.bits == 4 and N.stream_bytes(self.out_features, self.in_features, self.tile_cols) > 0

    def gemv_ok(self, M: int, nmat: int = 1) -> bool:
        """The tensor-core GEMV takes this layer at batch ``M`` (4-bit: M <= 8; 8-bit weights: batch 1)."""
        if self.bits == 8:
            return bool(N.call_ret("ll_q8_gemv_supported", self.out_features, self.in_features, self.tile_cols, nmat, M))
        return self.bits == 4 and N.gemv_supported(self.out_features, self.in_features, self.tile_cols, nmat, M)

    def _kind(self, M: int) -> int:
        return N.STREAM_W8 if self.bits == 8 else N.stream_kind(self.out_features, self.in_features, self.tile_cols, M)

    def _state_key(self):
        return (self.quant_weight.data_ptr(), self.scales.data_ptr(), self.zeros.data_ptr(),
                self.quant_weight._version, self.scales._version, self.zeros._version)

    def prepare(self, M: int = 1):
        """Build (once per stream kind) the fragment-ordered weight stream the int4 GEMV reads at batch ``M`` (a derived
        cache of ``quant_weight``/``scales``/``zeros``; the checkpoint-layout buffers stay the state): batch 1 with
        128-multiple scale groups takes the integer-MMA kernel's layout, everything else the bf16 kernel's
        (``ll_int4_stream_kind``).  Returns the stream, or ``None`` when the layer cannot take the tensor-core GEMV
        (shape, or scales / zeros not exactly representable in bf16)."""
        if not self._fast_eligible() or (self.bits == 8 and not self.gemv_ok(M)):
            return None
        return self._prepare_kind(self._kind(M))

    def _prepare_kind(self, kind: int):
        ent = self._streams.get(kind)
        if ent is not None and ent[0] == self._state_key():
            return ent[1] if ent[2] else None
        qw = self._packed_ptr_ok()
        dev = self.scales.device
        status = torch.zeros(1, dtype=torch.int32, device=dev)
        if kind == N.STREAM_W8:
            nbytes = N.call_ret("ll_q8_stream_bytes", self.out_features, self.in_features, self.tile_cols)
            buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            N.call("ll_q8_prepare", N.ptr(qw), N.ptr(self.scales.contiguous()), N.ptr(self.zeros.contiguous()),
                   N.dtype_code(self.scales.dtype), N.ptr(buf), self.out_features, self.in_features, self.tile_cols,
                   N.ptr(status), N.stream())
        else:
            nbytes = N.stream_bytes(self.out_features, self.in_features, self.tile_cols)
            buf = torch.empty(nbytes, dtype=torch.uint8, device=dev)
            N.call("ll_int4_prepare", N.ptr(qw), N.ptr(self.scales.contiguous()), N.ptr(self.zeros.contiguous()),
                   N.dtype_code(self.scales.dtype), N.ptr(buf), self.out_features, self.in_features, self.tile_cols, kind,
                   N.ptr(status), N.stream())
        ok = int(status.item()) == 0
        self._streams[kind] = (self._state_key(), buf, ok)
        return buf if ok else None

    FUSED_GEMM_MAX_M = 128       # rows one ll_int4_gemm_tc call takes
    # rows up to which the fused kernel is looped over 128-row blocks (each block streams the packed weights again) instead of
    # unpacking W to bf16 once and running the plain tcgen05 GEMM (measured cross-over: profiles/r02_fused_gemm.md)
    FUSED_GEMM_LOOP_MAX_M = int(__import__("os").environ.get("LLB200_FUSED_GEMM_MAX_M", 256))

    def _fused_gemm_eligible(self, x) -> bool:
        """9 <= M <= 128 bf16 rows on a layer whose scale groups are multiples of 128: the fused tcgen05 de-quantise + GEMM."""
        return (self.bits == 4 and x.dtype == torch.bfloat16 and self.scales.dtype == torch.bfloat16
                and self.out_features % 128 == 0 and self.tile_cols % 128 == 0 and self.in_features % self.tile_cols == 0
                and (self.bias is None or self.bias.dtype == torch.bfloat16))

    def stream_for(self, M: int):
        """(prepared stream, its kind) for batch ``M`` WITHOUT building it (safe under CUDA-graph capture), or (None, kind)."""
        kind = self._kind(M)
        ent = self._streams.get(kind)
        return ((ent[1] if ent is not None and ent[2] else None), kind)

    # ------------------------------------------------------------------ row F
    def forward(self, inp):
        """ref: quantization.py:411-421.  M <= 8 bf16 rows: TMA-streamed int4 GEMV (exact (q - zero) * scale,
        fp32 accumulation); small M in other dtypes / bit widths: reference-semantics SIMT kernel;
        large M: dequant-GEMM."""
        qw = self._packed_ptr_ok()
        N.require_cuda(inp, qw)
        K, Nout = self.in_features, self.out_features
        assert inp.shape[-1] == K, "incompatible dimensions"
        x = inp.reshape(-1, K)
        if not x.is_contiguous():
            x = x.contiguous()
        M = x.shape[0]
        out_shape = inp.shape[:-1] + (Nout,)
        if M == 0:
            return inp.new_empty(out_shape)
        y = torch.empty((M, Nout), device=inp.device, dtype=inp.dtype)
        fast = inp.dtype == torch.bfloat16 and M <= 8 and self.gemv_ok(M)
        if fast and torch.cuda.is_current_stream_capturing():
            fast = self.stream_for(M)[0] is not None   # the stream must have been prepared before capture
        elif fast:
            fast = self.prepare(M) is not None
        if fast:
            self._gemv(x, y)
        elif M <= max(self.FUSED_GEMM_MAX_M, self.FUSED_GEMM_LOOP_MAX_M) and self.FUSED_GEMM_MAX_M > 0 \
                and self._fused_gemm_eligible(x) and self._fused_gemm(x, y):
            pass
        elif M <= 16:
            assert inp.dtype in (torch.float32, torch.bfloat16)
            N.call("ll_qlinear_simt", N.ptr(x), N.dtype_code(x.dtype), M, N.ptr(qw), N.ptr(self.scales.contiguous()),
                   N.ptr(self.zeros.contiguous()), N.dtype_code(self.scales.dtype), N.ptr(self.bias), N.ptr(y),
                   Nout, K, self.bits, self.tile_cols, N.stream())
        else:
            y = dequant_gemm(self, x)
        return y.reshape(out_shape)

    def _fused_gemm(self, x, y) -> bool:
        """``ll_int4_gemm_tc``: weights streamed once in packed form, de-quantised in the main loop (False if the stream
        cannot be built here, e.g. under CUDA-graph capture before it was prepared)."""
        ent = self._streams.get(N.STREAM_U8)
        if ent is not None and ent[0] == self._state_key():
            w = ent[1] if ent[2] else None
        elif torch.cuda.is_current_stream_capturing():
            return False
        else:
            w = self._prepare_kind(N.STREAM_U8)
        if w is None:
            return False
        step = self.FUSED_GEMM_MAX_M
        for r0 in range(0, x.shape[0], step):
            xb, yb = x[r0:r0 + step], y[r0:r0 + step]
            M = xb.shape[0]
            nbytes = N.call_ret("ll_int4_gemm_tc_scratch_bytes", M, self.out_features, self.in_features)
            scratch = _int8_workspace(x.device, nbytes) if nbytes else None
            N.call("ll_int4_gemm_tc", N.ptr(w), N.ptr(xb), N.ptr(self.bias), N.ptr(yb), N.ptr(scratch), M, self.out_features,
                   self.in_features, self.tile_cols, N.stream())
        return True

    def _gemv(self, x, y, *, pdl=False, grid=0):
        a = N.GemvArgs()
        w, a.stream_kind = self.stream_for(x.shape[0])
        a.w = w.data_ptr()
        a.N, a.K, a.group, a.M = self.out_features, self.in_features, self.tile_cols, x.shape[0]
        a.x, a.y = x.data_ptr(), y.data_ptr()
        a.bias = 0 if self.bias is None else self.bias.data_ptr()
        a.epilogue, a.flags, a.grid = N.EPI_STORE, (N.GEMV_PDL if pdl else 0), grid
        N.gemv(a)


def dequant_gemm(lin: ColBlockQuantizedLinear, x: torch.Tensor) -> torch.Tensor:
    """Large-M path (prefill, batched decode): bit-exact de-quantisation of W into a bf16 scratch (the
    reference's ``get_weight``) followed by the tcgen05/TMEM tensor-core GEMM - the reference's fallback
    branch (quantization.py:420-421) on the 5th-generation tensor cores.  fp32 activations (tests scale
    only) use the library GEMM."""
    w = lin.get_weight(dtype=x.dtype)
    if x.dtype != torch.bfloat16 or lin.in_features % 8 != 0:
        return torch.nn.functional.linear(x, w, lin.bias)
    y = torch.empty((x.shape[0], lin.out_features), device=x.device, dtype=x.dtype)
    N.call("ll_gemm_bf16_tc", N.ptr(x), N.ptr(w), N.ptr(lin.bias), N.ptr(y), x.shape[0], lin.out_features,
           lin.in_features, N.stream())
    return y


def qlinear_4bit_weight(inp, weight, scales, zeros):
    """ref: quantization.py:282-331 (Triton launcher) - same signature: ``weight`` is the
    ``[N, K/2]`` uint8 buffer with strides (1, N), ``scales``/``zeros`` are ``[N, 1]``.  Bias is not
    applied, as in the reference."""
    assert scales.shape == (weight.shape[0], 1) and zeros.shape == (weight.shape[0], 1)
    K = weight.shape[1] * 2
    assert inp.shape[-1] == K, "incompatible dimensions"
    lin = ColBlockQuantizedLinear.__new__(ColBlockQuantizedLinear)
    torch.nn.Module.__init__(lin)
    lin.in_features, lin.out_features, lin.bits, lin.entries_per_byte, lin.tile_cols = K, weight.shape[0], 4, 2, K
    lin.quant_weight, lin.scales, lin.zeros, lin.bias = weight, scales, zeros, None
    lin._streams = {}
    return lin(inp)


class Linear8bitLt(torch.nn.Module):
    """LLM.int8() Linear (ref: quantization.py:36-75, which wraps ``bnb.nn.Linear8bitLt`` with
    ``has_fp16_weights=False, threshold=6.0``).  Same surface: called like ``nn.Linear(in, out, bias=True)``;
    ``weight`` is an int8 Parameter (row-major ``CB``) carrying the attributes ``CB`` and ``SCB`` (row absmax);
    loading a state dict takes the floating-point ``weight`` of an ordinary checkpoint and re-quantizes it
    (there is no int8 checkpoint format).  CUDA only.

    One deviation, on purpose: the reference's ``_load_from_state_dict`` picks the first key of the WHOLE
    state dict that ends in ``weight`` (quantization.py:52-55), which is only right for the first module
    loaded; this one looks up ``prefix + "weight"``."""

    threshold = 6.0

    def __init__(self, in_features, out_features, bias=True, device=None, dtype=None, **kwargs):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        dev = torch.device(device) if device is not None else torch.empty(0).device
        if dev.type != "cuda":
            raise RuntimeError("Linear8bitLt (llm.int8) needs a CUDA device: there is no CPU fallback")
        self.weight = torch.nn.Parameter(torch.zeros((out_features, in_features), dtype=torch.int8, device=dev),
                                         requires_grad=False)
        self.weight.CB = self.weight.data
        self.weight.SCB = torch.zeros(out_features, dtype=torch.float32, device=dev)
        self.bias = torch.nn.Parameter(torch.zeros(out_features, device=dev, dtype=dtype), requires_grad=False) if bias else None
        self._ws = None
        self._cbt = None

    def _load_from_state_dict(self, state_dict, prefix, local_metadata, strict, missing_keys, unexpected_keys, error_msgs):
        key = prefix + "weight"
        if key in state_dict:
            w = state_dict[key]
            if w.dtype == torch.int8:      # a state dict saved from this module: CB as is, SCB cannot be recovered
                raise RuntimeError("Linear8bitLt loads the floating-point weight of an ordinary checkpoint (ref quantization.py:50-65)")
            self._quantize_weight(w)
        elif strict:
            missing_keys.append(key)
        bkey = prefix + "bias"
        if self.bias is not None:
            if bkey in state_dict:
                self.bias.data.copy_(state_dict[bkey])
            elif strict:
                missing_keys.append(bkey)
        elif bkey in state_dict and strict:
            unexpected_keys.append(bkey)

    def _quantize_weight(self, weight: torch.Tensor) -> None:
        """ref: quantization.py:67-75 - ``weight.half()``, row absmax ``SCB``, ``CB = rint(127 W / SCB)``."""
        dev = self.weight.device
        w = weight.to(device=dev).contiguous()
        if w.dtype not in (torch.float32, torch.bfloat16, torch.float16):
            w = w.float()
        assert w.shape == (self.out_features, self.in_features)
        cb = torch.empty(w.shape, dtype=torch.int8, device=dev)
        scb = torch.empty(w.shape[0], dtype=torch.float32, device=dev)
        N.call("ll_int8_quantize_weight", N.ptr(w), N.dtype_code(w.dtype), N.ptr(cb), N.ptr(scb), w.shape[0], w.shape[1], N.stream())
        self.weight.data = cb
        self.weight.CB, self.weight.SCB = cb, scb
        # CB transposed ([K][N]): the outlier columns of a step are then contiguous rows (csrc/int8_linear.cu, finish kernel);
        # a second copy of the int8 weights, built once at load time (LLB200_INT8_TRANSPOSED=0: without it)
        self._cbt = cb.t().contiguous() if __import__("os").environ.get("LLB200_INT8_TRANSPOSED", "1") != "0" else None

    def forward(self, inp):
        N.require_cuda(inp, self.weight)
        K, Nout = self.in_features, self.out_features
        assert inp.shape[-1] == K, "incompatible dimensions"
        x = inp.reshape(-1, K).contiguous()
        M = x.shape[0]
        if M == 0:
            return inp.new_empty(inp.shape[:-1] + (Nout,))
        y = torch.empty((M, Nout), device=x.device, dtype=x.dtype)
        need = N._fn["ll_int8_workspace_bytes"](min(M, 64), Nout, K)
        ws = _int8_workspace(x.device, need)
        bias = None if self.bias is None else self.bias.to(x.dtype)
        cbt = getattr(self, "_cbt", None)
        if cbt is not None and (cbt.device != x.device or cbt.shape != (K, Nout)):
            cbt = None
        N.call("ll_int8_linear", N.ptr(x), N.dtype_code(x.dtype), N.ptr(self.weight.CB), N.ptr(cbt), N.ptr(self.weight.SCB), N.ptr(bias),
               N.ptr(y), M, Nout, K, float(self.threshold), N.ptr(ws), ws.numel(), N.stream())
        return y.reshape(inp.shape[:-1] + (Nout,))


_INT8_WS: dict = {}


def _int8_workspace(device, nbytes: int) -> torch.Tensor:
    """Per-(device, stream) scratch of the int8 linear (kernels on one stream are ordered)."""
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    ws = _INT8_WS.get(key)
    if ws is None or ws.numel() < nbytes:
        ws = _INT8_WS[key] = torch.empty(nbytes + 256, dtype=torch.uint8, device=device)
    return ws


# --------------------------------------------------------------------------------------- row Z
def find_params_weight(x: torch.Tensor, maxq: int):
    """ref: GPTQQuantizer.find_params_weight, quantization.py:474-511 (perchannel, asymmetric):
    per-row ``scale = (max(x,0) - min(x,0)) / maxq``, ``zero = round(-xmin / scale)``; an all-zero
    row uses (-1, +1).  Returns ``[rows, 1]`` tensors.  Device-agnostic torch ops (load-time)."""
    x = x.flatten(1)
    zero_t = torch.zeros(x.shape[0], device=x.device, dtype=x.dtype)
    xmin = torch.minimum(x.amin(1), zero_t)
    xmax = torch.maximum(x.amax(1), zero_t)
    dead = (xmin == 0) & (xmax == 0)
    xmin = torch.where(dead, torch.full_like(xmin, -1), xmin)
    xmax = torch.where(dead, torch.full_like(xmax, +1), xmax)
    # a tensor divisor: torch's CUDA kernels turn division by a Python scalar into a multiplication by its
    # reciprocal (1 ulp off the reference's CPU result); tensor / tensor is IEEE division
    scale = (xmax - xmin) / torch.full_like(xmax, float(maxq))
    zero = torch.round(-xmin / scale)
    return scale.reshape(-1, 1), zero.reshape(-1, 1)


def quantize_weight(x, scale, zero, maxq):
    """ref: GPTQQuantizer.quantize_weight, quantization.py:468-472 - snap to the grid, de-quantised."""
    q = torch.clamp(torch.round(x / scale) + zero, 0, maxq)
    return scale * (q - zero)


@torch.no_grad()
def rtn_quantize_(lin: ColBlockQuantizedLinear, weight: torch.Tensor) -> ColBlockQuantizedLinear:
    """Round-to-nearest quantisation of ``weight`` into ``lin`` (per group of ``lin.tile_cols``):
    parameters from ``find_params_weight``, on-grid weights from ``quantize_weight``, bytes from
    ``pack_weight`` - the grouped fixture of SURVEY 8c/8d, all on the module's device."""
    maxq = 2 ** lin.bits - 1
    w = weight.to(device=lin.quant_weight.device, dtype=torch.float32)
    n, k = w.shape
    t = lin.tile_cols
    assert k % t == 0
    wg = w.view(n, k // t, t)
    s, z = find_params_weight(wg.reshape(-1, t), maxq)
    s, z = s.view(n, k // t), z.view(n, k // t)
    lin.scales.copy_(s)
    lin.zeros.copy_(z)
    # de-quantised on-grid weights with the parameters as stored (possibly rounded to bf16)
    s_st, z_st = lin.scales.float().unsqueeze(-1), lin.zeros.float().unsqueeze(-1)
    q = quantize_weight(wg, s_st, z_st, maxq).view(n, k)
    lin.pack_weight(q.to(lin.scales.dtype))
    return lin


# --------------------------------------------------------------------------------------- row Z2: the GPTQ solver
@torch.no_grad()
def gptq_solve(W: torch.Tensor, Hinv: torch.Tensor, scales: torch.Tensor, zeros: torch.Tensor, *, bits: int,
               blocksize: int = 128, groupsize: int = -1):
    """The column walk of ``GPTQQuantizer.quantize`` (ref quantization.py:559-592) on the device.

    ``W`` [N, K] fp32 (modified in place), ``Hinv`` [K, K] fp32 upper Cholesky factor of the inverse Hessian,
    ``scales`` / ``zeros`` [N, G] fp32 (per-row parameters are read, group parameters are written).  Per block of
    ``blocksize`` columns one kernel (``ll_gptq_block``) walks the columns in the reference's order; the cross-block
    update is a plain GEMM.  Returns ``(Q, loss_per_row)``: on-grid weights [N, K] and sum of Losses per row."""
    N.require_cuda(W, Hinv, scales, zeros)
    assert W.dtype == Hinv.dtype == scales.dtype == zeros.dtype == torch.float32
    assert W.is_contiguous() and Hinv.is_contiguous() and scales.is_contiguous() and zeros.is_contiguous()
    rows, columns = W.shape
    if blocksize > 128:
        raise ValueError("gptq_solve: blocksize must be <= 128")
    if groupsize != -1 and (blocksize % groupsize != 0 or columns % groupsize != 0):
        raise ValueError("gptq_solve: groupsize must be -1 or divide blocksize and the column count")
    Q = torch.zeros_like(W)
    loss = torch.zeros(rows, dtype=torch.float32, device=W.device)
    G = scales.shape[1]
    for i1 in range(0, columns, blocksize):
        i2 = min(i1 + blocksize, columns)
        count = i2 - i1
        err = torch.empty((rows, count), dtype=torch.float32, device=W.device)
        N.call("ll_gptq_block", N.ptr(W), columns, N.ptr(Hinv), columns, N.ptr(scales), N.ptr(zeros), G, N.ptr(Q), columns,
               N.ptr(err), N.ptr(loss), rows, i1, count, groupsize, 2 ** bits - 1, N.stream())
        if i2 < columns:
            W[:, i2:] -= err @ Hinv[i1:i2, i2:]           # ref 592 (cuBLAS: a plain GEMM)
    return Q, loss / 2                                     # ref 590: Losses1 / 2


class GPTQQuantizer:
    """ref: lit_llama/quantization.py:424-614 (E. Frantar et al., GPTQ, arXiv:2210.17323) - same constructor,
    ``collect_input_stats`` forward-hook signature and ``quantize() -> (ColBlockQuantizedLinear, error)``.

    B200-native: the Hessian accumulates on the device (one GEMM per calibration batch), the damped inverse and its
    Cholesky factor come from cuSOLVER (torch.linalg), and the column walk runs as one kernel per 128-column block
    (``gptq_solve``) instead of one Python iteration per column.  The reference's ``groupsize != -1`` branch raises
    (573-577 assigns a [rows, 1] tensor into a [rows] slice); here the group's parameters are stored in column
    ``g`` as that code intends, so group-128 checkpoints can actually be written.  CUDA only."""

    def __init__(self, linear_module, *, bits, perchannel=True, sym=False, blocksize=128, percdamp=0.01, groupsize=-1,
                 actorder=False):
        assert isinstance(linear_module, torch.nn.Linear)
        self.linear_module = linear_module
        self.dev = linear_module.weight.device
        N.require_cuda(linear_module.weight)
        self.rows, self.columns = linear_module.weight.shape
        self.H = torch.zeros((self.columns, self.columns), device=self.dev)
        self.nsamples = 0
        self.bits = bits
        self.maxq = 2 ** bits - 1
        self.perchannel = perchannel
        self.sym = sym
        self.blocksize = blocksize
        self.percdamp = percdamp
        self.groupsize = groupsize
        self.actorder = actorder
        self.tile_cols = self.columns if groupsize == -1 else groupsize
        self.scales = torch.zeros((self.rows, (self.columns + self.tile_cols - 1) // self.tile_cols),
                                  dtype=linear_module.weight.dtype, device=self.dev)
        self.zeros = torch.zeros_like(self.scales)
        assert not (self.actorder and self.groupsize != -1), "The permutation trick does not work for grouped quantization"
        if not perchannel or sym:
            raise NotImplementedError("only the reference's defaults (perchannel=True, sym=False) are on the path")

    @staticmethod
    def quantize_weight(x, scale, zero, maxq):
        return quantize_weight(x, scale, zero, maxq)

    def find_params_weight(self, x):
        return find_params_weight(x, self.maxq)

    @torch.no_grad()
    def collect_input_stats(self, _1, inp, _2):
        """ref 513-527: running mean of 2 x x^T over the calibration batches (a forward hook)."""
        inp = inp[0].detach()
        self.last_inp = inp
        if inp.dim() == 2:
            inp = inp.unsqueeze(0)
        tmp = inp.shape[0]
        if inp.dim() == 3:
            inp = inp.reshape((-1, inp.shape[-1]))
        inp = inp.t()
        self.H *= self.nsamples / (self.nsamples + tmp)
        self.nsamples += tmp
        inp = math.sqrt(2 / self.nsamples) * inp.float()
        self.H += inp.matmul(inp.t())

    @torch.no_grad()
    def quantize(self):
        """ref 529-614."""
        W = self.linear_module.weight.detach().to(dtype=torch.float, copy=True)
        scale, zero = self.find_params_weight(W)
        scales = torch.zeros(self.scales.shape, dtype=torch.float32, device=self.dev)
        zeros = torch.zeros_like(scales)
        scales[:] = scale
        zeros[:] = zero

        H = self.H
        del self.H
        dead = torch.diag(H) == 0
        H[dead, dead] = 1
        W[:, dead] = 0
        if self.actorder:
            perm = torch.argsort(torch.diag(H), descending=True)
            W = W[:, perm].contiguous()
            H = H[perm][:, perm]
        damp = self.percdamp * torch.mean(torch.diag(H))
        diag = torch.arange(self.columns, device=self.dev)
        H[diag, diag] += damp
        H = torch.linalg.cholesky(H)
        H = torch.cholesky_inverse(H)
        Hinv = torch.linalg.cholesky(H, upper=True).contiguous()

        Q, loss = gptq_solve(W.contiguous(), Hinv, scales, zeros, bits=self.bits, blocksize=self.blocksize,
                             groupsize=self.groupsize)
        if self.actorder:
            Q = Q[:, torch.argsort(perm)]
        weight = Q.reshape(self.linear_module.weight.shape).to(self.linear_module.weight.data.dtype)
        error = torch.sum(loss).item()

        self.scales.copy_(scales)
        self.zeros.copy_(zeros)
        prev = torch.get_default_dtype()
        wd = self.linear_module.weight.dtype
        torch.set_default_dtype(wd if wd.is_floating_point else prev)
        try:
            q_module = ColBlockQuantizedLinear(self.linear_module.in_features, self.linear_module.out_features,
                                               self.linear_module.bias is not None, bits=self.bits,
                                               tile_cols=self.groupsize).to(self.dev)
        finally:
            torch.set_default_dtype(prev)
        q_module.scales = self.scales
        q_module.zeros = self.zeros
        q_module.pack_weight(weight)
        q_module.bias = self.linear_module.bias
        return q_module, error
