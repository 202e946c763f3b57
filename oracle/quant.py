"""Numpy restatement of the reference's quantized-Linear arithmetic.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Citations are into
``/root/reference/lit_llama/quantization.py`` unless another file is named.

dtype handling: the reference computes in whatever dtype the tensors carry
(fp32 under ``32-true``, bf16 under ``bf16-true``).  The oracle keeps every value
in a float32 numpy array and, when ``dtype == "bf16"``, rounds to bfloat16
(round-to-nearest-even) after each elementwise op exactly where torch would.
"""
from __future__ import annotations

import numpy as np

F32 = np.float32


# --------------------------------------------------------------------------- dtypes
def bf16_round(x) -> np.ndarray:
    """float32 -> nearest bfloat16 (RNE) -> float32.  NaN passes through."""
    x = np.ascontiguousarray(x, dtype=F32)
    u = x.view(np.uint32).astype(np.uint64)
    bias = ((u >> 16) & 1) + 0x7FFF
    r = ((u + bias) & 0xFFFF0000).astype(np.uint32).view(F32)
    return np.where(np.isnan(x), x, r).astype(F32)


def fp16_round(x) -> np.ndarray:
    with np.errstate(over="ignore"):
        return np.asarray(x, dtype=F32).astype(np.float16).astype(F32)


def rd(x, dtype: str) -> np.ndarray:
    """Round a float32 array to the storage dtype the reference tensor would have."""
    if dtype == "fp32":
        return np.asarray(x, dtype=F32)
    if dtype == "bf16":
        return bf16_round(x)
    if dtype == "fp16":
        return fp16_round(x)
    raise ValueError(dtype)


def to_bf16_bits(x) -> np.ndarray:
    """float32 (already bf16-representable or not) -> uint16 bf16 bit patterns (RNE)."""
    return (bf16_round(x).view(np.uint32) >> 16).astype(np.uint16)


def from_bf16_bits(b) -> np.ndarray:
    return (np.asarray(b, dtype=np.uint16).astype(np.uint32) << 16).view(F32)


# --------------------------------------------------------------------------- RTN params (row Z)
def find_params_weight(x: np.ndarray, maxq: int):
    """Per-row asymmetric min/max scale and zero.  quantization.py:474-511
    (perchannel=True, sym=False - the only configuration quantize/gptq.py uses).
    Returns (scale[rows,1], zero[rows,1]) float32."""
    x = np.asarray(x, dtype=F32).reshape(x.shape[0], -1)
    xmin = np.minimum(x.min(1), F32(0))
    xmax = np.maximum(x.max(1), F32(0))
    dead = (xmin == 0) & (xmax == 0)
    xmin = np.where(dead, F32(-1), xmin).astype(F32)
    xmax = np.where(dead, F32(+1), xmax).astype(F32)
    scale = ((xmax - xmin) / F32(maxq)).astype(F32)
    zero = np.round((-xmin / scale).astype(F32)).astype(F32)  # torch.round == half-to-even == np.round
    return scale.reshape(-1, 1), zero.reshape(-1, 1)


def quantize_weight(x, scale, zero, maxq: int) -> np.ndarray:
    """Round-to-grid and de-quantise.  quantization.py:468-472."""
    x = np.asarray(x, dtype=F32)
    q = np.clip(np.round((x / scale).astype(F32)) + zero, 0, maxq).astype(F32)
    return (scale * (q - zero)).astype(F32)


def rtn_quantize(W: np.ndarray, bits: int, groupsize: int = -1):
    """Round-to-nearest on-grid weights + params per group: the grouped fixture of
    SURVEY.md section 8c (find_params_weight per column slice, quantize_weight).
    Returns (W_on_grid[N,K], scales[N,G], zeros[N,G])."""
    W = np.asarray(W, dtype=F32)
    N, K = W.shape
    tile = K if groupsize == -1 else groupsize
    G = (K + tile - 1) // tile
    maxq = 2**bits - 1
    scales = np.empty((N, G), F32)
    zeros = np.empty((N, G), F32)
    Q = np.empty_like(W)
    for j in range(G):
        sl = slice(j * tile, (j + 1) * tile)
        s, z = find_params_weight(W[:, sl], maxq)
        scales[:, j : j + 1] = s
        zeros[:, j : j + 1] = z
        Q[:, sl] = quantize_weight(W[:, sl], s, z, maxq)
    return Q, scales, zeros


# --------------------------------------------------------------------------- pack / unpack (rows P, G)
def pack_weight(weight, scales, zeros, bits: int, tile_cols: int, weight_dtype: str = "fp32") -> np.ndarray:
    """ColBlockQuantizedLinear.pack_weight, quantization.py:374-388.

    weight [N,K] (on-grid floats), scales/zeros [N,G].  Per group, in the weight's
    dtype and in this order: ``/= scale``, ``+= zero``; then clamp to [0, 2^bits-1]
    and cast to uint8 (truncation toward zero, NOT rounding); entry ``nr`` of each
    byte is shifted by ``nr*bits`` (even k -> low nibble for bits=4).
    Returns uint8 [N, K/epb] (logical shape; the reference stores it column-major)."""
    w = rd(np.array(weight, dtype=F32, copy=True), weight_dtype)
    N, K = w.shape
    tile = K if tile_cols == -1 else tile_cols
    epb = 8 // bits
    assert epb > 0 and epb * bits == 8 and K % epb == 0
    scales = np.asarray(scales, dtype=F32)
    zeros = np.asarray(zeros, dtype=F32)
    for j in range(scales.shape[1]):
        sl = slice(j * tile, (j + 1) * tile)
        w[:, sl] = rd(w[:, sl] / scales[:, j : j + 1], weight_dtype)
        w[:, sl] = rd(w[:, sl] + zeros[:, j : j + 1], weight_dtype)
    w = np.clip(w, 0, 2**bits - 1)
    codes = np.trunc(w).astype(np.uint8)
    packed = np.zeros((N, K // epb), np.uint8)
    for nr in range(epb):
        packed += (codes[:, nr::epb] << np.uint8(nr * bits)).astype(np.uint8)
    return packed


def unpack_codes(qw: np.ndarray, bits: int) -> np.ndarray:
    """uint8 [N, K/epb] -> integer codes [N, K].  quantization.py:396-400."""
    epb = 8 // bits
    mask = (1 << bits) - 1
    N, Kp = qw.shape
    codes = np.empty((N, Kp * epb), np.uint8)
    for nr in range(epb):
        codes[:, nr::epb] = (qw >> np.uint8(nr * bits)) & np.uint8(mask)
    return codes


def get_weight(qw, scales, zeros, bits: int, tile_cols: int, dtype: str = "fp32") -> np.ndarray:
    """ColBlockQuantizedLinear.get_weight, quantization.py:390-409: codes as floats,
    then per group ``-= zero`` and ``*= scale`` in ``dtype`` (each op rounded)."""
    w = unpack_codes(np.asarray(qw), bits).astype(F32)
    N, K = w.shape
    tile = K if tile_cols == -1 else tile_cols
    scales = np.asarray(scales, dtype=F32)
    zeros = np.asarray(zeros, dtype=F32)
    for j in range(scales.shape[1]):
        sl = slice(j * tile, (j + 1) * tile)
        w[:, sl] = rd(w[:, sl] - zeros[:, j : j + 1], dtype)
        w[:, sl] = rd(w[:, sl] * scales[:, j : j + 1], dtype)
    return w


def linear_forward(inp, qw, scales, zeros, bias, bits: int, tile_cols: int, dtype: str = "fp32") -> np.ndarray:
    """ColBlockQuantizedLinear.forward, fallback branch, quantization.py:420-421:
    ``F.linear(inp, get_weight(inp.dtype), bias)`` - fp32 accumulation, one rounding
    to ``dtype`` on the way out."""
    w = get_weight(qw, scales, zeros, bits, tile_cols, dtype)
    x = rd(inp, dtype)
    y = x.reshape(-1, x.shape[-1]).astype(F32) @ w.T
    if bias is not None:
        y = y + np.asarray(bias, dtype=F32)
    return rd(y, dtype).reshape(*x.shape[:-1], w.shape[0])


def linear_forward_exact(inp, qw, scales, zeros, bits: int, tile_cols: int) -> np.ndarray:
    """The Triton branch's algebra, quantization.py:247-267, without TF32 rounding:
    ``C = A . ((nib - zero) * scale)`` in float64 - the tolerance anchor."""
    w = unpack_codes(np.asarray(qw), bits).astype(np.float64)
    N, K = w.shape
    tile = K if tile_cols == -1 else tile_cols
    s = np.repeat(np.asarray(scales, np.float64), tile, axis=1)[:, :K]
    z = np.repeat(np.asarray(zeros, np.float64), tile, axis=1)[:, :K]
    w = (w - z) * s
    x = np.asarray(inp, np.float64)
    return (x.reshape(-1, K) @ w.T).reshape(*x.shape[:-1], N)


# --------------------------------------------------------------------------- GPTQ solver (row Z2)
def gptq_hessian(inputs) -> tuple[np.ndarray, int]:
    """GPTQQuantizer.collect_input_stats, quantization.py:513-527, replayed over a
    list of [tokens, K] (or [B, tokens, K]) activations."""
    H = None
    nsamples = 0
    for inp in inputs:
        inp = np.asarray(inp, F32)
        if inp.ndim == 2:
            inp = inp[None]
        tmp = inp.shape[0]
        x = inp.reshape(-1, inp.shape[-1]).T
        if H is None:
            H = np.zeros((x.shape[0], x.shape[0]), F32)
        H *= F32(nsamples / (nsamples + tmp))
        nsamples += tmp
        x = (F32(np.sqrt(2 / nsamples)) * x).astype(F32)
        H += x @ x.T
    return H, nsamples


def gptq_hinv(H, percdamp: float = 0.01) -> np.ndarray:
    """quantization.py:549-555: damp the diagonal by ``percdamp * mean(diag H)``, then
    ``cholesky -> cholesky_inverse -> cholesky(upper=True)`` in float32 (LAPACK spotrf / spotri)."""
    from scipy.linalg import lapack

    H = np.array(H, dtype=F32, copy=True)
    columns = H.shape[0]
    damp = F32(percdamp) * np.mean(np.diag(H)).astype(F32)
    H[np.arange(columns), np.arange(columns)] += damp
    L, info = lapack.spotrf(H, lower=1)
    assert info == 0
    Hinv, info = lapack.spotri(L, lower=1)
    assert info == 0
    Hinv = np.tril(Hinv) + np.tril(Hinv, -1).T
    U, info = lapack.spotrf(Hinv.astype(F32), lower=0)
    assert info == 0
    return np.triu(U).astype(F32)


def gptq_quantize(W, H, bits: int, blocksize: int = 128, percdamp: float = 0.01,
                  groupsize: int = -1, actorder: bool = False, return_hinv: bool = False):
    """GPTQQuantizer.quantize, quantization.py:529-614, float32 throughout.

    Returns (Q_on_grid[N,K], scales[N,G], zeros[N,G], error) (+ the Hinv factor and the permuted working W the
    column walk started from, when ``return_hinv``).  For ``groupsize != -1``
    the reference raises at 573-577 (it assigns a [rows,1] tensor into a [rows] slice);
    the evident intent - store the group's params in column ``g`` - is restated here."""
    W = np.array(W, dtype=F32, copy=True)
    H = np.array(H, dtype=F32, copy=True)
    rows, columns = W.shape
    maxq = 2**bits - 1
    tile = columns if groupsize == -1 else groupsize
    G = (columns + tile - 1) // tile
    scales = np.zeros((rows, G), F32)
    zeros = np.zeros((rows, G), F32)
    scale, zero = find_params_weight(W, maxq)
    scales[:] = scale
    zeros[:] = zero

    d = np.diag(H).copy()
    dead = d == 0
    H[dead, dead] = 1
    W[:, dead] = 0
    if actorder:
        perm = np.argsort(-np.diag(H), kind="stable")
        W = W[:, perm]
        H = H[perm][:, perm]
    Losses = np.zeros_like(W)
    Q = np.zeros_like(W)
    Hinv = gptq_hinv(H, percdamp)
    W_start = W.copy()

    for i1 in range(0, columns, blocksize):
        i2 = min(i1 + blocksize, columns)
        W1 = W[:, i1:i2].copy()
        Q1 = np.zeros_like(W1)
        Err1 = np.zeros_like(W1)
        Losses1 = np.zeros_like(W1)
        Hinv1 = Hinv[i1:i2, i1:i2]
        for i in range(i2 - i1):
            w = W1[:, i]
            dd = Hinv1[i, i]
            if groupsize != -1 and (i1 + i) % groupsize == 0:
                scale, zero = find_params_weight(W[:, i1 + i : i1 + i + groupsize], maxq)
                scales[:, (i1 + i) // groupsize] = scale[:, 0]
                zeros[:, (i1 + i) // groupsize] = zero[:, 0]
            q = quantize_weight(w[:, None], scale, zero, maxq)[:, 0]
            Q1[:, i] = q
            Losses1[:, i] = (w - q) ** 2 / dd**2
            err1 = ((w - q) / dd).astype(F32)
            W1[:, i:] -= np.outer(err1, Hinv1[i, i:]).astype(F32)
            Err1[:, i] = err1
        Q[:, i1:i2] = Q1
        Losses[:, i1:i2] = Losses1 / 2
        W[:, i2:] -= Err1 @ Hinv[i1:i2, i2:]
    if actorder:
        Q = Q[:, np.argsort(perm, kind="stable")]
    if return_hinv:
        return Q, scales, zeros, float(Losses.sum()), Hinv, W_start
    return Q, scales, zeros, float(Losses.sum())


# --------------------------------------------------------------------------- llm.int8 (row I8) - parity unpinned
INT8_THRESHOLD = 6.0  # quantization.py:45


def int8_quantize_weight(W):
    """``Linear8bitLt._quantize_weight`` (quantization.py:67-75): ``W.half()`` then
    bitsandbytes ``double_quant`` row statistics: ``SCB = absmax_row`` (fp16-valued),
    ``CB = round(127 * W / SCB)`` int8.  bitsandbytes is un-vendored/unpinned in the
    reference (requirements.txt:7) - restated from arXiv:2208.07339 section 3."""
    Wh = fp16_round(W)
    scb = np.abs(Wh).max(axis=1)
    safe = np.where(scb == 0, F32(1), scb)
    cb = np.clip(np.rint(Wh * (F32(127) / safe[:, None])), -127, 127).astype(np.int8)
    return cb, scb.astype(F32)


def int8_linear_forward(x, cb, scb, bias=None, threshold: float = INT8_THRESHOLD, out_dtype: str = "bf16"):
    """bitsandbytes ``MatMul8bitLt`` forward with ``has_fp16_weights=False`` as called
    through quantization.py:36-45 - vector-wise int8 matmul plus fp16 outlier
    decomposition:

    1. ``A = x.half()`` flattened to [M,K]; outlier feature columns = any |A| >= threshold.
    2. row statistics ``SCA`` = absmax over the non-outlier ENTRIES of each row;
       ``CA = round(127 * A / SCA)`` with outlier columns zeroed.
    3. ``C32 = CA @ CB^T`` (int32); ``out = C32 * SCA * SCB / 127^2`` -> fp16 (+bias).
    4. ``out += A[:, idx] @ (CB[:, idx] * SCB / 127)^T`` in fp16 (fp32 accumulate).
    5. cast to the input dtype."""
    x = np.asarray(x, F32)
    A = fp16_round(x.reshape(-1, x.shape[-1]))
    big = np.abs(A) >= F32(threshold)
    idx = np.nonzero(big.any(axis=0))[0]
    A_in = np.where(big, F32(0), A)
    sca = np.abs(A_in).max(axis=1)
    safe = np.where(sca == 0, F32(1), sca)
    ca = np.clip(np.rint(A_in * (F32(127) / safe[:, None])), -127, 127).astype(np.int32)
    ca[:, idx] = 0
    c32 = ca @ cb.astype(np.int32).T
    out = c32.astype(F32) * sca[:, None] * scb[None, :] * F32(1.0 / (127.0 * 127.0))
    if bias is not None:
        out = out + fp16_round(bias)
    out = fp16_round(out)
    if idx.size:
        subB = fp16_round(cb[:, idx].astype(F32) * scb[:, None] / F32(127.0))
        out = fp16_round(out + fp16_round(A[:, idx] @ subB.T))
    return rd(out, out_dtype).reshape(*x.shape[:-1], cb.shape[0])
