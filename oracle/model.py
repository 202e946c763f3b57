"""Numpy restatement of the reference LLaMA decode path around the quantized Linears.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Citations are into
``/root/reference/lit_llama/model.py`` and ``/root/reference/generate.py``.
Values live in float32 arrays; ``dtype="bf16"`` rounds after each op where torch would.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field
from typing import Callable, Optional

import numpy as np

from . import quant as Q

F32 = np.float32


def find_multiple(n: int, k: int) -> int:
    """lit_llama/utils.py:39-42"""
    return n if n % k == 0 else n + k - (n % k)


@dataclass
class Config:
    """LLaMAConfig, model.py:23-36."""
    block_size: int = 2048
    vocab_size: int = 32000
    padded_vocab_size: Optional[int] = None
    n_layer: int = 32
    n_head: int = 32
    n_embd: int = 4096

    def __post_init__(self):
        if self.padded_vocab_size is None:
            self.padded_vocab_size = find_multiple(self.vocab_size, 64)

    @property
    def n_hidden(self) -> int:
        """MLP width, model.py:249-251."""
        return find_multiple(int(2 * (4 * self.n_embd) / 3), 256)

    @property
    def head_size(self) -> int:
        return self.n_embd // self.n_head


# --------------------------------------------------------------------------- rows N, R
def rmsnorm(x, scale, eps: float = 1e-5, dtype: str = "fp32", rsqrt_mode: str = "cuda") -> np.ndarray:
    """RMSNorm.forward, model.py:276-283, every intermediate in the activation dtype:
    ``norm = mean(x*x)``; ``x_normed = x * rsqrt(norm + eps)``; ``scale * x_normed``.

    ``rsqrt_mode``: torch's bf16 ``rsqrt`` differs by device - the CUDA kernel evaluates
    ``rsqrtf(float(x))`` and rounds once (the behaviour that matters: bf16-true is only
    selected on CUDA, generate.py:121), the CPU vectorised kernel evaluates
    ``sqrt`` then ``reciprocal`` and rounds to bf16 after each (probed, torch 2.11).
    "cpu" exists so the oracle can be pinned against goldens produced on the CPU."""
    x = Q.rd(x, dtype)
    sq = Q.rd(x * x, dtype)
    norm = Q.rd(sq.astype(F32).mean(axis=-1, keepdims=True, dtype=F32), dtype)
    r = Q.rd(norm + F32(eps), dtype)
    if rsqrt_mode == "cpu" and dtype != "fp32":
        r = Q.rd(F32(1) / Q.rd(np.sqrt(r, dtype=F32), dtype), dtype)
    else:
        r = Q.rd(F32(1) / np.sqrt(r, dtype=F32), dtype)
    xn = Q.rd(x * r, dtype)
    return Q.rd(Q.rd(scale, dtype) * xn, dtype)


def build_rope_cache(seq_len: int, n_elem: int, base: int = 10000) -> np.ndarray:
    """build_rope_cache, model.py:286-309, for an integer or fp32 ``dtype`` argument
    (what ``LLaMA.build_rope_cache`` passes in generate.py: ``idx.dtype`` is an integer
    type, so the cache stays float32).  Returns [seq_len, n_elem/2, 2] = (cos, sin)."""
    theta = (F32(1.0) / (F32(base) ** (np.arange(0, n_elem, 2, dtype=F32) / F32(n_elem)))).astype(F32)
    seq_idx = np.arange(seq_len, dtype=F32)
    idx_theta = np.outer(seq_idx, theta).astype(F32)
    return np.stack([np.cos(idx_theta), np.sin(idx_theta)], axis=-1).astype(F32)


def apply_rope(x, rope, dtype: str = "fp32") -> np.ndarray:
    """apply_rope, model.py:312-329.  x [B,T,nh,hs]; rope [T, hs/2, 2] float32.
    Rotation of interleaved pairs in float32, then cast back to x's dtype."""
    x = np.asarray(x, F32)
    B, T, nh, hs = x.shape
    xs = x.reshape(B, T, nh, hs // 2, 2)
    rc = np.asarray(rope[:T], F32).reshape(1, T, 1, hs // 2, 2)
    o0 = xs[..., 0] * rc[..., 0] - xs[..., 1] * rc[..., 1]
    o1 = xs[..., 1] * rc[..., 0] + xs[..., 0] * rc[..., 1]
    return Q.rd(np.stack([o0, o1], axis=-1).reshape(B, T, nh, hs), dtype)


def sdpa(q, k, v, mask, dtype: str = "fp32") -> np.ndarray:
    """F.scaled_dot_product_attention(q,k,v,attn_mask=bool mask), model.py:237.
    q [B,nh,T,hs], k/v [B,nh,S,hs], mask broadcastable bool [.., T, S]; float32 softmax."""
    hs = q.shape[-1]
    att = (np.asarray(q, F32) @ np.asarray(k, F32).transpose(0, 1, 3, 2)) * F32(1.0 / math.sqrt(hs))
    att = np.where(mask, att, F32(-np.inf))
    att = att - att.max(axis=-1, keepdims=True)
    p = np.exp(att, dtype=F32)
    p = p / p.sum(axis=-1, keepdims=True, dtype=F32)
    return Q.rd(p @ np.asarray(v, F32), dtype)


# --------------------------------------------------------------------------- linears
@dataclass
class QLinear:
    """State of one ColBlockQuantizedLinear (quantization.py:339-372)."""
    qw: np.ndarray        # uint8 [N, K/epb]
    scales: np.ndarray    # [N, G]
    zeros: np.ndarray     # [N, G]
    bits: int = 4
    tile_cols: int = -1
    bias: Optional[np.ndarray] = None
    cache: bool = False   # keep the de-quantised weight between calls (7B-size layers over many decode steps)
    _w: dict = field(default_factory=dict, repr=False)

    def __call__(self, x, dtype):
        if not self.cache:
            return Q.linear_forward(x, self.qw, self.scales, self.zeros, self.bias, self.bits, self.tile_cols, dtype)
        # same arithmetic as Q.linear_forward (quantization.py:420-421) with get_weight() evaluated once
        if dtype not in self._w:
            self._w[dtype] = np.ascontiguousarray(Q.get_weight(self.qw, self.scales, self.zeros, self.bits, self.tile_cols, dtype).T)
        xr = Q.rd(x, dtype)
        y = xr.reshape(-1, xr.shape[-1]).astype(F32) @ self._w[dtype]
        if self.bias is not None:
            y = y + np.asarray(self.bias, dtype=F32)
        return Q.rd(y, dtype).reshape(*xr.shape[:-1], self._w[dtype].shape[1])


@dataclass
class FLinear:
    """A plain nn.Linear (un-quantized reference model)."""
    weight: np.ndarray    # [N, K]
    bias: Optional[np.ndarray] = None

    def __call__(self, x, dtype):
        x = Q.rd(x, dtype)
        y = x.reshape(-1, x.shape[-1]) @ Q.rd(self.weight, dtype).T
        if self.bias is not None:
            y = y + self.bias
        return Q.rd(y, dtype).reshape(*x.shape[:-1], self.weight.shape[0])


@dataclass
class I8Linear:
    """State of one Linear8bitLt (quantization.py:36-75)."""
    cb: np.ndarray
    scb: np.ndarray
    bias: Optional[np.ndarray] = None

    def __call__(self, x, dtype):
        return Q.int8_linear_forward(x, self.cb, self.scb, self.bias, out_dtype=dtype)


# --------------------------------------------------------------------------- the model (rows M, A, K)
@dataclass
class LLaMA:
    """LLaMA.forward with and without KV cache, model.py:84-128, over a flat state dict
    whose keys are the reference's (``transformer.h.0.attn.c_attn`` ...)."""
    config: Config
    linears: dict          # name -> QLinear | FLinear | I8Linear
    wte: np.ndarray        # [padded_vocab, E]
    norms: dict            # name -> scale[E]
    dtype: str = "fp32"
    rsqrt_mode: str = "cuda"
    rope_cache: Optional[np.ndarray] = None
    kv_caches: list = field(default_factory=list)

    def reset_cache(self):
        self.kv_caches = []

    def _attn(self, i, x, rope, mask, max_seq_length, input_pos, kv):
        """CausalSelfAttention.forward, model.py:192-243."""
        c = self.config
        B, T, C = x.shape
        qkv = self.linears[f"transformer.h.{i}.attn.c_attn"](x, self.dtype)
        q, k, v = np.split(qkv, 3, axis=2)
        hs = c.head_size
        q = apply_rope(q.reshape(B, T, c.n_head, hs), rope, self.dtype).transpose(0, 2, 1, 3)
        k = apply_rope(k.reshape(B, T, c.n_head, hs), rope, self.dtype).transpose(0, 2, 1, 3)
        v = v.reshape(B, T, c.n_head, hs).transpose(0, 2, 1, 3)
        if kv is not None:
            ck, cv = kv
            pos = np.asarray(input_pos)
            if pos[-1] >= max_seq_length:            # model.py:221-225 sliding window
                pos = np.array([max_seq_length - 1])
                ck = np.roll(ck, -1, axis=2)
                cv = np.roll(cv, -1, axis=2)
            ck = ck.copy(); cv = cv.copy()
            ck[:, :, pos] = k
            cv[:, :, pos] = v
            k, v = ck, cv
            kv = (ck, cv)
        y = sdpa(q, k, v, mask, self.dtype)
        y = self._attn_extra(i, q, y)
        y = y.transpose(0, 2, 1, 3).reshape(B, T, C)
        return self.linears[f"transformer.h.{i}.attn.c_proj"](y, self.dtype), kv

    def _attn_extra(self, i, q, y):
        """Nothing in the plain model; oracle/adapter.py adds the prefix attention here (between SDPA and c_proj)."""
        return y

    def _mlp(self, i, x):
        """MLP.forward, model.py:257-260."""
        a = self.linears[f"transformer.h.{i}.mlp.c_fc1"](x, self.dtype)
        b = self.linears[f"transformer.h.{i}.mlp.c_fc2"](x, self.dtype)
        silu = Q.rd(a / (F32(1) + np.exp(-a, dtype=F32)), self.dtype)
        return self.linears[f"transformer.h.{i}.mlp.c_proj"](Q.rd(silu * b, self.dtype), self.dtype)

    def forward(self, idx, max_seq_length=None, input_pos=None):
        c = self.config
        idx = np.asarray(idx)
        B, T = idx.shape
        if max_seq_length is None:
            max_seq_length = c.block_size
        assert T <= max_seq_length <= c.block_size
        if self.rope_cache is None:
            self.rope_cache = build_rope_cache(c.block_size, c.head_size)
        tril = np.tril(np.ones((c.block_size, c.block_size), bool))
        if input_pos is not None:
            input_pos = np.asarray(input_pos)
            rope = self.rope_cache[input_pos]
            mask = tril[input_pos][None, None, :, :max_seq_length]
        else:
            rope = self.rope_cache[:T]
            mask = tril[None, None, :T, :T]
        x = Q.rd(self.wte[idx], self.dtype)
        if input_pos is not None and not self.kv_caches:
            shape = (B, c.n_head, max_seq_length, c.head_size)
            self.kv_caches = [(np.zeros(shape, F32), np.zeros(shape, F32)) for _ in range(c.n_layer)]
        for i in range(c.n_layer):
            # Block.forward, model.py:162-175
            kv = self.kv_caches[i] if input_pos is not None else None
            h, kv = self._attn(i, rmsnorm(x, self.norms[f"transformer.h.{i}.rms_1.scale"], dtype=self.dtype, rsqrt_mode=self.rsqrt_mode),
                               rope, mask, max_seq_length, input_pos, kv)
            if input_pos is not None:
                self.kv_caches[i] = kv
            x = Q.rd(x + h, self.dtype)
            x = Q.rd(x + self._mlp(i, rmsnorm(x, self.norms[f"transformer.h.{i}.rms_2.scale"], dtype=self.dtype, rsqrt_mode=self.rsqrt_mode)), self.dtype)
        x = rmsnorm(x, self.norms["transformer.ln_f.scale"], dtype=self.dtype, rsqrt_mode=self.rsqrt_mode)
        return self.linears["lm_head"](x, self.dtype)


# --------------------------------------------------------------------------- row S
def sample_argmax(probs: np.ndarray) -> int:
    return int(np.argmax(probs))


def generate(model: LLaMA, idx, max_new_tokens: int, *, max_seq_length: Optional[int] = None,
             temperature: float = 1.0, top_k: Optional[int] = None, eos_id: Optional[int] = None,
             sampler: Callable[[np.ndarray], int] = sample_argmax, logits_out: Optional[list] = None):
    """generate(), generate.py:18-89.  ``sampler`` stands in for ``torch.multinomial``
    (it receives the softmax probabilities after temperature / top-k)."""
    idx = np.asarray(idx)
    T = idx.shape[0]
    T_new = T + max_new_tokens
    if max_seq_length is None:
        max_seq_length = min(T_new, model.config.block_size)
    buf = np.zeros(T_new, dtype=idx.dtype)
    buf[:T] = idx
    input_pos = np.arange(0, T)
    for _ in range(max_new_tokens):
        x = buf[input_pos].reshape(1, -1)
        logits = model.forward(x, max_seq_length, input_pos)
        if logits_out is not None:
            logits_out.append(logits[0, -1].copy())
        logits = Q.rd(logits[0, -1] / F32(temperature), model.dtype)
        if top_k is not None:
            kth = np.sort(logits)[::-1][min(top_k, logits.shape[-1]) - 1]
            logits = np.where(logits < kth, F32(-np.inf), logits)
        m = logits.max()
        p = np.exp(logits - m, dtype=F32)
        probs = Q.rd(p / p.sum(dtype=F32), model.dtype)
        nxt = sampler(probs)
        input_pos = input_pos[-1:] + 1
        buf[input_pos] = nxt
        if eos_id is not None and nxt == eos_id:
            return buf[: input_pos[0]]  # generate.py:86-87 slices [:input_pos], i.e. WITHOUT the eos
    return buf
