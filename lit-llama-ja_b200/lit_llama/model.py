"""LLaMA for inference on B200: the reference's module tree and call signatures
(ref: lit_llama/model.py:23-329) over hand-written sm_100a kernels.

What differs from the reference, by design:
* every op between the Linears is one kernel call through the C-ABI (RMSNorm, RoPE + in-place
  KV-cache write, split-KV decode attention, SwiGLU product, residual add) instead of 5-16 ATen
  launches each;
* the KV cache is written IN PLACE; the reference's roll-when-full (model.py:221-225) becomes a
  ring slot ``pos % max_seq_length`` - attention is order-independent over the slots;
* single-token steps of an int4/bf16 model run through the fused, CUDA-graph-captured decode
  engine (``engine.py``);
* CUDA only: tensors on the CPU raise (no CPU fallback).
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Tuple

import torch
import torch.nn as nn
from torch.nn import functional as F

from . import _native as N
from .utils import find_multiple

KVCache = Tuple[torch.Tensor, torch.Tensor]

llama_configs = {
    "19M": dict(n_layer=6, n_head=8, n_embd=512, vocab_size=35000),
    "49M": dict(n_layer=10, n_head=10, n_embd=640, vocab_size=35000),
    "125M": dict(n_layer=12, n_head=10, n_embd=780, vocab_size=35000),
    "7B": dict(n_layer=32, n_head=32, n_embd=4096),
    "13B": dict(n_layer=40, n_head=40, n_embd=5120),
    "30B": dict(n_layer=60, n_head=52, n_embd=6656),
    "65B": dict(n_layer=80, n_head=64, n_embd=8192),
}


@dataclass
class LLaMAConfig:
    """ref: model.py:23-46"""
    block_size: int = 2048
    vocab_size: int = 32000
    padded_vocab_size: Optional[int] = None
    n_layer: int = 32
    n_head: int = 32
    n_embd: int = 4096

    def __post_init__(self):
        if self.padded_vocab_size is None:
            self.padded_vocab_size = find_multiple(self.vocab_size, 64)

    @classmethod
    def from_name(cls, name: str) -> "LLaMAConfig":
        return cls(**llama_configs[name])

    @property
    def head_size(self) -> int:
        return self.n_embd // self.n_head

    @property
    def n_hidden(self) -> int:
        return find_multiple(int(2 * (4 * self.n_embd) / 3), 256)


class RMSNorm(nn.Module):
    """ref: model.py:263-283 - one kernel; every intermediate rounded to x's dtype as torch does."""

    def __init__(self, size: int, dim: int = -1, eps: float = 1e-5) -> None:
        super().__init__()
        self.scale = nn.Parameter(torch.ones(size))
        self.eps = eps
        self.dim = dim

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        N.require_cuda(x, self.scale)
        assert self.dim in (-1, x.dim() - 1), "RMSNorm kernel normalises the last dimension"
        x = x.contiguous()
        scale = self.scale if self.scale.dtype == x.dtype else self.scale.to(x.dtype)
        y = torch.empty_like(x)
        N.call("ll_rmsnorm", N.ptr(x), N.ptr(scale), N.ptr(y), N.dtype_code(x.dtype), x.numel() // x.shape[-1],
               x.shape[-1], float(self.eps), N.stream())
        return y


def build_rope_cache(seq_len: int, n_elem: int, dtype: torch.dtype, device: torch.device, base: int = 10000) -> torch.Tensor:
    """ref: model.py:286-309.  ``[seq_len, n_elem/2, 2]`` = (cos, sin) of ``pos * base^(-2i/n_elem)``;
    fp32 unless ``dtype`` is a 16-bit float or int8 (then fp16, as the reference).  One-time table."""
    exponents = torch.arange(0, n_elem, 2, dtype=dtype, device=device) / n_elem
    theta = 1.0 / (base ** exponents)
    positions = torch.arange(seq_len, dtype=dtype, device=device)
    angles = torch.outer(positions, theta).float()
    table = torch.stack((angles.cos(), angles.sin()), dim=-1)
    return table.half() if dtype in (torch.float16, torch.bfloat16, torch.int8) else table


def apply_rope(x: torch.Tensor, rope_cache: torch.Tensor) -> torch.Tensor:
    """ref: model.py:312-329.  x ``[B, T, n_head, head_size]``, rope_cache ``[>=T, head_size/2, 2]``."""
    N.require_cuda(x, rope_cache)
    B, T, nh, hs = x.shape
    rope = rope_cache[:T].float().contiguous()
    x = x.contiguous()
    y = torch.empty_like(x)
    N.call("ll_apply_rope", N.ptr(x), N.ptr(rope), N.ptr(y), N.dtype_code(x.dtype), B, T, nh, hs, N.stream())
    return y


@dataclass
class _Step:
    """Per-forward positional state shared by all blocks."""
    rope: torch.Tensor                 # full fp32 rope table [block_size, hs/2, 2]
    pos_ids: torch.Tensor              # int32 [T] rope rows
    slots: torch.Tensor                # int32 [T] cache slots
    pos3: Optional[torch.Tensor]       # int32 [3] (pos, slot, n_valid) for single-token steps
    mask: Optional[torch.Tensor]       # bool mask for T > 1 with cache
    max_seq_length: int
    use_cache: bool


class CausalSelfAttention(nn.Module):
    """ref: model.py:178-243"""

    def __init__(self, config: LLaMAConfig) -> None:
        super().__init__()
        assert config.n_embd % config.n_head == 0
        self.c_attn = nn.Linear(config.n_embd, 3 * config.n_embd, bias=False)
        self.c_proj = nn.Linear(config.n_embd, config.n_embd, bias=False)
        self.n_head = config.n_head          # heads held by THIS rank (tensor_parallel.py lowers it)
        self.head_size = config.n_embd // config.n_head
        self.n_embd = config.n_embd
        self.block_size = config.block_size

    def forward(self, x: torch.Tensor, step: _Step, kv_cache: Optional[KVCache] = None):
        _, y, kv_cache = self._attend(x, step, kv_cache)
        return self.c_proj(y), kv_cache

    def _attend(self, x: torch.Tensor, step: _Step, kv_cache: Optional[KVCache] = None):
        """Everything up to the output projection: the rotated queries ``[B, nh, T, hs]``, the attention output
        ``[B, T, nh*hs]`` and the updated cache (the adapter model adds its prefix attention between this and c_proj)."""
        B, T, _ = x.shape
        nh, hs = self.n_head, self.head_size
        C = nh * hs                                             # width of this rank's attention output
        qkv = self.c_attn(x).contiguous()                       # [B, T, 3C], rows ordered [q; k; v]
        if kv_cache is None:                                    # no-cache forward: a T-slot scratch cache
            S = T
            kv_cache = (torch.empty((B, nh, S, hs), device=x.device, dtype=x.dtype),
                        torch.empty((B, nh, S, hs), device=x.device, dtype=x.dtype))
        k_cache, v_cache = kv_cache
        S = k_cache.shape[2]
        q = torch.empty((B, nh, T, hs), device=x.device, dtype=x.dtype)
        N.call("ll_rope_kv_append", N.ptr(qkv), N.ptr(step.rope), N.ptr(step.pos_ids), N.ptr(step.slots), N.ptr(q),
               N.ptr(k_cache), N.ptr(v_cache), N.dtype_code(x.dtype), B, T, nh, hs, S, N.stream())
        if step.use_cache and T == 1 and hs in (32, 64, 128):
            y = torch.empty((B, 1, C), device=x.device, dtype=x.dtype)
            splits = max(1, min(8, (2 * N._fn["ll_sm_count"]()) // max(1, B * nh), (S + 63) // 64))
            ws = N.Workspace.get(x.device).ensure(attn_bytes=N.attn_scratch_bytes(B, nh, hs, splits),
                                                  attn_counters=B * nh)
            N.call("ll_decode_attn", N.ptr(q), N.ptr(k_cache), N.ptr(v_cache), N.ptr(step.pos3), N.ptr(y),
                   N.dtype_code(x.dtype), B, nh, hs, S, splits, N.ptr(ws.attn_scratch), N.ptr(ws.attn_counters), 0,
                   N.stream())
        else:
            # multi-token (prefill / no-cache) attention: library SDPA, as the reference (model.py:237)
            if step.use_cache:
                y = F.scaled_dot_product_attention(q, k_cache, v_cache, attn_mask=step.mask, dropout_p=0.0)
            else:
                y = F.scaled_dot_product_attention(q, k_cache, v_cache, is_causal=True, dropout_p=0.0)
            y = y.transpose(1, 2).contiguous().view(B, T, C)
        return q, y, kv_cache


class MLP(nn.Module):
    """ref: model.py:246-260 (SwiGLU)"""

    def __init__(self, config: LLaMAConfig) -> None:
        super().__init__()
        self.c_fc1 = nn.Linear(config.n_embd, config.n_hidden, bias=False)
        self.c_fc2 = nn.Linear(config.n_embd, config.n_hidden, bias=False)
        self.c_proj = nn.Linear(config.n_hidden, config.n_embd, bias=False)

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        a, b = self.c_fc1(x).contiguous(), self.c_fc2(x).contiguous()
        h = torch.empty_like(a)
        N.call("ll_silu_mul", N.ptr(a), N.ptr(b), N.ptr(h), N.dtype_code(a.dtype), a.numel(), N.stream())
        return self.c_proj(h)


def _residual(x: torch.Tensor, h: torch.Tensor) -> torch.Tensor:
    h = h.contiguous()
    y = torch.empty_like(x)
    N.call("ll_add", N.ptr(x), N.ptr(h), N.ptr(y), N.dtype_code(x.dtype), x.numel(), N.stream())
    return y


class Block(nn.Module):
    """ref: model.py:154-175"""

    def __init__(self, config: LLaMAConfig) -> None:
        super().__init__()
        self.rms_1 = RMSNorm(config.n_embd)
        self.attn = CausalSelfAttention(config)
        self.rms_2 = RMSNorm(config.n_embd)
        self.mlp = MLP(config)

    def forward(self, x: torch.Tensor, step: _Step, kv_cache: Optional[KVCache] = None):
        h, kv_cache = self.attn(self.rms_1(x), step, kv_cache)
        x = _residual(x, h)
        x = _residual(x, self.mlp(self.rms_2(x)))
        return x, kv_cache


class LLaMA(nn.Module):
    """ref: model.py:59-151 - same submodule names (state-dict keys), same ``forward`` signature."""

    use_fused_decode = True   # single-token int4/bf16 steps go through engine.DecodeEngine
    use_step_graph = True     # other single-token steps with a KV cache (batch > 8, ...) replay one CUDA graph per (B, S)
    graph_every_batch = False  # model families the fused engine does not take (adapter.py) graph batch 1 as well

    def __init__(self, config: LLaMAConfig) -> None:
        super().__init__()
        assert config.padded_vocab_size is not None
        self.config = config
        self.lm_head = nn.Linear(config.n_embd, config.padded_vocab_size, bias=False)
        self.transformer = nn.ModuleDict(
            dict(
                wte=nn.Embedding(config.padded_vocab_size, config.n_embd),
                h=nn.ModuleList(Block(config) for _ in range(config.n_layer)),
                ln_f=RMSNorm(config.n_embd),
            )
        )
        self.rope_cache: Optional[torch.Tensor] = None
        self.mask_cache: Optional[torch.Tensor] = None
        self.kv_caches: List[KVCache] = []
        self._rope_f32: Optional[torch.Tensor] = None
        self._engine = None
        self._kv_pool = {}          # (shape, dtype, device) -> KV tensors: stable addresses across reset_cache()
        self._step_graphs = {}      # (B, max_seq_length) -> captured single-token step of the per-module path

    def _init_weights(self, module: nn.Module) -> None:
        """ref: model.py:78-82"""
        std = 0.02 / math.sqrt(2 * self.config.n_layer)
        if isinstance(module, (nn.Linear, nn.Embedding)) and hasattr(module, "weight"):
            torch.nn.init.normal_(module.weight, mean=0.0, std=std)

    @classmethod
    def from_name(cls, name: str) -> "LLaMA":
        return cls(LLaMAConfig.from_name(name))

    def build_rope_cache(self, idx: torch.Tensor) -> torch.Tensor:
        return build_rope_cache(seq_len=self.config.block_size, n_elem=self.config.head_size, dtype=idx.dtype,
                                device=idx.device)

    def build_mask_cache(self, idx: torch.Tensor) -> torch.Tensor:
        n = self.config.block_size
        return torch.ones((n, n), device=idx.device, dtype=torch.bool).tril_()[None, None]

    def _ensure_caches(self, idx: torch.Tensor) -> None:
        """ref: model.py:100-103 - the RoPE and mask tables are built on first use, from ``idx``'s dtype and device."""
        if self.rope_cache is None:
            self.rope_cache = self.build_rope_cache(idx)
            self._rope_f32 = self.rope_cache.float().contiguous()
        if self.mask_cache is None:
            self.mask_cache = self.build_mask_cache(idx)

    def reset_cache(self) -> None:
        self.kv_caches.clear()
        if self._engine is not None:
            self._engine.reset()

    # ------------------------------------------------------------------ forward (row M)
    def forward(self, idx: torch.Tensor, max_seq_length: Optional[int] = None,
                input_pos: Optional[torch.Tensor] = None) -> torch.Tensor:
        N.require_cuda(idx)
        B, T = idx.size()
        block_size = self.config.block_size
        if max_seq_length is None:
            max_seq_length = block_size
        assert T <= max_seq_length, f"Cannot forward sequence of length {T}, max seq length is only {max_seq_length}"
        assert max_seq_length <= block_size, f"Cannot attend to {max_seq_length}, block size is only {block_size}"
        assert T <= block_size, f"Cannot forward sequence of length {T}, block size is only {block_size}"

        self._ensure_caches(idx)

        use_cache = input_pos is not None
        if use_cache and T == 1 and self.use_fused_decode:
            eng = self._fused_engine(B, max_seq_length)
            if eng is not None:
                return eng.forward_token(idx, input_pos)
        if (use_cache and T == 1 and self.use_step_graph and ((self.use_fused_decode and B > 1) or self.graph_every_batch)
                and input_pos.device.type == "cuda"
                and not torch.cuda.is_current_stream_capturing() and __import__("os").environ.get("LLB200_STEP_GRAPH", "1") != "0"):
            out = self._graphed_step(idx, max_seq_length, input_pos)
            if out is not None:
                return out
        return self._forward_modules(idx, max_seq_length, input_pos)

    def _graphed_step(self, idx, max_seq_length, input_pos):
        """One decode step of the per-module path (batches the fused engine does not take) as ONE CUDA-graph replay: the
        step is ~9 kernels per block and, launched eagerly, costs more host time than device time.  The position is read
        from a device buffer by every kernel (ll_set_pos), so one capture per (batch, cache length) serves every token."""
        B = idx.size(0)
        key = (B, max_seq_length)
        ent = self._step_graphs.get(key)
        if ent is not None and (not self.kv_caches or self.kv_caches[0][0].data_ptr() != ent["kv0"]):
            if not self.kv_caches and ent["kv"][0][0].shape[0] == B:
                self.kv_caches = list(ent["kv"])          # after reset_cache(): the captured cache tensors are re-attached
            else:
                ent = None
        if ent is None:
            if not self.kv_caches or self.kv_caches[0][0].shape[2] != max_seq_length or self.kv_caches[0][0].shape[0] != B:
                return None                                # no cache of this shape yet (the eager path allocates it)
            s_idx, s_pos = idx.clone(), input_pos.clone()
            side = torch.cuda.Stream()
            side.wait_stream(torch.cuda.current_stream())
            with torch.cuda.stream(side):                  # warm-up: workspaces, weight streams, kernel attributes
                for _ in range(2):
                    self._forward_modules(s_idx, max_seq_length, s_pos)
            torch.cuda.current_stream().wait_stream(side)
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                out = self._forward_modules(s_idx, max_seq_length, s_pos)
            ent = self._step_graphs[key] = dict(graph=g, idx=s_idx, pos=s_pos, out=out, kv=list(self.kv_caches),
                                                kv0=self.kv_caches[0][0].data_ptr())
        ent["idx"].copy_(idx)
        ent["pos"].copy_(input_pos)
        ent["graph"].replay()
        return ent["out"].clone()

    def _forward_modules(self, idx: torch.Tensor, max_seq_length: int, input_pos: Optional[torch.Tensor]) -> torch.Tensor:
        B, T = idx.size()
        block_size = self.config.block_size
        use_cache = input_pos is not None
        dev = idx.device
        pos3 = mask = None
        if use_cache:
            pos_ids = input_pos.to(torch.int32)
            if T == 1:
                pos3 = torch.empty(3, dtype=torch.int32, device=dev)
                N.call("ll_set_pos", N.ptr(pos3), N.ptr(input_pos.to(torch.int64)), max_seq_length, block_size, N.stream())
                pos_ids, slots = pos3[0:1], pos3[1:2]
                if self.config.head_size not in (32, 64, 128):   # SDPA branch needs an explicit mask
                    mask = (torch.arange(max_seq_length, device=dev) < pos3[2]).view(1, 1, 1, -1)
            else:
                slots = pos_ids % max_seq_length
                mask = self.mask_cache.index_select(2, input_pos)[:, :, :, :max_seq_length]
        else:
            pos_ids = slots = torch.arange(T, dtype=torch.int32, device=dev)
        step = _Step(self._rope_f32, pos_ids.contiguous(), slots.contiguous(), pos3, mask, max_seq_length, use_cache)

        x = self.transformer.wte(idx)
        if use_cache and not self.kv_caches:
            shape = (B, self.transformer.h[0].attn.n_head, max_seq_length, self.config.head_size)
            self.kv_caches = self._kv_from_pool(shape, x.dtype, dev)
        x = self._run_blocks(x, step)
        x = self.transformer.ln_f(x)
        return self.lm_head(x)

    def _run_blocks(self, x: torch.Tensor, step: _Step) -> torch.Tensor:
        """The block loop of ref model.py:105-122 (the adapter model overrides it to carry its prefix caches along)."""
        if not step.use_cache:
            for block in self.transformer.h:
                x, _ = block(x, step)
        else:
            for i, block in enumerate(self.transformer.h):
                x, self.kv_caches[i] = block(x, step, self.kv_caches[i])
        return x

    def _kv_from_pool(self, shape, dtype, dev):
        """Zeroed KV tensors of this shape - the SAME tensors after every reset_cache(), so that captured step graphs
        (which hold their addresses) stay valid; one set per shape (ref model.py:112-118 allocates fresh zeros each time)."""
        key = (tuple(shape), dtype, str(dev))
        kv = self._kv_pool.get(key)
        if kv is None:
            if len(self._kv_pool) >= 2:            # keep at most two shapes alive
                self._kv_pool.pop(next(iter(self._kv_pool)))
                self._step_graphs.clear()
            kv = self._kv_pool[key] = [(torch.zeros(shape, device=dev, dtype=dtype), torch.zeros(shape, device=dev, dtype=dtype))
                                       for _ in range(self.config.n_layer)]
        else:
            for k, v in kv:
                k.zero_(); v.zero_()
        return list(kv)

    def _fused_engine(self, B: int, max_seq_length: int):
        """The fused decode engine for this (batch, cache length), or None if the model is not
        eligible (not bf16/int4, unsupported shapes) - then the per-module path above runs."""
        from .engine import DecodeEngine

        eng = self._engine
        if eng is not None and eng.matches(B, max_seq_length):
            return eng
        if not DecodeEngine.eligible(self, B):
            return None
        if not self.kv_caches:
            shape = (B, self.transformer.h[0].attn.n_head, max_seq_length, self.config.head_size)
            dev = self.transformer.wte.weight.device
            self.kv_caches = [(torch.zeros(shape, device=dev, dtype=torch.bfloat16),
                               torch.zeros(shape, device=dev, dtype=torch.bfloat16)) for _ in range(self.config.n_layer)]
        elif self.kv_caches[0][0].shape[2] != max_seq_length or self.kv_caches[0][0].shape[0] != B:
            return None
        self._engine = DecodeEngine(self, B, max_seq_length)
        return self._engine
