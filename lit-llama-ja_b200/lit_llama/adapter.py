"""LLaMA-Adapter for inference on B200 (ref: lit_llama/adapter.py:47-313): the reference's module tree,
state-dict keys and call signatures over this package's kernels.

What the adapter adds to an attention layer (``block_idx >= adapter_start_layer``): the queries also attend to a learned
prompt of ``adapter_prompt_length`` rows pushed through ``c_attn`` (keys/values NOT rotated), and the result, times a
per-head gating factor, is summed with the self-attention output before ``c_proj``.  The reference does this with a second
``scaled_dot_product_attention`` call, an all-true mask, a broadcast multiply and an add (adapter.py:145-165); here it is
one launch of ``ll_adapter_attn_add`` that updates the attention output in place.

Differences by design:
* the prompt's keys/values are kept as ``[1, n_head, prompt_len, head_size]`` (one copy, read by every batch row) instead
  of the reference's ``repeat(B, ...)`` copy; they are computed on the first forward after ``reset_cache()``;
* single-token steps with a KV cache replay one CUDA graph per (batch, cache length) of the per-module path (the fused
  decode engine of ``engine.py`` is the plain model's);
* CUDA only (no CPU fallback).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Optional

import torch
import torch.nn as nn

from . import _native as N
from . import model as llama
from .model import KVCache, MLP, RMSNorm, _residual, _Step, apply_rope, build_rope_cache  # noqa: F401  (re-exported as the reference does)


@dataclass
class LLaMAConfig(llama.LLaMAConfig):
    """ref: adapter.py:47-50"""
    adapter_prompt_length: int = 10
    adapter_start_layer: int = 2


class CausalSelfAttention(llama.CausalSelfAttention):
    """ref: adapter.py:53-186 - ``lit_llama.model.CausalSelfAttention`` plus the attention over the adaption prompt."""

    def __init__(self, config: LLaMAConfig, block_idx: int) -> None:
        super().__init__(config)
        if block_idx >= config.adapter_start_layer:
            self.adapter_wte = nn.Embedding(config.adapter_prompt_length, config.n_embd)
            # zero-initialised: an untrained adapter leaves the pretrained model's output unchanged (adapter.py:72-74)
            self.gating_factor = nn.Parameter(torch.zeros(1, config.n_head, 1, 1))
        self.block_idx = block_idx
        self.adapter_prompt_length = config.adapter_prompt_length
        self.adapter_start_layer = config.adapter_start_layer

    def prefix_kv(self) -> KVCache:
        """Keys and values of the adaption prompt, ``[1, n_head, prompt_len, head_size]`` each (adapter.py:151-156: the
        k and v thirds of ``c_attn(prompt)``; no rotation)."""
        aT, nh, hs = self.adapter_prompt_length, self.n_head, self.head_size
        prefix = self.adapter_wte.weight.reshape(1, aT, self.n_embd)
        _, ak, av = self.c_attn(prefix).split(nh * hs, dim=2)
        ak = ak.reshape(1, aT, nh, hs).transpose(1, 2).contiguous()
        av = av.reshape(1, aT, nh, hs).transpose(1, 2).contiguous()
        return ak, av

    def forward(self, x: torch.Tensor, step: _Step, kv_cache: Optional[KVCache] = None,
                adapter_kv_cache: Optional[KVCache] = None):
        q, y, kv_cache = self._attend(x, step, kv_cache)
        if self.block_idx >= self.adapter_start_layer:
            if adapter_kv_cache is None:
                adapter_kv_cache = self.prefix_kv()
            ak, av = adapter_kv_cache
            B, nh, T, hs = q.shape
            gate = self.gating_factor.detach().reshape(-1)
            if gate.dtype != y.dtype:
                gate = gate.to(y.dtype)
            if ak.dtype != y.dtype:
                ak, av = ak.to(y.dtype), av.to(y.dtype)
            N.call("ll_adapter_attn_add", N.ptr(q), N.ptr(ak), N.ptr(av), N.ptr(gate.contiguous()), N.ptr(y),
                   N.dtype_code(y.dtype), B, nh, T, hs, ak.shape[2], int(ak.shape[0] != 1), N.stream())
        return self.c_proj(y), kv_cache, adapter_kv_cache

    def _load_from_state_dict(self, state_dict, prefix, *args, **kwargs):
        """ref: adapter.py:171-186 - old checkpoints hold ONE gating value for all heads."""
        name = prefix + "gating_factor"
        if name in state_dict:
            tensor = state_dict[name]
            tensor = tensor._load_tensor() if hasattr(tensor, "_load_tensor") else tensor
            if tensor.dim() < 4:
                tensor = tensor.reshape(1, 1, 1, 1).repeat(1, self.n_head, 1, 1)
            state_dict[name] = tensor
        return super()._load_from_state_dict(state_dict, prefix, *args, **kwargs)


class Block(nn.Module):
    """ref: adapter.py:189-214"""

    def __init__(self, config: LLaMAConfig, block_idx: int) -> None:
        super().__init__()
        self.rms_1 = RMSNorm(config.n_embd)
        self.attn = CausalSelfAttention(config, block_idx)
        self.rms_2 = RMSNorm(config.n_embd)
        self.mlp = MLP(config)

    def forward(self, x: torch.Tensor, step: _Step, kv_cache: Optional[KVCache] = None,
                adapter_kv_cache: Optional[KVCache] = None):
        h, kv_cache, adapter_kv_cache = self.attn(self.rms_1(x), step, kv_cache, adapter_kv_cache)
        x = _residual(x, h)
        x = _residual(x, self.mlp(self.rms_2(x)))
        return x, kv_cache, adapter_kv_cache


class LLaMA(llama.LLaMA):
    """ref: adapter.py:217-301 - the same submodule names (state-dict keys) and ``forward`` signature; as in the
    reference ``lm_head`` and ``wte`` are sized by ``config.vocab_size``."""

    use_fused_decode = False      # the fused decode engine knows the plain model only
    graph_every_batch = True      # ... so batch-1 steps go through the graphed per-module step as well

    def __init__(self, config: LLaMAConfig) -> None:
        nn.Module.__init__(self)
        assert config.vocab_size is not None
        assert config.block_size is not None
        self.config = config
        self.lm_head = nn.Linear(config.n_embd, config.vocab_size, bias=False)
        self.transformer = nn.ModuleDict(
            dict(
                wte=nn.Embedding(config.vocab_size, config.n_embd),
                h=nn.ModuleList(Block(config, i) for i in range(config.n_layer)),
                ln_f=RMSNorm(config.n_embd),
            )
        )
        self.rope_cache: Optional[torch.Tensor] = None
        self.mask_cache: Optional[torch.Tensor] = None
        self.kv_caches: List[KVCache] = []
        self.adapter_kv_caches: List[Optional[KVCache]] = []
        self._rope_f32: Optional[torch.Tensor] = None
        self._engine = None
        self._kv_pool = {}
        self._step_graphs = {}

    @classmethod
    def from_name(cls, name: str) -> "LLaMA":
        return cls(LLaMAConfig.from_name(name))

    def reset_cache(self) -> None:
        """ref: adapter.py:245-247.  The captured step graphs read the prompt's keys/values at fixed addresses, so they
        go with them (the next generation captures again: one warm-up step's worth of time)."""
        super().reset_cache()
        self.adapter_kv_caches.clear()
        self._step_graphs.clear()

    def _run_blocks(self, x: torch.Tensor, step: _Step) -> torch.Tensor:
        """ref: adapter.py:276-294"""
        if not step.use_cache:
            for block in self.transformer.h:
                x, _, _ = block(x, step)
            return x
        if not self.adapter_kv_caches:
            self.adapter_kv_caches = [None for _ in range(self.config.n_layer)]
        for i, block in enumerate(self.transformer.h):
            x, self.kv_caches[i], self.adapter_kv_caches[i] = block(x, step, self.kv_caches[i], self.adapter_kv_caches[i])
        return x


def mark_only_adapter_as_trainable(model: LLaMA) -> None:
    """ref: adapter.py:304-307"""
    for name, param in model.named_parameters():
        param.requires_grad = "adapter_wte" in name or "gating_factor" in name


def adapter_state_from_state_dict(state_dict: dict) -> dict:
    """ref: adapter.py:310-313"""
    return {name: param for name, param in state_dict.items() if "adapter_wte" in name or "gating_factor" in name}
