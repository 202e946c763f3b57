"""Numpy restatement of the reference's LLaMA-Adapter (v1 and v2) forward.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).  Citations are into ``/root/reference/lit_llama/adapter.py`` and
``/root/reference/lit_llama/adapter_v2.py``.  Parity: **pinned** against outputs of the unmodified reference run in the
build container (``tests/golden/make_golden_adapter.py`` -> ``tests/golden/adapter_tiny.npz``).
"""
from __future__ import annotations

from dataclasses import dataclass, field

import numpy as np

from . import model as OM
from . import quant as Q

F32 = np.float32


@dataclass
class V2Linear:
    """adapter_v2_new_forward, adapter_v2.py:29-32: ``adapter_scale * (linear(x) + adapter_bias)``, each op in the
    activation dtype."""
    inner: object
    adapter_bias: np.ndarray
    adapter_scale: np.ndarray

    def __call__(self, x, dtype):
        y = self.inner(x, dtype)
        s = Q.rd(y + Q.rd(self.adapter_bias, dtype), dtype)
        return Q.rd(Q.rd(self.adapter_scale, dtype) * s, dtype)


@dataclass
class AdapterLLaMA(OM.LLaMA):
    """adapter.LLaMA.forward, adapter.py:249-301: the plain model with the prefix attention added in the layers
    ``block_idx >= adapter_start_layer``."""
    adapter_wte: dict = field(default_factory=dict)        # layer index -> [prompt_len, E]
    gating_factor: dict = field(default_factory=dict)      # layer index -> [n_head]
    adapter_start_layer: int = 2
    adapter_kv: dict = field(default_factory=dict)

    def reset_cache(self):
        """adapter.py:245-247"""
        super().reset_cache()
        self.adapter_kv = {}

    def _prefix_kv(self, i):
        """adapter.py:151-156: the k and v thirds of c_attn(prompt), split by head; no rotation."""
        c = self.config
        prefix = Q.rd(self.adapter_wte[i], self.dtype)[None]                      # [1, aT, E]
        aT = prefix.shape[1]
        qkv = self.linears[f"transformer.h.{i}.attn.c_attn"](prefix, self.dtype)
        _, ak, av = np.split(qkv, 3, axis=2)
        ak = ak.reshape(1, aT, c.n_head, c.head_size).transpose(0, 2, 1, 3)       # [1, nh, aT, hs]
        av = av.reshape(1, aT, c.n_head, c.head_size).transpose(0, 2, 1, 3)
        return ak, av

    def _attn_extra(self, i, q, y):
        """adapter.py:145-165: ``y + gating_factor * sdpa(q, ak, av, all-true mask)``."""
        if i < self.adapter_start_layer:
            return y
        if i not in self.adapter_kv:
            self.adapter_kv[i] = self._prefix_kv(i)
        ak, av = self.adapter_kv[i]
        B = q.shape[0]
        ak = np.repeat(ak, B, axis=0)
        av = np.repeat(av, B, axis=0)
        amask = np.ones((q.shape[-2], ak.shape[-2]), bool)
        ay = OM.sdpa(q, ak, av, amask, self.dtype)
        g = Q.rd(np.asarray(self.gating_factor[i], F32).reshape(1, -1, 1, 1), self.dtype)
        return Q.rd(y + Q.rd(g * ay, self.dtype), self.dtype)
