"""LLaMA-Adapter v2 for inference on B200 (ref: lit_llama/adapter_v2.py:1-45): every Linear gets a learned
per-output-column ``adapter_bias`` and ``adapter_scale`` applied after its matrix product,
``y = adapter_scale * (linear(x) + adapter_bias)``.

The reference patches ``forward`` of every ``nn.Linear`` to ``F.linear`` + two broadcast ops.  Here the layer's own forward
runs (the quantized Linears keep their int4 / int8 kernels - in the reference the patch cannot be combined with them: under
``quantization("gptq.int4")`` its ``isinstance(module, nn.Linear)`` test is made against a ``functools.partial``) and one
``ll_bias_scale`` launch applies bias and scale in place, with torch's roundings (the sum, then the product, each in the
activation dtype).
"""
from __future__ import annotations

import torch
import torch.nn as nn
from torch import Tensor

from . import _native as N
from .adapter import LLaMA  # noqa: F401  (the reference re-exports it)
from .quantization import ColBlockQuantizedLinear, Linear8bitLt


def get_adapter_substrings():
    """ref: adapter_v2.py:9-13"""
    substrings = ["adapter_wte", "gating_factor"]          # adapter v1 parameters
    substrings.extend(["adapter_scale", "adapter_bias"])   # v2: bias and scale of every Linear
    substrings.extend(["rms_1", "rms_2", "ln_f"])          # v2: the RMSNorm scales are trained too
    return substrings


def mark_only_adapter_v2_as_trainable(model: LLaMA) -> None:
    """ref: adapter_v2.py:16-19"""
    for name, param in model.named_parameters():
        param.requires_grad = any(s in name for s in get_adapter_substrings())


def adapter_v2_state_from_state_dict(state_dict: dict) -> dict:
    """ref: adapter_v2.py:22-26"""
    return {name: param for name, param in state_dict.items() if any(s in name for s in get_adapter_substrings())}


def adapter_v2_new_forward(self, input: Tensor) -> Tensor:
    """ref: adapter_v2.py:29-32 - ``adapter_scale * (linear(input) + adapter_bias)``."""
    y = self._adapter_v2_inner_forward(input)
    N.require_cuda(y, self.adapter_bias, self.adapter_scale)
    y = y.contiguous()
    bias, scale = self.adapter_bias.detach(), self.adapter_scale.detach()
    if bias.dtype != y.dtype:
        bias = bias.to(y.dtype)
    if scale.dtype != y.dtype:
        scale = scale.to(y.dtype)
    n = y.shape[-1]
    N.call("ll_bias_scale", N.ptr(y), N.ptr(bias.contiguous()), N.ptr(scale.contiguous()), N.ptr(y), N.dtype_code(y.dtype),
           y.numel() // n, n, N.stream())
    return y


def adapter_v2_linear_with_bias_and_scale(layer):
    """ref: adapter_v2.py:35-40 - zero bias, unit scale: an untrained v2 layer computes what the layer computed."""
    out_features = layer.out_features
    ref = getattr(layer, "scales", None)      # ColBlockQuantizedLinear: parameters in the dtype of its scales
    if ref is None:
        ref = layer.weight                    # nn.Linear; Linear8bitLt (int8 weight: the default dtype)
    dtype = ref.dtype if ref.dtype.is_floating_point else torch.get_default_dtype()
    layer.adapter_bias = nn.Parameter(torch.zeros(out_features, device=ref.device, dtype=dtype), requires_grad=True)
    layer.adapter_scale = nn.Parameter(torch.ones(out_features, device=ref.device, dtype=dtype), requires_grad=True)
    layer._adapter_v2_inner_forward = layer.forward
    setattr(layer, "forward", adapter_v2_new_forward.__get__(layer, layer.__class__))
    return layer


def add_adapter_v2_parameters_to_linear_layers(model):
    """ref: adapter_v2.py:43-45 (the quantized Linears of this package count as Linears)."""
    # inside ``with quantization(...)`` torch.nn.Linear is a functools.partial, not a type: only types go to isinstance
    kinds = tuple(t for t in (nn.Linear, ColBlockQuantizedLinear, Linear8bitLt) if isinstance(t, type))
    for module in model.modules():
        if isinstance(module, kinds):
            adapter_v2_linear_with_bias_and_scale(module)
