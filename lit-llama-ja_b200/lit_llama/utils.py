"""Class-swap boundary and small helpers - the same names the reference's ``lit_llama/utils.py``
exports on the inference path (ref: utils.py:19-42, 105-194, 364-376).  Training-time savers
(FSDP / DeepSpeed) are out of scope for this build.
"""
from __future__ import annotations

import functools
from contextlib import contextmanager

import torch
import torch.utils._device

# width of transformer.wte.weight -> model name (ref: utils.py:19-27)
llama_model_sizes = {512: "19M", 640: "49M", 780: "125M", 4096: "7B", 5120: "13B", 6656: "30B", 8192: "65B"}

QUANT_MODES = ("llm.int8", "gptq.int4", "gptq.int8")


def llama_model_lookup(checkpoint: dict) -> str:
    """ref: utils.py:30-36 - the embedding width identifies the model (wte is never quantized)."""
    return llama_model_sizes[checkpoint["transformer.wte.weight"].shape[1]]


def find_multiple(n: int, k: int) -> int:
    """ref: utils.py:39-42"""
    r = n % k
    return n if r == 0 else n + (k - r)


def _quantized_linear_factory(mode, groupsize: int = -1):
    """The class (or partial) that stands in for ``torch.nn.Linear`` under ``mode``; ``None`` for
    no quantization.  Unknown modes are the caller's error to raise (the two callers raise
    different exception types in the reference)."""
    if mode is None:
        return None
    if mode == "llm.int8":
        from .quantization import Linear8bitLt

        return Linear8bitLt
    if mode in ("gptq.int4", "gptq.int8"):
        from .quantization import ColBlockQuantizedLinear

        return functools.partial(ColBlockQuantizedLinear, bits=4 if mode == "gptq.int4" else 8, tile_cols=groupsize)
    return NotImplemented


@contextmanager
def quantization(mode: str = None, *, groupsize: int = -1):
    """ref: utils.py:173-194.  Inside the block ``torch.nn.Linear`` IS the quantized class, so
    ``LLaMA.from_name(...)`` builds quantized layers; restored on exit (also on error, which the
    reference does not guarantee).  ``groupsize`` is an extension: the reference always builds
    per-row (``tile_cols=-1``) modules; ours also re-shape themselves to a grouped checkpoint at
    ``load_state_dict`` time, so the one-argument call stays valid for group-128 checkpoints."""
    factory = _quantized_linear_factory(mode, groupsize)
    if factory is NotImplemented:
        raise ValueError(f"Unknown quantization mode: {mode}")
    if factory is None:
        yield
        return
    saved = torch.nn.Linear
    torch.nn.Linear = factory
    try:
        yield
    finally:
        torch.nn.Linear = saved


class EmptyInitOnDevice(torch.overrides.TorchFunctionMode):
    """ref: utils.py:105-170.  Tensor factories get ``device``/``dtype`` defaults, ``torch.nn.init.*``
    becomes a no-op, and (optionally) ``torch.nn.Linear`` is swapped for the quantized class.

    Example::
        with EmptyInitOnDevice("cuda", dtype=torch.bfloat16, quantization_mode="gptq.int4"):
            model = LLaMA.from_name("7B")
    """

    def __init__(self, device=None, dtype=None, quantization_mode=None):
        if quantization_mode == "llm.int8" and torch.device(device if device is not None else "cpu").type != "cuda":
            raise ValueError("Quantization is only supported on the GPU.")
        factory = _quantized_linear_factory(quantization_mode)
        if factory is NotImplemented:
            raise RuntimeError(f"unknown quantization mode {quantization_mode}")
        self.quantization_mode = quantization_mode
        self.quantized_linear_cls = factory
        self.device = device
        self.dtype = dtype
        self._saved_linear = None

    def __enter__(self):
        if self.quantized_linear_cls is not None:
            self._saved_linear = torch.nn.Linear
            torch.nn.Linear = self.quantized_linear_cls
        return super().__enter__()

    def __exit__(self, exc_type, exc_val, exc_tb):
        if self.quantized_linear_cls is not None:
            torch.nn.Linear = self._saved_linear
        return super().__exit__(exc_type, exc_val, exc_tb)

    def __torch_function__(self, func, types, args=(), kwargs=None):
        kwargs = dict(kwargs or {})
        if getattr(func, "__module__", None) == "torch.nn.init":
            return kwargs["tensor"] if "tensor" in kwargs else args[0]
        if func in torch.utils._device._device_constructors():
            if self.device is not None and kwargs.get("device") is None:
                kwargs["device"] = self.device
            if self.dtype is not None and kwargs.get("dtype") is None:
                kwargs["dtype"] = self.dtype
        return func(*args, **kwargs)


# ref: utils.py:198-376 (NotYetLoadedTensor, LazyLoadingUnpickler, lazy_load): the archive is indexed and tensors are read
# straight from their zip records on first use; checkpoint_stream.stream_load_state_dict adds the disk -> pinned -> HBM
# pipeline with the quantized layers' one-time repack under the read
from .checkpoint_stream import LazyLoadingUnpickler, NotYetLoadedTensor, lazy_load, stream_load_state_dict  # noqa: E402,F401
