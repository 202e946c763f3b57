"""Loads the unmodified reference package installed in baseline/_ref (baseline/install_ref.sh) beside this repo's own
``lit_llama`` package, in one process.

Both packages are called ``lit_llama`` and the reference imports itself by absolute name, so it is imported first with
baseline/_ref at the head of ``sys.path`` and its modules are then taken OUT of ``sys.modules`` (they keep working through
the references they hold to each other); whatever ``lit_llama`` modules were loaded before are put back.  ``lightning``
(absent from the image; the reference needs two class names from it, lit_llama/utils.py:13) is stubbed - SURVEY.md App. A.

Used only by tests/ and by bench.py's reference legs; nothing in the product imports it.
"""
from __future__ import annotations

import importlib
import sys
import types
from pathlib import Path

REF = Path(__file__).resolve().parent / "_ref"
_cache = None


def available() -> bool:
    return (REF / "lit_llama" / "quantization.py").exists()


def _stub_lightning():
    if "lightning" in sys.modules:
        return
    import torch

    L = types.ModuleType("lightning")
    F = types.ModuleType("lightning.fabric")
    S = types.ModuleType("lightning.fabric.strategies")
    S.DeepSpeedStrategy = type("DeepSpeedStrategy", (), {})
    S.FSDPStrategy = type("FSDPStrategy", (), {})
    F.strategies = S
    L.fabric = F
    L.Fabric = object
    L.seed_everything = lambda s: torch.manual_seed(s)
    sys.modules.update({"lightning": L, "lightning.fabric": F, "lightning.fabric.strategies": S})


def load():
    """-> namespace with .model, .quantization, .utils, .generate (the reference's modules)."""
    global _cache
    if _cache is not None:
        return _cache
    if not available():
        raise RuntimeError("baseline/_ref is missing: run baseline/install_ref.sh where /root/reference exists")
    _stub_lightning()
    mine = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "lit_llama" or k.startswith("lit_llama.")}
    sys.path.insert(0, str(REF))
    try:
        importlib.invalidate_caches()
        import lit_llama  # noqa: F401  (the reference)
        from lit_llama import model, quantization, utils
        spec = importlib.util.spec_from_file_location("ref_generate", REF / "ref_generate.py")
        gen = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(gen)
    finally:
        sys.path.remove(str(REF))
        theirs = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "lit_llama" or k.startswith("lit_llama.")}
        sys.modules.update(mine)
    assert str(REF) in str(Path(model.__file__).resolve()), model.__file__
    _cache = types.SimpleNamespace(model=model, quantization=quantization, utils=utils, generate=gen, modules=theirs)
    return _cache


def build_quantized_model(ref, config, *, bits=4, tile_cols=-1, device="cuda", dtype=None):
    """The reference's LLaMA with every nn.Linear replaced by its ColBlockQuantizedLinear - what
    ``lit_llama.utils.quantization("gptq.int4")`` does (utils.py:173-194), with the group size as a parameter (the
    reference's context manager is fixed at ``tile_cols=-1``; its module supports groups, quantization.py:339-372)."""
    import functools

    import torch

    saved = torch.nn.Linear
    saved_dtype = torch.get_default_dtype()
    torch.nn.Linear = functools.partial(ref.quantization.ColBlockQuantizedLinear, bits=bits, tile_cols=tile_cols)
    try:
        if dtype is not None:
            torch.set_default_dtype(dtype)
        with torch.device(device):
            m = ref.model.LLaMA(config)
    finally:
        torch.nn.Linear = saved
        torch.set_default_dtype(saved_dtype)
    return m.eval()
