"""GPTQ checkpoint writer for the B200-native LLaMA (ref: /root/reference/quantize/gptq.py).

Same entry points: ``llama_blockwise_quantization(model, sample_inputs, working_device, *, bits, groupsize)`` and
``main(checkpoint_path, output_path, tokenizer_path, n_samples, dtype, quantize)``; the written state dict has the
reference's layout (``quant_weight`` uint8 ``[N, K/2]`` strides ``(1, N)``, ``scales`` / ``zeros`` ``[N, G]``), so
either implementation loads it.  Differences, all deliberate:

* the solver runs on the device (``lit_llama.quantization.GPTQQuantizer``: one kernel per 128-column block instead
  of one Python iteration per column), and ``groupsize=128`` works (the reference's grouped branch raises);
* a B200 holds the whole model (180 GB), so blocks are not shuttled between host and device unless the model was
  handed over on the CPU (then they are, as in the reference, and the result is returned on the CPU);
* there is no network here: calibration text comes from ``--calibration_path`` (a text file, or a ``.pt`` tensor of
  token ids ``[n_samples, block_size]``) instead of a download of C4.
"""
from __future__ import annotations

import gc
import sys
import time
from pathlib import Path
from typing import Optional

import torch

sys.path.insert(0, str(Path(__file__).parent.parent.resolve()))

from lit_llama import LLaMA, Tokenizer  # noqa: E402
from lit_llama.model import _Step  # noqa: E402
from lit_llama.quantization import GPTQQuantizer  # noqa: E402
from lit_llama.utils import EmptyInitOnDevice, llama_model_lookup  # noqa: E402

SUBMODULES = ["attn.c_attn", "attn.c_proj", "mlp.c_fc1", "mlp.c_fc2", "mlp.c_proj"]   # ref gptq.py:64-70


def _block_forward(model, block, x):
    """``block(x, rope=..., mask=..., max_seq_length=block_size)`` of the reference (no KV cache)."""
    T = x.shape[1]
    ar = torch.arange(T, dtype=torch.int32, device=x.device)
    step = _Step(model._rope_f32, ar, ar, None, None, model.config.block_size, False)
    return block(x, step)[0]


@torch.no_grad()
def llama_blockwise_quantization(model, sample_inputs, working_device, *, bits=4, groupsize=-1, verbose=True):
    """ref: quantize/gptq.py:37-148.  Post-training quantisation of every Linear, in order: the statistics of a layer
    are collected on the outputs of the already-quantised layers before it."""
    log = print if verbose else (lambda *a, **k: None)
    dev = torch.device(working_device)
    home = model.transformer.wte.weight.device            # where the caller keeps the model
    model.transformer.wte.to(dev)
    sample_inputs = sample_inputs.to(dev)
    model.rope_cache = model._rope_f32 = model.mask_cache = None
    model._ensure_caches(sample_inputs)
    inps = model.transformer.wte(sample_inputs)
    model.transformer.wte.to(home)
    outs = torch.zeros_like(inps)

    def quantize_one(parent, attr, run):
        module = getattr(parent, attr)
        gptq = GPTQQuantizer(module, bits=bits, groupsize=groupsize, actorder=(groupsize == -1))
        handle = module.register_forward_hook(gptq.collect_input_stats)
        run()
        handle.remove()
        q_module, error = gptq.quantize()
        setattr(parent, attr, q_module)
        del gptq
        return error

    for i, block in enumerate(model.transformer.h):
        block.to(dev)

        def run_block():
            for j in range(inps.size(0)):
                outs[j:j + 1] = _block_forward(model, block, inps[j:j + 1])

        for name in SUBMODULES:
            t0 = time.perf_counter()
            pname, dname = name.rsplit(".", 1)
            error = quantize_one(block.get_submodule(pname), dname, run_block)
            log(f"{i} {name} time {time.perf_counter() - t0:.1f}s quantization error {error:.1f}")
        run_block()
        block.to(home)
        inps, outs = outs, inps                            # the outputs are the next block's inputs

    model.transformer.ln_f.to(dev)
    for j in range(inps.size(0)):
        outs[j:j + 1] = model.transformer.ln_f(inps[j:j + 1])
    model.transformer.ln_f.to(home)
    inps, outs = outs, inps

    model.lm_head.to(dev)

    def run_head():
        for j in range(inps.size(0)):
            model.lm_head(inps[j:j + 1])

    quantize_one(model, "lm_head", run_head)
    model.lm_head.to(home)
    model.kv_caches.clear()
    model._engine = None
    gc.collect()
    torch.cuda.empty_cache()
    return model


def get_sample_data(calibration_path: Path) -> str:
    """ref: quantize/gptq.py:22-33 downloads 1000 C4 documents; offline, the text comes from a local file."""
    return Path(calibration_path).read_text()


def main(
    *,
    checkpoint_path: Path = Path("checkpoints/lit-llama/7B/lit-llama.pth"),
    output_path: Optional[Path] = None,
    tokenizer_path: Path = Path("checkpoints/lit-llama/tokenizer.model"),
    n_samples: int = 128,
    dtype: str = "float32",
    quantize: Optional[str] = None,
    groupsize: int = -1,
    calibration_path: Optional[Path] = None,
) -> None:
    """ref: quantize/gptq.py:150-236 - same flags, plus ``groupsize`` and ``calibration_path``."""
    checkpoint_path, tokenizer_path = Path(checkpoint_path), Path(tokenizer_path)
    assert checkpoint_path.is_file()
    if output_path is None:
        output_path = checkpoint_path.parent / "llama-gptq.4bit.pth"
    output_path = Path(output_path)
    assert output_path.parent.is_dir() and (not output_path.exists() or output_path.is_file())
    if not torch.cuda.is_available():
        raise RuntimeError("lit-llama-ja_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    device = "cuda"
    dt = getattr(torch, dtype, None)
    if not isinstance(dt, torch.dtype):
        raise ValueError(f"{dtype} is not a valid dtype.")
    if quantize == "gptq.int4":
        bits = 4
    elif quantize == "gptq.int8":
        bits = 8
    else:
        raise RuntimeError(f"unknown/unsupported quantization mode {quantize}")
    if calibration_path is None:
        raise RuntimeError("no network: pass --calibration_path (text file, or .pt tensor of token ids)")

    print("Loading model ...", file=sys.stderr)
    t0 = time.time()
    checkpoint = torch.load(checkpoint_path, map_location="cpu")
    name = llama_model_lookup(checkpoint)
    with EmptyInitOnDevice(device=device, dtype=dt):        # the whole model fits one B200
        model = LLaMA.from_name(name)
    model.load_state_dict(checkpoint)
    model.eval()
    print(f"Time to load model: {time.time() - t0:.02f} seconds.", file=sys.stderr)

    block_size = 2048                                       # ref gptq.py:214
    calibration_path = Path(calibration_path)
    if calibration_path.suffix == ".pt":
        encoded = torch.load(calibration_path).reshape(-1)
    else:
        assert tokenizer_path.is_file()
        encoded = Tokenizer(tokenizer_path).encode(get_sample_data(calibration_path), bos=True, eos=False)
    n_samples = min(n_samples, encoded.numel() // block_size)
    assert n_samples >= 1, "calibration data shorter than one 2048-token sample"
    encoded = encoded[: n_samples * block_size].reshape(n_samples, block_size)

    t0 = time.perf_counter()
    llama_blockwise_quantization(model, encoded, device, bits=bits, groupsize=groupsize)
    print(f"\n\nTime for quantization: {time.perf_counter() - t0:.02f} sec total", file=sys.stderr)
    print(f"Memory used: {torch.cuda.max_memory_reserved() / 1e9:.02f} GB", file=sys.stderr)
    torch.save(model.state_dict(), output_path)


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("--checkpoint_path", type=Path, default=Path("checkpoints/lit-llama/7B/lit-llama.pth"))
    ap.add_argument("--output_path", type=Path, default=None)
    ap.add_argument("--tokenizer_path", type=Path, default=Path("checkpoints/lit-llama/tokenizer.model"))
    ap.add_argument("--n_samples", type=int, default=128)
    ap.add_argument("--dtype", default="float32")
    ap.add_argument("--quantize", default=None, choices=[None, "gptq.int4", "gptq.int8"])
    ap.add_argument("--groupsize", type=int, default=-1)
    ap.add_argument("--calibration_path", type=Path, default=None)
    main(**vars(ap.parse_args()))
