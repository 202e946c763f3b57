"""Sampling from a LLaMA-Adapter fine-tune (ref: /root/reference/generate/adapter.py:20-106): the same
flags; the pretrained checkpoint and the adapter checkpoint are loaded into ``lit_llama.adapter.LLaMA`` built under
``quantization(mode)``, the instruction goes through the Alpaca template, and the text after "### Response:" is printed."""
import importlib.util
import sys
import time
from pathlib import Path
from typing import Optional

import torch

wd = Path(__file__).absolute().parent.parent
sys.path.insert(0, str(wd))

from lit_llama import Tokenizer  # noqa: E402
from lit_llama.adapter import LLaMA  # noqa: E402
from lit_llama.utils import EmptyInitOnDevice, lazy_load, llama_model_lookup, stream_load_state_dict  # noqa: E402
import generate as _gen  # noqa: E402  (generate.py next to this directory: the module wins over the namespace package)


def _sibling(name: str):
    spec = importlib.util.spec_from_file_location(f"llb_generate_{name}", Path(__file__).absolute().parent / f"{name}.py")
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


generate_prompt = _sibling("full").generate_prompt      # ref scripts/prepare_alpaca.py:111-125


def load_adapter_model(pretrained_path: Path, adapter_path: Path, quantize: Optional[str], device, dtype, v2: bool = False) -> LLaMA:
    """ref generate/adapter.py:57-68 (v2: generate/adapter_v2.py:64-76): the model named by the pretrained checkpoint's
    embedding width, built under the quantization context; pretrained weights first, then the adapter's, both non-strict.
    GPTQ checkpoints stream disk -> pinned -> HBM with the repack under the read."""
    with lazy_load(pretrained_path) as pretrained_checkpoint:
        name = llama_model_lookup(pretrained_checkpoint)
        with EmptyInitOnDevice(device=device, dtype=dtype, quantization_mode=quantize):
            model = LLaMA.from_name(name)
            if v2:
                from lit_llama.adapter_v2 import add_adapter_v2_parameters_to_linear_layers

                add_adapter_v2_parameters_to_linear_layers(model)
        if quantize is not None and quantize.startswith("gptq") and torch.device(device).type == "cuda":
            stream_load_state_dict(model, pretrained_path, strict=False)
        else:
            model.load_state_dict(pretrained_checkpoint, strict=False)
    with lazy_load(adapter_path) as adapter_checkpoint:
        model.load_state_dict(adapter_checkpoint, strict=False)
    return model.eval()


def run(prompt, input, adapter_path, pretrained_path, tokenizer_path, quantize, max_new_tokens, top_k, temperature, v2):
    adapter_path, pretrained_path, tokenizer_path = Path(adapter_path), Path(pretrained_path), Path(tokenizer_path)
    assert adapter_path.is_file()
    assert pretrained_path.is_file()
    assert tokenizer_path.is_file()
    if not torch.cuda.is_available():
        raise RuntimeError("this implementation runs on a CUDA device (sm_100a): there is no CPU fallback")
    device, dtype = torch.device("cuda"), torch.bfloat16
    print("Loading model ...", file=sys.stderr)
    t0 = time.time()
    model = load_adapter_model(pretrained_path, adapter_path, quantize, device, dtype, v2)
    print(f"Time to load model: {time.time() - t0:.02f} seconds.", file=sys.stderr)
    tokenizer = Tokenizer(tokenizer_path)
    encoded = tokenizer.encode(generate_prompt({"instruction": prompt, "input": input}), bos=True, eos=False, device=device)
    prompt_length = encoded.size(0)
    t0 = time.perf_counter()
    y = _gen.generate(model, encoded, max_new_tokens, temperature=temperature, top_k=top_k, eos_id=tokenizer.eos_id)
    torch.cuda.synchronize()
    t = time.perf_counter() - t0
    model.reset_cache()
    output = tokenizer.decode(y)
    print(output.split("### Response:")[1].strip())
    tokens_generated = y.size(0) - prompt_length
    print(f"\n\nTime for inference: {t:.02f} sec total, {tokens_generated / t:.02f} tokens/sec", file=sys.stderr)
    print(f"Memory used: {torch.cuda.max_memory_reserved() / 1e9:.02f} GB", file=sys.stderr)


def main(
    prompt: str = "What food do lamas eat?",
    input: str = "",
    adapter_path: Path = Path("out/adapter/alpaca/lit-llama-adapter-finetuned.pth"),
    pretrained_path: Path = Path("checkpoints/lit-llama/7B/lit-llama.pth"),
    tokenizer_path: Path = Path("checkpoints/lit-llama/tokenizer.model"),
    quantize: Optional[str] = None,
    max_new_tokens: int = 100,
    top_k: int = 200,
    temperature: float = 0.8,
) -> None:
    """ref generate/adapter.py:20-106 - same flags and output."""
    run(prompt, input, adapter_path, pretrained_path, tokenizer_path, quantize, max_new_tokens, top_k, temperature, v2=False)


def cli(entry, default_adapter: str) -> None:
    import argparse

    ap = argparse.ArgumentParser(description=entry.__doc__)
    ap.add_argument("--prompt", default="What food do lamas eat?")
    ap.add_argument("--input", default="")
    ap.add_argument("--adapter_path", type=Path, default=Path(default_adapter))
    ap.add_argument("--pretrained_path", type=Path, default=Path("checkpoints/lit-llama/7B/lit-llama.pth"))
    ap.add_argument("--tokenizer_path", type=Path, default=Path("checkpoints/lit-llama/tokenizer.model"))
    ap.add_argument("--quantize", default=None)
    ap.add_argument("--max_new_tokens", type=int, default=100)
    ap.add_argument("--top_k", type=int, default=200)
    ap.add_argument("--temperature", type=float, default=0.8)
    a = ap.parse_args()
    entry(a.prompt, a.input, a.adapter_path, a.pretrained_path, a.tokenizer_path, a.quantize, a.max_new_tokens, a.top_k,
          a.temperature)


if __name__ == "__main__":
    cli(main, "out/adapter/alpaca/lit-llama-adapter-finetuned.pth")
