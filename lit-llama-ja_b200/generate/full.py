"""Sampling from a fully fine-tuned checkpoint (ref: /root/reference/generate/full.py:21-91): the same flags,
the Alpaca instruction template around the prompt (ref scripts/prepare_alpaca.py:111-125), the model built under
``quantization(mode)`` and decoded by this repository's ``generate()`` (device-side sampling loop, fused decode step)."""
import sys
import time
from pathlib import Path
from typing import Optional

import torch

wd = Path(__file__).absolute().parent.parent
sys.path.insert(0, str(wd))

from lit_llama import LLaMA, Tokenizer  # noqa: E402
from lit_llama.utils import EmptyInitOnDevice, lazy_load, stream_load_state_dict  # noqa: E402
import generate as _gen  # noqa: E402  (generate.py next to this directory: the module wins over the namespace package)


def generate_prompt(example: dict) -> str:
    """ref scripts/prepare_alpaca.py:111-125: instruction (+ optional input) -> the prompt the fine-tuned model was trained on."""
    if example.get("input"):
        return ("Below is an instruction that describes a task, paired with an input that provides further context. "
                "Write a response that appropriately completes the request.\n\n"
                f"### Instruction:\n{example['instruction']}\n\n### Input:\n{example['input']}\n\n### Response:")
    return ("Below is an instruction that describes a task. "
            "Write a response that appropriately completes the request.\n\n"
            f"### Instruction:\n{example['instruction']}\n\n### Response:")


def load_full_model(checkpoint_path: Path, model_size: str, quantize: Optional[str], device, dtype) -> LLaMA:
    """ref generate/full.py:59-66: ``LLaMA.from_name(model_size)`` under the quantization context, then the state dict.
    GPTQ checkpoints stream disk -> pinned -> HBM with the repack under the read; others load from lazy placeholders."""
    with EmptyInitOnDevice(device=device, dtype=dtype, quantization_mode=quantize):
        model = LLaMA.from_name(model_size)
    if quantize is not None and quantize.startswith("gptq") and torch.device(device).type == "cuda":
        stream_load_state_dict(model, checkpoint_path, strict=True)
    else:
        with lazy_load(checkpoint_path) as checkpoint:
            model.load_state_dict(checkpoint)
    return model.eval()


def main(
    prompt: str = "Hello, my name is",
    *,
    input: str = "",
    num_samples: int = 1,
    max_new_tokens: int = 50,
    top_k: int = 200,
    temperature: float = 0.8,
    checkpoint_path: Optional[Path] = None,
    tokenizer_path: Path = Path("checkpoints/lit-llama/tokenizer.model"),
    model_size: str = "7B",
    quantize: Optional[str] = None,
) -> None:
    """ref generate/full.py:21-91 - same flags and output (``input`` is the template's optional context: the reference
    passes the Python builtin there by accident, generate/full.py:74)."""
    if not checkpoint_path:
        checkpoint_path = Path(f"checkpoints/lit-llama/{model_size}/lit-llama.pth")
    checkpoint_path, tokenizer_path = Path(checkpoint_path), Path(tokenizer_path)
    assert checkpoint_path.is_file(), checkpoint_path
    assert tokenizer_path.is_file(), tokenizer_path
    if not torch.cuda.is_available():
        raise RuntimeError("this implementation runs on a CUDA device (sm_100a): there is no CPU fallback")
    device, dtype = torch.device("cuda"), torch.bfloat16
    print("Loading model ...", file=sys.stderr)
    t0 = time.time()
    model = load_full_model(checkpoint_path, model_size, quantize, device, dtype)
    print(f"Time to load model: {time.time() - t0:.02f} seconds.", file=sys.stderr)
    tokenizer = Tokenizer(tokenizer_path)
    encoded = tokenizer.encode(generate_prompt({"instruction": prompt, "input": input}), bos=True, eos=False, device=device)
    prompt_length = encoded.size(0)
    torch.manual_seed(1234)
    for i in range(num_samples):
        t0 = time.perf_counter()
        y = _gen.generate(model, encoded, max_new_tokens, temperature=temperature, top_k=top_k)
        torch.cuda.synchronize()
        t = time.perf_counter() - t0
        model.reset_cache()
        print(tokenizer.decode(y))
        tokens_generated = y.size(0) - prompt_length
        print(f"Time for inference {i + 1}: {t:.02f} sec total, {tokens_generated / t:.02f} tokens/sec", file=sys.stderr)
    print(f"Memory used: {torch.cuda.max_memory_reserved() / 1e9:.02f} GB", file=sys.stderr)


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("--prompt", default="Hello, my name is")
    ap.add_argument("--input", default="")
    ap.add_argument("--num_samples", type=int, default=1)
    ap.add_argument("--max_new_tokens", type=int, default=50)
    ap.add_argument("--top_k", type=int, default=200)
    ap.add_argument("--temperature", type=float, default=0.8)
    ap.add_argument("--checkpoint_path", type=Path, default=None)
    ap.add_argument("--tokenizer_path", type=Path, default=Path("checkpoints/lit-llama/tokenizer.model"))
    ap.add_argument("--model_size", default="7B")
    ap.add_argument("--quantize", default=None)
    a = ap.parse_args()
    main(a.prompt, input=a.input, num_samples=a.num_samples, max_new_tokens=a.max_new_tokens, top_k=a.top_k,
         temperature=a.temperature, checkpoint_path=a.checkpoint_path, tokenizer_path=a.tokenizer_path,
         model_size=a.model_size, quantize=a.quantize)
