"""Perplexity of a LLaMA-Adapter fine-tune (ref: /root/reference/evaluate/adapter.py:50-160): as evaluate/full.py, on
``lit_llama.adapter.LLaMA`` with the pretrained and the adapter checkpoint loaded non-strictly, the text wrapped in the Alpaca
instruction template first (the reference's ``instruction_tuning = True``, evaluate/adapter.py:24).  Text comes from local
files (``--data_paths a.txt,b.txt``): there is no network for ``datasets``."""
import importlib.util
import sys
import time
from pathlib import Path
from typing import Optional

import torch

wd = Path(__file__).absolute().parent.parent
sys.path.insert(0, str(wd))

from lit_llama import Tokenizer  # noqa: E402


def _load(rel: str, name: str):
    spec = importlib.util.spec_from_file_location(name, wd / rel)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


perplexity = _load("evaluate/full.py", "llb_evaluate_full").perplexity
_ga = _load("generate/adapter.py", "llb_generate_adapter")
load_adapter_model, generate_prompt = _ga.load_adapter_model, _ga.generate_prompt

instruction_tuning = True     # ref evaluate/adapter.py:24


def run(data_paths, accelerator, adapter_path, checkpoint_path, tokenizer_path, dtype, quantize, v2):
    adapter_path, checkpoint_path, tokenizer_path = Path(adapter_path), Path(checkpoint_path), Path(tokenizer_path)
    assert adapter_path.is_file()
    assert checkpoint_path.is_file()
    assert tokenizer_path.is_file()
    dt = getattr(torch, dtype, None)
    if not isinstance(dt, torch.dtype):
        raise ValueError(f"{dtype} is not a valid dtype.")
    if accelerator not in ("auto", "cuda", "gpu") or not torch.cuda.is_available():
        raise RuntimeError("this implementation runs on a CUDA device (sm_100a): there is no CPU fallback")
    device = torch.device("cuda")
    print("Loading model ...", file=sys.stderr)
    t0 = time.time()
    model = load_adapter_model(checkpoint_path, adapter_path, quantize, device, dt, v2)
    print(f"Time to load model: {time.time() - t0:.02f} seconds.", file=sys.stderr)
    tokenizer = Tokenizer(tokenizer_path)
    total_toks, t0 = 0, time.perf_counter()
    for path in data_paths.split(","):
        text = Path(path).read_text()
        if instruction_tuning:
            text = generate_prompt({"instruction": text, "input": ""})
        encoded = tokenizer.encode(text, bos=True, eos=False, device=device)[None]
        ppl, toks = perplexity(model, encoded)
        print(f"Perplexity on {path}: {ppl:.2f}")
        total_toks += toks
    torch.cuda.synchronize()
    t = time.perf_counter() - t0
    print(f"\n\nTime for inference: {t:.02f} sec total, {total_toks / t:.02f} tokens/sec", file=sys.stderr)
    print(f"Memory used: {torch.cuda.max_memory_reserved() / 1e9:.02f} GB", file=sys.stderr)


def main(
    data_paths: str = "data/wikitext-2-raw/wiki.test.raw",
    *,
    accelerator: str = "auto",
    adapter_path: Path = Path("out/adapter/alpaca/lit-llama-adapter-finetuned.pth"),
    checkpoint_path: Path = Path("checkpoints/lit-llama/7B/lit-llama.pth"),
    tokenizer_path: Path = Path("checkpoints/lit-llama/tokenizer.model"),
    dtype: str = "float32",
    quantize: Optional[str] = None,
) -> None:
    """ref evaluate/adapter.py:50-160 - one perplexity per text file (comma-separated ``data_paths`` in place of dataset names)."""
    run(data_paths, accelerator, adapter_path, checkpoint_path, tokenizer_path, dtype, quantize, v2=False)


def cli(entry, default_adapter: str) -> None:
    import argparse

    ap = argparse.ArgumentParser(description=entry.__doc__)
    ap.add_argument("--data_paths", default="data/wikitext-2-raw/wiki.test.raw")
    ap.add_argument("--accelerator", default="auto")
    ap.add_argument("--adapter_path", type=Path, default=Path(default_adapter))
    ap.add_argument("--checkpoint_path", type=Path, default=Path("checkpoints/lit-llama/7B/lit-llama.pth"))
    ap.add_argument("--tokenizer_path", type=Path, default=Path("checkpoints/lit-llama/tokenizer.model"))
    ap.add_argument("--dtype", default="float32")
    ap.add_argument("--quantize", default=None)
    a = ap.parse_args()
    entry(a.data_paths, accelerator=a.accelerator, adapter_path=a.adapter_path, checkpoint_path=a.checkpoint_path,
          tokenizer_path=a.tokenizer_path, dtype=a.dtype, quantize=a.quantize)


if __name__ == "__main__":
    cli(main, "out/adapter/alpaca/lit-llama-adapter-finetuned.pth")
