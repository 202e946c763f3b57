"""Perplexity of a (quantized) checkpoint, GPTQ-style (ref: /root/reference/evaluate/full.py:45-141): the
text is tokenised once, cut into windows of 2048 tokens, each window goes through ``model(inp)`` WITHOUT a KV cache - the
prefill path (large-M quantized Linear: fused de-quantise + tcgen05 GEMM up to 256 rows, unpack + tcgen05 GEMM beyond) -
and the summed next-token negative log-likelihood gives exp(nll / tokens).

The reference downloads wikitext / ptb / c4 through ``datasets``; there is no network here, so the text comes from local
files (``--data_paths a.txt,b.txt``).  ``perplexity()`` is the loop itself, usable on any token tensor."""
import math
import sys
import time
from pathlib import Path
from typing import Optional

import torch

wd = Path(__file__).absolute().parent.parent
sys.path.insert(0, str(wd))

from lit_llama import LLaMA, Tokenizer  # noqa: E402
from lit_llama.utils import EmptyInitOnDevice, lazy_load, stream_load_state_dict  # noqa: E402


@torch.inference_mode()
def perplexity(model: LLaMA, encoded_text: torch.Tensor, block_size: int = 2048):
    """ref evaluate/full.py:104-125.  ``encoded_text`` [1, n] token ids; windows of ``block_size`` (2048 as in GPTQ's
    evaluation, not the model's), trimmed to 256 * model.config.block_size tokens like the reference.  The cross entropy
    is taken in fp32 over the window's logits.  Returns (perplexity, tokens scored)."""
    encoded_text = encoded_text[:, : 256 * model.config.block_size]
    nlls, toks = 0.0, 0
    for i in range(0, encoded_text.shape[1], block_size):
        inp = encoded_text[:, i:i + block_size]
        if inp.size(1) < 2:
            break
        logits = model(inp)[0]
        nll = torch.nn.functional.cross_entropy(logits[:-1].float(), inp[0, 1:].to(dtype=torch.long), reduction="sum")
        toks += inp.size(1) - 1
        nlls += nll.item()
    return math.exp(nlls / max(toks, 1)), toks


def main(
    data_paths: str = "data/wikitext-2-raw/wiki.test.raw",
    *,
    checkpoint_path: Optional[Path] = None,
    tokenizer_path: Path = Path("checkpoints/lit-llama/tokenizer.model"),
    model_size: str = "7B",
    dtype: str = "bfloat16",
    quantize: Optional[str] = None,
) -> None:
    """ref evaluate/full.py:45-141 - one perplexity per text file (comma-separated ``data_paths`` in place of dataset names)."""
    if not checkpoint_path:
        checkpoint_path = Path(f"checkpoints/lit-llama/{model_size}/lit-llama.pth")
    checkpoint_path, tokenizer_path = Path(checkpoint_path), Path(tokenizer_path)
    assert checkpoint_path.is_file() and tokenizer_path.is_file()
    dt = getattr(torch, dtype, None)
    if not isinstance(dt, torch.dtype):
        raise ValueError(f"{dtype} is not a valid dtype.")
    if not torch.cuda.is_available():
        raise RuntimeError("this implementation runs on a CUDA device (sm_100a): there is no CPU fallback")
    device = torch.device("cuda")
    print("Loading model ...", file=sys.stderr)
    t0 = time.time()
    with EmptyInitOnDevice(device=device, dtype=dt, quantization_mode=quantize):
        model = LLaMA.from_name(model_size)
    if quantize is not None and quantize.startswith("gptq"):
        stream_load_state_dict(model, checkpoint_path, strict=True, prepare_batch=None)   # perplexity never decodes
    else:
        with lazy_load(checkpoint_path) as checkpoint:
            model.load_state_dict(checkpoint)
    print(f"Time to load model: {time.time() - t0:.02f} seconds.", file=sys.stderr)
    model.eval()
    tokenizer = Tokenizer(tokenizer_path)
    total_toks, t0 = 0, time.perf_counter()
    for path in data_paths.split(","):
        text = Path(path).read_text()
        encoded = tokenizer.encode(text, bos=True, eos=False, device=device)[None]
        ppl, toks = perplexity(model, encoded)
        print(f"Perplexity on {path}: {ppl:.2f}")
        total_toks += toks
    torch.cuda.synchronize()
    t = time.perf_counter() - t0
    print(f"\n\nTime for inference: {t:.02f} sec total, {total_toks / t:.02f} tokens/sec", file=sys.stderr)
    print(f"Memory used: {torch.cuda.max_memory_reserved() / 1e9:.02f} GB", file=sys.stderr)


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("--data_paths", default="data/wikitext-2-raw/wiki.test.raw")
    ap.add_argument("--checkpoint_path", type=Path, default=None)
    ap.add_argument("--tokenizer_path", type=Path, default=Path("checkpoints/lit-llama/tokenizer.model"))
    ap.add_argument("--model_size", default="7B")
    ap.add_argument("--dtype", default="bfloat16")
    ap.add_argument("--quantize", default=None)
    a = ap.parse_args()
    main(a.data_paths, checkpoint_path=a.checkpoint_path, tokenizer_path=a.tokenizer_path, model_size=a.model_size,
         dtype=a.dtype, quantize=a.quantize)
