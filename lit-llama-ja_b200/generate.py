"""Text generation with the B200-native quantized LLaMA (ref: /root/reference/generate.py).

Same ``generate(model, idx, max_new_tokens, *, max_seq_length, temperature, top_k, eos_id)``
signature and sampling semantics, same ``main`` flags (``--quantize gptq.int4|llm.int8`` ...).
Lightning Fabric is not needed: device and dtype are handled directly.
"""
from __future__ import annotations

import sys
import time
from pathlib import Path
from typing import Optional

import torch

sys.path.insert(0, str(Path(__file__).parent.resolve()))

from lit_llama import LLaMA, HFTokenizer, Tokenizer  # noqa: E402
from lit_llama.utils import EmptyInitOnDevice, lazy_load, llama_model_lookup  # noqa: E402


def _generate_on_device(model, idx, max_new_tokens, max_seq_length, temperature, top_k, eos_id):
    """The whole sampling loop on the device for the fused int4/bf16 engine: per new token ONE CUDA-graph replay of
    [top-k draw (csrc/sampling.cu) -> advance position -> fused decode step], no host synchronisation.  Returns None
    when the model is not eligible (then the per-op loop below runs).  Sampling semantics are the reference's
    (generate.py:62-75) step for step; the random stream is torch's generator consumed as one uniform per token
    instead of torch.multinomial's per-entry exponentials, so tokens differ from the reference for equal seeds."""
    if idx.device.type != "cuda" or not model.use_fused_decode or max_new_tokens < 1 or temperature <= 0:
        return None
    n_prompt = idx.size(0)
    if n_prompt + max_new_tokens > model.config.block_size + 1:
        return None
    model._ensure_caches(idx)
    eng = model._fused_engine(1, max_seq_length)
    if eng is not None:
        eng.ensure_sample_graph(temperature, top_k)      # before the prefill: its warm-up runs scribble on the cache
    input_pos = torch.arange(0, n_prompt, device=idx.device)
    logits = model(idx.view(1, -1), max_seq_length, input_pos)[0, -1]
    if eng is None or logits.dtype != torch.bfloat16 or not eng.begin_sampling(idx, logits, max_new_tokens):
        return _sample_loop(model, idx, logits, max_new_tokens, max_seq_length, temperature, top_k, eos_id)
    for _ in range(max_new_tokens - 1):
        eng.sample_and_step(temperature, top_k)
    eng.sample_last(temperature, top_k)
    out = eng.out_tokens[: n_prompt + max_new_tokens].to(idx.dtype)
    if eos_id is not None:   # ref generate.py:82-84: stop at the first eos; the returned slice ends before its slot
        hit = (out[n_prompt:] == eos_id).nonzero()
        if hit.numel():
            out = out[: n_prompt + int(hit[0])]
    return out


def _sample_loop(model, idx, logits, max_new_tokens, max_seq_length, temperature, top_k, eos_id):
    """ref: generate.py:56-89 with the prompt's logits already computed (per-op sampling, any dtype)."""
    n_prompt = idx.size(0)
    out = torch.empty(n_prompt + max_new_tokens, dtype=idx.dtype, device=idx.device)
    out[:n_prompt] = idx
    input_pos = torch.arange(0, n_prompt, device=idx.device)
    for i in range(max_new_tokens):
        if i > 0:
            logits = model(out.index_select(0, input_pos).view(1, -1), max_seq_length, input_pos)[0, -1]
        logits = logits / temperature
        if top_k is not None:
            kth = torch.topk(logits, min(top_k, logits.size(-1))).values[-1:]
            logits = torch.where(logits < kth, -float("Inf"), logits)
        probs = torch.nn.functional.softmax(logits, dim=-1)
        nxt = torch.multinomial(probs, num_samples=1).to(dtype=idx.dtype)
        input_pos = input_pos[-1:] + 1
        out = out.index_copy(0, input_pos, nxt)
        if nxt == eos_id:
            return out[:input_pos]
    return out


@torch.no_grad()
def generate(
    model: LLaMA,
    idx: torch.Tensor,
    max_new_tokens: int,
    *,
    max_seq_length: Optional[int] = None,
    temperature: float = 1.0,
    top_k: Optional[int] = None,
    eos_id: Optional[int] = None,
) -> torch.Tensor:
    """ref: generate.py:18-89.  ``idx`` is a 1-D prompt; returns prompt + generated ids.

    Iteration 0 feeds the whole prompt (prefill), every later iteration one token at the next
    position - which the model runs as one CUDA-graph replay of the fused decode step."""
    n_prompt = idx.size(0)
    total = n_prompt + max_new_tokens
    if max_seq_length is None:
        max_seq_length = min(total, model.config.block_size)
    fast = _generate_on_device(model, idx, max_new_tokens, max_seq_length, temperature, top_k, eos_id)
    if fast is not None:
        return fast
    input_pos = torch.arange(0, n_prompt, device=idx.device)
    logits = model(idx.view(1, -1), max_seq_length, input_pos)[0, -1]
    return _sample_loop(model, idx, logits, max_new_tokens, max_seq_length, temperature, top_k, eos_id)


def load_model(checkpoint_path: Path, quantize: Optional[str], device: torch.device, dtype: torch.dtype) -> LLaMA:
    """ref: generate.py:124-136 (lazy_load + quantization ctx + load_state_dict).  GPTQ checkpoints go through the
    streaming loader (disk -> pinned -> HBM, the GEMV stream of every layer built under the read); llm.int8 quantizes a
    floating-point checkpoint inside ``load_state_dict``, tensor by tensor, from the lazy placeholders."""
    from lit_llama.utils import stream_load_state_dict

    with lazy_load(checkpoint_path) as checkpoint:
        name = llama_model_lookup(checkpoint)
        with EmptyInitOnDevice(device=device, dtype=dtype, quantization_mode=quantize):
            model = LLaMA.from_name(name)
        if torch.device(device).type == "cuda" and quantize is not None and quantize.startswith("gptq"):
            stream_load_state_dict(model, checkpoint_path, strict=True)
        else:
            model.load_state_dict(checkpoint)
    return model.eval()


def main(
    prompt: str = "Hello, my name is",
    *,
    num_samples: int = 1,
    max_new_tokens: int = 50,
    top_k: int = 200,
    temperature: float = 0.8,
    checkpoint_path: Path = Path("checkpoints/lit-llama/7B/lit-llama.pth"),
    tokenizer_path: Path = Path("checkpoints/lit-llama/tokenizer.model"),
    quantize: Optional[str] = None,
) -> None:
    """ref: generate.py:92-155 - same flags and output."""
    checkpoint_path, tokenizer_path = Path(checkpoint_path), Path(tokenizer_path)
    assert checkpoint_path.is_file(), checkpoint_path
    assert tokenizer_path.is_file(), tokenizer_path
    if not torch.cuda.is_available():
        raise RuntimeError("lit-llama-ja_b200 needs a CUDA device (sm_100a); there is no CPU fallback")
    device = torch.device("cuda", torch.cuda.current_device())
    dtype = torch.bfloat16 if torch.cuda.is_bf16_supported() else torch.float32

    print("Loading model ...", file=sys.stderr)
    t0 = time.time()
    model = load_model(checkpoint_path, quantize, device, dtype)
    print(f"Time to load model: {time.time() - t0:.02f} seconds.", file=sys.stderr)

    tokenizer = HFTokenizer(tokenizer_path) if tokenizer_path.suffix == ".json" else Tokenizer(tokenizer_path)
    encoded = tokenizer.encode(prompt, bos=True, eos=False, device=device)
    prompt_length = encoded.size(0)

    torch.manual_seed(1234)
    for i in range(num_samples):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        y = generate(model, encoded, max_new_tokens, temperature=temperature, top_k=top_k)
        torch.cuda.synchronize()
        t = time.perf_counter() - t0
        model.reset_cache()
        print(tokenizer.decode(y))
        n_new = y.size(0) - prompt_length
        print(f"Time for inference {i + 1}: {t:.02f} sec total, {n_new / t:.02f} tokens/sec", file=sys.stderr)
    print(f"Memory used: {torch.cuda.max_memory_reserved() / 1e9:.02f} GB", file=sys.stderr)


if __name__ == "__main__":
    import argparse

    ap = argparse.ArgumentParser(description=main.__doc__)
    ap.add_argument("--prompt", default="Hello, my name is")
    ap.add_argument("--num_samples", type=int, default=1)
    ap.add_argument("--max_new_tokens", type=int, default=50)
    ap.add_argument("--top_k", type=int, default=200)
    ap.add_argument("--temperature", type=float, default=0.8)
    ap.add_argument("--checkpoint_path", type=Path, default=Path("checkpoints/lit-llama/7B/lit-llama.pth"))
    ap.add_argument("--tokenizer_path", type=Path, default=Path("checkpoints/lit-llama/tokenizer.model"))
    ap.add_argument("--quantize", default=None, choices=[None, "llm.int8", "gptq.int4", "gptq.int8"])
    a = ap.parse_args()
    main(a.prompt, num_samples=a.num_samples, max_new_tokens=a.max_new_tokens, top_k=a.top_k,
         temperature=a.temperature, checkpoint_path=a.checkpoint_path, tokenizer_path=a.tokenizer_path,
         quantize=a.quantize)
