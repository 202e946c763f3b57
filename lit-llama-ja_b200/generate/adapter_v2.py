"""Sampling from a LLaMA-Adapter v2 fine-tune (ref: /root/reference/generate/adapter_v2.py:22-108): as
generate/adapter.py, with the per-Linear ``adapter_bias`` / ``adapter_scale`` parameters added to the model before the two
checkpoints are loaded (ref :71-73)."""
import importlib.util
from pathlib import Path
from typing import Optional

_spec = importlib.util.spec_from_file_location("llb_generate_adapter", Path(__file__).absolute().parent / "adapter.py")
_v1 = importlib.util.module_from_spec(_spec)
_spec.loader.exec_module(_v1)
load_adapter_model, generate_prompt = _v1.load_adapter_model, _v1.generate_prompt


def main(
    prompt: str = "What food do lamas eat?",
    input: str = "",
    adapter_path: Path = Path("out/adapter_v2/alpaca/lit-llama-adapter-finetuned.pth"),
    pretrained_path: Path = Path("checkpoints/lit-llama/7B/lit-llama.pth"),
    tokenizer_path: Path = Path("checkpoints/lit-llama/tokenizer.model"),
    quantize: Optional[str] = None,
    max_new_tokens: int = 100,
    top_k: int = 200,
    temperature: float = 0.8,
) -> None:
    """ref generate/adapter_v2.py:22-108 - same flags and output."""
    _v1.run(prompt, input, adapter_path, pretrained_path, tokenizer_path, quantize, max_new_tokens, top_k, temperature, v2=True)


if __name__ == "__main__":
    _v1.cli(main, "out/adapter_v2/alpaca/lit-llama-adapter-finetuned.pth")
